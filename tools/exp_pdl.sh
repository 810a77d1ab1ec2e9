python -m pytest tests -m gpu -x -q > gpurun_out/v8e_gpu_tests.log 2>&1; echo tests rc=$?; tail -2 gpurun_out/v8e_gpu_tests.log
B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0"
for pdl in 1 0 1 0; do
  E2SEC_PDL=$pdl $B > gpurun_out/v8e_pdl$pdl.json 2>/dev/null
  python -c "
import json
d=json.load(open('gpurun_out/v8e_pdl$pdl.json')); print('pdl=$pdl', round(d['value']), d['ms_per_step'], d['stages_ms_last_step'])"
done
