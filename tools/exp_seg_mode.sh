python -m pytest tests -m gpu -x -q > gpurun_out/v12_gpu_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/v12_gpu_tests.log
for m in stream cluster stream cluster; do
  E2SEC_SEG_MODE=$m python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$m', round(d['value']), d['ms_per_step'], d['stages_ms_last_step'])"
done
