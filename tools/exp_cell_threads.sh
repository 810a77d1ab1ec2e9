set -x
python -m pytest tests -m gpu -x -q > gpurun_out/v8c_gpu_tests.log 2>&1; echo tests rc=$?; tail -2 gpurun_out/v8c_gpu_tests.log
B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0"
$B > gpurun_out/v8c_t512.json 2>/dev/null
for cfg in "256 3" "256 4" "384 2" "1024 1"; do
  set -- $cfg
  E2SEC_NVCC_FLAGS="-DE2_CELL_THREADS=$1 -DE2_CELL_CTAS=$2" python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
  $B > gpurun_out/v8c_t$1_$2.json 2>/dev/null
done
python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
for f in gpurun_out/v8c_t*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); print(d['value'], d['ms_per_step'], d['stages_ms_last_step'])"; done
