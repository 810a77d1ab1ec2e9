B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0"
for cfg in "384 3" "352 3" "416 3" "288 3" "256 3"; do
  set -- $cfg
  E2SEC_NVCC_FLAGS="-DE2_CELL_THREADS=$1 -DE2_CELL_CTAS=$2" python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
  $B | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('threads $1 ctas $2', round(d['value']), d['ms_per_step'], round(d['stages_ms_last_step']['radius'],3))"
done
python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
