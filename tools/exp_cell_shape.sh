# CTA shape of cell_filter_kernel: threads x resident CTAs per SM (shared memory per CTA follows from the cell table size)
B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0 --blocks 6"
for cfg in "384 3 1024 12" "384 3 768 11" "256 4 768 11" "288 4 768 11" "256 5 512 11" "320 3 1024 12"; do
  set -- $cfg
  E2SEC_NVCC_FLAGS="-DE2_CELL_THREADS=$1 -DE2_CELL_CTAS=$2 -DE2_CELL_MAX=$3 -DE2_CELL_HASH_LOG=$4" python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
  $B > gpurun_out/exp_shape.json 2>/dev/null
  python -c "
import json
d=json.load(open('gpurun_out/exp_shape.json')); s=d['stages_ms_solo_step']
print('threads $1 ctas $2 cells $3 hash 2^$4:', round(d['value']), 'radius', round(s['radius'],3), 'filter', round(s['filter'],3), 'exact', d['config']['exact_path_entities'])"
done
python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
