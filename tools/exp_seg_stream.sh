B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0"
for cfg in "3 16 7" "4 16 6" "2 16 8" "3 8 7" "3 32 7" "4 32 5" "3 16 5"; do
  set -- $cfg
  E2SEC_NVCC_FLAGS="-DE2_SS_STAGES=$1 -DE2_SS_UNIT_TILES=$2 -DE2_SS_CTAS=$3" python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
  $B | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('stages $1 unit_tiles $2 ctas $3', round(d['value']), 'segment', round(d['stages_ms_last_step']['segment'],3), 'stream', round(d['roofline']['streaming_kernel']['ms'],3), round(d['roofline']['streaming_kernel']['frac'],3))"
done
python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
