B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0"
for c in 8 4 2; do for b in 100 142 148; do
  E2SEC_SF_CLUSTER=$c $B --batch $b --pool $((3*b)) > gpurun_out/v8d_c${c}_b${b}.json 2>/dev/null
  python -c "
import json
d=json.load(open('gpurun_out/v8d_c${c}_b${b}.json')); s=d['stages_ms_last_step']; b=$b
print('C=$c batch=$b', round(d['value']), 'seg us/frame', round(1e3*s['segment']/b,3), 'radius', round(1e3*s['radius']/b,3), 'pairs', round(1e3*s['pairs']/b,3))"
done; done
