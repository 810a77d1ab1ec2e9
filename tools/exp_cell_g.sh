# cells along the largest robust extent: speed of the radius stage vs entities the cell table cannot decide
B="python bench.py --steps 10 --warmup 3 --no-e2e --cpu-frames 0 --blocks 6"
for g in 12 16 20 24; do
  E2SEC_NVCC_FLAGS="-DE2_CELL_G=$g" python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
  $B > gpurun_out/exp_g$g.json 2>/dev/null
  python -c "
import json
d=json.load(open('gpurun_out/exp_g$g.json')); s=d['stages_ms_solo_step']
print('G=$g', round(d['value']), 'radius', round(s['radius'],3), 'filter', round(s['filter'],3), 'exact', d['config']['exact_path_entities'], 'of', d['config']['filtered_entities'], 'cols', d['esec_columns'])"
done
python -m manipulationlibrary_b200.build --force > /dev/null 2>&1
