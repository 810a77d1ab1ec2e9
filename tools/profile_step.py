"""A few batches of the hot path on one context, for ncu:  python tools/profile_step.py [config] [frames] [batches]
(profiles/: `ncu --set full --import-source on --clock-control none -k regex:... --launch-skip N --launch-count M -o ... python tools/profile_step.py`)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from bench import CONFIGS, generate_pool, make_stream  # noqa: E402
from manipulationlibrary_b200 import _native as nat  # noqa: E402
from manipulationlibrary_b200.engine import Engine, Phase2  # noqa: E402

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 2
F = int(sys.argv[2]) if len(sys.argv) > 2 else CONFIGS[cfg]["batch"]
batches = int(sys.argv[3]) if len(sys.argv) > 3 else 3
s = make_stream(cfg)
xyz, lab = generate_pool(s, 0, F)
dx, dl = torch.from_numpy(xyz).cuda(), torch.from_numpy(lab).cuda()
p, keep = nat.make_params(**s.params())
eng = Engine(0)
eng.configure(p, F, s.N, keep)
p2 = Phase2(eng, max_columns=F * batches + 8)
for b in range(batches):
    ids = np.arange(b * F, (b + 1) * F, dtype=np.int32)
    eng.process_batch_device_async(dx.data_ptr(), dl.data_ptr(), F, ids, s.N)
    p2.run_engine_batch(eng, F)
eng.sync()
print("columns", p2.sync(), "stages", eng.stage_ms(), "stream kernel ms", eng.seg_stream_ms)
