"""Per-batch stage times of the benchmark stream (one synchronous batch at a time): shows which batches take the exact
O(n^2) fallback and what it costs.  python tools/stage_trace.py [batches]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from manipulationlibrary_b200 import _native as nat          # noqa: E402
from manipulationlibrary_b200.engine import Engine             # noqa: E402
from manipulationlibrary_b200.synthetic import SyntheticStream  # noqa: E402

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 6
B = 100
s = SyntheticStream(640, 480, 10, config=2)
p, keep = nat.make_params(**s.params(filter_mode="closure", filter_seed=20261104, radius_mode="bounded"))
eng = Engine(0)
eng.configure(p, B, s.N, keep)
K = int(p.n_entities)
for b in range(nb):
    ids = np.arange(b * B, (b + 1) * B)
    xyz, lab = s.fetch(ids)
    dx, dl = torch.from_numpy(xyz).cuda(), torch.from_numpy(lab).cuda()
    torch.cuda.synchronize()
    rec = np.empty((B, nat.record_doubles(K)))
    for rep in range(2):
        eng.reset_sequence()
        eng.process_batch_device(dx.data_ptr(), dl.data_ptr(), B, ids, s.N, records=rec)
    st = eng.stage_ms()
    ent = rec[:, nat.REC_HEADER: nat.REC_HEADER + nat.REC_ENTITY * K].reshape(B, K, nat.REC_ENTITY)
    n_exact = int(((ent[:, :, 2] >= 0) & ~np.isnan(ent[:, :, 3])).sum())
    print("batch %d exact=%d total=%.3f " % (b, n_exact, sum(st.values())), {k: round(v, 3) for k, v in st.items()}, flush=True)
eng.close()
