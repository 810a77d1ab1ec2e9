"""Lists the filtered entities of a synthetic stream block that needed the exact O(n^2) path in BOUNDED radius mode
(record radius field not NaN), with their cell-table bounds.  GPU only: python tools/exact_path_report.py [first] [count]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from manipulationlibrary_b200 import _native as nat            # noqa: E402
from manipulationlibrary_b200.engine import Engine, rec_entity, rec_radius_bounds   # noqa: E402
from manipulationlibrary_b200.synthetic import SyntheticStream  # noqa: E402

first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
count = int(sys.argv[2]) if len(sys.argv) > 2 else 200
s = SyntheticStream()
K = len(s.total_unique_labels)
eng = Engine(0)
res = {}
for rm in ("bounded", "exact"):
    p, keep = nat.make_params(**s.params(radius_mode=rm))
    eng.configure(p, 50, s.N, keep)
    out = []
    for b in range(first, first + count, 50):
        ids = np.arange(b, min(b + 50, first + count))
        xyz, lab = s.fetch(ids)
        out.append(eng.process_batch(xyz, lab, ids))
    res[rm] = np.concatenate(out)
nf = ne = 0
for rb, rx in zip(res["bounded"], res["exact"]):
    bb = rec_radius_bounds(rb, K)
    for e in range(K):
        a, x = rec_entity(rb, e), rec_entity(rx, e)
        if a["seed"] < 0:
            continue
        nf += 1
        if not np.isnan(a["radius"]):
            ne += 1
            print("frame %d entity %d n_raw %d n_out %d radius %.6f" % (rb[0], e, a["n_raw"], a["n_out"], a["radius"]))
        else:
            lo, hi = bb[e]
            assert np.float32(lo) <= np.float32(x["radius"]) <= np.float32(hi), (rb[0], e, lo, x["radius"], hi)
            if (hi - lo) / x["radius"] > 0.5:
                print("frame %d entity %d wide bounds lo %.4f r %.4f hi %.4f" % (rb[0], e, lo, x["radius"], hi))
        assert a["n_out"] == x["n_out"] and (a["lo"] == x["lo"]).all() and (a["hi"] == x["hi"]).all() and \
            (a["median"] == x["median"]).all(), ("bounded != exact", rb[0], e)
print("%d filtered entities, %d took the exact path; bounded == exact on counts, boxes and medians" % (nf, ne))
