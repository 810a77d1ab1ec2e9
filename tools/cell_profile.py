"""Phase split of cell_filter_kernel (clock ticks of thread 0, summed over CTAs).  Build with
E2SEC_NVCC_FLAGS=-DE2_CELL_PROFILE python -m manipulationlibrary_b200.build --force, run on the GPU, rebuild without."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from manipulationlibrary_b200 import _native as nat  # noqa: E402
from manipulationlibrary_b200.engine import Engine  # noqa: E402
from bench import CONFIGS, make_stream  # noqa: E402

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 2
F = int(sys.argv[2]) if len(sys.argv) > 2 else 100
s = make_stream(cfg)
p, keep = nat.make_params(**s.params())
eng = Engine(0)
eng.configure(p, F, s.N, keep)
ids = np.arange(30, 30 + F, dtype=np.int32)
xyz, lab = s.fetch(ids)
fn = nat.lib.e2sec_debug_cell_profile
fn.argtypes = [C.c_void_p, C.c_int]
out = np.zeros(16, np.uint64)
eng.process_batch(xyz, lab, ids)
fn(None, 1)
eng.process_batch(xyz, lab, ids)
fn(nat.ptr(out), 0)
names = ["M medians", "A seed+MAD", "B1/B2 hash", "B3 boxes", "C pair hist", "D rank", "E closure", "F candidates", "G compaction"]
tot = float(out[:9].sum())
print("config %d, %d frames; stages ms:" % (cfg, F), eng.stage_ms())
for n, v in zip(names, out[:9]):
    print("%-14s %6.2f %%  %.3e ticks" % (n, 100.0 * float(v) / tot, float(v)))
for n, v in zip(["M minmax", "M hist pass", "M rank bins", "M resolve pass", "M collect rank"], out[9:14]):
    print("   %-16s %6.2f %%  %.3e ticks" % (n, 100.0 * float(v) / tot, float(v)))
print("   extra median rounds:", int(out[15]), "over", F * (len(s.total_unique_labels)), "entities")
