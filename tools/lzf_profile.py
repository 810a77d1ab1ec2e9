"""Stage split of lzf_decode_kernel (clock ticks of lane 0, summed over the streams of a batch).  Build with
E2SEC_NVCC_FLAGS=-DE2_LZ_PROFILE python -m manipulationlibrary_b200.build --force, run on the GPU, rebuild without."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from manipulationlibrary_b200 import _native as nat  # noqa: E402
from manipulationlibrary_b200.engine import Engine  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
d = os.path.join(ROOT, "tests", "golden", "pcd")
pcd = open(os.path.join(d, "00580_pc.pcd"), "rb").read()
txt = open(os.path.join(d, "00580_left-labels.dat"), "rb").read()
fx = np.load(os.path.join(ROOT, "tests", "golden", "pushing06_fixture.npz"))
p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"],
                          replace=True, old=[7, 9], new=[6, 5])
eng = Engine(0)
eng.configure(p, B, 307200, keep)
pb, lb = np.frombuffer(pcd * B, np.uint8).copy(), np.frombuffer(txt * B, np.uint8).copy()
po, lo = np.arange(B + 1, dtype=np.int64) * len(pcd), np.arange(B + 1, dtype=np.int64) * len(txt)
ids = np.arange(B, dtype=np.int32)
fn = nat.lib.e2sec_debug_lzf_profile
fn.argtypes = [C.c_void_p, C.c_int]
out = np.zeros(16, np.uint64)
eng.process_batch_files(pb, po, lb, lo, ids, (256, 320), (480, 640))
fn(None, 1)
eng.process_batch_files(pb, po, lb, lo, ids, (256, 320), (480, 640))
fn(nat.ptr(out), 0)
names = ["load input", "lane parses", "sweep", "place tokens", "expand", "pointer jumping", "flush", "loop head"]
tot = float(out[:8].sum())
w = float(out[8])
print("%d streams, %.0f windows per stream, %.0f ticks per window (%.2f ms per stream at 1.9 GHz)" % (B, w / B, tot / w, tot / B / 1.9e6))
for n, v in zip(names, out[:8]):
    print("%-16s %6.2f %%  %8.0f ticks per window" % (n, 100.0 * float(v) / tot, float(v) / w))
