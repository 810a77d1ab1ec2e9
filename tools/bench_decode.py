"""Input decode on the device, measured on a bundled MANIAC recording (tests/golden/pcd/00580: 1.73 MB binary_compressed
cloud + 165 kB label text) replicated into batches:  python tools/bench_decode.py [frames_per_batch] [batches]
  host reader      e2sec_read_pcd_xyz + e2sec_read_labels, one thread
  arrays e2e       e2sec_process_batch_async from pinned host arrays (12 B xyz per point + the 256 x 320 int32 label image)
  files  e2e       e2sec_process_batch_pcd_async from the pinned file bytes (upload, LZF decode + text parse on the copy
                   stream under the previous batch's kernels, then the same pipeline)
Both e2e figures are wall clock over all batches, inputs in pinned host memory, records back on the host.  With a third
argument N > 1 the batches go round robin over N engines (contexts) of the GPU: a stream is decoded by ONE warp, so a
batch's decode has a fixed latency that only other batches in flight can hide -- what analyse_maniac_manipulation does
(E2SEC_CONTEXTS)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from manipulationlibrary_b200 import _native as nat  # noqa: E402
from manipulationlibrary_b200 import manipulation_analyser as ma  # noqa: E402
from manipulationlibrary_b200.engine import Engine  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 200
NB = int(sys.argv[2]) if len(sys.argv) > 2 else 6
NC = int(sys.argv[3]) if len(sys.argv) > 3 else 1
d = os.path.join(ROOT, "tests", "golden", "pcd")
pp, lp = os.path.join(d, "00580_pc.pcd"), os.path.join(d, "00580_left-labels.dat")
fx = np.load(os.path.join(ROOT, "tests", "golden", "pushing06_fixture.npz"))
p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"],
                          replace=True, old=[7, 9], new=[6, 5])
t0 = time.perf_counter()
for _ in range(5):
    xyz, lab = ma.read_frame(pp, lp)
host_fps = 5 / (time.perf_counter() - t0)
pcd, txt = open(pp, "rb").read(), open(lp, "rb").read()
pb = torch.from_numpy(np.frombuffer(pcd * B, np.uint8).copy()).pin_memory()
lb = torch.from_numpy(np.frombuffer(txt * B, np.uint8).copy()).pin_memory()
po = np.arange(B + 1, dtype=np.int64) * len(pcd)
lo = np.arange(B + 1, dtype=np.int64) * len(txt)
hx = torch.from_numpy(np.broadcast_to(xyz, (B,) + xyz.shape).copy()).pin_memory()
hl = torch.from_numpy(np.broadcast_to(lab.reshape(-1).astype(np.int32), (B, lab.size)).copy()).pin_memory()
eng = Engine(0)
eng.configure(p, B, 307200, keep)
rec = [torch.empty((B, nat.record_doubles(eng.K)), dtype=torch.float64).pin_memory() for _ in range(2)]
recs = {id(eng): rec}
ids = np.arange(B, dtype=np.int32)


def run(files, engines=None):
    engines = engines or [eng]
    for e in engines:
        e.reset_sequence()
    t = time.perf_counter()
    for b in range(NB):
        e = engines[b % len(engines)]
        r = recs[id(e)][(b // len(engines)) & 1].numpy()
        if files:
            e.process_batch_files(pb.numpy(), po, lb.numpy(), lo, ids, (256, 320), (480, 640), records=r, wait=False)
        else:
            e.process_batch_device_async(hx.data_ptr(), hl.data_ptr(), B, ids, lab.size, records=r, label_shape=(256, 320),
                                         image_shape=(480, 640), on_device=False)
    for e in engines:
        e.sync()
    return B * NB / (time.perf_counter() - t), recs[id(engines[(NB - 1) % len(engines)])][((NB - 1) // len(engines)) & 1].numpy().copy()


def single(files):
    eng.reset_sequence()
    t = time.perf_counter()
    if files:
        eng.process_batch_files(pb.numpy(), po, lb.numpy(), lo, ids, (256, 320), (480, 640), records=rec[0].numpy(), wait=True)
    else:
        eng.process_batch_device_async(hx.data_ptr(), hl.data_ptr(), B, ids, lab.size, records=rec[0].numpy(), label_shape=(256, 320),
                                       image_shape=(480, 640), on_device=False)
        eng.sync()
    return (time.perf_counter() - t) * 1e3


run(True); run(False)
print("one batch alone, ms: files %.1f / %.1f  arrays %.1f / %.1f; stages of it: %s" % (single(True), single(True), single(False), single(False), eng.stage_ms()), file=sys.stderr)
f_fps, f_rec = run(True)
a_fps, a_rec = run(False)
assert np.array_equal(f_rec, a_rec, equal_nan=True), "files and arrays give different records"
ent = f_rec[0, 4: 4 + 16 * eng.K].reshape(eng.K, 16)
print("entities of the frame (n_raw, n_out, exact radius or nan):", [(int(e[0]), int(e[1]), float(e[3])) for e in ent], file=sys.stderr)
multi = None
if NC > 1:
    more = [eng]
    for _ in range(NC - 1):
        e = Engine(0)
        e.configure(p, B, 307200, keep)
        recs[id(e)] = [torch.empty((B, nat.record_doubles(eng.K)), dtype=torch.float64).pin_memory() for _ in range(2)]
        more.append(e)
    run(True, more)
    m_fps, m_rec = run(True, more)
    assert np.array_equal(m_rec, f_rec, equal_nan=True), "several contexts give different records"
    ma_fps, _ = run(False, more)
    multi = {"contexts": NC, "files_e2e_frames_per_s": m_fps, "arrays_e2e_frames_per_s": ma_fps}
print(json.dumps({"recording": "Pushing_06 frame 00580 x %d per batch, %d batches" % (B, NB),
                  "host_reader_frames_per_s_one_thread": host_fps,
                  "arrays_e2e_frames_per_s": a_fps, "arrays_h2d_bytes_per_frame": 307200 * 12 + lab.size * 4,
                  "files_e2e_frames_per_s": f_fps, "files_h2d_bytes_per_frame": len(pcd) + len(txt),
                  "records_equal": True, "several_contexts": multi}))
