"""CPU-side checks (no GPU): the C-ABI library loads and exports every symbol include/e2sec.h declares, the host
readers agree with the oracle's, the product refuses to run without a device (no CPU fallback), the synthetic
stream is reproducible, and the multi-rank record gather works under gloo with world_size 2."""
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

from oracle import pcd_io

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from manipulationlibrary_b200 import _native as nat
    hdr = open(os.path.join(ROOT, "include", "e2sec.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(e2sec_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in sorted(declared):
        assert hasattr(nat.lib, name), "libe2sec.so does not export %s" % name
    assert declared == set(nat.EXPORTS), declared ^ set(nat.EXPORTS)
    assert nat.lib.e2sec_abi_version() == nat.ABI_VERSION
    assert nat.record_doubles(6) == 4 + 16 * 6 + 49 + 36 + 12


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device behaviour")
def test_no_cpu_fallback():
    from manipulationlibrary_b200._native import E2SecError
    from manipulationlibrary_b200.engine import Engine
    with pytest.raises(E2SecError, match="no CPU fallback"):
        Engine(0)
    from manipulationlibrary_b200.cython_filter_new import filter_cython_new as f
    pts = np.random.default_rng(0).normal(size=(10, 3))
    with pytest.raises(E2SecError):
        f.region_filter_cython(pts.mean(0), pts, 0, pts.mean(0), pts.mean(0), 0.1)
    # the host-only operator still answers (pure arithmetic, pyx:6-7)
    assert f.distance_cython(np.zeros(3), np.array([3.0, 4.0, 0.0])) == 5.0


def _lzf_literal(raw):
    """A valid LZF stream made of literal runs only (control byte < 32 -> that many + 1 literal bytes)."""
    out = bytearray()
    for i in range(0, len(raw), 32):
        chunk = raw[i: i + 32]
        out.append(len(chunk) - 1)
        out += chunk
    return bytes(out)


def _write_pcd(path, xyz, mode, width, height):
    n = len(xyz)
    rgba = np.arange(n, dtype=np.uint32)
    head = ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z rgba\nSIZE 4 4 4 4\nTYPE F F F U\n"
            "COUNT 1 1 1 1\nWIDTH %d\nHEIGHT %d\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA %s\n" % (width, height, n, mode))
    with open(path, "wb") as f:
        f.write(head.encode())
        if mode == "ascii":
            for p, c in zip(xyz, rgba):
                f.write(("%r %r %r %d\n" % (float(p[0]), float(p[1]), float(p[2]), c)).replace("nan", "nan").encode())
        elif mode == "binary":
            rec = np.zeros(n, dtype=[("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("c", "<u4")])
            rec["x"], rec["y"], rec["z"], rec["c"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], rgba
            f.write(rec.tobytes())
        else:
            raw = xyz[:, 0].tobytes() + xyz[:, 1].tobytes() + xyz[:, 2].tobytes() + rgba.tobytes()
            comp = _lzf_literal(raw)
            f.write(np.array([len(comp), len(raw)], np.uint32).tobytes())
            f.write(comp)


@pytest.mark.parametrize("mode", ["ascii", "binary", "binary_compressed"])
def test_pcd_reader_matches_oracle_reader(tmp_path, mode):
    from manipulationlibrary_b200 import _native as nat
    rng = np.random.default_rng(1)
    xyz = rng.normal(size=(24 * 10, 3)).astype(np.float32)
    xyz[::5] = np.nan
    path = str(tmp_path / "f_pc.pcd")
    _write_pcd(path, xyz, mode, 24, 10)
    got = np.empty((300, 3), np.float32)
    w, h = nat.C.c_int32(), nat.C.c_int32()
    assert nat.lib.e2sec_read_pcd_xyz(path.encode(), nat.ptr(got), 300, nat.C.byref(w), nat.C.byref(h)) == 0
    assert (w.value, h.value) == (24, 10)
    want, wh = pcd_io.read_pcd_xyz(path)
    assert wh == (24, 10)
    np.testing.assert_array_equal(got[:240], want)
    np.testing.assert_array_equal(want, xyz)
    assert nat.lib.e2sec_read_pcd_xyz(path.encode(), nat.ptr(got), 100, nat.C.byref(w), nat.C.byref(h)) == -3   # capacity
    assert nat.lib.e2sec_read_pcd_xyz(b"/nonexistent.pcd", nat.ptr(got), 300, None, None) == -4


def test_label_reader_matches_oracle_reader(tmp_path):
    from manipulationlibrary_b200 import _native as nat
    rng = np.random.default_rng(2)
    lab = rng.integers(0, 12, size=(16, 20))
    path = str(tmp_path / "f_left-labels.dat")
    with open(path, "w") as f:
        for row in lab:
            f.write(" ".join("%d" % v for v in row) + " \n")
    got = np.empty(400, np.int32)
    r, c = nat.C.c_int32(), nat.C.c_int32()
    assert nat.lib.e2sec_read_labels(path.encode(), nat.ptr(got), 400, nat.C.byref(r), nat.C.byref(c)) == 0
    assert (r.value, c.value) == (16, 20)
    np.testing.assert_array_equal(got[:320].reshape(16, 20), pcd_io.read_labels(path).astype(np.int32))
    np.testing.assert_array_equal(got[:320].reshape(16, 20), lab)


def test_synthetic_stream_is_reproducible_per_frame():
    from manipulationlibrary_b200.synthetic import SyntheticStream
    a = SyntheticStream(width=160, height=120, n_objects=3, config=7, cycle=16)
    b = SyntheticStream(width=160, height=120, n_objects=3, config=7, cycle=16)
    xa, la = a.frame(11)
    b.frame(3)                                    # history must not matter
    xb, lb = b.frame(11)
    np.testing.assert_array_equal(la, lb)
    np.testing.assert_array_equal(xa, xb)
    assert set(np.unique(la)) <= set(range(0, 6))
    valid = ~np.isnan(xa).any(1)
    assert 0.08 < 1 - valid.mean() < 0.25
    x2, l2 = a.fetch([11, 12])
    np.testing.assert_array_equal(x2[0], xa)
    # every frame of a cycle moves the hand
    assert not np.array_equal(a.frame(12)[1], la)


def test_shard_ranges_cover_the_sequence():
    from manipulationlibrary_b200.sharding import shard_range
    for n in (0, 1, 7, 100, 1001):
        for world in (1, 2, 3, 8):
            cuts = [shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            assert max(b - a for a, b in cuts) - min(b - a for a, b in cuts) <= 1


_GLOO_WORKER = r"""
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, %(root)r)
from manipulationlibrary_b200.sharding import gather_records, shard_range
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 11
full = np.arange(n * 5, dtype=np.float64).reshape(n, 5) * 1.5
a, b = shard_range(n, rank, world)
got = gather_records(full[a:b], n, world)
assert got.shape == full.shape and (got == full).all(), got
# rank 0 resolves the gathered Pushing_06 records: same matrix as the unsharded golden
from tests.helpers import load_golden
from manipulationlibrary_b200 import _native as nat
from manipulationlibrary_b200.engine import Resolver, esec_to_e2sec_native
rec = load_golden("pushing06_records_closure")["records"]
a, b = shard_range(len(rec), rank, world)
allrec = gather_records(rec[a:b], len(rec), world)
if rank == 0:
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_closure")
    p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"])
    r = Resolver(p, keep)
    r.feed(allrec)
    table, fc = r.table()
    e2, fc2 = esec_to_e2sec_native(table, 3, fc)
    assert (e2.astype("U5") == gold["e2sec"]).all()
    assert fc2 == list(gold["final_columns"])
# the sharded run's one collective: columns emitted by batches dealt round-robin come back merged in frame order
from manipulationlibrary_b200.sharding import batch_owner, gather_columns, ring_tickets
from manipulationlibrary_b200.engine import columns_table
n_batches, B = 7, 10
rng = np.random.default_rng(5)
all_frames = np.sort(rng.choice(n_batches * B, size=23, replace=False)).astype(np.int32)
all_cols = rng.integers(1, 19, size=(len(all_frames), 32)).astype(np.uint8)
mine = [i for i, f in enumerate(all_frames) if batch_owner(int(f) // B, world)[0] == rank]
cols, frames = gather_columns(all_cols[mine], all_frames[mine], world)
if rank == 0:
    assert (frames == all_frames).all() and (cols == all_cols).all()
    t, fc = columns_table(cols, frames, 3)
    assert t.shape == (30, len(all_frames)) and fc == all_frames.tolist()
else:
    assert cols is None
tk = [ring_tickets(g, n_batches, serial=2) for g in range(n_batches)]
assert tk[0][0] == 0 and tk[-1][1] == 0 and all(tk[g][1] == tk[g + 1][0] != 0 for g in range(n_batches - 1))
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_record_gather_gloo_world2(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER % {"root": ROOT})
    env = dict(os.environ, PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert out.stdout.count("ok") == 2


def test_bench_numa_binding_is_harmless_without_topology():
    """bench.bind_near_gpu keeps a rank near its GPU when the box says where that is; on a box that does not (no sysfs entry
    for the device, a single node) it changes nothing and says so."""
    import bench

    class _Props:
        pci_domain_id, pci_bus_id, pci_device_id = 0xffff, 0xfe, 0x1f            # no such device

    class _Cuda:
        @staticmethod
        def get_device_properties(i):
            return _Props()

    class _Torch:
        cuda = _Cuda()

    before = os.sched_getaffinity(0)
    assert bench.bind_near_gpu(_Torch(), 0) is None
    assert os.sched_getaffinity(0) == before
