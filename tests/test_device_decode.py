"""Input decode on the device (csrc/lzf.cuh, e2sec_process_batch_pcd): the reference reads every cloud with
o3d.io.read_point_cloud (manipulation_analyser.py:1441; PCD v0.7 binary_compressed = one LZF stream over the fields in SoA
order) and every label image with pd.read_csv (:1391).  Here the files travel as they are and are decoded on the GPU.
  * two bundled Pushing_06 recordings (tests/golden/pcd/, copied from the reference's example/): decoded on the device ==
    the library's host reader == the oracle's reader, bit for bit; the records equal those of the decoded arrays;
  * synthetic frames written as binary_compressed PCD + label text by the oracle-side writer (oracle/pcd_io.py);
  * malformed streams and label files are reported, not read out of bounds."""
import os

import numpy as np
import pytest

from oracle import pcd_io
from tests.helpers import GOLD, load_golden

PCD_DIR = os.path.join(GOLD, "pcd")
STEMS = ["00505", "00580"]


def _blob(parts):
    off = np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64)
    return np.frombuffer(b"".join(parts), np.uint8).copy(), off


def test_oracle_lzf_writer_roundtrip(tmp_path):
    rng = np.random.default_rng(3)
    xyz = rng.normal(size=(640 * 48, 3)).astype(np.float32)
    xyz[:, 2] = np.round(xyz[:, 2] * 50) / 50
    xyz[rng.random(len(xyz)) < 0.2] = np.nan
    raw = pcd_io.pcd_binary_compressed(xyz, 640, 48)
    p = tmp_path / "a.pcd"
    p.write_bytes(raw)
    got, (w, h) = pcd_io.read_pcd_xyz(str(p))
    assert (w, h) == (640, 48) and np.array_equal(got.view(np.uint32), xyz.view(np.uint32))
    body = raw[raw.index(b"DATA binary_compressed\n") + 23:]
    comp, rawlen = np.frombuffer(body[:8], np.uint32)
    assert pcd_io.lzf_decompress_py(body[8: 8 + int(comp)], int(rawlen))[: 12 * len(xyz)] == \
        xyz[:, 0].tobytes() + xyz[:, 1].tobytes() + xyz[:, 2].tobytes()
    # the library's host reader agrees
    from manipulationlibrary_b200 import _native as nat
    out = np.empty((len(xyz), 3), np.float32)
    ww, hh = nat.C.c_int32(), nat.C.c_int32()
    assert nat.lib.e2sec_read_pcd_xyz(os.fsencode(str(p)), nat.ptr(out), len(out), nat.C.byref(ww), nat.C.byref(hh)) == 0
    assert np.array_equal(out.view(np.uint32), xyz.view(np.uint32))


@pytest.fixture(scope="module")
def engine():
    from manipulationlibrary_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()


@pytest.mark.gpu
def test_bundled_recordings_decoded_on_device(engine):
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200 import manipulation_analyser as ma
    fx = load_golden("pushing06_fixture")
    p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"],
                              replace=True, old=[7, 9], new=[6, 5])
    engine.configure(p, 4, 307200, keep)
    pcd = [open(os.path.join(PCD_DIR, s + "_pc.pcd"), "rb").read() for s in STEMS]
    lab = [open(os.path.join(PCD_DIR, s + "_left-labels.dat"), "rb").read() for s in STEMS]
    pb, po = _blob(pcd)
    lb, lo = _blob(lab)
    ids = np.array([0, 75], np.int32)
    rec = engine.process_batch_files(pb, po, lb, lo, ids, (256, 320), (480, 640))
    xyz_d, lab_d = engine.staged_inputs(2, 256 * 320)
    for k, s in enumerate(STEMS):
        xyz_h, lab_h = ma.read_frame(os.path.join(PCD_DIR, s + "_pc.pcd"), os.path.join(PCD_DIR, s + "_left-labels.dat"))
        xyz_o, _ = pcd_io.read_pcd_xyz(os.path.join(PCD_DIR, s + "_pc.pcd"))
        assert np.array_equal(xyz_d[k].view(np.uint32), xyz_h.view(np.uint32)), "device LZF decode != host reader"
        assert np.array_equal(xyz_d[k].view(np.uint32), xyz_o.view(np.uint32)), "device LZF decode != oracle reader"
        assert np.array_equal(lab_d[k].reshape(256, 320), lab_h)
        assert np.array_equal(lab_h, pcd_io.read_labels(os.path.join(PCD_DIR, s + "_left-labels.dat")).astype(np.int32))
    engine.reset_sequence()
    want = engine.process_batch(xyz_d, lab_d, ids, label_shape=(256, 320), image_shape=(480, 640))
    assert np.array_equal(rec, want, equal_nan=True)
    assert rec[:, 4: 4 + 16 * 6: 16].sum() > 100000                      # the tracked labels hold real clouds


@pytest.mark.gpu
def test_synthetic_compressed_files_equal_arrays(engine):
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=320, height=240, n_objects=4, config=99, cycle=24)
    ids = np.arange(10, 16, dtype=np.int32)
    xyz, lab = s.fetch(ids)
    p, keep = nat.make_params(**s.params())
    engine.configure(p, len(ids), s.N, keep)
    want = engine.process_batch(xyz, lab, ids)
    pcd = [pcd_io.pcd_binary_compressed(x, s.W, s.H, rgba=(i % 2 == 0)) for i, x in enumerate(xyz)]
    txt = ["\n".join(" ".join(map(str, r)) for r in l.reshape(s.H, s.W).tolist()).encode() + b"\n" for l in lab]
    pb, po = _blob(pcd)
    lb, lo = _blob(txt)
    engine.reset_sequence()
    rec = engine.process_batch_files(pb, po, lb, lo, ids, (s.H, s.W), (s.H, s.W))
    assert np.array_equal(rec, want, equal_nan=True)
    xyz_d, lab_d = engine.staged_inputs(len(ids), s.N)
    assert np.array_equal(xyz_d.view(np.uint32), xyz.view(np.uint32)) and np.array_equal(lab_d, lab)
    # two batches of three: the second one's decode runs under the first one's kernels (copy stream), same records
    engine.reset_sequence()
    a = engine.process_batch_files(pb, po[:4], lb, lo[:4], ids[:3], (s.H, s.W), (s.H, s.W))
    b = engine.process_batch_files(pb, po[3:], lb, lo[3:], ids[3:], (s.H, s.W), (s.H, s.W))
    assert np.array_equal(np.concatenate([a, b]), want, equal_nan=True)


@pytest.mark.gpu
def test_malformed_files_are_reported(engine):
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200._native import E2SecError
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=160, height=120, n_objects=3, config=98, cycle=16)
    ids = np.arange(0, 2, dtype=np.int32)
    xyz, lab = s.fetch(ids)
    p, keep = nat.make_params(**s.params())
    engine.configure(p, 2, s.N, keep)
    good = [pcd_io.pcd_binary_compressed(x, s.W, s.H) for x in xyz]
    txt = ["\n".join(" ".join(map(str, r)) for r in l.reshape(s.H, s.W).tolist()).encode() + b"\n" for l in lab]
    lb, lo = _blob(txt)
    # a back-reference before the start of the output, a truncated stream, a label file one number short, binary (not compressed)
    body = good[0].index(b"DATA binary_compressed\n") + 23 + 8
    broken = bytearray(good[0]); broken[body] = 0xE0; broken[body + 1] = 0xFF; broken[body + 2] = 0xFF
    cut = good[1][: len(good[1]) // 2]
    cut = cut[: body - 8] + np.array([len(cut) - body, 12 * s.N + 4 * s.N], np.uint32).tobytes() + cut[body:]
    for files in ([bytes(broken), good[1]], [good[0], cut]):
        pb, po = _blob(files)
        engine.reset_sequence()
        with pytest.raises(E2SecError):
            engine.process_batch_files(pb, po, lb, lo, ids, (s.H, s.W), (s.H, s.W))
    pb, po = _blob(good)
    short = [txt[0][: txt[0].rstrip().rindex(b" ")] + b"\n", txt[1]]
    lb2, lo2 = _blob(short)
    engine.reset_sequence()
    with pytest.raises(E2SecError):
        engine.process_batch_files(pb, po, lb2, lo2, ids, (s.H, s.W), (s.H, s.W))
    plain = good[0].replace(b"DATA binary_compressed", b"DATA binary")
    pb3, po3 = _blob([plain, good[1]])
    with pytest.raises(E2SecError):
        engine.process_batch_files(pb3, po3, lb, lo, ids, (s.H, s.W), (s.H, s.W))
    # and the context still works afterwards
    engine.reset_sequence()
    rec = engine.process_batch_files(pb, po, lb, lo, ids, (s.H, s.W), (s.H, s.W))
    engine.reset_sequence()
    assert np.array_equal(rec, engine.process_batch(xyz, lab, ids), equal_nan=True)
