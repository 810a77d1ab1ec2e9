"""GPU parity: the CUDA path through the C ABI (libe2sec.so) against the CPU oracle and the committed goldens.

Bars (BASELINE.json north_star): relation codes / matrices / counts / indices bit-exact; float medians, AABBs,
radii and touching distances are compared for equality first (the kernels evaluate the reference's float64
expressions in the reference's operation order), means within 1e-12 relative (tolerance bar: 1e-5).
"""
import os
import random

import numpy as np
import pytest

from oracle import esec_oracle as eo
from tests.helpers import PUSHING06_ARGS, Pushing06Frames, load_golden, upgrade_records

pytestmark = pytest.mark.gpu

REC_HEADER, REC_ENTITY = 4, 16
SEED = 20261104


@pytest.fixture(scope="module")
def engine():
    from manipulationlibrary_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()


def compare_records(got, want, K, thresh, radius_exact=True):
    """Field by field.  In BOUNDED radius mode the record's radius is NaN wherever the cell table decided the entity;
    the bounds tail must then bracket the oracle's radius after float32 rounding (the only way the radius is used)."""
    got, want = np.asarray(got), upgrade_records(want, K)
    assert got.shape == want.shape
    stats = dict(filtered=0, exact=0)
    tail = REC_HEADER + REC_ENTITY * K + (K + 1) * (K + 1) + K * K
    for g, w in zip(got, want):
        f = int(w[0])
        assert g[0] == w[0] and g[1] == w[1] and g[2] == w[2] and g[3] == w[3], ("header", f, g[:4], w[:4])
        for e in range(K):
            a = g[REC_HEADER + REC_ENTITY * e: REC_HEADER + REC_ENTITY * (e + 1)]
            b = w[REC_HEADER + REC_ENTITY * e: REC_HEADER + REC_ENTITY * (e + 1)]
            lo, hi = g[tail + 2 * e], g[tail + 2 * e + 1]
            assert a[0] == b[0] and a[1] == b[1] and a[2] == b[2], ("counts/seed", f, e, a[:3], b[:3])
            if np.isnan(b[3]):
                assert np.isnan(a[3]) and np.isnan(lo) and np.isnan(hi)
            else:
                stats["filtered"] += 1
                assert np.float32(lo) <= np.float32(b[3]) <= np.float32(hi), ("radius bounds", f, e, lo, b[3], hi)
                if radius_exact or not np.isnan(a[3]):
                    stats["exact"] += 1
                    assert a[3] == b[3] and lo == b[3] and hi == b[3], ("radius", f, e, a[3], b[3])
            np.testing.assert_array_equal(a[4:7], b[4:7], err_msg="axis median %d %d" % (f, e))
            np.testing.assert_array_equal(a[7:13], b[7:13], err_msg="aabb %d %d" % (f, e))
            np.testing.assert_allclose(a[13:16], b[13:16], rtol=1e-12, atol=0, err_msg="mean %d %d" % (f, e))
        o = REC_HEADER + REC_ENTITY * K
        for name, size in (("pairs", (K + 1) * (K + 1)), ("cross", K * K)):
            a, b = g[o: o + size], w[o: o + size]
            near = b < thresh
            if name == "pairs":
                np.testing.assert_array_equal(a[near], b[near], err_msg="%s %d" % (name, f))
            else:                                   # contact flags: 0.0 where closer than thresh
                assert (a[near] == 0.0).all(), (name, f, a[near], b[near])
            assert (a[~near] >= thresh).all(), (name, f)
            o += size
    return stats


def _pushing06_engine(engine, mode, max_frames, radius_mode="exact"):
    from manipulationlibrary_b200 import _native as nat
    fx = load_golden("pushing06_fixture")
    p, keep = nat.make_params(fx["total_unique_labels"], PUSHING06_ARGS["hand_label"], PUSHING06_ARGS["ground_label"],
                              None, PUSHING06_ARGS["thresh"], fx["translation"], fx["roll"], fx["ground"],
                              filter_mode=mode, filter_seed=SEED, radius_mode=radius_mode)
    engine.configure(p, max_frames, 307200, keep)
    return p, keep


def _run_pushing06(engine, mode, batch, radius_mode="exact"):
    frames = Pushing06Frames()
    p, keep = _pushing06_engine(engine, mode, batch, radius_mode)
    recs = []
    for s in range(0, len(frames), batch):
        ks = range(s, min(s + batch, len(frames)))
        dense = [frames.dense(k) for k in ks]
        xyz = np.stack([d[0] for d in dense])
        lab = np.stack([d[1] for d in dense])
        recs.append(engine.process_batch(xyz, lab, frames.frame_ids[list(ks)]))
    return np.concatenate(recs), p, keep


# ---- operators at the reference's seams -----------------------------------------------------------
def _cloud(rng, n, outliers=True):
    pts = rng.normal(size=(n, 3)) * [0.05, 0.03, 0.02]
    if outliers and n >= 10:
        pts[: n // 10] += rng.normal(size=(n // 10, 3)) * 0.3
    return np.ascontiguousarray(pts)


@pytest.mark.parametrize("n", [4, 5, 31, 256, 257, 1000, 4097])
def test_axis_median(engine, n):
    rng = np.random.default_rng(n)
    pts = _cloud(rng, n)
    pts[::7, 0] = pts[0, 0]            # ties
    np.testing.assert_array_equal(engine.axis_median(pts), np.median(pts, axis=0))


@pytest.mark.parametrize("n", [4, 5, 16, 255, 256, 257, 700, 2500])
def test_pairwise_median_exact(engine, n):
    rng = np.random.default_rng(100 + n)
    pts = _cloud(rng, n)
    assert engine.pairwise_median(pts) == eo.pairwise_median(pts)


def test_pairwise_median_degenerate(engine):
    same = np.tile(np.array([[0.1, -0.2, 0.3]]), (50, 1))
    assert engine.pairwise_median(same) == 0.0
    grid = np.stack(np.meshgrid(np.arange(12.0), np.arange(9.0), [0.0], indexing="ij"), -1).reshape(-1, 3) * 0.01
    assert engine.pairwise_median(grid) == eo.pairwise_median(grid)      # massive ties


def _median_cases():
    rng = np.random.default_rng(4242)
    blob = lambda n, s=0.05: rng.normal(0.0, s, (n, 3)) + [0.3, -0.1, 1.2]
    two = np.concatenate([blob(1500, 0.01), blob(1500, 0.01) + [2.0, 0.0, 0.0]])            # the middle ranks sit at the jump between
    two_odd = two[:-1]                                                                    # within- and between-cluster distances
    outlier = np.concatenate([blob(3000, 0.002), [[40.0, -30.0, 25.0]]])                   # box diagonal >> median
    grid = np.stack(np.meshgrid(np.arange(60.0), np.arange(60.0), [0.0], indexing="ij"), -1).reshape(-1, 3) * 0.004   # one key holds
    line = np.column_stack([np.arange(2200.0) * 1e-3, np.zeros(2200), np.zeros(2200)])     # far more pairs than the candidate buffer
    return {"181 (all pairs fit the candidate buffer)": blob(181), "182 (one more than fits)": blob(182), "6000": blob(6000),
            "two clusters": two, "two clusters, odd pair count": two_odd, "outlier": outlier, "grid ties": grid, "line ties": line}


@pytest.mark.parametrize("name", list(_median_cases()))
def test_pairwise_median_sampled_interval(engine, name):
    """radius.cuh places the first histogram interval from a sample of pairs (and skips the histogram for small entities): the
    result stays the exact np.median(pdist()) whatever the distance distribution looks like."""
    pts = np.ascontiguousarray(_median_cases()[name])
    assert engine.pairwise_median(pts) == eo.pairwise_median(pts)


@pytest.mark.parametrize("start", [1, 2])
def test_pairwise_median_reopens_a_missed_interval(start, monkeypatch):
    """The sample only chooses where the first pass looks.  E2SEC_TEST_EXACT_START makes radius_init_kernel start under (1) /
    above (2) the median on purpose: radius_select_kernel has to reopen the interval upwards / downwards and the result is
    still exact -- through the operator and through a whole batch in exact radius mode."""
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine
    from manipulationlibrary_b200.synthetic import SyntheticStream
    monkeypatch.setenv("E2SEC_TEST_EXACT_START", str(start))
    eng = Engine(0)
    monkeypatch.delenv("E2SEC_TEST_EXACT_START")
    plain = Engine(0)
    try:
        rng = np.random.default_rng(77 + start)
        for n in (300, 2500):
            pts = _cloud(rng, n)
            assert eng.pairwise_median(pts) == eo.pairwise_median(pts)
        s = SyntheticStream(width=320, height=240, n_objects=4, config=97, cycle=24)
        ids = np.arange(4, 10, dtype=np.int32)
        xyz, lab = s.fetch(ids)
        rec = []
        for e in (eng, plain):
            p, keep = nat.make_params(**s.params(radius_mode="exact"))
            e.configure(p, len(ids), s.N, keep)
            rec.append(e.process_batch(xyz, lab, ids))
        assert np.array_equal(rec[0], rec[1], equal_nan=True)
        K = eng.K
        radius = rec[0][:, nat.REC_HEADER + 3: nat.REC_HEADER + nat.REC_ENTITY * K: nat.REC_ENTITY]
        assert np.isfinite(radius).sum() >= 4 * len(ids), "the batch should have gone through the exact selection"
    finally:
        eng.close()
        plain.close()


@pytest.mark.parametrize("n", [5, 64, 300, 1500])
def test_region_filter_closure(engine, n):
    rng = np.random.default_rng(200 + n)
    pts = _cloud(rng, n)
    center = np.median(pts, axis=0)
    radius = eo.pairwise_median(pts)
    want = eo.closure_filter(pts, center, radius)
    got = engine.region_filter(pts, center, radius, mode="closure")
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("n", [5, 64, 300, 1500])
def test_region_filter_reference_seeded(engine, n):
    rng = np.random.default_rng(300 + n)
    pts = _cloud(rng, n)
    center = np.median(pts, axis=0)
    radius = eo.pairwise_median(pts) * 0.5     # a small radius makes the random walk matter
    for seed in (0, 1, 12345, 2 ** 32 - 1):
        want = eo.region_filter_restated(pts, center, radius, random.Random(seed))
        got = engine.region_filter(pts, center, radius, mode="reference_seeded", seed=seed)
        np.testing.assert_array_equal(got, want)


def test_region_filter_matches_compiled_reference(engine):
    from oracle import build_ref
    try:
        ref = build_ref.load()
    except ImportError:
        pytest.skip("oracle/_ref/filter_cython_new*.so not built")
    rng = np.random.default_rng(7)
    for trial in range(4):
        pts = _cloud(rng, int(rng.integers(50, 900)))
        center = np.median(pts, axis=0)
        radius = eo.pairwise_median(pts) * float(rng.uniform(0.3, 1.0))
        random.seed(77 + trial)
        want = np.asarray(ref.region_filter_cython(pts.mean(0), pts, 0, center, center, radius))
        got = engine.region_filter(pts, center, radius, mode="reference_seeded", seed=77 + trial)
        np.testing.assert_array_equal(got, want)


def test_min_distance(engine):
    rng = np.random.default_rng(9)
    for trial in range(6):
        a = _cloud(rng, int(rng.integers(1, 3000)), outliers=False)
        b = _cloud(rng, int(rng.integers(1, 3000)), outliers=False) + [rng.uniform(0.0, 0.3), 0.0, 0.0]
        want = eo.min_distance(a, b)
        assert engine.min_distance(a, b) == want
        lim = 0.07
        got = engine.min_distance(a, b, limit=lim)
        assert (got == want) if want < lim else (got >= lim)
    assert np.isinf(engine.min_distance(np.zeros((0, 3)), np.ones((3, 3))))


def test_distance_operator():
    from manipulationlibrary_b200._native import lib, ptr
    p, q = np.array([0.1, 0.2, 0.3]), np.array([-1.0, 0.5, 2.0])
    assert lib.e2sec_distance(ptr(p), ptr(q)) == float(np.sqrt(((p - q) ** 2)[0] + ((p - q) ** 2)[1] + ((p - q) ** 2)[2]))


# ---- the per-frame pass ----------------------------------------------------------------------------
@pytest.mark.parametrize("mode,radius_mode", [("closure", "exact"), ("closure", "bounded"), ("reference_seeded", "exact"),
                                              ("reference_seeded", "bounded")])
def test_pushing06_records(engine, mode, radius_mode):
    got, p, keep = _run_pushing06(engine, mode, batch=16, radius_mode=radius_mode)
    want = load_golden("pushing06_records_" + mode)["records"]
    # REFERENCE_SEEDED needs the exact radius: the library forces EXACT there whatever was asked
    st = compare_records(got, want, 6, PUSHING06_ARGS["thresh"], radius_exact=(radius_mode == "exact" or mode != "closure"))
    if radius_mode == "bounded" and mode == "closure":
        print("pushing06 bounded: %d of %d filtered entities needed the exact path" % (st["exact"], st["filtered"]))
        assert st["exact"] < st["filtered"]


@pytest.mark.parametrize("mode,radius_mode", [("closure", "exact"), ("closure", "bounded"), ("reference_seeded", "exact")])
def test_pushing06_matrix_equals_reference_run(engine, mode, radius_mode):
    """CUDA records -> host resolver -> esec_to_e2sec == the matrix the reference's own code produced."""
    from manipulationlibrary_b200.engine import Resolver, esec_to_e2sec_native
    got, p, keep = _run_pushing06(engine, mode, batch=25, radius_mode=radius_mode)
    gold = load_golden("pushing06_" + mode)
    r = Resolver(p, keep)
    r.feed(got)
    table, fc = r.table()
    e2, fc2 = esec_to_e2sec_native(table, 3, fc)
    assert e2.shape == gold["e2sec"].shape
    assert (e2.astype("U5") == gold["e2sec"]).all()
    assert fc2 == list(gold["final_columns"])


def test_batch_size_invariance(engine):
    """Records do not depend on how the frames are cut into batches (the halo carries the previous kept frame)."""
    for rm in ("exact", "bounded"):
        a, _, _ = _run_pushing06(engine, "closure", batch=75, radius_mode=rm)
        b, _, _ = _run_pushing06(engine, "closure", batch=7, radius_mode=rm)
        np.testing.assert_array_equal(a, b)


def test_label_upscale_and_replace(engine):
    """cv2.resize INTER_NEAREST (:1433) + _replace_labels (:176-198) fused into the label gather."""
    from manipulationlibrary_b200 import _native as nat
    rng = np.random.default_rng(4)
    fx = load_golden("pushing06_fixture")
    raw = rng.choice([0, 2, 4, 5, 6, 7, 9, 11, 8, 3], size=(256, 320)).astype(np.float64)
    xyz = (rng.normal(size=(307200, 3)) * 0.2 + [0, 0, 1.2]).astype(np.float32)
    xyz[rng.random(307200) < 0.1] = np.nan
    p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"],
                              replace=True, old=[7, 9], new=[6, 5])
    engine.configure(p, 1, 307200, keep)
    rec = engine.process_batch(xyz[None], raw.astype(np.int32)[None], [0], label_shape=(256, 320), image_shape=(480, 640))[0]
    full = eo.replace_labels(eo.resize_nearest(raw, 640, 480), [7, 9], [6, 5]).flatten()
    valid = ~np.isnan(xyz).any(1)
    for e, v in enumerate(fx["total_unique_labels"]):
        assert rec[REC_HEADER + REC_ENTITY * e] == np.count_nonzero(valid & (full == v))


def test_empty_and_ragged_frames(engine):
    """All-NaN frame, frame without the hand label, entity with <= 3 points (unfiltered, :1520)."""
    from manipulationlibrary_b200 import _native as nat
    fx = load_golden("pushing06_fixture")
    p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"])
    N = 4096
    engine.configure(p, 3, N, keep)
    xyz = np.full((3, N, 3), np.nan, np.float32)
    lab = np.full((3, N), 2, np.int32)
    rng = np.random.default_rng(1)
    xyz[1, :100] = rng.normal(size=(100, 3)) * 0.05 + [0, 0, 1.0]      # frame 1: object 4 only, no hand
    lab[1, :100] = 4
    xyz[2, :50] = rng.normal(size=(50, 3)) * 0.05 + [0, 0, 1.0]        # frame 2: hand with 50 points, object with 3
    lab[2, :50] = 11
    xyz[2, 60:63] = rng.normal(size=(3, 3)) * 0.01 + [0.1, 0, 1.0]
    lab[2, 60:63] = 5
    rec = engine.process_batch(xyz, lab, [0, 3, 6])
    assert rec[0, 1] == 0 and rec[1, 1] == 0 and rec[2, 1] == 3
    ent = lambda f, e: rec[f, REC_HEADER + REC_ENTITY * e: REC_HEADER + REC_ENTITY * (e + 1)]
    obj4 = eo.transform_points(eo.transform_points(xyz[1, :100].astype(np.float64), fx["translation"]), fx["roll"])
    want4 = eo.closure_filter(obj4, np.median(obj4, axis=0), eo.pairwise_median(obj4))
    assert ent(0, 5)[0] == 0 and ent(1, 2)[0] == 100 and ent(1, 2)[1] == len(want4)
    np.testing.assert_array_equal(ent(1, 2)[7:10], want4.min(0))
    assert ent(2, 3)[0] == 3 and ent(2, 3)[1] == 3 and ent(2, 3)[2] == -1 and np.isnan(ent(2, 3)[3])
    obj = eo.transform_points(eo.transform_points(xyz[2, 60:63].astype(np.float64), fx["translation"]), fx["roll"])
    np.testing.assert_array_equal(ent(2, 3)[7:10], obj.min(0))
    np.testing.assert_array_equal(ent(2, 3)[10:13], obj.max(0))


@pytest.mark.parametrize("radius_mode", ["exact", "bounded"])
def test_synthetic_records(engine, radius_mode):
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.synthetic import SyntheticStream
    g = load_golden("synth_records_closure")
    first, count = int(g["first"]), int(g["count"])
    s = SyntheticStream()
    p, keep = nat.make_params(**s.params(radius_mode=radius_mode))
    engine.configure(p, count, s.N, keep)
    xyz, lab = s.frames(first, first + count)
    got = engine.process_batch(xyz, lab, np.arange(first, first + count))
    st = compare_records(got, g["records"], len(s.total_unique_labels), s.thresh, radius_exact=(radius_mode == "exact"))
    if radius_mode == "bounded":
        print("synthetic bounded: %d of %d filtered entities needed the exact path" % (st["exact"], st["filtered"]))
        assert st["exact"] * 4 <= st["filtered"]          # the cell table decides most of them


def test_bounded_equals_exact_on_hard_clouds(engine):
    """Entities built to defeat the cell table (points sitting right at the flood-fill radius, duplicated points, a
    line, far stragglers): BOUNDED must fall back and give EXACT's record, field for field."""
    from manipulationlibrary_b200 import _native as nat
    fx = load_golden("pushing06_fixture")
    rng = np.random.default_rng(42)
    N = 8192
    F = 6
    xyz = np.full((F, N, 3), np.nan, np.float32)
    lab = np.full((F, N), 2, np.int32)
    def put(f, label, pts):
        k = int(np.count_nonzero(~np.isnan(xyz[f, :, 0])))
        xyz[f, k: k + len(pts)] = pts
        lab[f, k: k + len(pts)] = label
    base = np.array([0.0, 0.0, 1.0])
    blob = rng.normal(size=(1500, 3)) * [0.05, 0.04, 0.01] + base
    for f in range(F):
        put(f, 11, rng.normal(size=(300, 3)) * 0.02 + base + [0.2, 0, 0])
    r = eo.pairwise_median(blob.astype(np.float32).astype(np.float64))
    put(0, 4, np.vstack([blob, blob[:40] + [0.0, 0.0, r * 1.0], blob[:40] + [0, 0, r * 1.5]]))      # a shell at ~the radius
    put(1, 4, np.tile(blob[:1], (200, 1)))                                                            # identical points
    put(2, 4, np.linspace(0, 1, 700)[:, None] * [0.3, 0.0, 0.0] + base)                              # a line
    put(3, 4, np.vstack([blob, rng.normal(size=(30, 3)) * 0.5 + base]))                              # scattered stragglers
    put(4, 4, np.vstack([blob[:5], blob[:2] + 0.3]))                                                  # tiny
    put(5, 4, rng.uniform(-0.2, 0.2, size=(3000, 3)) + base)                                          # a volume (many cells)
    put(5, 5, rng.uniform(-0.02, 0.02, size=(2000, 3)) * [1, 1, 20] + base + [0.5, 0, 0])            # a needle
    recs = {}
    for rm in ("exact", "bounded"):
        p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"],
                                  radius_mode=rm)
        engine.configure(p, F, N, keep)
        recs[rm] = engine.process_batch(xyz, lab, np.arange(F) * 3)
    st = compare_records(recs["bounded"], recs["exact"], 6, 0.07, radius_exact=False)
    assert st["filtered"] >= 2 * F


def test_synthetic_sharded_equals_whole(engine):
    """Frame sharding with a one-kept-frame halo: two shards run back to back on one device give
    the records, and therefore the matrix, of the unsharded run."""
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Resolver
    from manipulationlibrary_b200.sharding import run_shard
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream()
    p, keep = nat.make_params(**s.params())
    F = 24
    ids = np.arange(20, 20 + F)
    whole = run_shard(engine, p, keep, s, ids, rank=0, world=1, batch=8)
    parts = [run_shard(engine, p, keep, s, ids, rank=r, world=2, batch=8) for r in range(2)]
    np.testing.assert_array_equal(np.concatenate(parts), whole)
    r = Resolver(p, keep)
    r.feed(whole)
    assert r.n_columns >= 1


def _records_both_modes(engine, stream, ids, batch):
    from manipulationlibrary_b200 import _native as nat
    xyz, lab = stream.fetch(ids)
    out = {}
    for rm in ("exact", "bounded"):
        p, keep = nat.make_params(**stream.params(radius_mode=rm))
        engine.configure(p, batch, stream.N, keep)
        out[rm] = np.concatenate([engine.process_batch(xyz[i: i + batch], lab[i: i + batch], ids[i: i + batch])
                                  for i in range(0, len(ids), batch)])
    return out, xyz, lab


def test_config4_pair_stress_vs_oracle(engine):
    """BASELINE config 4 (20 objects -> 190 pairs, neighbours within +-5 mm of thresh) at a size the oracle finishes in
    seconds: every in-frame distance below thresh bit-exact, every contact flag right, both radius modes."""
    from manipulationlibrary_b200.synthetic import SyntheticStream
    from oracle.records import RecordOracle
    s = SyntheticStream(width=640, height=360, n_objects=20, config=4, near_thresh=True)
    ids = np.arange(40, 43)
    recs, xyz, lab = _records_both_modes(engine, s, ids, 3)
    orc = eo.E2SecOracle(s.ground_label, s.hand_label, None, s.thresh, s.translation, s.roll, s.ground,
                         s.total_unique_labels, filter_mode="closure")
    ro = RecordOracle(orc)
    want = np.stack([ro.frame_record(int(f), xyz[i], lab[i]) for i, f in enumerate(ids)])
    K = len(s.total_unique_labels)
    compare_records(recs["exact"], want, K, s.thresh)
    compare_records(recs["bounded"], want, K, s.thresh, radius_exact=False)
    pt = want[:, REC_HEADER + REC_ENTITY * K: REC_HEADER + REC_ENTITY * K + (K + 1) * (K + 1)]
    assert np.count_nonzero(np.abs(pt[np.isfinite(pt)] - s.thresh) < 0.006) >= 4      # the stress is really there


@pytest.mark.parametrize("shape", [(1280, 720, 20, 4), (2000, 1000, 30, 5)])
def test_large_configs_bounded_equals_exact(engine, shape):
    """BASELINE configs 4 and 5 at full frame size (1280x720 / 20 objects, 2M points / 30 objects): the bounded route
    gives the exact route's records, and the records satisfy the size-independent properties of the path."""
    from manipulationlibrary_b200.synthetic import SyntheticStream
    w, h, n_obj, cfg = shape
    s = SyntheticStream(width=w, height=h, n_objects=n_obj, config=cfg, near_thresh=(cfg == 4))
    ids = np.arange(30, 32)
    recs, xyz, lab = _records_both_modes(engine, s, ids, 2)
    K = len(s.total_unique_labels)
    st = compare_records(recs["bounded"], recs["exact"], K, s.thresh, radius_exact=False)
    assert st["filtered"] == 2 * (n_obj + 1)
    for f, rec in enumerate(recs["bounded"]):
        valid = ~np.isnan(xyz[f]).any(1)
        W = K + 1
        pt = rec[REC_HEADER + REC_ENTITY * K: REC_HEADER + REC_ENTITY * K + W * W].reshape(W, W)
        assert (pt == pt.T).all()                                    # min distance is symmetric
        for e in range(K):
            a = rec[REC_HEADER + REC_ENTITY * e: REC_HEADER + REC_ENTITY * (e + 1)]
            assert a[0] == np.count_nonzero(valid & (lab[f] == s.total_unique_labels[e]))     # segmentation counts
            if a[2] >= 0:
                assert 1 <= a[1] <= a[0] + 1                          # filter output is a subset (+ the doubled seed)
                assert (a[7:10] <= a[13:16]).all() and (a[13:16] <= a[10:13]).all()      # mean inside the AABB
                assert (a[7:10] <= a[4:7] + 1.0).all()


@pytest.mark.parametrize("cfg", [2, 4, 5])
def test_full_size_configs_equal_oracle_golden(engine, cfg):
    """BASELINE configs 4 and 5 at FULL frame size (and config 2 deep inside the benchmark's frame range) against records
    the CPU oracle produced (oracle/gen_records_golden.py --large; minutes of pdist per frame, hence committed): two
    consecutive frames, so the second one carries the cross table against the first.  Both radius modes."""
    from manipulationlibrary_b200.synthetic import SyntheticStream
    z = load_golden("synth_records_config%d" % cfg)
    w, h, n_obj, c = (int(v) for v in z["shape"])
    assert c == cfg
    s = SyntheticStream(width=w, height=h, n_objects=n_obj, config=cfg, near_thresh=(cfg == 4))
    ids = np.arange(int(z["first"]), int(z["first"]) + int(z["count"]))
    recs, _, _ = _records_both_modes(engine, s, ids, len(ids))
    K = len(s.total_unique_labels)
    st = compare_records(recs["exact"], z["records"], K, s.thresh)
    assert st["filtered"] == len(ids) * (n_obj + 1)
    compare_records(recs["bounded"], z["records"], K, s.thresh, radius_exact=False)


# ---- the drop-in Python surface ---------------------------------------------------------------------
def _write_sequence(tmp, frames, every=3):
    """Pushing_06 fixture frames as files named like the MANIAC recordings: <stem>_pc.pcd (binary PCD) and
    <stem>_left-labels.dat; the frames the reference skips (i % 3 != 0, :2272) are empty placeholders."""
    pcl, lab = tmp / "pcl", tmp / "lab"
    pcl.mkdir()
    lab.mkdir()
    head = ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\n"
            "WIDTH 640\nHEIGHT 480\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS 307200\nDATA binary\n").encode()
    by_id = {int(f): k for k, f in enumerate(frames.frame_ids)}
    last = int(frames.frame_ids[-1])
    for i in range(last + 1):
        stem = "%05d" % (505 + i)
        if i % every or i not in by_id:
            (pcl / (stem + "_pc.pcd")).write_bytes(b"")
            continue
        xyz, labels = frames.dense(by_id[i])
        with open(pcl / (stem + "_pc.pcd"), "wb") as f:
            f.write(head)
            f.write(np.ascontiguousarray(xyz, "<f4").tobytes())
        rows = labels.reshape(480, 640).copy()
        # the sparse fixture dropped border / ground pixels: put their labels back on point-less (NaN) pixels so that
        # np.unique(frame-0 labels) gives the reference's total_unique_labels [0, 2, 4, 5, 6, 11] (:1406-1426)
        rows[0][rows[0] == 255] = 0
        rows[1][rows[1] == 255] = 2
        with open(lab / (stem + "_left-labels.dat"), "w") as f:
            f.write("\n".join(" ".join(map(str, r)) for r in rows.tolist()) + "\n")
    return str(pcl) + "/", str(lab) + "/"


def test_dropin_analyse_maniac_manipulation(tmp_path, monkeypatch):
    """The reference's entry point, called like notebook cell 4 (icp=False): the saved e2sec_<savename>.npy equals the
    matrix the reference's own code produced, debug artefacts are pruned to final_columns."""
    from manipulationlibrary_b200 import manipulation_analyser as ma
    frames = Pushing06Frames()
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_closure")
    pcl, lab = _write_sequence(tmp_path, frames)
    monkeypatch.chdir(tmp_path)
    ma.set_frame0(fx["translation"], fx["roll"], fx["ground"])
    # the fixture's labels are already upscaled and replaced; 255 marks untracked pixels -> ignored_labels
    out = ma.analyse_maniac_manipulation(pcl, lab, 2, 11, None, 3, False, [], [], [255], 0.07, debug=True, icp=False,
                                         savename="Pushing_06")
    assert out is None
    e2 = np.load(tmp_path / "e2sec_Pushing_06.npy")
    assert e2.shape == gold["e2sec"].shape and e2.dtype.kind == "S"
    assert (e2.astype("U5") == gold["e2sec"]).all()
    assert list(ma.final_columns) == list(gold["final_columns"])
    kept = sorted(int(p.name.split(".")[0]) for p in (tmp_path / "debug_images" / "Pushing_06").iterdir())
    assert kept == sorted(set(int(c) for c in gold["final_columns"]))


def test_dropin_icp_model_tracking(tmp_path, monkeypatch):
    """icp=True, the notebook's default (:1677-1738): o1 / o2 / o3 are rigid models registered onto every frame.  The saved
    matrix equals the one the reference's own code produced with icp=True (oracle/gen_golden.py --icp)."""
    from manipulationlibrary_b200 import manipulation_analyser as ma
    frames = Pushing06Frames()
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_icp_closure")
    pcl, lab = _write_sequence(tmp_path, frames)
    monkeypatch.chdir(tmp_path)
    ma.set_frame0(fx["translation"], fx["roll"], fx["ground"])
    ma.analyse_maniac_manipulation(pcl, lab, 2, 11, None, 3, False, [], [], [255], 0.07, icp=True, savename="icp")
    e2 = np.load(tmp_path / "e2sec_icp.npy")
    assert e2.shape == gold["e2sec"].shape, (e2.shape, gold["e2sec"].shape, ma.final_columns, gold["final_columns"])
    assert (e2.astype("U5") == gold["e2sec"]).all()
    assert list(ma.final_columns) == list(gold["final_columns"])


@pytest.mark.parametrize("want_relations", [1, 2, 3])
def test_dropin_scripted_scene_relations_and_batches(tmp_path, monkeypatch, want_relations):
    """analyse_maniac_manipulation on a scripted scene of tests/golden/relation_kats.npz (the outputs of the reference's
    own analyse_maniac_manipulation): relations 1, 2 and 3, the hand-as-lowest-label scene among them, batches of 4 frames
    dealt to the device(s) behind their one-frame halo with the file reads running ahead."""
    from manipulationlibrary_b200 import manipulation_analyser as ma
    from oracle import pcd_io
    from oracle.gen_relation_kats import ROLL, THRESH, scene_from_golden
    z = load_golden("relation_kats")
    z = {k: z[k] for k in z.files}
    picked = [k for k in range(int(z["b_count"])) if int(z["b%d_params" % k][4]) == want_relations][:2]
    if want_relations == 3:
        picked += [k for k in range(int(z["b_count"])) if int(z["b%d_params" % k][1]) == 1][:1]      # the hand is label 1
    assert picked
    monkeypatch.setenv("E2SEC_BATCH_FRAMES", "4")
    head = ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\n"
            "WIDTH 640\nHEIGHT 480\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS 307200\nDATA binary\n").encode()
    for k in picked:
        sc, exp = scene_from_golden(z, k)
        d = tmp_path / ("scene%d" % k)
        pcl, lab = d / "pcl", d / "lab"
        pcl.mkdir(parents=True)
        lab.mkdir()
        for i in range(len(sc.frames) * 3):
            stem = "%05d" % i
            if i % 3:
                (pcl / (stem + "_pc.pcd")).write_bytes(b"")
                continue
            xyz, labels = sc.dense(i // 3)
            with open(pcl / (stem + "_pc.pcd"), "wb") as f:
                if want_relations == 1:                      # DATA binary: decoded by the host reader
                    f.write(head)
                    f.write(np.ascontiguousarray(xyz, "<f4").tobytes())
                else:                                        # binary_compressed like the recordings: decoded on the device
                    f.write(pcd_io.pcd_binary_compressed(xyz, 640, 480))
            np.savetxt(lab / (stem + "_left-labels.dat"), labels, fmt="%d")
        monkeypatch.chdir(d)
        ma.set_frame0(np.eye(4), ROLL, sc.ground_final)
        ma.analyse_maniac_manipulation(str(pcl) + "/", str(lab) + "/", sc.ground_label, sc.hand_label, sc.support, sc.relations,
                                       False, [], [], [], THRESH, cut_labels=[], icp=False, savename="kat")
        e2 = np.load(d / "e2sec_kat.npy")
        assert e2.shape == exp["e2sec"].shape and (e2.astype("U5") == exp["e2sec"]).all(), k
        if sc.relations == 3:
            assert list(ma.final_columns) == list(exp["final_columns"]), k
        assert list(np.asarray(ma.total_unique_labels, np.float64)) == list(exp["total_unique_labels"]), k


def test_dropin_process_frame_by_frame(tmp_path, monkeypatch):
    """_process called per frame like the reference's loop (:2271-2281) builds the same raw eSEC table."""
    from manipulationlibrary_b200 import manipulation_analyser as ma
    frames = Pushing06Frames()
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_closure")
    pcl, lab = _write_sequence(tmp_path, frames)
    ma.set_frame0(fx["translation"], fx["roll"], fx["ground"])
    table = np.chararray((30, 1), itemsize=5)
    table[:] = "-"
    T = R = 0
    files = sorted(os.listdir(pcl))
    for i, f in enumerate(files[:120]):
        if i % 3 == 0:
            T, R, table = ma._process(pcl + f[0:-7] + "_pc.pcd", lab + f[0:-7] + "_left-labels.dat", 2, 11, None, T, R, i,
                                      table, 3, ignored_labels=[255], thresh=0.07, icp=False)
    gcols = {int(c[0]): row for c, row in zip(gold["columns"], gold["raw_columns"])}
    assert table.shape[1] == len(ma.final_columns) >= 5
    for j, fr in enumerate(ma.final_columns):
        assert (np.char.upper(table[:, j].astype("U5")) == np.char.upper(gcols[fr])).all()


def test_dropin_region_filter_cython(engine):
    from manipulationlibrary_b200.cython_filter_new import filter_cython_new as f
    rng = np.random.default_rng(11)
    pts = _cloud(rng, 500)
    center = np.median(pts, axis=0)
    radius = eo.pairwise_median(pts)
    f.set_mode("closure")
    np.testing.assert_array_equal(np.asarray(f.region_filter_cython(pts.mean(0), pts, 1, center, center, radius)),
                                  eo.closure_filter(pts, center, radius))
    f.set_mode("reference_seeded", 99)
    np.testing.assert_array_equal(np.asarray(f.region_filter_cython(pts.mean(0), pts, 0, center, center, radius * 0.4)),
                                  eo.region_filter_restated(pts, center, radius * 0.4, random.Random(99)))
    f.set_mode("closure")
    assert f.region_filter_cython(pts.mean(0), pts[:2].copy(), 0, center, center, radius).shape == (1, 3)
    with pytest.raises(ValueError):
        f.region_filter_cython(pts.mean(0), pts[:, ::-1], 0, center, center, radius)       # not C-contiguous
    assert f.distance_cython(pts[0].copy(), pts[1].copy()) == float(
        np.sqrt((pts[0, 0] - pts[1, 0]) ** 2 + (pts[0, 1] - pts[1, 1]) ** 2 + (pts[0, 2] - pts[1, 2]) ** 2))


# ---- single-pass cluster segmentation (seg_fused.cuh) against the two-pass kernels (seg.cuh) -----------------------
@pytest.mark.parametrize("shape", [(320, 240, False), (640, 480, True), (64, 48, False), (252, 1, False), (250, 1, False), (2048, 3, False)])
def test_seg_fused_equals_two_pass(shape, monkeypatch):
    """The bulk-copy / cluster kernel and the two-pass fallback must give bit-identical records: same per-entity counts,
    same raster order inside an entity, same transformed coordinates (so every downstream field is equal)."""
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine
    from manipulationlibrary_b200.synthetic import SyntheticStream
    w, h, upscale = shape
    recs = []
    for two_pass, mode in (("0", "stream"), ("1", "stream"), ("0", "cluster")):     # unit stream, two-pass, cluster kernel
        monkeypatch.setenv("E2SEC_SEG_TWO_PASS", two_pass)
        monkeypatch.setenv("E2SEC_SEG_MODE", mode)
        eng = Engine(0)
        try:
            if h >= 48:
                s = SyntheticStream(width=w, height=h, n_objects=4, config=7, cycle=24)
                p, keep = nat.make_params(**s.params())
                ids = np.arange(8, 13)
                xyz, lab = s.fetch(ids)
                eng.configure(p, len(ids), s.N, keep)
                if upscale:          # labels arrive at half resolution and are upscaled like cv2.resize (:1433)
                    small = lab.reshape(len(ids), h, w)[:, ::2, ::2]
                    recs.append(eng.process_batch(xyz, small, ids, label_shape=(h // 2, w // 2), image_shape=(h, w)))
                else:
                    recs.append(eng.process_batch(xyz, lab, ids))
            else:                    # ragged sizes: N not a multiple of the tile, fewer tiles than the cluster has CTAs
                fx = load_golden("pushing06_fixture")
                p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"])
                N = w * h
                rng = np.random.default_rng(N)
                xyz = (rng.normal(size=(3, N, 3)) * 0.05 + [0, 0, 1.0]).astype(np.float32)
                xyz[rng.random((3, N)) < 0.1] = np.nan
                lab = rng.choice([0, 2, 4, 5, 6, 11], size=(3, N)).astype(np.int32)
                eng.configure(p, 3, N, keep)
                recs.append(eng.process_batch(xyz, lab, [0, 3, 6]))
        finally:
            eng.close()
    np.testing.assert_array_equal(recs[0], recs[1])
    np.testing.assert_array_equal(recs[0], recs[2])


# ---- uint8 label images (e2sec_process_batch_labels8) ----------------------------------------------------------------
@pytest.mark.parametrize("shape", [(320, 240, False), (640, 480, True), (250, 1, False), (253, 1, False)])
def test_labels8_equals_int32(shape):
    """Label images sent as uint8 (13 B/point over PCIe) give bit-identical records to the int32 entry point, for full-size
    images, for images that are upscaled on the device, and for sizes the 4-wide widening kernel has to finish one by one."""
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine
    from manipulationlibrary_b200.synthetic import SyntheticStream
    w, h, upscale = shape
    eng = Engine(0)
    try:
        if h >= 48:
            s = SyntheticStream(width=w, height=h, n_objects=4, config=7, cycle=24)
            p, keep = nat.make_params(**s.params())
            ids = np.arange(8, 13)
            xyz, lab = s.fetch(ids)
            eng.configure(p, len(ids), s.N, keep)
            kw = {}
            if upscale:
                lab = np.ascontiguousarray(lab.reshape(len(ids), h, w)[:, ::2, ::2])
                kw = dict(label_shape=(h // 2, w // 2), image_shape=(h, w))
        else:
            fx = load_golden("pushing06_fixture")
            p, keep = nat.make_params(fx["total_unique_labels"], 11, 2, None, 0.07, fx["translation"], fx["roll"], fx["ground"])
            N = w * h
            rng = np.random.default_rng(N)
            xyz = (rng.normal(size=(3, N, 3)) * 0.05 + [0, 0, 1.0]).astype(np.float32)
            xyz[rng.random((3, N)) < 0.1] = np.nan
            lab = rng.choice([0, 2, 4, 5, 6, 11, 200, 255], size=(3, N)).astype(np.int32)
            ids, kw = [0, 3, 6], {}
            eng.configure(p, 3, N, keep)
        eng.reset_sequence()
        a = eng.process_batch(xyz, lab.astype(np.int32), ids, **kw)
        eng.reset_sequence()
        b = eng.process_batch(xyz, lab.astype(np.uint8), ids, **kw)
    finally:
        eng.close()
    np.testing.assert_array_equal(a, b)


@pytest.mark.gpu
def test_batches_with_explicit_halo_equal_one_sequence():
    """bench.py --contexts 2: every batch is its own sequence on one of two contexts, with the frame before it in front
    as the halo.  The records (cross-frame fields included) must equal those of one context walking the whole sequence."""
    import numpy as np
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=320, height=240, n_objects=4, config=99, cycle=24)
    p, keep = nat.make_params(**s.params())
    B, nb = 6, 3
    ids = np.arange(0, B * nb)
    xyz, lab = s.fetch(ids)
    one = Engine(0)
    one.configure(p, B + 1, s.N, keep)
    one.reset_sequence()
    want = np.concatenate([one.process_batch(xyz[i * B:(i + 1) * B], lab[i * B:(i + 1) * B], ids[i * B:(i + 1) * B])
                           for i in range(nb)])
    one.close()
    engs = [Engine(0), Engine(0)]
    for e in engs:
        e.configure(p, B + 1, s.N, keep)
    got = []
    for i in range(nb):
        e = engs[i % 2]
        e.reset_sequence()
        lo = i * B - (1 if i else 0)
        rec = e.process_batch(xyz[lo:(i + 1) * B], lab[lo:(i + 1) * B], ids[lo:(i + 1) * B])
        if i:
            assert int(rec[0, 1]) == (nat.FLAG_HAND_LABEL | nat.FLAG_HAND_POINTS)      # the halo frame is a kept frame
        got.append(rec[1:] if i else rec)
    for e in engs:
        e.close()
    got = np.concatenate(got)
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
