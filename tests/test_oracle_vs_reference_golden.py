"""Pins oracle/esec_oracle.py (the CPU restatement) against goldens produced by running the
reference's OWN manipulation_analyser.py + compiled filter_cython_new.pyx (oracle/gen_golden.py)."""
import random

import numpy as np
import pytest

from oracle import esec_oracle as eo
from tests.helpers import PUSHING06_ARGS, Pushing06Frames, load_golden


def _run_oracle(mode, radius_from_golden=True):
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_" + mode)
    frames = Pushing06Frames()
    ov = {}
    if radius_from_golden:
        for row in gold["filter_calls"]:
            ov[(int(row[0]), int(row[1]))] = float(row[4])
    ref_filter = None
    if mode == "reference_seeded":
        from oracle import build_ref
        try:
            ref_filter = build_ref.load()
        except ImportError:
            ref_filter = None          # falls back to region_filter_restated (same stream of choices)
    orc = eo.E2SecOracle(PUSHING06_ARGS["ground_label"], PUSHING06_ARGS["hand_label"], None,
                         PUSHING06_ARGS["thresh"], fx["translation"], fx["roll"], fx["ground"],
                         fx["total_unique_labels"], replace=False, filter_mode=mode, seed=20261104,
                         ref_filter=ref_filter, keep_details=True, radius_override=ov)
    for k in range(len(frames)):
        xyz, lab = frames.dense(k)
        if not frames.hand_in_raw[k]:
            continue
        orc.process_frame(int(frames.frame_ids[k]), xyz, lab.astype(np.float64))
    return orc, gold


@pytest.mark.parametrize("mode", ["closure", "reference_seeded"])
def test_sequence_matches_reference_run(mode):
    orc, gold = _run_oracle(mode)
    e2, esec, fc = orc.finish()
    assert e2.shape == gold["e2sec"].shape
    assert (e2 == gold["e2sec"]).all()
    assert list(fc) == list(gold["final_columns"])   # thinned by esec_to_e2sec (:104-105)
    # every pre-dedup column the reference built
    raw = np.array([d["add"] for d in orc.details])
    assert raw.shape == gold["raw_columns"].shape
    assert (np.char.upper(raw) == np.char.upper(gold["raw_columns"])).all()
    # o1/o2/o3 label indices and the filtered hand (count, mean) per processed frame
    cols = gold["columns"]
    for d, row in zip(orc.details, cols):
        assert d["frame"] == int(row[0])
        for j, k in enumerate("123"):
            want = int(row[1 + j])
            got = d["labels"][k] if d["counts"][k] else -1
            assert got == want
        hand = d["ents"][orc.hand_idx]
        assert hand["n"] == int(row[4])
        np.testing.assert_allclose(hand["mean"], row[5:8], rtol=1e-12)


def test_filter_outputs_match_reference_calls():
    """Per region_filter_cython call of the reference run: n_out, AABB and mean of the output."""
    orc, gold = _run_oracle("reference_seeded")
    calls = {(int(r[0]), int(r[1])): r for r in gold["filter_calls"]}
    seen = 0
    for d in orc.details:
        for i, e in enumerate(d["ents"]):
            r = calls.get((d["frame"], i))
            if r is None:
                continue
            seen += 1
            assert e["n"] == int(r[3])
            np.testing.assert_allclose(e["lo"], r[5:8], rtol=0, atol=0)
            np.testing.assert_allclose(e["hi"], r[8:11], rtol=0, atol=0)
            np.testing.assert_allclose(e["mean"], r[11:14], rtol=1e-12)
    assert seen > 150


def test_pairwise_median_matches_reference_calls_on_small_objects():
    """np.median(pdist()) of the restated segmentation == the median_distance the reference passed."""
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_closure")
    frames = Pushing06Frames()
    tul = fx["total_unique_labels"]
    rows = sorted(gold["filter_calls"], key=lambda r: r[2])[:6]
    fid = {int(f): k for k, f in enumerate(frames.frame_ids)}
    for r in rows:
        xyz, lab = frames.dense(fid[int(r[0])])
        sel = (lab == int(tul[int(r[1])])) & ~np.isnan(xyz).any(1)
        obj = eo.transform_points(eo.transform_points(xyz[sel].astype(np.float64), fx["translation"]), fx["roll"])
        assert len(obj) == int(r[2])
        assert eo.pairwise_median(obj) == float(r[4])


def test_restated_filter_equals_compiled_reference_filter():
    from oracle import build_ref
    try:
        ref = build_ref.load()
    except ImportError:
        pytest.skip("compiled reference filter not available")
    rng = np.random.default_rng(5)
    for trial in range(6):
        n = int(rng.integers(5, 400))
        pts = rng.normal(size=(n, 3)) * [0.05, 0.03, 0.02]
        pts[: n // 10] += rng.normal(size=(n // 10, 3)) * 0.3
        center = np.median(pts, axis=0)
        radius = eo.pairwise_median(pts)
        random.seed(1234 + trial)
        want = np.asarray(ref.region_filter_cython(pts.mean(0), np.ascontiguousarray(pts), 0, center, center, radius))
        got = eo.region_filter_restated(pts, center, radius, random.Random(1234 + trial))
        assert want.shape == got.shape and (want == got).all()


def test_esec_to_e2sec_quirks():
    t = np.array([list("ab"), list("ab")], dtype="U5").T   # 2 rows? build a 30-row table instead
    table = np.full((30, 4), "n", dtype="U5")
    table[0, 2] = "t"
    table[0, 3] = "t"
    e2, esec, fc = eo.esec_to_e2sec(table, 3, [0, 3, 6, 9])
    assert e2.shape == (24, 2)           # column 0 dropped, duplicate column dropped
    assert (e2[0] == ["N", "T"]).all()
    assert list(fc) == [0, 3, 9]
