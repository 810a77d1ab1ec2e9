"""TEST INFRASTRUCTURE ONLY -- phase-1 records (include/e2sec.h layout) of the CPU oracle as committed goldens.

    python oracle/gen_records_golden.py            # tests/golden/pushing06_records_<mode>.npz
    python oracle/gen_records_golden.py --synth    # tests/golden/synth_records_closure.npz (16 frames, 640x480x10)
    python oracle/gen_records_golden.py --large    # tests/golden/synth_records_config{2,4,5}.npz (full-size frames of configs 4 / 5)

Pushing_06: oracle.records.RecordOracle over the 75 processed frames of tests/golden/pushing06_frames.npz with the
frame-0 fixture; the O(n^2) np.median(pdist()) values are the ones the reference's own run passed to
region_filter_cython (tests/golden/pushing06_<mode>.npz 'filter_calls'; test_oracle_vs_reference_golden.py checks
that the restated segmentation reproduces them).  Synthetic: manipulationlibrary_b200.synthetic.SyntheticStream
frames 28..43 (hand arriving at the first object), radii from scipy pdist.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import esec_oracle as eo  # noqa: E402
from oracle.records import RecordOracle  # noqa: E402
from tests.helpers import PUSHING06_ARGS, Pushing06Frames, load_golden  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SEED = 20261104


def pushing06(mode):
    fx = load_golden("pushing06_fixture")
    gold = load_golden("pushing06_" + mode)
    ov = {(int(r[0]), int(r[1])): float(r[4]) for r in gold["filter_calls"]}
    ref_filter = None
    if mode == "reference_seeded":
        from oracle import build_ref
        ref_filter = build_ref.load()
    orc = eo.E2SecOracle(PUSHING06_ARGS["ground_label"], PUSHING06_ARGS["hand_label"], None, PUSHING06_ARGS["thresh"],
                         fx["translation"], fx["roll"], fx["ground"], fx["total_unique_labels"], replace=False,
                         filter_mode=mode, seed=SEED, ref_filter=ref_filter, radius_override=ov)
    ro = RecordOracle(orc)
    frames = Pushing06Frames()
    recs = []
    for k in range(len(frames)):
        xyz, lab = frames.dense(k)
        recs.append(ro.frame_record(int(frames.frame_ids[k]), xyz, lab, bool(frames.hand_in_raw[k])))
    np.savez_compressed(os.path.join(GOLD, "pushing06_records_%s.npz" % mode), records=np.array(recs))
    print(mode, np.array(recs).shape)


def synth(first=28, count=16):
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream()
    orc = eo.E2SecOracle(s.ground_label, s.hand_label, None, s.thresh, s.translation, s.roll, s.ground,
                         s.total_unique_labels, filter_mode="closure", seed=SEED)
    ro = RecordOracle(orc)
    recs = []
    for f in range(first, first + count):
        xyz, lab = s.frame(f)
        recs.append(ro.frame_record(f, xyz, lab))
        print("synthetic frame", f, flush=True)
    np.savez_compressed(os.path.join(GOLD, "synth_records_closure.npz"), records=np.array(recs),
                        first=np.int64(first), count=np.int64(count))


def large(cfg):
    """BASELINE configs 4 / 5 at FULL frame size (1280x720 / 20 objects, 2000x1000 / 30 objects): two consecutive frames
    (the second one carries the cross table), and config 2 deep inside the benchmark's range (frames 777, 778)."""
    from manipulationlibrary_b200.synthetic import SyntheticStream
    import time
    w, h, n_obj, first = {4: (1280, 720, 20, 30), 5: (2000, 1000, 30, 30), 2: (640, 480, 10, 777)}[cfg]
    s = SyntheticStream(width=w, height=h, n_objects=n_obj, config=cfg, near_thresh=(cfg == 4))
    orc = eo.E2SecOracle(s.ground_label, s.hand_label, None, s.thresh, s.translation, s.roll, s.ground,
                         s.total_unique_labels, filter_mode="closure", seed=SEED)
    ro = RecordOracle(orc)
    recs = []
    for f in range(first, first + 2):
        t0 = time.time()
        xyz, lab = s.frame(f)
        recs.append(ro.frame_record(f, xyz, lab))
        print("config", cfg, "frame", f, "%.1f s" % (time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(GOLD, "synth_records_config%d.npz" % cfg), records=np.array(recs),
                        first=np.int64(first), count=np.int64(2), shape=np.array([w, h, n_obj, cfg], np.int64))


if __name__ == "__main__":
    if "--large" in sys.argv:
        for c in (2, 4, 5):
            large(c)
    elif "--synth" in sys.argv:
        synth()
    else:
        for m in ("closure", "reference_seeded"):
            pushing06(m)
