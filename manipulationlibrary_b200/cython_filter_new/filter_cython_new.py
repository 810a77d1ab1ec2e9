"""Drop-in for cython_filter_new/filter_cython_new.pyx (reference :6-93): same two entry points, same argument
meaning, computed by libe2sec.so on the GPU (e2sec_region_filter / e2sec_distance, include/e2sec.h).

    region_filter_cython(center, pcd, hand_cloud, hand_center, obj_center, median_distance) -> (m, 3) float64
    distance_cython(p1, p2) -> float

The reference's flood fill draws its pivots from the process-global `random` stream (pyx:78).  Here the mode is
explicit: set_mode("closure") (default) returns the deterministic fixed point of that flood fill -- the seed point
followed by its connected component at radius float32(median_distance), raster order; set_mode("reference_seeded",
seed) replays pyx:61-83 with CPython's MT19937 seeded by `seed` (what random.seed(seed) immediately before the
reference call would give, bit for bit).  Like the reference, `pcd` must be C-contiguous float64 (ValueError
otherwise) and the output starts with the seed point, which appears twice.
"""
import numpy as np

from .. import _native as nat
from ..engine import Engine

_engine = None
_mode = "closure"
_seed = 0


def set_mode(mode, seed=0):
    global _mode, _seed
    if mode not in nat.FILTER_MODES:
        raise ValueError("mode must be one of %s" % sorted(nat.FILTER_MODES))
    _mode, _seed = mode, int(seed)


def _eng():
    global _engine
    if _engine is None:
        _engine = Engine(0)          # raises E2SecError without a CUDA device: no CPU fallback
    return _engine


def _buf(a, ndim, name):
    a = np.asarray(a)
    if a.dtype != np.float64 or a.ndim != ndim or not a.flags.c_contiguous:
        raise ValueError("%s: Buffer dtype mismatch / not C-contiguous float64 with ndim=%d" % (name, ndim))
    return a


def distance_cython(p1, p2):
    p1, p2 = _buf(p1, 1, "p1"), _buf(p2, 1, "p2")
    return float(nat.lib.e2sec_distance(nat.ptr(p1), nat.ptr(p2)))


def region_filter_cython(center, pcd, hand_cloud, hand_center, obj_center, median_distance):
    _buf(center, 1, "center")
    pcd = _buf(pcd, 2, "pcd")
    target = _buf(hand_center if hand_cloud else obj_center, 1, "center")     # pyx:20-26
    if len(pcd) == 2:
        return np.zeros((1, 3))                                               # pyx:35-42
    if len(pcd) < 4:
        raise ValueError("region_filter_cython needs more than 3 points (the reference never calls it with fewer, "
                         "manipulation_analyser.py:1520)")
    return _eng().region_filter(pcd, target, float(median_distance), mode=_mode, seed=_seed)
