"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy / scipy) of the e2SEC extraction hot path of
tobi-code/ManipulationLibrary.  Nothing in the product path may import this module; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do, as the *checker*.

It follows, line by line, /root/reference/manipulation_analyser.py:
    _process                 :1355-1829, :2000-2007, :2179      -> E2SecOracle.process_frame
    _replace_labels          :176-198                           -> replace_labels
    _distance                :200-209                           -> _dist3
    _fillTN_absent           :363-487                           -> fill_tn
    _fillSSR_8 + predicates  :491-1061                          -> fill_ssr
    _fillDSR_new_2           :1064-1353                         -> fill_dsr
    esec_to_e2sec            :62-111                            -> esec_to_e2sec
    analyse_maniac_manipulation :2184-2291                      -> analyse_sequence
and /root/reference/cython_filter_new/filter_cython_new.pyx:6-93 -> region_filter_restated.

Open3D (pinned open3d==0.11.2, docs/requirements.txt:15) is absent; its calls are replaced by their
definitions (see oracle/fake_o3d.py header for the table).

PARITY PIN.  The reference has no tests, golden vectors or stored outputs, so this
restatement is pinned against the reference's own code *executed in the build container*:
oracle/gen_golden.py imports /root/reference/manipulation_analyser.py unmodified (with
oracle/fake_o3d.py standing in for Open3D and the reference's compiled filter_cython_new.pyx) and
writes tests/golden/*.npz; tests/test_oracle_vs_reference_golden.py checks this file against them.

Determinism pins that are ADDITIONS to the reference:
  * points of one label keep raster order (stable sort) instead of :1449's unstable argsort;
  * random.seed(filter_seed(base, frame, label_index)) immediately before every filter call;
  * translation / roll / ground cloud are inputs (a fixture), not recomputed with random.uniform
    (:250-252) and unseeded RANSAC (:353);
  * icp=False (:1669-1675).
Filter modes: "reference_seeded" replays the reference's randomized flood fill (pyx:61-83) under the
seed above; "closure" is the deterministic fixed point it converges to (the connected component of
the seed point in the radius-r disk graph), see closure_filter().
"""
import random

import numpy as np
from scipy.spatial import cKDTree
from scipy.spatial.distance import pdist

CENTER_DISTANCE = 0.005107632747118286   # manipulation_analyser.py:1121
STABLE_DIST = 0.0045302034127268085      # manipulation_analyser.py:1122
FILTER_PATIENCE = 30                     # filter_cython_new.pyx:61
NAN_SENTINEL = -100.0                    # manipulation_analyser.py:1445

# (a, b) entity pairs of the 10 rows of each relation family, manipulation_analyser.py:373-383
ROW_PAIRS = [("H", "1"), ("H", "2"), ("H", "3"), ("H", "G"), ("1", "2"),
             ("1", "3"), ("1", "G"), ("2", "3"), ("2", "G"), ("3", "G")]


def filter_seed(base, frame, label_index):
    """The oracle's per-call seed (an addition; the reference never seeds)."""
    return (int(base) + 1000003 * int(frame) + 7919 * int(label_index)) & 0xFFFFFFFF


# --------------------------------------------------------------------------------------------------
# a1: label ingest
# --------------------------------------------------------------------------------------------------
def resize_nearest(label, out_w, out_h):
    """cv2.resize(label, (out_w, out_h), interpolation=cv2.INTER_NEAREST)  (:1433).
    OpenCV's resizeNN: src_x = min(floor(dst_x * (1/fx)), src_w-1) with fx = out_w/src_w in double."""
    h, w = label.shape
    ifx = 1.0 / (float(out_w) / float(w))
    ify = 1.0 / (float(out_h) / float(h))
    sx = np.minimum(np.floor(np.arange(out_w) * ifx).astype(np.int64), w - 1)
    sy = np.minimum(np.floor(np.arange(out_h) * ify).astype(np.int64), h - 1)
    return label[sy][:, sx]


def replace_labels(label, old, new):
    """_replace_labels :176-198 -- sequential substitution; later pairs see earlier results."""
    out = label
    for i in range(len(old)):
        out = np.where(out == old[i], new[i], out)
    return out


def label_lut(old, new, size=256):
    """The sequential substitution composed into one table (what the CUDA path uploads)."""
    lut = np.arange(size, dtype=np.int64)
    for o, n in zip(old, new):
        lut = np.where(lut == o, n, lut)
    return lut


# --------------------------------------------------------------------------------------------------
# geometry definitions standing in for Open3D
# --------------------------------------------------------------------------------------------------
def transform_points(pts, M):
    """PointCloud.transform (:1484-1485) with the operation order pinned (see oracle/fake_o3d.py)."""
    pts = np.asarray(pts, np.float64)
    M = np.asarray(M, np.float64)
    x, y, z = pts[:, 0], pts[:, 1], pts[:, 2]
    q = [((M[r, 0] * x + M[r, 1] * y) + M[r, 2] * z) + M[r, 3] for r in range(4)]
    return np.stack([q[0] / q[3], q[1] / q[3], q[2] / q[3]], axis=1)


def _dist_rows(P, q):
    """||P_i - q|| exactly as distance_cython (pyx:6-7): sqrt((dx*dx + dy*dy) + dz*dz), float64."""
    dx = P[:, 0] - q[0]
    dy = P[:, 1] - q[1]
    dz = P[:, 2] - q[2]
    return np.sqrt((dx * dx + dy * dy) + dz * dz)


def min_distance(A, B):
    """np.min(A.compute_point_cloud_distance(B)) (:397-484): exact minimum point-to-point distance in
    float64, evaluated as sqrt((dx*dx + dy*dy) + dz*dz).  A KD-tree proposes the closest pairs, the
    pinned formula decides."""
    A = np.asarray(A, np.float64)
    B = np.asarray(B, np.float64)
    if len(A) == 0 or len(B) == 0:
        return np.inf
    if len(A) > len(B):
        A, B = B, A
    tree = cKDTree(B)
    d, _ = tree.query(A, k=1)
    dmin = d.min()
    cand = np.nonzero(d <= dmin * (1 + 1e-9) + 1e-300)[0]
    best = np.inf
    for i in cand:
        js = tree.query_ball_point(A[i], dmin * (1 + 1e-9) + 1e-300)
        if js:
            best = min(best, _dist_rows(B[js], A[i]).min())
    return float(best)


def pairwise_median(obj):
    """np.median(pdist(objects[i])) (:1524)."""
    return float(np.median(pdist(obj)))


# --------------------------------------------------------------------------------------------------
# a5: the region filter
# --------------------------------------------------------------------------------------------------
def find_seed(pcd, center):
    """pyx:18-31 -- first point whose float32 distance to `center` is strictly smallest."""
    d32 = _dist_rows(pcd, center).astype(np.float32)
    best = np.float32(50000)
    arg = -1
    m = d32.min()
    if m < best:
        arg = int(np.nonzero(d32 == m)[0][0])
    return arg


def region_filter_restated(pcd, center, median_distance, rng):
    """filter_cython_new.pyx:9-93 restated (hand_cloud / obj_center collapse: the caller passes the
    same array for both, :1525).  `rng` is a random.Random; choice() is the reference's
    random.choice(list(dicti.values())) (pyx:78).  Returns (m,3) float64: dicti[0]=seed, then
    accepted points pass by pass in scan order -- the seed is accepted again in pass 1."""
    pcd = np.ascontiguousarray(pcd, np.float64)
    n = len(pcd)
    if n == 2:                 # pyx:35-42: i == 0 after a one-iteration loop -> zeros((1,3))
        return np.zeros((1, 3))
    seed = find_seed(pcd, center)
    radius = float(np.float32(median_distance))      # cdef float mean_distance = median_distance
    out = [pcd[seed]]
    remaining = pcd
    pivot = pcd[seed]
    abort = 0
    while abort < FILTER_PATIENCE:
        hit = _dist_rows(remaining, pivot) <= radius
        if hit.any():
            out.extend(remaining[hit])
            remaining = remaining[~hit]
            abort = 0
        else:
            abort += 1
        pivot = rng.choice(out)                       # pyx:78, drawn every pass
    return np.array(out)


def closure_filter(pcd, center, median_distance, return_mask=False):
    """Deterministic fixed point of the reference's flood fill: every point reachable from the seed
    through hops of length <= float32(median_distance).  Output order: seed, then the component in
    input (raster) order -- same multiset the reference returns whenever its random pivots reach the
    closure (measured 87.5% of runs)."""
    pcd = np.ascontiguousarray(pcd, np.float64)
    n = len(pcd)
    if n == 2:
        out = np.zeros((1, 3))
        return (out, np.zeros(n, bool), -1) if return_mask else out
    seed = find_seed(pcd, center)
    radius = float(np.float32(median_distance))
    member = np.zeros(n, bool)
    frontier = np.array([seed])
    while len(frontier):
        rest = np.nonzero(~member)[0]
        if len(rest) == 0:
            break
        hit = np.zeros(len(rest), bool)
        for f in frontier:
            todo = ~hit
            if not todo.any():
                break
            hit[todo] = _dist_rows(pcd[rest[todo]], pcd[f]) <= radius
        new = rest[hit]
        member[new] = True
        frontier = new
    out = np.vstack([pcd[seed][None], pcd[member]])
    return (out, member, seed) if return_mask else out


# --------------------------------------------------------------------------------------------------
# a7-a9: relation rows.  An entity is None (undefined), or a dict(n, lo, hi, center, pts).
# --------------------------------------------------------------------------------------------------
def make_entity(pts, is_hand=False):
    pts = np.asarray(pts, np.float64).reshape(-1, 3)
    if len(pts) == 0:
        return dict(n=0, pts=pts, lo=None, hi=None, center=None, mean=None)
    lo, hi = pts.min(0), pts.max(0)
    mean = pts.mean(axis=0)
    # DSR centres (:1173-1353): hand -> cloud mean (hand.get_center()); objects/ground -> AABB centre
    return dict(n=len(pts), pts=pts, lo=lo, hi=hi, mean=mean,
                center=mean if is_hand else (lo + hi) * 0.5)


def fill_tn(ent, absent, thresh, mind):
    """_fillTN_absent :363-487.  ent: dict H,1,2,3,G; absent: dict '1','2','3' (mutated, sticky);
    mind(a, b) -> exact min distance.  Returns 10 codes."""
    out = [None] * 10
    H = ent["H"]
    for r, k in enumerate("123"):                       # rows 0-2 (:389-427)
        o = ent[k]
        if o is None:
            out[r] = "U"
        elif H["n"] > 0:
            if o["n"] == 0 or absent[k]:
                out[r] = "A"
                absent[k] = True
            elif mind("H", k) < thresh:
                out[r] = "T"
            else:
                out[r] = "N"
    if H["n"] > 0:                                      # row 3 (:429-433)
        out[3] = "T" if mind("H", "G") < thresh else "N"
    for r, (a, b) in ((4, "12"), (5, "13"), (7, "23")):  # (:435-442, :444-451, :462-469)
        if ent[a] is None or ent[b] is None:
            out[r] = "U"
        elif ent[a]["n"] == 0 or ent[b]["n"] == 0 or absent[a] or absent[b]:
            out[r] = "A"
        elif mind(a, b) < thresh:
            out[r] = "T"
        else:
            out[r] = "N"
    for r, a in ((6, "1"), (8, "2"), (9, "3")):           # (:453-460, :471-487)
        if ent[a] is None:
            out[r] = "U"
        elif ent[a]["n"] == 0 or absent[a]:
            out[r] = "A"
        elif mind(a, "G") < thresh:
            out[r] = "T"
        else:
            out[r] = "N"
    return out


def _box(e):
    """SSR axis naming (:660-675): x = data[:,0], z = data[:,1], y = data[:,2]."""
    lo, hi = e["lo"], e["hi"]
    return dict(min_x=lo[0], max_x=hi[0], min_z=lo[1], max_z=hi[1], min_y=lo[2], max_y=hi[2])


def _ov(a, b, u):
    return a["min_" + u] < b["max_" + u] and a["max_" + u] > b["min_" + u]


def _Ar(a, b):   # :491-568
    R = a["max_x"] > b["max_x"] and _ov(a, b, "y") and _ov(a, b, "z")
    L = a["min_x"] < b["min_x"] and _ov(a, b, "y") and _ov(a, b, "z")
    F = a["max_z"] > b["max_z"] and (a["min_x"] < b["max_x"] and b["min_x"] < a["max_x"]) and _ov(a, b, "y")
    Ba = a["min_z"] < b["min_z"] and (a["min_x"] < b["max_x"] and b["min_x"] < a["max_x"]) and _ov(a, b, "y")
    return R or L or F or Ba


def _tail(a, b, touching):
    """The elif chain shared by every SSR row (e.g. :729-745)."""
    if _Ar(a, b):
        return "HArT" if touching else "HAr"
    if a["max_y"] > b["max_y"] or a["min_y"] < b["min_y"]:       # _Ab or _Be (:570-599)
        return "VArT" if touching else "VAr"
    return "O"


def _between(m, a, b):   # one disjunct of _Bw :615-633
    return (m["min_x"] >= min(a["max_x"], b["max_x"]) and m["max_x"] <= max(a["min_x"], b["min_x"]) and
            m["min_y"] >= max(a["min_y"], b["min_y"]) and m["max_y"] <= min(a["max_y"], b["max_y"]) and
            m["min_z"] >= max(a["min_z"], b["min_z"]) and m["max_z"] <= min(a["max_z"], b["max_z"]))


def fill_ssr(ent, tn):
    """_fillSSR_8 :636-1061.  tn = the 10 T/N codes of the same column."""
    out = [None] * 10
    H = _box(ent["H"])
    G = _box(ent["G"])
    B = {}
    for r, k in enumerate("123"):
        if ent[k] is not None and tn[r] != "A":
            B[k] = _box(ent[k])
    all_there = all(ent[k] is not None for k in "123")
    for r, k in enumerate("123"):                       # rows 10-12 (:706-812)
        if ent[k] is None:
            out[r] = "U"
        elif tn[r] == "A":
            out[r] = "A"
        else:
            o = B[k]
            if ((o["min_z"] < H["min_z"] and o["max_z"] > H["max_z"]) and
                    ((o["max_x"] < H["min_x"] and o["max_x"] > H["max_x"]) or
                     (o["min_x"] < H["min_x"] and o["min_x"] > H["max_x"]) or
                     (o["min_x"] > H["min_x"] and o["max_x"] < H["max_x"])) and
                    (H["min_y"] < o["max_y"] <= H["max_y"])):
                out[r] = "Sa"
            else:
                out[r] = _tail(H, o, tn[r] == "T")
    out[3] = _tail(H, G, tn[3] == "T")                  # row 13 (:814-842)
    bw_gate = all_there and tn[0] != "A" and tn[1] != "A" and tn[2] != "A"
    for r, (j, k), bw_args in ((4, "12", 12), (5, "13", 13), (7, "23", 23)):   # rows 14,15,17
        if ent[j] is None or ent[k] is None:
            out[r] = "U"
        elif tn[r] == "A":
            out[r] = "A"
        else:
            a, b = B[j], B[k]
            if ((b["min_z"] < a["min_z"] and b["max_z"] > a["max_z"]) and
                    ((a["max_x"] > b["min_x"] and a["max_x"] < b["max_x"]) or
                     (a["min_x"] > b["min_x"] and a["min_x"] < b["max_x"]) or
                     (a["min_x"] < b["min_x"] and a["max_x"] > b["max_x"])) and
                    (a["min_y"] < b["max_y"] <= a["max_y"])):
                out[r] = "In"
            elif ((a["min_z"] < b["min_z"] and a["max_z"] > b["max_z"]) and
                  ((a["max_x"] < b["min_x"] and a["max_x"] > b["max_x"]) or
                   (a["min_x"] < b["min_x"] and a["min_x"] > b["max_x"]) or
                   (a["min_x"] > b["min_x"] and a["max_x"] < b["max_x"])) and
                  (b["min_y"] < a["max_y"] <= b["max_y"])):
                out[r] = "Sa"
            else:
                out[r] = _tail(a, b, tn[r] == "T")
        if bw_gate:                                     # :884-888, :925-929, :994-998
            B1 = _between(B["1"], B["2"], B["3"])
            B2 = _between(B["2"], B["1"], B["3"])
            B3 = _between(B["3"], B["1"], B["2"])
            # Python's `x == k and A or B` precedence in _Bw (:621-633), SURVEY.md 8a.2
            bw = (B2 or B3) if bw_args == 23 else (B1 or B2 or B3)
            if bw:
                out[r] = "Bw"
    for r, k in ((6, "1"), (8, "2"), (9, "3")):           # rows 16,18,19
        if ent[k] is None:
            out[r] = "U"
        elif tn[r] == "A":
            out[r] = "A"
        else:
            out[r] = _tail(B[k], G, tn[r] == "T")
    return out


def _dist3(p, q):   # _distance :200-209
    return np.sqrt((p[0] - q[0]) ** 2 + (p[1] - q[1]) ** 2 + (p[2] - q[2]) ** 2)


def fill_dsr(ent, prev, tn):
    """_fillDSR_new_2 :1064-1353.  prev = entity dict of the previous kept frame (previous_array)."""
    cd, sd = CENTER_DISTANCE, STABLE_DIST
    out = [None] * 10

    def moved(k):
        if k == "G":        # the same ground object is in previous_array (:1799, :1822)
            return _dist3(ent["G"]["center"], prev["G"]["center"]) > cd
        return _dist3(ent[k]["center"], prev[k]["center"]) > cd

    def gap(a, b, e):
        if b == "G":        # rows 23,26,28,29 use data axis 2 only (:1233, :1291, ...)
            return abs(e[a]["center"][2] - e["G"]["center"][2])
        return _dist3(e[a]["center"], e[b]["center"])

    for r, (a, b) in enumerate(ROW_PAIRS):
        objs = [k for k in (a, b) if k in "123"]
        if any(ent[k] is None or prev[k] is None for k in objs):
            out[r] = "U"
        elif tn[r] == "A":
            out[r] = "A"
        elif tn[r] == "T":
            out[r] = "MT" if (moved(a) or moved(b)) else "HT"
        else:
            d, pd_ = gap(a, b, ent), gap(a, b, prev)
            if tn[r] == "N" and abs(d - pd_) < sd:
                out[r] = "S"
            elif d > pd_:
                out[r] = "MA"
            elif d < pd_:
                out[r] = "GC"
            else:
                out[r] = "Q"
    return out


# --------------------------------------------------------------------------------------------------
# a11: eSEC -> e2SEC
# --------------------------------------------------------------------------------------------------
def esec_to_e2sec(table, relations, final_columns):
    """esec_to_e2sec :62-111.  table: (rows, c) array of str; returns (e2sec, esec_upper, final_columns).
    Quirk kept: the *first column is deleted unconditionally* (:83-84) although for relations==3 it is
    a real column (:1826-1829 replaces the '-' column), and final_columns is only thinned by the
    duplicate indices k, so it stays one longer than the matrix."""
    esec = np.char.upper(np.asarray(table, dtype="U5"))
    e2 = np.delete(esec, 0, 1)
    rows = {3: [3, 9, 13, 19, 23, 29], 2: [3, 9, 13, 19], 1: [3, 9]}[relations]
    e2 = np.delete(e2, rows, 0)
    k = [j + 1 for j in range(e2.shape[1] - 1) if np.array_equal(e2[:, j], e2[:, j + 1])]
    e2 = np.delete(e2, k, 1)
    fc = np.delete(np.asarray(final_columns), k, 0)
    return e2, esec, fc


# --------------------------------------------------------------------------------------------------
# the per-frame pass and the sequence driver
# --------------------------------------------------------------------------------------------------
class E2SecOracle:
    """Sequential restatement of _process for relations==3 / cython=True / icp=False.
    State mirrors the module globals manipulation_analyser.py:32-59."""

    def __init__(self, ground_label, hand_label, support_hand, thresh, translation, roll, ground_pts,
                 total_unique_labels, replace=False, old=(), new=(), relations=3,
                 filter_mode="closure", seed=0, ref_filter=None, keep_details=False,
                 radius_override=None):
        self.ground_label, self.hand_label, self.support_hand = ground_label, hand_label, support_hand
        self.thresh = float(thresh)
        self.T = np.asarray(translation, np.float64)
        self.R = np.asarray(roll, np.float64)
        self.ground = make_entity(ground_pts)
        self.tul = np.asarray(total_unique_labels, np.float64)
        self.hand_idx = int(np.where(self.tul == hand_label)[0][0])
        self.ground_idx = int(np.where(self.tul == ground_label)[0][0])
        self.support_idx = None if support_hand is None else int(np.where(self.tul == support_hand)[0][0])
        self.replace, self.old, self.new = replace, list(old), list(new)
        self.relations = relations
        self.filter_mode, self.seed, self.ref_filter = filter_mode, seed, ref_filter
        self.keep_details = keep_details
        # test speed-up only: {(frame, label_index): median_distance} replaces the O(n^2) pdist call
        self.radius_override = radius_override or {}
        # globals :32-59
        self.o = {"1": None, "2": None, "3": None}
        self.count = {"1": 0, "2": 0, "3": 0}
        self.olabel = {"1": 0, "2": 0, "3": 0}
        self.absent = {"1": False, "2": False, "3": False}
        self.previous = None
        self.internal_count = 0
        self.table = None
        self.final_columns = []
        self.details = []

    @staticmethod
    def total_unique_labels_from_frame0(label0, hand_label, replace, old, ignored_labels, cut_labels=()):
        """:1406-1426."""
        tul = np.unique(label0)
        if replace:
            for v in old:
                tul = np.delete(tul, np.where(tul == v))
        for v in ignored_labels:
            if v in tul:
                tul = np.delete(tul, np.where(tul == v))
        if hand_label not in tul:
            if len(cut_labels):
                tul = np.append(tul, cut_labels)
            tul = np.append(tul, hand_label)
        elif len(cut_labels):
            tul = np.append(tul, cut_labels)
        return tul

    # -- B/C: segment + filter every entity of one frame ------------------------------------------
    def entities_of_frame(self, frame, xyz32, label_full_flat):
        """:1441-1546 -> list over total_unique_labels of filtered (m,3) arrays (empty if absent),
        plus per-entity details."""
        cloud = np.nan_to_num(np.asarray(xyz32).astype(np.float64), nan=NAN_SENTINEL)
        valid = ~((cloud == NAN_SENTINEL).any(axis=1) | (label_full_flat == NAN_SENTINEL))
        filtered, det = [], []
        for i, value in enumerate(self.tul):
            idx = np.nonzero(valid & (label_full_flat == float(value)))[0]   # raster order (pin)
            d = dict(n_raw=len(idx), radius=np.nan, seed=-1, filtered=False)
            if len(idx) == 0:
                filtered.append(np.zeros((0, 3)))
                det.append(d)
                continue
            obj = transform_points(transform_points(cloud[idx], self.T), self.R)
            if i != self.ground_idx and i != 0 and len(obj) > 3:             # :1520
                center = np.median(obj, axis=0)                              # :1523
                radius = self.radius_override.get((frame, i))
                if radius is None:
                    radius = pairwise_median(obj)                            # :1524
                sd = find_seed(obj, center)
                d.update(radius=radius, seed=sd, seed_raster=int(idx[sd]), filtered=True, median=center)
                if self.filter_mode == "closure":
                    out = closure_filter(obj, center, radius)
                else:
                    s = filter_seed(self.seed, frame, i)
                    if self.ref_filter is not None:       # the reference's compiled pyx, unchanged
                        random.seed(s)
                        out = np.asarray(self.ref_filter.region_filter_cython(
                            np.ascontiguousarray(obj.mean(0)), np.ascontiguousarray(obj),
                            int(i == self.hand_idx), center, center, radius))
                    else:
                        out = region_filter_restated(obj, center, radius, random.Random(s))
                filtered.append(out)
            else:
                filtered.append(obj)
            det.append(d)
        return filtered, det

    # -- D-F ---------------------------------------------------------------------------------------
    def process_frame(self, frame, xyz32, label):
        """One call of _process (:1355).  `label` is the raw (rows, cols) label image of the file."""
        label = np.asarray(label, np.float64)
        if self.hand_label not in np.unique(label):                          # :1429 early-out #1
            return False
        h = 480 if len(xyz32) == 307200 else None
        if label.size != len(xyz32):
            assert h is not None, "label upscaling is only defined for the 640x480 organised cloud"
            label = resize_nearest(label, 640, 480)                          # :1433
        if self.replace:
            label = replace_labels(label, self.old, self.new)                # :1434-1435
        flat = label.flatten()
        filtered, det = self.entities_of_frame(frame, xyz32, flat)
        if len(filtered[self.hand_idx]) == 0:                                # :1475, :1551 early-outs #2/#3
            return False
        ents = [make_entity(p, is_hand=(i == self.hand_idx)) for i, p in enumerate(filtered)]
        hand = ents[self.hand_idx]
        cache = {}

        def md(pa, pb):
            key = (id(pa), id(pb))
            if key not in cache:
                cache[key] = min_distance(pa, pb)
            return cache[key]

        # ---- D: o1/o2/o3 selection :1562-1658 (support_hand variant :1563-1598 has fewer o3 branches)
        o, cnt, lab = self.o, self.count, self.olabel
        for i in range(len(self.tul)):
            if i == self.hand_idx or i == self.ground_idx or i <= 0 or ents[i]["n"] == 0:
                continue
            if self.support_idx is not None and i == self.support_idx:
                continue
            ci = ents[i]
            near = lambda e: md(e["pts"], ci["pts"]) < self.thresh
            if cnt["1"] == 0 and near(hand):
                lab["1"], cnt["1"], o["1"] = i, 1, ci
            elif o["1"] is not None and cnt["2"] == 0 and i != lab["1"]:
                if near(hand) or near(o["1"]):
                    lab["2"], cnt["2"], o["2"] = i, 1, ci
            elif self.support_idx is not None:
                if o["2"] is not None and cnt["3"] == 0 and i != lab["1"] and i != lab["2"]:
                    if near(hand) or near(o["1"]) or near(o["2"]):
                        lab["3"], cnt["3"], o["3"] = i, 1, ci
            elif o["2"] is not None and cnt["3"] == 0 and i != lab["1"] and i != lab["2"]:
                n1, n2 = o["1"]["n"], o["2"]["n"]
                if n2 > 0 and n1 > 0:
                    ok = near(hand) or near(o["1"]) or near(o["2"])
                elif n2 > 0 and n1 == 0:
                    ok = near(hand) or near(o["2"])
                elif n1 > 0 and n2 == 0:
                    ok = near(hand) or near(o["1"])
                else:
                    ok = False
                if ok:
                    lab["3"], cnt["3"], o["3"] = i, 1, ci
        for k in "123":                                                      # icp == False :1669-1675
            if cnt[k] == 1:
                o[k] = ents[lab[k]]
        ent = {"H": hand, "G": self.ground, "1": o["1"], "2": o["2"], "3": o["3"]}
        if self.relations != 3:
            # relations 1 / 2 (:1744-1794): no priming frame, no previous_array; the '-' column the table was created
            # with (:2252-2260) is replaced only when file index 0 itself yields a column, final_columns is never filled
            pts = {k: (None if v is None else v["pts"]) for k, v in ent.items()}
            rows = fill_tn(ent, self.absent, self.thresh, lambda a, b: md(pts[a], pts[b]))
            if self.relations == 2:
                rows = rows + fill_ssr(ent, rows)
            add = np.array(rows, dtype="U5").reshape(-1, 1)
            if self.table is None:
                self.table = np.full((10 * self.relations, 1), "-", dtype="U5")
            if frame == 0:
                self.table = add
            elif np.array_equal(add[:, 0], self.table[:, -1]):
                return False
            else:
                self.table = np.column_stack((self.table, add))
            return True
        if self.internal_count == 0:                                         # :1798-1801
            self.previous = dict(ent)
            self.internal_count = 1
            return False
        pts = {k: (None if v is None else v["pts"]) for k, v in ent.items()}
        tn = fill_tn(ent, self.absent, self.thresh, lambda a, b: md(pts[a], pts[b]))
        ssr = fill_ssr(ent, tn)
        dsr = fill_dsr(ent, self.previous, tn)
        add = np.array(tn + ssr + dsr, dtype="U5").reshape(30, 1)
        self.previous = dict(ent)                                            # :1822
        emitted = False
        if self.internal_count == 1:                                         # :1826-1829
            self.final_columns.append(frame)
            self.table = add
            self.internal_count = 2
            emitted = True
        elif np.array_equal(add[:, 0], self.table[:, -1]):                   # :2001-2002
            pass
        else:                                                                # :2005-2007
            self.final_columns.append(frame)
            self.table = np.column_stack((self.table, add))
            emitted = True
        if self.keep_details:
            self.details.append(dict(frame=frame, add=add[:, 0].copy(), emitted=emitted, det=det,
                                     ents=ents, labels=dict(lab), counts=dict(cnt)))
        return emitted

    def finish(self):
        table = self.table if self.table is not None else np.full((10 * self.relations, 1), "-", dtype="U5")
        return esec_to_e2sec(table, self.relations, self.final_columns)


def analyse_sequence(frames, oracle, fps=10):
    """analyse_maniac_manipulation's loop :2266-2281: every int(30/fps)-th file, frame = file index.
    frames: sequence of (xyz32, label) or a callable index -> (xyz32, label)."""
    step = int(30 / fps)
    n = len(frames)
    for i in range(n):
        if i % step == 0:
            xyz, label = frames[i]
            oracle.process_frame(i, xyz, label)
    return oracle.finish()
