"""icp=True: rigid-model tracking of o1 / o2 / o3 (manipulation_analyser.py:1677-1738, the notebook's default).

With icp=True the reference does not relate the frame's own filtered clouds of o1 / o2 / o3 but MODEL clouds it carries from
frame to frame: every frame each model is registered onto the label's unfiltered cloud with point-to-point ICP
(max correspondence distance 0.02) and moved by the result; a frame whose filtered cloud is larger than the model replaces
it.  That makes the frames sequentially dependent through whole clouds, so this mode runs frame by frame:

    phase 1 of the frame on the device (records, e2sec_process_batch)         -- as always
    the clouds behind the record (e2sec_entity_cloud)                         -- models live on the host between frames
    selection (:1562-1658) with the MODELS as o1 / o2 (e2sec_min_distance)
    ICP per bound object (e2sec_icp_correspondences per step, 3 x 3 SVD here)  -- frame0.registration_icp
    relation rows of hand / ground / models (e2sec_min_distance, e2sec_relation_rows = csrc/relations.h)
    column emit / dedup (:1744-1829, :2000-2007)

Open3D's calls follow the definitions of oracle/fake_o3d.py (exact nearest neighbours, <= 30 ICP iterations, relative
tolerances 1e-6); tests/golden/pushing06_icp_closure.npz holds what the reference's own code gives with them.
"""
import numpy as np

from . import _native as nat
from . import frame0 as f0
from .engine import rec_pairs

ICP_DISTANCE = 0.02                 # thresh_registration :1678
PAIRS = [(0, 2), (0, 3), (0, 4), (0, 1), (2, 3), (2, 4), (2, 1), (3, 4), (3, 1), (4, 1)]     # rows of a family (:373-383)


def _entity(pts, is_hand=False):
    """what the relation rows read of a cloud (None stays None)"""
    if pts is None:
        return None
    if len(pts) == 0:
        z = np.zeros(3)
        return dict(n=0, lo=z, hi=z, center=z)
    lo, hi = pts.min(0), pts.max(0)
    return dict(n=len(pts), lo=lo, hi=hi, center=pts.mean(0) if is_hand else (lo + hi) * 0.5)


class Tracker:
    """The sequential state of _process with icp=True (globals :32-59): models, bindings, sticky flags, table."""

    def __init__(self, engine, total_unique_labels, hand_label, ground_label, support_hand, thresh, relations, ground_pts):
        tul = np.asarray(total_unique_labels, np.float64)
        self.eng, self.K = engine, len(tul)
        self.hand_idx = int(np.where(tul == hand_label)[0][0])
        self.ground_idx = int(np.where(tul == ground_label)[0][0])
        self.support_idx = None if support_hand is None else int(np.where(tul == support_hand)[0][0])
        self.thresh, self.relations = float(thresh), int(relations)
        self.ground_pts = np.ascontiguousarray(ground_pts, np.float64)
        self.ground = _entity(self.ground_pts)
        self.model = [None, None, None]          # o1..o3: None = the Python None, (0, 3) array = an empty cloud
        self.label = [0, 0, 0]
        self.count = [0, 0, 0]
        self.absent = [False, False, False]
        self.previous = None
        self.internal_count = 0
        self.columns, self.frames = [], []       # emitted columns (lists of codes) and their frame ids
        if self.relations < 3:
            self.columns.append(["-"] * (10 * self.relations))
            self.frames.append(-1)

    # -- distances ------------------------------------------------------------------------------------------------------
    def _near(self, a, b):
        if a is None or b is None or len(a) == 0 or len(b) == 0:
            return False                          # the reference's np.min over nothing raises; "far" here
        return self.eng.min_distance(a, b, limit=self.thresh) < self.thresh

    def _dist(self, a, b):
        if a is None or b is None or len(a) == 0 or len(b) == 0:
            return np.inf
        return self.eng.min_distance(a, b, limit=self.thresh)

    # -- one frame ------------------------------------------------------------------------------------------------------
    def frame(self, frame_id, slot, rec):
        """rec: the frame's record; slot: its position in the batch the engine processed last."""
        flags = int(rec[1])
        if flags & 3 != 3:                        # :1429, :1475 / :1551
            return
        K, th = self.K, self.thresh
        n_out = rec[nat.REC_HEADER + 1: nat.REC_HEADER + nat.REC_ENTITY * K: nat.REC_ENTITY].astype(np.int64)
        cache = {}

        def filt(i):
            if i not in cache:
                cache[i] = self.eng.entity_cloud(slot, i, True)
            return cache[i]

        hand = filt(self.hand_idx)
        pairs = rec_pairs(rec, K)
        # ---- selection :1562-1658, the models standing in for o1 / o2 ----------------------------------------------------
        for i in range(1, K):
            if i == self.hand_idx or i == self.ground_idx or i == self.support_idx or n_out[i] == 0:
                continue
            near_hand = pairs[min(self.hand_idx, i), max(self.hand_idx, i)] < th
            o1, o2 = self.model[0], self.model[1]
            take = -1
            if self.count[0] == 0 and near_hand:
                take = 0
            elif o1 is not None and self.count[1] == 0 and i != self.label[0]:
                if near_hand or self._near(o1, filt(i)):
                    take = 1
            elif o2 is not None and self.count[2] == 0 and i != self.label[0] and i != self.label[1]:
                if self.support_idx is not None or (len(o2) > 0 and len(o1) > 0):
                    ok = near_hand or self._near(o1, filt(i)) or self._near(o2, filt(i))
                elif len(o2) > 0 and len(o1) == 0:
                    ok = near_hand or self._near(o2, filt(i))
                elif len(o1) > 0 and len(o2) == 0:
                    ok = near_hand or self._near(o1, filt(i))
                else:
                    ok = False
                if ok:
                    take = 2
            if take >= 0:
                self.label[take], self.count[take], self.model[take] = i, 1, filt(i).copy()
        # ---- model tracking :1677-1738 -----------------------------------------------------------------------------------
        for k in range(3):
            if self.model[k] is None:
                continue
            f_k = filt(self.label[k])
            if len(f_k) > len(self.model[k]):      # a larger view of the object becomes the model (:1681-1683)
                self.model[k] = f_k.copy()
            raw = self.eng.entity_cloud(slot, self.label[k], False)
            if len(raw) != 0:
                T, _, _ = f0.registration_icp(self.eng, self.model[k], raw, ICP_DISTANCE)
                self.model[k] = f0.transform_points(self.model[k], T)
            else:
                self.model[k] = np.zeros((0, 3))
        # ---- relation rows ----------------------------------------------------------------------------------------------
        clouds = [hand, self.ground_pts] + self.model
        cur = [_entity(hand, True), self.ground] + [_entity(m) for m in self.model]
        if self.relations == 3 and self.internal_count == 0:     # :1798-1801 the first kept frame only primes
            self.previous, self.internal_count = cur, 1
            return
        d = [self._dist(clouds[a], clouds[b]) for a, b in PAIRS]
        codes, absent = nat.relation_rows(cur, self.previous if self.previous is not None else [None] * 5, d, th, self.absent,
                                          self.relations)
        self.absent = absent
        self.previous = cur
        if self.relations == 3:
            if self.internal_count == 1:
                self.internal_count, emit = 2, True
            else:
                emit = codes != self.columns[-1]
        elif frame_id == 0:                                       # :1753 / :1778 the '-' table is replaced
            self.columns, self.frames, emit = [], [], True
        else:
            emit = codes != self.columns[-1]
        if emit:
            self.columns.append(codes)
            self.frames.append(int(frame_id))

    def table(self):
        """-> (codes (10 * relations, n) 'S5', frame id per column) like Resolver.table()"""
        rows, n = 10 * self.relations, len(self.columns)
        t = np.empty((rows, n), "S5")
        for c, col in enumerate(self.columns):
            t[:, c] = [x.encode() for x in col]
        return t, list(self.frames)

    def state(self):
        return [self.label[k] if self.count[k] else -1 for k in range(3)], list(self.absent)
