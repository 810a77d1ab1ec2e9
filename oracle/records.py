"""TEST INFRASTRUCTURE ONLY -- phase-1 records (include/e2sec.h layout) computed by the CPU oracle.

RecordOracle runs oracle.esec_oracle's segmentation + filter on each frame and packs, per frame, exactly the
record the CUDA path emits: entity stats, the in-frame pair table and the cross table against the previous
kept frame (the stale o1/o2/o3 clouds of manipulation_analyser.py:1562-1658, see the header comment there).
Used by tests to (a) compare CUDA records field by field and (b) drive the host resolver without a GPU.
"""
import numpy as np

from . import esec_oracle as eo

REC_HEADER, REC_ENTITY = 4, 16


def record_doubles(K):
    return REC_HEADER + REC_ENTITY * K + (K + 1) * (K + 1) + K * K + 2 * K


class RecordOracle:
    def __init__(self, oracle, exact_far=False):
        """oracle: an eo.E2SecOracle (used for its parameters and entities_of_frame only)."""
        self.o = oracle
        self.K = len(oracle.tul)
        self.prev = None          # (frame_id, filtered clouds) of the previous kept frame
        self.exact_far = exact_far

    def _md(self, a, b):
        d = eo.min_distance(a, b)
        return d

    def frame_record(self, frame, xyz32, label_flat, hand_in_raw=None):
        o, K = self.o, self.K
        label_flat = np.asarray(label_flat, np.float64).reshape(-1)
        if hand_in_raw is None:
            hand_in_raw = bool((label_flat == o.hand_label).any())
        filtered, det = o.entities_of_frame(frame, xyz32, label_flat)
        rec = np.zeros(record_doubles(K))
        flags = 1 if hand_in_raw else 0
        stored = [(i != 0 or i == o.hand_idx) and i != o.ground_idx for i in range(K)]   # index 0 is only ever read as the hand (:1518-1520, :1549)
        if stored[o.hand_idx] and len(filtered[o.hand_idx]) > 0:
            flags |= 2
        rec[0], rec[1], rec[2] = frame, flags, K
        rec[3] = self.prev[0] if self.prev is not None else -1
        W = K + 1
        pt = np.full((W, W), np.inf)
        ct = np.full((K, K), np.inf)
        for e in range(K):
            r = rec[REC_HEADER + REC_ENTITY * e: REC_HEADER + REC_ENTITY * (e + 1)]
            d = det[e]
            r[0] = d["n_raw"]
            r[2], r[3], r[4:7] = -1, np.nan, np.nan
            if not stored[e]:
                continue
            p = filtered[e]
            r[1] = len(p)
            if d["filtered"]:
                r[3] = d["radius"]
                r[4:7] = d["median"]
                r[2] = d["seed_raster"]
            if len(p):
                r[7:10], r[10:13] = p.min(0), p.max(0)
                s = np.zeros(3)
                for row in p:          # sequential sum, like np.mean(axis=0) on an (m, 3) array
                    s += row
                r[13:16] = s / len(p)
        idx = [e for e in range(K) if stored[e]]
        for a in idx:
            for b in idx:
                if a < b:
                    pt[a, b] = pt[b, a] = self._md(filtered[a], filtered[b])
            pt[a, K] = pt[K, a] = self._md(filtered[a], o.ground["pts"])
            if self.prev is not None:
                for b in idx:
                    if a != b:
                        ct[a, b] = self._md(self.prev[1][a], filtered[b])
        off = REC_HEADER + REC_ENTITY * K
        rec[off: off + W * W] = pt.reshape(-1)
        rec[off + W * W: off + W * W + K * K] = ct.reshape(-1)
        rb = rec[off + W * W + K * K:].reshape(K, 2)          # radius bounds: the oracle knows the exact value
        for e in range(K):
            rb[e] = det[e]["radius"] if (stored[e] and det[e]["filtered"]) else np.nan
        if flags == 3:
            self.prev = (frame, filtered)
        return rec
