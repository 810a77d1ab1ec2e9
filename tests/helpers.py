"""Shared test helpers: golden fixtures -> dense frames."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

PUSHING06_ARGS = dict(ground_label=2, hand_label=11, support_hand=None, relations=3, replace=True,
                      old=[7, 9], new=[6, 5], ignored_labels=[8, 3], thresh=0.07)


def load_golden(name):
    return np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False)


class Pushing06Frames:
    """The sparsified Pushing_06 frames (tests/golden/pushing06_frames.npz) as dense organised clouds:
    xyz (307200,3) float32 with NaN where the fixture holds no point, labels (307200,) int32 with 255
    (an untracked label) there.  Labels are already upscaled and replaced (7->6, 9->5)."""

    def __init__(self):
        z = load_golden("pushing06_frames")
        self.frame_ids = z["frame_ids"]
        self.hand_in_raw = z["hand_in_raw"]
        self.offsets = z["offsets"]
        self.n_points = int(z["n_points"])
        self._idx_delta = z["idx_delta"]
        self._xyz = np.stack([z["x"], z["y"], z["z"]], axis=1)
        self._label = z["label"]

    def __len__(self):
        return len(self.frame_ids)

    def sparse(self, k):
        a, b = self.offsets[k], self.offsets[k + 1]
        idx = np.cumsum(self._idx_delta[a:b].astype(np.int64))
        return idx, self._xyz[a:b], self._label[a:b]

    def dense(self, k):
        idx, xyz, lab = self.sparse(k)
        full = np.full((self.n_points, 3), np.nan, np.float32)
        full[idx] = xyz
        labels = np.full(self.n_points, 255, np.int32)
        labels[idx] = lab
        if self.hand_in_raw[k] and not (lab == 11).any():
            # the raw label image held the hand label but every such point was NaN: keep the label visible
            labels[0] = 11
        return full, labels


REC_HEADER, REC_ENTITY = 4, 16


def upgrade_records(records, K):
    """Goldens written before ABI 2 lack the radius-bounds tail (K x 2): append lo = hi = the exact radius."""
    records = np.asarray(records)
    old = REC_HEADER + REC_ENTITY * K + (K + 1) * (K + 1) + K * K
    if records.shape[1] == old + 2 * K:
        return records
    assert records.shape[1] == old, records.shape
    rad = records[:, REC_HEADER + 3: REC_HEADER + REC_ENTITY * K: REC_ENTITY]
    return np.concatenate([records, np.repeat(rad, 2, axis=1)], axis=1)
