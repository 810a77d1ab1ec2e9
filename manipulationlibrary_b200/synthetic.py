"""Seeded synthetic labelled point-cloud streams (SURVEY.md 8d, configs 2-5 of BASELINE.json).

A stream is an organised W x H cloud per frame as the MANIAC recordings are (manipulation_analyser.py:1441 reads
640 x 480 clouds): a pinhole camera (f = 525 px at 640 columns, scaled with the width) looks obliquely at a
table.  Per frame: the table plane (ground label), a border ring (label 0, total_unique_labels index 0, which
the reference never relates, :1520), `n_objects` slabs standing on the table and a hand blob that runs a scripted
approach -> push -> pull back -> retreat cycle against one object per cycle, so the T/N, SSR and DSR rows all
change.  ~14% of the pixels are NaN (all-or-nothing per point, like the real data), depth noise sigma = 1 mm,
0.2-1% of every object's pixels are outliers displaced >= 0.3 m towards the camera so that the region filter
(filter_cython_new.pyx:9-93) rejects something.

Every frame depends only on (seed, config, frame index): any rank regenerates its own shard.  Pure numpy; this is
input synthesis for tests and bench.py, not part of the measured path.
"""
import math

import numpy as np

BORDER_LABEL = 0
GROUND_LABEL = 1
THRESH = 0.07


def _rng(seed, config, frame):
    # frame -2 / -1 are the stream-level draws (ground jitter, layout)
    return np.random.Generator(np.random.PCG64(np.random.SeedSequence([int(seed), int(config), int(frame) + 2])))


class SyntheticStream:
    def __init__(self, width=640, height=480, n_objects=10, seed=20261104, config=2, cycle=96, near_thresh=False):
        self.W, self.H, self.n_objects = int(width), int(height), int(n_objects)
        self.N = self.W * self.H
        self.seed, self.config, self.cycle = int(seed), int(config), int(cycle)
        self.hand_label = self.n_objects + 2
        self.ground_label = GROUND_LABEL
        self.object_labels = list(range(2, 2 + self.n_objects))
        self.total_unique_labels = np.arange(0, self.n_objects + 3, dtype=np.float64)
        self.thresh = THRESH
        W, H = self.W, self.H
        self.f = 525.0 * W / 640.0
        cx, cy = W / 2.0, H / 2.0
        u = (np.arange(W, dtype=np.float32) - np.float32(cx)) / np.float32(self.f)
        v = (np.arange(H, dtype=np.float32) - np.float32(cy)) / np.float32(self.f)
        self.rx = np.broadcast_to(u[None, :], (H, W)).astype(np.float32)
        self.ry = np.broadcast_to(v[:, None], (H, W)).astype(np.float32)
        phi = math.radians(55.0)
        self.a, self.b, self.c = math.cos(phi), math.sin(phi), 1.0
        # plane a*y + b*z = c  ->  z = c / (a*ry + b);  a slab of height h: z = (c - h) / (a*ry + b)
        self.inv_den = (1.0 / (np.float32(self.a) * self.ry + np.float32(self.b))).astype(np.float32)
        # camera -> table frame (height = axis 2), composed with the reference's roll = diag(1, -1, -1, 1)
        a, b, c = self.a, self.b, self.c
        self.translation = np.array([[1, 0, 0, 0], [0, -b, a, 0], [0, a, b, -c], [0, 0, 0, 1]], np.float64)
        self.roll = np.diag([1.0, -1.0, -1.0, 1.0])
        self._layout(near_thresh)
        self._static()
        self.ground = self._ground_cloud()

    # ---- scene -----------------------------------------------------------------------------------
    def _layout(self, near_thresh):
        W, H, n = self.W, self.H, self.n_objects
        r = _rng(self.seed, self.config, -1)
        self.ring = max(2, int(round(9 * W / 640.0)))          # border-label ring (~6.5% of the pixels)
        self.nan_band = int(round(40 * W / 640.0))             # NaN columns on the left (~6%)
        x0, x1 = self.nan_band + self.ring, W - self.ring
        y0, y1 = self.ring, H - self.ring
        cols = max(2, int(math.ceil(math.sqrt(n * 1.2 * (x1 - x0) / float(y1 - y0)))))
        rows = int(math.ceil((n + 1) / float(cols)))
        while cols * rows < n + 1:
            rows += 1
        cw, ch = (x1 - x0) / float(cols), (y1 - y0) / float(rows)
        cells = [(cc, rr) for rr in range(rows) for cc in range(cols)]
        self.cols, self.rows, self.cw, self.ch = cols, rows, cw, ch
        self.home_cell = cells[-1]                                # the hand hovers over the last cell
        self.obj = []
        for k in range(n):
            cc, rr = cells[k]
            fw, fh = r.uniform(0.38, 0.70), r.uniform(0.38, 0.70)
            w, h = int(cw * fw), int(ch * fh)
            ox = int(x0 + cc * cw + (cw - w) * r.uniform(0.35, 0.65))
            oy = int(y0 + rr * ch + (ch - h) * r.uniform(0.35, 0.65))
            height = float(r.uniform(0.04, 0.15))
            if near_thresh and k % 2 == 1:
                # config 4: neighbours whose gap is within +-5 mm of thresh (exercises the exact recheck)
                height = self.obj[k - 1]["height"] + float(r.uniform(-0.004, 0.004))
            self.obj.append(dict(x=ox, y=oy, w=w, h=h, height=height, ellipse=bool(r.random() < 0.4),
                                 outlier_p=float(r.uniform(0.002, 0.01)), label=2 + k, cell=(cc, rr)))
        if near_thresh:
            # pull odd objects towards their left neighbour until the lateral gap is ~thresh in 3-D
            for k in range(1, n, 2):
                A, B = self.obj[k - 1], self.obj[k]
                if A["cell"][1] != B["cell"][1]:
                    continue
                z = (self.c - A["height"]) / (self.a * ((A["y"] + A["h"] / 2 - H / 2) / self.f) + self.b)
                gap_px = int(round((THRESH + float(r.uniform(-0.005, 0.005))) * self.f / z))
                B["x"] = min(A["x"] + A["w"] + gap_px, W - self.ring - B["w"] - 1)
        self.hand_rx, self.hand_ry = max(4, int(35 * W / 640.0)), max(3, int(25 * W / 640.0))
        x0h = x0 + self.home_cell[0] * cw + cw / 2
        y0h = y0 + self.home_cell[1] * ch + ch / 2
        self.hand_home = (x0h, y0h, 0.28)
        self.interior = (x0, x1, y0, y1)

    def _mask(self, o):
        if not o["ellipse"]:
            return np.ones((o["h"], o["w"]), bool)
        yy = (np.arange(o["h"]) + 0.5) / o["h"] * 2 - 1
        xx = (np.arange(o["w"]) + 0.5) / o["w"] * 2 - 1
        return (yy[:, None] ** 2 + xx[None, :] ** 2) <= 1.0

    def _static(self):
        W, H = self.W, self.H
        lab = np.full((H, W), GROUND_LABEL, np.int32)
        hgt = np.zeros((H, W), np.float32)
        outp = np.zeros((H, W), np.float32)
        for o in self.obj:
            o["mask"] = self._mask(o)
        rg = self.ring
        self._ring_mask = np.zeros((H, W), bool)
        self._ring_mask[:rg] = self._ring_mask[-rg:] = True
        self._ring_mask[:, :rg] = self._ring_mask[:, -rg:] = True
        self._base = (lab, hgt, outp)

    def _draw(self, lab, hgt, outp, o, dx=0, dy=0):
        x, y = o["x"] + dx, o["y"] + dy
        xa, ya = max(x, 0), max(y, 0)
        xb, yb = min(x + o["w"], self.W), min(y + o["h"], self.H)
        if xb <= xa or yb <= ya:
            return
        m = o["mask"][ya - y: yb - y, xa - x: xb - x]
        lab[ya:yb, xa:xb][m] = o["label"]
        hgt[ya:yb, xa:xb][m] = o["height"]
        outp[ya:yb, xa:xb][m] = o["outlier_p"]

    def _ground_cloud(self):
        """Stands in for the frame-0 ground of _getGroundiNewNew (:278-361): plane points on a 0.02 grid (the
        reference voxel-downsamples at 0.02, :359) over the visible table, already in the final frame."""
        a, b, c = self.a, self.b, self.c
        x0, x1, y0, y1 = self.interior
        corners = []
        for (uu, vv) in ((x0, y0), (x1, y0), (x0, y1), (x1, y1)):
            ry = (vv - self.H / 2.0) / self.f
            z = c / (a * ry + b)
            corners.append(((uu - self.W / 2.0) / self.f * z, ry * z, z))
        cam = np.array(corners)
        fin = self.to_final(cam)
        lo, hi = fin.min(0), fin.max(0)
        gx = np.arange(lo[0], hi[0], 0.02)
        gy = np.arange(lo[1], hi[1], 0.02)
        g = np.stack(np.meshgrid(gx, gy, indexing="ij"), -1).reshape(-1, 2)
        r = _rng(self.seed, self.config, -2)
        z = r.normal(0.0, 0.0005, len(g))
        return np.ascontiguousarray(np.column_stack([g, z]))

    def to_final(self, cam_pts):
        p = np.asarray(cam_pts, np.float64)
        q = np.column_stack([p, np.ones(len(p))]) @ self.translation.T
        q = q @ self.roll.T
        return q[:, :3] / q[:, 3:4]

    # ---- script ----------------------------------------------------------------------------------
    def script(self, frame):
        """-> (target object index, dx, dy of the pushed object, hand centre (x, y) px, hand height m)."""
        L = self.cycle
        cyc, t = divmod(int(frame), L)
        t = t / float(L)
        k = cyc % self.n_objects
        o = self.obj[k]
        cc, rr = o["cell"]
        # push towards the right neighbour if there is one in the same row, else to the left
        right = (k + 1 < self.n_objects and self.obj[k + 1]["cell"][1] == rr)
        sgn = 1 if right else -1
        nb = self.obj[k + 1] if right else (self.obj[k - 1] if k > 0 and self.obj[k - 1]["cell"][1] == rr else None)
        if nb is not None:
            room = (nb["x"] - (o["x"] + o["w"])) if right else (o["x"] - (nb["x"] + nb["w"]))
            push = max(0, room - 2)
        else:
            push = int(self.cw * 0.2)
        push = min(push, int(self.cw * 0.45))
        ocx, ocy = o["x"] + o["w"] / 2.0, o["y"] + o["h"] / 2.0
        contact_x = ocx - sgn * (o["w"] / 2.0 + self.hand_rx * 0.6)      # hand sits on the far side of the object
        hx0, hy0, hh0 = self.hand_home
        touch_h = o["height"] + 0.03

        def lerp(p, q, s):
            return p + (q - p) * s

        if t < 0.30:                                   # approach
            s = t / 0.30
            return k, 0, 0, (lerp(hx0, contact_x, s), lerp(hy0, ocy, s)), lerp(hh0, touch_h, s)
        if t < 0.55:                                   # push
            s = (t - 0.30) / 0.25
            d = int(round(sgn * push * s))
            return k, d, 0, (contact_x + d, ocy), touch_h
        if t < 0.75:                                   # pull back
            s = (t - 0.55) / 0.20
            d = int(round(sgn * push * (1 - s)))
            return k, d, 0, (contact_x + d, ocy), touch_h
        s = (t - 0.75) / 0.25                          # retreat
        return k, 0, 0, (lerp(contact_x, hx0, s), lerp(ocy, hy0, s)), lerp(touch_h, hh0, s)

    # ---- frames ----------------------------------------------------------------------------------
    def frame(self, frame, hand=True):
        """-> xyz (N, 3) float32 with NaN rows, labels (N,) int32."""
        W, H = self.W, self.H
        lab, hgt, outp = (x.copy() for x in self._base)
        k, dx, dy, (hx, hy), hh = self.script(frame)
        for j, o in enumerate(self.obj):
            if j != k:
                self._draw(lab, hgt, outp, o)
        self._draw(lab, hgt, outp, self.obj[k], dx, dy)
        if hand:
            rxh, ryh = self.hand_rx, self.hand_ry
            xa, xb = int(max(hx - rxh, 0)), int(min(hx + rxh + 1, W))
            ya, yb = int(max(hy - ryh, 0)), int(min(hy + ryh + 1, H))
            if xb > xa and yb > ya:
                yy = (np.arange(ya, yb) - hy) / ryh
                xx = (np.arange(xa, xb) - hx) / rxh
                m = (yy[:, None] ** 2 + xx[None, :] ** 2) <= 1.0
                lab[ya:yb, xa:xb][m] = self.hand_label
                hgt[ya:yb, xa:xb][m] = np.float32(hh)
                outp[ya:yb, xa:xb][m] = 0.004
        lab[self._ring_mask] = BORDER_LABEL
        r = _rng(self.seed, self.config, frame)
        noise = r.standard_normal((H, W), dtype=np.float32) * np.float32(0.001)
        uni = r.random((H, W), dtype=np.float32)
        z = (np.float32(self.c) - hgt) * self.inv_den + noise
        out = uni < outp                                  # outliers: >= 0.3 m towards the camera
        z = np.where(out, z - np.float32(0.3) - (uni / np.maximum(outp, np.float32(1e-9))) * np.float32(0.3), z)
        nan = (uni > np.float32(0.915))                   # ~8.5% random holes
        nan[:, : self.nan_band] = True                    # + the band = ~14%
        xyz = np.empty((H, W, 3), np.float32)
        xyz[..., 0] = self.rx * z
        xyz[..., 1] = self.ry * z
        xyz[..., 2] = z
        xyz[nan] = np.nan
        return xyz.reshape(-1, 3), lab.reshape(-1)

    def frames(self, start, stop, hand_every=None):
        n = stop - start
        xyz = np.empty((n, self.N, 3), np.float32)
        lab = np.empty((n, self.N), np.int32)
        for i in range(n):
            xyz[i], lab[i] = self.frame(start + i)
        return xyz, lab

    def fetch(self, frame_ids):
        """Frames by id -> xyz (F, N, 3) float32, labels (F, N) int32 (the `source` protocol of sharding.run_shard)."""
        ids = [int(i) for i in frame_ids]
        xyz = np.empty((len(ids), self.N, 3), np.float32)
        lab = np.empty((len(ids), self.N), np.int32)
        for i, f in enumerate(ids):
            xyz[i], lab[i] = self.frame(f)
        return xyz, lab

    def params(self, filter_mode="closure", filter_seed=0, radius_mode="bounded"):
        """Keyword arguments of _native.make_params for this stream."""
        return dict(total_unique_labels=self.total_unique_labels, hand_label=self.hand_label,
                    ground_label=self.ground_label, support_hand=None, thresh=self.thresh,
                    translation=self.translation, roll=self.roll, ground=self.ground,
                    filter_mode=filter_mode, filter_seed=filter_seed, radius_mode=radius_mode)
