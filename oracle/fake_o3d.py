"""TEST INFRASTRUCTURE ONLY -- a duck-typed stand-in for the subset of Open3D 0.11.2
(docs/requirements.txt:15 of the reference) that manipulation_analyser.py calls.

Open3D is a third-party dependency that is absent from /root/reference and not
installable here (no network).  This module restates the *published definitions*
of the calls on the e2SEC path so that the reference's own Python
(manipulation_analyser.py, unmodified, imported from /root/reference by
oracle/gen_golden.py) can be executed in this container to produce golden vectors:

  call site (manipulation_analyser.py)                 definition used here
  ---------------------------------------------------  -------------------------------------------
  io.read_point_cloud :1441                            oracle/pcd_io.read_pcd_xyz, float32 -> float64
  PointCloud.transform :1484-1485                      q = M @ [p,1]; p' = q[:3] / q[3]  (float64), see
                                                       transform_points() for the pinned operation order
  PointCloud.get_center :1522, :1173-1353              arithmetic mean of the points (duplicates count)
  compute_point_cloud_distance :397-484, :1570-1653    d[i] = min_j ||A_i - B_j||_2 in float64 (exact NN)
  get_axis_aligned_bounding_box/get_box_points :672-   per-axis min / max; 8 corners
  AxisAlignedBoundingBox.get_center :1173-1353         (min_bound + max_bound) * 0.5
  voxel_down_sample :224, :359                         mean of the points in each voxel, origin min_bound - voxel/2
  segment_plane :353                                   RANSAC (ransac_n points, num_iterations), seeded here
  OrientedBoundingBox.create_from_points :227          PCA box, extent sorted by decreasing variance
  registration_icp :260                                point-to-point ICP, <=30 iterations, rel. tol 1e-6

The frame-0 functions (last four rows) are random in the reference itself
(unseeded RANSAC, random.uniform plane); they are only used to make a fixture
(translation, roll, ground cloud) that every implementation then takes as an INPUT.
"""
import types

import numpy as np
from scipy.spatial import cKDTree

from . import pcd_io

_RNG = np.random.default_rng(20261104)


def seed(s):
    global _RNG
    _RNG = np.random.default_rng(s)


def transform_points(pts, M):
    """Pinned restatement of PointCloud::Transform: for every point
    q_r = ((M[r,0]*x + M[r,1]*y) + M[r,2]*z) + M[r,3] (separate IEEE mul/add, no FMA),
    p' = q[:3] / q[3].  The CUDA path uses exactly this order (__dmul_rn/__dadd_rn)."""
    pts = np.asarray(pts, np.float64)
    M = np.asarray(M, np.float64)
    x, y, z = pts[:, 0], pts[:, 1], pts[:, 2]
    q = [((M[r, 0] * x + M[r, 1] * y) + M[r, 2] * z) + M[r, 3] for r in range(4)]
    return np.stack([q[0] / q[3], q[1] / q[3], q[2] / q[3]], axis=1)


class AxisAlignedBoundingBox:
    def __init__(self, lo, hi):
        self.min_bound = np.asarray(lo, np.float64)
        self.max_bound = np.asarray(hi, np.float64)

    def get_box_points(self):
        lo, hi = self.min_bound, self.max_bound
        return np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])])

    def get_center(self):
        return (self.min_bound + self.max_bound) * 0.5

    def get_extent(self):
        return self.max_bound - self.min_bound


class OrientedBoundingBox:
    def __init__(self):
        self.center = np.zeros(3)
        self.R = np.eye(3)
        self.extent = np.zeros(3)

    @staticmethod
    def create_from_points(points):
        p = np.asarray(points, np.float64)
        mean = p.mean(0)
        cov = np.cov((p - mean).T, bias=True)
        ev, R = np.linalg.eigh(cov)
        order = np.argsort(ev)[::-1]
        R = R[:, order]
        R[:, 2] = np.cross(R[:, 0], R[:, 1])
        q = (p - mean) @ R
        box = OrientedBoundingBox()
        box.R = R
        box.extent = q.max(0) - q.min(0)
        box.center = R @ ((q.max(0) + q.min(0)) * 0.5) + mean
        return box


class PointCloud:
    def __init__(self, pts=None):
        self._p = np.zeros((0, 3)) if pts is None else np.array(pts, np.float64).reshape(-1, 3)

    @property
    def points(self):
        return self._p

    @points.setter
    def points(self, v):
        self._p = np.array(np.asarray(v), np.float64).reshape(-1, 3)

    def transform(self, M):
        if len(self._p):
            self._p = transform_points(self._p, M)
        return self

    def get_center(self):
        if len(self._p) == 0:
            return np.zeros(3)
        return self._p.mean(axis=0)

    def compute_point_cloud_distance(self, other):
        if len(self._p) == 0 or len(other._p) == 0:
            return np.zeros(0)
        d, _ = cKDTree(other._p).query(self._p, k=1)
        return d

    def get_axis_aligned_bounding_box(self):
        return AxisAlignedBoundingBox(self._p.min(0), self._p.max(0))

    def voxel_down_sample(self, voxel_size):
        p = self._p
        origin = p.min(0) - voxel_size * 0.5
        idx = np.floor((p - origin) / voxel_size).astype(np.int64)
        _, inv, cnt = np.unique(idx, axis=0, return_inverse=True, return_counts=True)
        inv = inv.reshape(-1)
        out = np.zeros((len(cnt), 3))
        for a in range(3):
            out[:, a] = np.bincount(inv, weights=p[:, a], minlength=len(cnt)) / cnt
        return PointCloud(out)

    def select_by_index(self, idx):
        return PointCloud(self._p[np.asarray(idx, np.int64)])

    def segment_plane(self, distance_threshold, ransac_n, num_iterations):
        p = self._p
        best, best_cnt = None, -1
        for _ in range(num_iterations):
            s = p[_RNG.choice(len(p), ransac_n, replace=False)]
            n = np.cross(s[1] - s[0], s[2] - s[0])
            nn = np.linalg.norm(n)
            if nn == 0:
                continue
            n = n / nn
            d = -n.dot(s[0])
            cnt = int((np.abs(p @ n + d) < distance_threshold).sum())
            if cnt > best_cnt:
                best_cnt, best = cnt, np.array([n[0], n[1], n[2], d])
        inl = np.nonzero(np.abs(p @ best[:3] + best[3]) < distance_threshold)[0]
        return best, list(inl)


class _RegistrationResult:
    def __init__(self, T):
        self.transformation = T


def registration_icp(source, target, max_correspondence_distance, init, estimation=None,
                     max_iteration=30, relative_fitness=1e-6, relative_rmse=1e-6):
    src0 = np.asarray(source.points)
    tgt = np.asarray(target.points)
    tree = cKDTree(tgt)
    T = np.array(init, np.float64)
    prev_fit, prev_rmse = None, None
    for _ in range(max_iteration + 1):
        src = transform_points(src0, T)
        d, j = tree.query(src, k=1, distance_upper_bound=max_correspondence_distance)
        ok = np.isfinite(d)
        if ok.sum() < 3:
            break
        fit = ok.mean()
        rmse = np.sqrt((d[ok] ** 2).mean())
        if prev_fit is not None and abs(fit - prev_fit) < relative_fitness and abs(rmse - prev_rmse) < relative_rmse:
            break
        prev_fit, prev_rmse = fit, rmse
        a, b = src[ok], tgt[j[ok]]
        ca, cb = a.mean(0), b.mean(0)
        H = (a - ca).T @ (b - cb)
        U, _, Vt = np.linalg.svd(H)
        D = np.diag([1, 1, np.sign(np.linalg.det(Vt.T @ U.T))])
        R = Vt.T @ D @ U.T
        upd = np.eye(4)
        upd[:3, :3] = R
        upd[:3, 3] = cb - R @ ca
        T = upd @ T
    return _RegistrationResult(T)


def _read_point_cloud(path, remove_nan_points=False, **_):
    xyz, _wh = pcd_io.read_pcd_xyz(path)
    pts = xyz.astype(np.float64)
    if remove_nan_points:
        pts = pts[~np.isnan(pts).any(1)]
    return PointCloud(pts)


def make_module():
    """-> an object usable as `import open3d as o3d` by the reference."""
    o3d = types.ModuleType("open3d")
    o3d.geometry = types.SimpleNamespace(PointCloud=PointCloud, OrientedBoundingBox=OrientedBoundingBox,
                                         AxisAlignedBoundingBox=AxisAlignedBoundingBox)
    o3d.utility = types.SimpleNamespace(Vector3dVector=lambda a: np.array(np.asarray(a), np.float64).reshape(-1, 3))
    o3d.io = types.SimpleNamespace(read_point_cloud=_read_point_cloud)
    reg = types.SimpleNamespace(registration_icp=registration_icp,
                                TransformationEstimationPointToPoint=lambda: None)
    o3d.pipelines = types.SimpleNamespace(registration=reg)
    o3d.__version__ = "0.11.2-definitions"
    return o3d
