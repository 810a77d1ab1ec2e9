"""Frame-0 setup of a sequence: the ground plane, the levelling transform and the ground cloud that
the reference computes once, at frame 0, with Open3D (manipulation_analyser.py):

    _getGroundiNewNew :278-361   ground-label points -> RANSAC plane (0.01, 3 points, 19 iterations) -> voxel grid 0.02
    _getTranslation   :212-276   voxel grid again -> oriented box extent -> a synthetic flat plane of random points ->
                                 point-to-point ICP (0.2) of the ground onto it; roll = diag(1, -1, -1, 1)
    :1397-1404                   ground = ground.transform(translation).transform(roll)

The geometry that touches clouds runs on the device through the operators of include/e2sec.h (csrc/frame0.cuh): plane
scoring for all RANSAC hypotheses at once, the inlier mask, both voxel grids, and every correspondence step of the ICP.
The host keeps what is a handful of numbers: the three sample points of each hypothesis, the 3 x 3 eigen / singular value
decompositions, the convergence test.

The reference is random here (unseeded RANSAC, random.uniform plane).  This version is deterministic: `seed` drives
numpy's default_rng for the RANSAC samples and Python's random for the plane -- exactly what oracle/gen_golden.py seeded
when the reference's own code produced tests/golden/pushing06_fixture.npz, so that fixture is reproduced from the raw
files of frame 0 (tests/test_frame0_device.py).
"""
import ctypes as C
import random

import numpy as np

from . import _native as nat
from ._native import E2SecError, lib, ptr

ROLL = np.array([[1.0, 0.0, 0.0, 0.0], [0.0, -1.0, 0.0, 0.0], [0.0, 0.0, -1.0, 0.0], [0.0, 0.0, 0.0, 1.0]])   # roll @ roll2 of :262-271
DEFAULT_SEED = 20261104


def _check(engine, rc, what):
    if rc != 0:
        raise E2SecError("%s failed (%d): %s" % (what, rc, lib.e2sec_last_error(engine._h).decode()))


def resize_nearest(label, width, height):
    """cv2.resize(label, (width, height), interpolation=cv2.INTER_NEAREST) (:300): source index floor(dst * src / dst_size)."""
    label = np.asarray(label)
    h, w = label.shape
    cols = np.minimum(np.floor(np.arange(width) * (1.0 / (width / w))).astype(np.int64), w - 1)
    rows = np.minimum(np.floor(np.arange(height) * (1.0 / (height / h))).astype(np.int64), h - 1)
    return label[rows][:, cols]


def transform_points(pts, M):
    """PointCloud.transform with the operation order the device uses (csrc/seg.cuh: transform_point)."""
    x, y, z = pts[:, 0], pts[:, 1], pts[:, 2]
    q = [((M[r, 0] * x + M[r, 1] * y) + M[r, 2] * z) + M[r, 3] for r in range(4)]
    return np.stack([q[0] / q[3], q[1] / q[3], q[2] / q[3]], axis=1)


def ground_label_points(xyz32, label_image, ground_label, image_shape=(480, 640)):
    """The cloud _getGroundiNewNew hands to segment_plane (:290-350): valid points sorted by (raw, upscaled) label -- raster
    order inside a label --, then the slice [last index of the next lower label present, last index of the ground label):
    the last point of the label below the ground's, followed by all ground points but the last one."""
    lab = resize_nearest(np.asarray(label_image), image_shape[1], image_shape[0]).reshape(-1)
    cloud = np.nan_to_num(np.asarray(xyz32, np.float64), nan=-100.0)
    valid = ~(cloud == -100.0).any(axis=1)
    uniq = np.unique(np.asarray(label_image))
    gi = int(np.where(uniq == ground_label)[0][0])                  # IndexError like the reference when the label is absent
    lower = None
    i = 1
    while lower is None:
        cand = uniq[gi - i]                                         # wraps around like the reference's negative index
        if (valid & (lab == cand)).any():
            lower = cand
        else:
            i += 1
            if i > len(uniq) + 1:
                raise ValueError("no label with valid points below the ground label")
    g = cloud[valid & (lab == ground_label)]
    low = cloud[valid & (lab == lower)]
    if float(lower) > float(ground_label) or len(g) == 0:
        # the reference's slice is empty (or runs backwards) in these cases: there is no ground cloud to fit
        raise ValueError("the ground label has no valid points in frame 0")
    return np.ascontiguousarray(np.concatenate([low[-1:], g[:-1]]))


def segment_plane(engine, pts, threshold=0.01, ransac_n=3, iterations=19, rng=None):
    """pcd.segment_plane (:353): hypotheses from `ransac_n` random points each, scored on the device all at once."""
    rng = rng if rng is not None else np.random.default_rng(DEFAULT_SEED)
    planes = []
    for _ in range(iterations):
        s = pts[rng.choice(len(pts), ransac_n, replace=False)]
        n = np.cross(s[1] - s[0], s[2] - s[0])
        nn = np.linalg.norm(n)
        if nn == 0:
            continue
        n = n / nn
        planes.append([n[0], n[1], n[2], -n.dot(s[0])])
    if not planes:
        raise ValueError("every RANSAC sample was degenerate")
    planes = np.ascontiguousarray(planes, np.float64)
    counts = np.zeros(len(planes), np.int32)
    _check(engine, lib.e2sec_plane_inlier_counts(engine._h, ptr(pts), len(pts), ptr(planes), len(planes), float(threshold),
                                                 ptr(counts)), "e2sec_plane_inlier_counts")
    best = planes[int(np.argmax(counts))]                           # the first maximum, like `cnt > best_cnt`
    mask = np.zeros(len(pts), np.uint8)
    _check(engine, lib.e2sec_plane_inlier_mask(engine._h, ptr(pts), len(pts), ptr(best), float(threshold), ptr(mask)),
           "e2sec_plane_inlier_mask")
    return best, np.nonzero(mask)[0]


def voxel_down_sample(engine, pts, voxel):
    pts = np.ascontiguousarray(pts, np.float64)
    out = np.empty_like(pts)
    n = C.c_int32()
    _check(engine, lib.e2sec_voxel_down_sample(engine._h, ptr(pts), len(pts), float(voxel), ptr(out), C.byref(n)),
           "e2sec_voxel_down_sample")
    return out[: n.value].copy()


def registration_icp(engine, source, target, max_distance, init=None, max_iteration=30, relative_fitness=1e-6,
                     relative_rmse=1e-6):
    """o3d.pipelines.registration.registration_icp, point to point (:258-260): correspondences and their sums on the
    device, the Kabsch update (3 x 3 SVD) here.  -> (4 x 4 transformation, fitness, inlier rmse)."""
    source = np.ascontiguousarray(source, np.float64)
    target = np.ascontiguousarray(target, np.float64)
    T = np.eye(4) if init is None else np.array(init, np.float64)
    prev = None
    fit = rmse = 0.0
    for _ in range(max_iteration + 1):
        st = np.zeros(17)
        Tc = np.ascontiguousarray(T)
        _check(engine, lib.e2sec_icp_correspondences(engine._h, ptr(source), len(source), ptr(target), len(target), ptr(Tc),
                                                     float(max_distance), ptr(st)), "e2sec_icp_correspondences")
        n_ok = int(st[0])
        if n_ok < 3:
            break
        fit, rmse = n_ok / len(source), float(np.sqrt(st[1] / n_ok))
        if prev is not None and abs(fit - prev[0]) < relative_fitness and abs(rmse - prev[1]) < relative_rmse:
            break
        prev = (fit, rmse)
        ca, cb, H = st[2:5], st[5:8], st[8:17].reshape(3, 3)
        U, _, Vt = np.linalg.svd(H)
        D = np.diag([1.0, 1.0, np.sign(np.linalg.det(Vt.T @ U.T))])
        R = Vt.T @ D @ U.T
        upd = np.eye(4)
        upd[:3, :3] = R
        upd[:3, 3] = cb - R @ ca
        T = upd @ T
    return T, fit, rmse


def oriented_extent(points):
    """OrientedBoundingBox.create_from_points(...).extent (:227-228): PCA box, axes by decreasing variance."""
    p = np.asarray(points, np.float64)
    mean = p.mean(0)
    ev, R = np.linalg.eigh(np.cov((p - mean).T, bias=True))
    R = R[:, np.argsort(ev)[::-1]]
    R[:, 2] = np.cross(R[:, 0], R[:, 1])
    q = (p - mean) @ R
    return q.max(0) - q.min(0)


def get_ground(engine, xyz32, label_image, ground_label, rng=None):
    """_getGroundiNewNew (:278-361) -> the voxelised ground plane of frame 0, camera frame."""
    pts = ground_label_points(xyz32, label_image, ground_label)
    _, inliers = segment_plane(engine, pts, 0.01, 3, 19, rng)
    return voxel_down_sample(engine, pts[inliers], 0.02)


def get_translation(engine, ground, py_random=None):
    """_getTranslation (:212-276) -> (translation, roll)."""
    rnd = py_random if py_random is not None else random.Random(DEFAULT_SEED)
    vox = voxel_down_sample(engine, ground, 0.02)
    ext = oriented_extent(vox)
    middle = vox.mean(0)
    pcd_z = float(np.mean(vox[:, 2]))
    xs = [middle[0] + ext[0] / 2, middle[0] - ext[0] / 2]
    ys = [middle[1] + ext[1] / 2, middle[1] - ext[1] / 2]
    min_x, max_x, min_y, max_y = min(xs), max(xs), min(ys), max(ys)
    plain = np.empty_like(vox)
    for i in range(len(vox)):                                       # :249-253, the draw order of the reference
        plain[i] = [rnd.uniform(min_x, max_x), rnd.uniform(min_y, max_y), rnd.uniform(pcd_z - 0.001, pcd_z + 0.001)]
    T, _, _ = registration_icp(engine, vox, plain, 0.2, np.eye(4))
    return T, ROLL.copy()


def frame0_setup(engine, xyz32, label_image, ground_label, seed=DEFAULT_SEED):
    """The frame-0 branch of _process (:1397-1404): -> (translation, roll, ground cloud in the levelled frame)."""
    ground = get_ground(engine, xyz32, label_image, ground_label, np.random.default_rng(seed))
    translation, roll = get_translation(engine, ground, random.Random(seed))
    return translation, roll, transform_points(transform_points(ground, translation), roll)
