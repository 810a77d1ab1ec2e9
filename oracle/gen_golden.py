"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/* by running the reference's OWN code.

Runs only in the build container (needs /root/reference).  It imports
/root/reference/manipulation_analyser.py *unmodified* with these stand-ins for what is not
installed here:
    open3d               -> oracle/fake_o3d.py (definitions of the Open3D calls, see its header)
    matplotlib, progressbar -> no-op stubs (debug plotting is out of scope, debug=False)
    cython_filter_new.filter_cython_new -> the reference's own .pyx compiled by oracle/build_ref.py,
                            wrapped so that random.seed(filter_seed(base, frame, i)) runs before each call
                            (mode reference_seeded), or replaced by oracle.esec_oracle.closure_filter
                            (mode closure)
    pd.read_csv          -> same call with delim_whitespace=True spelled sep=r"\\s+" (pandas 3 removed it)
    np.column_stack      -> returns an ndarray subclass whose argsort() is stable (raster-order pin)
and then calls analyse_maniac_manipulation(...) with the notebook's arguments
(example/manipulation_analyser.ipynb cell 3) except icp=False.

Outputs (small, committed):
    tests/golden/pushing06_fixture.npz      translation, roll, ground cloud, total_unique_labels
    tests/golden/pushing06_<mode>.npz       e2sec matrix, raw eSEC table, final_columns, per-column frames,
                                            per-filter-call (frame, i, n_in, n_out, radius, aabb, mean)
    tests/golden/pushing06_frames.npz       sparsified input frames (tracked labels only) so that the GPU
                                            parity test can run where /root/reference does not exist
"""
import importlib.util
import os
import random
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import build_ref, esec_oracle, fake_o3d, pcd_io  # noqa: E402

REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")
ARGS = dict(ground_label=2, hand_label=11, support_hand=None, relations=3, replace=True, old=[7, 9],
            new=[6, 5], ignored_labels=[8, 3], thresh=0.07, cut_labels=[], debug=False, cython=True,
            icp=False)
BASE_SEED = 20261104


class _StableSortArray(np.ndarray):
    def argsort(self, *a, **k):
        k.setdefault("kind", "stable")
        return np.asarray(self).argsort(*a, **k)


class _NumpyProxy:
    def __getattr__(self, name):
        return getattr(np, name)

    @staticmethod
    def column_stack(tup):
        return np.column_stack(tup).view(_StableSortArray)


class _PandasProxy:
    def __init__(self):
        import pandas
        self._pd = pandas

    def __getattr__(self, name):
        return getattr(self._pd, name)

    def read_csv(self, path, delim_whitespace=False, **kw):
        if delim_whitespace:
            kw["sep"] = r"\s+"
        return self._pd.read_csv(path, **kw)


class FilterProxy:
    """Stands where `filter_cython` is in the reference module; logs every call."""

    def __init__(self, mode, real):
        self.mode, self.real, self.calls = mode, real, []

    def region_filter_cython(self, center, pcd, hand_cloud, hand_center, obj_center, median_distance):
        f = sys._getframe(1)
        frame, i = int(f.f_locals["frame"]), int(f.f_locals["i"])
        pcd = np.ascontiguousarray(pcd)
        if self.mode == "closure":
            out = esec_oracle.closure_filter(pcd, np.asarray(hand_center), float(median_distance))
        else:
            random.seed(esec_oracle.filter_seed(BASE_SEED, frame, i))
            out = np.asarray(self.real.region_filter_cython(center, pcd, hand_cloud, hand_center,
                                                            obj_center, median_distance))
        self.calls.append((frame, i, len(pcd), len(out), float(median_distance),
                           *out.min(0), *out.max(0), *out.mean(0)))
        return out


def load_reference(filter_proxy):
    for name in ("open3d", "matplotlib", "matplotlib.pyplot", "progressbar", "cython_filter_new"):
        sys.modules.pop(name, None)
    sys.modules["open3d"] = fake_o3d.make_module()
    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")
    plt.close = lambda *a, **k: None
    mpl.pyplot = plt
    sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    pb = types.ModuleType("progressbar")
    pb.progressbar = lambda it, *a, **k: it
    sys.modules["progressbar"] = pb
    pkg = types.ModuleType("cython_filter_new")
    pkg.filter_cython_new = filter_proxy
    sys.modules["cython_filter_new"] = pkg
    spec = importlib.util.spec_from_file_location("ref_manipulation_analyser",
                                                  os.path.join(REF, "manipulation_analyser.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.np = _NumpyProxy()
    mod.pd = _PandasProxy()
    return mod


def run_reference(mode, pcl_path, label_path, savename, workdir):
    proxy = FilterProxy(mode, build_ref.load())
    ref = load_reference(proxy)
    random.seed(BASE_SEED)          # frame-0 synthetic plane (random.uniform :250-252)
    fake_o3d.seed(BASE_SEED)        # frame-0 RANSAC (:353)
    columns = []
    orig_tn = ref._fillTN_absent

    def spy_tn(hand, ground, thresh, table):
        orig_tn(hand, ground, thresh, table)
        f = sys._getframe(1)
        columns.append((int(f.f_locals["frame"]), int(ref.o1_label) if ref.count1 else -1,
                        int(ref.o2_label) if ref.count2 else -1, int(ref.o3_label) if ref.count3 else -1,
                        len(hand.points), *np.asarray(hand.points).mean(0)))

    ref._fillTN_absent = spy_tn
    raw_cols = []
    orig_dsr = ref._fillDSR_new_2

    def spy_dsr(hand, ground, previous_array, thresh, table, savename):
        orig_dsr(hand, ground, previous_array, thresh, table, savename)
        raw_cols.append([x.decode() for x in table[:, 0]])

    ref._fillDSR_new_2 = spy_dsr
    cwd = os.getcwd()
    os.chdir(workdir)
    try:
        ref.analyse_maniac_manipulation(pcl_path, label_path, savename=savename, **ARGS)
        e2sec = np.load(os.path.join(workdir, "e2sec_%s.npy" % savename))
    finally:
        os.chdir(cwd)
    out = dict(
        e2sec=np.asarray(e2sec).astype("U5"),
        final_columns=np.asarray(ref.final_columns, np.int64),
        filter_calls=np.asarray(proxy.calls, np.float64),
        columns=np.asarray(columns, np.float64),            # one row per frame that reached the fillers
        raw_columns=np.asarray(raw_cols, dtype="U5"),       # the 30 codes of each such frame, pre-dedup
        translation=np.asarray(ref._last_translation if hasattr(ref, "_last_translation") else 0),
    )
    fixture = dict(total_unique_labels=np.asarray(ref.total_unique_labels, np.float64),
                   ground=np.asarray(ref.ground.points, np.float64),
                   hand_label_inarray=np.int64(ref.hand_label_inarray))
    return out, fixture, ref


def main():
    import tempfile
    os.makedirs(GOLD, exist_ok=True)
    pcl = os.path.join(REF, "example", "Pushing_06") + "/"
    lab = os.path.join(REF, "example", "Pushing_06_labels") + "/"
    work = tempfile.mkdtemp(prefix="e2sec_gold_")
    # capture translation/roll: _process returns them; wrap it
    for mode in ("closure", "reference_seeded"):
        tr = {}
        proxy_out, fixture, ref = None, None, None

        def run():
            nonlocal proxy_out, fixture, ref
            proxy = FilterProxy(mode, build_ref.load())
            return proxy

        out, fixture, ref = run_reference_capture(mode, pcl, lab, "Pushing_06", work, tr)
        fixture.update(translation=tr["translation"], roll=tr["roll"])
        np.savez_compressed(os.path.join(GOLD, "pushing06_%s.npz" % mode), **out)
        if mode == "closure":
            np.savez_compressed(os.path.join(GOLD, "pushing06_fixture.npz"), **fixture)
        print(mode, "e2sec", out["e2sec"].shape, "final_columns", out["final_columns"])
        print(out["e2sec"])


def run_reference_capture(mode, pcl, lab, savename, work, tr):
    """run_reference + capture of the (translation, roll) pair _process hands back."""
    proxy = FilterProxy(mode, build_ref.load())
    ref = load_reference(proxy)
    orig_process = ref._process

    def spy_process(*a, **k):
        t, r, table = orig_process(*a, **k)
        tr["translation"], tr["roll"] = np.asarray(t, np.float64), np.asarray(r, np.float64)
        return t, r, table

    ref._process = spy_process
    # re-use run_reference's spies by inlining it on this module instance
    random.seed(BASE_SEED)
    fake_o3d.seed(BASE_SEED)
    columns, raw_cols = [], []
    orig_tn, orig_dsr = ref._fillTN_absent, ref._fillDSR_new_2

    def spy_tn(hand, ground, thresh, table):
        orig_tn(hand, ground, thresh, table)
        f = sys._getframe(1)
        columns.append((int(f.f_locals["frame"]), int(ref.o1_label) if ref.count1 else -1,
                        int(ref.o2_label) if ref.count2 else -1, int(ref.o3_label) if ref.count3 else -1,
                        len(hand.points), *np.asarray(hand.points).mean(0)))

    def spy_dsr(hand, ground, previous_array, thresh, table, savename):
        orig_dsr(hand, ground, previous_array, thresh, table, savename)
        raw_cols.append([x.decode() for x in table[:, 0]])

    ref._fillTN_absent, ref._fillDSR_new_2 = spy_tn, spy_dsr
    cwd = os.getcwd()
    os.chdir(work)
    try:
        ref.analyse_maniac_manipulation(pcl, lab, savename=savename, **ARGS)
        e2sec = np.load(os.path.join(work, "e2sec_%s.npy" % savename))
    finally:
        os.chdir(cwd)
    out = dict(e2sec=np.asarray(e2sec).astype("U5"),
               final_columns=np.asarray(ref.final_columns, np.int64),
               filter_calls=np.asarray(proxy.calls, np.float64),
               columns=np.asarray(columns, np.float64),
               raw_columns=np.asarray(raw_cols, dtype="U5"))
    fixture = dict(total_unique_labels=np.asarray(ref.total_unique_labels, np.float64),
                   ground=np.asarray(ref.ground.points, np.float64),
                   hand_label_inarray=np.int64(ref.hand_label_inarray))
    return out, fixture, ref


def main_icp():
    """icp=True (rigid-model tracking, :1677-1738 -- the notebook's default): the reference's own code on Pushing_06 with
    the same seeds -> tests/golden/pushing06_icp_closure.npz (e2sec, final_columns, per-frame raw columns and bindings)."""
    import tempfile
    ARGS["icp"] = True
    pcl = os.path.join(REF, "example", "Pushing_06") + "/"
    lab = os.path.join(REF, "example", "Pushing_06_labels") + "/"
    work = tempfile.mkdtemp(prefix="e2sec_gold_icp_")
    tr = {}
    out, fixture, ref = run_reference_capture("closure", pcl, lab, "Pushing_06_icp", work, tr)
    np.savez_compressed(os.path.join(GOLD, "pushing06_icp_closure.npz"), e2sec=out["e2sec"], final_columns=out["final_columns"],
                        columns=out["columns"], raw_columns=out["raw_columns"])
    print("icp e2sec", out["e2sec"].shape, "final_columns", out["final_columns"])
    print(out["e2sec"])


if __name__ == "__main__" and "--icp" in sys.argv:
    main_icp()
elif __name__ == "__main__" and "--frames" not in sys.argv:
    main()


def write_frames_fixture(path=None):
    """tests/golden/pushing06_frames.npz -- the 75 processed Pushing_06 frames (file index i % 3 == 0)
    reduced to the points the path can use: valid (non-NaN) points whose upscaled + replaced label is a
    tracked non-ground entity (4, 5, 6, 11).  Ground-label and border points are never read after frame 0
    (manipulation_analyser.py:1514-1546 skips them, the relations use the frame-0 `ground`), so dropping
    them changes no output.  Stored sparse: raster index (delta coded), float32 xyz, uint8 label."""
    path = path or os.path.join(GOLD, "pushing06_frames.npz")
    pcl = os.path.join(REF, "example", "Pushing_06")
    lab = os.path.join(REF, "example", "Pushing_06_labels")
    files = sorted(os.listdir(pcl))
    fid, hand_raw, offs, idx, xs, labs = [], [], [0], [], [], []
    for i, f in enumerate(files):
        if i % 3:
            continue
        stem = f[:-7]
        xyz, _ = pcd_io.read_pcd_xyz(os.path.join(pcl, stem + "_pc.pcd"))
        raw = pcd_io.read_labels(os.path.join(lab, stem + "_left-labels.dat"))
        full = esec_oracle.replace_labels(esec_oracle.resize_nearest(raw, 640, 480),
                                          ARGS["old"], ARGS["new"]).flatten()
        keep = np.isin(full, [4, 5, 6, 11]) & ~np.isnan(xyz).any(1)
        k = np.nonzero(keep)[0]
        fid.append(i)
        hand_raw.append(bool((raw == ARGS["hand_label"]).any()))
        offs.append(offs[-1] + len(k))
        idx.append(np.diff(k, prepend=0).astype(np.uint32))
        xs.append(xyz[k])
        labs.append(full[k].astype(np.uint8))
    xyz = np.concatenate(xs)
    np.savez_compressed(path, frame_ids=np.asarray(fid, np.int32), hand_in_raw=np.asarray(hand_raw),
                        offsets=np.asarray(offs, np.int64), idx_delta=np.concatenate(idx),
                        x=xyz[:, 0].copy(), y=xyz[:, 1].copy(), z=xyz[:, 2].copy(),
                        label=np.concatenate(labs), n_points=np.int64(307200))
    return path


if __name__ == "__main__" and "--frames" in sys.argv:
    print(write_frames_fixture())
