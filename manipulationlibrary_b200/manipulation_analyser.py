"""Drop-in surface of the reference's manipulation_analyser.py for the e2SEC extraction path.

Same entry points, arguments and outputs as /root/reference/manipulation_analyser.py:
    analyse_maniac_manipulation(...)   :2184-2291   writes e2sec_<savename>.npy in the CWD, returns None
    _process(...)                      :1355-2182   one frame -> (translation, roll, ESEC_table)
    esec_to_e2sec(esec_array, relations) :62-111    -> (e2sec_array, esec_array), thins the module's final_columns
Everything that touches points runs in libe2sec.so (sm_100a CUDA through the C ABI of include/e2sec.h); the
sequential o1/o2/o3 + relation logic is the library's host resolver.  There is no Open3D, no Cython and no CPU
fallback: without a CUDA device the first call raises.

Scope notes:
  * several GPUs: analyse_maniac_manipulation deals batches of BATCH_FRAMES frames to every visible device in turn, each
    behind its one-frame halo, while reader threads decode the files ahead; one resolver consumes the records in order.
  * frame-0 setup (_getGroundiNewNew :278-361, _getTranslation :212-276) runs on the device (frame0.py, csrc/frame0.cuh),
    seeded and therefore repeatable (the reference draws unseeded random numbers); set_frame0(translation, roll, ground)
    supplies the products instead, clear_frame0() goes back to computing them.
  * icp=True (:1677-1738, rigid-model tracking) runs frame by frame (tracking.py): the models are registered with ICP steps
    on the device.  relations = 1, 2, 3 are all there.
  * debug=True writes the data behind the reference's debug PNGs (:1830-1998) as
    debug_images/<savename>/<frame>.json (AABBs, centres, relation codes) and prunes them to final_columns like
    _clean_debug_images (:160-173); matplotlib rendering is unavailable on this box.
State lives in module globals like the reference's (:32-59): one sequence at a time per process.
"""
import json
import os
import shutil
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import _native as nat
from .engine import Engine, Resolver, esec_to_e2sec_native, rec_entity

IMAGE_SHAPE = (480, 640)            # :1390, :1433
BATCH_FRAMES = 25                   # frames per GPU batch (E2SEC_BATCH_FRAMES overrides)
CONTEXTS_PER_GPU = 3                # engines per GPU that take batches in turn (E2SEC_CONTEXTS overrides)
READ_THREADS = 4                    # file decode runs ahead of the GPUs on this many threads

final_columns = []
column_frames = []                  # frame id of every raw eSEC column (-1: the initial '-' column), all relation modes
total_unique_labels = 0
hand_label_inarray = -1
_frame0 = None                      # (translation, roll, ground) in use
_frame0_given = False               # set by set_frame0: kept for every following sequence until clear_frame0()
_state = None                       # per-sequence: engine, resolver, params


def set_frame0(translation, roll, ground):
    """The frame-0 products the reference computes at :1397-1404 (4x4, 4x4, (G,3) ground cloud already transformed)."""
    global _frame0, _frame0_given
    _frame0_given = True
    _frame0 = (np.asarray(translation, np.float64).reshape(4, 4), np.asarray(roll, np.float64).reshape(4, 4),
               np.ascontiguousarray(ground, np.float64).reshape(-1, 3))


def clear_frame0():
    """Forget the products given with set_frame0: the next sequence computes them from its frame 0 on the device."""
    global _frame0, _frame0_given
    _frame0, _frame0_given = None, False


def frame0_products():
    """(translation, roll, ground) of the current sequence (given, or computed from frame 0)."""
    return _frame0


def _total_unique_labels(label0, hand_label, replace, old, ignored_labels, cut_labels):
    """:1406-1426"""
    tul = np.unique(label0)
    if replace:
        for v in old:
            tul = np.delete(tul, np.where(tul == v))
    for v in ignored_labels:
        if v in tul:
            tul = np.delete(tul, np.where(tul == v))
    if hand_label not in tul:
        if len(cut_labels) > 0:
            tul = np.append(tul, cut_labels)
        tul = np.append(tul, hand_label)
    elif len(cut_labels) > 0:
        tul = np.append(tul, cut_labels)
    return tul


def read_frame(pcd_file, label_file):
    """o3d.io.read_point_cloud (:1441) + pd.read_csv (:1391) through the library's host readers."""
    xyz = np.empty((IMAGE_SHAPE[0] * IMAGE_SHAPE[1], 3), np.float32)
    w, h = nat.C.c_int32(), nat.C.c_int32()
    rc = nat.lib.e2sec_read_pcd_xyz(os.fsencode(pcd_file), nat.ptr(xyz), xyz.shape[0], nat.C.byref(w), nat.C.byref(h))
    if rc != 0:
        raise IOError("cannot read %s (e2sec_read_pcd_xyz -> %d)" % (pcd_file, rc))
    lab = np.empty(IMAGE_SHAPE[0] * IMAGE_SHAPE[1], np.int32)
    rows, cols = nat.C.c_int32(), nat.C.c_int32()
    rc = nat.lib.e2sec_read_labels(os.fsencode(label_file), nat.ptr(lab), lab.shape[0], nat.C.byref(rows), nat.C.byref(cols))
    if rc != 0:
        raise IOError("cannot read %s (e2sec_read_labels -> %d)" % (label_file, rc))
    if w.value * h.value != xyz.shape[0]:
        raise ValueError("%s holds %d x %d points, this path reads organised %d x %d clouds (:1433)"
                         % (pcd_file, w.value, h.value, IMAGE_SHAPE[1], IMAGE_SHAPE[0]))
    if rows.value < 1 or cols.value < 1 or rows.value > IMAGE_SHAPE[0] or cols.value > IMAGE_SHAPE[1]:
        raise ValueError("%s: label image %d x %d does not fit %d x %d" % (label_file, rows.value, cols.value, *IMAGE_SHAPE))
    return xyz, lab[: rows.value * cols.value].reshape(rows.value, cols.value)


def _label_shape(label_file):
    """(rows, cols) of a label file: lines and numbers per line (host look at one file; the batch is parsed on the device)"""
    with open(label_file, "rb") as fh:
        lines = [ln for ln in fh.read().split(b"\n") if ln.strip()]
    return len(lines), len(lines[0].split())


def _check_args(relations, icp, cython):
    if relations not in (1, 2, 3):
        raise ValueError("relations must be 1 (T/N), 2 (+ SSR) or 3 (+ DSR), got %r" % (relations,))
    if not cython:
        raise NotImplementedError("cython=False selects Open3D's statistical outlier branch (:1534-1541), not built")


def _start_sequence(label0, ground_label, hand_label, support_hand, replace, old, new, ignored_labels, cut_labels,
                    thresh, max_frames, filter_mode, filter_seed, relations=3, n_devices=1, xyz0=None, frame0_seed=None,
                    icp=False):
    global _state, total_unique_labels, hand_label_inarray, final_columns, _frame0
    if not _frame0_given:
        # the frame-0 branch of _process (:1397-1404) on the device: RANSAC ground plane, voxel grids, levelling ICP
        # (frame0.py; deterministic, seeded -- the reference draws unseeded random numbers here)
        if xyz0 is None:
            raise ValueError("frame-0 setup needs the cloud of frame 0")
        from . import frame0 as f0
        tmp = Engine(0)
        try:
            _frame0 = f0.frame0_setup(tmp, xyz0, label0, ground_label, f0.DEFAULT_SEED if frame0_seed is None else frame0_seed)
        finally:
            tmp.close()
    tul = _total_unique_labels(label0, hand_label, replace, old, ignored_labels, cut_labels)
    total_unique_labels = tul
    hand_label_inarray = int(np.where(tul == hand_label)[0][0])      # IndexError like :1520 when absent
    int(np.where(tul == ground_label)[0][0])
    T, R, ground = _frame0
    p, keep = nat.make_params(tul, hand_label, ground_label, support_hand, thresh, T, R, ground, replace=replace,
                              old=old, new=new, filter_mode=filter_mode, filter_seed=filter_seed, relations=relations)
    if _state is not None:
        for e in _state["engines"]:
            e.close()
        _state["resolver"].close()
    # every visible GPU takes batches of the sequence in turn (E2SEC_DEVICES limits how many), each through CONTEXTS_PER_GPU
    # engines (E2SEC_CONTEXTS): a compressed cloud is decoded by one warp, a batch's decode has a fixed latency, and only
    # other batches in flight on the same GPU hide it.  _process, frame by frame, and model tracking use one engine.
    n_gpu = max(1, min(int(nat.lib.e2sec_device_count()), int(os.environ.get("E2SEC_DEVICES", "64")), 1 if icp else n_devices))
    per_gpu = 1 if (icp or n_devices == 1) else max(1, int(os.environ.get("E2SEC_CONTEXTS", CONTEXTS_PER_GPU)))
    engines = []
    for i in range(n_gpu * per_gpu):
        eng = Engine(i % n_gpu)
        eng.configure(p, max_frames + 1, IMAGE_SHAPE[0] * IMAGE_SHAPE[1], keep)      # + the halo frame in front of a batch
        engines.append(eng)
    _state = dict(engine=engines[0], engines=engines, resolver=Resolver(p, keep), params=p, keep=keep, K=len(tul), records={},
                  tracker=None)
    if icp:
        # rigid-model tracking (:1677-1738) makes the frames depend on each other through whole clouds: tracking.py walks
        # them one by one over the same device operators
        from .tracking import Tracker
        _state["tracker"] = Tracker(engines[0], tul, hand_label, ground_label, support_hand, thresh, relations, ground)
    final_columns = []


def _table_from_resolver(relations):
    """The raw eSEC table so far.  final_columns follows the reference: frame ids of the columns with relations == 3
    (:1813-1829), left empty with relations 1 / 2 (those branches never append, :1744-1794); column_frames has them always."""
    global final_columns, column_frames
    table, fc = (_state["tracker"] or _state["resolver"]).table()
    column_frames = list(fc)
    final_columns = list(fc) if relations == 3 else []
    if table.shape[1] == 0:
        t = np.chararray((10 * relations, 1), itemsize=5)
        t[:] = "-"
        return t
    return table.view(np.chararray)


def _run_batch(frames, xyz, labels, label_shape):
    eng = _state["engine"]
    kw = {}
    if labels.shape[1] != xyz.shape[1]:
        kw = dict(label_shape=label_shape, image_shape=IMAGE_SHAPE)
    rec = eng.process_batch(xyz, labels, frames, **kw)
    if _state["tracker"] is not None:
        for slot, (f, r) in enumerate(zip(frames, rec)):
            _state["tracker"].frame(int(f), slot, r)
    else:
        _state["resolver"].feed(rec)
    for f, r in zip(frames, rec):
        _state["records"][int(f)] = r


def _process(pcd_file, label_file, ground_label, hand_label, support_hand, translation, roll, frame, ESEC_table,
             relations, replace=False, old=[], new=[], ignored_labels=[], cut_labels=[], thresh=0.1, debug=False,
             cython=True, icp=True, savename="", filter_mode="closure", filter_seed=0, frame0_seed=None):
    """One frame of the sequence (the reference's per-frame entry point, :1355-1357)."""
    _check_args(relations, icp, cython)
    xyz, label = read_frame(pcd_file, label_file)
    if frame == 0 or _state is None:
        _start_sequence(label, ground_label, hand_label, support_hand, replace, old, new, ignored_labels, cut_labels,
                        thresh, 1, filter_mode, filter_seed, relations, xyz0=xyz, frame0_seed=frame0_seed, icp=icp)
    _run_batch(np.array([frame], np.int32), xyz[None], label.reshape(1, -1).astype(np.int32), label.shape)
    T, R, _ = _frame0
    return T, R, _table_from_resolver(relations)


def esec_to_e2sec(esec_array, relations):
    """:62-111 -> (e2sec_array, esec_array upper-cased); thins the module-level final_columns like the reference."""
    global final_columns
    e2, fc = esec_to_e2sec_native(np.asarray(esec_array), relations, final_columns)
    final_columns = fc
    up = np.char.upper(np.asarray(esec_array)).view(np.chararray)
    return e2.view(np.chararray), up


def _debug_artefact(frame, rec, column, K):
    ents = {}
    for e in range(K):
        d = rec_entity(rec, e)
        ents[str(int(total_unique_labels[e]))] = dict(points=d["n_out"], aabb_min=list(d["lo"]), aabb_max=list(d["hi"]),
                                                      mean=list(d["mean"]))
    return dict(frame=int(frame), relations=[c.decode() if isinstance(c, bytes) else str(c) for c in column], entities=ents)


def _clean_debug_images(final_columns, savename):
    """:160-173"""
    path = "debug_images/" + savename + "/"
    for item in sorted(os.listdir(path)):
        if int(item.split(".")[0]) not in final_columns:
            os.remove(path + item)


def analyse_maniac_manipulation(pcl_path, label_path, ground_label, hand_label, support_hand, relations, replace, old,
                                new, ignored_labels, thresh, cut_labels=[], debug=False, cython=True, icp=True,
                                savename="", filter_mode="closure", filter_seed=0, frame0_seed=None):
    """:2184-2291.  Extra keyword arguments (additions): filter_mode 'closure' | 'reference_seeded', filter_seed, and
    frame0_seed for the frame-0 setup (RANSAC samples, synthetic plane), which runs on the device unless set_frame0 gave
    its products."""
    global final_columns
    _check_args(relations, icp, cython)
    if debug:
        for d in ("debug_images/", "event_images/"):
            if os.path.exists(d + savename + "/"):
                shutil.rmtree(d + savename + "/")
            os.makedirs(d + savename + "/")
    fps = 10
    step = int(30 / fps)                                                 # :2266-2268
    files = sorted(os.listdir(pcl_path))
    todo = [(i, f) for i, f in enumerate(files) if i % step == 0]
    batch = max(1, int(os.environ.get("E2SEC_BATCH_FRAMES", BATCH_FRAMES)))
    chunks = [todo[s: s + batch] for s in range(0, len(todo), batch)]

    def paths(f):
        return pcl_path + f[0:-7] + "_pc.pcd", label_path + f[0:-7] + "_left-labels.dat"

    def read_chunk(chunk):
        """The files of a chunk as they are (clouds that are PCD binary_compressed with x, y, z in front are decoded on the
        device, csrc/lzf.cuh); any other flavour goes through the host reader frame by frame."""
        frames = []
        for i, f in chunk:
            pp, lp = paths(f)
            with open(pp, "rb") as fh:
                pb = fh.read()
            head = pb[:1024]
            if b"DATA binary_compressed" in head and b"FIELDS x y z" in head:
                with open(lp, "rb") as fh:
                    frames.append(dict(id=i, pcd=pb, lab=fh.read(), paths=(pp, lp)))
            else:
                xyz, label = read_frame(pp, lp)
                frames.append(dict(id=i, xyz=xyz, labels=label.reshape(-1).astype(np.int32), lshape=label.shape, paths=(pp, lp)))
        if "lshape" not in frames[0]:
            frames[0]["lshape"] = _label_shape(frames[0]["paths"][1])
        return frames

    def as_arrays(fr):
        if "xyz" not in fr:
            xyz, label = read_frame(*fr["paths"])
            fr.update(xyz=xyz, labels=label.reshape(-1).astype(np.int32), lshape=label.shape)
        return fr

    def run_on(dev, frames, halo):
        """One batch as its own little sequence on GPU `dev`: the frame in front of it (halo = the last frame of the batch
        before) goes first and its record is dropped -- the one-frame halo of the frame sharding."""
        eng = _state["engines"][dev]
        eng.reset_sequence()
        seq = ([halo] if halo is not None else []) + frames
        ids = np.array([fr["id"] for fr in seq], np.int32)
        lshape = frames[0].get("lshape") or _label_shape(frames[0]["paths"][1])
        if all("pcd" in fr for fr in seq):
            po = np.concatenate([[0], np.cumsum([len(fr["pcd"]) for fr in seq])]).astype(np.int64)
            lo = np.concatenate([[0], np.cumsum([len(fr["lab"]) for fr in seq])]).astype(np.int64)
            rec = eng.process_batch_files(np.frombuffer(b"".join(fr["pcd"] for fr in seq), np.uint8),
                                          po, np.frombuffer(b"".join(fr["lab"] for fr in seq), np.uint8), lo, ids, lshape,
                                          IMAGE_SHAPE)
        else:
            seq = [as_arrays(fr) for fr in seq]
            xyz, lab = np.stack([fr["xyz"] for fr in seq]), np.stack([fr["labels"] for fr in seq])
            kw = dict(label_shape=lshape, image_shape=IMAGE_SHAPE) if lab.shape[1] != xyz.shape[1] else {}
            rec = eng.process_batch(xyz, lab, ids, **kw)
        return (rec[0], rec[1:]) if halo is not None else (None, rec)

    KEPT = nat.FLAG_HAND_LABEL | nat.FLAG_HAND_POINTS
    last_kept = None                     # the last kept frame fed to the resolver so far
    with ThreadPoolExecutor(max_workers=READ_THREADS) as readers:
        ahead = [readers.submit(read_chunk, c) for c in chunks[: 2]]
        first = ahead[0].result() if ahead else None
        if first is not None:
            rows, cols = first[0]["lshape"]
            fr0 = as_arrays(dict(first[0]))
            label0 = fr0["labels"].reshape(rows, cols)
            _start_sequence(label0, ground_label, hand_label, support_hand, replace, old, new,
                            ignored_labels, cut_labels, thresh, batch, filter_mode, filter_seed, relations, n_devices=64,
                            xyz0=fr0["xyz"], frame0_seed=frame0_seed, icp=icp)
        D = len(_state["engines"]) if first is not None else 1
        workers = [ThreadPoolExecutor(max_workers=1) for _ in range(D)]
        window = 2 * D + 2                                               # chunks read ahead / in flight
        ahead += [readers.submit(read_chunk, c) for c in chunks[2: window]]
        jobs = []                                                        # (frames, halo, future of (halo record, records))
        prev_last = None                                                 # last frame of the chunk in front (speculative halo)
        for c in range(len(chunks)):
            frames = ahead[c].result()
            ahead[c] = None
            if c + window < len(chunks):
                ahead.append(readers.submit(read_chunk, chunks[c + window]))
            jobs.append((frames, prev_last, workers[c % D].submit(run_on, c % D, frames, prev_last)))
            prev_last = frames[-1]
            in_flight = 0 if _state["tracker"] is not None else D        # model tracking reads the batch's clouds back: no overlap
            while jobs and (len(jobs) > in_flight or c == len(chunks) - 1):   # collect in order, D batches stay in flight
                frames0, halo, fut = jobs.pop(0)
                hrec, rec = fut.result()
                want = last_kept["id"] if last_kept is not None else -1
                if halo is not None and (int(hrec[1]) != KEPT or halo["id"] != want):
                    # the frame in front was not a kept frame (no hand in it): redo the batch behind the real one
                    _, rec = workers[0].submit(run_on, 0, frames0, last_kept).result()
                    halo = last_kept
                if _state["tracker"] is not None:
                    # (one device, D == 1: the batch's clouds are still in the engine's buffers; a halo'd batch holds them
                    # one slot further back)
                    for slot, (fr, r) in enumerate(zip(frames0, rec)):
                        _state["tracker"].frame(int(fr["id"]), slot + (1 if halo is not None else 0), r)
                else:
                    _state["resolver"].feed(rec)
                for fr, r in zip(frames0, rec):
                    _state["records"][int(fr["id"])] = r
                kept = np.nonzero(rec[:, 1].astype(np.int64) == KEPT)[0]
                if len(kept):
                    last_kept = frames0[int(kept[-1])]
        for w in workers:
            w.shutdown()
    table = _table_from_resolver(relations)
    raw_columns = list(column_frames)
    if debug:
        for j, f in enumerate(raw_columns):
            if f < 0:
                continue
            with open("debug_images/%s/%d.json" % (savename, f), "w") as fh:
                json.dump(_debug_artefact(f, _state["records"][f], table[:, j], _state["K"]), fh)
    e2sec, esec = esec_to_e2sec(table, relations)
    np.save("e2sec_%s.npy" % savename.replace("/", "_"), e2sec)
    if debug:
        _clean_debug_images(final_columns, savename)
