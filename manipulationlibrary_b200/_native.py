"""ctypes binding of libe2sec.so (include/e2sec.h). There is no fallback: if the library is missing or
cannot be loaded, importing this module raises."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libe2sec.so")

ABI_VERSION = 6            # E2SEC_ABI_VERSION of include/e2sec.h this binding was written against
MAX_ENTITIES = 64
REC_HEADER = 4
REC_ENTITY = 16
FLAG_HAND_LABEL = 1
FLAG_HAND_POINTS = 2
PHASE2_STATE_BYTES = 512
PHASE2_COLUMN_BYTES = 32
IPC_HANDLE_BYTES = 64
STATIC_ICP_MAX_SOURCE = 65536     # E2SEC_STATIC_ICP_MAX_SOURCE: points the no-allocation ICP step has scratch for
FILTER_CLOSURE = 0
FILTER_REFERENCE_SEEDED = 1
FILTER_MODES = {"closure": FILTER_CLOSURE, "reference_seeded": FILTER_REFERENCE_SEEDED}
RADIUS_EXACT = 0
RADIUS_BOUNDED = 1
RADIUS_MODES = {"exact": RADIUS_EXACT, "bounded": RADIUS_BOUNDED}


class Params(C.Structure):
    _fields_ = [("n_entities", C.c_int32), ("entity_label", C.c_int32 * MAX_ENTITIES),
                ("hand_label", C.c_int32), ("ground_label", C.c_int32), ("support_hand", C.c_int32),
                ("n_replace", C.c_int32), ("replace_old", C.c_int32 * 16), ("replace_new", C.c_int32 * 16),
                ("thresh", C.c_double), ("translation", C.c_double * 16), ("roll", C.c_double * 16),
                ("n_ground", C.c_int32), ("ground_xyz", C.POINTER(C.c_double)),
                ("filter_mode", C.c_int32), ("filter_seed", C.c_uint32), ("radius_mode", C.c_int32),
                ("relations", C.c_int32)]


class E2SecError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s not found: build it with `python -m manipulationlibrary_b200.build` "
                          "(this path has no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    sig = {
        "e2sec_abi_version": (C.c_int, []),
        "e2sec_record_doubles": (C.c_size_t, [i32]),
        "e2sec_device_count": (C.c_int, []),
        "e2sec_create": (C.c_int, [C.c_int, C.POINTER(vp)]),
        "e2sec_destroy": (None, [vp]),
        "e2sec_last_error": (C.c_char_p, [vp]),
        "e2sec_stream": (vp, [vp]),
        "e2sec_configure": (C.c_int, [vp, C.POINTER(Params), i32, i32]),
        "e2sec_reset_sequence": (C.c_int, [vp]),
        "e2sec_process_batch": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, vp, C.c_int, vp]),
        "e2sec_process_batch_async": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, vp, C.c_int, vp]),
        "e2sec_process_batch_labels8": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, vp, C.c_int, vp]),
        "e2sec_process_batch_labels8_async": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, vp, C.c_int, vp]),
        "e2sec_process_batch_pcd": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
        "e2sec_process_batch_pcd_async": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
        "e2sec_staged_inputs_dev": (C.c_int, [vp, vp, vp]),
        "e2sec_copy_staged_inputs": (C.c_int, [vp, vp, vp, i32, i32]),
        "e2sec_entity_cloud": (C.c_int, [vp, i32, i32, i32, vp, i32, vp]),
        "e2sec_records_dev": (vp, [vp]),
        "e2sec_copy_records": (C.c_int, [vp, vp, i32, i32]),
        "e2sec_sync": (C.c_int, [vp]),
        "e2sec_launch_count": (i64, [vp]),
        "e2sec_stage_ms": (C.c_int, [vp, vp]),
        "e2sec_region_filter": (C.c_int, [vp, vp, i32, vp, dbl, i32, C.c_uint32, vp, vp]),
        "e2sec_distance": (dbl, [vp, vp]),
        "e2sec_pairwise_median": (C.c_int, [vp, vp, i32, vp]),
        "e2sec_axis_median": (C.c_int, [vp, vp, i32, vp]),
        "e2sec_min_distance": (C.c_int, [vp, vp, i32, vp, i32, dbl, vp]),
        "e2sec_plane_inlier_counts": (C.c_int, [vp, vp, i32, vp, i32, dbl, vp]),
        "e2sec_plane_inlier_mask": (C.c_int, [vp, vp, i32, vp, dbl, vp]),
        "e2sec_voxel_down_sample": (C.c_int, [vp, vp, i32, dbl, vp, vp]),
        "e2sec_icp_correspondences": (C.c_int, [vp, vp, i32, vp, i32, vp, dbl, vp]),
        "e2sec_icp_correspondences_static": (C.c_int, [vp, vp, i32, vp, i32, vp, dbl, vp]),
        "e2sec_dissimilarity": (C.c_int, [vp, vp, vp, vp, i32, i32, dbl, vp, vp]),
        "e2sec_dissimilarity_ms": (C.c_int, [vp, vp]),
        "e2sec_resolver_create": (C.c_int, [C.POINTER(Params), C.POINTER(vp)]),
        "e2sec_resolver_destroy": (None, [vp]),
        "e2sec_resolver_feed": (C.c_int, [vp, vp, i32]),
        "e2sec_resolver_columns": (i32, [vp]),
        "e2sec_resolver_error": (C.c_char_p, [vp]),
        "e2sec_resolver_table": (C.c_int, [vp, vp, vp]),
        "e2sec_resolver_state": (C.c_int, [vp, vp, vp]),
        "e2sec_esec_to_e2sec": (i32, [vp, i32, i32, i32, vp, vp, vp]),
        "e2sec_relation_rows": (C.c_int, [vp, i32, vp, vp]),
        "e2sec_phase2_create": (C.c_int, [vp, i32, C.POINTER(vp)]),
        "e2sec_phase2_destroy": (None, [vp]),
        "e2sec_phase2_reset": (C.c_int, [vp]),
        "e2sec_phase2_run": (C.c_int, [vp, vp, vp, i32]),
        "e2sec_phase2_mailbox_handle": (C.c_int, [vp, vp]),
        "e2sec_phase2_connect": (C.c_int, [vp, vp]),
        "e2sec_phase2_connect_local": (C.c_int, [vp, vp]),
        "e2sec_phase2_run_ring": (C.c_int, [vp, vp, vp, i32, C.c_uint32, C.c_uint32]),
        "e2sec_phase2_state_get": (C.c_int, [vp, vp]),
        "e2sec_phase2_state_set": (C.c_int, [vp, vp]),
        "e2sec_phase2_sync": (C.c_int, [vp, vp]),
        "e2sec_phase2_fetch": (C.c_int, [vp, vp, vp, i32, vp]),
        "e2sec_phase2_export": (C.c_int, [vp, vp, i32]),
        "e2sec_phase2_columns_dev": (C.c_int, [vp, vp, vp, vp]),
        "e2sec_phase2_bindings": (C.c_int, [vp, vp, vp]),
        "e2sec_phase2_stream": (vp, [vp]),
        "e2sec_columns_table": (i32, [vp, vp, i32, i32, vp, vp]),
        "e2sec_read_pcd_xyz": (C.c_int, [C.c_char_p, vp, i64, vp, vp]),
        "e2sec_read_labels": (C.c_int, [C.c_char_p, vp, i64, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)        # AttributeError here = the library does not export what the header declares
        fn.restype = res
        fn.argtypes = args
    return lib, sorted(sig)


lib, EXPORTS = _load()
if lib.e2sec_abi_version() != ABI_VERSION:
    raise ImportError("libe2sec.so has ABI %d, this binding needs %d: rebuild with `python -m manipulationlibrary_b200.build --force`"
                      % (lib.e2sec_abi_version(), ABI_VERSION))


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def make_params(total_unique_labels, hand_label, ground_label, support_hand, thresh, translation, roll, ground,
                replace=False, old=(), new=(), filter_mode="closure", filter_seed=0, radius_mode="bounded", relations=3):
    """-> (Params, keepalive). Labels must be integer valued, as the MANIAC label images are."""
    tul = np.asarray(total_unique_labels, np.float64)
    if len(tul) > MAX_ENTITIES:
        raise ValueError("more than %d entities" % MAX_ENTITIES)
    if not np.all(tul == np.round(tul)):
        raise ValueError("labels must be integer valued")
    p = Params()
    p.n_entities = len(tul)
    for i, v in enumerate(tul):
        p.entity_label[i] = int(v)
    p.hand_label, p.ground_label = int(hand_label), int(ground_label)
    p.support_hand = -1 if support_hand is None else int(support_hand)
    old, new = list(old) if replace else [], list(new) if replace else []
    if len(old) != len(new) or len(old) > 16:
        raise ValueError("old/new must have equal length <= 16")
    p.n_replace = len(old)
    for i, (o, n) in enumerate(zip(old, new)):
        p.replace_old[i], p.replace_new[i] = int(o), int(n)
    p.thresh = float(thresh)
    T = np.ascontiguousarray(translation, np.float64).reshape(16)
    R = np.ascontiguousarray(roll, np.float64).reshape(16)
    for i in range(16):
        p.translation[i], p.roll[i] = T[i], R[i]
    g = np.ascontiguousarray(ground, np.float64).reshape(-1, 3)
    p.n_ground = len(g)
    p.ground_xyz = g.ctypes.data_as(C.POINTER(C.c_double))
    p.filter_mode = FILTER_MODES[filter_mode] if isinstance(filter_mode, str) else int(filter_mode)
    p.filter_seed = int(filter_seed) & 0xFFFFFFFF
    p.radius_mode = RADIUS_MODES[radius_mode] if isinstance(radius_mode, str) else int(radius_mode)
    if int(relations) not in (1, 2, 3):
        raise ValueError("relations must be 1, 2 or 3")
    p.relations = int(relations)
    return p, g


class RelationEntity(C.Structure):
    _fields_ = [("defined", C.c_int32), ("n_points", C.c_int32), ("aabb_min", C.c_double * 3), ("aabb_max", C.c_double * 3),
                ("center", C.c_double * 3)]


class RelationInput(C.Structure):
    _fields_ = [("cur", RelationEntity * 5), ("prev", RelationEntity * 5), ("min_distance", C.c_double * 10),
                ("thresh", C.c_double), ("absent", C.c_int32 * 3)]


def relation_rows(cur, prev, min_distance, thresh, absent, relations=3):
    """_fillTN_absent / _fillSSR_8 / _fillDSR_new_2 (manipulation_analyser.py:363-1353) through e2sec_relation_rows.
    cur / prev: 5 entries [hand, ground, o1, o2, o3], each None or dict(n, lo, hi, center).  -> (codes, absent_out)."""
    inp = RelationInput()
    for dst, src in ((inp.cur, cur), (inp.prev, prev)):
        for s in range(5):
            e = src[s] if src is not None else None
            dst[s].defined = 0 if e is None else 1
            if e is None:
                continue
            dst[s].n_points = int(e["n"])
            for k in range(3):
                dst[s].aabb_min[k] = float(e["lo"][k]) if e["n"] else 0.0
                dst[s].aabb_max[k] = float(e["hi"][k]) if e["n"] else 0.0
                dst[s].center[k] = float(e["center"][k]) if e["n"] else 0.0
    for r in range(10):
        inp.min_distance[r] = float(min_distance[r])
    inp.thresh = float(thresh)
    for k in range(3):
        inp.absent[k] = int(bool(absent[k]))
    codes = np.zeros((10 * relations, 6), np.uint8)
    out = np.zeros(3, np.int32)
    rc = lib.e2sec_relation_rows(C.byref(inp), int(relations), ptr(codes), ptr(out))
    if rc != 0:
        raise E2SecError("e2sec_relation_rows failed (%d)" % rc)
    return [bytes(row).split(b"\0")[0].decode() for row in codes], out.astype(bool).tolist()


def record_doubles(K):
    return int(lib.e2sec_record_doubles(int(K)))
