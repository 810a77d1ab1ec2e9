"""Thin Python object model over the C ABI: Engine (device phase 1 + operators), Resolver (host phase 2),
and record accessors. Mirrors nothing in the reference by itself -- manipulation_analyser.py in this
package is the drop-in surface; this is what it calls."""
import ctypes as C

import numpy as np

from . import _native as nat
from ._native import E2SecError, lib, ptr


class Engine:
    """One CUDA device + stream. Raises E2SecError when no device is present (no CPU fallback)."""

    def __init__(self, device=0):
        h = C.c_void_p()
        rc = lib.e2sec_create(int(device), C.byref(h))
        if rc != 0:
            raise E2SecError("e2sec_create failed (%d): %s" % (rc, lib.e2sec_last_error(None).decode()))
        self._h = h
        self.K = None
        self.n_points = None
        self.max_frames = None
        self._keep = None

    def close(self):
        if self._h:
            lib.e2sec_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise E2SecError("%s failed (%d): %s" % (what, rc, lib.e2sec_last_error(self._h).decode()))

    # ---- sequence -------------------------------------------------------------------------------
    def configure(self, params, max_frames, n_points, keepalive=None):
        self._check(lib.e2sec_configure(self._h, C.byref(params), int(max_frames), int(n_points)), "e2sec_configure")
        self.K, self.n_points, self.max_frames = int(params.n_entities), int(n_points), int(max_frames)
        self._keep = (params, keepalive)

    def reset_sequence(self):
        self._check(lib.e2sec_reset_sequence(self._h), "e2sec_reset_sequence")

    def process_batch(self, xyz, labels, frame_ids, label_shape=None, image_shape=None):
        """xyz (F, N, 3) float32, labels (F, N) or (F, h, w) int32 (host numpy). -> records (F, rec) float64."""
        xyz = np.ascontiguousarray(xyz, np.float32)
        labels = np.asarray(labels)
        fid = np.ascontiguousarray(frame_ids, np.int32).reshape(-1)
        F = self._check_batch(xyz.shape, labels.shape, len(fid))
        rec = np.empty((F, nat.record_doubles(self.K)), np.float64)
        if labels.dtype == np.uint8:                   # uint8 label images travel as they are (13 B/point over PCIe)
            labels = np.ascontiguousarray(labels).reshape(F, -1)
            lw, lh, w, h = self._shapes(labels.shape[1], label_shape, image_shape)
            self._check(lib.e2sec_process_batch_labels8(self._h, ptr(xyz), ptr(labels), F, lw, lh, w, h, ptr(fid), 0,
                                                        ptr(rec)), "e2sec_process_batch_labels8")
            return rec
        labels = np.ascontiguousarray(labels, np.int32).reshape(F, -1)
        lw, lh, w, h = self._shapes(labels.shape[1], label_shape, image_shape)
        self._check(lib.e2sec_process_batch(self._h, ptr(xyz), ptr(labels), F, lw, lh, w, h, ptr(fid), 0, ptr(rec)),
                    "e2sec_process_batch")
        return rec

    def process_batch_files(self, pcd_blob, pcd_offsets, label_blob, label_offsets, frame_ids, label_shape, image_shape,
                            records=None, wait=True):
        """The FILES as input: pcd_blob / label_blob are uint8 arrays holding the PCD (binary_compressed) and label files back
        to back, *_offsets (F + 1) their boundaries.  Clouds are LZF-decoded and the label text is parsed on the device.
        With wait=False the buffers (pinned) must stay untouched until sync() and `records` must be given (pinned)."""
        pcd_blob = np.ascontiguousarray(pcd_blob, np.uint8).reshape(-1)
        label_blob = np.ascontiguousarray(label_blob, np.uint8).reshape(-1)
        po = np.ascontiguousarray(pcd_offsets, np.int64).reshape(-1)
        lo = np.ascontiguousarray(label_offsets, np.int64).reshape(-1)
        fid = np.ascontiguousarray(frame_ids, np.int32).reshape(-1)
        F = len(fid)
        if len(po) != F + 1 or len(lo) != F + 1 or po[-1] > len(pcd_blob) or lo[-1] > len(label_blob) or po[0] < 0 or lo[0] < 0:
            raise ValueError("offsets must hold frames + 1 boundaries inside the blobs")
        if F > self.max_frames:
            raise ValueError("%d frames exceed max_frames=%d of configure()" % (F, self.max_frames))
        lh, lw, h, w = int(label_shape[0]), int(label_shape[1]), int(image_shape[0]), int(image_shape[1])
        if h * w != self.n_points:
            raise ValueError("image_shape %r does not match the %d points per frame" % ((h, w), self.n_points))
        rec = np.empty((F, nat.record_doubles(self.K)), np.float64) if records is None else records
        fn = lib.e2sec_process_batch_pcd if wait else lib.e2sec_process_batch_pcd_async
        self._check(fn(self._h, ptr(pcd_blob), ptr(po), ptr(label_blob), ptr(lo), F, lw, lh, w, h, ptr(fid), ptr(rec)),
                    "e2sec_process_batch_pcd")
        self._keep_files = (pcd_blob, label_blob, po, lo, fid)
        return rec

    def staged_inputs(self, n_frames, label_size):
        """The decoded inputs of the batch queued last, copied back: xyz (F, N, 3) float32, labels (F, label_size) int32."""
        xyz = np.empty((n_frames, self.n_points, 3), np.float32)
        lab = np.empty((n_frames, label_size), np.int32)
        self._check(lib.e2sec_copy_staged_inputs(self._h, ptr(xyz), ptr(lab), int(n_frames), int(label_size)), "e2sec_copy_staged_inputs")
        return xyz, lab

    def process_batch_device(self, xyz_ptr, labels_ptr, n_frames, frame_ids, label_size, records=None,
                             label_shape=None, image_shape=None):
        """Device-resident inputs (raw pointers, e.g. torch tensors' data_ptr())."""
        fid = np.ascontiguousarray(frame_ids, np.int32)
        lw, lh, w, h = self._shapes(label_size, label_shape, image_shape)
        self._check(lib.e2sec_process_batch(self._h, C.c_void_p(xyz_ptr), C.c_void_p(labels_ptr), int(n_frames), lw, lh,
                                            w, h, ptr(fid), 1, ptr(records) if records is not None else None),
                    "e2sec_process_batch")
        return records

    def process_batch_device_async(self, xyz_ptr, labels_ptr, n_frames, frame_ids, label_size, records=None,
                                   label_shape=None, image_shape=None, on_device=True, labels8=False):
        """Queues a batch without waiting (e2sec_process_batch_async): `records` must be pinned host memory that stays
        alive, inputs must stay untouched until sync(); device-side errors surface at the next sync()."""
        fid = np.ascontiguousarray(frame_ids, np.int32)
        lw, lh, w, h = self._shapes(label_size, label_shape, image_shape)
        fn = lib.e2sec_process_batch_labels8_async if labels8 else lib.e2sec_process_batch_async
        self._check(fn(self._h, C.c_void_p(xyz_ptr), C.c_void_p(labels_ptr), int(n_frames), lw, lh, w, h, ptr(fid),
                       1 if on_device else 0, ptr(records) if records is not None else None), "e2sec_process_batch_async")
        return records

    def _check_batch(self, xyz_shape, labels_shape, n_ids):
        """The C side copies n_frames * n_points * 3 floats from a raw pointer: every shape is checked here (the reference
        fails with a shape error at np.column_stack, manipulation_analyser.py:1448, when cloud and labels disagree)."""
        if self.n_points is None:
            raise E2SecError("configure() has not been called")
        if len(xyz_shape) != 3 or xyz_shape[2] != 3 or xyz_shape[1] != self.n_points:
            raise ValueError("xyz must be (frames, %d, 3), got %r" % (self.n_points, tuple(xyz_shape)))
        F = int(xyz_shape[0])
        if F > self.max_frames:
            raise ValueError("%d frames exceed max_frames=%d of configure()" % (F, self.max_frames))
        if n_ids != F:
            raise ValueError("%d frame ids for %d frames" % (n_ids, F))
        if len(labels_shape) < 2 or labels_shape[0] != F:
            raise ValueError("labels must be (frames, ...) with %d frames, got %r" % (F, tuple(labels_shape)))
        return F

    def _shapes(self, label_size, label_shape, image_shape):
        if label_size == self.n_points and label_shape is None:
            return label_size, 1, self.n_points, 1
        if label_shape is None or image_shape is None:
            raise ValueError("label image smaller than the cloud: pass label_shape=(rows, cols) and image_shape=(rows, cols)")
        lh, lw, h, w = int(label_shape[0]), int(label_shape[1]), int(image_shape[0]), int(image_shape[1])
        if lh * lw != label_size:
            raise ValueError("label_shape %r does not match the %d labels per frame" % ((lh, lw), label_size))
        if h * w != self.n_points:
            raise ValueError("image_shape %r does not match the %d points per frame" % ((h, w), self.n_points))
        if label_size == self.n_points:
            return label_size, 1, self.n_points, 1
        return lw, lh, w, h

    def entity_cloud(self, frame, entity, filtered):
        """(n, 3) float64: entity's cloud of frame `frame` of the batch processed last -- pcd[i] or filtered_pcd_voxel[i]."""
        n = C.c_int32()
        self._check(lib.e2sec_entity_cloud(self._h, int(frame), int(entity), 1 if filtered else 0, None, 0, C.byref(n)), "e2sec_entity_cloud")
        out = np.empty((n.value, 3), np.float64)
        if n.value:
            self._check(lib.e2sec_entity_cloud(self._h, int(frame), int(entity), 1 if filtered else 0, ptr(out), n.value, C.byref(n)),
                        "e2sec_entity_cloud")
        return out

    def copy_records(self, dst_ptr, first, count):
        """Stream-ordered copy of records [first, first + count) of the batch queued last to host or device memory."""
        self._check(lib.e2sec_copy_records(self._h, C.c_void_p(int(dst_ptr)), int(first), int(count)), "e2sec_copy_records")

    @property
    def records_dev(self):
        return lib.e2sec_records_dev(self._h)

    def sync(self):
        self._check(lib.e2sec_sync(self._h), "e2sec_sync")

    @property
    def stream(self):
        return lib.e2sec_stream(self._h)

    @property
    def launch_count(self):
        return int(lib.e2sec_launch_count(self._h))

    def stage_ms(self):
        out = np.zeros(8, np.float32)
        self._check(lib.e2sec_stage_ms(self._h, ptr(out)), "e2sec_stage_ms")
        st = dict(zip(["segment", "medians", "radius", "filter", "stats", "pairs", "assemble"], out[:7].tolist()))
        self.seg_stream_ms = float(out[7])          # the streaming kernel alone (inside "segment")
        return st

    # ---- operators ------------------------------------------------------------------------------
    def region_filter(self, pcd, center, median_distance, mode="closure", seed=0):
        pcd = np.ascontiguousarray(pcd, np.float64)
        if pcd.ndim != 2 or pcd.shape[1] != 3:
            raise ValueError("pcd must be (n, 3)")
        center = np.ascontiguousarray(center, np.float64)
        out = np.empty((len(pcd) + 1, 3), np.float64)
        n_out = C.c_int32()
        self._check(lib.e2sec_region_filter(self._h, ptr(pcd), len(pcd), ptr(center), float(median_distance),
                                            nat.FILTER_MODES[mode], int(seed) & 0xFFFFFFFF, ptr(out), C.byref(n_out)),
                    "e2sec_region_filter")
        return out[:n_out.value].copy()

    def pairwise_median(self, pts):
        pts = np.ascontiguousarray(pts, np.float64)
        out = C.c_double()
        self._check(lib.e2sec_pairwise_median(self._h, ptr(pts), len(pts), C.byref(out)), "e2sec_pairwise_median")
        return out.value

    def dissimilarity(self, tables, divisor=10.0, want_counts=False):
        """tables: sequence of 2-D arrays of relation codes (rows x n_cols, the same row count, a multiple of 3).
        -> D (M, M) float64 [, counts (M, M, 3) uint32]; eSEC_analyser._calc_D_shaped (:819-841) on the device."""
        tabs = [np.asarray(t) for t in tables]
        if not tabs:
            raise ValueError("no tables")
        rows = tabs[0].shape[0]
        enc = []
        for t in tabs:
            if t.ndim != 2 or t.shape[0] != rows:
                raise ValueError("every table needs the same number of rows")
            if t.dtype.kind == "O":                    # pandas hands out object arrays
                t = t.astype(str)
            if t.dtype.kind not in "US":
                raise ValueError("tables hold strings")
            # code units of every entry as integers (no per-element Python work): code points for str, bytes for bytes
            unit = np.uint32 if t.dtype.kind == "U" else np.uint8
            width = t.dtype.itemsize // np.dtype(unit).itemsize
            cp = np.ascontiguousarray(t).view(unit).reshape(t.shape + (width,))
            if cp.size and (int(cp.max()) > 127 or (width > 5 and cp[..., 5:].any())):
                raise ValueError("relation codes are at most 5 ASCII characters")
            b = np.zeros(t.shape + (6,), np.uint8)
            b[..., :min(width, 5)] = cp[..., :5]
            enc.append(b)
        n_cols = np.array([t.shape[1] for t in enc], np.int32)
        sizes = np.array([t.size for t in enc], np.int64)
        offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)
        blob = np.concatenate([t.reshape(-1) for t in enc]) if sizes.sum() else np.zeros(1, np.uint8)
        M = len(enc)
        D = np.empty((M, M), np.float64)
        counts = np.empty((M, M, 3), np.uint32) if want_counts else None
        self._check(lib.e2sec_dissimilarity(self._h, ptr(blob), ptr(offsets), ptr(n_cols), M, int(rows), float(divisor),
                                            ptr(D), ptr(counts) if want_counts else None), "e2sec_dissimilarity")
        return (D, counts) if want_counts else D

    def dissimilarity_ms(self):
        ms = C.c_float()
        self._check(lib.e2sec_dissimilarity_ms(self._h, C.byref(ms)), "e2sec_dissimilarity_ms")
        return ms.value

    def axis_median(self, pts):
        pts = np.ascontiguousarray(pts, np.float64)
        out = np.empty(3, np.float64)
        self._check(lib.e2sec_axis_median(self._h, ptr(pts), len(pts), ptr(out)), "e2sec_axis_median")
        return out

    def min_distance(self, a, b, limit=0.0):
        a = np.ascontiguousarray(a, np.float64)
        b = np.ascontiguousarray(b, np.float64)
        out = C.c_double()
        self._check(lib.e2sec_min_distance(self._h, ptr(a), len(a), ptr(b), len(b), float(limit), C.byref(out)),
                    "e2sec_min_distance")
        return out.value


class Resolver:
    """Host phase 2 (sequential): records in frame order -> eSEC table."""

    def __init__(self, params, keepalive=None):
        h = C.c_void_p()
        rc = lib.e2sec_resolver_create(C.byref(params), C.byref(h))
        if rc != 0:
            raise E2SecError("e2sec_resolver_create failed (%d)" % rc)
        self._h = h
        self.K = int(params.n_entities)
        self.rows = 10 * (int(params.relations) or 3)

    def close(self):
        if self._h:
            lib.e2sec_resolver_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def feed(self, records):
        records = np.ascontiguousarray(records, np.float64).reshape(-1, nat.record_doubles(self.K))
        rc = lib.e2sec_resolver_feed(self._h, ptr(records), len(records))
        if rc != 0:
            raise E2SecError("e2sec_resolver_feed failed (%d): %s" % (rc, lib.e2sec_resolver_error(self._h).decode()))

    @property
    def n_columns(self):
        return int(lib.e2sec_resolver_columns(self._h))

    def table(self):
        """-> (codes (10 * relations, n) 'S5' array as the reference's chararray, frame id of every column)."""
        n, rows = self.n_columns, self.rows
        codes = np.zeros((rows, max(n, 1), 6), np.uint8)
        fc = np.zeros(max(n, 1), np.int32)
        if n:
            lib.e2sec_resolver_table(self._h, ptr(codes), ptr(fc))
        arr = codes[:, :n].reshape(rows, n, 6).view("S6").reshape(rows, n).astype("S5")
        return arr, fc[:n].tolist()

    def state(self):
        lab = np.zeros(3, np.int32)
        ab = np.zeros(3, np.int32)
        lib.e2sec_resolver_state(self._h, ptr(lab), ptr(ab))
        return lab.tolist(), ab.astype(bool).tolist()


def columns_table(columns, frames, relations):
    """Packed phase-2 columns (n, 32) uint8 + frame ids -> (codes (10 * relations, total) 'S5', frame id per column) in the
    layout Resolver.table() gives (e2sec_columns_table: the initial '-' column of relations 1 / 2 included)."""
    columns = np.ascontiguousarray(columns, np.uint8).reshape(-1, nat.PHASE2_COLUMN_BYTES)
    frames = np.ascontiguousarray(frames, np.int32).reshape(-1)
    n, rows = len(frames), 10 * int(relations)
    total = int(lib.e2sec_columns_table(ptr(columns), ptr(frames), n, int(relations), None, None))
    if total < 0:
        raise E2SecError("e2sec_columns_table failed (%d)" % total)
    codes = np.zeros((rows, max(total, 1), 6), np.uint8)
    fc = np.zeros(max(total, 1), np.int32)
    if total:
        lib.e2sec_columns_table(ptr(columns), ptr(frames), n, int(relations), ptr(codes), ptr(fc))
    arr = codes[:, :total].reshape(rows, total, 6).view("S6").reshape(rows, total).astype("S5")
    return arr, fc[:total].tolist()


class Phase2:
    """Device phase 2 (csrc/phase2.cuh): selection, relation rows and column dedup as one kernel over records that stay in
    device memory.  Belongs to a configured Engine; same outputs as Resolver."""

    def __init__(self, engine, max_columns=65536):
        h = C.c_void_p()
        rc = lib.e2sec_phase2_create(engine._h, int(max_columns), C.byref(h))
        if rc != 0:
            raise E2SecError("e2sec_phase2_create failed (%d): %s" % (rc, lib.e2sec_last_error(engine._h).decode()))
        self._h, self._eng = h, engine
        self.K = engine.K
        self.relations = int(engine._keep[0].relations) or 3
        self.max_columns = int(max_columns)

    def close(self):
        if self._h:
            lib.e2sec_phase2_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            if self._eng._h:
                self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise E2SecError("%s failed (%d): %s" % (what, rc, lib.e2sec_last_error(self._eng._h).decode()))

    def reset(self):
        self._check(lib.e2sec_phase2_reset(self._h), "e2sec_phase2_reset")

    def run(self, records_dev_ptr, n_frames, src=None):
        """records_dev_ptr: device address of n_frames records; src: the Engine whose stream produces them (or None)."""
        self._check(lib.e2sec_phase2_run(self._h, src._h if src is not None else None, C.c_void_p(int(records_dev_ptr)),
                                         int(n_frames)), "e2sec_phase2_run")

    def run_engine_batch(self, engine, n_frames, skip=0):
        """Phase 2 of the batch `engine` has just queued (its device record buffer), dropping the first `skip` records."""
        base = lib.e2sec_records_dev(engine._h)
        self.run(base + skip * nat.record_doubles(self.K) * 8, n_frames - skip, src=engine)

    def mailbox_handle(self):
        """CUDA IPC handle (64 bytes) of this object's mailbox, for the rank that runs the frames in front."""
        h = np.zeros(nat.IPC_HANDLE_BYTES, np.uint8)
        self._check(lib.e2sec_phase2_mailbox_handle(self._h, ptr(h)), "e2sec_phase2_mailbox_handle")
        return h

    def connect(self, next_handle):
        """Opens the mailbox of the rank that runs the NEXT frames (peer memory over NVLink)."""
        h = np.ascontiguousarray(next_handle, np.uint8)
        self._check(lib.e2sec_phase2_connect(self._h, ptr(h)), "e2sec_phase2_connect")

    def connect_local(self, nxt):
        """The same for a Phase2 object of this very process."""
        self._check(lib.e2sec_phase2_connect_local(self._h, nxt._h), "e2sec_phase2_connect_local")

    def run_ring(self, records_dev_ptr, n_frames, wait_ticket, post_ticket, src=None):
        self._check(lib.e2sec_phase2_run_ring(self._h, src._h if src is not None else None, C.c_void_p(int(records_dev_ptr)),
                                              int(n_frames), int(wait_ticket), int(post_ticket)), "e2sec_phase2_run_ring")

    def run_ring_engine_batch(self, engine, n_frames, skip, wait_ticket, post_ticket):
        """run_engine_batch whose kernel takes the state from its mailbox (wait_ticket) and posts it on (post_ticket)."""
        base = lib.e2sec_records_dev(engine._h)
        self._check(lib.e2sec_phase2_run_ring(self._h, engine._h, C.c_void_p(base + skip * nat.record_doubles(self.K) * 8),
                                              int(n_frames - skip), int(wait_ticket), int(post_ticket)), "e2sec_phase2_run_ring")

    def state_get(self, dst_ptr=None):
        if dst_ptr is None:
            buf = np.zeros(nat.PHASE2_STATE_BYTES, np.uint8)
            self._check(lib.e2sec_phase2_state_get(self._h, ptr(buf)), "e2sec_phase2_state_get")
            return buf
        self._check(lib.e2sec_phase2_state_get(self._h, C.c_void_p(int(dst_ptr))), "e2sec_phase2_state_get")

    def state_set(self, src):
        if isinstance(src, np.ndarray):
            src = np.ascontiguousarray(src, np.uint8)
            self._check(lib.e2sec_phase2_state_set(self._h, ptr(src)), "e2sec_phase2_state_set")
        else:
            self._check(lib.e2sec_phase2_state_set(self._h, C.c_void_p(int(src))), "e2sec_phase2_state_set")

    def sync(self):
        n = C.c_int32()
        self._check(lib.e2sec_phase2_sync(self._h, C.byref(n)), "e2sec_phase2_sync")
        return n.value

    def fetch(self):
        """-> (packed columns (n, 32) uint8, frame ids (n,) int32) on the host."""
        n = self.sync()
        cols = np.zeros((max(n, 1), nat.PHASE2_COLUMN_BYTES), np.uint8)
        fr = np.zeros(max(n, 1), np.int32)
        got = C.c_int32()
        self._check(lib.e2sec_phase2_fetch(self._h, ptr(cols), ptr(fr), max(n, 1), C.byref(got)), "e2sec_phase2_fetch")
        return cols[:got.value], fr[:got.value]

    def export(self, dst_ptr, capacity):
        """Columns, frame ids and their count into one fixed-size buffer (16 + 36 * capacity bytes, host or device)."""
        self._check(lib.e2sec_phase2_export(self._h, C.c_void_p(int(dst_ptr)), int(capacity)), "e2sec_phase2_export")

    @staticmethod
    def parse_export(buf, capacity):
        """The buffer export() filled (uint8 array) -> (columns (n, 32) uint8, frames (n,) int32)."""
        buf = np.ascontiguousarray(buf, np.uint8).reshape(-1)
        n = int(buf[:4].view(np.int32)[0])
        cols = buf[16: 16 + capacity * nat.PHASE2_COLUMN_BYTES].reshape(capacity, nat.PHASE2_COLUMN_BYTES)[:n]
        fr = buf[16 + capacity * nat.PHASE2_COLUMN_BYTES: 16 + capacity * (nat.PHASE2_COLUMN_BYTES + 4)].view(np.int32)[:n]
        return cols.copy(), fr.copy()

    def columns_dev(self):
        c, f, n = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._check(lib.e2sec_phase2_columns_dev(self._h, C.byref(c), C.byref(f), C.byref(n)), "e2sec_phase2_columns_dev")
        return c.value, f.value, n.value

    def table(self):
        """Same as Resolver.table()."""
        cols, fr = self.fetch()
        return columns_table(cols, fr, self.relations)

    def state(self):
        lab = np.zeros(3, np.int32)
        ab = np.zeros(3, np.int32)
        self._check(lib.e2sec_phase2_bindings(self._h, ptr(lab), ptr(ab)), "e2sec_phase2_bindings")
        return lab.tolist(), ab.astype(bool).tolist()

    @property
    def stream(self):
        return lib.e2sec_phase2_stream(self._h)


def esec_to_e2sec_native(table, relations, final_columns):
    """esec_to_e2sec (manipulation_analyser.py:62-111) through the C ABI. table: (rows, n) bytes/str array."""
    t = np.asarray(table)
    rows, n = t.shape
    codes = np.zeros((rows, n, 6), np.uint8)
    as_bytes = np.char.encode(t, "ascii") if t.dtype.kind == "U" else t
    flat = as_bytes.astype("S6").reshape(rows, n)
    codes[:] = np.frombuffer(flat.tobytes(), np.uint8).reshape(rows, n, 6)
    out = np.zeros((rows, max(n, 1), 6), np.uint8)
    fc = np.ascontiguousarray(final_columns, np.int32).copy()
    nf = C.c_int32(len(fc))
    fcbuf = np.zeros(max(len(fc), 1), np.int32)
    fcbuf[:len(fc)] = fc
    c2 = lib.e2sec_esec_to_e2sec(ptr(codes), rows, n, int(relations), ptr(out), ptr(fcbuf), C.byref(nf))
    if c2 < 0:
        raise E2SecError("e2sec_esec_to_e2sec failed (%d)" % c2)
    r2 = rows - {3: 6, 2: 4, 1: 2}[int(relations)]
    e2 = out.reshape(-1)[: r2 * c2 * 6].reshape(r2, c2, 6).view("S6").reshape(r2, c2).astype("S5")
    return e2, fcbuf[: nf.value].tolist()


# ---- record accessors ------------------------------------------------------------------------------
def rec_entity(rec, e):
    r = rec[nat.REC_HEADER + nat.REC_ENTITY * e: nat.REC_HEADER + nat.REC_ENTITY * (e + 1)]
    return dict(n_raw=int(r[0]), n_out=int(r[1]), seed=int(r[2]), radius=r[3], median=r[4:7], lo=r[7:10], hi=r[10:13],
                mean=r[13:16])


def rec_pairs(rec, K):
    o = nat.REC_HEADER + nat.REC_ENTITY * K
    return rec[o: o + (K + 1) * (K + 1)].reshape(K + 1, K + 1)


def rec_cross(rec, K):
    o = nat.REC_HEADER + nat.REC_ENTITY * K + (K + 1) * (K + 1)
    return rec[o: o + K * K].reshape(K, K)


def rec_radius_bounds(rec, K):
    """(K, 2) [lo, hi] with float32(lo) <= float32(np.median(pdist())) <= float32(hi); NaN rows: not filtered."""
    o = nat.REC_HEADER + nat.REC_ENTITY * K + (K + 1) * (K + 1) + K * K
    return rec[o: o + 2 * K].reshape(K, 2)
