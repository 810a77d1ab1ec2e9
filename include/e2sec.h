/* e2sec.h -- C ABI of the B200-native e2SEC extraction hot path.
 *
 * Drop-in boundary for ONE path of tobi-code/ManipulationLibrary: the per-frame pass of
 * manipulation_analyser.py (_process :1355-2182) and the native operator it calls
 * (cython_filter_new/filter_cython_new.pyx:6-93).  Every entry point below names the reference
 * interface it replaces.  Plain pointers and sizes only; no C++ exceptions cross this boundary.
 *
 * Conventions
 *   - every function returns 0 on success, a negative E2SEC_ERR_* otherwise; e2sec_last_error(ctx)
 *     gives the message (e2sec_last_error(NULL) for failures of e2sec_create itself);
 *   - the caller allocates every output buffer; pointer arguments are HOST pointers unless the
 *     argument name ends in _dev or the function says otherwise;
 *   - one context = one CUDA device + one stream; not thread-safe (the reference keeps its state in
 *     module globals, manipulation_analyser.py:32-59, and is not re-entrant either);
 *   - there is NO CPU fallback: without a CUDA device e2sec_create fails with E2SEC_ERR_CUDA.
 */
#ifndef E2SEC_H
#define E2SEC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define E2SEC_ABI_VERSION 6
#define E2SEC_MAX_ENTITIES 64      /* len(total_unique_labels) upper bound                           */
#define E2SEC_LABEL_LUT 4096       /* raw labels must be integers in [0, E2SEC_LABEL_LUT)            */

#define E2SEC_OK 0
#define E2SEC_ERR_ARG (-1)
#define E2SEC_ERR_CUDA (-2)
#define E2SEC_ERR_CAPACITY (-3)
#define E2SEC_ERR_IO (-4)
#define E2SEC_ERR_STATE (-5)

/* Filter modes.  CLOSURE = deterministic fixed point of the reference's flood fill
 * (connected component of the seed in the radius graph).  REFERENCE_SEEDED replays pyx:61-83 pass by
 * pass with CPython's MT19937 (random.seed(s); random.choice(list)) for the per-call seed
 * filter_seed(base, frame, label_index) = (base + 1000003*frame + 7919*label_index) mod 2^32.        */
#define E2SEC_FILTER_CLOSURE 0
#define E2SEC_FILTER_REFERENCE_SEEDED 1

/* How the flood-fill radius np.median(pdist(obj)) (:1524) is obtained in CLOSURE mode.
 * EXACT: selection over all n(n-1)/2 float64 pair distances.  BOUNDED: guaranteed bounds lo <= median <= hi from a
 * cell partition of the entity; when the closure provably is the same set for every radius in [lo, hi] it is taken
 * from the cell table, otherwise that entity falls back to EXACT.  Both give the same filtered clouds; BOUNDED leaves
 * the record's radius field NaN where it did not need the exact value.  REFERENCE_SEEDED always uses EXACT.       */
#define E2SEC_RADIUS_EXACT 0
#define E2SEC_RADIUS_BOUNDED 1

typedef struct e2sec_ctx e2sec_ctx;

/* Sequence parameters = the arguments of analyse_maniac_manipulation (:2184-2185) that the per-frame
 * pass reads, plus the frame-0 products (:1397-1426) which are INPUTS here.           */
typedef struct e2sec_params {
    int32_t n_entities;                          /* len(total_unique_labels) (:1406-1426)              */
    int32_t entity_label[E2SEC_MAX_ENTITIES];    /* total_unique_labels, ascending as np.unique gives  */
    int32_t hand_label, ground_label;
    int32_t support_hand;                        /* label, or -1 for None                              */
    int32_t n_replace;                           /* replace=True: old[i] -> new[i], sequential (:176)  */
    int32_t replace_old[16], replace_new[16];
    double thresh;                               /* touching threshold                                 */
    double translation[16], roll[16];            /* row-major 4x4 (:1484-1485)                         */
    int32_t n_ground;                            /* the frame-0 ground cloud, already transformed      */
    const double *ground_xyz;                    /* n_ground x 3, row-major, host                      */
    int32_t filter_mode;                         /* E2SEC_FILTER_*                                     */
    uint32_t filter_seed;                        /* base seed for REFERENCE_SEEDED                     */
    int32_t radius_mode;                         /* E2SEC_RADIUS_*                                     */
    int32_t relations;                           /* 1: T/N, 2: + SSR, 3 (or 0): + DSR (:1744-1829)     */
} e2sec_params;

/* ---- records: phase-1 output, one per frame, all doubles (integers are exact in a double) --------
 * record[0] frame id   record[1] flags (bit0: hand label occurs in the frame's labels (:1429),
 *                                       bit1: filtered hand non-empty (:1475/:1551))
 * record[2] n_entities K   record[3] frame id of the previous KEPT frame (bit0 and bit1 set) whose clouds the
 *                                       cross table below was computed against, -1 if there is none
 * entity e at record[4 + 16*e ..]:
 *    +0 n_raw (valid points of the label)   +1 n_out (points of the filtered cloud, seed counted twice)
 *    +2 seed raster index (-1: not filtered) +3 radius = np.median(pdist()) (:1524), NaN if unfiltered
 *    +4..6 per-axis median (:1523)   +7..9 AABB min   +10..12 AABB max   +13..15 mean of the cloud
 * pair table at record[4 + 16*K ..], (K+1) x (K+1) row-major, index K = the ground cloud:
 *    exact minimum point-to-point distance when it is < thresh; otherwise some value >= thresh
 *    (+inf when either cloud is empty).  Entities 0 (border) and ground_label's own index hold +inf.
 * cross table after it, K x K row-major: [a][b] = contact flag between entity a's filtered cloud in the previous
 *    kept frame and entity b's filtered cloud in this frame: 0.0 when their minimum distance is < thresh, otherwise
 *    some value >= thresh (+inf when either is empty; only `< thresh` is ever asked of these, :1600-1653).  _process keeps o1/o2/o3 as the clouds
 *    of the previous frame that passed :1551 until :1669-1675 re-assigns them, so the o2 / o3 search of
 *    :1562-1658 measures exactly these cross-frame distances.
 * radius bounds after it, K x 2: [lo, hi] with float32(lo) <= float32(np.median(pdist())) <= float32(hi) for every
 *    filtered entity (lo == hi == radius where the exact value was computed), NaN for unfiltered entities.       */
#define E2SEC_REC_HEADER 4
#define E2SEC_REC_ENTITY 16
#define E2SEC_FLAG_HAND_LABEL 1
#define E2SEC_FLAG_HAND_POINTS 2
size_t e2sec_record_doubles(int32_t n_entities);

/* ---- context ----------------------------------------------------------------------------------- */
int e2sec_abi_version(void);
int e2sec_device_count(void);        /* CUDA devices visible to the process (0: none -- and no CPU fallback) */
int e2sec_create(int device, e2sec_ctx **out);
void e2sec_destroy(e2sec_ctx *ctx);
const char *e2sec_last_error(const e2sec_ctx *ctx);
/* cudaStream_t the context launches on (for callers that time with CUDA events) */
void *e2sec_stream(e2sec_ctx *ctx);

/* Sets the sequence parameters and (re)allocates device buffers for batches of up to max_frames
 * frames of n_points points.  Replaces the global reset of analyse_maniac_manipulation :2224-2250
 * and the frame-0 branch of _process :1397-1426 (whose products arrive in `p`).                     */
int e2sec_configure(e2sec_ctx *ctx, const e2sec_params *p, int32_t max_frames, int32_t n_points);
/* Forgets the previous kept frame (the halo the cross-frame distances are measured against): call it before
 * the first batch of a new sequence or shard.  Frames of one sequence must then arrive in order; a shard
 * that starts mid-sequence feeds its halo frame(s) first and drops their records.                     */
int e2sec_reset_sequence(e2sec_ctx *ctx);

/* Phase 1 of _process (:1433-1546 + the geometry behind :363-1353) for a batch of frames: label
 * upscale/replace, NaN mask, segmentation, transform, per-entity filter, AABB / mean, all-pairs
 * touching distances.  Stream-ordered; returns after the records are on the host unless
 * records is NULL (then read them with e2sec_records_dev / your own copy after e2sec_sync).
 *   xyz        n_frames x n_points x 3 float32 (PCD order, NaN = invalid), host or device
 *   labels     n_frames x (label_w*label_h) int32, host or device; if label_w*label_h != n_points the
 *              image is upscaled nearest-neighbour to (width x height) like cv2.resize (:1433)
 *   frame_ids  n_frames int32 host (file index i of :2271; seeds REFERENCE_SEEDED)
 *   on_device  non-zero: xyz and labels are device pointers                                         */
int e2sec_process_batch(e2sec_ctx *ctx, const float *xyz, const int32_t *labels, int32_t n_frames,
                        int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                        const int32_t *frame_ids, int on_device, double *records);
/* The same, stream-ordered and WITHOUT waiting: `records` (pinned host memory, or NULL) is filled by an asynchronous
 * copy, host inputs must stay untouched and device-side errors are reported by the next e2sec_sync / e2sec_process_batch.
 * Lets a caller queue several batches back to back (bench.py) -- the reference's loop (:2271-2281) has no equivalent.
 * `records` may also be DEVICE memory here (a shard that keeps all its records for e2sec_phase2_run).                  */
int e2sec_process_batch_async(e2sec_ctx *ctx, const float *xyz, const int32_t *labels, int32_t n_frames,
                              int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                              const int32_t *frame_ids, int on_device, double *records);
/* The same two calls for label images stored as uint8 (every label of the sequence < 256, as in the MANIAC label files
 * the reference reads at :1391): 13 instead of 16 bytes per point cross PCIe; the image is widened on the device.     */
int e2sec_process_batch_labels8(e2sec_ctx *ctx, const float *xyz, const uint8_t *labels, int32_t n_frames,
                                int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                const int32_t *frame_ids, int on_device, double *records);
int e2sec_process_batch_labels8_async(e2sec_ctx *ctx, const float *xyz, const uint8_t *labels, int32_t n_frames,
                                      int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                      const int32_t *frame_ids, int on_device, double *records);
/* The same pass with the FILES as input (input decode on the device): what the reference reads with
 * o3d.io.read_point_cloud (:1441) and pd.read_csv (:1391).  pcd_files holds the n_frames PCD v0.7 files back to back, file f
 * at [pcd_offsets[f], pcd_offsets[f + 1]); they must be DATA binary_compressed with float32 x, y, z as their first three
 * fields and width * height = n_points points (any other cloud: e2sec_read_pcd_xyz + e2sec_process_batch).  label_files
 * likewise holds the label files -- label_h lines of label_w whitespace-separated non-negative integers each.  Headers are
 * read on the host (a few hundred bytes per file); the compressed payloads and the text are uploaded as they are, LZF-decoded
 * and parsed on the device (csrc/lzf.cuh) on the context's copy stream, under the kernels of the batch before.  Host memory
 * (pinned for the _async form, and untouched until e2sec_sync); records as in e2sec_process_batch.                         */
int e2sec_process_batch_pcd(e2sec_ctx *ctx, const uint8_t *pcd_files, const int64_t *pcd_offsets, const uint8_t *label_files,
                            const int64_t *label_offsets, int32_t n_frames, int32_t label_w, int32_t label_h, int32_t width,
                            int32_t height, const int32_t *frame_ids, double *records);
int e2sec_process_batch_pcd_async(e2sec_ctx *ctx, const uint8_t *pcd_files, const int64_t *pcd_offsets, const uint8_t *label_files,
                                  const int64_t *label_offsets, int32_t n_frames, int32_t label_w, int32_t label_h, int32_t width,
                                  int32_t height, const int32_t *frame_ids, double *records);
/* device addresses of the decoded inputs of the batch queued last from host memory or files: xyz n_frames x n_points x 3
 * float32, labels n_frames x (label_w * label_h) int32 (valid until the batch after the next one is queued)              */
int e2sec_staged_inputs_dev(e2sec_ctx *ctx, const float **xyz_dev, const int32_t **labels_dev);
/* ... and copied to the host (waits for the context's stream): what the device decoders made of the files */
int e2sec_copy_staged_inputs(e2sec_ctx *ctx, float *xyz, int32_t *labels, int32_t n_frames, int32_t label_size);
/* The clouds behind a record of the batch processed last, for callers that track rigid models (icp=True, :1677-1738):
 * entity (index into total_unique_labels) of frame (position in the batch).  which = 0: every valid point of the label,
 * transformed = pcd[i] (:1516); which = 1: filtered_pcd_voxel[i] (:1519-1546; the seed point first -- it is in the
 * reference's cloud twice, pyx:48 --, then the kept points in raster order).  out: capacity x 3 (may be NULL to ask for
 * the size); entities that are never stored (index 0 unless it is the hand, the ground label) give 0 points.            */
int e2sec_entity_cloud(e2sec_ctx *ctx, int32_t frame, int32_t entity, int32_t which, double *out, int32_t capacity, int32_t *n_out);
const double *e2sec_records_dev(e2sec_ctx *ctx);
/* Stream-ordered copy of records [first, first + count) of the batch queued last to dst (host or device memory): a shard
 * that keeps its records on the device for e2sec_phase2_run, minus the halo frame's.  Does not wait.                  */
int e2sec_copy_records(e2sec_ctx *ctx, double *dst, int32_t first, int32_t count);
/* Waits for the context's stream; returns the first device-side error of the batches queued since the last check. */
int e2sec_sync(e2sec_ctx *ctx);
/* number of kernels launched by this context so far (bench.py's gpu_launches) */
int64_t e2sec_launch_count(const e2sec_ctx *ctx);
/* per-stage device time of the last batch in ms: segment, medians, radius, filter, stats, pairs, assemble;
 * out[7] = the streaming segmentation kernel alone (part of "segment")                              */
int e2sec_stage_ms(e2sec_ctx *ctx, float out[8]);

/* ---- operators at the reference's own seams ------------------------------------------------------ */
/* filter_cython_new.region_filter_cython (pyx:9).  pcd: n x 3 float64 C-contiguous.  `center` is the
 * seed target (the reference passes hand_center == obj_center, :1525).  out must hold (n+1) x 3.     */
int e2sec_region_filter(e2sec_ctx *ctx, const double *pcd, int32_t n, const double center[3],
                        double median_distance, int32_t mode, uint32_t seed, double *out,
                        int32_t *n_out);
/* filter_cython_new.distance_cython (pyx:6) -- host arithmetic, kept for API completeness.           */
double e2sec_distance(const double p1[3], const double p2[3]);
/* np.median(scipy.spatial.distance.pdist(pts)) (:1524), exact.                                       */
int e2sec_pairwise_median(e2sec_ctx *ctx, const double *pts, int32_t n, double *median);
/* np.median(pts, axis=0) (:1523), exact.                                                            */
int e2sec_axis_median(e2sec_ctx *ctx, const double *pts, int32_t n, double median[3]);
/* np.min(a.compute_point_cloud_distance(b)) (:397-484): exact when < limit, else a value >= limit.
 * limit <= 0 means "always exact".                                                                   */
int e2sec_min_distance(e2sec_ctx *ctx, const double *a, int32_t na, const double *b, int32_t nb,
                       double limit, double *dist);

/* ---- frame-0 setup (csrc/frame0.cuh): the Open3D calls of _getGroundiNewNew (:278-361) and _getTranslation (:212-276) as
 * device operators; the host side (manipulationlibrary_b200/frame0.py) strings them together deterministically.
 * segment_plane (:353): counts[h] = number of points with |p . n_h + d_h| < threshold for m <= 64 plane hypotheses
 * (planes: m x 4, host), and the inlier mask of one plane.                                                              */
int e2sec_plane_inlier_counts(e2sec_ctx *ctx, const double *pts, int32_t n, const double *planes, int32_t m, double threshold,
                              int32_t *counts);
int e2sec_plane_inlier_mask(e2sec_ctx *ctx, const double *pts, int32_t n, const double plane[4], double threshold, uint8_t *mask);
/* voxel_down_sample (:224, :359): one point per occupied voxel of size `voxel` = the mean of its points (summed in their
 * order of appearance), voxels in ascending (ix, iy, iz) with origin min_bound - voxel / 2.  out: up to n x 3.            */
int e2sec_voxel_down_sample(e2sec_ctx *ctx, const double *pts, int32_t n, double voxel, double *out, int32_t *n_out);
/* One correspondence step of registration_icp (:260, point to point): every source point, moved by `transform` (row-major
 * 4 x 4), is matched to its exact nearest target point if that lies within max_distance.  stats[0] = matched pairs,
 * [1] = sum of squared distances, [2..4] / [5..7] = centroids of the matched source / target points, [8..16] = the centred
 * cross-covariance sum (a - ca)(b - cb)^T row-major.  The 3 x 3 SVD of the update stays with the caller.                  */
int e2sec_icp_correspondences(e2sec_ctx *ctx, const double *source, int32_t n_source, const double *target, int32_t n_target,
                              const double transform[16], double max_distance, double stats[17]);
/* The same correspondence step for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * `source` (n_source x 3), `target` (n_target x 3) and `stats` (17 doubles) are DEVICE pointers the caller owns on ctx's
 * device; `transform` is host memory.  The kernels' scratch (moved points, nearest index and distance per source point,
 * 36 bytes each) lives in a file-scope __device__ array of the library sized for E2SEC_STATIC_ICP_MAX_SOURCE points
 * (2.25 MiB on every GPU that loads the library, until the process exits); n_source beyond that returns E2SEC_ERR_CAPACITY
 * with a message in e2sec_last_error.  The call makes no cudaMalloc / cudaFree and no allocating library call, runs on
 * ctx's stream and returns after the stream has drained.  All calls on a device share the one array (it keeps the last
 * call's contents; every element a call reads it has written first, nothing has to start at zero): two calls of one
 * process that overlap on a GPU (from other contexts or threads) race and must be serialised by the caller.
 * Results are bit-identical to e2sec_icp_correspondences.                                                                */
#define E2SEC_STATIC_ICP_MAX_SOURCE 65536
int e2sec_icp_correspondences_static(e2sec_ctx *ctx, const double *source_dev, int32_t n_source, const double *target_dev,
                                     int32_t n_target, const double transform[16], double max_distance, double *stats_dev);

/* ---- phase 2: the sequential part of _process (:1558-1675 selection, :363-1353 relation rows,
 * :1795-1829 / :2000-2007 column emit) over records in frame order.  Host code. -------------------- */
typedef struct e2sec_resolver e2sec_resolver;
int e2sec_resolver_create(const e2sec_params *p, e2sec_resolver **out);
void e2sec_resolver_destroy(e2sec_resolver *r);
/* Feeds n_frames records (frame order).  Emitted 30-row columns accumulate inside the resolver.      */
int e2sec_resolver_feed(e2sec_resolver *r, const double *records, int32_t n_frames);
/* Number of eSEC columns so far (0 before the first column; the reference's '-' column is implicit). */
int32_t e2sec_resolver_columns(const e2sec_resolver *r);
const char *e2sec_resolver_error(const e2sec_resolver *r);
/* Copies the raw eSEC table: codes as (10 * relations) x n_cols x 6 bytes (NUL-terminated, lower/mixed case as the
 * reference writes them) and the frame id of every column (n_cols).  With relations 1 / 2 column 0 is the '-' column
 * the reference's table starts with (frame id -1) unless frame 0 itself produced a column (:1753, :1778).            */
int e2sec_resolver_table(const e2sec_resolver *r, char *codes, int32_t *final_columns);
/* o1/o2/o3 entity indices (-1 while unassigned) and sticky absent flags (:389-427).                  */
int e2sec_resolver_state(const e2sec_resolver *r, int32_t labels[3], int32_t absent[3]);
/* esec_to_e2sec (:62-111) on a rows x n_cols table of 6-byte codes: upper-cases, drops column 0 and
 * the H,G / 3,G rows, removes repeated columns; returns the new column count, thins final_columns.   */
int32_t e2sec_esec_to_e2sec(const char *codes, int32_t rows, int32_t n_cols, int32_t relations,
                            char *out_codes, int32_t *final_columns, int32_t *n_final);

/* _fillTN_absent (:363-487), _fillSSR_8 (:636-1061) and _fillDSR_new_2 (:1064-1353) on explicit inputs -- the
 * reference's own seam for the relation rows (the resolver and the device classifier run the same code, csrc/relations.h).
 * Slots of cur / prev follow previous_array (:1799): 0 hand, 1 ground, 2..4 o1..o3.  defined = 0 means the Python variable
 * is None.  center = what the DSR rows read: the cloud mean for the hand, the AABB centre for objects and ground.
 * min_distance[r] = np.min(a.compute_point_cloud_distance(b)) of row r's pair in the order (H,1) (H,2) (H,3) (H,G) (1,2)
 * (1,3) (1,G) (2,3) (2,G) (3,G) (:373-383); ignored where a cloud is None or empty.  absent = absent_o1..3 on entry;
 * absent_out (may be NULL) receives them after the call (sticky flags, :394-396).
 * codes: 10 * relations entries of 6 bytes, NUL-terminated, spelled as the reference writes them.               */
typedef struct e2sec_relation_entity {
    int32_t defined, n_points;
    double aabb_min[3], aabb_max[3], center[3];
} e2sec_relation_entity;
typedef struct e2sec_relation_input {
    e2sec_relation_entity cur[5], prev[5];
    double min_distance[10];
    double thresh;
    int32_t absent[3];
} e2sec_relation_input;
int e2sec_relation_rows(const e2sec_relation_input *in, int32_t relations, char *codes, int32_t absent_out[3]);

/* ---- phase 2 on the device: the same selection / relation rows / column dedup as the resolver above, as ONE kernel over a
 * run of records that are already in device memory (csrc/phase2.cuh; one thread per frame, the sequential state -- three
 * bindings, three sticky bits, the previous kept frame, the last column -- carried by scans).  Only the emitted columns
 * (E2SEC_PHASE2_COLUMN_BYTES code bytes + the frame id each) ever leave the GPU.
 * The object belongs to a configured context (same device, the context's parameters) and works on its own stream.        */
#define E2SEC_PHASE2_STATE_BYTES 512
#define E2SEC_PHASE2_COLUMN_BYTES 32
typedef struct e2sec_phase2 e2sec_phase2;
int e2sec_phase2_create(e2sec_ctx *ctx, int32_t max_columns, e2sec_phase2 **out);
void e2sec_phase2_destroy(e2sec_phase2 *p2);
/* a new sequence: no bindings, no previous frame, no columns */
int e2sec_phase2_reset(e2sec_phase2 *p2);
/* Classifies n_frames records (frame order, device memory, e2sec_record_doubles(K) doubles each) behind everything fed
 * before.  Does not wait.  src (may be NULL) = the context whose stream produces the records, e.g.
 * e2sec_phase2_run(p2, ctx, e2sec_records_dev(ctx), n) right after e2sec_process_batch_async(ctx, ..., NULL): phase 2 waits
 * for that batch, and the context's next batch waits until its record buffer has been read.                               */
int e2sec_phase2_run(e2sec_phase2 *p2, e2sec_ctx *src, const double *records_dev, int32_t n_frames);
/* The state a shard of a frame-sharded sequence hands to the next one (opaque, E2SEC_PHASE2_STATE_BYTES; dst / src may be
 * host or device memory, e.g. the buffer of an NCCL send / recv).  Both wait for the object's stream.                     */
int e2sec_phase2_state_get(e2sec_phase2 *p2, void *dst);
int e2sec_phase2_state_set(e2sec_phase2 *p2, const void *src);
/* The same handing-over WITHOUT the host, for one process per GPU on one box: every object owns a mailbox in device memory;
 * e2sec_phase2_mailbox_handle gives its CUDA IPC handle (E2SEC_IPC_HANDLE_BYTES, exchanged once between the processes),
 * e2sec_phase2_connect opens the mailbox of the rank that runs the NEXT frames as peer memory (NVLink / NVSwitch).
 * e2sec_phase2_run_ring is e2sec_phase2_run whose kernel first waits -- on the device -- until the state posted under
 * wait_ticket has arrived in its own mailbox (0: start from the object's own state), and at the end writes its state and
 * then post_ticket straight into the next rank's mailbox (0: post nothing).  Tickets are any non-zero numbers that differ
 * from hop to hop, e.g. 1 + the running index of the batch in the sequence.  A ticket that does not arrive within 20 s
 * is reported by e2sec_phase2_sync as E2SEC_ERR_STATE.                                                                    */
#define E2SEC_IPC_HANDLE_BYTES 64
int e2sec_phase2_mailbox_handle(e2sec_phase2 *p2, void *handle);
int e2sec_phase2_connect(e2sec_phase2 *p2, const void *next_handle);
/* the same for two objects of ONE process (several GPUs driven by one process, or two objects on one GPU) */
int e2sec_phase2_connect_local(e2sec_phase2 *p2, e2sec_phase2 *next);
int e2sec_phase2_run_ring(e2sec_phase2 *p2, e2sec_ctx *src, const double *records_dev, int32_t n_frames,
                          uint32_t wait_ticket, uint32_t post_ticket);
/* Waits; reports device-side errors (records out of order, more than max_columns columns); n_columns (may be NULL) = the
 * columns this object holds.                                                                                              */
int e2sec_phase2_sync(e2sec_phase2 *p2, int32_t *n_columns);
/* The packed columns (n x E2SEC_PHASE2_COLUMN_BYTES relation codes as csrc/relations.h numbers them, rows 0..10*relations-1
 * used) and the frame id of each, to the host; and their device addresses for a gather between GPUs.                      */
int e2sec_phase2_fetch(e2sec_phase2 *p2, uint8_t *columns, int32_t *frames, int32_t capacity, int32_t *n_columns);
/* The same in ONE fixed-size buffer (host or device) for a gather between GPUs: [int32 n][12 bytes unused][capacity x
 * E2SEC_PHASE2_COLUMN_BYTES codes][capacity x int32 frame ids] = 16 + 36 * capacity bytes.  Waits for the object's stream. */
int e2sec_phase2_export(e2sec_phase2 *p2, void *dst, int32_t capacity);
int e2sec_phase2_columns_dev(e2sec_phase2 *p2, const uint8_t **columns_dev, const int32_t **frames_dev, const int32_t **n_dev);
/* o1/o2/o3 entity indices (-1 while unassigned) and the sticky absent flags, like e2sec_resolver_state */
int e2sec_phase2_bindings(e2sec_phase2 *p2, int32_t labels[3], int32_t absent[3]);
void *e2sec_phase2_stream(e2sec_phase2 *p2);
/* Packed columns (host) -> the table of e2sec_resolver_table: codes (10*relations) x total x 6 bytes and the frame id per
 * column; returns total = n (+ 1 for the initial '-' column with relations 1 / 2, :2252-2260, :1753).  With codes and
 * final_columns both NULL only the count is returned.                                                                    */
int32_t e2sec_columns_table(const uint8_t *columns, const int32_t *frames, int32_t n, int32_t relations, char *codes,
                            int32_t *final_columns);

/* ---- the consumer of the tables: eSEC_analyser.getDissimilarityMatrix (eSEC_analyser.py:782-841, with
 * _make_triple :866-889 and _compare_triple_matrics :915-944) and similarity_manipulations (:843-864).
 * n_tables tables of `rows` (= 3 blocks of rows/3: T/N, SSR, DSR) x n_cols[m] codes of 6 bytes each
 * (NUL-terminated, e2sec_resolver_table's / e2sec_esec_to_e2sec's layout) start at codes + offsets[m].
 * Every pair is compared on the device, the shorter table continued by its last column.
 * D (n_tables x n_tables, host) = sum over triples of sqrt(l)/sqrt(3) / (max(n_i, n_j) * divisor), l = the
 * components of the triple that differ; divisor is 10 in getDissimilarityMatrix, 8 in
 * similarity_manipulations (which returns (1 - D) * 100).  counts (may be NULL): n_tables x n_tables x 3
 * triples with l = 1, 2, 3 -- the integers D is made of.  The counts are exact; D is formed from them as
 * (c1/sqrt(3) + c2*sqrt(2)/sqrt(3) + c3) / (n * divisor) and agrees with the reference's pairwise float sum to <= 1e-13
 * absolute, not bit for bit: callers that break ties on D should compare the counts.  All tables share `rows`.        */
int e2sec_dissimilarity(e2sec_ctx *ctx, const char *codes, const int64_t *offsets, const int32_t *n_cols,
                        int32_t n_tables, int32_t rows, double divisor, double *D, uint32_t *counts);
/* device time of the two kernels of the last e2sec_dissimilarity call, in ms */
int e2sec_dissimilarity_ms(e2sec_ctx *ctx, float *ms);

/* ---- input decode on the host (any PCD v0.7 flavour; e2sec_process_batch_pcd decodes the common one on the device) ---- */
int e2sec_read_pcd_xyz(const char *path, float *xyz, int64_t capacity_points, int32_t *width,
                       int32_t *height);
int e2sec_read_labels(const char *path, int32_t *labels, int64_t capacity, int32_t *rows, int32_t *cols);

#ifdef __cplusplus
}
#endif
#endif /* E2SEC_H */
