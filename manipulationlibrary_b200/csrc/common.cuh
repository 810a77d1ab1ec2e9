// common.cuh -- shared device structures and helpers of the e2SEC hot path (sm_100a).
//
// Compiled with --fmad=false: every float64 expression below is evaluated with separate IEEE
// multiplies and adds in source order, which is what numpy / scipy / Cython do on the reference side
// (manipulation_analyser.py and cython_filter_new/filter_cython_new.pyx). Fused operations are
// requested explicitly (fma / fmaf) where a kernel wants them.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/e2sec.h"

#define E2_SEG 1024          // points per warp segment in the segmentation kernels
#define E2_SEG_WARPS 8       // warps per CTA there
#define E2_TILE 256          // points per tile in the pair-distance kernels
#ifndef E2_NBINS
#define E2_NBINS 192         // bins per refinement pass of the radius selection (a multiple of 32; the thread-private
                             // counters of a pass are (E2_NBINS + 2) x 256 bytes per CTA: 192 bins let four CTAs share an SM)
#endif
#define E2_CT_SUB 256        // ground points per pruning box (contact.cuh)
#define E2_CAND_CAP 16384    // candidate keys kept for the final exact selection (one bin of a pass that starts from the
                             // cell table's radius band holds ~0.1 % of an entity's pairs)

typedef unsigned long long u64;

// Per-(frame, entity) state of the radius selection (np.median(pdist()), manipulation_analyser.py:1524).
struct RadiusState {
    u64 klo;        // interval [klo, klo + width) in key space (key = bit pattern of the float64 d^2)
    u64 width;
    u64 below;      // exact number of pairs with key < klo
    u64 k1, k2;     // 0-based ranks of the two middle pair distances (equal when the pair count is odd)
    int shift;      // bin = (key - klo) >> shift
    int phase;      // 0 = refine (histogram pass), 1 = collect, 2 = done, 3 = inactive (not filtered)
    int n;          // points of the entity
    int pad;
    u64 ktop;       // one past the key of the largest possible d^2 (the box diagonal): the interval never has to reach beyond
};

struct CellState;

// Device view of one context. Entity-level arrays are indexed fe = frame * K + e.
struct Dev {
    int F, N, K, cap;                 // frames in this batch, points per frame, entities, stored-point capacity
    int maxF;                         // frame slot maxF of every per-frame buffer is the HALO: a copy of the last kept
                                      // frame of the previous batches (its clouds feed the cross-frame distances)
    int S;                            // segments per frame = ceil(N / E2_SEG)
    int hand_idx, ground_idx;
    int ent_shift;                    // 1 when the hand is index 0 of total_unique_labels: the device then works on K + 1 entities (a
                                      // phantom in front, so that "index 0 is never stored" keeps holding) and the hand is left
                                      // unfiltered as :1518-1520 does for i == 0; assemble_kernel writes the caller's K-entity record
    int hand_label;
    int n_ground;
    int filter_mode;
    int radius_mode;                  // E2SEC_RADIUS_* actually in force (REFERENCE_SEEDED forces EXACT)
    unsigned filter_seed;
    double thresh;
    double T[16], R[16];
    // inputs
    const float *xyz;                 // F x N x 3
    const int *labels;                // F x label_size
    int label_size;                   // label_w * label_h
    const int *rowmap, *colmap;       // nearest-neighbour source row / column per destination row / column (or null)
    int width, label_w;
    const int *lut;                   // raw label -> entity index, -1 untracked
    const int *frame_ids;             // F
    // segmentation
    unsigned *seg_cnt;                // F x S x K   valid points of entity e in segment s (becomes offsets)
    unsigned *ent_box;                // F x K x 6   float32 box of every entity's TRANSFORMED points rounded outward, as order-
                                      //             preserving keys (min x,y,z, max x,y,z): written by seg_emit_kernel, null when
                                      //             the entities were produced another way
    int *ent_raw;                     // F x K       valid points per entity
    int *ent_off;                     // F x K       offset of the entity's points in the frame's storage (-1: not stored)
    int *flags;                       // F           E2SEC_FLAG_*
    int *frame_used;                  // F+1         stored points of the frame
    int *prev_slot;                   // F           frame slot of the previous kept frame (maxF = halo, -1 = none)
    int *prev_id;                     // F           its frame id (-1 = none)
    int *halo_info;                   // [0] valid, [1] frame id, [2] slot of the last kept frame of this batch (-1 none)
    unsigned *seg_len;                // F x units   entries of every unit's scratch list (seg_stream.cuh)
    uint4 *seg_list;                  // F x N       scratch of seg_fused_kernel: compacted (x, y, z, raster index << 6 | entity)
    int affine;                       // translation and roll both end in the row (0, 0, 0, 1): the perspective divide is by exactly 1.0
    double *px, *py, *pz;             // F x cap     transformed points, grouped by entity, raster order inside
    unsigned *pidx;                   // F x cap     raster index of each stored point
    // filter inputs
    double *med;                      // F x K x 3   per-axis median
    double *radius;                   // F x K       median pairwise distance
    RadiusState *rs;                  // F x K
    u64 *hist;                        // F x K x E2_NBINS
    u64 *cand;                        // F x K x E2_CAND_CAP
    unsigned *cand_n;                 // F x K
    // filter outputs
    int *seed_pos;                    // F x K       position of the seed inside the entity (-1: unfiltered)
    unsigned char *member;            // F x cap     closure membership
    unsigned *front_a, *front_b;      // F x cap     frontier lists (positions inside the entity)
    unsigned *flist;                  // F x cap     filtered cloud as positions, output order (seed duplicate excluded)
    int *n_list;                      // F x K       entries of flist
    // bounded-radius route (cells.cuh)
    unsigned short *pcell;            // F x cap     cell of each stored point (hash slot, then compact index)
    double *rad_lo, *rad_hi;          // F x K       float32-rounded bounds of the median pair distance
    int *amb;                         // F x K       1 = the cell table could not decide: exact radius + point-level flood fill;
                                      //             2 = only a point inside the radius band is open: exact radius, then the cell
                                      //                 closure is finished by cell_filter_kernel<true>
    struct CellState *cell_state;     // F x K       closure state of the entities with amb == 2 (cells.cuh)
    int *amb_count;                   // number of such entities in the batch (reset by halo_commit_kernel)
    int *amb_list;                    // F x K       their fe, in the order they turned up: what the exact path walks
    unsigned *exact_ticket;           // CTAs of an exact-path kernel that are done (the last one plans the next pass)
    int test_exact_start;             // tests only (E2SEC_TEST_EXACT_START): 1 / 2 = start the selection under / above the median
    // stats + pairs
    double *aabb;                     // F x K x 6
    double *mean;                     // F x K x 3
    const double *gx, *gy, *gz;       // ground cloud (Morton order)
    const double *gsub;               // ceil(n_ground / E2_CT_SUB) x 6   tight box of every run of E2_CT_SUB ground points
    double gaabb[6];
    u64 *pair_key;                    // F x (K+1) x (K+1)   min d^2 key per in-frame pair
    u64 *cross_key;                   // F x K x K           min d^2 key, entity a of the previous kept frame vs b of this one
    double *records;                  // F x rec_doubles
    int rec_doubles;
    int *error;                       // device error flag
};

// Programmatic dependent launch (e2sec_api.cu: launch_pdl): wait until the grids this one depends on have completed and
// their writes are visible, then let the NEXT grid of the stream start its own launch; it parks at the same wait.
// A no-op for a kernel launched without the attribute.  Must run before the first global-memory access.
__device__ __forceinline__ void pdl_enter()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
// The two halves on their own: a kernel that is launched in plain stream order (e2sec_api.cu: launch_plain) still lets the
// grid after it start its launch early, once its own work is done.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// BOUNDED radius mode and the cell table decided every entity of the batch: the kernels of the exact O(n^2) radius have
// nothing to do (they are launched unconditionally -- the host never waits for the device inside a batch)
__device__ __forceinline__ bool exact_path_idle(const Dev &d)
{
    return d.radius_mode == E2SEC_RADIUS_BOUNDED && *((volatile int *)d.amb_count) == 0;
}

// The entities the exact path works on: in BOUNDED mode the few the cell table left open (amb_list, filled by
// cell_filter_kernel<false>), otherwise every (frame, entity).  Its kernels run small fixed grids that stride over this list.
__device__ __forceinline__ int exact_count(const Dev &d)
{
    return d.radius_mode == E2SEC_RADIUS_BOUNDED ? *((volatile int *)d.amb_count) : d.F * d.K;
}
__device__ __forceinline__ int exact_entity(const Dev &d, int li)
{
    return d.radius_mode == E2SEC_RADIUS_BOUNDED ? d.amb_list[li] : li;
}

// :1518-1520: an entity is filtered when it has more than 3 points and is neither the ground nor index 0 (not stored, or the
// hand itself when ent_shift is set)
__device__ __forceinline__ bool entity_filtered(const Dev &d, int fe, int n)
{
    return n > 3 && !(d.ent_shift && fe % d.K == d.hand_idx);
}

__device__ __forceinline__ u64 d2key(double v) { return (u64)__double_as_longlong(v); }
__device__ __forceinline__ double key2d(u64 k) { return __longlong_as_double((long long)k); }

// sqrt((dx*dx + dy*dy) + dz*dz) squared part, in the operation order of scipy's pdist / distance_cython
__device__ __forceinline__ double dist2(double ax, double ay, double az, double bx, double by, double bz)
{
    double dx = ax - bx, dy = ay - by, dz = az - bz;
    return (dx * dx + dy * dy) + dz * dz;
}

// order-preserving 32-bit key of a float32 and back
__device__ __forceinline__ unsigned f32_ord(float f)
{
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord_f32(unsigned u)
{
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// total order key of a float64 for radix selection
__device__ __forceinline__ u64 f64_sortkey(double v)
{
    u64 b = (u64)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_from_sortkey(u64 k)
{
    u64 b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

__device__ __forceinline__ double warp_min_d(double v)
{
    for (int o = 16; o; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// the same for values that are never negative (squared distances): their bit patterns order like unsigned integers, so
// the minimum is two 32-bit warp reductions (high words, then the low words of the lanes that hold the smallest high word)
// instead of five 64-bit shuffle + compare steps
__device__ __forceinline__ double warp_min_nonneg_d(double v)
{
    const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
    const unsigned mhi = __reduce_min_sync(0xffffffffu, hi);
    const unsigned mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
    return __hiloint2double((int)mhi, (int)mlo);
}
__device__ __forceinline__ double warp_max_d(double v)
{
    for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ u64 warp_sum_u64(u64 v)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ u64 warp_min_u64(u64 v)
{
    for (int o = 16; o; o >>= 1) { u64 w = __shfl_xor_sync(0xffffffffu, v, o); v = w < v ? w : v; }
    return v;
}
