// cells.cuh -- the bounded-radius route of the region filter, one CTA per (frame, entity).
//
// The reference filters every entity with a flood fill of radius float32(np.median(pdist(obj)))
// (manipulation_analyser.py:1523-1525, cython_filter_new/filter_cython_new.pyx:9-93).  The median of the
// n(n-1)/2 pair distances is O(n^2); only the CLOSURE it produces is used downstream.  This kernel partitions
// the entity's points into cells and works on the cell table:
//   A  seed (pyx:18-31, float32 strict minimum, first index), mean absolute deviation per axis (robust grid
//      range, so that far outliers land in overflow cells of their own), AABB of the entity
//   B  every point -> cell; shared-memory hash of occupied cells; per cell: count + float32 box rounded OUTWARD
//   C  every cell pair (a <= b): all count_a * count_b point pairs have their float64 squared distance inside
//      [dmin2(a,b), dmax2(a,b)] (box-box minimum / maximum, evaluated with the reference's own operation order
//      (dx*dx + dy*dy) + dz*dz, which is monotone in |dx|, |dy|, |dz| under round-to-nearest) -> two weighted
//      histograms over d^2
//   D  rank selection: lo <= median <= hi, r_lo = float32(lo) <= float32(median) <= float32(hi) = r_hi
//   E  cell closure at r_lo: a cell joins when its dmax to a member cell (or to the seed point) is <= r_lo --
//      then every one of its points is reachable at every radius in [r_lo, r_hi]
//   F  if no non-member cell has dmin <= r_hi to a member cell, no further point can be reached at any radius in
//      [r_lo, r_hi]: the closure of the TRUE radius is exactly the union of the member cells.  Otherwise the
//      points of the cells that could still be reached (few) are settled one by one with exact float64 distances
//      to the member points: <= r_lo joins, > r_hi stays out; a point in between makes the entity ambiguous: the
//      exact radius is selected (radius.cuh) and a SECOND VISIT of this kernel (cell_filter_kernel<true>) settles the
//      same candidate points at that radius from the saved closure (CellState) -- exact by the argument of E / F, and
//      no flood fill over the whole cloud.  Only a cell table that overflows or more than E2_CELL_UMAX candidate points
//      send the entity to the point-level flood fill of filter.cuh.
//   G  stable compaction of the member points (raster order, like the exact path) + AABB / mean / count
#pragma once
#include "common.cuh"
#include "filter.cuh"

#ifndef E2_CELL_THREADS
#define E2_CELL_THREADS 384        // 384 x 3 CTAs/SM beat 512 x 2 by 9 % (tools/exp_cell_threads.sh): more, smaller CTAs hide the barriers
#endif
#ifndef E2_CELL_CTAS
#define E2_CELL_CTAS 3              // resident CTAs per SM the register budget is cut for
#endif
#define E2_CELL_WARPS (E2_CELL_THREADS / 32)
#ifndef E2_CELL_MAX
#define E2_CELL_MAX 1024          // occupied cells per entity
#endif
#ifndef E2_CELL_HASH_LOG
#define E2_CELL_HASH_LOG 12
#endif
#define E2_CELL_HASH (1 << E2_CELL_HASH_LOG)   // hash slots
#ifndef E2_CELL_G
#define E2_CELL_G 24              // cells along the largest robust extent; ids use 5 bits per axis
#endif
#define E2_CELL_BINS 1024         // histogram bins over d^2
#define E2_CELL_EMPTY 0xffffffffu
#define E2_CELL_JOIN (E2_CELL_MAX - 1)   // virtual cell of candidate points that were settled as members one by one
#define E2_CELL_UMAX 256          // candidate points settled at point level before giving up
#define E2_CELL_MCAP (E2_CELL_MAX < 1024 ? E2_CELL_MAX : 1024)   // median selection: values ranked in shared memory once the middle bin holds no more than this
#define E2_CELL_MSMALL 64         // ... directly once a sub-bin of it holds no more than this
#define E2_CELL_HPAD(b) ((b) + ((b) >> 5))   // histogram slot of bin b: a lane that walks 32 consecutive bins meets 32 banks
#define E2_CELL_HSIZE (E2_CELL_BINS + E2_CELL_BINS / 32)
#define E2_M_DONE 0
#define E2_M_HIST 1               // needs a histogram round over [m_lo, m_hi]
#define E2_M_COLLECT 2            // both middle ranks in bin b1, few values: collect them
#define E2_M_REFINE 3             // both in bin b1, many values: shrink the range to that bin's values
#define E2_M_SPLIT 4              // the ranks straddle two bins: a = max of bin b1, b = min of bin b2

struct CellSmem {
    union {
        unsigned mhist[3][E2_CELL_HSIZE];     // M: per-axis value histograms of the median selection
        unsigned keys[E2_CELL_HASH];          // B: cell ids
        u64 hist[2][E2_CELL_HSIZE];           // C, D: weights by dmin2 / dmax2 bin
    };
    unsigned short cidx[E2_CELL_HASH];        // slot -> compact cell index
    unsigned cnt[E2_CELL_MAX];
    union {
        unsigned box[6][E2_CELL_MAX];         // lo x,y,z, hi x,y,z: order-preserving uint of a float during B, floats after
        double mcand[3][E2_CELL_MCAP];        // M: values of the bin that holds the middle rank(s)
    };
    // M: state of the three selections
    double m_lo[3], m_hi[3], m_scale[3], m_a[3], m_b[3];
    unsigned m_below[3], m_before[3], m_ccnt[3], m_scnt[3], m_sbef[3][2];
    int m_fb[3];                              // round-0 bin an axis was refined into (-1: none): its values are the only live ones
    int m_sb[3][2];
    double msmall[3][E2_CELL_MSMALL];         // M: values of the sub-bin(s) that hold the middle rank(s)
    u64 m_bef64[3];
    int m_mode[3], m_b1[3], m_b2[3];
    double ebox[6];                           // value range of the entity per axis (exact, or its outward float32 box)
    unsigned ebk[6];                          // union of the cell boxes as order-preserving keys
    unsigned char member[E2_CELL_MAX];
    unsigned short front[2][E2_CELL_MAX];
    unsigned short rest[E2_CELL_MAX];
    double red[9][E2_CELL_WARPS];
    float best_d[E2_CELL_WARPS];
    int best_i[E2_CELL_WARPS];
    unsigned wcnt[E2_CELL_WARPS];
    unsigned char cand[E2_CELL_MAX];
    unsigned ulist[E2_CELL_UMAX];
    unsigned char joined[E2_CELL_UMAX];
    double udist[E2_CELL_UMAX];
    unsigned m, n_front, n_new, n_cand, n_rest, overflow, amb;
    int seed;
    double lo[3], inv_c, scale, d2max, t_lo, t_hi, r_lo, r_hi, seedp[3];
};

__device__ __forceinline__ double next_down(double v) { return v > 0.0 ? __longlong_as_double(__double_as_longlong(v) - 1) : 0.0; }
__device__ __forceinline__ double next_up(double v) { return __longlong_as_double(__double_as_longlong(v) + 1); }

__device__ __forceinline__ int cell_bin(double d2, double scale)
{
    double t = d2 * scale;
    return t >= (double)(E2_CELL_BINS - 1) ? E2_CELL_BINS - 1 : (int)t;
}

// box-box squared distance bounds from the float boxes, in float32 with directed rounding: every step rounds the lower
// bound down and the upper bound up, so  lo <= (exact real d^2 of any point pair) <= hi;  the final factor (1 -+ 2^-23)
// absorbs the rounding of the reference's own float64 evaluation (dx*dx + dy*dy) + dz*dz (relative error < 2^-50).
// float32 halves the pipe time of these O(cells^2) loops against float64 and needs no conversions.
__device__ __forceinline__ float cell_pair_lo32(const CellSmem &s, int a, int b)
{
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float alo = __uint_as_float(s.box[k][a]), ahi = __uint_as_float(s.box[3 + k][a]);
        const float blo = __uint_as_float(s.box[k][b]), bhi = __uint_as_float(s.box[3 + k][b]);
        const float g = fmaxf(fmaxf(__fsub_rd(alo, bhi), __fsub_rd(blo, ahi)), 0.0f);
        acc = __fadd_rd(acc, __fmul_rd(g, g));
    }
    return __fmul_rd(acc, 0.99999988f);
}
__device__ __forceinline__ float cell_pair_hi32(const CellSmem &s, int a, int b)
{
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float alo = __uint_as_float(s.box[k][a]), ahi = __uint_as_float(s.box[3 + k][a]);
        const float blo = __uint_as_float(s.box[k][b]), bhi = __uint_as_float(s.box[3 + k][b]);
        const float f = fmaxf(__fsub_ru(ahi, blo), __fsub_ru(bhi, alo));
        acc = __fadd_ru(acc, __fmul_ru(f, f));
    }
    return __fmul_ru(acc, 1.00000012f);
}
__device__ __forceinline__ void cell_pair_d2(const CellSmem &s, int a, int b, double &dmin2, double &dmax2)
{
    float lo = 0.0f, hi = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float alo = __uint_as_float(s.box[k][a]), ahi = __uint_as_float(s.box[3 + k][a]);
        const float blo = __uint_as_float(s.box[k][b]), bhi = __uint_as_float(s.box[3 + k][b]);
        const float g = fmaxf(fmaxf(__fsub_rd(alo, bhi), __fsub_rd(blo, ahi)), 0.0f);
        const float f = fmaxf(__fsub_ru(ahi, blo), __fsub_ru(bhi, alo));
        lo = __fadd_rd(lo, __fmul_rd(g, g));
        hi = __fadd_ru(hi, __fmul_ru(f, f));
    }
    dmin2 = (double)__fmul_rd(lo, 0.99999988f);
    dmax2 = (double)__fmul_ru(hi, 1.00000012f);
}
// the same between a point and a box
__device__ __forceinline__ double cell_point_dmax2(const CellSmem &s, const double *p, int b)
{
    double f[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double blo = (double)__uint_as_float(s.box[k][b]), bhi = (double)__uint_as_float(s.box[3 + k][b]);
        f[k] = fmax(p[k] - blo, bhi - p[k]);
    }
    return (f[0] * f[0] + f[1] * f[1]) + f[2] * f[2];
}

// block-wide reductions over E2_CELL_THREADS threads; `slot` picks a row of s.red.  Every thread reads the
// E2_CELL_WARPS partials with one shuffle tree instead of a serial loop.
__device__ __forceinline__ double cell_block_sum(CellSmem &s, int slot, double v)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s.red[slot][threadIdx.x >> 5] = v;
    __syncthreads();
    const int l = threadIdx.x & 31;
    double t = (l < E2_CELL_WARPS) ? s.red[slot][l] : 0.0;
    // fixed summation order (warp 0, 1, 2, ...) regardless of the lane: the mean must not depend on the thread that asks
    double acc = 0.0;
    for (int w = 0; w < E2_CELL_WARPS; ++w) acc += __shfl_sync(0xffffffffu, t, w);
    return acc;
}
__device__ __forceinline__ double cell_block_min(CellSmem &s, int slot, double v)
{
    v = warp_min_d(v);
    if ((threadIdx.x & 31) == 0) s.red[slot][threadIdx.x >> 5] = v;
    __syncthreads();
    const int l = threadIdx.x & 31;
    return warp_min_d(s.red[slot][l < E2_CELL_WARPS ? l : 0]);
}
__device__ __forceinline__ double cell_block_max(CellSmem &s, int slot, double v)
{
    v = warp_max_d(v);
    if ((threadIdx.x & 31) == 0) s.red[slot][threadIdx.x >> 5] = v;
    __syncthreads();
    const int l = threadIdx.x & 31;
    return warp_max_d(s.red[slot][l < E2_CELL_WARPS ? l : 0]);
}

// One warp finds the bin holding 0-based rank k of a weighted histogram (cumulative weight > k). Returns
// E2_CELL_BINS - 1 when the rank lies beyond the total (cannot happen: the weights sum to the pair count).
template <class T, int NB = E2_CELL_BINS>
__device__ __forceinline__ int cell_rank_bin(const T *h, u64 k, u64 *before_out = nullptr)
{
    const int lane = threadIdx.x & 31;
    const int per = NB / 32;
    static_assert(per <= 32, "one lane per bin in the second level");
    u64 mine = 0;
#pragma unroll 8
    for (int j = 0; j < per; ++j) mine += h[E2_CELL_HPAD(lane * per + j)];
    u64 incl = mine;
    for (int o = 1; o < 32; o <<= 1) { u64 t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const unsigned has = __ballot_sync(0xffffffffu, incl > k);
    if (!has) return NB - 1;
    // second level: the bins of the first lane whose running total passes k, one bin per lane
    const int L = __ffs(has) - 1;
    const u64 before = __shfl_sync(0xffffffffu, incl - mine, L);
    const u64 w = lane < per ? (u64)h[E2_CELL_HPAD(L * per + lane)] : 0ull;
    u64 inc2 = w;
    for (int o = 1; o < 32; o <<= 1) { u64 t = __shfl_up_sync(0xffffffffu, inc2, o); if (lane >= o) inc2 += t; }
    const unsigned has2 = __ballot_sync(0xffffffffu, lane < per && before + inc2 > k);
    const int J = has2 ? __ffs(has2) - 1 : per - 1;
    if (before_out && lane == J) *before_out = before + inc2 - w;      // weight strictly before the bin (written by one lane)
    return L * per + J;
}

// -DE2_CELL_PROFILE: thread 0 of every CTA adds the clock ticks of each phase to g_cell_prof (tools/cell_profile.py)
#ifdef E2_CELL_PROFILE
__device__ unsigned long long g_cell_prof[16];
#define CELL_TICK(slot) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_cell_prof[slot], (unsigned long long)(t_ - prof_t)); prof_t = t_; } } while (0)
#else
#define CELL_TICK(slot) do { } while (0)
#endif

__device__ __forceinline__ int median_bin(double v, double lo, double scale)
{
    const double t = (v - lo) * scale;                       // v >= lo: monotone non-decreasing in v
    return t >= (double)(E2_CELL_BINS - 1) ? E2_CELL_BINS - 1 : (int)t;
}

// np.median(objects[i], axis=0) (:1523) for the three axes at once, exact.  Each axis narrows a value range
// [m_lo, m_hi] that holds both middle order statistics (ranks k1 <= k2 = k1 or k1 + 1): histogram of the values in the
// range -> bins of the two ranks -> either few enough values to rank directly, or the bin's own value span becomes
// the next range, or the two ranks sit in different bins (then they are that bin's maximum / the other's minimum).
// Two passes over the points per round; planar clouds (one axis with a tiny spread next to far outliers) take two
// rounds on that axis, everything else one.  Also leaves the entity's AABB in s.ebox.  Returns false (after raising
// the device error flag) if a selection does not converge -- only non-finite coordinate spans can do that.
// pbin: one scratch word per point.  Round 0 leaves every point's three bins there (10 bits each); the resolve pass and any
// later round read that word first and touch a coordinate only where its bin is the one in question.
__device__ bool cell_axis_medians(CellSmem &s, const Dev &d, const double *X, const double *Y, const double *Z, unsigned *pbin, int n, double med[3])
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double *A[3] = {X, Y, Z};
#ifdef E2_CELL_PROFILE
    long long prof_t = clock64();
#endif
    {
        // the value range of every axis: the outward-rounded float32 box seg_emit_kernel left behind (Dev::ent_box), or a
        // pass over the points when the entity came another way
        const unsigned *eb = d.ent_box ? d.ent_box + (size_t)blockIdx.x * 6 : nullptr;
        double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
        if (!eb)
            for (int i = tid; i < n; i += E2_CELL_THREADS)
                for (int a = 0; a < 3; ++a) { const double v = A[a][i]; mn[a] = fmin(mn[a], v); mx[a] = fmax(mx[a], v); }
        for (int a = 0; a < 3; ++a) {
            const double l = eb ? (double)ord_f32(eb[a]) : cell_block_min(s, a, mn[a]);
            const double h = eb ? (double)ord_f32(eb[3 + a]) : cell_block_max(s, 3 + a, mx[a]);
            if (tid == 0) {
                s.ebox[a] = l; s.ebox[3 + a] = h;
                s.m_lo[a] = l; s.m_hi[a] = h; s.m_below[a] = 0; s.m_fb[a] = -1;
                if (h > l) { s.m_scale[a] = (double)E2_CELL_BINS / (h - l); s.m_mode[a] = E2_M_HIST; }
                else { s.m_a[a] = s.m_b[a] = l; s.m_mode[a] = E2_M_DONE; }
            }
        }
    }
    __syncthreads();
    CELL_TICK(9);
    const unsigned k2 = (unsigned)(n / 2), k1 = (n & 1) ? k2 : k2 - 1;
    for (int round = 0; round < 64; ++round) {
#ifdef E2_CELL_PROFILE
        if (tid == 0 && round > 0) atomicAdd(&g_cell_prof[15], 1ull);
#endif
        int mode[3];
        double lo[3], hi[3], sc[3];
        for (int a = 0; a < 3; ++a) { mode[a] = s.m_mode[a]; lo[a] = s.m_lo[a]; hi[a] = s.m_hi[a]; sc[a] = s.m_scale[a]; }
        if (mode[0] == E2_M_DONE && mode[1] == E2_M_DONE && mode[2] == E2_M_DONE) {
            for (int a = 0; a < 3; ++a) med[a] = (n & 1) ? s.m_a[a] : (s.m_a[a] + s.m_b[a]) / 2.0;     // numpy: mean of the two middle values
            return true;
        }
        for (int i = tid; i < 3 * E2_CELL_HSIZE; i += E2_CELL_THREADS) (&s.mhist[0][0])[i] = 0;
        if (tid < 3) s.m_ccnt[tid] = 0;
        __syncthreads();
        // ---- pass H: histograms (one shared atomic per distinct bin of a warp) ----------------------------------
        // (values spread over 1024 bins: plain shared atomics beat a match_any aggregation; the three loads of a point are
        // issued together, two points per thread in flight)
        int fb[3];
        for (int a = 0; a < 3; ++a) fb[a] = s.m_fb[a];
        if (round == 0) {
            for (int base = 0; base < n; base += 2 * E2_CELL_THREADS) {
                const int i0 = base + tid, i1 = i0 + E2_CELL_THREADS;
                double v0[3], v1[3];
#pragma unroll
                for (int a = 0; a < 3; ++a) { v0[a] = (i0 < n) ? A[a][i0] : 0.0; v1[a] = (i1 < n) ? A[a][i1] : 0.0; }
                unsigned p0 = 0, p1 = 0;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    if (mode[a] != E2_M_HIST) continue;             // a degenerate axis: bin 0 everywhere, never asked for
                    // (the range is the entity's own box: every value lies inside)
                    if (i0 < n) { const int b = median_bin(v0[a], lo[a], sc[a]); atomicAdd(&s.mhist[a][E2_CELL_HPAD(b)], 1u); p0 |= (unsigned)b << (10 * a); }
                    if (i1 < n) { const int b = median_bin(v1[a], lo[a], sc[a]); atomicAdd(&s.mhist[a][E2_CELL_HPAD(b)], 1u); p1 |= (unsigned)b << (10 * a); }
                }
                if (i0 < n) pbin[i0] = p0;
                if (i1 < n) pbin[i1] = p1;
            }
        } else {
            for (int i = tid; i < n; i += E2_CELL_THREADS) {
                const unsigned pk = pbin[i];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    if (mode[a] != E2_M_HIST || (int)((pk >> (10 * a)) & 1023u) != fb[a]) continue;
                    const double v = A[a][i];
                    if (v >= lo[a] && v <= hi[a]) { const int b = median_bin(v, lo[a], sc[a]); atomicAdd(&s.mhist[a][E2_CELL_HPAD(b)], 1u); }
                }
            }
        }
        __syncthreads();
        CELL_TICK(10);
        if (warp < 3 && mode[warp] == E2_M_HIST) {
            const int a = warp;
            const int b1 = cell_rank_bin(s.mhist[a], (u64)(k1 - s.m_below[a]), &s.m_bef64[a]);
            const int b2 = (k2 == k1) ? b1 : cell_rank_bin(s.mhist[a], (u64)(k2 - s.m_below[a]));
            if (lane == 0) { s.m_b1[a] = b1; s.m_b2[a] = b2; }
        }
        __syncthreads();
        if (tid < 3 && mode[tid] == E2_M_HIST) {
            const int a = tid, b1 = s.m_b1[a], b2 = s.m_b2[a];
            s.m_before[a] = (unsigned)s.m_bef64[a];                 // values of the range below bin b1
            s.m_mode[a] = (b1 != b2) ? E2_M_SPLIT : (s.mhist[a][E2_CELL_HPAD(b1)] <= E2_CELL_MCAP ? E2_M_COLLECT : E2_M_REFINE);
        }
        __syncthreads();
        CELL_TICK(11);
        // ---- pass R: resolve ---------------------------------------------------------------------------------------
        double tmin[3] = {1e300, 1e300, 1e300}, tmax[3] = {-1e300, -1e300, -1e300};
        int m2[3], b1[3], b2[3];
        for (int a = 0; a < 3; ++a) { m2[a] = (mode[a] == E2_M_HIST) ? s.m_mode[a] : E2_M_DONE; b1[a] = s.m_b1[a]; b2[a] = s.m_b2[a]; }
        for (int i = tid; i < n; i += E2_CELL_THREADS) {
            const unsigned pk = pbin[i];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                if (m2[a] == E2_M_DONE) continue;
                const int bin0 = (int)((pk >> (10 * a)) & 1023u);
                int bin;
                double v;
                if (round == 0) {                                 // the cached bins are this round's bins
                    if (bin0 != b1[a] && !(m2[a] == E2_M_SPLIT && bin0 == b2[a])) continue;
                    bin = bin0;
                    v = A[a][i];
                } else {
                    if (bin0 != fb[a]) continue;                  // outside the refined bin: outside [lo, hi]
                    v = A[a][i];
                    if (!(v >= lo[a] && v <= hi[a])) continue;
                    bin = median_bin(v, lo[a], sc[a]);
                }
                if (m2[a] == E2_M_COLLECT) {
                    // one counter bump per warp and axis: the middle bin of a flat face holds hundreds of raster neighbours,
                    // and a returning atomic on ONE shared word serialises them across the whole CTA
                    if (bin == b1[a]) {
                        const unsigned peers = __activemask();
                        const int leader = __ffs(peers) - 1;
                        unsigned base = 0;
                        if (lane == leader) base = atomicAdd(&s.m_ccnt[a], (unsigned)__popc(peers));
                        base = __shfl_sync(peers, base, leader);
                        s.mcand[a][base + __popc(peers & ((1u << lane) - 1u))] = v;
                    }
                }
                else if (m2[a] == E2_M_REFINE) { if (bin == b1[a]) { tmin[a] = fmin(tmin[a], v); tmax[a] = fmax(tmax[a], v); } }
                else { if (bin == b1[a]) tmax[a] = fmax(tmax[a], v); else if (bin == b2[a]) tmin[a] = fmin(tmin[a], v); }
            }
        }
        for (int a = 0; a < 3; ++a) {                            // uniform: m2 comes from shared memory
            if (m2[a] != E2_M_REFINE && m2[a] != E2_M_SPLIT) continue;
            const double l = cell_block_min(s, a, tmin[a]), h = cell_block_max(s, 3 + a, tmax[a]);
            if (tid == 0) {
                if (m2[a] == E2_M_SPLIT) { s.m_a[a] = h; s.m_b[a] = l; s.m_mode[a] = E2_M_DONE; }
                else {
                    if (round == 0) s.m_fb[a] = b1[a];
                    s.m_below[a] += s.m_before[a];
                    s.m_lo[a] = l; s.m_hi[a] = h;
                    if (h > l && (double)E2_CELL_BINS / (h - l) < 1e300) { s.m_scale[a] = (double)E2_CELL_BINS / (h - l); s.m_mode[a] = E2_M_HIST; }
                    else { s.m_a[a] = l; s.m_b[a] = h; s.m_mode[a] = E2_M_DONE; }      // a single value (or two adjacent ones)
                }
            }
        }
        __syncthreads();
        CELL_TICK(12);
        // ---- rank the collected values (shared memory only).  Up to E2_CELL_MSMALL of them directly; more go through 256
        // sub-bins of the bin first -- monotone in v like the bins themselves -- and only the sub-bin(s) of the two middle
        // ranks are ranked directly.
        auto subbin = [&](int a, double v) {
            const double t = ((v - lo[a]) * sc[a] - (double)b1[a]) * 256.0;
            return t >= 255.0 ? 255 : (t > 0.0 ? (int)t : 0);
        };
        bool big[3], any_big = false;
        unsigned cc[3], r1[3], r2[3];
        for (int a = 0; a < 3; ++a) {
            cc[a] = (m2[a] == E2_M_COLLECT) ? s.m_ccnt[a] : 0u;
            r1[a] = k1 - s.m_below[a] - s.m_before[a];
            r2[a] = k2 - s.m_below[a] - s.m_before[a];
            big[a] = cc[a] > E2_CELL_MSMALL;
            any_big |= big[a];
        }
        if (any_big) {                                           // uniform: everything above comes from shared memory
            __syncthreads();                                     // the bins' histogram has been read by everybody
            for (int a = 0; a < 3; ++a)
                if (big[a]) for (int i = tid; i < E2_CELL_HPAD(256) + 1; i += E2_CELL_THREADS) s.mhist[a][i] = 0;
            if (tid < 3) s.m_scnt[tid] = 0;
            __syncthreads();
            for (int a = 0; a < 3; ++a)
                if (big[a]) for (unsigned t = tid; t < cc[a]; t += E2_CELL_THREADS) atomicAdd(&s.mhist[a][E2_CELL_HPAD(subbin(a, s.mcand[a][t]))], 1u);
            __syncthreads();
            if (warp < 3 && big[warp]) {
                const int a = warp;
                u64 bef1 = 0, bef2 = 0;
                const int sb1 = cell_rank_bin<unsigned, 256>(s.mhist[a], (u64)r1[a], &s.m_bef64[a]);
                __syncwarp();
                bef1 = s.m_bef64[a];
                __syncwarp();
                const int sb2 = cell_rank_bin<unsigned, 256>(s.mhist[a], (u64)r2[a], &s.m_bef64[a]);
                __syncwarp();
                bef2 = s.m_bef64[a];
                if (lane == 0) { s.m_sb[a][0] = sb1; s.m_sb[a][1] = sb2; s.m_sbef[a][0] = (unsigned)bef1; s.m_sbef[a][1] = (unsigned)bef2; }
            }
            __syncthreads();
            for (int a = 0; a < 3; ++a)
                if (big[a]) for (unsigned t = tid; t < cc[a]; t += E2_CELL_THREADS) {
                    const double v = s.mcand[a][t];
                    const int sb = subbin(a, v);
                    if (sb == s.m_sb[a][0] || sb == s.m_sb[a][1]) {
                        const unsigned q = atomicAdd(&s.m_scnt[a], 1u);
                        if (q < E2_CELL_MSMALL) s.msmall[a][q] = v;
                    }
                }
            __syncthreads();
        }
        for (int a = 0; a < 3; ++a) {
            if (m2[a] != E2_M_COLLECT) continue;
            const unsigned c = cc[a];
            if (big[a] && s.m_scnt[a] <= E2_CELL_MSMALL) {
                const unsigned cs = s.m_scnt[a];
                for (unsigned t = tid; t < cs; t += E2_CELL_THREADS) {
                    const double v = s.msmall[a][t];
                    const int sb = subbin(a, v);
                    unsigned rank = s.m_sbef[a][sb == s.m_sb[a][0] ? 0 : 1];
                    for (unsigned j = 0; j < cs; ++j) { const double w = s.msmall[a][j]; rank += (subbin(a, w) == sb) && ((w < v) || (w == v && j < t)); }
                    if (rank == r1[a]) s.m_a[a] = v;
                    if (rank == r2[a]) s.m_b[a] = v;
                }
            } else {
                for (unsigned t = tid; t < c; t += E2_CELL_THREADS) {
                    const double v = s.mcand[a][t];
                    unsigned rank = 0;
                    for (unsigned j = 0; j < c; ++j) { const double w = s.mcand[a][j]; rank += (w < v) || (w == v && j < t); }
                    if (rank == r1[a]) s.m_a[a] = v;
                    if (rank == r2[a]) s.m_b[a] = v;
                }
            }
            if (tid == 0) s.m_mode[a] = E2_M_DONE;
        }
        __syncthreads();
        CELL_TICK(13);
    }
    if (tid == 0) atomicExch(d.error, E2SEC_ERR_STATE);
    return false;
}

// What an entity with a candidate point inside the band (r_lo, r_hi] leaves behind for cell_filter_kernel<true>: once the
// exact radius is known, the same point-level settlement finishes it -- no flood fill over the whole cloud.
struct CellState {
    unsigned char member[E2_CELL_MAX];
    unsigned ulist[E2_CELL_UMAX];
    double udist[E2_CELL_UMAX];      // distance^2 of every open candidate to the member set of the first visit
    unsigned n_u;
    int seed;
};

// Point-level settlement of the nU candidate points against the member POINTS (exact float64 distances): a candidate whose
// distance to a member is <= t_join becomes a member itself (virtual cell E2_CELL_JOIN); repeated until nothing changes,
// so s.udist ends up measured against the final member set.  Only the first round looks at the whole cloud: afterwards the
// member set grows by the candidates that have just joined, and udist = min(udist, distance to those).  HAVE_UDIST: udist
// already holds the distances to the current member set (the second visit of an entity, cell_filter_kernel<true>).
template <bool HAVE_UDIST>
__device__ __forceinline__ void cell_settle(CellSmem &s, const double *X, const double *Y, const double *Z, unsigned short *pcell,
                                            int n, unsigned nU, double t_join)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (!HAVE_UDIST) {
        for (unsigned u = warp; u < nU; u += E2_CELL_WARPS) {
            if (s.joined[u]) continue;
            const unsigned p = s.ulist[u];
            const double x = X[p], y = Y[p], z = Z[p];
            double best = 1e300;
            for (int q = lane; q < n; q += 32)
                if (s.member[pcell[q]]) best = fmin(best, dist2(x, y, z, X[q], Y[q], Z[q]));
            best = warp_min_nonneg_d(best);
            if (lane == 0) s.udist[u] = best;
        }
    }
    unsigned short *fresh = s.front[0];                 // the candidates that joined in this round (the closure's lists are free)
    for (;;) {
        if (tid == 0) s.n_new = 0;
        __syncthreads();
        for (unsigned u = tid; u < nU; u += E2_CELL_THREADS)
            if (!s.joined[u] && s.udist[u] <= t_join) {
                s.joined[u] = 1;
                pcell[s.ulist[u]] = (unsigned short)E2_CELL_JOIN;
                fresh[atomicAdd(&s.n_new, 1u)] = (unsigned short)u;
            }
        __syncthreads();
        const unsigned grew = s.n_new;
        if (!grew) break;
        for (unsigned u = tid; u < nU; u += E2_CELL_THREADS) {
            if (s.joined[u]) continue;
            const unsigned p = s.ulist[u];
            const double x = X[p], y = Y[p], z = Z[p];
            double best = s.udist[u];
            for (unsigned k = 0; k < grew; ++k) {
                const unsigned q = s.ulist[fresh[k]];
                best = fmin(best, dist2(x, y, z, X[q], Y[q], Z[q]));
            }
            s.udist[u] = best;
        }
        __syncthreads();
    }
}

extern __shared__ __align__(16) unsigned char cell_smem_raw[];


// grid F*K CTAs of E2_CELL_THREADS. Needs d.med (axis_median_kernel). Writes, for every filtered entity that it can
// decide: seed_pos, flist, n_list, aabb, mean, rad_lo / rad_hi, amb = 0; for the others amb = 1 (exact path follows).
// Entities that are not stored or have <= 3 points are finished here too (raw cloud, :1543-1546).
// FINISH = true is the second visit (after the exact radius selection): only the entities that left a CellState (amb == 2).
template <bool FINISH>
__device__ __forceinline__ void cell_filter_entity(const Dev &d, CellSmem &s, const int fe)
{
    const int f = fe / d.K;
    const int off = d.ent_off[fe], n = d.ent_raw[fe];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double NaN = __longlong_as_double(0x7ff8000000000000ll);
    if (!FINISH && tid == 0) { d.amb[fe] = 0; d.rad_lo[fe] = NaN; d.rad_hi[fe] = NaN; d.radius[fe] = NaN; }
    double *aabb = d.aabb + (size_t)fe * 6, *mean = d.mean + (size_t)fe * 3;
    if (!FINISH && (off < 0 || n == 0)) {
        if (tid == 0) { d.n_list[fe] = 0; d.seed_pos[fe] = -1; }
        if (tid < 6) aabb[tid] = 0.0;          // Open3D: an empty cloud has a zero box and a zero centre
        if (tid < 3) mean[tid] = 0.0;
        return;
    }
    const size_t pb = (size_t)f * d.cap + off;
    const double *X = d.px + pb, *Y = d.py + pb, *Z = d.pz + pb;
    unsigned *flist = d.flist + pb;
    unsigned short *pcell = d.pcell + pb;
    const bool filtered = entity_filtered(d, fe, n);               // :1520
#ifdef E2_CELL_PROFILE
    long long prof_t = clock64();
#endif

    if (FINISH) {
        // the cell closure of the first visit + the candidate points, settled at the exact float32 radius (pyx:35)
        const CellState &cs = d.cell_state[fe];
        for (int i = tid; i < E2_CELL_MAX; i += E2_CELL_THREADS) s.member[i] = cs.member[i];
        const unsigned nU = cs.n_u;
        for (unsigned u = tid; u < nU; u += E2_CELL_THREADS) { const unsigned p = cs.ulist[u]; s.ulist[u] = p; s.joined[u] = pcell[p] == E2_CELL_JOIN; s.udist[u] = cs.udist[u]; }
        if (tid == 0) { s.seed = cs.seed; s.overflow = 0; s.amb = 0; s.t_lo = sq_threshold((double)(float)d.radius[fe]); }
        __syncthreads();
        cell_settle<true>(s, X, Y, Z, pcell, n, nU, s.t_lo);
    }
    if (!FINISH && filtered) {
        // ---- A: seed, mean absolute deviation, AABB ------------------------------------------------------------
        // ---- M: per-axis medians = the seed target (:1523) -----------------------------------------------------------
        double med[3];
        if (!cell_axis_medians(s, d, X, Y, Z, flist, n, med)) return;      // flist is free until phase G
        if (tid < 3) d.med[(size_t)fe * 3 + tid] = med[tid];
        CELL_TICK(0);
        const double cx = med[0], cy = med[1], cz = med[2];
        float bd = 50000.0f;
        int bi = 0x7fffffff;
        double bd2_guard = 2.5e9 * 1.0000003;         // (float)sqrt(d2) < bd implies d2 < bd^2 * (1 + 2^-22)
        double ad[3] = {0.0, 0.0, 0.0};
        for (int i = tid; i < n; i += E2_CELL_THREADS) {
            const double x = X[i], y = Y[i], z = Z[i];
            const double d2 = dist2(x, y, z, cx, cy, cz);
            if (d2 <= bd2_guard) {                    // only then can float32(sqrt(d2)) be < the current minimum
                const float dt = (float)sqrt(d2);
                if (dt < bd) { bd = dt; bi = i; bd2_guard = (double)bd * (double)bd * 1.0000003; }
            }
            ad[0] += fabs(x - cx); ad[1] += fabs(y - cy); ad[2] += fabs(z - cz);
        }
        for (int o = 16; o; o >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, bd, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; }
        }
        if (lane == 0) { s.best_d[warp] = bd; s.best_i[warp] = bi; }
        double half[3];
        for (int k = 0; k < 3; ++k) {
            const double a = cell_block_sum(s, k, ad[k]);
            half[k] = 3.5 * (a / (double)n);
        }
        if (tid == 0) {
            for (int w = 1; w < E2_CELL_WARPS; ++w)
                if (s.best_d[w] < bd || (s.best_d[w] == bd && s.best_i[w] < bi)) { bd = s.best_d[w]; bi = s.best_i[w]; }
            s.seed = (bi == 0x7fffffff) ? 0 : bi;
            double c = fmax(fmax(half[0], half[1]), half[2]) * 2.0 / (double)E2_CELL_G;
            if (!(c > 1e-300) || !(c < 1e300)) c = 1.0;          // degenerate cloud: one cell
            s.inv_c = 1.0 / c;
            s.lo[0] = cx - half[0]; s.lo[1] = cy - half[1]; s.lo[2] = cz - half[2];
            for (int k = 0; k < 3; ++k) { s.ebk[k] = 0xffffffffu; s.ebk[3 + k] = 0u; }
            s.m = 0; s.overflow = 0; s.amb = 0;
        }
        for (int i = tid; i < E2_CELL_HASH; i += E2_CELL_THREADS) s.keys[i] = E2_CELL_EMPTY;
        __syncthreads();

        CELL_TICK(1);
        // ---- B1: insert the cell ids ------------------------------------------------------------------------------
        // (which cell a point lands in is free -- the boxes are built from the points themselves -- so the grid coordinate
        // is float32 arithmetic: a third of the instructions of the float64 floor / clamp)
        const double l0 = s.lo[0], l1 = s.lo[1], l2 = s.lo[2];
        const float inv_cf = (float)s.inv_c;
        for (int i = tid; i < n; i += E2_CELL_THREADS) {
            const float tx = fminf(fmaxf(floorf((float)(X[i] - l0) * inv_cf), -1.0f), (float)E2_CELL_G);
            const float ty = fminf(fmaxf(floorf((float)(Y[i] - l1) * inv_cf), -1.0f), (float)E2_CELL_G);
            const float tz = fminf(fmaxf(floorf((float)(Z[i] - l2) * inv_cf), -1.0f), (float)E2_CELL_G);
            const unsigned id = ((unsigned)((int)tx + 1) << 10) | ((unsigned)((int)ty + 1) << 5) | (unsigned)((int)tz + 1);
            unsigned slot = (id * 2654435761u) >> (32 - E2_CELL_HASH_LOG);
            int probes = 0;
            for (;;) {
                // a plain look first: almost every point finds its cell already there, and compare-and-swaps of a whole
                // face on one word would queue up
                unsigned cur = ((volatile unsigned *)s.keys)[slot];
                if (cur == E2_CELL_EMPTY) cur = atomicCAS(&s.keys[slot], E2_CELL_EMPTY, id);
                if (cur == E2_CELL_EMPTY || cur == id) break;
                slot = (slot + 1) & (E2_CELL_HASH - 1);
                if (++probes >= E2_CELL_HASH) { s.overflow = 1; break; }
            }
            pcell[i] = (unsigned short)slot;
        }
        __syncthreads();
        // ---- B2: compact indices in slot order ------------------------------------------------------------------
        {
            const int per = (E2_CELL_HASH + E2_CELL_THREADS - 1) / E2_CELL_THREADS;      // consecutive slots per thread
            const int j0 = tid * per, j1 = min(E2_CELL_HASH, j0 + per);
            unsigned occ = 0;
            for (int j = j0; j < j1; ++j) occ += (s.keys[j] != E2_CELL_EMPTY);
            unsigned incl = occ;
            for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) s.wcnt[warp] = incl;
            __syncthreads();
            unsigned base = incl - occ;
            for (int w = 0; w < warp; ++w) base += s.wcnt[w];
            for (int j = j0; j < j1; ++j)
                if (s.keys[j] != E2_CELL_EMPTY) { s.cidx[j] = (unsigned short)(base < E2_CELL_MAX ? base : 0); ++base; }
            if (tid == E2_CELL_THREADS - 1) { s.m = base; if (base > E2_CELL_JOIN) s.overflow = 1; }
        }
        for (int i = tid; i < E2_CELL_MAX; i += E2_CELL_THREADS) {
            s.cnt[i] = 0;
            s.member[i] = 0;
            for (int k = 0; k < 3; ++k) { s.box[k][i] = 0xffffffffu; s.box[3 + k][i] = 0u; }
        }
        __syncthreads();
    }

    bool decided = true;
    if (!FINISH && filtered && !s.overflow) {
        CELL_TICK(2);
        // ---- B3: per-cell count and outward-rounded float box ----------------------------------------------------
        for (int i = tid; i < 2 * E2_CELL_HSIZE; i += E2_CELL_THREADS) (&s.hist[0][0])[i] = 0;     // keys are dead now
        // one native shared atomic per quantity and point: the points of a warp (raster neighbours) fall into ~8 cells, a
        // 4-way conflict costs far fewer issue slots than the match / reduce loop it replaces (21 % of the kernel before)
        for (int i = tid; i < n; i += E2_CELL_THREADS) {
            const unsigned c = s.cidx[pcell[i]];
            pcell[i] = (unsigned short)c;
            const double x = X[i], y = Y[i], z = Z[i];
            atomicAdd(&s.cnt[c], 1u);
            atomicMin(&s.box[0][c], f32_ord(__double2float_rd(x)));
            atomicMin(&s.box[1][c], f32_ord(__double2float_rd(y)));
            atomicMin(&s.box[2][c], f32_ord(__double2float_rd(z)));
            atomicMax(&s.box[3][c], f32_ord(__double2float_ru(x)));
            atomicMax(&s.box[4][c], f32_ord(__double2float_ru(y)));
            atomicMax(&s.box[5][c], f32_ord(__double2float_ru(z)));
        }
        __syncthreads();
        const int m = (int)s.m;
        // (the union of the cell boxes = the entity's outward float32 box: the same numbers whichever way the entity got
        // here, so the bounds below do not depend on the segmentation path)
        for (int i = tid; i < m; i += E2_CELL_THREADS)
            for (int k = 0; k < 6; ++k) {
                const unsigned o = s.box[k][i];
                if (k < 3) atomicMin(&s.ebk[k], o); else atomicMax(&s.ebk[k], o);
                s.box[k][i] = __float_as_uint(ord_f32(o));
            }
        __syncthreads();
        if (tid == 0) {
            double ext2 = 0.0;
            for (int k = 0; k < 3; ++k) { const double e = (double)ord_f32(s.ebk[3 + k]) - (double)ord_f32(s.ebk[k]); ext2 += e * e; }
            double top = ext2 * 1.000001;                        // d2 of any pair <= squared AABB diagonal (+ slack)
            if (!(top > 0.0) || !(top < 1e300)) top = 1.0;
            s.d2max = top;
            s.scale = (double)E2_CELL_BINS / top;
        }
        __syncthreads();

        CELL_TICK(3);
        // ---- C: cell-pair histograms over d^2 ----------------------------------------------------------------------
        // Weights are pair counts: while the entity has fewer than 2^32 pairs the bins are 32-bit words (native shared
        // atomics; a 64-bit shared atomicAdd is a compare-and-swap loop and collapses under the contention of the
        // dmin2 = 0 bin), otherwise 64-bit.
        const double scale = s.scale;
        const u64 P = (u64)n * (u64)(n - 1) / 2;
        const bool small = P < 0x100000000ull;
        unsigned *h32 = reinterpret_cast<unsigned *>(&s.hist[0][0]);
        for (int a = warp; a < m; a += E2_CELL_WARPS) {
            const u64 ca = s.cnt[a];
            for (int b = a + lane; b < m; b += 32) {
                double dmin2, dmax2;
                cell_pair_d2(s, a, b, dmin2, dmax2);
                const u64 w = (a == b) ? ca * (ca - 1) / 2 : ca * (u64)s.cnt[b];
                if (a == b) dmin2 = 0.0;
                if (w) {
                    const int b0 = cell_bin(dmin2, scale), b1 = cell_bin(dmax2, scale);
                    if (small) { atomicAdd(&h32[E2_CELL_HPAD(b0)], (unsigned)w); atomicAdd(&h32[E2_CELL_HSIZE + E2_CELL_HPAD(b1)], (unsigned)w); }
                    else { atomicAdd(&s.hist[0][E2_CELL_HPAD(b0)], w); atomicAdd(&s.hist[1][E2_CELL_HPAD(b1)], w); }
                }
            }
        }
        __syncthreads();

        CELL_TICK(4);
        // ---- D: lo <= median <= hi ---------------------------------------------------------------------------------
        if (warp < 2) {
            const u64 k2 = P / 2, k1 = (P & 1) ? k2 : k2 - 1;
            const int B = small ? cell_rank_bin(h32 + warp * E2_CELL_HSIZE, warp == 0 ? k1 : k2)
                                : cell_rank_bin(s.hist[warp], warp == 0 ? k1 : k2);
            if (lane == 0) {
                if (warp == 0) {
                    double lo2 = (double)B / scale;
                    while (lo2 > 0.0 && cell_bin(lo2, scale) >= B) lo2 = next_down(lo2);      // bin(lo2) < B <= bin(true d2)
                    if (B == 0) lo2 = 0.0;
                    const double r = (double)(float)sqrt(lo2);
                    s.r_lo = r;
                    s.t_lo = sq_threshold(r);
                } else {
                    double hi2;
                    if (B >= E2_CELL_BINS - 1) hi2 = s.d2max;
                    else {
                        hi2 = (double)(B + 1) / scale;
                        while (cell_bin(hi2, scale) <= B) hi2 = next_up(hi2);                   // bin(hi2) > B >= bin(true d2)
                    }
                    const double r = (double)(float)sqrt(hi2);
                    s.r_hi = r;
                    s.t_hi = sq_threshold(r);
                }
            }
        }
        if (tid == 64) { s.seedp[0] = X[s.seed]; s.seedp[1] = Y[s.seed]; s.seedp[2] = Z[s.seed]; s.n_front = 0; s.n_new = 0; s.n_cand = 0; }
        __syncthreads();

        CELL_TICK(5);
        // ---- E: cell closure at r_lo ---------------------------------------------------------------------------------
        const double t_lo = s.t_lo, t_hi = s.t_hi;
        const float t_lo32 = __double2float_rd(t_lo), t_hi32 = __double2float_ru(t_hi);     // conservative float32 copies of the thresholds
        {
            const double sp[3] = {s.seedp[0], s.seedp[1], s.seedp[2]};
            for (int b = tid; b < m; b += E2_CELL_THREADS)
                if (cell_point_dmax2(s, sp, b) <= t_lo) { s.member[b] = 1; s.front[0][atomicAdd(&s.n_new, 1u)] = (unsigned short)b; }
        }
        __syncthreads();
        // later rounds: one warp per remaining cell, lanes over the frontier of the previous round
        if (tid == 0) s.n_rest = 0;
        __syncthreads();
        for (int b = tid; b < m; b += E2_CELL_THREADS)
            if (!s.member[b]) s.rest[atomicAdd(&s.n_rest, 1u)] = (unsigned short)b;
        __syncthreads();
        const unsigned n_rest = s.n_rest;
        int cur = 0;
        for (;;) {
            const unsigned nf = s.n_new;
            __syncthreads();
            if (nf == 0) break;
            if (tid == 0) s.n_new = 0;
            __syncthreads();
            for (unsigned r = warp; r < n_rest; r += E2_CELL_WARPS) {
                const int b = s.rest[r];
                if (s.member[b]) continue;
                bool hit = false;
                for (unsigned q = lane; q < nf && !hit; q += 32) hit = cell_pair_hi32(s, (int)s.front[cur][q], b) <= t_lo32;
                if (__any_sync(0xffffffffu, hit) && lane == 0) { s.member[b] = 1; s.front[cur ^ 1][atomicAdd(&s.n_new, 1u)] = (unsigned short)b; }
            }
            cur ^= 1;
            __syncthreads();
        }
        CELL_TICK(6);
        // ---- F: can anything else be reached at r_hi? --------------------------------------------------------------
        // Candidate cells: non-members within r_hi of a member or of another candidate (transitively).  Cells that are
        // not candidates cannot be reached at any radius <= r_hi, whatever happens to the candidates.
        if (tid == 0 && !s.member[pcell[s.seed]]) s.amb = 2;      // the seed's own cell must be certain
        for (int b = tid; b < E2_CELL_MAX; b += E2_CELL_THREADS) s.cand[b] = 0;
        __syncthreads();
        for (;;) {
            if (tid == 0) s.n_new = 0;
            __syncthreads();
            for (int b = warp; b < m; b += E2_CELL_WARPS) {
                if (s.member[b] || s.cand[b]) continue;
                bool hit = false;
                for (int a = lane; a < m && !hit; a += 32) {
                    if (!s.member[a] && !s.cand[a]) continue;
                    hit = cell_pair_lo32(s, a, b) <= t_hi32;
                }
                if (__any_sync(0xffffffffu, hit) && lane == 0) { s.cand[b] = 1; atomicAdd(&s.n_new, s.cnt[b]); atomicAdd(&s.n_cand, s.cnt[b]); }
            }
            __syncthreads();
            const unsigned grew = s.n_new;
            __syncthreads();
            if (!grew) break;
        }
        // ---- F2: settle the candidate points one by one against the member POINTS (exact float64 distances) ------
        const unsigned nU = s.n_cand;
        if (nU > E2_CELL_UMAX) { if (tid == 0) s.amb = 3; }
        else if (nU > 0 && !s.amb) {
            for (int i = tid; i < n; i += E2_CELL_THREADS)
                if (s.cand[pcell[i]]) s.ulist[atomicAdd(&s.n_front, 1u)] = (unsigned)i;
            for (int u = tid; u < E2_CELL_UMAX; u += E2_CELL_THREADS) s.joined[u] = 0;
            if (tid == 0) s.member[E2_CELL_JOIN] = 1;            // virtual cell of the points that join here
            __syncthreads();
            cell_settle<false>(s, X, Y, Z, pcell, n, nU, t_lo);    // reachable at every radius >= r_lo
            // the last round changed nothing: udist is measured against the final member set
            for (unsigned u = tid; u < nU; u += E2_CELL_THREADS)
                if (!s.joined[u] && s.udist[u] <= t_hi) s.amb = 4;      // reachable at r_hi but not at r_lo
        }
        __syncthreads();
        CELL_TICK(7);
        decided = !s.amb;
    } else if (!FINISH && filtered) {
        decided = false;                                           // too many cells for the table
    }
    if (!decided) {
        // a point inside the band is the only thing in the way: leave the closure so far for the second visit (amb = 2);
        // everything else (table overflow, too many candidates) goes to the point-level flood fill of filter.cuh (amb = 1)
        const bool resume = filtered && !s.overflow && s.amb == 4;
        if (resume) {
            CellState &cs = d.cell_state[fe];
            for (int i = tid; i < E2_CELL_MAX; i += E2_CELL_THREADS) cs.member[i] = s.member[i];
            for (unsigned u = tid; u < s.n_cand; u += E2_CELL_THREADS) { cs.ulist[u] = s.ulist[u]; cs.udist[u] = s.udist[u]; }
            if (tid == 0) { cs.n_u = s.n_cand; cs.seed = s.seed; }
        }
        if (tid == 0) {
            d.amb[fe] = resume ? 2 : 1;
            d.amb_list[atomicAdd(d.amb_count, 1)] = fe;
            // the bounds of phase D, when it was reached, are where the exact selection starts (radius.cuh)
            if (filtered && !s.overflow && s.amb >= 2) { d.rad_lo[fe] = s.r_lo; d.rad_hi[fe] = s.r_hi; }
#ifdef E2_CELL_DEBUG
            printf("amb fe=%d n=%d reason=%d (1 table overflow, 2 seed cell, 3 candidates > %d, 4 point in the band) m=%u n_cand=%u r_lo=%g r_hi=%g\n", fe, n,
                   filtered && s.overflow ? 1 : (int)s.amb, E2_CELL_UMAX, s.m, s.n_cand, s.r_lo, s.r_hi);
#endif
        }
        return;
    }

    // ---- G: stable compaction of the members + AABB / mean / count (unfiltered clouds: every point) ------------
    // Every warp owns a contiguous run of the entity's points: count its members, one block-wide exchange of the twelve
    // counts, then place them from the warp's own offset -- no barrier per 384 points.
    double sum[3] = {0.0, 0.0, 0.0}, mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    const int chunk = ((n + E2_CELL_WARPS - 1) / E2_CELL_WARPS + 31) & ~31;
    const int c0 = min(n, warp * chunk), c1 = min(n, c0 + chunk);
    unsigned mine = (unsigned)(c1 - c0);
    if (filtered) {
        mine = 0;
        for (int i = c0 + lane; i < c1; i += 32) mine += s.member[pcell[i]] ? 1u : 0u;
        mine = __reduce_add_sync(0xffffffffu, mine);
    }
    __syncthreads();                                   // s.wcnt may still be read by phase B2 of a slow warp
    if (lane == 0) s.wcnt[warp] = mine;
    __syncthreads();
    unsigned run = 0, before = 0;
    for (int w = 0; w < E2_CELL_WARPS; ++w) { const unsigned c = s.wcnt[w]; if (w < warp) before += c; run += c; }
    for (int base = c0; base < c1; base += 32) {
        const int i = base + lane;
        const bool mem = (i < c1) && (!filtered || s.member[pcell[i]]);
        const unsigned bal = __ballot_sync(0xffffffffu, mem);
        if (mem) {
            flist[before + __popc(bal & ((1u << lane) - 1u))] = (unsigned)i;
            const double x = X[i], y = Y[i], z = Z[i];
            sum[0] += x; sum[1] += y; sum[2] += z;
            mn[0] = fmin(mn[0], x); mx[0] = fmax(mx[0], x);
            mn[1] = fmin(mn[1], y); mx[1] = fmax(mx[1], y);
            mn[2] = fmin(mn[2], z); mx[2] = fmax(mx[2], z);
        }
        before += __popc(bal);
    }
    // nine block reductions behind ONE barrier: warp partials to shared memory, thread 0 folds them in warp order
    for (int k = 0; k < 3; ++k) {
        double v = sum[k];
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        const double lo_k = warp_min_d(mn[k]), hi_k = warp_max_d(mx[k]);
        if (lane == 0) { s.red[k][warp] = v; s.red[3 + k][warp] = lo_k; s.red[6 + k][warp] = hi_k; }
    }
    __syncthreads();
    CELL_TICK(8);
    if (tid == 0) {
        double tot[3], l[3], h[3];
        for (int k = 0; k < 3; ++k) {
            tot[k] = 0.0; l[k] = 1e300; h[k] = -1e300;
            for (int w = 0; w < E2_CELL_WARPS; ++w) { tot[k] += s.red[k][w]; l[k] = fmin(l[k], s.red[3 + k][w]); h[k] = fmax(h[k], s.red[6 + k][w]); }
        }
        d.n_list[fe] = (int)run;
        d.seed_pos[fe] = filtered ? s.seed : -1;
        double cntp = (double)run;
        if (filtered) {                                            // dicti[0] = first_point (pyx:48): the seed counts twice
            tot[0] += X[s.seed]; tot[1] += Y[s.seed]; tot[2] += Z[s.seed];
            cntp += 1.0;
            if (!FINISH) { d.rad_lo[fe] = s.r_lo; d.rad_hi[fe] = s.r_hi; }
        }
        for (int k = 0; k < 3; ++k) {
            aabb[k] = run ? l[k] : 0.0;
            aabb[3 + k] = run ? h[k] : 0.0;
            mean[k] = run ? tot[k] / cntp : 0.0;
        }
    }
}

// FINISH = false: grid F*K, one entity per CTA.  FINISH = true: any grid, striding over the exact path's list.
template <bool FINISH>
__global__ void __launch_bounds__(E2_CELL_THREADS, E2_CELL_CTAS) cell_filter_kernel(Dev d)
{
    pdl_enter();
    CellSmem &s = *reinterpret_cast<CellSmem *>(cell_smem_raw);
    if (!FINISH) {
        cell_filter_entity<false>(d, s, blockIdx.x);
        return;
    }
    if (exact_path_idle(d)) return;
    const int count = exact_count(d);
    for (int li = blockIdx.x; li < count; li += gridDim.x) {
        const int fe = exact_entity(d, li);
        if (d.amb[fe] == 2) cell_filter_entity<true>(d, s, fe);
        __syncthreads();
    }
}
