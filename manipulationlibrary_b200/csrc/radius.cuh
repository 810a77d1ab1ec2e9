// radius.cuh -- exact np.median(scipy.spatial.distance.pdist(objects[i])) (manipulation_analyser.py:1524)
// without materialising the n(n-1)/2 distances.
//
// Selection runs on the float64 SQUARED distance d2 = (dx*dx + dy*dy) + dz*dz (pdist's own expression
// before its sqrt; sqrt is monotone, so the k-th smallest distance is the sqrt of the k-th smallest d2).
// Keys are the bit patterns of d2 (non-negative doubles order like unsigned integers).
//
//   radius_init_kernel     per entity: pair count, the two middle ranks, an upper bound of d2 from the
//                          AABB, first key interval: two order statistics of a fixed sample of 16384 pairs around the
//                          median, cut to the cell table's band [rad_lo^2, rad_hi^2] when there is one (BOUNDED mode)
//   plan_scan              work list: one item = a slice of one 256 x 256 tile pair (ti <= tj) of an active entity.  Run
//                          by the LAST CTA to finish of radius_init_kernel / radius_select_kernel (no launch of its own)
//   pair_pass_kernel<0>    every pair of the item -> bin = (key - klo) >> shift, counted in thread-private
//                          uint16 shared-memory counters [bin][thread] (no atomics in the pair loop)
//   radius_select_kernel   picks the bin holding the lower middle rank, shrinks the interval E2_NBINS x; when the
//                          interval holds <= E2_CAND_CAP pairs switches the entity to "collect"
//   pair_pass_kernel<1>    appends the keys inside the interval, tracks the smallest key above it
//   radius_final_kernel    block_select2 over the candidates -> (sqrt(a) + sqrt(b)) / 2
//
// The kernels walk the exact path's entity LIST (common.cuh: exact_count / exact_entity) with small fixed grids: in BOUNDED
// mode that is the handful of entities the cell table left open, not F x K CTAs that find out they have nothing to do.
#pragma once
#include "common.cuh"
#include "select.cuh"

#ifndef E2_PAIR_IPT
#define E2_PAIR_IPT 2           // i-points per thread
#endif
#define E2_PAIR_THREADS (E2_TILE / E2_PAIR_IPT)
#define E2_FLUSH_ITEMS (200 / E2_PAIR_IPT)      // uint16 counters: <= 256 * E2_PAIR_IPT increments per thread per item
#define E2_PAIR_MAX_SPLIT 8     // a tile pair is cut into up to 8 slices of its j range when there are few of them
#ifndef E2_PAIR_CTAS
#define E2_PAIR_CTAS 4          // CTAs of a histogram pass per SM (their counters: (E2_NBINS + 2) x 256 bytes each)
#endif
#ifndef E2_SAMPLE_SPAN
#define E2_SAMPLE_SPAN 1.5      // the sampled interval spans the ranks M / 2 -+ E2_SAMPLE_SPAN sqrt(M): 1.5 = 3 sigma of the
                                // median's rank (one entity in 370 reopens its interval and takes two or three cheap rounds more)
#endif

struct PairPlan {
    unsigned *off;      // count + 1 exclusive prefix of TILE PAIRS over the exact path's list
    unsigned *total;    // items = off[count] * split
    unsigned *split;    // slices per tile pair (1, 2, 4 or 8): enough items to occupy the grid
};

__device__ __forceinline__ int shift_for_width(u64 width)
{
    int s = 0;
    while (s < 63 && ((width - 1) >> s) >= (u64)E2_NBINS) ++s;
    return s;
}

// The calling CTA alone (any multiple of 32 threads): tile pairs of every listed entity in phase `want_phase`, exclusive
// scan.  RadiusState is read around L1 -- other CTAs of the same grid have just written it.
__device__ void plan_scan(const Dev &d, PairPlan plan, int want_phase, int grid_hint)
{
    __shared__ unsigned warp_tot[32];
    __shared__ unsigned carry;
    const int count = exact_count(d), nt = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < count; base += nt) {
        const int li = base + tid;
        unsigned items = 0;
        if (li < count) {
            const RadiusState *rs = d.rs + exact_entity(d, li);
            if (__ldcg(&rs->phase) == want_phase) {
                const unsigned T = (unsigned)((__ldcg(&rs->n) + E2_TILE - 1) / E2_TILE);
                items = T * (T + 1) / 2;
            }
        }
        unsigned v = items;
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
        if (lane == 31) warp_tot[warp] = v;
        __syncthreads();
        if (warp == 0) {
            unsigned w = (lane < (nt >> 5)) ? warp_tot[lane] : 0u;
            for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += t; }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const unsigned excl = carry + v - items + (warp ? warp_tot[warp - 1] : 0u);
        if (li < count) plan.off[li] = excl;
        __syncthreads();
        if (tid == nt - 1) carry = excl + items;
        __syncthreads();
    }
    if (tid == 0) {
        const unsigned pairs = carry;
        unsigned split = 1;
        while (split < E2_PAIR_MAX_SPLIT && pairs * split < 4u * (unsigned)grid_hint) split <<= 1;
        plan.off[count] = pairs;
        *plan.split = split;
        *plan.total = pairs * split;
    }
}

// true in exactly one CTA of the grid: the last one to get here.  Everything the other CTAs wrote before is visible to it.
__device__ __forceinline__ bool exact_last_cta(const Dev &d)
{
    __shared__ int last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(d.exact_ticket, 1u);
        last = (t == gridDim.x * gridDim.y - 1u);
        if (last) *d.exact_ticket = 0u;              // ready for the next kernel (stream order)
    }
    __syncthreads();
    if (last) __threadfence();
    return last != 0;
}

#define E2_SAMPLE_MAX 16384      // pairs sampled per entity to place the first interval (<= E2_CAND_CAP: they are parked in d.cand)

__device__ __forceinline__ unsigned mix32(unsigned x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// grid: any number of CTAs of 256 threads, striding over the list.
// Where the first histogram pass looks: a fixed pseudo-random sample of M pairs puts the median between two of its order
// statistics, M / 2 -+ 1.5 sqrt(M) (3 sigma of the median's rank in the sample): a fraction 3 / sqrt(M) of all pairs (2.3 %
// at M = 16384), so that one of the 96 .. 192 bins of that pass holds fewer than E2_CAND_CAP pairs up to ~7 * 10^7 pairs per
// entity -- the collect pass can follow at once.  In BOUNDED mode the interval is cut to the cell table's guaranteed band, and the band alone is used when it is
// narrow enough for the entity's pair count.  The sample only chooses where to look: the ranks are counted exactly, and
// radius_select_kernel reopens the interval on the side that is short should the guess ever miss.
// Entities with at most E2_CAND_CAP pairs skip the histogram altogether.
__global__ void __launch_bounds__(E2_SEL_THREADS) radius_init_kernel(Dev d, PairPlan plan, u64 *below_acc, u64 *min_above, u64 *single_count, int grid_hint)
{
    pdl_enter();
    if (exact_path_idle(d)) return;
    __shared__ double red[6][8];
    __shared__ SelectSmem sm;
    __shared__ int mode;                   // 0: settled here, otherwise the sample size
    const int count = exact_count(d);
    for (int li = blockIdx.x; li < count; li += gridDim.x) {
        const int fe = exact_entity(d, li), f = fe / d.K;
        const int off = d.ent_off[fe], n = d.ent_raw[fe];
        RadiusState &rs = d.rs[fe];
        if (threadIdx.x == 0) {
            d.radius[fe] = __longlong_as_double(0x7ff8000000000000ll);
            d.cand_n[fe] = 0;
            min_above[fe] = ~0ull;
            below_acc[fe] = 0;
            single_count[fe] = 0;
        }
        for (int b = threadIdx.x; b < E2_NBINS; b += blockDim.x) d.hist[(size_t)fe * E2_NBINS + b] = 0;
        if (off < 0 || !entity_filtered(d, fe, n)) {                 // nothing to select for this entity
            if (threadIdx.x == 0) { rs.phase = 3; rs.n = n; }
            continue;
        }
        const double *x = d.px + (size_t)f * d.cap + off, *y = d.py + (size_t)f * d.cap + off, *z = d.pz + (size_t)f * d.cap + off;
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            lo[0] = fmin(lo[0], x[i]); hi[0] = fmax(hi[0], x[i]);
            lo[1] = fmin(lo[1], y[i]); hi[1] = fmax(hi[1], y[i]);
            lo[2] = fmin(lo[2], z[i]); hi[2] = fmax(hi[2], z[i]);
        }
        for (int a = 0; a < 3; ++a) {
            lo[a] = warp_min_d(lo[a]);
            hi[a] = warp_max_d(hi[a]);
            if ((threadIdx.x & 31) == 0) { red[a][threadIdx.x >> 5] = lo[a]; red[3 + a][threadIdx.x >> 5] = hi[a]; }
        }
        __syncthreads();
        const u64 P = (u64)n * (u64)(n - 1) / 2;
        if (threadIdx.x == 0) {
            double ext2 = 0.0;
            for (int a = 0; a < 3; ++a) {
                double l = red[a][0], h = red[3 + a][0];
                for (int w = 1; w < 8; ++w) { l = fmin(l, red[a][w]); h = fmax(h, red[3 + a][w]); }
                ext2 += (h - l) * (h - l);
            }
            rs.k2 = P / 2;
            rs.k1 = (P & 1) ? rs.k2 : rs.k2 - 1;
            rs.n = n;
            rs.below = 0;
            mode = 0;
            if (!(ext2 > 0.0)) {           // every point identical: all distances are 0
                d.radius[fe] = 0.0;
                rs.phase = 2;
            } else {
                const double top = ext2 * 1.000001;  // d2 of any pair <= squared AABB diagonal (+ rounding slack)
                const u64 ktop = d2key(top) + 1;
                u64 khi = ktop;
                u64 klo = 0;                         // the sample below narrows it
                if (d.radius_mode == E2SEC_RADIUS_BOUNDED) {
                    // the cell table's bounds of the two middle distances (cells.cuh phase D; float32-rounded, hence the slack)
                    const double rl = d.rad_lo[fe], rh = d.rad_hi[fe];
                    if (rl >= 0.0 && rh >= rl && rh < 1e150) {
                        const u64 bl = d2key(rl * rl * (1.0 - 1.0 / 1048576.0));
                        const u64 bh = d2key(rh * rh * (1.0 + 1.0 / 1048576.0)) + 1;
                        if (bh < khi) khi = bh;
                        if (bl < khi) klo = bl;
                        if (klo >= khi) klo = 0;
                    }
                }
                rs.ktop = ktop;
                rs.klo = klo;
                rs.width = khi - klo;
                rs.shift = shift_for_width(rs.width);
                rs.phase = 0;
                if (P <= (u64)E2_CAND_CAP) {          // few pairs: collect them all, no histogram
                    rs.klo = 0;
                    rs.width = ktop;
                    rs.phase = 1;
                } else if (klo == 0 || P > 12000000ull) {
                    // (a band holds ~30 % of the pairs: below 1.2e7 pairs one of its bins usually fits the candidate buffer as it is, and
                    // a second histogram pass over such an entity costs less than the sample)
                    // sqrt(M) >= 4 P / (128 E2_CAND_CAP) would do (a pass uses 128 .. 256 of its bins), but a narrower
                    // interval also means fewer candidates for the collect pass to append: the largest sample that fits
                    // d.cand, except for small entities
                    const int M = P > 2000000ull ? E2_SAMPLE_MAX : 4096;
                    mode = M;
                }
            }
        }
        __syncthreads();                   // red[] is reused by the next entity of this CTA; mode is set
        if (mode) {
            const int M = mode, half = (int)ceil(E2_SAMPLE_SPAN * sqrt((double)M));
            u64 *smp = d.cand + (size_t)fe * E2_CAND_CAP;
            for (int t = threadIdx.x; t < M; t += blockDim.x) {
                const unsigned h1 = mix32((unsigned)t * 2u + 1u + (unsigned)fe * 0x9e3779b9u), h2 = mix32(h1 + 0x85ebca6bu + (unsigned)t);
                const int i = (int)(h1 % (unsigned)n);
                int j = (int)(h2 % (unsigned)(n - 1));
                j += (j >= i);
                smp[t] = d2key(dist2(x[i], y[i], z[i], x[j], y[j], z[j]));
            }
            __syncthreads();
            u64 qlo, qhi, unused;
            block_select2([&](int t) { return smp[t]; }, M, (u64)(M / 2 - half), (u64)(M / 2 - half), qlo, unused, sm);
            __syncthreads();
            block_select2([&](int t) { return smp[t]; }, M, (u64)(M / 2 + half), (u64)(M / 2 + half), qhi, unused, sm);
            if (threadIdx.x == 0) {
                const u64 klo = rs.klo, khi = rs.klo + rs.width;
                const u64 nlo = qlo > klo ? qlo : klo, nhi = (qhi + 1 < khi) ? qhi + 1 : khi;
                if (nlo < nhi) {
                    rs.klo = nlo;
                    rs.width = nhi - nlo;
                    rs.shift = shift_for_width(rs.width);
                }
                if (d.test_exact_start) {          // tests: a first interval that misses the median on purpose
                    const double top = key2d(rs.ktop - 1);
                    rs.klo = d.test_exact_start == 1 ? 0 : d2key(top * 0.9999);
                    rs.width = d.test_exact_start == 1 ? d2key(top * 1e-12) + 1 : rs.ktop - rs.klo;
                    rs.shift = shift_for_width(rs.width);
                }
            }
            __syncthreads();
        }
    }
    if (exact_last_cta(d)) plan_scan(d, plan, 0, grid_hint);
}

// single CTA: the plan for one phase on its own (operator calls that decide the number of rounds on the host)
__global__ void __launch_bounds__(1024) radius_plan_kernel(Dev d, PairPlan plan, int want_phase, int grid_hint)
{
    pdl_enter();
    if (exact_path_idle(d)) return;
    plan_scan(d, plan, want_phase, grid_hint);
}

struct PairItem { int fe, ti, tj, slice; };

// item index -> (entity, tile pair, slice of its j range). Row ti of the upper triangle holds tiles tj = ti .. T-1.
__device__ __forceinline__ PairItem decode_item(const Dev &d, const PairPlan &plan, unsigned item, int count, unsigned split)
{
    const unsigned pair = item / split;
    int lo = 0, hi = count;         // last list position with off[li] <= pair
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (plan.off[mid] <= pair) lo = mid; else hi = mid;
    }
    PairItem it;
    it.fe = exact_entity(d, lo);
    it.slice = (int)(item - pair * split);
    unsigned local = pair - plan.off[lo];
    int T = (d.rs[it.fe].n + E2_TILE - 1) / E2_TILE;
    int ti = 0;
    unsigned row = (unsigned)T;
    while (local >= row) { local -= row; --row; ++ti; }
    it.ti = ti;
    it.tj = ti + (int)local;
    return it;
}

extern __shared__ __align__(16) unsigned char pair_smem[];

#define E2_PAIR_ROWS (E2_NBINS + 2)     // the bins, then one row for keys under the interval and one for everything else
#define E2_PAIR_CNT_BYTES (sizeof(unsigned short) * E2_PAIR_ROWS * E2_PAIR_THREADS)
#define E2_PAIR_SMEM_HIST (E2_PAIR_CNT_BYTES + 3 * E2_TILE * sizeof(double))
#define E2_PAIR_SMEM_COLLECT (3 * E2_TILE * sizeof(double))

struct PairLoopState {
    u64 klo, width;
    int shift;
    u64 above_ta;       // COLLECT: smallest (key - klo) >= width seen so far
};

// All pairs (i, j) of one item for the E2_PAIR_IPT i-points of this thread, j over the shared-memory slice.  No
// data-dependent branch: a pair outside the interval counts into the row "below" or "rest" (keys under klo wrap around to
// >= 2^63), a pair that is not one (DIAG: j <= i on the diagonal tile; i beyond the entity: the caller gives it x = +inf, a
// key above every interval) into "rest".  The FP64 pipe -- 8 operations per pair -- is the floor; everything else has to
// fit under it.
template <bool COLLECT, bool DIAG>
__device__ __forceinline__ void pair_loop(const Dev &d, int fe, const double *sx, const double *sy, const double *sz, int jn, int jb,
                                          const int (&gi)[E2_PAIR_IPT], const double (&xi)[E2_PAIR_IPT], const double (&yi)[E2_PAIR_IPT],
                                          const double (&zi)[E2_PAIR_IPT], unsigned short *cnt_slot, PairLoopState &st)
{
    const u64 klo = st.klo, width = st.width;
    const int shift = st.shift;
#pragma unroll 2
    for (int j = 0; j < jn; ++j) {
        const double xj = sx[j], yj = sy[j], zj = sz[j];
        const int gj = jb + j;
        u64 k[E2_PAIR_IPT], t[E2_PAIR_IPT];
        bool h[E2_PAIR_IPT];
#pragma unroll
        for (int q = 0; q < E2_PAIR_IPT; ++q) {
            k[q] = d2key(dist2(xi[q], yi[q], zi[q], xj, yj, zj));
            t[q] = k[q] - klo;
            h[q] = t[q] < width;
        }
        if (!COLLECT) {
#pragma unroll
            for (int q = 0; q < E2_PAIR_IPT; ++q) {
                unsigned r = h[q] ? (unsigned)(t[q] >> shift) : (unsigned)(E2_NBINS + 1) - (unsigned)(t[q] >> 63);
                if (DIAG) r = (gj > gi[q]) ? r : (unsigned)(E2_NBINS + 1);
                cnt_slot[r * E2_PAIR_THREADS] += 1;
            }
        } else {
            unsigned bal[E2_PAIR_IPT], any = 0, tot = 0;
#pragma unroll
            for (int q = 0; q < E2_PAIR_IPT; ++q) {
                if (DIAG) h[q] = h[q] && gj > gi[q];
                bal[q] = __ballot_sync(0xffffffffu, h[q]);
                any |= bal[q];
                tot += __popc(bal[q]);
            }
            // candidates are appended with one global atomic per warp (the j loop is uniform across the CTA)
            if (any) {
                const int lane = threadIdx.x & 31;
                unsigned base = 0;
                if (lane == 0) base = atomicAdd(d.cand_n + fe, tot);
                base = __shfl_sync(0xffffffffu, base, 0);
                const unsigned lt = (1u << lane) - 1u;
#pragma unroll
                for (int q = 0; q < E2_PAIR_IPT; ++q) {
                    if (h[q]) { const unsigned pos = base + __popc(bal[q] & lt); if (pos < E2_CAND_CAP) d.cand[(size_t)fe * E2_CAND_CAP + pos] = k[q]; }
                    base += __popc(bal[q]);
                }
            }
#pragma unroll
            for (int q = 0; q < E2_PAIR_IPT; ++q) {
                u64 c = h[q] ? ~0ull : t[q];
                if (DIAG) c = (gj > gi[q]) ? c : ~0ull;
                st.above_ta = c < st.above_ta ? c : st.above_ta;
            }
        }
    }
}

// Shared layout: uint16 cnt[E2_PAIR_ROWS][E2_PAIR_THREADS] (histogram passes only) | double sx[256], sy[256], sz[256]
template <bool COLLECT>
__global__ void __launch_bounds__(E2_PAIR_THREADS) pair_pass_kernel(Dev d, PairPlan plan, u64 *below_acc, u64 *min_above)
{
    pdl_wait();                                    // launched in plain stream order (launch_plain); the next grid is released at the end
    if (exact_path_idle(d)) return;
    unsigned short *cnt = reinterpret_cast<unsigned short *>(pair_smem);
    double *sx = reinterpret_cast<double *>(pair_smem + (COLLECT ? 0 : E2_PAIR_CNT_BYTES));
    double *sy = sx + E2_TILE, *sz = sy + E2_TILE;
    __shared__ PairItem cur;
    const int tid = threadIdx.x;
    // counter slot of this thread inside a row: lanes of one warp hit 32 distinct 4-byte words (banks);
    // threads tid and tid + E2_PAIR_THREADS / 2 share a word but own different halves of it
    const int slot = ((tid % (E2_PAIR_THREADS / 2)) << 1) | (tid / (E2_PAIR_THREADS / 2));
    const unsigned total = *plan.total;
    if (total == 0) return;
    const int count = exact_count(d);
    const unsigned split = *plan.split;
    const int jw = E2_TILE / (int)split;           // j points per item
    const unsigned per = (total + gridDim.x - 1) / gridDim.x;
    const unsigned first = blockIdx.x * per;
    const unsigned last = min(total, first + per);
    if (first >= last) return;
    constexpr int CNT_WORDS = E2_PAIR_ROWS * E2_PAIR_THREADS / 2;
    if (!COLLECT)
        for (int i = tid; i < CNT_WORDS; i += E2_PAIR_THREADS) reinterpret_cast<unsigned *>(cnt)[i] = 0;
    int cur_fe = -1, since_flush = 0;
    PairLoopState st;
    st.klo = 0; st.width = 0; st.shift = 0; st.above_ta = ~0ull;

    auto flush = [&](int fe) {
        __syncthreads();
        if (!COLLECT) {
            for (int b = tid; b <= E2_NBINS; b += E2_PAIR_THREADS) {           // the bins and the row "below"
                const unsigned *row = reinterpret_cast<const unsigned *>(cnt + b * E2_PAIR_THREADS);
                unsigned s = 0;
                for (int w = 0; w < E2_PAIR_THREADS / 2; ++w) {
                    unsigned v = row[(w + tid) & (E2_PAIR_THREADS / 2 - 1)];
                    s += (v & 0xffffu) + (v >> 16);
                }
                if (s) {
                    if (b < E2_NBINS) atomicAdd(d.hist + (size_t)fe * E2_NBINS + b, (u64)s);
                    else atomicAdd(below_acc + fe, (u64)s);
                }
            }
            __syncthreads();
            for (int i = tid; i < CNT_WORDS; i += E2_PAIR_THREADS) reinterpret_cast<unsigned *>(cnt)[i] = 0;
        } else {
            const u64 m = warp_min_u64(st.above_ta);
            if ((tid & 31) == 0 && m < (1ull << 63)) atomicMin(min_above + fe, st.klo + m);
            st.above_ta = ~0ull;
        }
        __syncthreads();
        since_flush = 0;
    };

    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    for (unsigned item = first; item < last; ++item) {
        if (tid == 0) cur = decode_item(d, plan, item, count, split);
        __syncthreads();
        const PairItem it = cur;
        if (it.fe != cur_fe || since_flush >= E2_FLUSH_ITEMS) {
            if (cur_fe >= 0) flush(cur_fe);
            cur_fe = it.fe;
            const RadiusState &rs = d.rs[it.fe];
            st.klo = rs.klo; st.width = rs.width; st.shift = rs.shift;
        }
        ++since_flush;
        const int f = it.fe / d.K;
        const int n = d.rs[it.fe].n;
        const size_t pb = (size_t)f * d.cap + d.ent_off[it.fe];
        const double *X = d.px + pb, *Y = d.py + pb, *Z = d.pz + pb;
        const int jb = it.tj * E2_TILE + it.slice * jw, ib = it.ti * E2_TILE;
        const int jn = min(jw, n - jb);              // <= 0: the slice lies beyond the entity's last point
        __syncthreads();      // previous item's readers of sx/sy/sz are done
        for (int j = tid; j < jn; j += E2_PAIR_THREADS) { sx[j] = X[jb + j]; sy[j] = Y[jb + j]; sz[j] = Z[jb + j]; }
        // an i beyond the entity sits at x = +inf: its d2 is +inf, above every interval, and counts nowhere
        int gi[E2_PAIR_IPT];
        double xi[E2_PAIR_IPT], yi[E2_PAIR_IPT], zi[E2_PAIR_IPT];
#pragma unroll
        for (int q = 0; q < E2_PAIR_IPT; ++q) {
            gi[q] = ib + tid + q * E2_PAIR_THREADS;
            xi[q] = inf; yi[q] = 0; zi[q] = 0;
            if (gi[q] < n) { xi[q] = X[gi[q]]; yi[q] = Y[gi[q]]; zi[q] = Z[gi[q]]; }
        }
        __syncthreads();
        if (it.ti == it.tj) pair_loop<COLLECT, true>(d, it.fe, sx, sy, sz, jn, jb, gi, xi, yi, zi, cnt + slot, st);
        else pair_loop<COLLECT, false>(d, it.fe, sx, sy, sz, jn, jb, gi, xi, yi, zi, cnt + slot, st);
    }
    pdl_release();
    if (cur_fe >= 0) flush(cur_fe);
}

// One warp per listed entity (lanes own E2_NBINS / 32 consecutive bins each), any number of CTAs of 256 threads striding over the
// list; the last CTA to finish plans the next pass (the items of phase `next_phase`).
__global__ void __launch_bounds__(256) radius_select_kernel(Dev d, PairPlan plan, u64 *below_acc, u64 *single_count, int next_phase, int grid_hint)
{
    pdl_enter();
    if (exact_path_idle(d)) return;
    const int count = exact_count(d), lane = threadIdx.x & 31;
    const int warps = (blockDim.x >> 5) * gridDim.x;
    constexpr int PER = E2_NBINS / 32;
    for (int li = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); li < count; li += warps) {
        const int fe = exact_entity(d, li);
        RadiusState &rs = d.rs[fe];
        if (rs.phase != 0) continue;
        u64 *h = d.hist + (size_t)fe * E2_NBINS;
        u64 hb[PER], mine = 0;
        for (int k = 0; k < PER; ++k) { hb[k] = h[lane * PER + k]; mine += hb[k]; h[lane * PER + k] = 0; }
        u64 incl = mine;
        for (int o = 1; o < 32; o <<= 1) { const u64 t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        const u64 total = __shfl_sync(0xffffffffu, incl, 31);
        const u64 below = below_acc[fe];
        const u64 r1 = rs.k1;
        // the first bin whose cumulative count exceeds the rank (the last bin when none does: cannot happen once
        // r1 < below + total holds)
        const bool holds = r1 >= below && r1 < below + incl;
        const unsigned who = __ballot_sync(0xffffffffu, holds);
        __syncwarp();
        if (lane == 0) below_acc[fe] = 0;
        if (r1 < below) {                         // the rank lies under the interval: reopen downwards
            if (lane == 0) {
                rs.width += rs.klo;
                rs.klo = 0;
                rs.shift = shift_for_width(rs.width);
            }
        } else if (r1 >= below + total) {
            if (lane == 0) {
                const u64 end = rs.klo + rs.width;
                if (end < rs.ktop) {              // the rank lies above an interval that started from bounds: reopen upwards
                    rs.klo = end;
                    rs.width = rs.ktop - end;
                    rs.shift = shift_for_width(rs.width);
                } else {                          // cannot happen while the interval reaches the d2 upper bound
                    atomicExch(d.error, E2SEC_ERR_STATE);
                    rs.phase = 2;
                }
            }
        } else {
            const int owner = __ffs(who) - 1;
            if (lane == owner) {
                u64 acc = below + incl - mine;
                int k = 0;
                for (; k < PER - 1; ++k) {
                    if (r1 < acc + hb[k]) break;
                    acc += hb[k];
                }
                const int b1 = lane * PER + k;
                const u64 in_bin = hb[k];
                u64 nklo = rs.klo + ((u64)b1 << rs.shift);
                u64 nend = rs.klo + ((u64)(b1 + 1) << rs.shift);
                if (nend > rs.klo + rs.width || nend < nklo) nend = rs.klo + rs.width;
                rs.klo = nklo;
                rs.width = nend - nklo;
                rs.below = acc;
                if (rs.width == 1) {                  // a single key: every pair of the bin has exactly this d2
                    single_count[fe] = in_bin;
                    rs.phase = 1;
                } else if (in_bin <= (u64)E2_CAND_CAP) {
                    single_count[fe] = 0;
                    rs.phase = 1;
                } else {
                    rs.shift = shift_for_width(rs.width);
                }
            }
        }
    }
    if (exact_last_cta(d)) plan_scan(d, plan, next_phase, grid_hint);
}

// any number of CTAs of E2_SEL_THREADS, striding over the list
__global__ void __launch_bounds__(E2_SEL_THREADS) radius_final_kernel(Dev d, const u64 *min_above, const u64 *single_count)
{
    pdl_enter();
    if (exact_path_idle(d)) return;
    __shared__ SelectSmem sm;
    const int count = exact_count(d);
    for (int li = blockIdx.x; li < count; li += gridDim.x) {
        const int fe = exact_entity(d, li);
        RadiusState &rs = d.rs[fe];
        if (rs.phase == 0 && threadIdx.x == 0) atomicExch(d.error, E2SEC_ERR_STATE);   // the launched refinement rounds did not converge
        if (rs.phase != 1) continue;
        const u64 r1 = rs.k1 - rs.below, r2 = rs.k2 - rs.below;
        u64 v1, v2;
        const u64 single = single_count[fe];
        if (single) {
            v1 = rs.klo;
            v2 = (r2 < single) ? rs.klo : min_above[fe];
        } else {
            const unsigned m = min(d.cand_n[fe], (unsigned)E2_CAND_CAP);
            const u64 *c = d.cand + (size_t)fe * E2_CAND_CAP;
            if (r2 < m) {
                block_select2([&](int i) { return c[i]; }, (int)m, r1, r2, v1, v2, sm);
            } else {
                block_select2([&](int i) { return c[i]; }, (int)m, r1, r1, v1, v2, sm);
                v2 = min_above[fe];
            }
        }
        __syncthreads();                          // every thread has read rs before thread 0 moves it on
        if (threadIdx.x == 0) {
            double a = sqrt(key2d(v1)), b = sqrt(key2d(v2));
            d.radius[fe] = (rs.k1 == rs.k2) ? a : (a + b) / 2.0;
            rs.phase = 2;
        }
        __syncthreads();
    }
}
