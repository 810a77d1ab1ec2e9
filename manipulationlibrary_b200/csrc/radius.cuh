// radius.cuh -- exact np.median(scipy.spatial.distance.pdist(objects[i])) (manipulation_analyser.py:1524)
// without materialising the n(n-1)/2 distances.
//
// Selection runs on the float64 SQUARED distance d2 = (dx*dx + dy*dy) + dz*dz (pdist's own expression
// before its sqrt; sqrt is monotone, so the k-th smallest distance is the sqrt of the k-th smallest d2).
// Keys are the bit patterns of d2 (non-negative doubles order like unsigned integers).
//
//   radius_init_kernel     per entity: pair count, the two middle ranks, an upper bound of d2 from the
//                          AABB, first key interval: the cell table's band [rad_lo^2, rad_hi^2] when there is one
//                          (BOUNDED mode), else [key(D2max * 2^-16), key(D2max)]
//   plan_scan              work list: one item = a slice of one 256 x 256 tile pair (ti <= tj) of an active entity.  Run
//                          by the LAST CTA to finish of radius_init_kernel / radius_select_kernel (no launch of its own)
//   pair_pass_kernel<0>    every pair of the item -> bin = (key - klo) >> shift, counted in thread-private
//                          uint16 shared-memory counters [bin][thread] (no atomics in the pair loop)
//   radius_select_kernel   picks the bin holding the lower middle rank, shrinks the interval 256x; when the
//                          interval holds <= E2_CAND_CAP pairs switches the entity to "collect"
//   pair_pass_kernel<1>    appends the keys inside the interval, tracks the smallest key above it
//   radius_final_kernel    block_select2 over the candidates -> (sqrt(a) + sqrt(b)) / 2
//
// The kernels walk the exact path's entity LIST (common.cuh: exact_count / exact_entity) with small fixed grids: in BOUNDED
// mode that is the handful of entities the cell table left open, not F x K CTAs that find out they have nothing to do.
#pragma once
#include "common.cuh"
#include "select.cuh"

#define E2_PAIR_THREADS 128     // 2 i-points per thread
#define E2_FLUSH_ITEMS 100      // uint16 counters: <= 512 increments per thread per item
#define E2_PAIR_MAX_SPLIT 8     // a tile pair is cut into up to 8 slices of its j range when there are few of them

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

// grid: any number of CTAs of 256 threads, striding over the list
__global__ void __launch_bounds__(256) radius_init_kernel(Dev d, PairPlan plan, u64 *below_acc, u64 *min_above, u64 *single_count, int grid_hint)
{
    pdl_enter();
    if (exact_path_idle(d)) return;
    __shared__ double red[6][8];
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
        if (threadIdx.x == 0) {
            double ext2 = 0.0;
            for (int a = 0; a < 3; ++a) {
                double l = red[a][0], h = red[3 + a][0];
                for (int w = 1; w < 8; ++w) { l = fmin(l, red[a][w]); h = fmax(h, red[3 + a][w]); }
                ext2 += (h - l) * (h - l);
            }
            u64 P = (u64)n * (u64)(n - 1) / 2;
            rs.k2 = P / 2;
            rs.k1 = (P & 1) ? rs.k2 : rs.k2 - 1;
            rs.n = n;
            rs.below = 0;
            if (!(ext2 > 0.0)) {           // every point identical: all distances are 0
                d.radius[fe] = 0.0;
                rs.phase = 2;
            } else {
                const double top = ext2 * 1.000001;  // d2 of any pair <= squared AABB diagonal (+ rounding slack)
                const u64 ktop = d2key(top) + 1;
                u64 khi = ktop;
                u64 klo = d2key(top * (1.0 / 65536.0));
                if (d.radius_mode == E2SEC_RADIUS_BOUNDED) {
                    // the cell table's bounds of the two middle distances (cells.cuh phase D; float32-rounded, hence the
                    // slack).  Should they ever miss, radius_select_kernel reopens the interval on the side that is short.
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
            }
        }
        __syncthreads();                   // red[] is reused by the next entity of this CTA
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

// All pairs (i, j) of one item for the two i-points of this thread, j over the shared-memory slice.  No data-dependent
// branch: a pair outside the interval counts into the row "below" or "rest" (keys under klo wrap around to >= 2^63), a
// pair that is not one (DIAG: j <= i on the diagonal tile; i beyond the entity: the caller gives it x = +inf, a key above
// every interval) into "rest".  The FP64 pipe -- 8 operations per pair -- is the floor; everything else has to fit under it.
template <bool COLLECT, bool DIAG>
__device__ __forceinline__ void pair_loop(const Dev &d, int fe, const double *sx, const double *sy, const double *sz, int jn, int jb,
                                          int gi0, int gi1, double x0, double y0, double z0, double x1, double y1, double z1,
                                          unsigned short *cnt_slot, PairLoopState &st)
{
    const u64 klo = st.klo, width = st.width;
    const int shift = st.shift;
#pragma unroll 2
    for (int j = 0; j < jn; ++j) {
        const double xj = sx[j], yj = sy[j], zj = sz[j];
        const int gj = jb + j;
        const u64 ka = d2key(dist2(x0, y0, z0, xj, yj, zj));
        const u64 kb = d2key(dist2(x1, y1, z1, xj, yj, zj));
        const u64 ta = ka - klo, tb = kb - klo;
        bool ha = ta < width, hb = tb < width;
        if (!COLLECT) {
            unsigned ra = ha ? (unsigned)(ta >> shift) : (unsigned)(E2_NBINS + 1) - (unsigned)(ta >> 63);
            unsigned rb = hb ? (unsigned)(tb >> shift) : (unsigned)(E2_NBINS + 1) - (unsigned)(tb >> 63);
            if (DIAG) { ra = (gj > gi0) ? ra : (unsigned)(E2_NBINS + 1); rb = (gj > gi1) ? rb : (unsigned)(E2_NBINS + 1); }
            cnt_slot[ra * E2_PAIR_THREADS] += 1;
            cnt_slot[rb * E2_PAIR_THREADS] += 1;
        } else {
            if (DIAG) { ha = ha && gj > gi0; hb = hb && gj > gi1; }
            // candidates are appended with one global atomic per warp (the j loop is uniform across the CTA)
            const unsigned ba = __ballot_sync(0xffffffffu, ha), bb = __ballot_sync(0xffffffffu, hb);
            if (ba | bb) {
                const int lane = threadIdx.x & 31;
                unsigned base = 0;
                if (lane == 0) base = atomicAdd(d.cand_n + fe, (unsigned)(__popc(ba) + __popc(bb)));
                base = __shfl_sync(0xffffffffu, base, 0);
                const unsigned lt = (1u << lane) - 1u;
                if (ha) { const unsigned pos = base + __popc(ba & lt); if (pos < E2_CAND_CAP) d.cand[(size_t)fe * E2_CAND_CAP + pos] = ka; }
                if (hb) { const unsigned pos = base + __popc(ba) + __popc(bb & lt); if (pos < E2_CAND_CAP) d.cand[(size_t)fe * E2_CAND_CAP + pos] = kb; }
            }
            u64 ca = ha ? ~0ull : ta, cb = hb ? ~0ull : tb;
            if (DIAG) { ca = (gj > gi0) ? ca : ~0ull; cb = (gj > gi1) ? cb : ~0ull; }
            const u64 c = ca < cb ? ca : cb;
            st.above_ta = c < st.above_ta ? c : st.above_ta;
        }
    }
}

// Shared layout: uint16 cnt[E2_PAIR_ROWS][E2_PAIR_THREADS] (histogram passes only) | double sx[256], sy[256], sz[256]
template <bool COLLECT>
__global__ void __launch_bounds__(E2_PAIR_THREADS) pair_pass_kernel(Dev d, PairPlan plan, u64 *below_acc, u64 *min_above)
{
    pdl_enter();
    if (exact_path_idle(d)) return;
    unsigned short *cnt = reinterpret_cast<unsigned short *>(pair_smem);
    double *sx = reinterpret_cast<double *>(pair_smem + (COLLECT ? 0 : E2_PAIR_CNT_BYTES));
    double *sy = sx + E2_TILE, *sz = sy + E2_TILE;
    __shared__ PairItem cur;
    const int tid = threadIdx.x;
    // counter slot of this thread inside a row: lanes of one warp hit 32 distinct 4-byte words (banks);
    // threads tid and tid + 64 share a word but own different halves of it
    const int slot = ((tid & 63) << 1) | (tid >> 6);
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
        const int gi0 = ib + tid, gi1 = ib + tid + E2_PAIR_THREADS;
        // an i beyond the entity sits at x = +inf: its d2 is +inf, above every interval, and counts nowhere
        double x0 = inf, y0 = 0, z0 = 0, x1 = inf, y1 = 0, z1 = 0;
        if (gi0 < n) { x0 = X[gi0]; y0 = Y[gi0]; z0 = Z[gi0]; }
        if (gi1 < n) { x1 = X[gi1]; y1 = Y[gi1]; z1 = Z[gi1]; }
        __syncthreads();
        if (it.ti == it.tj) pair_loop<COLLECT, true>(d, it.fe, sx, sy, sz, jn, jb, gi0, gi1, x0, y0, z0, x1, y1, z1, cnt + slot, st);
        else pair_loop<COLLECT, false>(d, it.fe, sx, sy, sz, jn, jb, gi0, gi1, x0, y0, z0, x1, y1, z1, cnt + slot, st);
    }
    if (cur_fe >= 0) flush(cur_fe);
}

// One warp per listed entity (lanes own 8 consecutive bins each), any number of CTAs of 256 threads striding over the
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
