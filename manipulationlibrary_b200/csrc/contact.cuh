// contact.cuh -- all-pairs touching distances, work-list version (replaces one-CTA-per-pair for the per-frame pass).
//
//   np.min(a.compute_point_cloud_distance(b)) (manipulation_analyser.py:397-484 T/N rows, :1570-1653 o1/o2/o3 search)
//   for every pair of stored entities of a frame, every (entity, ground) pair and every cross pair (entity a of the
//   previous kept frame, entity b of this frame).
//
//   contact_plan_kernel  one thread per (frame, pair slot): empty clouds -> +inf; boxes further apart than thresh -> the
//                        box gap (a value >= thresh, which is all the relation rows need); otherwise the slot is
//                        initialised to max(gap, thresh)^2 and gets ceil(n_a / E2_CT_CHUNK) work items
//   contact_scan_kernel  exclusive prefix of the item counts (single CTA)
//   contact_work_kernel  persistent CTAs pull items from an atomic counter.  An item = 1024 consecutive points of A:
//                        those within thresh of B's box are compacted and get their own tight box; B's points within
//                        thresh of THAT box are staged tile by tile as float32 coordinates relative to the box centre.
//                        The O(n_a * n_b) sweep runs in float32 with 4 A points per thread; a pair is re-evaluated in
//                        float64 (the reference's own expression) only when its float32 d^2 is within a proven guard
//                        band of the thread's current minimum, so the result is the exact float64 minimum.
//                        Results merge with atomicMin on the float64 bit pattern (order-preserving for d^2 >= 0).
//   In-frame pairs report the exact minimum whenever it is < thresh.  Cross pairs only feed `< thresh` tests of the
//   o2 / o3 search (:1600-1653), so they stop at the first witness and the slot is a contact flag: 0.0 iff some
//   pair of points is closer than thresh, otherwise a value >= thresh (deterministic whatever the item order).
#pragma once
#include "common.cuh"
#include "pairs.cuh"

#define E2_CT_THREADS 256
#define E2_CT_CHUNK 1024          // A points per work item (4 per thread)
#define E2_CT_TILE 1024           // B entries staged per tile
#define E2_CT_RUNS 1024           // ground runs listed per item; a larger ground cloud is scanned tile by tile
static_assert(E2_CT_SUB == E2_CT_THREADS, "a thread slice of a B tile must be exactly one ground pruning run");

struct ContactPlan {
    unsigned *cnt;        // F * slots      items of the slot (0 = decided by the plan kernel)
    unsigned *off;        // F * slots + 1  exclusive prefix
    unsigned *next;       // work counter
    int slots;            // pair slots per frame: K(K-1)/2 + K + K*K
    unsigned *map;        // item -> index of its (frame, slot), filled by the scan kernel while the items fit map_cap
    unsigned map_cap;
};

struct ContactSlot { int a, b, fa; bool cross; u64 *out; };

__device__ __forceinline__ ContactSlot contact_decode(const Dev &d, int f, int slot)
{
    const int K = d.K, npairs = K * (K - 1) / 2;
    ContactSlot s;
    s.fa = f;
    s.cross = false;
    if (slot < npairs) {
        int a = 0, row = K - 1;
        while (slot >= row) { slot -= row; --row; ++a; }
        s.a = a;
        s.b = a + 1 + slot;
        s.out = d.pair_key + ((size_t)f * (K + 1) + s.a) * (K + 1) + s.b;
    } else if (slot < npairs + K) {
        s.a = slot - npairs;
        s.b = K;
        s.out = d.pair_key + ((size_t)f * (K + 1) + s.a) * (K + 1) + K;
    } else {
        slot -= npairs + K;
        s.a = slot / K;
        s.b = slot - s.a * K;
        s.out = d.cross_key + ((size_t)f * K + s.a) * K + s.b;
        s.fa = d.prev_slot[f];
        s.cross = true;
    }
    return s;
}

__global__ void contact_plan_kernel(Dev d, ContactPlan plan)
{
    pdl_enter();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx == 0) *plan.next = 0;
    if (idx >= d.F * plan.slots) return;
    const int f = idx / plan.slots, slot = idx - f * plan.slots, K = d.K;
    const u64 INF = 0x7ff0000000000000ull;
    const ContactSlot s = contact_decode(d, f, slot);
    unsigned items = 0;
    u64 init = INF;
    if (!(s.cross && (s.fa < 0 || s.a == s.b))) {
        const int offa = d.ent_off[s.fa * K + s.a];
        const int na = (offa < 0) ? 0 : d.n_list[s.fa * K + s.a];
        int nb;
        if (s.b == K) nb = d.n_ground;
        else { const int offb = d.ent_off[f * K + s.b]; nb = (offb < 0) ? 0 : d.n_list[f * K + s.b]; }
        if (na > 0 && nb > 0) {
            const double *ba = d.aabb + (size_t)(s.fa * K + s.a) * 6;
            const double *bb = (s.b == K) ? d.gaabb : d.aabb + (size_t)(f * K + s.b) * 6;
            const double lim2 = d.thresh * d.thresh * 1.000001 + 1e-300;
            const double gap2 = box_box_d2(ba, bb);
            init = d2key(fmax(gap2, d.thresh * d.thresh));
            if (!(gap2 > lim2)) items = 1u;          // one work item per pair: its chunks of A run one after the other
        }
    }
    *s.out = init;
    plan.cnt[idx] = items;
}

// single CTA of 1024 threads: exclusive scan of plan.cnt -> plan.off, 4096 entries per round (one coalesced uint4 per
// thread), then the item -> entry map the work kernel reads instead of searching plan.off
__global__ void __launch_bounds__(1024) contact_scan_kernel(ContactPlan plan, int n)
{
    pdl_enter();
    __shared__ unsigned warp_tot[32];
    __shared__ unsigned carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 4096) {
        const int i0 = base + 4 * tid;
        unsigned c[4] = {0, 0, 0, 0};
        if (i0 + 3 < n) { const uint4 q = *reinterpret_cast<const uint4 *>(plan.cnt + i0); c[0] = q.x; c[1] = q.y; c[2] = q.z; c[3] = q.w; }
        else for (int k = 0; k < 4; ++k) if (i0 + k < n) c[k] = plan.cnt[i0 + k];
        const unsigned mine = c[0] + c[1] + c[2] + c[3];
        unsigned v = mine;
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
        if (lane == 31) warp_tot[warp] = v;
        __syncthreads();
        if (warp == 0) {
            unsigned w = warp_tot[lane];
            for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += t; }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const unsigned carry = carry_s;
        unsigned off = carry + v - mine + (warp ? warp_tot[warp - 1] : 0u);
        for (int k = 0; k < 4; ++k)
            if (i0 + k < n) {
                plan.off[i0 + k] = off;
                for (unsigned j = 0; j < c[k]; ++j)
                    if (off + j < plan.map_cap) plan.map[off + j] = (unsigned)(i0 + k);
                off += c[k];
            }
        const unsigned block_total = warp_tot[31];
        __syncthreads();
        if (tid == 0) carry_s = carry + block_total;
        __syncthreads();
    }
    if (tid == 0) plan.off[n] = carry_s;
}

struct ContactSmem {
    double dx[E2_CT_TILE], dy[E2_CT_TILE], dz[E2_CT_TILE];     // compacted candidates (A chunk, then B tile), float64
    float fx[E2_CT_TILE], fy[E2_CT_TILE], fz[E2_CT_TILE];      // B tile relative to the chunk box centre, float32
    double red[6][E2_CT_THREADS / 32];
    unsigned wcnt[E2_CT_THREADS / 32];
    unsigned item, skip;
    double box[6];
    unsigned n_run;                                            // ground runs within thresh of the item's box ...
    unsigned short run[E2_CT_RUNS];                            // ... and their indices (contact_work_kernel)
};

// float32 acceptance bound for a float64 minimum `best` <= lim2: every pair whose float64 d^2 is < best has a float32 d^2
// (coordinates rounded to float32 relative to an origin, |coordinate| <= M, e = M * 2^-22) below
//   best * (1 + 1e-5) + 4 e sqrt(best) + 12 e^2  <=  best * 1.00001 + c0,   c0 = 4 e sqrt(lim2) + 12 e^2
__device__ __forceinline__ float contact_guard(double best, double c0)
{
    return __double2float_ru(best * 1.00001 + c0);
}

// Block-wide stable-enough compaction step: thread contributes (x, y, z) when `take`; returns the new count.
__device__ __forceinline__ unsigned contact_push(ContactSmem &s, bool take, double x, double y, double z, unsigned count,
                                                 bool with_f32, double ox, double oy, double oz)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned bal = __ballot_sync(0xffffffffu, take);
    if (lane == 0) s.wcnt[warp] = __popc(bal);
    __syncthreads();
    unsigned before = count, tot = 0;
    for (int w = 0; w < E2_CT_THREADS / 32; ++w) { const unsigned c = s.wcnt[w]; if (w < warp) before += c; tot += c; }
    if (take) {
        const unsigned p = before + __popc(bal & ((1u << lane) - 1u));
        s.dx[p] = x; s.dy[p] = y; s.dz[p] = z;
        if (with_f32) { s.fx[p] = (float)(x - ox); s.fy[p] = (float)(y - oy); s.fz[p] = (float)(z - oz); }
    }
    __syncthreads();
    return count + tot;
}

// The O(n_a * m_b) sweep of one staged B tile: NK compacted A points per thread in float32, entries jstart, jstart + jstep,
// ... of the tile; a pair is re-evaluated in float64 only inside the guard band of the thread's current minimum.
template <int NK>
__device__ __forceinline__ void contact_sweep(const ContactSmem &s, unsigned mB, unsigned jstart, unsigned jstep, const float *rx,
                                              const float *ry, const float *rz, const double *ax, const double *ay,
                                              const double *az, double &best, float &guard, double c0)
{
    for (unsigned j = jstart; j < mB; j += jstep) {
        const float bx = s.fx[j], by = s.fy[j], bz = s.fz[j];
        float d32[NK], lo = 3.0e38f;
#pragma unroll
        for (int k = 0; k < NK; ++k) {
            const float ux = rx[k] - bx, uy = ry[k] - by, uz = rz[k] - bz;
            d32[k] = fmaf(uz, uz, fmaf(uy, uy, ux * ux));
            lo = fminf(lo, d32[k]);
        }
        // the loop is warp-uniform (same j in every lane): the lanes share their minimum whenever one of them improves it,
        // so the guard follows the WARP's best pair and a "record" is a rare event (per-lane minima would send some lane of
        // the warp into the float64 recheck at a third of the entries)
        if (__any_sync(0xffffffffu, lo <= guard)) {
#pragma unroll
            for (int k = 0; k < NK; ++k)
                if (d32[k] <= guard) best = fmin(best, dist2(ax[k], ay[k], az[k], s.dx[j], s.dy[j], s.dz[j]));
            best = warp_min_nonneg_d(best);
            guard = contact_guard(best, c0);
        }
    }
}

// Four candidates per thread at once: all loads of a 1024-point tile are in flight together and the tile costs one
// pair of block barriers instead of four.  Order inside the tile is warp-major (irrelevant for a minimum).
__device__ __forceinline__ unsigned contact_push4(ContactSmem &s, const bool take[4], const double x[4], const double y[4],
                                                  const double z[4], bool with_f32, double ox, double oy, double oz)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned bal[4], wtot = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { bal[k] = __ballot_sync(0xffffffffu, take[k]); wtot += __popc(bal[k]); }
    if (lane == 0) s.wcnt[warp] = wtot;
    __syncthreads();
    unsigned before = 0, tot = 0;
    for (int w = 0; w < E2_CT_THREADS / 32; ++w) { const unsigned c = s.wcnt[w]; if (w < warp) before += c; tot += c; }
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (take[k]) {
            const unsigned p = before + __popc(bal[k] & lt);
            s.dx[p] = x[k]; s.dy[p] = y[k]; s.dz[p] = z[k];
            if (with_f32) { s.fx[p] = (float)(x[k] - ox); s.fy[p] = (float)(y[k] - oy); s.fz[p] = (float)(z[k] - oz); }
        }
        before += __popc(bal[k]);
    }
    __syncthreads();
    return tot;
}

__global__ void __launch_bounds__(E2_CT_THREADS, 3) contact_work_kernel(Dev d, ContactPlan plan)
{
    pdl_enter();
    __shared__ ContactSmem s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, K = d.K;
    const int n_idx = d.F * plan.slots;
    const unsigned total = plan.off[n_idx];
    const double lim2 = d.thresh * d.thresh * 1.000001 + 1e-300;            // anything with sqrt(d2) < thresh is below this
    for (;;) {
        __syncthreads();
        if (tid == 0) s.item = atomicAdd(plan.next, 1u);
        __syncthreads();
        const unsigned item = s.item;
        if (item >= total) break;
        int lo = 0, hi = n_idx;                               // last idx with off[idx] <= item
        if (total <= plan.map_cap) lo = (int)plan.map[item];
        else
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (plan.off[mid] <= item) lo = mid; else hi = mid;
            }
        const int f = lo / plan.slots;
        const ContactSlot cs = contact_decode(d, f, lo - f * plan.slots);
        const int offa = d.ent_off[cs.fa * K + cs.a], na = d.n_list[cs.fa * K + cs.a];
        const size_t pa = (size_t)cs.fa * d.cap + offa;
        const double *AX = d.px + pa, *AY = d.py + pa, *AZ = d.pz + pa;
        const unsigned *fla = d.flist + pa;
        const double *BX, *BY, *BZ;
        const unsigned *flb = nullptr;
        const double *gsub = nullptr;
        int nb;
        double bb[6];
        if (cs.b == K) {
            BX = d.gx; BY = d.gy; BZ = d.gz; nb = d.n_ground;
            gsub = d.gsub;
            for (int k = 0; k < 6; ++k) bb[k] = d.gaabb[k];
        } else {
            const int offb = d.ent_off[f * K + cs.b];
            const size_t pbb = (size_t)f * d.cap + offb;
            BX = d.px + pbb; BY = d.py + pbb; BZ = d.pz + pbb;
            flb = d.flist + pbb;
            nb = d.n_list[f * K + cs.b];
            for (int k = 0; k < 6; ++k) bb[k] = d.aabb[(size_t)(f * K + cs.b) * 6 + k];
        }
        // The chunks of A (1024 consecutive points of its raster order) run one after the other, LAST chunk first, and share
        // the minimum found so far: only points closer to B's box than that minimum can still improve it.  The last rows of an
        // object in the image are the ones next to what it stands on, so the first chunk usually finds (almost) the minimum
        // and the others are skimmed -- an (object, table) pair costs one sweep instead of one per chunk.
        double best = lim2;                                      // block-uniform at every tile and chunk boundary
        for (int chunk = (na + E2_CT_CHUNK - 1) / E2_CT_CHUNK - 1; chunk >= 0; --chunk) {
        // ---- A chunk: points within `best` of B's box, and their own tight box ---------------------------------
        __syncthreads();                                         // the previous chunk's staging is consumed
        unsigned nA = 0;
        double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
        {
            bool take[4];
            double x[4], y[4], z[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = chunk * E2_CT_CHUNK + k * E2_CT_THREADS + tid;
                take[k] = false;
                x[k] = y[k] = z[k] = 0.0;
                if (i < na) { const unsigned p = fla[i]; x[k] = AX[p]; y[k] = AY[p]; z[k] = AZ[p]; }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = chunk * E2_CT_CHUNK + k * E2_CT_THREADS + tid;
                take[k] = (i < na) && point_box_d2(x[k], y[k], z[k], bb) <= best;
                if (take[k]) {
                    mn[0] = fmin(mn[0], x[k]); mx[0] = fmax(mx[0], x[k]);
                    mn[1] = fmin(mn[1], y[k]); mx[1] = fmax(mx[1], y[k]);
                    mn[2] = fmin(mn[2], z[k]); mx[2] = fmax(mx[2], z[k]);
                }
            }
            nA = contact_push4(s, take, x, y, z, false, 0, 0, 0);
        }
        if (nA == 0) continue;
        for (int k = 0; k < 3; ++k) {
            const double l = warp_min_d(mn[k]), h = warp_max_d(mx[k]);
            if (lane == 0) { s.red[k][warp] = l; s.red[3 + k][warp] = h; }
        }
        __syncthreads();
        if (tid < 3) {
            double l = s.red[tid][0], h = s.red[3 + tid][0];
            for (int w = 1; w < E2_CT_THREADS / 32; ++w) { l = fmin(l, s.red[tid][w]); h = fmax(h, s.red[3 + tid][w]); }
            s.box[tid] = l; s.box[3 + tid] = h;
        }
        __syncthreads();
        double cb[6];
        for (int k = 0; k < 6; ++k) cb[k] = s.box[k];
        const double ox = 0.5 * (cb[0] + cb[3]), oy = 0.5 * (cb[1] + cb[4]), oz = 0.5 * (cb[2] + cb[5]);
        const double hx = cb[3] - cb[0], hy = cb[4] - cb[1], hz = cb[5] - cb[2];
        const double M = 0.5 * sqrt(hx * hx + hy * hy + hz * hz) + sqrt(lim2);       // |relative coordinate| bound
        const double e = M * 2.4e-7;
        // this thread's A points: float64 for the recheck, float32 relative for the sweep.  Most items keep few A points
        // (only the part of a chunk that lies within thresh of B's box): with n_a <= 256 every thread holds at most one
        // point, P = the power of two >= n_a (>= 32) threads share the points and the 256 / P thread groups split the B tile.
        const int nk = (int)((nA + E2_CT_THREADS - 1) / E2_CT_THREADS);
        unsigned P = E2_CT_THREADS;
        if (nk == 1) { P = 32; while (P < nA) P <<= 1; }
        const unsigned jstart = (nk == 1) ? (unsigned)tid / P : 0u, jstep = (nk == 1) ? E2_CT_THREADS / P : 1u;
        double ax[4], ay[4], az[4];
        float rx[4], ry[4], rz[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned j = (nk == 1) ? ((unsigned)tid & (P - 1)) + (k ? nA : 0u) : (unsigned)(tid + k * E2_CT_THREADS);
            if (j < nA) {
                ax[k] = s.dx[j]; ay[k] = s.dy[j]; az[k] = s.dz[j];
                rx[k] = (float)(ax[k] - ox); ry[k] = (float)(ay[k] - oy); rz[k] = (float)(az[k] - oz);
            } else {
                ax[k] = ay[k] = az[k] = 1e300;
                rx[k] = ry[k] = rz[k] = 1e18f;                  // float32 d^2 overflows to +inf: never accepted
            }
        }
        const double c0 = 4.0 * e * sqrt(lim2) + 12.0 * e * e;
        float guard = contact_guard(best, c0);
        // ---- B tiles -----------------------------------------------------------------------------------------------
        // Ground (Morton order, one tight box per run of E2_CT_SUB points): warp 0 lists the runs that lie within thresh
        // of the item's box once; the tiles below are made of listed runs only, the others are never loaded.
        const int n_sub = (nb + E2_CT_SUB - 1) / E2_CT_SUB;
        const bool listed = gsub && n_sub <= E2_CT_RUNS;
        if (listed) {
            __syncthreads();                                     // s.box readers are done, s.n_run of the last item is dead
            if (warp == 0) {
                unsigned n = 0;
                for (int b0 = 0; b0 < n_sub; b0 += 32) {
                    const int sub = b0 + lane;
                    const bool ok = sub < n_sub && box_box_d2(gsub + (size_t)6 * sub, cb) <= lim2;
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    if (ok) s.run[n + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)sub;
                    n += __popc(bal);
                }
                if (lane == 0) s.n_run = n;
            }
            __syncthreads();
        }
        const int n_tiles = listed ? (int)((s.n_run + 3u) / 4u) : (nb + E2_CT_TILE - 1) / E2_CT_TILE;
        for (int it = 0; it < n_tiles; ++it) {
            __syncthreads();                                     // the previous tile (or the A staging) is consumed
            unsigned mB = 0;
            {
                int jb[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (listed) jb[k] = (4u * it + k < s.n_run) ? (int)s.run[4 * it + k] * E2_CT_SUB + tid : nb;
                    else jb[k] = it * E2_CT_TILE + k * E2_CT_THREADS + tid;
                }
                bool take[4];
                double x[4], y[4], z[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int j = jb[k];
                    x[k] = y[k] = z[k] = 0.0;
                    if (j < nb) { const unsigned q = flb ? flb[j] : (unsigned)j; x[k] = BX[q]; y[k] = BY[q]; z[k] = BZ[q]; }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    take[k] = (jb[k] < nb) && point_box_d2(x[k], y[k], z[k], cb) <= best;   // nothing farther than the current minimum can improve it
                mB = contact_push4(s, take, x, y, z, true, ox, oy, oz);
            }
            switch (nk) {                                        // block-uniform
            case 1: contact_sweep<1>(s, mB, jstart, jstep, rx, ry, rz, ax, ay, az, best, guard, c0); break;
            case 2: contact_sweep<2>(s, mB, 0u, 1u, rx, ry, rz, ax, ay, az, best, guard, c0); break;
            case 3: contact_sweep<3>(s, mB, 0u, 1u, rx, ry, rz, ax, ay, az, best, guard, c0); break;
            default: contact_sweep<4>(s, mB, 0u, 1u, rx, ry, rz, ax, ay, az, best, guard, c0); break;
            }
            // share the minimum: every thread continues with the block's best, which tightens guard and the B filter
            const double wb = warp_min_nonneg_d(best);
            if (lane == 0) s.red[0][warp] = wb;
            __syncthreads();
            best = s.red[0][0];
            for (int w = 1; w < E2_CT_THREADS / 32; ++w) best = fmin(best, s.red[0][w]);
            guard = contact_guard(best, c0);
            if (cs.cross && sqrt(best) < d.thresh) break;        // a witness is enough for a cross pair (best is uniform)
        }
        if (cs.cross && sqrt(best) < d.thresh) break;
        }                                                        // chunks of A
        if (tid == 0) {
            if (cs.cross) { if (sqrt(best) < d.thresh) atomicMin(cs.out, 0ull); }     // contact flag: 0.0 = closer than thresh
            else if (best < lim2) atomicMin(cs.out, d2key(best));
        }
    }
}
