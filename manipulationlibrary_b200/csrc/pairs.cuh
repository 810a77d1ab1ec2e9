// pairs.cuh -- per-entity reductions, all-pairs touching distances and record assembly.
//
//   entity_stats_kernel   AABB (get_axis_aligned_bounding_box :672-702), cloud mean (get_center :1173,
//                         summed sequentially in output order like numpy / Open3D do, seed copy first) and
//                         count of every filtered cloud
//   pair_distance_kernel  np.min(a.compute_point_cloud_distance(b)) (:397-484, :1570-1653) for every pair of
//                         stored entities and every (entity, ground) pair, exact whenever it is < thresh
//   assemble_kernel       packs the per-frame record of include/e2sec.h
#pragma once
#include "common.cuh"

#define E2_STATS_THREADS 256
#define E2_PD_THREADS 256
#define E2_PD_TILE 1024

__global__ void __launch_bounds__(E2_STATS_THREADS) entity_stats_kernel(Dev d)
{
    pdl_enter();
    __shared__ double red[6][E2_STATS_THREADS / 32];
    const int fe = blockIdx.x, f = fe / d.K;
    const int off = d.ent_off[fe];
    const int tid = threadIdx.x;
    if (d.radius_mode == E2SEC_RADIUS_BOUNDED && d.amb[fe] != 1) return;  // cell_filter_kernel wrote the stats
    double *aabb = d.aabb + (size_t)fe * 6, *mean = d.mean + (size_t)fe * 3;
    const int m = (off < 0) ? 0 : d.n_list[fe];
    if (m == 0) {
        if (tid < 6) aabb[tid] = 0.0;      // Open3D: an empty cloud has a zero box and a zero centre
        if (tid < 3) mean[tid] = 0.0;
        return;
    }
    const size_t pb = (size_t)f * d.cap + off;
    const double *X = d.px + pb, *Y = d.py + pb, *Z = d.pz + pb;
    const unsigned *fl = d.flist + pb;
    const int seed = d.seed_pos[fe];
    // sequential sums, one thread per axis in three different warps
    if (tid == 0 || tid == 32 || tid == 64) {
        const double *A = (tid == 0) ? X : (tid == 32) ? Y : Z;
        double s = 0.0;
        int cnt = m;
        if (seed >= 0) { s = A[seed]; ++cnt; }     // dicti[0] = first_point (pyx:48)
        for (int i = 0; i < m; ++i) s += A[fl[i]];
        mean[tid >> 5] = s / (double)cnt;
    }
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int i = tid; i < m; i += E2_STATS_THREADS) {
        const unsigned j = fl[i];
        lo[0] = fmin(lo[0], X[j]); hi[0] = fmax(hi[0], X[j]);
        lo[1] = fmin(lo[1], Y[j]); hi[1] = fmax(hi[1], Y[j]);
        lo[2] = fmin(lo[2], Z[j]); hi[2] = fmax(hi[2], Z[j]);
    }
    for (int a = 0; a < 3; ++a) {
        lo[a] = warp_min_d(lo[a]);
        hi[a] = warp_max_d(hi[a]);
        if ((tid & 31) == 0) { red[a][tid >> 5] = lo[a]; red[3 + a][tid >> 5] = hi[a]; }
    }
    __syncthreads();
    if (tid < 3) {
        double l = red[tid][0], h = red[3 + tid][0];
        for (int w = 1; w < E2_STATS_THREADS / 32; ++w) { l = fmin(l, red[tid][w]); h = fmax(h, red[3 + tid][w]); }
        aabb[tid] = l;
        aabb[3 + tid] = h;
    }
}

// squared distance from a point to a box (0 inside)
__device__ __forceinline__ double point_box_d2(double x, double y, double z, const double *b)
{
    double dx = fmax(fmax(b[0] - x, 0.0), x - b[3]);
    double dy = fmax(fmax(b[1] - y, 0.0), y - b[4]);
    double dz = fmax(fmax(b[2] - z, 0.0), z - b[5]);
    return dx * dx + dy * dy + dz * dz;
}
__device__ __forceinline__ double box_box_d2(const double *a, const double *b)
{
    double s = 0.0;
    for (int k = 0; k < 3; ++k) {
        double g = fmax(fmax(a[k] - b[3 + k], 0.0), b[k] - a[3 + k]);
        s += g * g;
    }
    return s;
}

// One thread: which frame slot holds the previous KEPT frame of every frame of the batch (kept = hand label in
// the labels (:1429) and a non-empty filtered hand (:1475, :1551)); slot maxF is the halo carried over from
// the previous batches.
__global__ void __launch_bounds__(1024) prev_kept_kernel(Dev d)
{
    pdl_enter();
    __shared__ int warp_max[32];
    __shared__ int slot_s, id_s, last_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        slot_s = d.halo_info[0] ? d.maxF : -1;
        id_s = d.halo_info[0] ? d.halo_info[1] : -1;
        last_s = -1;
    }
    __syncthreads();
    for (int base = 0; base < d.F; base += 1024) {           // running maximum of the kept frame indices (a block scan)
        const int f = base + tid;
        const bool kept = f < d.F && (d.flags[f] & E2SEC_FLAG_HAND_LABEL) && d.ent_off[f * d.K + d.hand_idx] >= 0 &&
                          d.n_list[f * d.K + d.hand_idx] > 0;
        int incl = kept ? f : -1;
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl = max(incl, t); }
        if (lane == 31) warp_max[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = warp_max[lane];
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w = max(w, t); }
            warp_max[lane] = w;
        }
        __syncthreads();
        int excl = __shfl_up_sync(0xffffffffu, incl, 1);      // last kept frame strictly before f inside this block
        if (lane == 0) excl = -1;
        if (warp) excl = max(excl, warp_max[warp - 1]);
        if (f < d.F) {
            d.prev_slot[f] = excl >= 0 ? excl : slot_s;
            d.prev_id[f] = excl >= 0 ? d.frame_ids[excl] : id_s;
        }
        const int block_last = warp_max[31];
        __syncthreads();
        if (tid == 0 && block_last >= 0) { slot_s = block_last; id_s = d.frame_ids[block_last]; last_s = block_last; }
        __syncthreads();
    }
    if (tid == 0) d.halo_info[2] = last_s;
}

// Copies the last kept frame of the batch into the halo slot (grid-stride over its stored points).
__global__ void halo_save_kernel(Dev d)
{
    pdl_enter();
    const int src = d.halo_info[2];
    if (src < 0) return;
    const int K = d.K, H = d.maxF;
    const int used = d.frame_used[src];
    const size_t sb = (size_t)src * d.cap, hb = (size_t)H * d.cap;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < used; i += gridDim.x * blockDim.x) {
        d.px[hb + i] = d.px[sb + i];
        d.py[hb + i] = d.py[sb + i];
        d.pz[hb + i] = d.pz[sb + i];
        d.flist[hb + i] = d.flist[sb + i];
    }
    if (blockIdx.x == 0) {
        for (int e = threadIdx.x; e < K; e += blockDim.x) {
            d.ent_off[H * K + e] = d.ent_off[src * K + e];
            d.n_list[H * K + e] = d.n_list[src * K + e];
        }
        for (int i = threadIdx.x; i < K * 6; i += blockDim.x) d.aabb[(size_t)H * K * 6 + i] = d.aabb[(size_t)src * K * 6 + i];
        if (threadIdx.x == 0) { d.frame_used[H] = used; }
    }
}
__global__ void halo_commit_kernel(Dev d)     // after halo_save_kernel: publish the new halo
{
    pdl_enter();
    *d.amb_count = 0;                        // next batch
    const int src = d.halo_info[2];
    if (src < 0) return;
    d.halo_info[0] = 1;
    d.halo_info[1] = d.frame_ids[src];
}

// grid (pair slot, F). Slots: a < b over the K entities of the frame, then (a, ground) for a < K, then the K x K
// cross pairs (entity a of the previous kept frame, entity b of this frame).
__global__ void __launch_bounds__(E2_PD_THREADS) pair_distance_kernel(Dev d)
{
    pdl_enter();
    __shared__ double sx[E2_PD_TILE], sy[E2_PD_TILE], sz[E2_PD_TILE];
    __shared__ unsigned s_n;
    __shared__ double red[E2_PD_THREADS / 32];
    __shared__ double boxa[6], boxb[6];
    const int f = blockIdx.y, K = d.K, tid = threadIdx.x;
    const u64 INF = 0x7ff0000000000000ull;
    int a, b, fa = f;
    u64 *out;
    {
        int slot = blockIdx.x;
        const int npairs = K * (K - 1) / 2;
        if (slot < npairs) {
            a = 0;
            int row = K - 1;
            while (slot >= row) { slot -= row; --row; ++a; }
            b = a + 1 + slot;
            out = d.pair_key + ((size_t)f * (K + 1) + a) * (K + 1) + b;
        } else if (slot < npairs + K) {
            a = slot - npairs;
            b = K;
            out = d.pair_key + ((size_t)f * (K + 1) + a) * (K + 1) + b;
        } else {
            slot -= npairs + K;
            a = slot / K;
            b = slot - a * K;
            out = d.cross_key + ((size_t)f * K + a) * K + b;
            fa = d.prev_slot[f];
            if (fa < 0 || a == b) { if (tid == 0) *out = INF; return; }
        }
    }
    const int offa = d.ent_off[fa * K + a];
    const int na = (offa < 0) ? 0 : d.n_list[fa * K + a];
    int nb, offb = 0;
    if (b == K) nb = d.n_ground;
    else { offb = d.ent_off[f * K + b]; nb = (offb < 0) ? 0 : d.n_list[f * K + b]; }
    if (na == 0 || nb == 0) { if (tid == 0) *out = INF; return; }
    if (tid < 6) {
        boxa[tid] = d.aabb[(size_t)(fa * K + a) * 6 + tid];
        boxb[tid] = (b == K) ? d.gaabb[tid] : d.aabb[(size_t)(f * K + b) * 6 + tid];
    }
    __syncthreads();
    // anything closer than thresh survives this (slightly inflated) squared bound
    const double lim2 = d.thresh * d.thresh * 1.000001 + 1e-300;
    const double gap2 = box_box_d2(boxa, boxb);
    if (gap2 > lim2) { if (tid == 0) *out = d2key(gap2); return; }
    const size_t pa = (size_t)fa * d.cap + offa;
    const double *AX = d.px + pa, *AY = d.py + pa, *AZ = d.pz + pa;
    const unsigned *fla = d.flist + pa;
    const double *BX, *BY, *BZ;
    const unsigned *flb = nullptr;
    if (b == K) { BX = d.gx; BY = d.gy; BZ = d.gz; }
    else {
        const size_t pbb = (size_t)f * d.cap + offb;
        BX = d.px + pbb; BY = d.py + pbb; BZ = d.pz + pbb;
        flb = d.flist + pbb;
    }
    double best = __longlong_as_double((long long)INF);
    for (int base = 0; base < nb; base += E2_PD_TILE) {
        if (tid == 0) s_n = 0;
        __syncthreads();
        for (int j = base + tid; j < min(nb, base + E2_PD_TILE); j += E2_PD_THREADS) {
            const unsigned q = flb ? flb[j] : (unsigned)j;
            const double x = BX[q], y = BY[q], z = BZ[q];
            if (point_box_d2(x, y, z, boxa) <= lim2) {
                const unsigned p = atomicAdd(&s_n, 1u);
                sx[p] = x; sy[p] = y; sz[p] = z;
            }
        }
        __syncthreads();
        const unsigned m = s_n;
        if (m) {
            for (int i = tid; i < na; i += E2_PD_THREADS) {
                const unsigned p = fla[i];
                const double x = AX[p], y = AY[p], z = AZ[p];
                if (point_box_d2(x, y, z, boxb) > lim2) continue;
                for (unsigned j = 0; j < m; ++j) best = fmin(best, dist2(x, y, z, sx[j], sy[j], sz[j]));
            }
        }
        __syncthreads();
    }
    best = warp_min_nonneg_d(best);
    if ((tid & 31) == 0) red[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < E2_PD_THREADS / 32; ++w) best = fmin(best, red[w]);
        if (isinf(best)) best = fmax(gap2, d.thresh * d.thresh);   // no candidate pair: the clouds are >= thresh apart
        *out = d2key(best);
    }
}

// one CTA per frame
__global__ void assemble_kernel(Dev d)
{
    pdl_enter();
    // d.K entities on the device, Ke = d.K - d.ent_shift in the caller's record (the phantom in front is dropped)
    const int f = blockIdx.x, K = d.K, sh = d.ent_shift, Ke = K - sh;
    double *rec = d.records + (size_t)f * d.rec_doubles;
    const double NaN = __longlong_as_double(0x7ff8000000000000ll);
    const double INF = __longlong_as_double(0x7ff0000000000000ll);
    if (threadIdx.x == 0) {
        int flags = d.flags[f] & E2SEC_FLAG_HAND_LABEL;
        const int hoff = d.ent_off[f * K + d.hand_idx];
        if (hoff >= 0 && d.n_list[f * K + d.hand_idx] > 0) flags |= E2SEC_FLAG_HAND_POINTS;
        rec[0] = (double)d.frame_ids[f];
        rec[1] = (double)flags;
        rec[2] = (double)Ke;
        rec[3] = (double)d.prev_id[f];
    }
    for (int e = threadIdx.x; e < Ke; e += blockDim.x) {
        const int fe = f * K + e + sh;
        double *r = rec + E2SEC_REC_HEADER + E2SEC_REC_ENTITY * e;
        const bool stored = d.ent_off[fe] >= 0;
        const int m = stored ? d.n_list[fe] : 0;
        const int seed = stored ? d.seed_pos[fe] : -1;
        r[0] = (double)d.ent_raw[fe];
        r[1] = (double)(m + (seed >= 0 ? 1 : 0));
        r[2] = (seed >= 0) ? (double)d.pidx[(size_t)f * d.cap + d.ent_off[fe] + seed] : -1.0;
        r[3] = (seed >= 0) ? d.radius[fe] : NaN;
        for (int k = 0; k < 3; ++k) r[4 + k] = (seed >= 0) ? d.med[(size_t)fe * 3 + k] : NaN;
        for (int k = 0; k < 6; ++k) r[7 + k] = stored ? d.aabb[(size_t)fe * 6 + k] : 0.0;
        for (int k = 0; k < 3; ++k) r[13 + k] = stored ? d.mean[(size_t)fe * 3 + k] : 0.0;
    }
    double *pt = rec + E2SEC_REC_HEADER + E2SEC_REC_ENTITY * Ke;
    const int W = Ke + 1, Wd = K + 1;
    for (int i = threadIdx.x; i < W * W; i += blockDim.x) {
        const int a = i / W + sh, b = i % W + sh;                 // the ground cloud: Ke -> K
        double v = INF;
        if (a != b) {
            const int lo = a < b ? a : b, hi = a < b ? b : a;
            const bool ok = (lo != 0 && lo != d.ground_idx && (hi == K || (hi != 0 && hi != d.ground_idx)));
            if (ok) v = sqrt(key2d(d.pair_key[((size_t)f * Wd + lo) * Wd + hi]));
        }
        pt[i] = v;
    }
    double *ct = pt + W * W;
    for (int i = threadIdx.x; i < Ke * Ke; i += blockDim.x) {
        const int a = i / Ke + sh, b = i % Ke + sh;
        const bool ok = a != b && a != 0 && b != 0 && a != d.ground_idx && b != d.ground_idx;
        ct[i] = ok ? sqrt(key2d(d.cross_key[(size_t)f * K * K + (size_t)a * K + b])) : INF;
    }
    double *rb = ct + Ke * Ke;
    for (int e = threadIdx.x; e < Ke; e += blockDim.x) {
        const int fe = f * K + e + sh;
        const bool filt = d.ent_off[fe] >= 0 && d.seed_pos[fe] >= 0;
        const double r = d.radius[fe];
        const bool exact = filt && !(r != r);
        rb[2 * e] = !filt ? NaN : exact ? r : d.rad_lo[fe];
        rb[2 * e + 1] = !filt ? NaN : exact ? r : d.rad_hi[fe];
    }
}
