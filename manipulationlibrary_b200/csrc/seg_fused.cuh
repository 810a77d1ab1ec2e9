// seg_fused.cuh -- single-pass label filter + segment compaction (manipulation_analyser.py:1433-1504), the streaming
// kernel of the path: every input byte (12 B xyz + 4 B label per point) is read from HBM exactly once.
//
// Same results as the two-pass kernels of seg.cuh (which stay as the fallback for unaligned inputs): label image
// upscaled nearest-neighbour (cv2.resize :1433), label replacement folded into a LUT (:176-198), NaN / -100 rows dropped
// (:1444-1453), points grouped per entry of total_unique_labels (:1456-1472) in raster order, `translation` then `roll`
// applied in float64 (:1482-1485).
//
// One thread-block CLUSTER per frame; CTA r of the cluster owns a contiguous run of 1024-point tiles.
//   phase 1  tiles arrive in shared memory through a 3-stage ring of 1-D bulk asynchronous copies (cp.async.bulk +
//            mbarrier, tma.cuh) issued by one thread, so 2-3 tiles per CTA are always in flight.  Each thread classifies
//            4 consecutive points; points of stored entities are compacted in raster order into a scratch list
//            (16 B: float x, y, z + raster index / entity) that stays in L2; per-entity counts accumulate in shared memory.
//   phase 2  cluster barrier; every CTA reads the other CTAs' counts through distributed shared memory and derives the
//            entity offsets of the frame and its own first output position per entity.
//   phase 3  the CTA re-reads its own scratch list (13 % of the input on the 640x480 stream), ranks the points per entity
//            (stable) and writes the transformed float64 SoA points + raster indices.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "seg.cuh"
#include "tma.cuh"

#define E2_SF_THREADS 256
#define E2_SF_WARPS (E2_SF_THREADS / 32)
#define E2_SF_TILE 1024            // points per tile = 4 per thread
#define E2_SF_STAGES 3
#define E2_SF_CLUSTER 8

struct SegFusedSmem {
    float4 xs[E2_SF_STAGES][E2_SF_TILE * 3 / 4];       // 12 KB per stage: xyz of one tile
    int4 ls[E2_SF_STAGES][E2_SF_TILE / 4];             //  4 KB per stage: labels of one tile
    uint64_t full[E2_SF_STAGES];
    unsigned cnt[E2SEC_MAX_ENTITIES];                  // valid points per entity in this CTA's tiles
    unsigned tot[E2SEC_MAX_ENTITIES], bef[E2SEC_MAX_ENTITIES], base[E2SEC_MAX_ENTITIES];
    unsigned wc[E2_SF_WARPS][E2SEC_MAX_ENTITIES];
    unsigned wt[2][E2_SF_WARPS];
};

// translation / roll whose last row is (0, 0, 0, 1): q3 == 1.0 exactly for finite points, and v / 1.0 == v
__device__ __forceinline__ void transform_point_affine(const double *M, double &x, double &y, double &z)
{
    const double q0 = ((M[0] * x + M[1] * y) + M[2] * z) + M[3];
    const double q1 = ((M[4] * x + M[5] * y) + M[6] * z) + M[7];
    const double q2 = ((M[8] * x + M[9] * y) + M[10] * z) + M[11];
    x = q0;
    y = q1;
    z = q2;
}

extern __shared__ __align__(128) unsigned char seg_fused_smem_raw[];

template <bool UPSCALE>
__global__ void __launch_bounds__(E2_SF_THREADS) seg_fused_kernel(Dev d)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    SegFusedSmem &s = *reinterpret_cast<SegFusedSmem *>(seg_fused_smem_raw);
    const unsigned C = cl.num_blocks(), r = cl.block_rank();
    const int f = blockIdx.x / C;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = d.N, K = d.K;
    const int tpf = (N + E2_SF_TILE - 1) / E2_SF_TILE;
    const int t0 = (int)((long long)tpf * r / C), t1 = (int)((long long)tpf * (r + 1) / C);
    const int nt = t1 - t0;
    const float *xyz = d.xyz + (size_t)f * N * 3;
    const int *lab = d.labels + (size_t)f * d.label_size;
    uint4 *list = d.seg_list + (size_t)f * N + (size_t)t0 * E2_SF_TILE;      // this CTA's scratch list
    const int gi = d.ground_idx > 0 ? d.ground_idx : -2;

    if (tid == 0) {
        for (int i = 0; i < E2_SF_STAGES; ++i) mbar_init(&s.full[i], 1);
        mbar_fence_init();
    }
    for (int e = tid; e < E2SEC_MAX_ENTITIES; e += E2_SF_THREADS) s.cnt[e] = 0;
    __syncthreads();

    auto issue = [&](int i) {
        const int st = i % E2_SF_STAGES;
        const int start = (t0 + i) * E2_SF_TILE;
        const int npts = min(E2_SF_TILE, N - start);
        const unsigned xb = (unsigned)npts * 12u, lb = UPSCALE ? 0u : (unsigned)npts * 4u;
        mbar_arrive_expect_tx(&s.full[st], xb + lb);
        bulk_g2s(&s.xs[st][0], xyz + (size_t)start * 3, xb, &s.full[st]);
        if (!UPSCALE) bulk_g2s(&s.ls[st][0], lab + start, lb, &s.full[st]);
    };
    if (tid == 0)
        for (int i = 0; i < min(E2_SF_STAGES, nt); ++i) issue(i);

    // ---- phase 1 ---------------------------------------------------------------------------------------------------
    unsigned run = 0, c_zero = 0, c_ground = 0;
    bool hand_seen = false;
    for (int i = 0; i < nt; ++i) {
        const int st = i % E2_SF_STAGES;
        mbar_wait(&s.full[st], (unsigned)(i / E2_SF_STAGES) & 1u);
        const float4 a = s.xs[st][3 * tid], b = s.xs[st][3 * tid + 1], c = s.xs[st][3 * tid + 2];
        const float px[4] = {a.x, a.w, b.z, c.y}, py[4] = {a.y, b.x, b.w, c.z}, pz[4] = {a.z, b.y, c.x, c.w};
        int raw[4];
        const int p0 = (t0 + i) * E2_SF_TILE + 4 * tid;
        if (UPSCALE) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int p = p0 + j;
                raw[j] = -1;
                if (p < N) {
                    const int rr = p / d.width, cc = p - rr * d.width;
                    raw[j] = __ldg(lab + __ldg(d.rowmap + rr) * d.label_w + __ldg(d.colmap + cc));
                }
            }
        } else {
            const int4 l = s.ls[st][tid];
            raw[0] = l.x; raw[1] = l.y; raw[2] = l.z; raw[3] = l.w;
        }
        int ent[4];
        unsigned mine = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool in = (p0 + j) < N;
            const int rw = raw[j];
            hand_seen |= in && (rw == d.hand_label);
            int e = (in && rw >= 0 && rw < E2SEC_LABEL_LUT) ? __ldg(d.lut + rw) : -1;
            if (!(coord_valid(px[j]) && coord_valid(py[j]) && coord_valid(pz[j]))) e = -1;
            c_zero += __popc(__ballot_sync(0xffffffffu, e == 0));
            c_ground += __popc(__ballot_sync(0xffffffffu, e == gi));
            const bool stored = e > 0 && e != gi;
            ent[j] = stored ? e : -1;
            mine += stored;
            unsigned todo = __ballot_sync(0xffffffffu, stored);
            while (todo) {
                const int leader = __ffs(todo) - 1;
                const int ev = __shfl_sync(0xffffffffu, e, leader);
                const unsigned m = __ballot_sync(0xffffffffu, stored && e == ev);
                if (lane == leader) atomicAdd(&s.cnt[ev], (unsigned)__popc(m));
                todo &= ~m;
            }
        }
        unsigned incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) s.wt[i & 1][warp] = incl;
        __syncthreads();                                   // the stage is consumed, the warp totals are visible
        if (tid == 0 && i + E2_SF_STAGES < nt) issue(i + E2_SF_STAGES);
        unsigned before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < E2_SF_WARPS; ++w) { const unsigned t = s.wt[i & 1][w]; if (w < warp) before += t; total += t; }
        unsigned pos = run + before + incl - mine;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (ent[j] >= 0) {
                uint4 rec;
                rec.x = __float_as_uint(px[j]); rec.y = __float_as_uint(py[j]); rec.z = __float_as_uint(pz[j]);
                rec.w = ((unsigned)(p0 + j) << 6) | (unsigned)ent[j];
                list[pos++] = rec;
            }
        run += total;
    }
    if (lane == 0) {
        if (c_zero) atomicAdd(&s.cnt[0], c_zero);
        if (c_ground) atomicAdd(&s.cnt[gi], c_ground);
    }
    if (__any_sync(0xffffffffu, hand_seen) && lane == 0) atomicOr(d.flags + f, E2SEC_FLAG_HAND_LABEL);
    __syncthreads();

    // ---- phase 2: entity totals of the frame through distributed shared memory -------------------------------------
    cl.sync();
    if (tid < K) {
        unsigned tot = 0, bef = 0;
        for (unsigned c = 0; c < C; ++c) {
            const unsigned v = *cl.map_shared_rank(&s.cnt[tid], c);
            tot += v;
            if (c < r) bef += v;
        }
        s.tot[tid] = tot;
        s.bef[tid] = bef;
    }
    cl.sync();                                             // nobody reads a peer's shared memory after this point
    if (tid == 0) {
        unsigned off = 0;
        for (int k = 0; k < K; ++k) {
            const bool stored = (k != 0 && k != d.ground_idx);     // :1520 -- index 0 and the ground label are never filtered nor related
            s.base[k] = stored ? off + s.bef[k] : 0xffffffffu;
            if (r == 0) {
                d.ent_raw[f * K + k] = (int)s.tot[k];
                d.ent_off[f * K + k] = stored ? (int)off : -1;
            }
            if (stored) off += s.tot[k];
        }
        if (r == 0) {
            d.frame_used[f] = (int)off < d.cap ? (int)off : d.cap;
            if ((int)off > d.cap) atomicExch(d.error, E2SEC_ERR_CAPACITY);
        }
    }
    for (int e = tid; e < E2_SF_WARPS * E2SEC_MAX_ENTITIES; e += E2_SF_THREADS) (&s.wc[0][0])[e] = 0;
    __syncthreads();

    // ---- phase 3: stable per-entity ranks over the CTA's own list, transform, scatter ------------------------------
    const unsigned M = run;
    const unsigned rows = (M + 31u) / 32u, rpw = (rows + E2_SF_WARPS - 1) / E2_SF_WARPS;
    const unsigned w_lo = min(M, warp * rpw * 32u), w_hi = min(M, (warp + 1) * rpw * 32u);
    const unsigned lt = (1u << lane) - 1u;
    for (unsigned q = w_lo; q < w_hi; q += 32) {
        const unsigned i = q + lane;
        const int e = (i < w_hi) ? (int)(__ldcg(&list[i].w) & 63u) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, e);
        if (e >= 0 && lane == __ffs(peers) - 1) s.wc[warp][e] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    if (tid < K) {
        unsigned acc = s.base[tid];
        for (int w = 0; w < E2_SF_WARPS; ++w) { const unsigned t = s.wc[w][tid]; s.wc[w][tid] = acc; acc += t; }
    }
    __syncthreads();
    const size_t fb = (size_t)f * d.cap;
    for (unsigned q = w_lo; q < w_hi; q += 32) {
        const unsigned i = q + lane;
        uint4 rec = make_uint4(0, 0, 0, 0);
        int e = -1;
        if (i < w_hi) { rec = __ldcg(&list[i]); e = (int)(rec.w & 63u); }
        const unsigned peers = __match_any_sync(0xffffffffu, e);
        unsigned pos = 0;
        if (e >= 0) pos = s.wc[warp][e] + __popc(peers & lt);
        __syncwarp();
        if (e >= 0 && lane == __ffs(peers) - 1) s.wc[warp][e] += __popc(peers);
        __syncwarp();
        if (e >= 0 && pos < (unsigned)d.cap) {
            double X = widen(__uint_as_float(rec.x)), Y = widen(__uint_as_float(rec.y)), Z = widen(__uint_as_float(rec.z));
            if (d.affine) {
                transform_point_affine(d.T, X, Y, Z);
                transform_point_affine(d.R, X, Y, Z);
            } else {
                transform_point(d.T, X, Y, Z);
                transform_point(d.R, X, Y, Z);
            }
            d.px[fb + pos] = X;
            d.py[fb + pos] = Y;
            d.pz[fb + pos] = Z;
            d.pidx[fb + pos] = rec.w >> 6;
        }
    }
}
