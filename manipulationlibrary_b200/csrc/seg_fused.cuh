// seg_fused.cuh -- single-pass label filter + segment compaction (manipulation_analyser.py:1433-1504), the streaming
// kernel of the path: every input byte (12 B xyz + 4 B label per point) is read from HBM exactly once.
//
// Same results as the two-pass kernels of seg.cuh (which stay as the fallback for unaligned inputs): label image
// upscaled nearest-neighbour (cv2.resize :1433), label replacement folded into a LUT (:176-198), NaN / -100 rows dropped
// (:1444-1453), points grouped per entry of total_unique_labels (:1456-1472) in raster order, `translation` then `roll`
// applied in float64 (:1482-1485).
//
// One thread-block CLUSTER per frame.  CTA r of the cluster owns a contiguous run of 128-point warp tiles and every
// WARP of the CTA owns a contiguous part of that run, so raster order = (CTA, warp, position in the warp's list) and no
// block-wide barrier is needed while streaming.
//   phase 1  each warp streams its tiles through its OWN ring of 1-D bulk asynchronous copies (cp.async.bulk + mbarrier,
//            tma.cuh; lane 0 re-arms a stage as soon as the warp has read it), so E2_SF_STAGES tiles per warp are always
//            in flight.  Each lane classifies 4 consecutive points (label -> entity through a byte LUT in shared memory);
//            points of stored entities are compacted in raster order into the warp's scratch list (16 B: float x, y, z +
//            raster index / entity) that stays in L2.  Warps whose tile holds no stored point (ground, border, NaN:
//            87 % of the 640x480 stream) skip the ranking altogether.
//   phase 2  per-entity counts of the warp lists -> CTA counts; cluster barrier; every CTA reads the other CTAs' counts
//            through distributed shared memory and derives the entity offsets of the frame and its own first output
//            position per entity.
//   phase 3  each warp re-reads its own scratch list (13 % of the input on the 640x480 stream), ranks the points per
//            entity (stable) and writes the transformed float64 SoA points + raster indices.
#pragma once
#include <cooperative_groups.h>
#include <limits.h>

#include "common.cuh"
#include "seg.cuh"
#include "tma.cuh"

#define E2_SF_THREADS 256
#define E2_SF_WARPS (E2_SF_THREADS / 32)
#define E2_SF_TILE 128             // points per warp tile = 4 per lane
#define E2_SF_STAGES 3
#define E2_SF_CLUSTER 8
#define E2_SF_ZERO 0x40u           // class byte: bits 0-5 = stored entity (0: none), bit 6 = index 0 of total_unique_labels,
#define E2_SF_GROUND 0x80u         //             bit 7 = the ground entity (both are counted, never stored, :1520)

struct SegFusedSmem {
    float4 xs[E2_SF_WARPS][E2_SF_STAGES][E2_SF_TILE * 3 / 4];     // 1536 B per stage: xyz of one warp tile
    int4 ls[E2_SF_WARPS][E2_SF_STAGES][E2_SF_TILE / 4];           //  512 B per stage: its labels
    uint64_t full[E2_SF_WARPS][E2_SF_STAGES];
    unsigned char lut8[E2SEC_LABEL_LUT + 16];          // + the empty entry that out-of-range labels clamp to
    unsigned cnt[E2SEC_MAX_ENTITIES];                  // valid points per entity in this CTA's tiles
    unsigned tot[E2SEC_MAX_ENTITIES], bef[E2SEC_MAX_ENTITIES], base[E2SEC_MAX_ENTITIES];
    unsigned wc[E2_SF_WARPS][E2SEC_MAX_ENTITIES];
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
__global__ void __launch_bounds__(E2_SF_THREADS, 4) seg_fused_kernel(Dev d)
{
    pdl_enter();
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    SegFusedSmem &s = *reinterpret_cast<SegFusedSmem *>(seg_fused_smem_raw);
    const unsigned C = cl.num_blocks(), r = cl.block_rank();
    const int f = blockIdx.x / C;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = d.N, K = d.K;
    const int tpf = (N + E2_SF_TILE - 1) / E2_SF_TILE;
    const int c0 = (int)((long long)tpf * r / C), c1 = (int)((long long)tpf * (r + 1) / C);
    const int t0 = c0 + (int)((long long)(c1 - c0) * warp / E2_SF_WARPS);           // this warp's tiles
    const int t1 = c0 + (int)((long long)(c1 - c0) * (warp + 1) / E2_SF_WARPS);
    const int nt = t1 - t0;
    const float *xyz = d.xyz + (size_t)f * N * 3;
    const int *lab = d.labels + (size_t)f * d.label_size;
    uint4 *list = d.seg_list + (size_t)f * N + (size_t)t0 * E2_SF_TILE;            // this warp's scratch list
    const int gi = d.ground_idx > 0 ? d.ground_idx : -2;
    uint64_t *full = s.full[warp];

    auto issue = [&](int i) {
        const int st = i % E2_SF_STAGES;
        const int start = (t0 + i) * E2_SF_TILE;
        const int npts = min(E2_SF_TILE, N - start);
        const unsigned xb = (unsigned)npts * 12u, lb = UPSCALE ? 0u : (unsigned)npts * 4u;
        mbar_arrive_expect_tx(&full[st], xb + lb);
        bulk_g2s(&s.xs[warp][st][0], xyz + (size_t)start * 3, xb, &full[st]);
        if (!UPSCALE) bulk_g2s(&s.ls[warp][st][0], lab + start, lb, &full[st]);
    };
    if (lane == 0) {
        for (int i = 0; i < E2_SF_STAGES; ++i) mbar_init(&full[i], 1);
        mbar_fence_init();
        for (int i = 0; i < min(E2_SF_STAGES, nt); ++i) issue(i);                    // the first tiles fly during the prologue
    }
    for (int e = tid; e < E2SEC_MAX_ENTITIES; e += E2_SF_THREADS) s.cnt[e] = 0;
    for (int e = tid; e < E2_SF_WARPS * E2SEC_MAX_ENTITIES; e += E2_SF_THREADS) (&s.wc[0][0])[e] = 0;
    for (int raw = tid; raw <= E2SEC_LABEL_LUT; raw += E2_SF_THREADS) {
        const int e = raw < E2SEC_LABEL_LUT ? __ldg(d.lut + raw) : -1;
        s.lut8[raw] = (unsigned char)(e < 0 ? 0u : e == 0 ? E2_SF_ZERO : e == gi ? E2_SF_GROUND : (unsigned)e);
    }
    __syncthreads();

    // ---- phase 1 ---------------------------------------------------------------------------------------------------
    unsigned run = 0, c_zero = 0, c_ground = 0;
    bool hand_seen = false;
    const int hand_label = d.hand_label;
    for (int i = 0; i < nt; ++i) {
        const int st = i % E2_SF_STAGES;
        mbar_wait(&full[st], (unsigned)(i / E2_SF_STAGES) & 1u);
        const float4 a = s.xs[warp][st][3 * lane], b = s.xs[warp][st][3 * lane + 1], c = s.xs[warp][st][3 * lane + 2];
        int4 l = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
        if (!UPSCALE) l = s.ls[warp][st][lane];
        __syncwarp();                                          // every lane has read the stage
        if (lane == 0 && i + E2_SF_STAGES < nt) issue(i + E2_SF_STAGES);
        const float px[4] = {a.x, a.w, b.z, c.y}, py[4] = {a.y, b.x, b.w, c.z}, pz[4] = {a.z, b.y, c.x, c.w};
        const int p0 = (t0 + i) * E2_SF_TILE + 4 * lane;
        const bool in = p0 < N;                                // N % 4 == 0: a lane's four points are inside together
        if (!UPSCALE && !in) l = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);     // stale shared memory past the frame's end
        int raw[4] = {l.x, l.y, l.z, l.w};
        if (UPSCALE && in) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int p = p0 + j;
                const int rr = p / d.width, cc = p - rr * d.width;
                raw[j] = __ldg(lab + __ldg(d.rowmap + rr) * d.label_w + __ldg(d.colmap + cc));
            }
        }
        // one class byte per point, packed: bits 0-5 stored entity (0 = none), 0x40 = index 0, 0x80 = ground
        unsigned w4 = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int rw = raw[j];
            hand_seen |= (rw == hand_label);                   // :1429 looks at the raw label image, valid point or not
            unsigned code = s.lut8[min((unsigned)rw, (unsigned)E2SEC_LABEL_LUT)];      // out of range -> the empty entry
            if (!(coord_valid(px[j]) && coord_valid(py[j]) && coord_valid(pz[j]))) code = 0;
            w4 |= code << (8 * j);
        }
        c_zero += __popc(w4 & 0x40404040u);
        c_ground += __popc(w4 & 0x80808080u);
        const unsigned e4 = w4 & 0x3f3f3f3fu;
        if (__any_sync(0xffffffffu, e4 != 0)) {                // ground / border / NaN tiles never get here
            unsigned mine = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) mine += ((e4 >> (8 * j)) & 63u) != 0;
            unsigned incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            unsigned pos = run + incl - mine;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned e = (e4 >> (8 * j)) & 63u;
                if (e) {
                    uint4 rec;
                    rec.x = __float_as_uint(px[j]); rec.y = __float_as_uint(py[j]); rec.z = __float_as_uint(pz[j]);
                    rec.w = ((unsigned)(p0 + j) << 6) | e;
                    list[pos++] = rec;
                }
            }
            run += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    c_zero = __reduce_add_sync(0xffffffffu, c_zero);
    c_ground = __reduce_add_sync(0xffffffffu, c_ground);
    if (lane == 0) {
        if (c_zero) atomicAdd(&s.cnt[0], c_zero);
        if (c_ground) atomicAdd(&s.cnt[gi], c_ground);
    }
    if (__any_sync(0xffffffffu, hand_seen) && lane == 0) atomicOr(d.flags + f, E2SEC_FLAG_HAND_LABEL);
    __syncwarp();                                              // the warp's list entries are visible to all its lanes

    // ---- phase 2: entity counts of the warp lists, then the frame's totals through distributed shared memory ------
    const unsigned M = run;
    for (unsigned q = 0; q < M; q += 128) {
        int e[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned i = q + 32u * k + lane;
            e[k] = (i < M) ? (int)(__ldcg(&list[i].w) & 63u) : -1;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned peers = __match_any_sync(0xffffffffu, e[k]);
            if (e[k] >= 0 && lane == __ffs(peers) - 1) s.wc[warp][e[k]] += __popc(peers);
            __syncwarp();
        }
    }
    __syncthreads();
    if (tid < K) {
        unsigned acc = s.cnt[tid];                             // index 0 and the ground were counted on the fly
        for (int w = 0; w < E2_SF_WARPS; ++w) acc += s.wc[w][tid];
        s.cnt[tid] = acc;
    }
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
    __syncthreads();
    if (tid < K) {                                         // first output position of every warp, per entity
        unsigned acc = s.base[tid];
        for (int w = 0; w < E2_SF_WARPS; ++w) { const unsigned t = s.wc[w][tid]; s.wc[w][tid] = acc; acc += t; }
    }
    __syncthreads();

    // ---- phase 3: stable per-entity ranks over the warp's own list, transform, scatter -----------------------------
    const unsigned lt = (1u << lane) - 1u;
    const size_t fb = (size_t)f * d.cap;
    uint4 nxt = make_uint4(0, 0, 0, 0);
    if (lane < M) nxt = __ldcg(&list[lane]);
    for (unsigned q = 0; q < M; q += 32) {
        const unsigned i = q + lane;
        const uint4 rec = nxt;
        if (i + 32 < M) nxt = __ldcg(&list[i + 32]);       // the next row is in flight while this one is ranked
        const int e = (i < M) ? (int)(rec.w & 63u) : -1;
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
