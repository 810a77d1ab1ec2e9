// seg_stream.cuh -- label filter + segment compaction (manipulation_analyser.py:1433-1504) as two stream-ordered
// kernels that read the input exactly once and never make a warp wait for another one:
//
//   seg_stream_kernel   persistent warps pull UNITS (E2_SS_UNIT_TILES tiles of 128 points = 2048 consecutive points of one
//                       frame) from an atomic counter.  A warp streams its unit through its own 3-stage ring of 1-D bulk
//                       asynchronous copies (cp.async.bulk + mbarrier, tma.cuh), classifies 4 points per lane (label ->
//                       class byte through a LUT in shared memory, NaN / -100 rows dropped :1444-1453), compacts the points
//                       of stored entities in raster order into the unit's scratch list (16 B: float x, y, z + raster index
//                       / entity; stays in L2) and leaves the unit's per-entity counts in seg_cnt[frame][unit][entity].
//                       Units are small and handed out dynamically, so object-heavy image rows do not stall anybody
//                       (the cluster kernel of seg_fused.cuh spends 38 % of its stall samples at barriers).
//   seg_prefix_kernel   one CTA per frame: the small count table becomes the exclusive prefix over the units; entity totals
//                       and offsets (:1456-1472).
//   seg_emit_kernel     one warp per unit: re-reads the unit's list (13 % of the input on the 640x480 stream), ranks the
//                       points per entity (stable) from the unit's prefix and writes `translation` then `roll` applied in
//                       float64 (:1482-1485) as SoA + raster indices.
// Same results as seg_fused.cuh and as the two-pass kernels of seg.cuh (tests compare all three bit for bit).
#pragma once
#include <limits.h>

#include "common.cuh"
#include "seg.cuh"
#include "seg_fused.cuh"
#include "tma.cuh"

#define E2_SS_THREADS 128
#define E2_SS_WARPS (E2_SS_THREADS / 32)
#define E2_SS_TILE 128             // points per tile = 4 per lane
#ifndef E2_SS_STAGES
#define E2_SS_STAGES 3
#endif
#ifndef E2_SS_UNIT_TILES
#define E2_SS_UNIT_TILES 16
#endif
#ifndef E2_SS_CTAS
#define E2_SS_CTAS 7               // resident CTAs per SM (shared memory: 29 KB each)
#endif
#define E2_SS_UNIT (E2_SS_TILE * E2_SS_UNIT_TILES)

struct SegStreamSmem {
    float4 xs[E2_SS_WARPS][E2_SS_STAGES][E2_SS_TILE * 3 / 4];
    int4 ls[E2_SS_WARPS][E2_SS_STAGES][E2_SS_TILE / 4];
    uint64_t full[E2_SS_WARPS][E2_SS_STAGES];
    unsigned char lut8[E2SEC_LABEL_LUT + 16];
    unsigned wc[E2_SS_WARPS][E2SEC_MAX_ENTITIES];
};

extern __shared__ __align__(128) unsigned char seg_stream_smem_raw[];

// d.S = units per frame; `next` = the unit counter (zero on entry, reset by seg_emit_kernel)
template <bool UPSCALE>
__global__ void __launch_bounds__(E2_SS_THREADS, E2_SS_CTAS) seg_stream_kernel(Dev d, unsigned *next)
{
    pdl_enter();
    SegStreamSmem &s = *reinterpret_cast<SegStreamSmem *>(seg_stream_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = d.N, K = d.K, S = d.S;
    const int tpf = (N + E2_SS_TILE - 1) / E2_SS_TILE;
    const int gi = d.ground_idx > 0 ? d.ground_idx : -2;
    const int hand_label = d.hand_label;
    const unsigned total_units = (unsigned)d.F * (unsigned)S;
    uint64_t *full = s.full[warp];
    const uint64_t stream_once = l2_evict_first_policy();      // the input is read once: keep the scratch lists in L2 instead
    if (lane == 0) {
        for (int i = 0; i < E2_SS_STAGES; ++i) mbar_init(&full[i], 1);
        mbar_fence_init();
    }
    for (int raw = tid; raw <= E2SEC_LABEL_LUT; raw += E2_SS_THREADS) {
        const int e = raw < E2SEC_LABEL_LUT ? __ldg(d.lut + raw) : -1;
        s.lut8[raw] = (unsigned char)(e < 0 ? 0u : e == 0 ? E2_SF_ZERO : e == gi ? E2_SF_GROUND : (unsigned)e);
    }
    if (d.ent_box)                                             // the boxes seg_emit_kernel accumulates: min keys up, max keys down
        for (size_t i = (size_t)blockIdx.x * E2_SS_THREADS + tid; i < (size_t)d.F * K * 6; i += (size_t)gridDim.x * E2_SS_THREADS)
            d.ent_box[i] = (i % 6 < 3) ? 0xffffffffu : 0u;
    __syncthreads();

    unsigned phase = 0;                                        // tiles this warp has consumed so far (ring position / parity)
    for (;;) {
        unsigned unit = 0;
        if (lane == 0) unit = atomicAdd(next, 1u);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= total_units) break;
        const int f = (int)(unit / (unsigned)S), u = (int)(unit - (unsigned)f * (unsigned)S);
        const int t0 = u * E2_SS_UNIT_TILES, nt = min(E2_SS_UNIT_TILES, tpf - t0);
        const float *xyz = d.xyz + (size_t)f * N * 3;
        const int *lab = d.labels + (size_t)f * d.label_size;
        uint4 *list = d.seg_list + (size_t)f * N + (size_t)t0 * E2_SS_TILE;
        for (int e = lane; e < E2SEC_MAX_ENTITIES; e += 32) s.wc[warp][e] = 0;

        auto issue = [&](int i) {
            const int st = (int)((phase + (unsigned)i) % E2_SS_STAGES);
            const int start = (t0 + i) * E2_SS_TILE;
            const int npts = min(E2_SS_TILE, N - start);
            const unsigned xb = (unsigned)npts * 12u, lb = UPSCALE ? 0u : (unsigned)npts * 4u;
            mbar_arrive_expect_tx(&full[st], xb + lb);
            bulk_g2s_hint(&s.xs[warp][st][0], xyz + (size_t)start * 3, xb, &full[st], stream_once);
            if (!UPSCALE) bulk_g2s_hint(&s.ls[warp][st][0], lab + start, lb, &full[st], stream_once);
        };
        if (lane == 0)
            for (int i = 0; i < min(E2_SS_STAGES, nt); ++i) issue(i);

        unsigned run = 0, c_zero = 0, c_ground = 0;
        bool hand_seen = false;
        for (int i = 0; i < nt; ++i) {
            const unsigned g = phase + (unsigned)i;
            const int st = (int)(g % E2_SS_STAGES);
            mbar_wait(&full[st], (g / E2_SS_STAGES) & 1u);
            const float4 a = s.xs[warp][st][3 * lane], b = s.xs[warp][st][3 * lane + 1], c = s.xs[warp][st][3 * lane + 2];
            int4 l = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
            if (!UPSCALE) l = s.ls[warp][st][lane];
            __syncwarp();                                      // every lane has read the stage
            if (lane == 0 && i + E2_SS_STAGES < nt) issue(i + E2_SS_STAGES);
            const float px[4] = {a.x, a.w, b.z, c.y}, py[4] = {a.y, b.x, b.w, c.z}, pz[4] = {a.z, b.y, c.x, c.w};
            const int p0 = (t0 + i) * E2_SS_TILE + 4 * lane;
            const bool in = p0 < N;                            // N % 4 == 0: a lane's four points are inside together
            if (!UPSCALE && !in) l = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
            int raw[4] = {l.x, l.y, l.z, l.w};
            if (UPSCALE && in) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int p = p0 + j;
                    const int rr = p / d.width, cc = p - rr * d.width;
                    raw[j] = __ldg(lab + __ldg(d.rowmap + rr) * d.label_w + __ldg(d.colmap + cc));
                }
            }
            unsigned w4 = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int rw = raw[j];
                hand_seen |= (rw == hand_label);               // :1429 looks at the raw label image, valid point or not
                unsigned code = s.lut8[min((unsigned)rw, (unsigned)E2SEC_LABEL_LUT)];
                if (!(coord_valid(px[j]) && coord_valid(py[j]) && coord_valid(pz[j]))) code = 0;
                w4 |= code << (8 * j);
            }
            c_zero += __popc(w4 & 0x40404040u);
            c_ground += __popc(w4 & 0x80808080u);
            const unsigned e4 = w4 & 0x3f3f3f3fu;
            if (__any_sync(0xffffffffu, e4 != 0)) {            // ground / border / NaN tiles never get here
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
        phase += (unsigned)nt;
        __syncwarp();                                          // the unit's list entries are visible to all lanes
        // per-entity counts of the unit's list
        for (unsigned q = 0; q < run; q += 128) {
            int e[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned i = q + 32u * k + lane;
                e[k] = (i < run) ? (int)(__ldcg(&list[i].w) & 63u) : -1;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned peers = __match_any_sync(0xffffffffu, e[k]);
                if (e[k] >= 0 && lane == __ffs(peers) - 1) s.wc[warp][e[k]] += __popc(peers);
                __syncwarp();
            }
        }
        c_zero = __reduce_add_sync(0xffffffffu, c_zero);
        c_ground = __reduce_add_sync(0xffffffffu, c_ground);
        unsigned *cnt = d.seg_cnt + ((size_t)f * S + u) * K;
        for (int e = lane; e < K; e += 32) cnt[e] = s.wc[warp][e] + (e == 0 ? c_zero : 0u) + (e == gi ? c_ground : 0u);
        if (lane == 0) d.seg_len[(size_t)f * S + u] = run;
        if (__any_sync(0xffffffffu, hand_seen) && lane == 0) atomicOr(d.flags + f, E2SEC_FLAG_HAND_LABEL);
        __syncwarp();                                          // wc is zeroed again at the top
    }
}

// grid F, 256 threads: the frame's count table (S units x K entities) becomes, in place, the exclusive prefix over the
// units -- where every unit's points of an entity start inside the entity -- and the totals give the entity sizes and
// offsets (:1456-1472; index 0 and the ground are counted, never stored, :1520).  One warp per entity, 32 units per step.
__global__ void __launch_bounds__(256) seg_prefix_kernel(Dev d)
{
    pdl_enter();
    __shared__ unsigned total[E2SEC_MAX_ENTITIES];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, f = blockIdx.x, S = d.S, K = d.K;
    unsigned *c = d.seg_cnt + (size_t)f * S * K;
    for (int e = warp; e < K; e += 8) {
        unsigned carry = 0;
        for (int u0 = 0; u0 < S; u0 += 32) {
            const int u = u0 + lane;
            const unsigned v = u < S ? c[(size_t)u * K + e] : 0u;
            unsigned incl = v;
            for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            if (u < S) c[(size_t)u * K + e] = carry + incl - v;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) total[e] = carry;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int off = 0;
        for (int k = 0; k < K; ++k) {
            const bool stored = (k != 0 && k != d.ground_idx);
            d.ent_raw[f * K + k] = (int)total[k];
            d.ent_off[f * K + k] = stored ? off : -1;
            if (stored) off += (int)total[k];
        }
        d.frame_used[f] = off < d.cap ? off : d.cap;
        if (off > d.cap) atomicExch(d.error, E2SEC_ERR_CAPACITY);
    }
}

// grid (ceil(S / E2_SS_WARPS), F): one warp per unit (after seg_prefix_kernel): re-reads the unit's list, ranks the points
// per entity (stable) from the unit's first positions and writes the transformed SoA.
__global__ void __launch_bounds__(E2_SS_THREADS) seg_emit_kernel(Dev d, unsigned *next)
{
    pdl_enter();
    __shared__ unsigned wc[E2_SS_WARPS][E2SEC_MAX_ENTITIES];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = blockIdx.y, u = blockIdx.x * E2_SS_WARPS + warp, S = d.S, K = d.K;
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) *next = 0;      // the counter of the next batch's stream kernel
    const unsigned M = u < S ? d.seg_len[(size_t)f * S + u] : 0u;
    if (M == 0) return;                                          // a unit without stored points has nothing to place
    for (int e = lane; e < K; e += 32) {
        const int off = d.ent_off[f * K + e];
        wc[warp][e] = off < 0 ? 0xffffffffu : (unsigned)off + d.seg_cnt[((size_t)f * S + u) * K + e];   // first output position of the unit, per entity
    }
    __syncwarp();
    const uint4 *list = d.seg_list + (size_t)f * d.N + (size_t)u * E2_SS_UNIT;
    const unsigned lt = (1u << lane) - 1u;
    const size_t fb = (size_t)f * d.cap;
    // three rows of the list in flight while one is ranked
    uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0, r2 = r0;
    if ((unsigned)lane < M) r0 = __ldcg(&list[lane]);
    if ((unsigned)lane + 32 < M) r1 = __ldcg(&list[lane + 32]);
    if ((unsigned)lane + 64 < M) r2 = __ldcg(&list[lane + 64]);
    for (unsigned q = 0; q < M; q += 32) {
        const unsigned i = q + lane;
        const uint4 rec = r0;
        r0 = r1;
        r1 = r2;
        if (i + 96 < M) r2 = __ldcg(&list[i + 96]);
        const int e = (i < M) ? (int)(rec.w & 63u) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, e);
        unsigned pos = 0;
        if (e >= 0) pos = wc[warp][e] + __popc(peers & lt);
        __syncwarp();
        if (e >= 0 && lane == __ffs(peers) - 1) wc[warp][e] += __popc(peers);
        __syncwarp();
        const bool place = e >= 0 && pos < (unsigned)d.cap;
        double X = 0.0, Y = 0.0, Z = 0.0;
        if (place) {
            X = (double)__uint_as_float(rec.x); Y = (double)__uint_as_float(rec.y); Z = (double)__uint_as_float(rec.z);
            if (((rec.x & 0x7fffffffu) == 0x7f800000u) | ((rec.y & 0x7fffffffu) == 0x7f800000u) | ((rec.z & 0x7fffffffu) == 0x7f800000u)) {
                X = widen(__uint_as_float(rec.x)); Y = widen(__uint_as_float(rec.y)); Z = widen(__uint_as_float(rec.z));   // +-inf -> +-DBL_MAX (:1445), rare
            }
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
        if (d.ent_box) {
            // float32 box of the entity, rounded outward: the value range cell_filter_kernel's median selection starts from
            // (saves it a pass over the points).  One reduction per entity present in the row, six fire-and-forget atomics.
            unsigned k[6];
            k[0] = f32_ord(__double2float_rd(X)); k[1] = f32_ord(__double2float_rd(Y)); k[2] = f32_ord(__double2float_rd(Z));
            k[3] = f32_ord(__double2float_ru(X)); k[4] = f32_ord(__double2float_ru(Y)); k[5] = f32_ord(__double2float_ru(Z));
            unsigned todo = __ballot_sync(0xffffffffu, place);
            while (todo) {
                const int leader = __ffs(todo) - 1;
                const int el = __shfl_sync(0xffffffffu, e, leader);
                const unsigned mask = __ballot_sync(0xffffffffu, place && e == el);
                if (place && e == el) {
                    unsigned r[6];
#pragma unroll
                    for (int q = 0; q < 3; ++q) { r[q] = __reduce_min_sync(mask, k[q]); r[3 + q] = __reduce_max_sync(mask, k[3 + q]); }
                    if (lane == leader) {
                        unsigned *b = d.ent_box + ((size_t)f * K + el) * 6;
#pragma unroll
                        for (int q = 0; q < 3; ++q) { atomicMin(b + q, r[q]); atomicMax(b + 3 + q, r[3 + q]); }
                    }
                }
                todo &= ~mask;
            }
        }
    }
}
