// seg.cuh -- label filter + segment compaction (manipulation_analyser.py:1433-1504).
//
// Reference behaviour restated:  label image upscaled nearest-neighbour (cv2.resize :1433), sequential
// label replacement (_replace_labels :176-198, folded into a LUT), NaN -> -100 sentinel and removal of
// every row holding the sentinel (:1444-1453), points grouped per entry of total_unique_labels
// (:1456-1472), homogeneous transform by `translation` then `roll` (:1482-1485).
// The reference orders points with an unstable argsort (:1449); this path pins raster order inside
// each entity (a stable compaction), the same pin the oracle adds.
//
// Two passes over the frame (16 B/point each: 12 B xyz + 4 B label), one warp per 1024-point segment:
//   seg_count_kernel   per-(segment, entity) counts of valid points, hand-label presence flag
//   seg_scan_kernel    per-frame exclusive scans -> segment offsets, entity counts / offsets
//   seg_scatter_kernel recomputes the classification, ranks points inside the warp with ballots and
//                      writes transformed float64 SoA points + raster indices
#pragma once
#include "common.cuh"

__device__ __forceinline__ bool coord_valid(float v) { return !(v != v) && v != -100.0f; }

// np.nan_to_num also maps +-inf to +-DBL_MAX (:1445)
__device__ __forceinline__ double widen(float v)
{
    double w = (double)v;
    if (isinf(w)) w = copysign(1.7976931348623157e308, w);
    return w;
}

struct PointClass {
    int e;        // entity index, or -1 (untracked label or invalid point)
    bool hand;    // the raw label equals hand_label (:1429 looks at the un-replaced label image)
};

__device__ __forceinline__ PointClass classify_point(const Dev &d, int f, int p, float x, float y, float z)
{
    int src = p;
    if (d.rowmap) {
        int r = p / d.width, c = p - r * d.width;
        src = __ldg(d.rowmap + r) * d.label_w + __ldg(d.colmap + c);
    }
    int raw = __ldg(d.labels + (size_t)f * d.label_size + src);
    PointClass pc;
    pc.hand = (raw == d.hand_label);
    int e = (raw >= 0 && raw < E2SEC_LABEL_LUT) ? __ldg(d.lut + raw) : -1;
    bool valid = coord_valid(x) && coord_valid(y) && coord_valid(z);
    pc.e = valid ? e : -1;
    return pc;
}

// q = M * [p, 1]; p' = q[:3] / q[3]   (Open3D PointCloud::Transform, operation order pinned in the oracle)
__device__ __forceinline__ void transform_point(const double *M, double &x, double &y, double &z)
{
    double q0 = ((M[0] * x + M[1] * y) + M[2] * z) + M[3];
    double q1 = ((M[4] * x + M[5] * y) + M[6] * z) + M[7];
    double q2 = ((M[8] * x + M[9] * y) + M[10] * z) + M[11];
    double q3 = ((M[12] * x + M[13] * y) + M[14] * z) + M[15];
    x = q0 / q3;
    y = q1 / q3;
    z = q2 / q3;
}

__global__ void __launch_bounds__(32 * E2_SEG_WARPS) seg_count_kernel(Dev d)
{
    pdl_enter();
    __shared__ unsigned wcnt[E2_SEG_WARPS][E2SEC_MAX_ENTITIES];
    const int f = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * E2_SEG_WARPS + warp;
    for (int e = lane; e < d.K; e += 32) wcnt[warp][e] = 0;
    __syncwarp();
    if (s >= d.S) return;
    const float *xyz = d.xyz + (size_t)f * d.N * 3;
    bool hand_seen = false;
    const int base = s * E2_SEG;
#pragma unroll 4
    for (int it = 0; it < E2_SEG / 32; ++it) {
        int p = base + it * 32 + lane;
        int e = -1;
        if (p < d.N) {
            float x = __ldg(xyz + 3 * (size_t)p), y = __ldg(xyz + 3 * (size_t)p + 1), z = __ldg(xyz + 3 * (size_t)p + 2);
            PointClass pc = classify_point(d, f, p, x, y, z);
            e = pc.e;
            hand_seen |= pc.hand;
        }
        unsigned todo = __ballot_sync(0xffffffffu, e >= 0);
        while (todo) {
            int leader = __ffs(todo) - 1;
            int ev = __shfl_sync(0xffffffffu, e, leader);
            unsigned m = __ballot_sync(0xffffffffu, e == ev);
            if (lane == leader) wcnt[warp][ev] += __popc(m);
            todo &= ~m;
        }
    }
    __syncwarp();
    unsigned *out = d.seg_cnt + ((size_t)f * d.S + s) * d.K;
    for (int e = lane; e < d.K; e += 32) out[e] = wcnt[warp][e];
    if (__any_sync(0xffffffffu, hand_seen) && lane == 0) atomicOr(d.flags + f, E2SEC_FLAG_HAND_LABEL);
}

// one CTA per frame, thread e scans entity e over the S segments
__global__ void seg_scan_kernel(Dev d)
{
    pdl_enter();
    __shared__ int tot[E2SEC_MAX_ENTITIES];
    const int f = blockIdx.x, e = threadIdx.x;
    if (e < d.K) {
        unsigned *c = d.seg_cnt + (size_t)f * d.S * d.K + e;
        unsigned run = 0;
        for (int s = 0; s < d.S; ++s) {
            unsigned v = c[(size_t)s * d.K];
            c[(size_t)s * d.K] = run;
            run += v;
        }
        tot[e] = (int)run;
        d.ent_raw[f * d.K + e] = (int)run;
    }
    __syncthreads();
    if (e == 0) {
        int off = 0;
        for (int k = 0; k < d.K; ++k) {
            bool stored = (k != 0 && k != d.ground_idx);   // :1520 -- index 0 and the ground label are never filtered nor related
            d.ent_off[f * d.K + k] = stored ? off : -1;
            if (stored) off += tot[k];
        }
        d.frame_used[f] = off < d.cap ? off : d.cap;
        if (off > d.cap) atomicExch(d.error, E2SEC_ERR_CAPACITY);
    }
}

__global__ void __launch_bounds__(32 * E2_SEG_WARPS) seg_scatter_kernel(Dev d)
{
    pdl_enter();
    __shared__ unsigned wcnt[E2_SEG_WARPS][E2SEC_MAX_ENTITIES];
    const int f = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * E2_SEG_WARPS + warp;
    if (s >= d.S) return;
    {
        const unsigned *so = d.seg_cnt + ((size_t)f * d.S + s) * d.K;
        const int *eo = d.ent_off + f * d.K;
        for (int e = lane; e < d.K; e += 32) wcnt[warp][e] = (eo[e] < 0) ? 0xffffffffu : (unsigned)eo[e] + so[e];
    }
    __syncwarp();
    const float *xyz = d.xyz + (size_t)f * d.N * 3;
    const size_t fb = (size_t)f * d.cap;
    const unsigned lt = (1u << lane) - 1u;
    const int base = s * E2_SEG;
    for (int it = 0; it < E2_SEG / 32; ++it) {
        int p = base + it * 32 + lane;
        int e = -1;
        float x = 0.f, y = 0.f, z = 0.f;
        if (p < d.N) {
            x = __ldg(xyz + 3 * (size_t)p);
            y = __ldg(xyz + 3 * (size_t)p + 1);
            z = __ldg(xyz + 3 * (size_t)p + 2);
            e = classify_point(d, f, p, x, y, z).e;
            if (e >= 0 && (e == 0 || e == d.ground_idx)) e = -1;   // counted in pass 1, never stored
        }
        unsigned todo = __ballot_sync(0xffffffffu, e >= 0);
        unsigned pos = 0;
        while (todo) {
            int leader = __ffs(todo) - 1;
            int ev = __shfl_sync(0xffffffffu, e, leader);
            unsigned m = __ballot_sync(0xffffffffu, e == ev);
            unsigned b = 0;
            if (lane == leader) {
                b = wcnt[warp][ev];
                wcnt[warp][ev] = b + __popc(m);
            }
            b = __shfl_sync(0xffffffffu, b, leader);
            if (e == ev) pos = b + __popc(m & lt);
            todo &= ~m;
        }
        if (e >= 0 && pos < (unsigned)d.cap) {
            double X = widen(x), Y = widen(y), Z = widen(z);
            transform_point(d.T, X, Y, Z);
            transform_point(d.R, X, Y, Z);
            d.px[fb + pos] = X;
            d.py[fb + pos] = Y;
            d.pz[fb + pos] = Z;
            d.pidx[fb + pos] = (unsigned)p;
        }
    }
}
