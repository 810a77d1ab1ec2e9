// frame0.cuh -- the geometry of the frame-0 setup on the device: what the reference does once per
// sequence with Open3D (manipulation_analyser.py):
//   _getGroundiNewNew :278-361   pcd.segment_plane(0.01, 3, 19)  RANSAC plane of the ground label's points
//                                .voxel_down_sample(0.02)        one point per occupied voxel: the mean of its points
//   _getTranslation   :212-276   voxel_down_sample again, registration_icp(ground, synthetic plane, 0.2) point to point
// Open3D's definitions are restated in oracle/fake_o3d.py (the version the golden fixture was produced with); the kernels
// here follow the same arithmetic so that the fixture is reproduced (tests/test_frame0_device.py):
//   plane_count_kernel      |p . n + d| < threshold for every point and all hypotheses of the RANSAC at once
//   plane_mask_kernel       the inliers of the winning plane
//   voxel_key_kernel + thrust::stable_sort_by_key + voxel_mean_kernel
//                           voxel index floor((p - origin) / voxel), origin = min bound - voxel / 2; the points of a voxel are
//                           summed in their original order (what np.bincount does), voxels come out in ascending (ix, iy, iz)
//   icp_nn_kernel, icp_accumulate_kernel
//                           exact nearest neighbour of every transformed source point within max_distance (brute force, the
//                           clouds are a few thousand voxel centres), then count, sum of d^2, both centroids and the
//                           centred 3 x 3 cross-covariance of the matched pairs; the 3 x 3 SVD stays on the host
#pragma once
#include <thrust/device_ptr.h>
#include <thrust/execution_policy.h>
#include <thrust/sort.h>

#include "common.cuh"

#define E2_F0_THREADS 256
#define E2_F0_MAXPLANES 64

__global__ void __launch_bounds__(E2_F0_THREADS) plane_count_kernel(const double *pts, int n, const double *planes, int m, double thr,
                                                                    int *counts)
{
    __shared__ double pl[E2_F0_MAXPLANES][4];
    __shared__ int cnt[E2_F0_MAXPLANES];
    for (int i = threadIdx.x; i < m * 4; i += E2_F0_THREADS) pl[i / 4][i % 4] = planes[i];
    for (int i = threadIdx.x; i < m; i += E2_F0_THREADS) cnt[i] = 0;
    __syncthreads();
    for (int i = blockIdx.x * E2_F0_THREADS + threadIdx.x; i < n; i += gridDim.x * E2_F0_THREADS) {
        const double x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
        for (int h = 0; h < m; ++h) {
            const double dist = fabs(((x * pl[h][0] + y * pl[h][1]) + z * pl[h][2]) + pl[h][3]);
            if (dist < thr) atomicAdd(&cnt[h], 1);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < m; i += E2_F0_THREADS)
        if (cnt[i]) atomicAdd(&counts[i], cnt[i]);
}

__global__ void plane_mask_kernel(const double *pts, int n, double a, double b, double c, double d, double thr, unsigned char *mask)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    mask[i] = fabs(((x * a + y * b) + z * c) + d) < thr ? 1 : 0;
}

// key = (ix * ny + iy) * nz + iz: ascending keys = ascending (ix, iy, iz), the order np.unique(axis=0) gives
__global__ void voxel_key_kernel(const double *pts, int n, double ox, double oy, double oz, double voxel, long long ny, long long nz,
                                 unsigned long long *keys, int *index)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long ix = (long long)floor((pts[3 * (size_t)i] - ox) / voxel);
    const long long iy = (long long)floor((pts[3 * (size_t)i + 1] - oy) / voxel);
    const long long iz = (long long)floor((pts[3 * (size_t)i + 2] - oz) / voxel);
    keys[i] = (unsigned long long)((ix * ny + iy) * nz + iz);
    index[i] = i;
}

// one thread per sorted position that starts a voxel: sequential sum of the voxel's points in original order
__global__ void voxel_mean_kernel(const double *pts, int n, const unsigned long long *keys, const int *index, double *out, int *n_out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i > 0 && keys[i] == keys[i - 1]) return;
    // rank of this voxel = number of run starts before it: counted with a second pass by the caller (see voxel_rank_kernel)
    double sx = 0.0, sy = 0.0, sz = 0.0;
    int c = 0;
    for (int j = i; j < n && keys[j] == keys[i]; ++j) {
        const size_t p = (size_t)index[j];
        sx += pts[3 * p]; sy += pts[3 * p + 1]; sz += pts[3 * p + 2];
        ++c;
    }
    const int slot = n_out[i];                       // exclusive count of run starts before i (thrust::exclusive_scan)
    out[3 * (size_t)slot] = sx / c;
    out[3 * (size_t)slot + 1] = sy / c;
    out[3 * (size_t)slot + 2] = sz / c;
}

__global__ void voxel_start_kernel(const unsigned long long *keys, int n, int *start)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) start[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

// q = T p with the reference's operation order (PointCloud::Transform as oracle/fake_o3d.transform_points pins it)
__device__ __forceinline__ void f0_transform(const double *T, double x, double y, double z, double &qx, double &qy, double &qz)
{
    const double w = ((T[12] * x + T[13] * y) + T[14] * z) + T[15];
    qx = (((T[0] * x + T[1] * y) + T[2] * z) + T[3]) / w;
    qy = (((T[4] * x + T[5] * y) + T[6] * z) + T[7]) / w;
    qz = (((T[8] * x + T[9] * y) + T[10] * z) + T[11]) / w;
}

struct IcpParams { double T[16]; double max_d2; };

// one warp per source point: nearest target point (exact, float64), -1 when none within max distance
__global__ void __launch_bounds__(E2_F0_THREADS) icp_nn_kernel(const double *src, int ns, const double *tgt, int nt, IcpParams p,
                                                               int *nn, double *nn_d2, double *moved)
{
    const int w = (blockIdx.x * E2_F0_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= ns) return;
    double x, y, z;
    f0_transform(p.T, src[3 * (size_t)w], src[3 * (size_t)w + 1], src[3 * (size_t)w + 2], x, y, z);
    double best = 1e300;
    int bi = -1;
    for (int j = lane; j < nt; j += 32) {
        const double d2 = dist2(x, y, z, tgt[3 * (size_t)j], tgt[3 * (size_t)j + 1], tgt[3 * (size_t)j + 2]);
        if (d2 < best) { best = d2; bi = j; }
    }
    for (int o = 16; o; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ob < best || (ob == best && oi >= 0 && (bi < 0 || oi < bi))) { best = ob; bi = oi; }
    }
    if (lane == 0) {
        const bool ok = bi >= 0 && best <= p.max_d2;
        nn[w] = ok ? bi : -1;
        nn_d2[w] = best;
        moved[3 * (size_t)w] = x; moved[3 * (size_t)w + 1] = y; moved[3 * (size_t)w + 2] = z;
    }
}

// single CTA (the clouds are small): stats[0] matched pairs, [1] sum d^2, [2..4] centroid of the matched source points,
// [5..7] of their targets, [8..16] H = sum (a - ca)(b - cb)^T row-major.  Two passes: centroids, then the centred products.
__global__ void __launch_bounds__(1024) icp_accumulate_kernel(const double *moved, const double *tgt, const int *nn, const double *nn_d2, int ns,
                                                              double *stats)
{
    __shared__ double red[17][32];
    __shared__ double cen[7];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = tid; i < ns; i += 1024) {
        const int j = nn[i];
        if (j < 0) continue;
        v[0] += 1.0; v[1] += nn_d2[i];
        for (int k = 0; k < 3; ++k) { v[2 + k] += moved[3 * (size_t)i + k]; v[5 + k] += tgt[3 * (size_t)j + k]; }
    }
    for (int q = 0; q < 8; ++q) {
        double s = v[q];
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) red[q][warp] = s;
    }
    __syncthreads();
    if (tid < 8) {
        double s = 0.0;
        for (int w = 0; w < 32; ++w) s += red[tid][w];
        if (tid < 2) stats[tid] = s;
        else cen[tid - 2] = s;
        if (tid == 0) cen[6] = s;
    }
    __syncthreads();
    const double cnt = cen[6];
    double ca[3], cb[3];
    for (int k = 0; k < 3; ++k) { ca[k] = cnt > 0 ? cen[k] / cnt : 0.0; cb[k] = cnt > 0 ? cen[3 + k] / cnt : 0.0; }
    double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = tid; i < ns; i += 1024) {
        const int j = nn[i];
        if (j < 0) continue;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) h[3 * r + c] += (moved[3 * (size_t)i + r] - ca[r]) * (tgt[3 * (size_t)j + c] - cb[c]);
    }
    __syncthreads();
    for (int q = 0; q < 9; ++q) {
        double s = h[q];
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) red[q][warp] = s;
    }
    __syncthreads();
    if (tid < 9) {
        double s = 0.0;
        for (int w = 0; w < 32; ++w) s += red[tid][w];
        stats[8 + tid] = s;
    }
    if (tid < 3) { stats[2 + tid] = ca[tid]; stats[5 + tid] = cb[tid]; }
}
