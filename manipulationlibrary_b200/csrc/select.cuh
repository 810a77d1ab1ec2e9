// select.cuh -- exact order statistics inside one CTA, no sorting and no atomics.
//
// block_select2 finds the k1-th and k2-th smallest (0-based, k2 == k1 or k1 + 1) of n 64-bit keys by a
// most-significant-digit descent with 4-bit digits.  Each thread counts its elements in thread-private
// shared-memory counters laid out [digit][thread] (bank = thread, conflict free, plain LDS/STS -- shared
// atomics cost ~2 cycles per lane on this SM), the 16 rows are then summed and the digit holding the rank
// is chosen.  Used for np.median(objects[i], axis=0) (manipulation_analyser.py:1523) and for the final
// step of np.median(pdist()) (:1524) over the collected candidates.
#pragma once
#include "common.cuh"

#define E2_SEL_THREADS 256

struct SelectSmem {
    unsigned cnt[16 * E2_SEL_THREADS];
    u64 row[16];
    u64 prefix, rank, aux;
    unsigned long long red[E2_SEL_THREADS / 32];
    unsigned long long red2[E2_SEL_THREADS / 32];
};

template <class Get>
__device__ void block_select2(Get get, int n, u64 k1, u64 k2, u64 &v1, u64 &v2, SelectSmem &sm)
{
    const int tid = threadIdx.x;
    if (tid == 0) { sm.prefix = 0; sm.rank = k1; }
    __syncthreads();
    for (int shift = 60; shift >= 0; shift -= 4) {
        for (int dgt = 0; dgt < 16; ++dgt) sm.cnt[dgt * E2_SEL_THREADS + tid] = 0;
        const u64 prefix = sm.prefix;
        const u64 himask = (shift == 60) ? 0ull : (~0ull << (shift + 4));
        for (int i = tid; i < n; i += E2_SEL_THREADS) {
            u64 k = get(i);
            if ((k & himask) == prefix) {
                unsigned dgt = (unsigned)(k >> shift) & 15u;
                sm.cnt[dgt * E2_SEL_THREADS + tid] += 1;
            }
        }
        __syncthreads();
        if (tid < 16 * 8) {   // 8 threads per digit row, rotated start to spread banks
            int dgt = tid >> 3, part = tid & 7;
            u64 s = 0;
            for (int j = 0; j < E2_SEL_THREADS / 8; ++j) s += sm.cnt[dgt * E2_SEL_THREADS + part * (E2_SEL_THREADS / 8) + ((j + tid) & (E2_SEL_THREADS / 8 - 1))];
            for (int o = 4; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, 8);
            if (part == 0) sm.row[dgt] = s;
        }
        __syncthreads();
        if (tid == 0) {
            u64 r = sm.rank, acc = 0;
            int dgt = 0;
            for (; dgt < 15; ++dgt) {
                if (r < acc + sm.row[dgt]) break;
                acc += sm.row[dgt];
            }
            sm.rank = r - acc;
            sm.prefix = prefix | ((u64)dgt << shift);
        }
        __syncthreads();
    }
    v1 = sm.prefix;
    if (k2 == k1) { v2 = v1; return; }
    // k2 = k1 + 1: v2 = v1 if more than k2 keys are <= v1, else the smallest key above v1
    u64 le = 0, mn = ~0ull;
    for (int i = tid; i < n; i += E2_SEL_THREADS) {
        u64 k = get(i);
        if (k <= v1) ++le;
        else if (k < mn) mn = k;
    }
    le = warp_sum_u64(le);
    mn = warp_min_u64(mn);
    if ((tid & 31) == 0) { sm.red[tid >> 5] = le; sm.red2[tid >> 5] = mn; }
    __syncthreads();
    if (tid == 0) {
        u64 a = 0, b = ~0ull;
        for (int w = 0; w < E2_SEL_THREADS / 32; ++w) { a += sm.red[w]; b = sm.red2[w] < b ? sm.red2[w] : b; }
        sm.aux = (a > k2) ? v1 : b;
    }
    __syncthreads();
    v2 = sm.aux;
    __syncthreads();
}

// np.median(objects[i], axis=0) for every filtered entity: grid (3, F*K), E2_SEL_THREADS threads.
__global__ void __launch_bounds__(E2_SEL_THREADS) axis_median_kernel(Dev d)
{
    pdl_enter();
    __shared__ SelectSmem sm;
    const int fe = blockIdx.y, axis = blockIdx.x;
    const int f = fe / d.K, e = fe - f * d.K;
    const int off = d.ent_off[fe], n = d.ent_raw[fe];
    if (off < 0 || !entity_filtered(d, fe, n)) return;    // :1520 -- only clouds with more than 3 points are filtered
    const double *src = (axis == 0 ? d.px : axis == 1 ? d.py : d.pz) + (size_t)f * d.cap + off;
    u64 v1, v2;
    u64 k2 = (u64)(n / 2), k1 = (n & 1) ? k2 : k2 - 1;
    block_select2([&](int i) { return f64_sortkey(src[i]); }, n, k1, k2, v1, v2, sm);
    if (threadIdx.x == 0) {
        double a = f64_from_sortkey(v1), b = f64_from_sortkey(v2);
        // numpy: mean of the two middle elements = (a + b) / 2; a single middle element is returned as is
        d.med[(size_t)fe * 3 + axis] = (n & 1) ? a : (a + b) / 2.0;
    }
    (void)e;
}
