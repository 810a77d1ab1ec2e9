// Dissimilarity of e2SEC / eSEC tables: eSEC_analyser.py _make_triple (:866-889),
// _compare_triple_matrics (:915-944), _calc_D_shaped (:819-841), similarity_manipulations (:843-864).
//
// A table has rows = 3 * opal relation rows (T/N, SSR, DSR blocks of opal rows each) and n columns of short codes.  The
// reference turns every column into opal triples (T/N_r, SSR_r, DSR_r), pads the shorter of two tables by repeating its last
// column, and scores a pair of tables as  sum over triples of sqrt(l)/sqrt(3) / (n * divisor),  l = components that differ.
// Here a table is packed once into 64-bit words (one per code, columns past the end already holding the last column), and one
// warp per pair of tables counts the triples with l = 1, 2, 3; the score follows from the three integers.
#pragma once
#include "common.cuh"

#define E2_DS_THREADS 256
#define E2_DS_WARPS (E2_DS_THREADS / 32)
#define E2_DS_JTILE 32            // tables per CTA on the j side (4 per warp)
#define E2_DS_CODE 6              // bytes per code in a table (e2sec_resolver_table's format)

struct DissimDev {
    const unsigned char *codes;   // concatenated tables; table m holds rows x ncols[m] x 6 bytes from off[m]
    const long long *off;
    const int *ncols;
    int M, rows, opal, Lmax, E;   // E = Lmax * opal words per component
    u64 *packed;                  // [M][3][E]: component k of triple row r at column c sits at k * E + c * opal + r
    unsigned *counts;             // [M][M][3]: triples differing in exactly 1 / 2 / 3 components
    double *D;                    // [M][M]
    double divisor;
};

// numpy's chararray compares with trailing whitespace removed (the triples live in a chararray, :877)
__device__ __forceinline__ bool dissim_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }

__global__ void __launch_bounds__(E2_DS_THREADS) dissim_pack_kernel(DissimDev d)
{
    const int m = blockIdx.x;
    const int n = d.ncols[m];
    const unsigned char *tab = d.codes + d.off[m];
    u64 *out = d.packed + (size_t)m * 3 * d.E;
    for (int idx = threadIdx.x; idx < 3 * d.E; idx += E2_DS_THREADS) {
        const int k = idx / d.E, e = idx - k * d.E;
        const int c = e / d.opal, r = e - c * d.opal;
        const int cc = c < n ? c : n - 1;                                    // :931-932 / :939-940: the last column repeats
        const unsigned char *src = tab + ((size_t)(k * d.opal + r) * n + cc) * E2_DS_CODE;
        int len = 0;
        while (len < E2_DS_CODE && src[len]) ++len;
        while (len > 0 && dissim_space(src[len - 1])) --len;
        u64 w = 0;
        for (int b = 0; b < len; ++b) w |= (u64)src[b] << (8 * b);
        out[idx] = w;
    }
}

// grid (ceil(M / JTILE), M): CTA (jt, i) keeps table i in shared memory, its warps walk the tables j >= i of the tile.
__global__ void __launch_bounds__(E2_DS_THREADS) dissim_pairs_kernel(DissimDev d)
{
    extern __shared__ u64 ds_tab[];            // [3][E] of table i
    const int i = blockIdx.y;
    const int j0 = blockIdx.x * E2_DS_JTILE;
    const int j1 = min(j0 + E2_DS_JTILE, d.M);
    if (j1 <= i) return;                       // the score is symmetric: only j >= i is computed, both halves are written
    const u64 *ti = d.packed + (size_t)i * 3 * d.E;
    for (int idx = threadIdx.x; idx < 3 * d.E; idx += E2_DS_THREADS) ds_tab[idx] = ti[idx];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ni = d.ncols[i];
    const double a1 = sqrt(1.0) / sqrt(3.0), a2 = sqrt(2.0) / sqrt(3.0);       // :837
    for (int j = j0 + warp; j < j1; j += E2_DS_WARPS) {
        if (j < i) continue;
        const int n = max(ni, d.ncols[j]);
        const u64 *tj = d.packed + (size_t)j * 3 * d.E;
        unsigned c1 = 0, c2 = 0, c3 = 0;
        for (int e = lane; e < n * d.opal; e += 32) {
            const int l = (int)(ds_tab[e] != tj[e]) + (int)(ds_tab[d.E + e] != tj[d.E + e]) +
                          (int)(ds_tab[2 * d.E + e] != tj[2 * d.E + e]);
            c1 += l == 1; c2 += l == 2; c3 += l == 3;
        }
        c1 = __reduce_add_sync(0xffffffffu, c1);
        c2 = __reduce_add_sync(0xffffffffu, c2);
        c3 = __reduce_add_sync(0xffffffffu, c3);
        if (lane == 0) {
            const double s = ((double)c1 * a1 + (double)c2 * a2) + (double)c3;
            const double v = (1.0 / ((double)n * d.divisor)) * s;              // :838 / :860
            const size_t ij = (size_t)i * d.M + j, ji = (size_t)j * d.M + i;
            d.D[ij] = v; d.D[ji] = v;
            if (d.counts) {
                d.counts[3 * ij] = c1; d.counts[3 * ij + 1] = c2; d.counts[3 * ij + 2] = c3;
                d.counts[3 * ji] = c1; d.counts[3 * ji + 1] = c2; d.counts[3 * ji + 2] = c3;
            }
        }
    }
}
