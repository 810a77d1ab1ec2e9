// lzf.cuh -- input decode on the device: what o3d.io.read_point_cloud (manipulation_analyser.py:1441)
// and pd.read_csv (:1391) do on the host in the reference.
//
//   lzf_decode_kernel    PCD v0.7 "binary_compressed": the payload is ONE LZF stream (Marc Lehmann's format as PCL uses it:
//                        control byte c < 32 -> c + 1 literal bytes; otherwise a back-reference of length (c >> 5) + 2 -- 7
//                        escapes to one more byte -- at distance ((c & 31) << 8 | next) + 1, window 8 KB) over the fields in
//                        SoA order: all x, all y, all z, (rgba ...).  A stream is byte-serial -- every token's place depends on
//                        all tokens before it -- so the parallelism is across the frames of a batch: one warp per frame walks
//                        its tokens, the lanes copy a token's bytes together.  The last 16 KB of output live in a shared-memory
//                        ring (back-references never leave it), the compressed bytes come through a second small ring, and every
//                        completed 4 KB of the x / y / z planes is written out as interleaved float32 xyz -- the layout the
//                        segmentation kernels read -- so no second pass touches the cloud.  Decoding stops after the z plane.
//   label_text_kernel    the label image as whitespace-separated numbers: one CTA per frame, threads take slices of the text,
//                        count the numbers that START in their slice, a block scan gives every number its index.
#pragma once
#include "common.cuh"

#define E2_LZ_RING 16384              // output history per warp (the format's window is 8192)
#define E2_LZ_FLUSH 4096              // bytes written out at once
#define E2_LZ_IN 2048                 // compressed-input ring
#define E2_LZ_IN_HALF (E2_LZ_IN / 2)

struct LzfFrame {
    long long payload;                // offset of the LZF stream inside the blob
    unsigned comp;                    // its length
    unsigned need;                    // uncompressed bytes wanted: 3 planes of 4 * n_points
};

// grid = frames, 32 threads.  xyz: F x n_points x 3 float32.  status[f]: 0 ok, 1 malformed stream.
__global__ void __launch_bounds__(32) lzf_decode_kernel(const unsigned char *blob, const LzfFrame *frames, float *xyz, int n_points,
                                                        int *status, int *error)
{
    pdl_enter();
    __shared__ __align__(16) unsigned char ring[E2_LZ_RING];
    __shared__ __align__(16) unsigned char inb[E2_LZ_IN];
    const int lane = threadIdx.x, f = blockIdx.x;
    const LzfFrame fr = frames[f];
    const unsigned char *src = blob + fr.payload;
    // the input ring is indexed by ABSOLUTE address so that refills are aligned 16-byte loads whatever the header length
    const unsigned long long abs0 = (unsigned long long)src;
    const unsigned long long abs_end = abs0 + fr.comp;
    unsigned long long loaded = abs0 & ~(unsigned long long)(E2_LZ_IN_HALF - 1);      // everything below is in the ring (or past)
    auto refill = [&](unsigned long long upto) {                  // make [.., upto) resident, one half ring at a time
        while (loaded < upto && loaded < abs_end + 64) {
            const uint4 *g = reinterpret_cast<const uint4 *>(loaded);
            uint4 *s = reinterpret_cast<uint4 *>(inb + (loaded & (E2_LZ_IN - 1)));
            for (int i = lane; i < E2_LZ_IN_HALF / 16; i += 32) s[i] = __ldcs(g + i);
            loaded += E2_LZ_IN_HALF;
        }
        __syncwarp();
    };
    float *out = xyz + (size_t)f * n_points * 3;
    const unsigned need = fr.need;
    auto flush = [&](unsigned from, unsigned to) {                // plane bytes [from, to) -> interleaved xyz (4-byte words)
        for (unsigned w = from / 4 + lane; w < to / 4; w += 32) {
            const unsigned v = *reinterpret_cast<const unsigned *>(ring + ((w * 4) & (E2_LZ_RING - 1)));
            const unsigned plane = w / (unsigned)n_points, idx = w - plane * (unsigned)n_points;
            reinterpret_cast<unsigned *>(out)[(size_t)idx * 3 + plane] = v;
        }
    };
    unsigned long long ip = abs0;
    unsigned op = 0, flushed = 0;
    bool bad = false;
    // (the ring holds [loaded - 2048, loaded); refilling to at most ip + 552 + 1023 keeps the unread bytes from ip on inside it)
    refill(ip + 552);
    while (op < need && ip < abs_end) {
        if (ip + 40 > loaded) refill(ip + 552);                   // a token is at most 34 bytes long
        // the three header bytes a token can have, loaded together (one shared-memory latency instead of three in a row)
        const unsigned ctrl = inb[ip & (E2_LZ_IN - 1)];
        const unsigned h1 = inb[(ip + 1) & (E2_LZ_IN - 1)], h2 = inb[(ip + 2) & (E2_LZ_IN - 1)];
        if (ctrl < 32u) {
            const unsigned n = ctrl + 1u;
            if (ip + 1 + n > abs_end) { bad = true; break; }
            if (lane < n) ring[(op + lane) & (E2_LZ_RING - 1)] = inb[(ip + 1 + lane) & (E2_LZ_IN - 1)];
            ip += 1 + n;
            op += n;
        } else {
            unsigned len = ctrl >> 5;
            const bool ext = len == 7u;
            if (ext) len += h1;
            const unsigned long long q = ip + (ext ? 2 : 1);
            if (q >= abs_end) { bad = true; break; }
            const unsigned dist = (((ctrl & 31u) << 8) | (ext ? h2 : h1)) + 1u;
            len += 2u;
            if (dist > op) { bad = true; break; }
            const unsigned ref = op - dist;
            // out[op + i] = out[ref + i], byte after byte: with dist < len the source repeats with period dist
            if (dist >= len) {                                    // the usual case: no overlap, no division
                for (unsigned i = lane; i < len; i += 32) ring[(op + i) & (E2_LZ_RING - 1)] = ring[(ref + i) & (E2_LZ_RING - 1)];
            } else {
                for (unsigned i = lane; i < len; i += 32) ring[(op + i) & (E2_LZ_RING - 1)] = ring[(ref + i % dist) & (E2_LZ_RING - 1)];
            }
            ip = q + 1;
            op += len;
        }
        __syncwarp();
        if (op - flushed >= E2_LZ_FLUSH + 264u) {                 // keep 8 KB + the longest token of history behind the flush point
            const unsigned to = min((op - 264u) & ~(unsigned)(E2_LZ_FLUSH - 1), need);
            if (to > flushed) { flush(flushed, to); flushed = to; }
            __syncwarp();
        }
    }
    if (!bad && op < need) bad = true;                            // the stream ended early
    if (!bad) flush(flushed, need);
    if (lane == 0) {
        status[f] = bad ? 1 : 0;
        if (bad) atomicExch(error, E2SEC_ERR_IO);
    }
}

// grid = frames, 256 threads.  text: the frames' label files back to back, off[f] .. off[f + 1]; labels: F x (rows * cols)
// int32.  Numbers are non-negative integers, optionally with a fraction of zeros ("7" or "7.0"); anything else, or a
// number count different from rows * cols, raises E2SEC_ERR_IO.
#define E2_LT_THREADS 256
__global__ void __launch_bounds__(E2_LT_THREADS) label_text_kernel(const unsigned char *text, const long long *off, int *labels, int n_labels,
                                                                   int *status, int *error)
{
    pdl_enter();
    __shared__ unsigned wtot[E2_LT_THREADS / 32];
    __shared__ int bad_s;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned char *t = text + off[f];
    const long long n = off[f + 1] - off[f];
    const long long per = (n + E2_LT_THREADS - 1) / E2_LT_THREADS;
    const long long a = min(n, per * tid), b = min(n, a + per);
    if (tid == 0) bad_s = 0;
    __syncthreads();
    auto is_num = [](unsigned char c) { return (c >= '0' && c <= '9') || c == '.'; };
    unsigned starts = 0;
    for (long long i = a; i < b; ++i) {
        const unsigned char c = t[i];
        if (is_num(c) && (i == 0 || !is_num(t[i - 1]))) ++starts;
        else if (!is_num(c) && c != ' ' && c != '\n' && c != '\r' && c != '\t') bad_s = 1;
    }
    unsigned incl = starts;
    for (int o = 1; o < 32; o <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    unsigned idx = incl - starts, total = 0;
    for (int w = 0; w < E2_LT_THREADS / 32; ++w) { if (w < warp) idx += wtot[w]; total += wtot[w]; }
    int *out = labels + (size_t)f * n_labels;
    for (long long i = a; i < b; ++i) {
        if (!(is_num(t[i]) && (i == 0 || !is_num(t[i - 1])))) continue;
        // the number that starts here may run past the slice: follow it to its end
        long long v = 0, j = i;
        while (j < n && t[j] >= '0' && t[j] <= '9') { v = v * 10 + (t[j] - '0'); if (v > 0x7fffffff) v = 0x7fffffff; ++j; }
        if (j < n && t[j] == '.') { ++j; while (j < n && is_num(t[j])) { if (t[j] != '0') bad_s = 1; ++j; } }
        if (idx < (unsigned)n_labels) out[idx] = (int)v;
        ++idx;
    }
    __syncthreads();
    if (tid == 0) {
        const bool bad = bad_s || total != (unsigned)n_labels;
        status[f] = bad ? 1 : 0;
        if (bad) atomicExch(error, E2SEC_ERR_IO);
    }
}
