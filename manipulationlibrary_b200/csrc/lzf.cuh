// lzf.cuh -- input decode on the device: what o3d.io.read_point_cloud (manipulation_analyser.py:1441)
// and pd.read_csv (:1391) do on the host in the reference.
//
//   lzf_decode_kernel    PCD v0.7 "binary_compressed": the payload is ONE LZF stream (Marc Lehmann's format as PCL uses it:
//                        control byte c < 32 -> c + 1 literal bytes; otherwise a back-reference of length (c >> 5) + 2 -- 7
//                        escapes to one more byte -- at distance ((c & 31) << 8 | next) + 1, window 8 KB) over the fields in
//                        SoA order: all x, all y, all z, (rgba ...).  A stream is byte-serial on paper -- every token's place
//                        depends on all tokens before it, every copied byte on earlier output -- and one warp owns one stream;
//                        but inside a WINDOW of 2 KB of input / 8 KB of output the warp works in parallel:
//                          1. every lane parses its own 64 input bytes as if they began on a token boundary; token streams
//                             re-synchronise within a few tokens, so a sweep over the lanes only has to repair the head of a
//                             lane's parse where it did not start on the true boundary;
//                          2. a scan of the tokens' output lengths places every token; literals and references into earlier
//                             windows are written at once, a reference into the window itself leaves a POINTER;
//                          3. pointer jumping resolves those (front to back, so a reference to anything 128 bytes or more
//                             behind is final after one round; run-length style references take a few more).
//                        The last 16 KB of output live in a shared-memory ring (back-references never leave it) and every
//                        completed part of the x / y / z planes is written out as interleaved float32 xyz -- the layout the
//                        segmentation kernels read -- so no second pass touches the cloud.  Decoding stops after the z plane.
//   label_text_kernel    the label image as whitespace-separated numbers: one CTA per frame, threads take slices of the text,
//                        count the numbers that START in their slice, a block scan gives every number its index.
#pragma once
#include "common.cuh"

#define E2_LZ_RING 16384              // output history per warp (the format's window is 8192) + the window being decoded
#define E2_LZ_OUTW 8192               // output bytes per window
#define E2_LZ_INW 2048                // input bytes per window: 32 lanes x E2_LZ_SEG
#define E2_LZ_SEG 64
#define E2_LZ_INBUF (E2_LZ_INW + 64 + 32)   // + the longest token that can start on the last byte, + alignment slack
#define E2_LZ_RES 0xffffu             // pointer value of a byte that is final

// -DE2_LZ_PROFILE: lane 0 adds the clock ticks of each stage of a window to g_lz_prof (tools/lzf_profile.py)
#ifdef E2_LZ_PROFILE
__device__ unsigned long long g_lz_prof[16];
#define LZ_TICK(slot) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_lz_prof[slot], (unsigned long long)(t_ - prof_t)); prof_t = t_; } } while (0)
#else
#define LZ_TICK(slot) do { } while (0)
#endif

struct LzfFrame {
    long long payload;                // offset of the LZF stream inside the blob
    unsigned comp;                    // its length
    unsigned need;                    // uncompressed bytes wanted: 3 planes of 4 * n_points
};

// input bytes of the token that starts at p
__device__ __forceinline__ unsigned lz_token_size(const unsigned char *in, unsigned p)
{
    const unsigned c = in[p];
    return c < 32u ? c + 2u : ((c >> 5) == 7u ? 3u : 2u);
}
// output bytes of the token that starts at p
__device__ __forceinline__ unsigned lz_token_out(const unsigned char *in, unsigned p)
{
    const unsigned c = in[p];
    return c < 32u ? c + 1u : ((c >> 5) == 7u ? 9u + in[p + 1] : (c >> 5) + 2u);
}

// grid = frames, 32 threads.  xyz: F x n_points x 3 float32.  status[f]: 0 ok, 1 malformed stream.
__global__ void __launch_bounds__(32) lzf_decode_kernel(const unsigned char *blob, const LzfFrame *frames, float *xyz, int n_points,
                                                        int *status, int *error)
{
    pdl_enter();
    __shared__ __align__(16) unsigned char ring[E2_LZ_RING];
    __shared__ __align__(16) unsigned char inb[E2_LZ_INBUF];
    __shared__ __align__(16) unsigned short src[E2_LZ_OUTW];
    __shared__ unsigned short tok_r[E2_LZ_INW / 2 + 2], tok_q[E2_LZ_INW / 2 + 2], tok_d[E2_LZ_INW / 2 + 2];   // the window's tokens: output position, input position, distance
    const int lane = threadIdx.x, f = blockIdx.x;
    const LzfFrame fr = frames[f];
    const unsigned long long abs0 = (unsigned long long)(blob + fr.payload), abs_end = abs0 + fr.comp;
    float *out = xyz + (size_t)f * n_points * 3;
    const unsigned need = fr.need;
    auto flush = [&](unsigned from, unsigned to) {                // plane bytes [from, to) -> interleaved xyz (4-byte words)
        for (unsigned w = from / 4 + lane; w < to / 4; w += 32) {
            const unsigned v = *reinterpret_cast<const unsigned *>(ring + ((w * 4) & (E2_LZ_RING - 1)));
            const unsigned plane = w / (unsigned)n_points, idx = w - plane * (unsigned)n_points;
            reinterpret_cast<unsigned *>(out)[(size_t)idx * 3 + plane] = v;
        }
    };
    unsigned long long ip = abs0;                                 // always the address of a token's control byte
    unsigned op = 0, flushed = 0;
    bool bad = false;
#ifdef E2_LZ_PROFILE
    long long prof_t = clock64();
#endif
    while (op < need) {
        if (ip >= abs_end) { bad = true; break; }                 // the stream ended early
        LZ_TICK(7);
        // ---- the window's input: aligned 16-byte loads, in[0] = the byte at ip -----------------------------------------
        const unsigned long long base = ip & ~15ull;
        for (int i = lane; i < E2_LZ_INBUF / 16; i += 32) {
            const unsigned long long a = base + 16ull * i;
            reinterpret_cast<uint4 *>(inb)[i] = a < abs_end ? __ldcs(reinterpret_cast<const uint4 *>(a)) : make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
        LZ_TICK(0);
        const unsigned char *in = inb + (unsigned)(ip - base);
        const unsigned avail = (unsigned)min((unsigned long long)(E2_LZ_INW + 64), abs_end - ip);    // bytes of the stream in the buffer
        const unsigned wlen = min((unsigned)E2_LZ_INW, avail);                                        // tokens of this window start below
        // ---- 1. token boundaries: every lane parses its segment from its first byte, then the sweep repairs the heads ----
        const unsigned seg0 = lane * E2_LZ_SEG, seg1 = min(seg0 + E2_LZ_SEG, wlen);
        unsigned long long mask = 0;                              // token starts of the segment, bit = offset
        unsigned exitp = seg0;                                    // where the first token after the segment starts
        if (seg0 < wlen) {
            unsigned q = seg0;
            while (q < seg1) { mask |= 1ull << (q - seg0); q += lz_token_size(in, q); }
            exitp = q;
        }
        LZ_TICK(1);
        for (int i = 1; i < 32; ++i) {
            const unsigned t = __shfl_sync(0xffffffffu, exitp, i - 1);       // the true boundary that lane i has to start from
            if (lane != i) continue;
            if (seg0 >= wlen || t >= seg1) { mask = 0; exitp = t; continue; }
            const unsigned rel = t - seg0;
            if ((mask >> rel) & 1ull) { mask &= ~((1ull << rel) - 1ull); continue; }     // already a boundary of this lane's parse
            unsigned long long nm = 0;
            unsigned q = t;
            for (;;) {
                nm |= 1ull << (q - seg0);
                q += lz_token_size(in, q);
                if (q >= seg1) { exitp = q; break; }
                if ((mask >> (q - seg0)) & 1ull) { nm |= mask & ~((1ull << (q - seg0)) - 1ull); break; }   // back in step
            }
            mask = nm;
        }
        __syncwarp();
        const unsigned next_window = __shfl_sync(0xffffffffu, exitp, 31);
        LZ_TICK(2);
        // ---- 2. place the tokens: output lengths, exclusive scans over the lanes, the window ends at E2_LZ_OUTW bytes -----
        unsigned lsum = 0, lcnt = 0;
        for (unsigned long long m = mask; m; m &= m - 1) { lsum += lz_token_out(in, seg0 + (unsigned)__ffsll((long long)m) - 1u); ++lcnt; }
        unsigned incl = lsum, tincl = lcnt;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, incl, o), w = __shfl_up_sync(0xffffffffu, tincl, o);
            if (lane >= o) { incl += v; tincl += w; }
        }
        // the window's token table, in stream order: output position, input position, distance (0: literal run)
        unsigned at0 = incl - lsum, t = tincl - lcnt, drop = 0xffffffffu, drop_t = 0xffffffffu;
        bool lbad = false;
        for (unsigned long long m = mask; m; m &= m - 1) {
            const unsigned q = seg0 + (unsigned)__ffsll((long long)m) - 1u;
            const unsigned c = in[q], size = lz_token_size(in, q), len = lz_token_out(in, q);
            if (at0 + len > E2_LZ_OUTW) { drop = q; drop_t = t; break; }   // this token opens the next window
            if (q + size > avail) { lbad = true; break; }                  // the token runs past the end of the stream
            const unsigned dist = c < 32u ? 0u : (((c & 31u) << 8) | in[q + size - 1]) + 1u;
            if (dist > op + at0) { lbad = true; break; }                   // a reference to before the start of the output
            tok_r[t] = (unsigned short)at0;
            tok_q[t] = (unsigned short)q;
            tok_d[t] = (unsigned short)dist;
            at0 += len;
            ++t;
        }
        if (__any_sync(0xffffffffu, lbad)) { bad = true; break; }
        const unsigned first_drop = __reduce_min_sync(0xffffffffu, drop);
        const unsigned first_drop_t = __reduce_min_sync(0xffffffffu, drop_t);
        // the end of the kept output: where the dropping lane stopped, else the total
        const unsigned total_out = __shfl_sync(0xffffffffu, incl, 31), total_tok = __shfl_sync(0xffffffffu, tincl, 31);
        const unsigned produced = __reduce_min_sync(0xffffffffu, drop != 0xffffffffu ? at0 : total_out);
        const unsigned ntok = first_drop_t != 0xffffffffu ? first_drop_t : total_tok;
        ip += (first_drop != 0xffffffffu) ? first_drop : next_window;
        LZ_TICK(3);
        // ---- the token of every output byte: token starts are marked in src ------------------------------------------------------
        for (unsigned i = lane; i < (produced + 7) / 8; i += 32) reinterpret_cast<uint4 *>(src)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
        __syncwarp();
        for (unsigned k = lane; k < ntok; k += 32) src[tok_r[k]] = (unsigned short)k;
        __syncwarp();
        // ---- the bytes, 128 per step.  The token of a byte is the last mark at or before it: a max-scan over the 32 bytes of a
        // row (token indices grow along the stream) plus the carry of the rows before.  Literals and references into earlier
        // windows are final at once, a reference into the window leaves a pointer (src is rewritten in place: mark -> pointer)
        unsigned carry = 1;                                       // token index + 1 of the last mark so far (byte 0 starts token 0)
        for (unsigned r0 = 0; r0 < produced; r0 += 128) {
            unsigned tr[4], tq[4], td[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned at = r0 + 32u * j + lane;
                const unsigned mk = at < produced ? (unsigned)src[at] : E2_LZ_RES;
                unsigned v = mk == E2_LZ_RES ? 0u : mk + 1u;
                for (int o = 1; o < 32; o <<= 1) { const unsigned w = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v = max(v, w); }
                v = max(v, carry);
                carry = __shfl_sync(0xffffffffu, v, 31);
                const unsigned k = min(v - 1u, (unsigned)(E2_LZ_INW / 2));     // (the clamp never acts; it keeps a logic error from faulting)
                tr[j] = tok_r[k]; tq[j] = tok_q[k]; td[j] = tok_d[k];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned at = r0 + 32u * j + lane;
                if (at >= produced) continue;
                if (td[j] == 0u) { ring[(op + at) & (E2_LZ_RING - 1)] = in[tq[j] + 1u + (at - tr[j])]; src[at] = E2_LZ_RES; }
                else if (td[j] > at) { ring[(op + at) & (E2_LZ_RING - 1)] = ring[(op + at - td[j]) & (E2_LZ_RING - 1)]; src[at] = E2_LZ_RES; }
                else src[at] = (unsigned short)(at - td[j]);
            }
        }
        __syncwarp();
        LZ_TICK(4);
        // ---- 3. pointer jumping, front to back, 128 bytes per step: everything before a step is final, so a step is
        // repeated until its own bytes are (read everything, then write; a pointer only ever points backwards)
        for (unsigned r0 = 0; r0 < produced; r0 += 128) {
            for (;;) {
                unsigned sv[4], s2[4];
                unsigned char bv[4];
                bool open = false;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const unsigned at = r0 + 32u * j + lane;
                    sv[j] = at < produced ? (unsigned)src[at] : E2_LZ_RES;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    s2[j] = E2_LZ_RES; bv[j] = 0;
                    if (sv[j] != E2_LZ_RES) { s2[j] = src[sv[j] & (E2_LZ_OUTW - 1)]; bv[j] = ring[(op + sv[j]) & (E2_LZ_RING - 1)]; }
                }
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (sv[j] == E2_LZ_RES) continue;
                    const unsigned at = r0 + 32u * j + lane;
                    if (s2[j] == E2_LZ_RES) { ring[(op + at) & (E2_LZ_RING - 1)] = bv[j]; src[at] = E2_LZ_RES; }
                    else { src[at] = (unsigned short)s2[j]; open = true; }
                }
                __syncwarp();
                if (!__any_sync(0xffffffffu, open)) break;
            }
        }
        LZ_TICK(5);
        op += produced;
        const unsigned to = min(op & ~3u, need);
        if (to > flushed) { flush(flushed, to); flushed = to; }
        __syncwarp();
        LZ_TICK(6);
#ifdef E2_LZ_PROFILE
        if (lane == 0) atomicAdd(&g_lz_prof[8], 1ull);
#endif
    }
    if (!bad && flushed < need) { flush(flushed, need); }
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
