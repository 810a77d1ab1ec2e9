/* TEST INFRASTRUCTURE ONLY (oracle side): LZF block decoder used by oracle/pcd_io.py
 * to read PCD v0.7 "binary_compressed" files, i.e. what
 * o3d.io.read_point_cloud does at manipulation_analyser.py:1441.
 * Restates the published LZF format (Marc Lehmann's liblzf, as used by PCL's
 * pcl::lzfDecompress): control byte c < 32 -> literal run of c+1 bytes;
 * otherwise a back-reference with length (c>>5)+2 (length 7 escapes to an
 * extra byte) at distance ((c&0x1f)<<8 | next)+1.
 * The product's reader is a separate implementation in
 * manipulationlibrary_b200/csrc/pcd_reader.cpp; tests cross-check the two.
 */
#include <stddef.h>
#include <stdint.h>

long oracle_lzf_decompress(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_len)
{
    const uint8_t *ip = in, *in_end = in + in_len;
    uint8_t *op = out, *out_end = out + out_len;
    while (ip < in_end) {
        unsigned ctrl = *ip++;
        if (ctrl < 32) {
            ctrl++;
            if (op + ctrl > out_end || ip + ctrl > in_end) return -1;
            while (ctrl--) *op++ = *ip++;
        } else {
            unsigned len = ctrl >> 5;
            if (ip >= in_end) return -2;
            if (len == 7) { len += *ip++; if (ip >= in_end) return -2; }
            const uint8_t *ref = op - ((ctrl & 0x1f) << 8) - 1 - *ip++;
            if (ref < out) return -3;
            len += 2;
            if (op + len > out_end) return -1;
            while (len--) *op++ = *ref++;
        }
    }
    return (long)(op - out);
}

/* TEST INFRASTRUCTURE ONLY: an LZF *compressor* so that tests and bench.py can write PCD v0.7 "binary_compressed" files of
 * synthetic clouds (what PCL's writer does for the MANIAC recordings).  Greedy, 3-byte hash, 8 KB window, maximum match
 * 264: the published format only -- any valid LZF stream decodes the same, it need not equal liblzf's choices. */
long oracle_lzf_compress(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap)
{
    enum { HLOG = 16, HSIZE = 1 << HLOG, MAX_OFF = 1 << 13, MAX_REF = (1 << 8) + (1 << 3), MAX_LIT = 1 << 5 };
    static long htab_store[HSIZE];
    long *htab = htab_store;
    for (int i = 0; i < HSIZE; ++i) htab[i] = -1;
    size_t ip = 0, op = 0, lit_start = 0;
#define FLUSH_LITS(upto)                                                            \
    do {                                                                            \
        size_t s_ = lit_start, e_ = (upto);                                         \
        while (s_ < e_) {                                                           \
            size_t n_ = e_ - s_ < MAX_LIT ? e_ - s_ : MAX_LIT;                      \
            if (op + 1 + n_ > out_cap) return -1;                                   \
            out[op++] = (uint8_t)(n_ - 1);                                          \
            for (size_t k_ = 0; k_ < n_; ++k_) out[op++] = in[s_ + k_];             \
            s_ += n_;                                                               \
        }                                                                           \
    } while (0)
    while (ip + 2 < in_len) {
        unsigned h = ((unsigned)in[ip] << 16) | ((unsigned)in[ip + 1] << 8) | in[ip + 2];
        unsigned slot = ((h * 2654435761u) >> (32 - HLOG)) & (HSIZE - 1);
        long ref = htab[slot];
        htab[slot] = (long)ip;
        size_t len = 0;
        if (ref >= 0 && ip - (size_t)ref <= MAX_OFF && in[ref] == in[ip] && in[ref + 1] == in[ip + 1] && in[ref + 2] == in[ip + 2]) {
            len = 3;
            while (ip + len < in_len && len < MAX_REF && in[ref + len] == in[ip + len]) ++len;
        }
        if (len >= 3) {
            FLUSH_LITS(ip);
            size_t off = ip - (size_t)ref - 1, l = len - 2;
            if (op + 3 > out_cap) return -1;
            if (l < 7) out[op++] = (uint8_t)((l << 5) | (off >> 8));
            else { out[op++] = (uint8_t)((7 << 5) | (off >> 8)); out[op++] = (uint8_t)(l - 7); }
            out[op++] = (uint8_t)(off & 0xff);
            ip += len;
            lit_start = ip;
        } else
            ++ip;
    }
    FLUSH_LITS(in_len);
#undef FLUSH_LITS
    return (long)op;
}
