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
