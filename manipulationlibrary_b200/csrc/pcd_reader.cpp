// pcd_reader.cpp -- host-side input decode for config 1 (the bundled MANIAC files).
//   e2sec_read_pcd_xyz   o3d.io.read_point_cloud(file, remove_nan_points=False)  manipulation_analyser.py:1441
//                        PCD v0.7, float32 x/y/z fields, DATA ascii | binary | binary_compressed (LZF, fields
//                        stored one after another: all x, all y, all z, ...)
//   e2sec_read_labels    pd.read_csv(label_file, delim_whitespace=True, header=None)                  :1391
// Not part of the timed path (SURVEY.md 8d excludes file decode on both sides).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/e2sec.h"

namespace {

bool slurp(const char *path, std::vector<unsigned char> &buf)
{
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (n < 0) { fclose(f); return false; }
    buf.resize((size_t)n);
    size_t got = n ? fread(buf.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == (size_t)n;
}

// LZF: a stream of chunks. Control byte c: c < 32 -> c + 1 literal bytes follow; otherwise a back reference of
// length (c >> 5) + 2 (a length field of 7 is extended by the next byte) to offset ((c & 31) << 8 | next) + 1.
bool lzf_expand(const unsigned char *src, size_t n_src, std::vector<unsigned char> &dst, size_t n_dst)
{
    dst.resize(n_dst);
    size_t i = 0, o = 0;
    while (i < n_src) {
        const unsigned c = src[i++];
        if (c < 32u) {
            const size_t run = c + 1u;
            if (i + run > n_src || o + run > n_dst) return false;
            memcpy(&dst[o], &src[i], run);
            i += run;
            o += run;
            continue;
        }
        size_t len = c >> 5;
        if (len == 7) {
            if (i >= n_src) return false;
            len += src[i++];
        }
        if (i >= n_src) return false;
        const size_t back = (((size_t)(c & 31u)) << 8 | src[i++]) + 1u;
        len += 2;
        if (back > o || o + len > n_dst) return false;
        for (size_t k = 0; k < len; ++k, ++o) dst[o] = dst[o - back];   // may overlap: byte by byte
    }
    return o == n_dst;
}

struct Header {
    std::vector<std::string> fields, types;
    std::vector<int> sizes, counts;
    long width = 0, height = 0, points = 0;
    std::string data;
    size_t body = 0;
};

std::vector<std::string> split(const std::string &s)
{
    std::vector<std::string> out;
    size_t i = 0;
    while (i < s.size()) {
        while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\r')) ++i;
        size_t j = i;
        while (j < s.size() && s[j] != ' ' && s[j] != '\t' && s[j] != '\r') ++j;
        if (j > i) out.push_back(s.substr(i, j - i));
        i = j;
    }
    return out;
}

bool parse_header(const std::vector<unsigned char> &buf, Header &h)
{
    size_t pos = 0;
    while (pos < buf.size()) {
        size_t end = pos;
        while (end < buf.size() && buf[end] != '\n') ++end;
        std::string line((const char *)&buf[pos], end - pos);
        pos = end + 1;
        auto tok = split(line);
        if (tok.empty() || tok[0][0] == '#') continue;
        const std::string &key = tok[0];
        if (key == "FIELDS") h.fields.assign(tok.begin() + 1, tok.end());
        else if (key == "SIZE") for (size_t i = 1; i < tok.size(); ++i) h.sizes.push_back(atoi(tok[i].c_str()));
        else if (key == "TYPE") h.types.assign(tok.begin() + 1, tok.end());
        else if (key == "COUNT") for (size_t i = 1; i < tok.size(); ++i) h.counts.push_back(atoi(tok[i].c_str()));
        else if (key == "WIDTH" && tok.size() > 1) h.width = atol(tok[1].c_str());
        else if (key == "HEIGHT" && tok.size() > 1) h.height = atol(tok[1].c_str());
        else if (key == "POINTS" && tok.size() > 1) h.points = atol(tok[1].c_str());
        else if (key == "DATA" && tok.size() > 1) { h.data = tok[1]; h.body = pos; return true; }
    }
    return false;
}

}   // namespace

// Where the LZF stream of a binary_compressed cloud lies inside its file (for the device decoder, lzf.cuh): the header is
// read here, nothing else.  Only the layout the device decoder handles is accepted: x, y, z as the first three fields,
// float32 each, so that the first 12 * POINTS decoded bytes are the x, y and z planes.
int e2sec_pcd_locate(const unsigned char *file, size_t len, long long *payload, unsigned *comp, long long *points, int *width, int *height)
{
    std::vector<unsigned char> head(file, file + (len < 4096 ? len : 4096));       // a header is a few hundred bytes
    Header h;
    if (!parse_header(head, h)) return E2SEC_ERR_IO;
    const size_t nf = h.fields.size();
    if (nf < 3 || h.sizes.size() != nf || h.types.size() != nf) return E2SEC_ERR_IO;
    if (h.counts.empty()) h.counts.assign(nf, 1);
    if (h.counts.size() != nf) return E2SEC_ERR_IO;
    if (h.data != "binary_compressed") return E2SEC_ERR_ARG;
    for (int a = 0; a < 3; ++a)
        if (h.fields[a] != std::string(1, "xyz"[a]) || h.sizes[a] != 4 || h.types[a] != "F" || h.counts[a] != 1) return E2SEC_ERR_ARG;
    if (h.points <= 0) h.points = h.width * h.height;
    if (h.points <= 0 || h.points > (1ll << 28)) return E2SEC_ERR_IO;
    if (h.body + 8 > len) return E2SEC_ERR_IO;
    uint32_t c, raw;
    memcpy(&c, file + h.body, 4);
    memcpy(&raw, file + h.body + 4, 4);
    if ((uint64_t)c > len - h.body - 8 || (uint64_t)raw < (uint64_t)12 * (uint64_t)h.points) return E2SEC_ERR_IO;
    *payload = (long long)h.body + 8;
    *comp = c;
    *points = h.points;
    if (width) *width = (int)h.width;
    if (height) *height = (int)h.height;
    return E2SEC_OK;
}

extern "C" int e2sec_read_pcd_xyz(const char *path, float *xyz, int64_t capacity_points, int32_t *width, int32_t *height)
{
    if (!path || !xyz) return E2SEC_ERR_ARG;
    std::vector<unsigned char> buf;
    if (!slurp(path, buf)) return E2SEC_ERR_IO;
    Header h;
    if (!parse_header(buf, h)) return E2SEC_ERR_IO;
    const size_t nf = h.fields.size();
    if (h.sizes.size() != nf || h.types.size() != nf) return E2SEC_ERR_IO;
    if (h.counts.empty()) h.counts.assign(nf, 1);
    // the header is untrusted: one COUNT per FIELD, positive sizes that cannot make stride * POINTS wrap
    if (h.counts.size() != nf || nf == 0 || nf > 64) return E2SEC_ERR_IO;
    for (size_t k = 0; k < nf; ++k)
        if (h.sizes[k] <= 0 || h.sizes[k] > 8 || h.counts[k] <= 0 || h.counts[k] > 4096) return E2SEC_ERR_IO;
    if (h.width < 0 || h.height < 0 || h.width > (1 << 24) || h.height > (1 << 24)) return E2SEC_ERR_IO;
    if (h.points <= 0) h.points = h.width * h.height;
    if (h.points <= 0 || h.points > (1ll << 31)) return E2SEC_ERR_IO;
    if (h.points > capacity_points) return E2SEC_ERR_CAPACITY;
    int fi[3] = {-1, -1, -1};
    std::vector<size_t> foff(nf);
    size_t stride = 0;
    for (size_t k = 0; k < nf; ++k) {
        foff[k] = stride;
        stride += (size_t)h.sizes[k] * h.counts[k];
        for (int a = 0; a < 3; ++a)
            if (h.fields[k] == std::string(1, "xyz"[a])) fi[a] = (int)k;
    }
    for (int a = 0; a < 3; ++a)
        if (fi[a] < 0 || h.sizes[fi[a]] != 4 || h.types[fi[a]] != "F" || h.counts[fi[a]] != 1) return E2SEC_ERR_IO;
    const size_t P = (size_t)h.points;
    if (h.data == "binary_compressed") {
        if (h.body + 8 > buf.size()) return E2SEC_ERR_IO;
        uint32_t comp, raw;
        memcpy(&comp, &buf[h.body], 4);
        memcpy(&raw, &buf[h.body + 4], 4);
        if ((uint64_t)comp > buf.size() - h.body - 8 || (uint64_t)raw < (uint64_t)stride * P) return E2SEC_ERR_IO;
        std::vector<unsigned char> plain;
        if (!lzf_expand(&buf[h.body + 8], comp, plain, raw)) return E2SEC_ERR_IO;
        for (int a = 0; a < 3; ++a) {
            const unsigned char *col = plain.data() + foff[fi[a]] * P;
            for (size_t i = 0; i < P; ++i) memcpy(&xyz[3 * i + a], col + 4 * i, 4);
        }
    } else if (h.data == "binary") {
        if (h.body > buf.size() || (uint64_t)stride * P > buf.size() - h.body) return E2SEC_ERR_IO;
        for (size_t i = 0; i < P; ++i)
            for (int a = 0; a < 3; ++a) memcpy(&xyz[3 * i + a], &buf[h.body + i * stride + foff[fi[a]]], 4);
    } else if (h.data == "ascii") {
        std::vector<int> col(nf);
        int c = 0;
        for (size_t k = 0; k < nf; ++k) { col[k] = c; c += h.counts[k]; }
        const char *p = (const char *)buf.data() + h.body, *e = (const char *)buf.data() + buf.size();
        std::string text(p, e - p);
        char *cur = &text[0];
        for (size_t i = 0; i < P; ++i)
            for (int j = 0; j < c; ++j) {
                char *next = nullptr;
                double v = strtod(cur, &next);
                if (next == cur) return E2SEC_ERR_IO;
                cur = next;
                for (int a = 0; a < 3; ++a)
                    if (col[fi[a]] == j) xyz[3 * i + a] = (float)v;
            }
    } else
        return E2SEC_ERR_IO;
    if (width) *width = (int32_t)h.width;
    if (height) *height = (int32_t)h.height;
    return E2SEC_OK;
}

extern "C" int e2sec_read_labels(const char *path, int32_t *labels, int64_t capacity, int32_t *rows, int32_t *cols)
{
    if (!path || !labels) return E2SEC_ERR_ARG;
    std::vector<unsigned char> buf;
    if (!slurp(path, buf)) return E2SEC_ERR_IO;
    buf.push_back('\n');
    buf.push_back(0);
    int64_t n = 0;
    int r = 0, c_first = -1, c = 0;
    char *cur = (char *)buf.data();
    while (*cur) {
        while (*cur == ' ' || *cur == '\t' || *cur == '\r') ++cur;
        if (*cur == '\n') {
            if (c > 0) {
                if (c_first < 0) c_first = c;
                else if (c != c_first) return E2SEC_ERR_IO;     // ragged rows: pandas would pad with NaN; not a label image
                ++r;
            }
            c = 0;
            ++cur;
            continue;
        }
        if (!*cur) break;
        char *next = nullptr;
        double v = strtod(cur, &next);
        if (next == cur) return E2SEC_ERR_IO;
        if (n >= capacity) return E2SEC_ERR_CAPACITY;
        labels[n++] = (int32_t)v;
        ++c;
        cur = next;
    }
    if (rows) *rows = r;
    if (cols) *cols = c_first < 0 ? 0 : c_first;
    return E2SEC_OK;
}
