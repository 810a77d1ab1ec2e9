// resolver.cpp -- phase 2 of the e2SEC path: the strictly sequential part of _process, on the host.
//
// Consumes the per-frame records of phase 1 (include/e2sec.h) in frame order and reproduces
//   early-outs                      manipulation_analyser.py:1429, :1475, :1551
//   o1 / o2 / o3 selection          :1562-1658 (sticky, first come, ascending label index) and :1669-1675
//   _fillTN_absent                  :363-487    rows 0-9
//   _fillSSR_8 + _R.._Bw            :491-1061   rows 10-19
//   _fillDSR_new_2 + _distance      :1064-1353, :200-209  rows 20-29
//   column emit / dedup             :1795-1829, :2000-2007
//   esec_to_e2sec                   :62-111
// State mirrors the module globals :32-59. Scalar work on five entities per frame; no device code.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/e2sec.h"
#include "relations.h"

struct e2sec_resolver {
    e2sec_params p;
    int K = 0, hand_idx = -1, ground_idx = -1, support_idx = -1;
    int relations = 3;
    size_t rec = 0;
    RelEnt ground;
    // globals :32-59
    RelEnt o[3];
    int oindex[3] = {-1, -1, -1};   // entity index bound to o1..o3 (o1_label etc. index total_unique_labels)
    int count[3] = {0, 0, 0};
    int absent[3] = {0, 0, 0};
    int internal_count = 0;
    bool have_prev = false;
    RelEnt prev[5];              // previous_array = [hand, ground, o1, o2, o3] (:1799, :1822)
    int prev_frame_id = -1;      // frame id of the previous kept frame
    std::vector<uint8_t> table;  // rows codes per column, column-major (RelCode)
    std::vector<int> final_columns;   // frame id of every column (the reference only fills it for relations == 3)
    std::string err;
};

static int rows_of(int relations) { return 10 * relations; }

static RelEnt undefined_ent() { return rel_ent_undefined(); }
static RelEnt ent_from_record(const double *rec, int e, bool is_hand) { return rel_ent_from_record(rec, e, is_hand); }
static_assert(REL_REC_HEADER == E2SEC_REC_HEADER && REL_REC_ENTITY == E2SEC_REC_ENTITY, "record layout");

extern "C" int e2sec_resolver_create(const e2sec_params *p, e2sec_resolver **out)
{
    if (!p || !out) return E2SEC_ERR_ARG;
    *out = nullptr;
    if (p->n_entities < 2 || p->n_entities > E2SEC_MAX_ENTITIES) return E2SEC_ERR_ARG;
    if (p->relations != 0 && (p->relations < 1 || p->relations > 3)) return E2SEC_ERR_ARG;
    e2sec_resolver *r = new e2sec_resolver();
    r->p = *p;
    r->p.ground_xyz = nullptr;
    r->K = p->n_entities;
    r->relations = p->relations ? p->relations : 3;
    r->rec = e2sec_record_doubles(r->K);
    for (int e = 0; e < r->K; ++e) {
        if (p->entity_label[e] == p->hand_label) r->hand_idx = e;
        if (p->entity_label[e] == p->ground_label) r->ground_idx = e;
        if (p->support_hand >= 0 && p->entity_label[e] == p->support_hand) r->support_idx = e;
    }
    if (r->hand_idx < 0 || r->ground_idx < 0 || (p->support_hand >= 0 && r->support_idx < 0)) {
        delete r;
        return E2SEC_ERR_ARG;
    }
    for (auto &x : r->o) x = undefined_ent();
    for (auto &x : r->prev) x = undefined_ent();
    RelEnt &g = r->ground;
    g = undefined_ent();
    g.defined = 1;
    g.n = p->n_ground;
    for (int i = 0; i < p->n_ground; ++i)
        for (int a = 0; a < 3; ++a) {
            const double v = p->ground_xyz[3 * (size_t)i + a];
            if (i == 0 || v < g.lo[a]) g.lo[a] = v;
            if (i == 0 || v > g.hi[a]) g.hi[a] = v;
        }
    for (int a = 0; a < 3; ++a) g.center[a] = (g.lo[a] + g.hi[a]) * 0.5;
    if (r->relations < 3) {                                  // the '-' column the table starts with (:2252-2260) stays column 0
        r->table.assign(rows_of(r->relations), (uint8_t)REL_DASH);
        r->final_columns.push_back(-1);
    }
    *out = r;
    return E2SEC_OK;
}

extern "C" void e2sec_resolver_destroy(e2sec_resolver *r) { delete r; }

namespace {

struct FrameView {
    const e2sec_resolver *R;
    const double *rec;        // this frame
    double pair(int a, int b) const   // in-frame distance, b may be K (ground)
    {
        const int W = R->K + 1;
        const double *pt = rec + E2SEC_REC_HEADER + E2SEC_REC_ENTITY * R->K;
        return a < b ? pt[a * W + b] : pt[b * W + a];
    }
    double cross(int a_prev, int b) const   // entity a of the previous kept frame vs entity b of this frame
    {
        const int W = R->K + 1;
        const double *ct = rec + E2SEC_REC_HEADER + E2SEC_REC_ENTITY * R->K + W * W;
        return ct[a_prev * R->K + b];
    }
};

}   // namespace

// one call of _process (:1355) on one record
static int process_record(e2sec_resolver *R, const double *rec)
{
    const int K = R->K;
    const int flags = (int)rec[1];
    const int frame = (int)rec[0];
    if ((int)rec[2] != K) { R->err = "record has a different entity count"; return E2SEC_ERR_ARG; }
    if (!(flags & E2SEC_FLAG_HAND_LABEL)) return E2SEC_OK;      // :1429
    if (!(flags & E2SEC_FLAG_HAND_POINTS)) return E2SEC_OK;     // :1475, :1551
    // the cross table must have been measured against the frame this resolver saw last
    const int rec_prev = (int)rec[3];
    if (rec_prev != R->prev_frame_id) {
        R->err = "record " + std::to_string(frame) + " was computed against previous kept frame " + std::to_string(rec_prev) +
                 " but the resolver's previous kept frame is " + std::to_string(R->prev_frame_id) +
                 " (feed frames in order; a shard must start with its halo frame)";
        return E2SEC_ERR_STATE;
    }
    FrameView v{R, rec};
    const double th = R->p.thresh;
    const RelEnt hand = ent_from_record(rec, R->hand_idx, true);
    // ---- selection :1562-1658 (relations.h: rel_select, shared with the device classifier) --------
    {
        RelSel S;
        int stale_n[3];
        for (int k = 0; k < 3; ++k) { S.oindex[k] = R->oindex[k]; S.count[k] = R->count[k]; stale_n[k] = R->o[k].n; }
        rel_select(rec, K, R->hand_idx, R->ground_idx, R->support_idx, th, S, stale_n);
        for (int k = 0; k < 3; ++k) { R->oindex[k] = S.oindex[k]; R->count[k] = S.count[k]; }
    }
    for (int k = 0; k < 3; ++k)                              // icp == False :1669-1675
        if (R->count[k] == 1) R->o[k] = ent_from_record(rec, R->oindex[k], false);
    const RelEnt cur[5] = {hand, R->ground, R->o[0], R->o[1], R->o[2]};
    auto remember = [&]() {
        for (int s = 0; s < 5; ++s) R->prev[s] = cur[s];
        R->have_prev = true;
        R->prev_frame_id = frame;
    };
    const int rows = rows_of(R->relations);
    if (R->relations == 3 && R->internal_count == 0) {       // :1798-1801 the first kept frame only primes
        remember();
        R->internal_count = 1;
        return E2SEC_OK;
    }
    // np.min(a.compute_point_cloud_distance(b)) of the ten row pairs (entity index K = ground)
    double d[10];
    {
        const int slot_idx[5] = {R->hand_idx, K, R->oindex[0], R->oindex[1], R->oindex[2]};
        const int pairs[10][2] = REL_ROW_PAIRS;
        for (int r = 0; r < 10; ++r) {
            const int a = slot_idx[pairs[r][0]], b = slot_idx[pairs[r][1]];
            const bool ok = cur[pairs[r][0]].defined && cur[pairs[r][1]].defined && a >= 0 && b >= 0;
            d[r] = ok ? v.pair(a, b) : INFINITY;
        }
    }
    uint8_t add[30];
    rel_fill_column(cur, R->prev, d, th, R->relations, R->absent, add);
    remember();                                              // :1822 (relations 1 / 2 keep no previous_array; harmless)
    bool emit = true, replace_first = false;
    if (R->relations == 3) {
        if (R->internal_count == 1) R->internal_count = 2;   // :1826-1829
        else emit = memcmp(R->table.data() + R->table.size() - rows, add, rows) != 0;   // :2001-2002
    } else {
        // :1744-1794: `if frame == 0` the '-' table is replaced by the column; otherwise compare with the last column
        if (frame == 0) replace_first = true;
        else emit = memcmp(R->table.data() + R->table.size() - rows, add, rows) != 0;
    }
    if (replace_first) {
        R->table.assign(add, add + rows);
        R->final_columns.assign(1, frame);
    } else if (emit) {
        R->table.insert(R->table.end(), add, add + rows);
        R->final_columns.push_back(frame);
    }
    return E2SEC_OK;
}

extern "C" int e2sec_resolver_feed(e2sec_resolver *r, const double *records, int32_t n_frames)
{
    if (!r || (!records && n_frames > 0)) return E2SEC_ERR_ARG;
    for (int f = 0; f < n_frames; ++f) {
        int rc = process_record(r, records + (size_t)f * r->rec);
        if (rc) return rc;
    }
    return E2SEC_OK;
}

extern "C" int32_t e2sec_resolver_columns(const e2sec_resolver *r) { return r ? (int32_t)r->final_columns.size() : 0; }

extern "C" const char *e2sec_resolver_error(const e2sec_resolver *r) { return r ? r->err.c_str() : ""; }

extern "C" int e2sec_resolver_table(const e2sec_resolver *r, char *codes, int32_t *final_columns)
{
    if (!r) return E2SEC_ERR_ARG;
    const int n = (int)r->final_columns.size(), rows = rows_of(r->relations);
    if (codes)
        for (int c = 0; c < n; ++c)
            for (int row = 0; row < rows; ++row) {
                char *dst = codes + ((size_t)row * n + c) * 6;
                memset(dst, 0, 6);
                strncpy(dst, rel_code_name(r->table[(size_t)c * rows + row]), 5);
            }
    if (final_columns)
        for (int c = 0; c < n; ++c) final_columns[c] = r->final_columns[c];
    return E2SEC_OK;
}

extern "C" int e2sec_resolver_state(const e2sec_resolver *r, int32_t labels[3], int32_t absent[3])
{
    if (!r) return E2SEC_ERR_ARG;
    for (int k = 0; k < 3; ++k) {
        if (labels) labels[k] = r->count[k] ? r->oindex[k] : -1;
        if (absent) absent[k] = r->absent[k] ? 1 : 0;
    }
    return E2SEC_OK;
}

// _fillTN_absent / _fillSSR_8 / _fillDSR_new_2 on explicit inputs: the reference's own seam (include/e2sec.h)
extern "C" int e2sec_relation_rows(const e2sec_relation_input *in, int32_t relations, char *codes, int32_t absent_out[3])
{
    if (!in || !codes || relations < 1 || relations > 3) return E2SEC_ERR_ARG;
    RelEnt cur[5], prev[5];
    for (int s = 0; s < 5; ++s) {
        const e2sec_relation_entity *src[2] = {&in->cur[s], &in->prev[s]};
        RelEnt *dst[2] = {&cur[s], &prev[s]};
        for (int w = 0; w < 2; ++w) {
            dst[w]->defined = src[w]->defined;
            dst[w]->n = src[w]->n_points;
            for (int k = 0; k < 3; ++k) { dst[w]->lo[k] = src[w]->aabb_min[k]; dst[w]->hi[k] = src[w]->aabb_max[k]; dst[w]->center[k] = src[w]->center[k]; }
        }
    }
    int absent[3] = {in->absent[0], in->absent[1], in->absent[2]};
    uint8_t out[30];
    rel_fill_column(cur, prev, in->min_distance, in->thresh, relations, absent, out);
    for (int r = 0; r < 10 * relations; ++r) {
        memset(codes + 6 * r, 0, 6);
        strncpy(codes + 6 * r, rel_code_name(out[r]), 5);
    }
    if (absent_out) for (int k = 0; k < 3; ++k) absent_out[k] = absent[k];
    return E2SEC_OK;
}

// esec_to_e2sec :62-111
extern "C" int32_t e2sec_esec_to_e2sec(const char *codes, int32_t rows, int32_t n_cols, int32_t relations, char *out_codes,
                                       int32_t *final_columns, int32_t *n_final)
{
    if (!codes || !out_codes || rows < 1 || n_cols < 0) return E2SEC_ERR_ARG;
    std::vector<int> drop;
    if (relations == 3) drop = {3, 9, 13, 19, 23, 29};
    else if (relations == 2) drop = {3, 9, 13, 19};
    else if (relations == 1) drop = {3, 9};
    else return E2SEC_ERR_ARG;
    std::vector<int> keep_rows;
    for (int r = 0; r < rows; ++r) {
        bool d = false;
        for (int x : drop) d |= (x == r);
        if (!d) keep_rows.push_back(r);
    }
    const int R2 = (int)keep_rows.size();
    const int C1 = n_cols > 0 ? n_cols - 1 : 0;               // column 0 is deleted unconditionally (:83-84)
    auto at = [&](int r, int c) { return codes + ((size_t)keep_rows[r] * n_cols + (c + 1)) * 6; };
    auto eq_upper = [&](const char *a, const char *b) {
        for (int i = 0; i < 6; ++i) {
            char x = a[i], y = b[i];
            if (x >= 'a' && x <= 'z') x -= 32;
            if (y >= 'a' && y <= 'z') y -= 32;
            if (x != y) return false;
            if (!x) break;
        }
        return true;
    };
    std::vector<int> keep_cols;
    std::vector<int> deleted;                                 // indices k = j + 1 removed (:96-103)
    for (int c = 0; c < C1; ++c) {
        bool dup = false;
        if (c > 0) {
            dup = true;
            for (int r = 0; r < R2; ++r)
                if (!eq_upper(at(r, c - 1), at(r, c))) { dup = false; break; }
        }
        if (dup) deleted.push_back(c); else keep_cols.push_back(c);
    }
    const int C2 = (int)keep_cols.size();
    for (int r = 0; r < R2; ++r)
        for (int c = 0; c < C2; ++c) {
            const char *src = at(r, keep_cols[c]);
            char *dst = out_codes + ((size_t)r * C2 + c) * 6;
            memset(dst, 0, 6);
            for (int i = 0; i < 5 && src[i]; ++i) dst[i] = (src[i] >= 'a' && src[i] <= 'z') ? src[i] - 32 : src[i];
        }
    // final_columns is thinned by the same indices, applied to the un-shifted list (the reference quirk: it
    // stays one longer than the matrix, :104-105)
    if (final_columns && n_final) {
        std::vector<int> fc(final_columns, final_columns + *n_final), outv;
        for (int i = 0; i < (int)fc.size(); ++i) {
            bool d = false;
            for (int x : deleted) d |= (x == i);
            if (!d) outv.push_back(fc[i]);
        }
        for (size_t i = 0; i < outv.size(); ++i) final_columns[i] = outv[i];
        *n_final = (int)outv.size();
    }
    return C2;
}

// packed phase-2 columns (device classifier, possibly gathered from several shards) -> the table layout of
// e2sec_resolver_table.  With relations 1 / 2 the table starts with the '-' column (:2252-2260) unless frame 0 itself
// produced the first column (`if frame == 0: ESEC_table = add`, :1753 / :1778).
extern "C" int32_t e2sec_columns_table(const uint8_t *columns, const int32_t *frames, int32_t n, int32_t relations, char *codes,
                                       int32_t *final_columns)
{
    if (n < 0 || relations < 1 || relations > 3 || (n > 0 && (!columns || !frames))) return E2SEC_ERR_ARG;
    const int rows = 10 * relations;
    const int dash = (relations < 3 && !(n > 0 && frames[0] == 0)) ? 1 : 0;
    const int total = n + dash;
    if (!codes && !final_columns) return total;
    for (int c = 0; c < total; ++c) {
        for (int row = 0; codes && row < rows; ++row) {
            char *dst = codes + ((size_t)row * total + c) * 6;
            memset(dst, 0, 6);
            const uint8_t code = (c < dash) ? (uint8_t)REL_DASH : columns[(size_t)(c - dash) * E2SEC_PHASE2_COLUMN_BYTES + row];
            strncpy(dst, rel_code_name(code), 5);
        }
        if (final_columns) final_columns[c] = (c < dash) ? -1 : frames[c - dash];
    }
    return total;
}
