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

namespace {

const double kCenterDistance = 0.005107632747118286;   // :1121
const double kStableDist = 0.0045302034127268085;      // :1122

struct Ent {              // what the relation rows read of one cloud
    bool defined = false; // false = the Python variable is None
    int n = 0;            // len(points)
    double lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    double center[3] = {0, 0, 0};   // hand: cloud mean; objects and ground: AABB centre (:1173-1353)
    int index = -1;       // entity index in total_unique_labels (K = ground)
};

struct Code {
    char c[6];
};

Code mk(const char *s)
{
    Code c;
    memset(c.c, 0, 6);
    strncpy(c.c, s, 5);
    return c;
}
bool same(const Code &a, const char *s) { return strcmp(a.c, s) == 0; }

struct Box {   // SSR axis naming (:660-675): x = data[:,0], z = data[:,1], y = data[:,2]
    double min_x, max_x, min_z, max_z, min_y, max_y;
};
Box box_of(const Ent &e) { return Box{e.lo[0], e.hi[0], e.lo[1], e.hi[1], e.lo[2], e.hi[2]}; }

bool ov_x(const Box &a, const Box &b) { return a.min_x < b.max_x && a.max_x > b.min_x; }
bool ov_y(const Box &a, const Box &b) { return a.min_y < b.max_y && a.max_y > b.min_y; }
bool ov_z(const Box &a, const Box &b) { return a.min_z < b.max_z && a.max_z > b.min_z; }

bool around(const Box &a, const Box &b)   // _R or _L or _F or _Ba (:491-550)
{
    const bool R = a.max_x > b.max_x && ov_y(a, b) && ov_z(a, b);
    const bool L = a.min_x < b.min_x && ov_y(a, b) && ov_z(a, b);
    const bool F = a.max_z > b.max_z && (a.min_x < b.max_x && b.min_x < a.max_x) && ov_y(a, b);
    const bool Ba = a.min_z < b.min_z && (a.min_x < b.max_x && b.min_x < a.max_x) && ov_y(a, b);
    return R || L || F || Ba;
}

Code tail(const Box &a, const Box &b, bool touching)   // the elif chain every SSR row ends with (:729-745)
{
    if (around(a, b)) return mk(touching ? "HArT" : "HAr");
    if (a.max_y > b.max_y || a.min_y < b.min_y) return mk(touching ? "VArT" : "VAr");   // _Ab or _Be (:570-599)
    return mk("O");
}

bool between(const Box &m, const Box &a, const Box &b)   // one disjunct of _Bw (:615-633)
{
    return m.min_x >= std::fmin(a.max_x, b.max_x) && m.max_x <= std::fmax(a.min_x, b.min_x) &&
           m.min_y >= std::fmax(a.min_y, b.min_y) && m.max_y <= std::fmin(a.max_y, b.max_y) &&
           m.min_z >= std::fmax(a.min_z, b.min_z) && m.max_z <= std::fmin(a.max_z, b.max_z);
}

double dist3(const double *p, const double *q)   // _distance :200-209
{
    return std::sqrt((p[0] - q[0]) * (p[0] - q[0]) + (p[1] - q[1]) * (p[1] - q[1]) + (p[2] - q[2]) * (p[2] - q[2]));
}

}   // namespace

struct e2sec_resolver {
    e2sec_params p;
    int K = 0, hand_idx = -1, ground_idx = -1, support_idx = -1;
    size_t rec = 0;
    Ent ground;
    // globals :32-59
    Ent o[3];
    int count[3] = {0, 0, 0};
    int olabel[3] = {0, 0, 0};
    bool absent[3] = {false, false, false};
    int internal_count = 0;
    bool have_prev = false;
    Ent prev[5];                 // previous_array = [hand, ground, o1, o2, o3] (:1799, :1822)
    int prev_frame_id = -1;      // frame id of the previous kept frame
    std::vector<double> prev_rec;   // its record (for the stale o1/o2/o3 clouds of the selection loop)
    std::vector<Code> table;     // 30 codes per column, column-major
    std::vector<int> final_columns;
    std::string err;
};

static Ent ent_from_record(const double *rec, int e, bool is_hand)
{
    const double *r = rec + E2SEC_REC_HEADER + E2SEC_REC_ENTITY * e;
    Ent x;
    x.defined = true;
    x.index = e;
    x.n = (int)r[1];
    for (int k = 0; k < 3; ++k) {
        x.lo[k] = r[7 + k];
        x.hi[k] = r[10 + k];
        x.center[k] = is_hand ? r[13 + k] : (x.lo[k] + x.hi[k]) * 0.5;
    }
    return x;
}

extern "C" int e2sec_resolver_create(const e2sec_params *p, e2sec_resolver **out)
{
    if (!p || !out) return E2SEC_ERR_ARG;
    *out = nullptr;
    if (p->n_entities < 2 || p->n_entities > E2SEC_MAX_ENTITIES) return E2SEC_ERR_ARG;
    e2sec_resolver *r = new e2sec_resolver();
    r->p = *p;
    r->p.ground_xyz = nullptr;
    r->K = p->n_entities;
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
    Ent &g = r->ground;
    g.defined = true;
    g.index = r->K;
    g.n = p->n_ground;
    for (int i = 0; i < p->n_ground; ++i)
        for (int a = 0; a < 3; ++a) {
            const double v = p->ground_xyz[3 * (size_t)i + a];
            if (i == 0 || v < g.lo[a]) g.lo[a] = v;
            if (i == 0 || v > g.hi[a]) g.hi[a] = v;
        }
    for (int a = 0; a < 3; ++a) g.center[a] = (g.lo[a] + g.hi[a]) * 0.5;
    *out = r;
    return E2SEC_OK;
}

extern "C" void e2sec_resolver_destroy(e2sec_resolver *r) { delete r; }

namespace {

struct FrameView {
    const e2sec_resolver *R;
    const double *rec;        // this frame
    const double *prev_rec;   // previous kept frame (may be null)
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

// distance between the cloud currently bound to an o-variable and the frame's entity i.
// `fresh` = the variable was (re)bound during this frame's loop, so it is this frame's cloud; otherwise it is
// still the cloud of the previous kept frame (:1669-1675 rebinds only after the loop).
double dist_to(const FrameView &v, const Ent &o, bool fresh, int i)
{
    if (o.n == 0) return INFINITY;   // compute_point_cloud_distance on an empty cloud: np.min raises in the reference
    return fresh ? v.pair(o.index, i) : v.cross(o.index, i);
}

}   // namespace

static void fill_tn(e2sec_resolver *R, const FrameView &v, const Ent &H, const Ent *o, Code *out)
{
    const double th = R->p.thresh;
    for (int k = 0; k < 3; ++k) {                         // rows 0-2 (:389-427)
        if (!o[k].defined) out[k] = mk("U");
        else if (o[k].n == 0 || R->absent[k]) { out[k] = mk("A"); R->absent[k] = true; }
        else out[k] = mk(v.pair(H.index, o[k].index) < th ? "T" : "N");
    }
    out[3] = mk(v.pair(H.index, R->K) < th ? "T" : "N");   // row 3 (:429-433)
    const int pr[3][3] = {{4, 0, 1}, {5, 0, 2}, {7, 1, 2}};
    for (auto &q : pr) {                                   // rows 4,5,7 (:435-451, :462-469)
        const Ent &a = o[q[1]], &b = o[q[2]];
        if (!a.defined || !b.defined) out[q[0]] = mk("U");
        else if (a.n == 0 || b.n == 0 || R->absent[q[1]] || R->absent[q[2]]) out[q[0]] = mk("A");
        else out[q[0]] = mk(v.pair(a.index, b.index) < th ? "T" : "N");
    }
    const int gr[3][2] = {{6, 0}, {8, 1}, {9, 2}};
    for (auto &q : gr) {                                   // rows 6,8,9 (:453-460, :471-487)
        const Ent &a = o[q[1]];
        if (!a.defined) out[q[0]] = mk("U");
        else if (a.n == 0 || R->absent[q[1]]) out[q[0]] = mk("A");
        else out[q[0]] = mk(v.pair(a.index, R->K) < th ? "T" : "N");
    }
}

static void fill_ssr(const e2sec_resolver *R, const Ent &Hh, const Ent *o, const Code *tn, Code *out)
{
    const Box H = box_of(Hh), G = box_of(R->ground);
    Box B[3];
    bool have[3];
    for (int k = 0; k < 3; ++k) {
        have[k] = o[k].defined && !same(tn[k], "A");
        if (o[k].defined) B[k] = box_of(o[k]);
    }
    const bool all_there = o[0].defined && o[1].defined && o[2].defined;
    for (int k = 0; k < 3; ++k) {                          // rows 10-12 (:706-812)
        if (!o[k].defined) out[k] = mk("U");
        else if (same(tn[k], "A")) out[k] = mk("A");
        else {
            const Box &b = B[k];
            const bool sa = (b.min_z < H.min_z && b.max_z > H.max_z) &&
                            ((b.max_x < H.min_x && b.max_x > H.max_x) || (b.min_x < H.min_x && b.min_x > H.max_x) ||
                             (b.min_x > H.min_x && b.max_x < H.max_x)) &&
                            (H.min_y < b.max_y && b.max_y <= H.max_y);
            out[k] = sa ? mk("Sa") : tail(H, b, same(tn[k], "T"));
        }
    }
    out[3] = tail(H, G, same(tn[3], "T"));                 // row 13 (:814-842)
    const bool bw_gate = all_there && !same(tn[0], "A") && !same(tn[1], "A") && !same(tn[2], "A");
    const int pr[3][4] = {{4, 0, 1, 12}, {5, 0, 2, 13}, {7, 1, 2, 23}};
    for (auto &q : pr) {                                   // rows 14,15,17 (:844-1000)
        const int r = q[0], j = q[1], k = q[2];
        if (!o[j].defined || !o[k].defined) out[r] = mk("U");
        else if (same(tn[r], "A")) out[r] = mk("A");
        else {
            const Box &a = B[j], &b = B[k];
            const bool in = (b.min_z < a.min_z && b.max_z > a.max_z) &&
                            ((a.max_x > b.min_x && a.max_x < b.max_x) || (a.min_x > b.min_x && a.min_x < b.max_x) ||
                             (a.min_x < b.min_x && a.max_x > b.max_x)) &&
                            (a.min_y < b.max_y && b.max_y <= a.max_y);
            const bool sa = (a.min_z < b.min_z && a.max_z > b.max_z) &&
                            ((a.max_x < b.min_x && a.max_x > b.max_x) || (a.min_x < b.min_x && a.min_x > b.max_x) ||
                             (a.min_x > b.min_x && a.max_x < b.max_x)) &&
                            (b.min_y < a.max_y && a.max_y <= b.max_y);
            out[r] = in ? mk("In") : sa ? mk("Sa") : tail(a, b, same(tn[r], "T"));
        }
        if (bw_gate) {                                     // :884-888, :925-929, :994-998
            const bool B1 = between(B[0], B[1], B[2]), B2 = between(B[1], B[0], B[2]), B3 = between(B[2], B[0], B[1]);
            // Python's `x == k and A or B` precedence inside _Bw (:621-633): row 17 loses the B1 disjunct
            const bool bw = (q[3] == 23) ? (B2 || B3) : (B1 || B2 || B3);
            if (bw) out[r] = mk("Bw");
        }
    }
    const int gr[3][2] = {{6, 0}, {8, 1}, {9, 2}};
    for (auto &q : gr) {                                   // rows 16,18,19
        if (!o[q[1]].defined) out[q[0]] = mk("U");
        else if (same(tn[q[0]], "A")) out[q[0]] = mk("A");
        else out[q[0]] = tail(B[q[1]], G, same(tn[q[0]], "T"));
    }
    (void)have;
}

static void fill_dsr(const e2sec_resolver *R, const Ent *cur /*H,G,1,2,3*/, const Ent *prev, const Code *tn, Code *out)
{
    // entity slots: 0 = H, 1 = G, 2..4 = o1..o3 (previous_array order, :1799)
    const int rows[10][2] = {{0, 2}, {0, 3}, {0, 4}, {0, 1}, {2, 3}, {2, 4}, {2, 1}, {3, 4}, {3, 1}, {4, 1}};
    for (int r = 0; r < 10; ++r) {
        const int a = rows[r][0], b = rows[r][1];
        bool undef = false;
        for (int s : {a, b})
            if (s >= 2 && (!cur[s].defined || !prev[s].defined)) undef = true;
        if (undef) { out[r] = mk("U"); continue; }
        if (same(tn[r], "A")) { out[r] = mk("A"); continue; }
        if (same(tn[r], "T")) {
            const bool moved = dist3(cur[a].center, prev[a].center) > kCenterDistance ||
                               dist3(cur[b].center, prev[b].center) > kCenterDistance;
            out[r] = mk(moved ? "MT" : "HT");
            continue;
        }
        double d, pd;
        if (b == 1) {        // ground rows use data axis 2 only (:1233, :1291, :1330, :1347)
            d = std::fabs(cur[a].center[2] - cur[1].center[2]);
            pd = std::fabs(prev[a].center[2] - prev[1].center[2]);
        } else {
            d = dist3(cur[a].center, cur[b].center);
            pd = dist3(prev[a].center, prev[b].center);
        }
        if (same(tn[r], "N") && std::fabs(d - pd) < kStableDist) out[r] = mk("S");
        else if (d > pd) out[r] = mk("MA");
        else if (d < pd) out[r] = mk("GC");
        else out[r] = mk("Q");
    }
    (void)R;
}

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
    FrameView v{R, rec, R->prev_rec.empty() ? nullptr : R->prev_rec.data()};
    const double th = R->p.thresh;
    const Ent hand = ent_from_record(rec, R->hand_idx, true);
    bool fresh[3] = {false, false, false};
    // ---- selection :1562-1658 ---------------------------------------------------------------------
    for (int i = 0; i < K; ++i) {
        if (i == R->hand_idx || i == R->ground_idx || i <= 0) continue;
        if (R->support_idx >= 0 && i == R->support_idx) continue;
        const Ent ci = ent_from_record(rec, i, false);
        if (ci.n == 0) continue;
        auto near_hand = [&]() { return v.pair(R->hand_idx, i) < th; };
        auto near_o = [&](int k) { return dist_to(v, R->o[k], fresh[k], i) < th; };
        auto take = [&](int k) { R->olabel[k] = i; R->count[k] = 1; R->o[k] = ci; fresh[k] = true; };
        if (R->count[0] == 0 && near_hand()) take(0);
        else if (R->o[0].defined && R->count[1] == 0 && i != R->olabel[0]) {
            if (near_hand() || near_o(0)) take(1);
        } else if (R->support_idx >= 0) {
            if (R->o[1].defined && R->count[2] == 0 && i != R->olabel[0] && i != R->olabel[1])
                if (near_hand() || near_o(0) || near_o(1)) take(2);
        } else if (R->o[1].defined && R->count[2] == 0 && i != R->olabel[0] && i != R->olabel[1]) {
            const int n1 = R->o[0].n, n2 = R->o[1].n;
            bool ok = false;
            if (n2 > 0 && n1 > 0) ok = near_hand() || near_o(0) || near_o(1);
            else if (n2 > 0 && n1 == 0) ok = near_hand() || near_o(1);
            else if (n1 > 0 && n2 == 0) ok = near_hand() || near_o(0);
            if (ok) take(2);
        }
    }
    for (int k = 0; k < 3; ++k)                              // icp == False :1669-1675
        if (R->count[k] == 1) R->o[k] = ent_from_record(rec, R->olabel[k], false);
    const Ent cur[5] = {hand, R->ground, R->o[0], R->o[1], R->o[2]};
    auto remember = [&]() {
        for (int s = 0; s < 5; ++s) R->prev[s] = cur[s];
        R->have_prev = true;
        R->prev_frame_id = frame;
        R->prev_rec.assign(rec, rec + R->rec);
    };
    if (R->internal_count == 0) {                            // :1798-1801 the first kept frame only primes
        remember();
        R->internal_count = 1;
        return E2SEC_OK;
    }
    Code add[30];
    fill_tn(R, v, hand, R->o, add);
    fill_ssr(R, hand, R->o, add, add + 10);
    fill_dsr(R, cur, R->prev, add, add + 20);
    remember();                                              // :1822
    bool emit = true;
    if (R->internal_count == 1) R->internal_count = 2;       // :1826-1829
    else {
        const Code *last = R->table.data() + R->table.size() - 30;
        emit = false;
        for (int r = 0; r < 30; ++r)
            if (strcmp(last[r].c, add[r].c) != 0) { emit = true; break; }   // :2001-2002
    }
    if (emit) {
        R->table.insert(R->table.end(), add, add + 30);
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
    const int n = (int)r->final_columns.size();
    if (codes)
        for (int c = 0; c < n; ++c)
            for (int row = 0; row < 30; ++row) memcpy(codes + ((size_t)row * n + c) * 6, r->table[(size_t)c * 30 + row].c, 6);
    if (final_columns)
        for (int c = 0; c < n; ++c) final_columns[c] = r->final_columns[c];
    return E2SEC_OK;
}

extern "C" int e2sec_resolver_state(const e2sec_resolver *r, int32_t labels[3], int32_t absent[3])
{
    if (!r) return E2SEC_ERR_ARG;
    for (int k = 0; k < 3; ++k) {
        if (labels) labels[k] = r->count[k] ? r->olabel[k] : -1;
        if (absent) absent[k] = r->absent[k] ? 1 : 0;
    }
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
