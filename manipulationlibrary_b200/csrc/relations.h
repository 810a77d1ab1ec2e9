// relations.h -- the three relation classifiers of the e2SEC column, ONE implementation for host and device:
//   rel_fill_tn    _fillTN_absent                 manipulation_analyser.py:363-487    rows 0-9
//   rel_fill_ssr   _fillSSR_8 + _R.._Bw           :491-1061                           rows 10-19
//   rel_fill_dsr   _fillDSR_new_2 + _distance     :1064-1353, :200-209                rows 20-29
// Scalar work on five entities (hand, ground, o1, o2, o3).  Used by the host resolver (resolver.cpp), by the device
// classifier (phase2.cuh: one thread per frame) and by the operator e2sec_relation_rows (the reference's own seam, which
// the known-answer tests generated from the reference's three functions are compared at).
// Compiled with --fmad=false / -ffp-contract=off: every float64 expression is evaluated as written.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define REL_HD __host__ __device__ __forceinline__
#else
#define REL_HD inline
#endif

// relation codes as small integers; rel_code_name() gives the reference's spelling (:389-1353)
enum RelCode : uint8_t {
    REL_NONE = 0, REL_U, REL_A, REL_T, REL_N, REL_SA, REL_IN, REL_HAR, REL_HART, REL_VAR, REL_VART, REL_O, REL_BW,
    REL_MT, REL_HT, REL_S, REL_MA, REL_GC, REL_Q,
    REL_DASH,      // '-': the column the table is created with (:2252-2260)
    REL_CODE_COUNT
};

REL_HD const char *rel_code_name(uint8_t c)
{
    switch (c) {
    case REL_U: return "U";
    case REL_A: return "A";
    case REL_T: return "T";
    case REL_N: return "N";
    case REL_SA: return "Sa";
    case REL_IN: return "In";
    case REL_HAR: return "HAr";
    case REL_HART: return "HArT";
    case REL_VAR: return "VAr";
    case REL_VART: return "VArT";
    case REL_O: return "O";
    case REL_BW: return "Bw";
    case REL_MT: return "MT";
    case REL_HT: return "HT";
    case REL_S: return "S";
    case REL_MA: return "MA";
    case REL_GC: return "GC";
    case REL_Q: return "Q";
    case REL_DASH: return "-";
    default: return "";
    }
}

#define REL_CENTER_DISTANCE 0.005107632747118286   /* :1121 */
#define REL_STABLE_DIST 0.0045302034127268085      /* :1122 */

// what the relation rows read of one cloud
struct RelEnt {
    int defined;          // 0 = the Python variable is None
    int n;                // len(points)
    double lo[3], hi[3];  // AABB (get_axis_aligned_bounding_box, :672-702)
    double center[3];     // DSR centre: hand = cloud mean (hand.get_center()), objects / ground = AABB centre (:1173-1353)
};

// slots of cur[] / prev[]: previous_array order (:1799, :1822)
#define REL_H 0
#define REL_G 1
#define REL_O1 2

// row r of every 10-row family relates slots REL_ROW_A[r], REL_ROW_B[r] (:373-383)
#define REL_ROW_PAIRS {{0, 2}, {0, 3}, {0, 4}, {0, 1}, {2, 3}, {2, 4}, {2, 1}, {3, 4}, {3, 1}, {4, 1}}

// ---- T/N rows 0-9 ------------------------------------------------------------------------------------------------------
// d[r] = np.min(a.compute_point_cloud_distance(b)) of row r's pair (only `< thresh` is asked of it; unused where a cloud is
// None or empty).  absent[3] = absent_o1..3, read and (rows 0-2 only) set: sticky (:394-396 etc.).
REL_HD void rel_fill_tn(const RelEnt *cur, const double *d, double thresh, int *absent, uint8_t *out)
{
    for (int k = 0; k < 3; ++k) {                                 // rows 0-2 (:389-427)
        const RelEnt &o = cur[REL_O1 + k];
        if (!o.defined) out[k] = REL_U;
        else if (o.n == 0 || absent[k]) { out[k] = REL_A; absent[k] = 1; }
        else out[k] = d[k] < thresh ? REL_T : REL_N;
    }
    out[3] = d[3] < thresh ? REL_T : REL_N;                       // row 3 (:429-433)
    const int pr[3][3] = {{4, 0, 1}, {5, 0, 2}, {7, 1, 2}};
    for (int q = 0; q < 3; ++q) {                                 // rows 4,5,7 (:435-451, :462-469)
        const int r = pr[q][0], j = pr[q][1], k = pr[q][2];
        const RelEnt &a = cur[REL_O1 + j], &b = cur[REL_O1 + k];
        if (!a.defined || !b.defined) out[r] = REL_U;
        else if (a.n == 0 || b.n == 0 || absent[j] || absent[k]) out[r] = REL_A;
        else out[r] = d[r] < thresh ? REL_T : REL_N;
    }
    const int gr[3][2] = {{6, 0}, {8, 1}, {9, 2}};
    for (int q = 0; q < 3; ++q) {                                 // rows 6,8,9 (:453-460, :471-487)
        const int r = gr[q][0], k = gr[q][1];
        const RelEnt &a = cur[REL_O1 + k];
        if (!a.defined) out[r] = REL_U;
        else if (a.n == 0 || absent[k]) out[r] = REL_A;
        else out[r] = d[r] < thresh ? REL_T : REL_N;
    }
}

// ---- SSR rows 10-19 ----------------------------------------------------------------------------------------------------
struct RelBox {   // SSR axis naming (:660-675): x = data[:,0], z = data[:,1], y = data[:,2]
    double min_x, max_x, min_z, max_z, min_y, max_y;
};
REL_HD RelBox rel_box_of(const RelEnt &e)
{
    RelBox b;
    b.min_x = e.lo[0]; b.max_x = e.hi[0];
    b.min_z = e.lo[1]; b.max_z = e.hi[1];
    b.min_y = e.lo[2]; b.max_y = e.hi[2];
    return b;
}
REL_HD bool rel_ov_y(const RelBox &a, const RelBox &b) { return a.min_y < b.max_y && a.max_y > b.min_y; }
REL_HD bool rel_ov_z(const RelBox &a, const RelBox &b) { return a.min_z < b.max_z && a.max_z > b.min_z; }

REL_HD bool rel_around(const RelBox &a, const RelBox &b)          // _R or _L or _F or _Ba (:491-550)
{
    const bool R = a.max_x > b.max_x && rel_ov_y(a, b) && rel_ov_z(a, b);
    const bool L = a.min_x < b.min_x && rel_ov_y(a, b) && rel_ov_z(a, b);
    const bool F = a.max_z > b.max_z && (a.min_x < b.max_x && b.min_x < a.max_x) && rel_ov_y(a, b);
    const bool Ba = a.min_z < b.min_z && (a.min_x < b.max_x && b.min_x < a.max_x) && rel_ov_y(a, b);
    return R || L || F || Ba;
}

REL_HD uint8_t rel_tail(const RelBox &a, const RelBox &b, bool touching)   // the elif chain every SSR row ends with (:729-745)
{
    if (rel_around(a, b)) return touching ? REL_HART : REL_HAR;
    if (a.max_y > b.max_y || a.min_y < b.min_y) return touching ? REL_VART : REL_VAR;   // _Ab or _Be (:570-599)
    return REL_O;
}

REL_HD bool rel_between(const RelBox &m, const RelBox &a, const RelBox &b)   // one disjunct of _Bw (:615-633)
{
    return m.min_x >= fmin(a.max_x, b.max_x) && m.max_x <= fmax(a.min_x, b.min_x) &&
           m.min_y >= fmax(a.min_y, b.min_y) && m.max_y <= fmin(a.max_y, b.max_y) &&
           m.min_z >= fmax(a.min_z, b.min_z) && m.max_z <= fmin(a.max_z, b.max_z);
}

// tn = rows 0-9 of the same column (SSR reads them: `table[k][0] == b'T'` / `== b'A'`, e.g. :729)
REL_HD void rel_fill_ssr(const RelEnt *cur, const uint8_t *tn, uint8_t *out)
{
    const RelBox H = rel_box_of(cur[REL_H]), G = rel_box_of(cur[REL_G]);
    RelBox B[3];
    for (int k = 0; k < 3; ++k) B[k] = rel_box_of(cur[REL_O1 + k]);
    const bool all_there = cur[REL_O1].defined && cur[REL_O1 + 1].defined && cur[REL_O1 + 2].defined;
    for (int k = 0; k < 3; ++k) {                                 // rows 10-12 (:706-812)
        if (!cur[REL_O1 + k].defined) out[k] = REL_U;
        else if (tn[k] == REL_A) out[k] = REL_A;
        else {
            const RelBox &b = B[k];
            const bool sa = (b.min_z < H.min_z && b.max_z > H.max_z) &&
                            ((b.max_x < H.min_x && b.max_x > H.max_x) || (b.min_x < H.min_x && b.min_x > H.max_x) ||
                             (b.min_x > H.min_x && b.max_x < H.max_x)) &&
                            (H.min_y < b.max_y && b.max_y <= H.max_y);
            out[k] = sa ? (uint8_t)REL_SA : rel_tail(H, b, tn[k] == REL_T);
        }
    }
    out[3] = rel_tail(H, G, tn[3] == REL_T);                      // row 13 (:814-842)
    const bool bw_gate = all_there && tn[0] != REL_A && tn[1] != REL_A && tn[2] != REL_A;
    const int pr[3][4] = {{4, 0, 1, 12}, {5, 0, 2, 13}, {7, 1, 2, 23}};
    for (int q = 0; q < 3; ++q) {                                 // rows 14,15,17 (:844-1000)
        const int r = pr[q][0], j = pr[q][1], k = pr[q][2];
        if (!cur[REL_O1 + j].defined || !cur[REL_O1 + k].defined) out[r] = REL_U;
        else if (tn[r] == REL_A) out[r] = REL_A;
        else {
            const RelBox &a = B[j], &b = B[k];
            const bool in = (b.min_z < a.min_z && b.max_z > a.max_z) &&
                            ((a.max_x > b.min_x && a.max_x < b.max_x) || (a.min_x > b.min_x && a.min_x < b.max_x) ||
                             (a.min_x < b.min_x && a.max_x > b.max_x)) &&
                            (a.min_y < b.max_y && b.max_y <= a.max_y);
            const bool sa = (a.min_z < b.min_z && a.max_z > b.max_z) &&
                            ((a.max_x < b.min_x && a.max_x > b.max_x) || (a.min_x < b.min_x && a.min_x > b.max_x) ||
                             (a.min_x > b.min_x && a.max_x < b.max_x)) &&
                            (b.min_y < a.max_y && a.max_y <= b.max_y);
            out[r] = in ? (uint8_t)REL_IN : sa ? (uint8_t)REL_SA : rel_tail(a, b, tn[r] == REL_T);
        }
        if (bw_gate) {                                            // :884-888, :925-929, :994-998
            const bool B1 = rel_between(B[0], B[1], B[2]), B2 = rel_between(B[1], B[0], B[2]), B3 = rel_between(B[2], B[0], B[1]);
            // Python's `x == k and A or B` precedence inside _Bw (:621-633): row 17 loses the B1 disjunct
            const bool bw = (pr[q][3] == 23) ? (B2 || B3) : (B1 || B2 || B3);
            if (bw) out[r] = REL_BW;
        }
    }
    const int gr[3][2] = {{6, 0}, {8, 1}, {9, 2}};
    for (int q = 0; q < 3; ++q) {                                 // rows 16,18,19
        const int r = gr[q][0], k = gr[q][1];
        if (!cur[REL_O1 + k].defined) out[r] = REL_U;
        else if (tn[r] == REL_A) out[r] = REL_A;
        else out[r] = rel_tail(B[k], G, tn[r] == REL_T);
    }
}

// ---- DSR rows 20-29 ----------------------------------------------------------------------------------------------------
REL_HD double rel_dist3(const double *p, const double *q)         // _distance :200-209
{
    return sqrt((p[0] - q[0]) * (p[0] - q[0]) + (p[1] - q[1]) * (p[1] - q[1]) + (p[2] - q[2]) * (p[2] - q[2]));
}

REL_HD void rel_fill_dsr(const RelEnt *cur, const RelEnt *prev, const uint8_t *tn, uint8_t *out)
{
    const int rows[10][2] = REL_ROW_PAIRS;
    for (int r = 0; r < 10; ++r) {
        const int a = rows[r][0], b = rows[r][1];
        bool undef = false;
        if (a >= REL_O1 && (!cur[a].defined || !prev[a].defined)) undef = true;
        if (b >= REL_O1 && (!cur[b].defined || !prev[b].defined)) undef = true;
        if (undef) { out[r] = REL_U; continue; }
        if (tn[r] == REL_A) { out[r] = REL_A; continue; }
        if (tn[r] == REL_T) {
            const bool moved = rel_dist3(cur[a].center, prev[a].center) > REL_CENTER_DISTANCE ||
                               rel_dist3(cur[b].center, prev[b].center) > REL_CENTER_DISTANCE;
            out[r] = moved ? REL_MT : REL_HT;
            continue;
        }
        double d, pd;
        if (b == REL_G) {        // ground rows use data axis 2 only (:1233, :1291, :1330, :1347)
            d = fabs(cur[a].center[2] - cur[REL_G].center[2]);
            pd = fabs(prev[a].center[2] - prev[REL_G].center[2]);
        } else {
            d = rel_dist3(cur[a].center, cur[b].center);
            pd = rel_dist3(prev[a].center, prev[b].center);
        }
        if (tn[r] == REL_N && fabs(d - pd) < REL_STABLE_DIST) out[r] = REL_S;
        else if (d > pd) out[r] = REL_MA;
        else if (d < pd) out[r] = REL_GC;
        else out[r] = REL_Q;
    }
}

// One column: rows 0-9, and with relations >= 2 rows 10-19, with relations == 3 rows 20-29 (:1744-1829).
REL_HD void rel_fill_column(const RelEnt *cur, const RelEnt *prev, const double *d, double thresh, int relations, int *absent,
                            uint8_t *out)
{
    rel_fill_tn(cur, d, thresh, absent, out);
    if (relations >= 2) rel_fill_ssr(cur, out, out + 10);
    if (relations >= 3) rel_fill_dsr(cur, prev, out, out + 20);
}


// ---- records (include/e2sec.h) -> what the rows read --------------------------------------------------------------------
#define REL_REC_HEADER 4
#define REL_REC_ENTITY 16
REL_HD RelEnt rel_ent_undefined()
{
    RelEnt x;
    x.defined = 0; x.n = 0;
    for (int k = 0; k < 3; ++k) x.lo[k] = x.hi[k] = x.center[k] = 0.0;
    return x;
}
REL_HD RelEnt rel_ent_from_record(const double *rec, int e, bool is_hand)
{
    const double *r = rec + REL_REC_HEADER + REL_REC_ENTITY * e;
    RelEnt x;
    x.defined = 1;
    x.n = (int)r[1];
    for (int k = 0; k < 3; ++k) {
        x.lo[k] = r[7 + k];
        x.hi[k] = r[10 + k];
        x.center[k] = is_hand ? r[13 + k] : (x.lo[k] + x.hi[k]) * 0.5;
    }
    return x;
}
REL_HD double rel_rec_pair(const double *rec, int K, int a, int b)     // in-frame minimum distance, index K = the ground cloud
{
    const int W = K + 1;
    const double *pt = rec + REL_REC_HEADER + REL_REC_ENTITY * K;
    return a < b ? pt[a * W + b] : pt[b * W + a];
}
REL_HD double rel_rec_cross(const double *rec, int K, int a_prev, int b)   // entity a of the previous kept frame vs b of this frame
{
    const int W = K + 1;
    return rec[REL_REC_HEADER + REL_REC_ENTITY * K + W * W + a_prev * K + b];
}

// ---- o1 / o2 / o3 selection (:1562-1658), one kept frame -----------------------------------------------------------------
// The bindings are sticky and first come, first served in ascending entity index.  While the loop runs, o_k is still the
// cloud of the previous kept frame unless it was bound in this very frame (:1669-1675 rebinds after the loop), so the
// distances to o_k come from the cross table (stale) or the in-frame table (fresh).  stale_n[k] = len(o_k.points) of the
// stale cloud.  np.min over an empty distance array raises in the reference; here an empty stale cloud is "far".
struct RelSel {
    int oindex[3];      // entity index bound to o1..o3 (-1: none)
    int count[3];       // count1..3
};
REL_HD bool rel_select(const double *rec, int K, int hand_idx, int ground_idx, int support_idx, double th, RelSel &S,
                       const int *stale_n)
{
    bool fresh[3] = {false, false, false}, changed = false;
    int on[3] = {stale_n[0], stale_n[1], stale_n[2]};
    if (S.count[0] && S.count[1] && S.count[2]) return false;
    for (int i = 1; i < K; ++i) {
        if (i == hand_idx || i == ground_idx || i == support_idx) continue;
        const int ni = (int)rec[REL_REC_HEADER + REL_REC_ENTITY * i + 1];
        if (ni == 0) continue;
        const bool near_hand = rel_rec_pair(rec, K, hand_idx, i) < th;
        bool near_k[2];
        for (int k = 0; k < 2; ++k)
            near_k[k] = S.count[k] && on[k] != 0 &&
                        (fresh[k] ? rel_rec_pair(rec, K, S.oindex[k], i) : rel_rec_cross(rec, K, S.oindex[k], i)) < th;
        int take = -1;
        if (S.count[0] == 0 && near_hand) take = 0;
        else if (S.count[0] && S.count[1] == 0 && i != S.oindex[0]) {
            if (near_hand || near_k[0]) take = 1;
        } else if (S.count[1] && S.count[2] == 0 && i != S.oindex[0] && i != S.oindex[1]) {
            bool ok;
            if (support_idx >= 0) ok = near_hand || near_k[0] || near_k[1];
            else if (on[1] > 0 && on[0] > 0) ok = near_hand || near_k[0] || near_k[1];
            else if (on[1] > 0 && on[0] == 0) ok = near_hand || near_k[1];
            else if (on[0] > 0 && on[1] == 0) ok = near_hand || near_k[0];
            else ok = false;
            if (ok) take = 2;
        }
        if (take >= 0) { S.oindex[take] = i; S.count[take] = 1; on[take] = ni; fresh[take] = true; changed = true; }
    }
    return changed;
}
