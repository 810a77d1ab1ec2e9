// phase2.cuh -- the relation classifiers with their frame-to-frame deltas ON THE DEVICE: phase 2 of the e2SEC path.
//
// The reference runs, per kept frame and strictly in order (manipulation_analyser.py):
//   o1 / o2 / o3 selection   :1562-1658, :1669-1675      sticky bindings, at most three changes in a whole sequence
//   _fillTN_absent           :363-487                    + sticky absent_o1..3
//   _fillSSR_8               :636-1061
//   _fillDSR_new_2           :1064-1353                  against previous_array = the previous kept frame
//   column emit / dedup      :1744-1829, :2000-2007      a column is kept when it differs from the last one
// What makes that sequential is tiny: the three bindings, three sticky bits, the previous kept frame and the last
// column.  Here ONE kernel classifies a whole run of phase-1 records (include/e2sec.h) in parallel, one thread per frame,
// 1024 frames per round, carrying that state (Phase2State) from round to round and from call to call:
//   1  previous kept frame of every frame: exclusive max-scan over the kept flags
//   2  selection by speculation: every kept frame evaluates rel_select() from the CURRENT bindings; the first frame that
//      changes them wins (block minimum), the frames behind it re-evaluate -- at most three winners per sequence
//   3  sticky absent bits: inclusive OR-scan over the frames
//   4  rel_fill_column() per frame (relations.h -- the very code the host resolver and e2sec_relation_rows run)
//   5  dedup = compare with the previous classified frame's column (equality with the last EMITTED column is transitive),
//      exclusive sum-scan -> compacted columns (32 bytes + frame id each), the only thing that ever leaves the GPU
// A shard of a frame-sharded sequence starts from the state its predecessor ended in (e2sec_phase2_state_get / _set).
#pragma once
#include <limits.h>

#include "common.cuh"
#include "relations.h"

#define E2_P2_THREADS 1024
#define E2_P2_COL 32               // bytes per packed column (30 codes used)

struct Phase2State {               // POD; e2sec.h: E2SEC_PHASE2_STATE_BYTES
    RelSel sel;                    // bindings after the last kept frame
    int absent[3];                 // absent_o1..3
    int internal_count;            // relations == 3: 0 no kept frame yet, 1 primed (:1798-1801), 2 columns exist
    int have_last;                 // `last` holds the previous classified column
    int prev_frame_id;             // the previous kept frame, -1: none
    int n_columns;                 // columns emitted along the whole chain of shards so far
    int error;                     // E2SEC_ERR_STATE: a record was measured against another previous kept frame
    RelEnt prev[5];                // previous_array (:1799, :1822)
    unsigned char last[E2_P2_COL];
};

// A rank's inbox for the state of a frame-sharded sequence (e2sec_phase2_run_ring): the kernel of the rank that ran the
// batch in front writes the state and then the ticket straight into this memory over NVLink (peer access through a CUDA IPC
// handle); the kernel that needs it polls the ticket.  No host, no collective on the way.
struct Phase2Mailbox {
    Phase2State state;
    unsigned ticket;
    unsigned pad[3];
};

struct Phase2Dev {
    const double *records;
    int n_frames, rec_doubles;
    int K, hand_idx, ground_idx, support_idx, relations;
    double thresh;
    RelEnt ground;
    Phase2State *state;
    unsigned char *cols;           // max_columns x E2_P2_COL
    int *col_frame;                // max_columns
    int *n_cols;                   // columns this object holds (a shard's own, unlike state->n_columns)
    int max_columns;
    Phase2Mailbox *inbox;          // this rank's mailbox (local memory, written by the previous rank)
    Phase2Mailbox *outbox;         // the next rank's mailbox (peer memory)
    unsigned wait_ticket;          // != 0: start from the state posted under this ticket instead of *state
    unsigned post_ticket;          // != 0: post the final state to the outbox under this ticket
    unsigned long long timeout_ns; // give up waiting after this long (E2SEC_ERR_STATE)
};

// inclusive block scans over E2_P2_THREADS threads; `tot` (shared, 32 ints) is scratch; every thread also gets the total
template <class Op>
__device__ __forceinline__ int p2_scan(int v, int ident, Op op, int *tot, int &total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v = op(v, t); }
    __syncthreads();                               // the previous use of tot is over
    if (lane == 31) tot[warp] = v;
    __syncthreads();
    int w = lane < E2_P2_THREADS / 32 ? tot[lane] : ident;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w = op(w, t); }
    total = __shfl_sync(0xffffffffu, w, E2_P2_THREADS / 32 - 1);
    const int before = __shfl_sync(0xffffffffu, w, warp > 0 ? warp - 1 : 0);
    return warp > 0 ? op(before, v) : v;
}

__global__ void __launch_bounds__(E2_P2_THREADS) phase2_kernel(Phase2Dev a)
{
    pdl_enter();
    __shared__ Phase2State st;
    __shared__ int tot[32];
    __shared__ RelSel s_sel[4];
    __shared__ int s_chg[4], s_nchg;
    __shared__ unsigned char s_col[E2_P2_THREADS][E2_P2_COL];
    const int tid = threadIdx.x, K = a.K, rows = 10 * a.relations;
    auto op_max = [](int x, int y) { return x > y ? x : y; };
    auto op_or = [](int x, int y) { return x | y; };
    auto op_add = [](int x, int y) { return x + y; };
    if (a.wait_ticket) {
        // the state of everything in front of these frames comes from the rank that ran it: wait for its ticket
        __shared__ int arrived;
        if (tid == 0) {
            unsigned long long t0;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
            int ok = 1;
            while (*((volatile unsigned *)&a.inbox->ticket) != a.wait_ticket) {
                __nanosleep(100);
                unsigned long long t1;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
                if (t1 - t0 > a.timeout_ns) { ok = 0; break; }
            }
            __threadfence_system();
            arrived = ok;
        }
        __syncthreads();
        const volatile unsigned *src = (const volatile unsigned *)&a.inbox->state;
        unsigned *dst = (unsigned *)&st;
        if (arrived) for (int i = tid; i < (int)(sizeof(Phase2State) / 4); i += E2_P2_THREADS) dst[i] = src[i];
        else for (int i = tid; i < (int)(sizeof(Phase2State) / 4); i += E2_P2_THREADS) dst[i] = ((const unsigned *)a.state)[i];
        __syncthreads();
        if (!arrived && tid == 0) st.error = E2SEC_ERR_STATE;
    } else if (tid == 0) st = *a.state;
    __syncthreads();
    for (int base = 0; base < a.n_frames; base += E2_P2_THREADS) {
        const int f = base + tid;
        const bool live = f < a.n_frames;
        const double *rec = a.records + (size_t)(live ? f : base) * a.rec_doubles;
        const bool kept = live && ((int)rec[1] & 3) == 3;                     // :1429, :1475 / :1551
        const int frame_id = (int)rec[0];
        // ---- 1: previous kept frame inside the round (-1: the carried one) ------------------------------------------
        int last_kept;
        const int incl = p2_scan(kept ? tid : -1, -1, op_max, tot, last_kept);
        int pk = __shfl_up_sync(0xffffffffu, incl, 1);
        if ((tid & 31) == 0) { pk = -1; for (int w = 0; w < (tid >> 5); ++w) pk = max(pk, tot[w]); }   // tot = the warps' own maxima
        const double *recp = pk >= 0 ? a.records + (size_t)(base + pk) * a.rec_doubles : nullptr;
        if (kept && (int)rec[3] != (pk >= 0 ? (int)recp[0] : st.prev_frame_id)) st.error = E2SEC_ERR_STATE;   // benign race: any writer
        // ---- 2: selection ------------------------------------------------------------------------------------------
        if (tid == 0) { s_sel[0] = st.sel; s_chg[0] = -1; s_nchg = 1; }
        __syncthreads();
        for (int it = 0; it < 4; ++it) {
            const int n = s_nchg;
            RelSel S = s_sel[n - 1];
            bool ch = false;
            if (kept && tid > s_chg[n - 1] && !(S.count[0] && S.count[1] && S.count[2])) {
                int stale_n[3];
                for (int k = 0; k < 3; ++k)
                    stale_n[k] = !S.count[k] ? 0 : pk >= 0 ? (int)recp[REL_REC_HEADER + REL_REC_ENTITY * S.oindex[k] + 1] : st.prev[REL_O1 + k].n;
                ch = rel_select(rec, K, a.hand_idx, a.ground_idx, a.support_idx, a.thresh, S, stale_n);
            }
            int first;
            p2_scan(ch ? -tid : INT_MIN, INT_MIN, op_max, tot, first);         // max of -tid = the first changing frame
            if (first == INT_MIN) break;
            if (tid == -first) { s_sel[n] = S; s_chg[n] = tid; s_nchg = n + 1; }
            __syncthreads();
        }
        __syncthreads();
        RelSel Sf = s_sel[0], Sp = s_sel[0];
        for (int j = 1; j < s_nchg; ++j) {
            if (s_chg[j] <= tid) Sf = s_sel[j];
            if (s_chg[j] <= pk) Sp = s_sel[j];
        }
        // ---- 3: who is classified, sticky absent bits ---------------------------------------------------------------
        const bool priming_round = a.relations == 3 && st.internal_count == 0;  // the first kept frame only primes (:1798-1801)
        const bool classified = kept && !(priming_round && pk < 0);
        RelEnt cur[5];
        if (kept) {
            cur[REL_H] = rel_ent_from_record(rec, a.hand_idx, true);
            cur[REL_G] = a.ground;
            for (int k = 0; k < 3; ++k) cur[REL_O1 + k] = Sf.count[k] ? rel_ent_from_record(rec, Sf.oindex[k], false) : rel_ent_undefined();
        }
        int az = 0;
        if (classified)
            for (int k = 0; k < 3; ++k) az |= (cur[REL_O1 + k].defined && cur[REL_O1 + k].n == 0) ? (1 << k) : 0;
        int az_total;
        const int az_incl = p2_scan(az, 0, op_or, tot, az_total) | (st.absent[0] ? 1 : 0) | (st.absent[1] ? 2 : 0) | (st.absent[2] ? 4 : 0);
        // ---- 4: the column ------------------------------------------------------------------------------------------
        unsigned char *col = s_col[tid];
        if (classified) {
            RelEnt prev[5];
            if (pk >= 0) {
                prev[REL_H] = rel_ent_from_record(recp, a.hand_idx, true);
                prev[REL_G] = a.ground;
                for (int k = 0; k < 3; ++k) prev[REL_O1 + k] = Sp.count[k] ? rel_ent_from_record(recp, Sp.oindex[k], false) : rel_ent_undefined();
            } else {
                for (int s = 0; s < 5; ++s) prev[s] = st.prev[s];
            }
            double d[10];
            const int slot_idx[5] = {a.hand_idx, K, Sf.oindex[0], Sf.oindex[1], Sf.oindex[2]};
            const int pairs[10][2] = REL_ROW_PAIRS;
            for (int r = 0; r < 10; ++r) {
                const int x = slot_idx[pairs[r][0]], y = slot_idx[pairs[r][1]];
                const bool ok = cur[pairs[r][0]].defined && cur[pairs[r][1]].defined && x >= 0 && y >= 0;
                d[r] = ok ? rel_rec_pair(rec, K, x, y) : __longlong_as_double(0x7ff0000000000000ll);
            }
            int absent[3] = {az_incl & 1, (az_incl >> 1) & 1, (az_incl >> 2) & 1};
            uint8_t out[E2_P2_COL];
            for (int r = 0; r < E2_P2_COL; ++r) out[r] = 0;
            rel_fill_column(cur, prev, d, a.thresh, a.relations, absent, out);
            for (int r = 0; r < E2_P2_COL; ++r) col[r] = out[r];
        }
        __syncthreads();
        // ---- 5: dedup against the previous classified column, compaction -------------------------------------------
        bool emit = false;
        // the priming frame (if any) is the first kept frame of the round: its index
        int first_kept;
        p2_scan(kept ? -tid : INT_MIN, INT_MIN, op_max, tot, first_kept);
        const int priming_idx = (priming_round && first_kept != INT_MIN) ? -first_kept : -1;
        if (classified) {
            const unsigned char *pc = nullptr;
            if (pk >= 0 && pk != priming_idx) pc = s_col[pk];
            else if (pk < 0 && st.have_last) pc = st.last;
            if (!pc) emit = true;                                  // the first column ever (:1826-1829); relations 1 / 2: differs from '-'
            else {
                for (int r = 0; r < rows; ++r) emit |= pc[r] != col[r];
            }
        }
        int n_emit;
        const int pos = p2_scan(emit ? 1 : 0, 0, op_add, tot, n_emit) - (emit ? 1 : 0);
        const int have = *a.n_cols;                                 // written by thread 0 after the barrier below
        if (emit) {
            if (have + pos < a.max_columns) {
                unsigned char *dst = a.cols + (size_t)(have + pos) * E2_P2_COL;
                for (int r = 0; r < E2_P2_COL; ++r) dst[r] = col[r];
                a.col_frame[have + pos] = frame_id;
            } else st.error = E2SEC_ERR_CAPACITY;
        }
        int last_classified;
        p2_scan(classified ? tid : -1, -1, op_max, tot, last_classified);
        __syncthreads();                                            // every read of st / *a.n_cols of this round is done
        // ---- carry ---------------------------------------------------------------------------------------------------
        if (kept && tid == last_kept) {
            for (int s = 0; s < 5; ++s) st.prev[s] = cur[s];
            st.prev_frame_id = frame_id;
            st.sel = s_sel[s_nchg - 1];
            if (a.relations == 3 && st.internal_count == 0) st.internal_count = 1;
        }
        if (classified && tid == last_classified) {
            for (int r = 0; r < E2_P2_COL; ++r) st.last[r] = col[r];
            st.have_last = 1;
        }
        __syncthreads();
        if (tid == 0) {
            for (int k = 0; k < 3; ++k) st.absent[k] = (st.absent[k] || ((az_total >> k) & 1)) ? 1 : 0;
            if (last_classified >= 0 && a.relations == 3) st.internal_count = 2;
            st.n_columns += n_emit;
            *a.n_cols = min(have + n_emit, a.max_columns);
        }
        __syncthreads();
    }
    if (tid == 0) *a.state = st;
    if (a.post_ticket) {
        // hand the state to the rank that runs the next frames: payload, system-wide fence, then the ticket
        __syncthreads();
        volatile unsigned *dst = (volatile unsigned *)&a.outbox->state;
        const unsigned *src = (const unsigned *)&st;
        for (int i = tid; i < (int)(sizeof(Phase2State) / 4); i += E2_P2_THREADS) dst[i] = src[i];
        __threadfence_system();
        __syncthreads();
        if (tid == 0) { *((volatile unsigned *)&a.outbox->ticket) = a.post_ticket; __threadfence_system(); }
    }
}
