// e2sec_api.cu -- context, buffers and the C ABI of include/e2sec.h (device side).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -lineinfo (see build.py).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <cmath>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "seg.cuh"
#include "seg_fused.cuh"
#include "seg_stream.cuh"
#include "select.cuh"
#include "radius.cuh"
#include "filter.cuh"
#include "cells.cuh"
#include "pairs.cuh"
#include "contact.cuh"
#include "dissim.cuh"
#include "phase2.cuh"
#include "lzf.cuh"
#include "frame0.cuh"

int e2sec_pcd_locate(const unsigned char *file, size_t len, long long *payload, unsigned *comp, long long *points, int *width, int *height);   // pcd_reader.cpp

#define E2_MAX_REFINE 6        // histogram rounds launched per batch (each narrows the key interval 256x)

static std::string g_create_error;

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
};

struct e2sec_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;      // H2D of host inputs, double-buffered against the kernels of the previous batch
    cudaEvent_t copy_done[2], slot_free[2];
    bool slot_used[2] = {false, false};
    unsigned stage_idx = 0;
    std::string err;
    e2sec_params params;
    std::vector<double> ground;
    bool configured = false;
    int maxF = 0, N = 0, K = 0;
    size_t cap_total = 0;          // stored-point capacity over the whole batch (maxF * N)
    Dev d;
    std::vector<DevBuf *> bufs;
    DevBuf lab8_stage, gsub, amb_count, cell_state, amb_list, exact_ticket, plan_split;
    DevBuf pcd_stage, pcd_frames, pcd_status, ltext_stage, ltext_off;       // e2sec_process_batch_pcd: file bytes, per-frame jobs
    struct PcdJob {                                                          // the batch e2sec_process_batch_pcd is handing over
        const unsigned char *files = nullptr, *ltext = nullptr;
        size_t files_bytes = 0, ltext_bytes = 0;
        std::vector<LzfFrame> frames;
        std::vector<long long> loff;
    } pcd;
    DevBuf ds_codes, ds_off, ds_ncols, ds_packed, ds_counts, ds_D;     // e2sec_dissimilarity
    cudaEvent_t ds_ev[2];
    bool ds_timed = false;
    DevBuf xyz_stage, lab_stage, rowmap, colmap, lut, frame_ids, seg_cnt, ent_raw, ent_off, flags, px, py, pz,
        pidx, med, radius, rs, hist, cand, cand_n, seed_pos, member, front_a, front_b, flist, n_list, aabb, mean,
        gx, gy, gz, pair_key, cross_key, records, error, plan_off, plan_total, below_acc, min_above, single_count,
        frame_used, prev_slot, prev_id, halo_info, ent_box, pcell, rad_lo, rad_hi, amb, ct_cnt, ct_off, ct_next, ct_map, seg_list, seg_len, seg_next;
    int map_cfg[4] = {0, 0, 0, 0};
    int64_t launches = 0;
    cudaEvent_t ev[10];            // [0..7] stage boundaries, [8] / [9] around the streaming segmentation kernel
    bool ev_stream = false;
    bool ev_valid = false;
    int sm_count = 148;
    int sf_cluster = 8;
    bool seg_cluster = false;      // E2SEC_SEG_MODE=cluster: the one-kernel cluster segmentation instead of the unit stream
    bool pdl = true;
    bool pdl_once_off = false;  // the next launch_pdl launches in plain stream order (launch_plain)
    cudaError_t launch_err = cudaSuccess;
    bool force_two_pass = false;   // E2SEC_SEG_TWO_PASS=1: always take the two-pass segmentation (tests compare both)
};

#define CU(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);               \
            return E2SEC_ERR_CUDA;                                                       \
        }                                                                                \
    } while (0)

static int fail(e2sec_ctx *ctx, int code, const std::string &msg)
{
    if (ctx) ctx->err = msg; else g_create_error = msg;
    return code;
}

static int ensure(e2sec_ctx *ctx, DevBuf &b, size_t bytes)
{
    if (b.bytes >= bytes && b.p) return 0;
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.bytes = 0;
    CU(cudaMalloc(&b.p, bytes ? bytes : 16));
    b.bytes = bytes ? bytes : 16;
    return 0;
}

extern "C" int e2sec_abi_version(void) { return E2SEC_ABI_VERSION; }

extern "C" int e2sec_device_count(void)
{
    int count = 0;
    return cudaGetDeviceCount(&count) == cudaSuccess ? count : 0;
}

extern "C" size_t e2sec_record_doubles(int32_t K)
{
    return (size_t)E2SEC_REC_HEADER + (size_t)E2SEC_REC_ENTITY * K + (size_t)(K + 1) * (K + 1) + (size_t)K * K + (size_t)2 * K;
}

extern "C" const char *e2sec_last_error(const e2sec_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

extern "C" int e2sec_create(int device, e2sec_ctx **out)
{
    if (!out) return fail(nullptr, E2SEC_ERR_ARG, "e2sec_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, E2SEC_ERR_CUDA, std::string("e2sec_create: no CUDA device (") + cudaGetErrorString(e) +
                                                 "); this path has no CPU fallback");
    if (device < 0 || device >= count) return fail(nullptr, E2SEC_ERR_ARG, "e2sec_create: bad device index");
    e2sec_ctx *ctx = new e2sec_ctx();
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, E2SEC_ERR_CUDA, "e2sec_create: cannot create a stream");
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    cudaError_t ce = cudaSuccess;
    auto keep = [&](cudaError_t r) { if (ce == cudaSuccess) ce = r; };
    for (auto &ev : ctx->ev) keep(cudaEventCreate(&ev));
    for (auto &ev : ctx->ds_ev) keep(cudaEventCreate(&ev));
    keep(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        keep(cudaEventCreateWithFlags(&ctx->copy_done[i], cudaEventDisableTiming));
        keep(cudaEventCreateWithFlags(&ctx->slot_free[i], cudaEventDisableTiming));
    }
    if (ce != cudaSuccess) {
        const std::string msg = std::string("e2sec_create: cannot create events / copy stream: ") + cudaGetErrorString(ce);
        cudaStreamDestroy(ctx->stream);
        delete ctx;
        return fail(nullptr, E2SEC_ERR_CUDA, msg);
    }
    ctx->ev_valid = true;
    cudaFuncSetAttribute(pair_pass_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)E2_PAIR_SMEM_HIST);
    cudaFuncSetAttribute(cell_filter_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CellSmem));
    cudaFuncSetAttribute(cell_filter_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CellSmem));
    cudaFuncSetAttribute(seg_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegFusedSmem));
    cudaFuncSetAttribute(seg_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegFusedSmem));
    cudaFuncSetAttribute(seg_stream_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegStreamSmem));
    cudaFuncSetAttribute(seg_stream_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegStreamSmem));
    { const char *e = getenv("E2SEC_SEG_TWO_PASS"); ctx->force_two_pass = e && e[0] == '1'; }
    { const char *e = getenv("E2SEC_SEG_MODE"); ctx->seg_cluster = e && !strcmp(e, "cluster"); }
    { const char *e = getenv("E2SEC_PDL"); ctx->pdl = !(e && e[0] == '0'); }
    { const char *e = getenv("E2SEC_SF_CLUSTER"); int c = e ? atoi(e) : E2_SF_CLUSTER; ctx->sf_cluster = (c == 1 || c == 2 || c == 4 || c == 8) ? c : E2_SF_CLUSTER; }
    cudaFuncSetAttribute(cell_filter_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    // The exact path's kernels follow each other under programmatic launch: a pass's CTAs (72 KB each, three per SM) arrive
    // while the small kernel before it is still resident, and an SM cannot change its shared-memory carve-out with CTAs on it
    // -- with the default preference the pass ran with fewer CTAs per SM, 1.5x slower.  Same carve-out for the whole chain.
    cudaFuncSetAttribute(cell_filter_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(radius_init_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(radius_select_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(radius_final_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(radius_plan_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(pair_pass_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(pair_pass_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    memset(&ctx->d, 0, sizeof(Dev));
    { const char *e = getenv("E2SEC_TEST_EXACT_START"); ctx->d.test_exact_start = e ? atoi(e) : 0; }     // radius.cuh: forces the reopening paths
    ctx->bufs = {&ctx->pcd_stage, &ctx->pcd_frames, &ctx->pcd_status, &ctx->ltext_stage, &ctx->ltext_off, &ctx->lab8_stage, &ctx->gsub, &ctx->amb_count, &ctx->cell_state, &ctx->amb_list, &ctx->exact_ticket, &ctx->plan_split, &ctx->xyz_stage, &ctx->lab_stage, &ctx->rowmap, &ctx->colmap, &ctx->lut, &ctx->frame_ids, &ctx->seg_cnt,
                 &ctx->ent_raw, &ctx->ent_off, &ctx->flags, &ctx->px, &ctx->py, &ctx->pz, &ctx->pidx, &ctx->med,
                 &ctx->radius, &ctx->rs, &ctx->hist, &ctx->cand, &ctx->cand_n, &ctx->seed_pos, &ctx->member,
                 &ctx->front_a, &ctx->front_b, &ctx->flist, &ctx->n_list, &ctx->aabb, &ctx->mean, &ctx->gx, &ctx->gy,
                 &ctx->gz, &ctx->pair_key, &ctx->records, &ctx->error, &ctx->plan_off, &ctx->plan_total,
                 &ctx->below_acc, &ctx->min_above, &ctx->single_count, &ctx->cross_key, &ctx->frame_used, &ctx->prev_slot,
                 &ctx->prev_id, &ctx->halo_info, &ctx->ent_box, &ctx->pcell, &ctx->rad_lo, &ctx->rad_hi, &ctx->amb, &ctx->ct_cnt, &ctx->ct_off, &ctx->ct_next, &ctx->ct_map, &ctx->seg_list, &ctx->seg_len, &ctx->seg_next,
                 &ctx->ds_codes, &ctx->ds_off, &ctx->ds_ncols, &ctx->ds_packed, &ctx->ds_counts, &ctx->ds_D};
    *out = ctx;
    return E2SEC_OK;
}

extern "C" void e2sec_destroy(e2sec_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (DevBuf *b : ctx->bufs)
        if (b->p) cudaFree(b->p);
    if (ctx->ev_valid)
        for (auto &ev : ctx->ev) cudaEventDestroy(ev);
    if (ctx->ev_valid)
        for (auto &ev : ctx->ds_ev) cudaEventDestroy(ev);
    if (ctx->copy_stream) {
        cudaStreamSynchronize(ctx->copy_stream);
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(ctx->copy_done[i]); cudaEventDestroy(ctx->slot_free[i]); }
        cudaStreamDestroy(ctx->copy_stream);
    }
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" void *e2sec_stream(e2sec_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
extern "C" int64_t e2sec_launch_count(const e2sec_ctx *ctx) { return ctx ? ctx->launches : 0; }
extern "C" const double *e2sec_records_dev(e2sec_ctx *ctx) { return ctx ? (const double *)ctx->records.p : nullptr; }

static int check_device_error(e2sec_ctx *ctx);

extern "C" int e2sec_sync(e2sec_ctx *ctx)
{
    if (!ctx) return E2SEC_ERR_ARG;
    if (!ctx->error.p) { CU(cudaStreamSynchronize(ctx->stream)); return E2SEC_OK; }
    return check_device_error(ctx);       // waits for the stream, reports (and clears) a device-side error of any batch since
}

extern "C" int e2sec_reset_sequence(e2sec_ctx *ctx)
{
    if (!ctx) return E2SEC_ERR_ARG;
    if (!ctx->halo_info.p) return E2SEC_OK;
    // stream-ordered: the batches queued before it still see the old halo, the ones after it start a sequence
    CU(cudaMemsetAsync(ctx->halo_info.p, 0, 4 * sizeof(int), ctx->stream));
    return E2SEC_OK;
}

// Allocates every work buffer for F frames x N points x K entities and points ctx->d at them.
static int allocate(e2sec_ctx *ctx, int F, int N, int K)
{
    const size_t cap = (size_t)(F + 1) * N;      // + the halo slot
    const int S = (N + E2_SEG - 1) / E2_SEG;
    const size_t FK = (size_t)(F + 1) * K;
    int rc = 0;
    if ((rc = ensure(ctx, ctx->xyz_stage, (size_t)2 * F * N * 3 * sizeof(float)))) return rc;     // two staging slots
    if ((rc = ensure(ctx, ctx->lab_stage, (size_t)2 * F * N * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->cross_key, (size_t)F * K * K * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->frame_used, (size_t)(F + 1) * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->prev_slot, (size_t)F * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->prev_id, (size_t)F * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->halo_info, 4 * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->lut, E2SEC_LABEL_LUT * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->frame_ids, F * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->seg_cnt, (size_t)F * S * K * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->ent_raw, FK * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->ent_box, FK * 6 * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->ent_off, FK * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->flags, F * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->px, cap * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->py, cap * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->pz, cap * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->pidx, cap * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->med, FK * 3 * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->radius, FK * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->rs, FK * sizeof(RadiusState)))) return rc;
    if ((rc = ensure(ctx, ctx->hist, FK * E2_NBINS * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->cand, FK * E2_CAND_CAP * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->cand_n, FK * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->seed_pos, FK * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->member, cap))) return rc;
    if ((rc = ensure(ctx, ctx->front_a, cap * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->front_b, cap * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->flist, cap * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->n_list, FK * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->aabb, FK * 6 * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->mean, FK * 3 * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->pair_key, (size_t)F * (K + 1) * (K + 1) * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->records, (size_t)F * e2sec_record_doubles(K) * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->error, sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->plan_off, (FK + 1) * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->plan_total, sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->below_acc, FK * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->min_above, FK * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->single_count, FK * sizeof(u64)))) return rc;
    if ((rc = ensure(ctx, ctx->pcell, cap * sizeof(unsigned short)))) return rc;
    if ((rc = ensure(ctx, ctx->rad_lo, FK * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->rad_hi, FK * sizeof(double)))) return rc;
    if ((rc = ensure(ctx, ctx->amb, FK * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->amb_count, sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->amb_list, FK * sizeof(int)))) return rc;
    if ((rc = ensure(ctx, ctx->exact_ticket, sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->plan_split, sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->cell_state, FK * sizeof(CellState)))) return rc;
    const size_t slots = (size_t)K * (K - 1) / 2 + K + (size_t)K * K;
    if ((rc = ensure(ctx, ctx->ct_cnt, (size_t)F * slots * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->ct_off, ((size_t)F * slots + 1) * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->ct_next, sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->ct_map, std::max((size_t)F * slots * 4, (size_t)65536) * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->seg_list, (size_t)F * N * sizeof(uint4)))) return rc;
    if ((rc = ensure(ctx, ctx->seg_len, (size_t)F * ((N + E2_SS_UNIT - 1) / E2_SS_UNIT) * sizeof(unsigned)))) return rc;
    if ((rc = ensure(ctx, ctx->seg_next, sizeof(unsigned)))) return rc;
    CU(cudaMemset(ctx->error.p, 0, sizeof(int)));
    CU(cudaMemset(ctx->seg_next.p, 0, sizeof(unsigned)));
    CU(cudaMemset(ctx->amb_count.p, 0, sizeof(int)));
    CU(cudaMemset(ctx->exact_ticket.p, 0, sizeof(unsigned)));
    ctx->maxF = F;
    ctx->N = N;
    ctx->K = K;
    ctx->cap_total = ctx->px.bytes / sizeof(double);
    Dev &d = ctx->d;
    d.maxF = F;
    d.cross_key = (u64 *)ctx->cross_key.p;
    d.frame_used = (int *)ctx->frame_used.p;
    d.prev_slot = (int *)ctx->prev_slot.p;
    d.prev_id = (int *)ctx->prev_id.p;
    d.halo_info = (int *)ctx->halo_info.p;
    d.lut = (const int *)ctx->lut.p;
    d.frame_ids = (const int *)ctx->frame_ids.p;
    d.seg_cnt = (unsigned *)ctx->seg_cnt.p;
    d.ent_raw = (int *)ctx->ent_raw.p;
    d.ent_off = (int *)ctx->ent_off.p;
    d.flags = (int *)ctx->flags.p;
    d.px = (double *)ctx->px.p;
    d.py = (double *)ctx->py.p;
    d.pz = (double *)ctx->pz.p;
    d.pidx = (unsigned *)ctx->pidx.p;
    d.med = (double *)ctx->med.p;
    d.radius = (double *)ctx->radius.p;
    d.rs = (RadiusState *)ctx->rs.p;
    d.hist = (u64 *)ctx->hist.p;
    d.cand = (u64 *)ctx->cand.p;
    d.cand_n = (unsigned *)ctx->cand_n.p;
    d.seed_pos = (int *)ctx->seed_pos.p;
    d.member = (unsigned char *)ctx->member.p;
    d.front_a = (unsigned *)ctx->front_a.p;
    d.front_b = (unsigned *)ctx->front_b.p;
    d.flist = (unsigned *)ctx->flist.p;
    d.n_list = (int *)ctx->n_list.p;
    d.aabb = (double *)ctx->aabb.p;
    d.mean = (double *)ctx->mean.p;
    d.pair_key = (u64 *)ctx->pair_key.p;
    d.records = (double *)ctx->records.p;
    d.error = (int *)ctx->error.p;
    d.pcell = (unsigned short *)ctx->pcell.p;
    d.rad_lo = (double *)ctx->rad_lo.p;
    d.rad_hi = (double *)ctx->rad_hi.p;
    d.amb = (int *)ctx->amb.p;
    d.amb_count = (int *)ctx->amb_count.p;
    d.amb_list = (int *)ctx->amb_list.p;
    d.exact_ticket = (unsigned *)ctx->exact_ticket.p;
    d.cell_state = (CellState *)ctx->cell_state.p;
    d.seg_list = (uint4 *)ctx->seg_list.p;
    d.seg_len = (unsigned *)ctx->seg_len.p;
    return E2SEC_OK;
}

extern "C" int e2sec_configure(e2sec_ctx *ctx, const e2sec_params *p, int32_t max_frames, int32_t n_points)
{
    if (!ctx || !p) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (p->n_entities < 2 || p->n_entities > E2SEC_MAX_ENTITIES) return fail(ctx, E2SEC_ERR_ARG, "n_entities out of range");
    if (max_frames < 1 || n_points < 1) return fail(ctx, E2SEC_ERR_ARG, "max_frames / n_points must be positive");
    if (p->n_replace < 0 || p->n_replace > 16) return fail(ctx, E2SEC_ERR_ARG, "n_replace out of range");
    int hand_idx = -1, ground_idx = -1;
    for (int e = 0; e < p->n_entities; ++e) {
        if (p->entity_label[e] == p->hand_label) hand_idx = e;
        if (p->entity_label[e] == p->ground_label) ground_idx = e;
    }
    // the reference raises IndexError at :1520 / :1406-1426 when either label is missing
    if (hand_idx < 0) return fail(ctx, E2SEC_ERR_ARG, "hand_label is not in total_unique_labels");
    if (ground_idx < 0) return fail(ctx, E2SEC_ERR_ARG, "ground_label is not in total_unique_labels");
    if (p->n_ground < 0 || (p->n_ground > 0 && !p->ground_xyz)) return fail(ctx, E2SEC_ERR_ARG, "ground cloud missing");
    // The hand as the lowest tracked label sits at index 0 of total_unique_labels, which the reference never filters or
    // relates EXCEPT as the hand (:1456-1475, :1518-1520, :1549).  The device keeps "index 0 is never stored" by working on
    // one phantom entity in front (Dev::ent_shift); records come out in the caller's K-entity layout.
    const int shift = hand_idx == 0 ? 1 : 0;
    const int Kd = p->n_entities + shift;
    if (Kd > E2SEC_MAX_ENTITIES) return fail(ctx, E2SEC_ERR_ARG, "n_entities out of range (the hand at index 0 takes one more slot)");
    int rc = allocate(ctx, max_frames, n_points, Kd);
    if (rc) return rc;
    ctx->params = *p;
    ctx->ground.assign(p->ground_xyz, p->ground_xyz + (size_t)3 * p->n_ground);
    ctx->params.ground_xyz = ctx->ground.data();
    // label LUT: sequential replacement (:176-198) composed with the lookup in total_unique_labels (:1456-1472)
    std::vector<int> lut(E2SEC_LABEL_LUT);
    for (int raw = 0; raw < E2SEC_LABEL_LUT; ++raw) {
        int v = raw;
        for (int i = 0; i < p->n_replace; ++i)
            if (v == p->replace_old[i]) v = p->replace_new[i];
        int idx = -1;
        for (int e = 0; e < p->n_entities; ++e)
            if (p->entity_label[e] == v) { idx = e + shift; break; }
        lut[raw] = idx;
    }
    CU(cudaMemcpyAsync(ctx->lut.p, lut.data(), lut.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    // ground cloud as SoA + its AABB.  The cloud only ever enters minimum distances, so its order is free: the points
    // are sorted along a Morton curve and every run of E2_CT_SUB points gets a tight box -- contact_work_kernel skips the
    // runs that lie farther from a work item than its current minimum without loading them.
    const int G = p->n_ground;
    std::vector<double> soa((size_t)3 * (G ? G : 1));
    double box[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < G; ++i)
        for (int a = 0; a < 3; ++a) {
            double v = p->ground_xyz[3 * (size_t)i + a];
            if (i == 0 || v < box[a]) box[a] = v;
            if (i == 0 || v > box[3 + a]) box[3 + a] = v;
        }
    std::vector<int> order(G);
    {
        std::vector<unsigned long long> code(G);
        auto spread = [](unsigned long long v) {            // 21 bits -> every third bit
            v &= 0x1fffffull;
            v = (v | v << 32) & 0x1f00000000ffffull;
            v = (v | v << 16) & 0x1f0000ff0000ffull;
            v = (v | v << 8) & 0x100f00f00f00f00full;
            v = (v | v << 4) & 0x10c30c30c30c30c3ull;
            v = (v | v << 2) & 0x1249249249249249ull;
            return v;
        };
        double ext = 0.0;
        for (int a = 0; a < 3; ++a) ext = std::max(ext, box[3 + a] - box[a]);
        const double q = (ext > 0.0 && std::isfinite(ext)) ? 2097151.0 / ext : 0.0;
        for (int i = 0; i < G; ++i) {
            unsigned long long c = 0;
            for (int a = 0; a < 3; ++a) {
                double t = (p->ground_xyz[3 * (size_t)i + a] - box[a]) * q;
                if (!(t >= 0.0)) t = 0.0;
                if (t > 2097151.0) t = 2097151.0;
                c |= spread((unsigned long long)t) << a;
            }
            code[i] = c;
            order[i] = i;
        }
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return code[x] < code[y]; });
    }
    const int n_sub = (G + E2_CT_SUB - 1) / E2_CT_SUB;
    std::vector<double> sub((size_t)6 * (n_sub ? n_sub : 1), 0.0);
    for (int i = 0; i < G; ++i)
        for (int a = 0; a < 3; ++a) {
            const double v = p->ground_xyz[3 * (size_t)order[i] + a];
            soa[(size_t)a * G + i] = v;
            double *b = &sub[(size_t)6 * (i / E2_CT_SUB)];
            if (i % E2_CT_SUB == 0 || v < b[a]) b[a] = v;
            if (i % E2_CT_SUB == 0 || v > b[3 + a]) b[3 + a] = v;
        }
    if (ensure(ctx, ctx->gsub, sizeof(double) * 6 * (n_sub ? n_sub : 1))) return E2SEC_ERR_CUDA;
    CU(cudaMemcpyAsync(ctx->gsub.p, sub.data(), sizeof(double) * 6 * (n_sub ? n_sub : 1), cudaMemcpyHostToDevice, ctx->stream));
    if (ensure(ctx, ctx->gx, sizeof(double) * (G ? G : 1)) || ensure(ctx, ctx->gy, sizeof(double) * (G ? G : 1)) ||
        ensure(ctx, ctx->gz, sizeof(double) * (G ? G : 1)))
        return E2SEC_ERR_CUDA;
    if (G) {
        CU(cudaMemcpyAsync(ctx->gx.p, soa.data(), sizeof(double) * G, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->gy.p, soa.data() + G, sizeof(double) * G, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->gz.p, soa.data() + 2 * (size_t)G, sizeof(double) * G, cudaMemcpyHostToDevice, ctx->stream));
    }
    CU(cudaStreamSynchronize(ctx->stream));
    Dev &d = ctx->d;
    d.K = Kd;
    d.N = n_points;
    d.hand_idx = hand_idx + shift;
    d.ground_idx = ground_idx + shift;
    d.ent_shift = shift;
    d.hand_label = p->hand_label;
    d.n_ground = G;
    d.filter_mode = p->filter_mode;
    d.filter_seed = p->filter_seed;
    if (p->radius_mode != E2SEC_RADIUS_EXACT && p->radius_mode != E2SEC_RADIUS_BOUNDED) return fail(ctx, E2SEC_ERR_ARG, "bad radius_mode");
    d.radius_mode = (p->filter_mode == E2SEC_FILTER_CLOSURE) ? p->radius_mode : E2SEC_RADIUS_EXACT;
    d.thresh = p->thresh;
    memcpy(d.T, p->translation, sizeof(d.T));
    memcpy(d.R, p->roll, sizeof(d.R));
    d.affine = 1;
    for (int i = 12; i < 16; ++i)
        if (d.T[i] != (i == 15 ? 1.0 : 0.0) || d.R[i] != (i == 15 ? 1.0 : 0.0)) d.affine = 0;
    d.gsub = (const double *)ctx->gsub.p;
    d.gx = (const double *)ctx->gx.p;
    d.gy = (const double *)ctx->gy.p;
    d.gz = (const double *)ctx->gz.p;
    memcpy(d.gaabb, box, sizeof(box));
    d.rec_doubles = (int)e2sec_record_doubles(p->n_entities);
    ctx->configured = true;
    return e2sec_reset_sequence(ctx);
}

// cv2.resize(..., INTER_NEAREST) source index tables (:1433)
static int upload_maps(e2sec_ctx *ctx, int lw, int lh, int width, int height)
{
    if (ctx->map_cfg[0] == lw && ctx->map_cfg[1] == lh && ctx->map_cfg[2] == width && ctx->map_cfg[3] == height) return 0;
    std::vector<int> rows(height), cols(width);
    const double ifx = 1.0 / ((double)width / (double)lw), ify = 1.0 / ((double)height / (double)lh);
    for (int x = 0; x < width; ++x) { int s = (int)std::floor(x * ifx); cols[x] = s < lw - 1 ? s : lw - 1; }
    for (int y = 0; y < height; ++y) { int s = (int)std::floor(y * ify); rows[y] = s < lh - 1 ? s : lh - 1; }
    if (ensure(ctx, ctx->rowmap, height * sizeof(int)) || ensure(ctx, ctx->colmap, width * sizeof(int))) return E2SEC_ERR_CUDA;
    CU(cudaMemcpyAsync(ctx->rowmap.p, rows.data(), height * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->colmap.p, cols.data(), width * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->map_cfg[0] = lw; ctx->map_cfg[1] = lh; ctx->map_cfg[2] = width; ctx->map_cfg[3] = height;
    return 0;
}

#define LAUNCHED() do { ++ctx->launches; } while (0)

// Kernels of the per-batch pipeline are launched with programmatic stream serialization: every one of them starts with
// griddepcontrol.wait (common.cuh: pdl_enter), so the next grid's launch latency and prologue hide behind the tail of the
// previous one instead of adding a gap per launch (30 launches per batch).  E2SEC_PDL=0 falls back to plain ordering.
template <class... P, class... A>
static inline void launch_pdl(e2sec_ctx *ctx, void (*k)(P...), dim3 g, dim3 b, size_t smem, cudaStream_t st, A... a)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = g;
    cfg.blockDim = b;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = (ctx->pdl && !ctx->pdl_once_off) ? 1 : 0;
    ctx->pdl_once_off = false;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, k, P(a)...);
    if (e != cudaSuccess && ctx->launch_err == cudaSuccess) ctx->launch_err = e;     // the first failing launch is the one reported
    ++ctx->launches;
}

// The same without the attribute, for the persistent kernels whose whole grid is one resident wave (pair_pass_kernel,
// contact_work_kernel): launched early, their CTAs take their places while the kernel before them still holds part of the SMs,
// and the pass then runs 1.45x slower than when it starts on an empty machine (measured: 36 vs 25 ms for the two passes of a
// 200-frame batch of the bundled recording; 0.8 % of the step for contact_work_kernel).  A few microseconds of launch gap
// are cheap against that.
template <class... P, class... A>
static inline void launch_plain(e2sec_ctx *ctx, void (*k)(P...), dim3 g, dim3 b, size_t smem, cudaStream_t st, A... a)
{
    ctx->pdl_once_off = true;
    launch_pdl(ctx, k, g, b, smem, st, a...);
}

// every launch is checked; the pipeline functions report the first failure once they have queued their launches
static int launch_status(e2sec_ctx *ctx, const char *where)
{
    if (ctx->launch_err == cudaSuccess) return E2SEC_OK;
    const cudaError_t e = ctx->launch_err;
    ctx->launch_err = cudaSuccess;
    cudaGetLastError();
    return fail(ctx, E2SEC_ERR_CUDA, std::string(where) + ": kernel launch failed: " + cudaGetErrorString(e));
}

// Everything after segmentation: medians, radius, filter, stats, pairs, records. `d` already describes the batch.
static int run_entity_stages(e2sec_ctx *ctx, const Dev &d, bool need_median, bool need_radius, bool do_pairs, bool do_records)
{
    cudaStream_t st = ctx->stream;
    const int FK = d.F * d.K;
    PairPlan plan{(unsigned *)ctx->plan_off.p, (unsigned *)ctx->plan_total.p, (unsigned *)ctx->plan_split.p};
    u64 *below_acc = (u64 *)ctx->below_acc.p, *min_above = (u64 *)ctx->min_above.p, *single = (u64 *)ctx->single_count.p;
    // BOUNDED mode selects the medians inside cell_filter_kernel
    if (need_median && d.radius_mode != E2SEC_RADIUS_BOUNDED) { launch_pdl(ctx, axis_median_kernel, dim3(dim3(3, FK)), dim3(E2_SEL_THREADS), 0, st, d); }
    cudaEventRecord(ctx->ev[2], st);
    if (d.radius_mode == E2SEC_RADIUS_BOUNDED) {
        launch_pdl(ctx, cell_filter_kernel<false>, dim3(FK), dim3(E2_CELL_THREADS), sizeof(CellSmem), st, d);
    }
    if (need_radius) {
        // The exact O(n^2) selection for the entities of the exact path's list (BOUNDED: the few the cell table left open --
        // small fixed grids that stride over the list; otherwise every entity).  The host never waits inside a batch, so the
        // rounds are launched unconditionally; a round with nothing left returns at once.
        const bool bounded = d.radius_mode == E2SEC_RADIUS_BOUNDED;
        const int grid = ctx->sm_count * E2_PAIR_CTAS;
        const int small_grid = bounded ? std::min(FK, ctx->sm_count * 2) : FK;
        const size_t smem = E2_PAIR_SMEM_HIST;
        launch_pdl(ctx, radius_init_kernel, dim3(small_grid), dim3(256), 0, st, d, plan, below_acc, min_above, single, grid);
        for (int round = 0; round < E2_MAX_REFINE; ++round) {
            launch_plain(ctx, pair_pass_kernel<false>, dim3(grid), dim3(E2_PAIR_THREADS), smem, st, d, plan, below_acc, min_above);
            const bool last = round == E2_MAX_REFINE - 1;         // the plan that follows the last round is the collect pass's
            launch_pdl(ctx, radius_select_kernel, dim3(bounded ? std::min((FK + 7) / 8, ctx->sm_count) : (FK + 7) / 8), dim3(256), 0, st, d, plan, below_acc, single,
                       last ? 1 : 0, last ? ctx->sm_count * 8 : grid);
        }
        launch_plain(ctx, pair_pass_kernel<true>, dim3(ctx->sm_count * 8), dim3(E2_PAIR_THREADS), E2_PAIR_SMEM_COLLECT, st, d, plan, below_acc, min_above);
        launch_pdl(ctx, radius_final_kernel, dim3(small_grid), dim3(E2_SEL_THREADS), 0, st, d, min_above, single);
        // entities that only had a point inside the radius band: the cell closure is finished at the exact radius
        if (bounded)
            launch_pdl(ctx, cell_filter_kernel<true>, dim3(small_grid), dim3(E2_CELL_THREADS), sizeof(CellSmem), st, d);
    }
    cudaEventRecord(ctx->ev[3], st);
    launch_pdl(ctx, region_filter_kernel, dim3(FK), dim3(E2_FILTER_THREADS), 0, st, d);
    cudaEventRecord(ctx->ev[4], st);
    launch_pdl(ctx, entity_stats_kernel, dim3(FK), dim3(E2_STATS_THREADS), 0, st, d);
    cudaEventRecord(ctx->ev[5], st);
    if (do_pairs) {
        launch_pdl(ctx, prev_kept_kernel, dim3(1), dim3(1024), 0, st, d);
        ContactPlan cp{(unsigned *)ctx->ct_cnt.p, (unsigned *)ctx->ct_off.p, (unsigned *)ctx->ct_next.p,
                       d.K * (d.K - 1) / 2 + d.K + d.K * d.K, (unsigned *)ctx->ct_map.p,
                       (unsigned)(ctx->ct_map.bytes / sizeof(unsigned))};
        const int n_idx = d.F * cp.slots;
        launch_pdl(ctx, contact_plan_kernel, dim3((n_idx + 255) / 256), dim3(256), 0, st, d, cp);
        launch_pdl(ctx, contact_scan_kernel, dim3(1), dim3(1024), 0, st, cp, n_idx);
        launch_plain(ctx, contact_work_kernel, dim3(ctx->sm_count * 4), dim3(E2_CT_THREADS), 0, st, d, cp);
        launch_pdl(ctx, halo_save_kernel, dim3(64), dim3(256), 0, st, d);
        launch_pdl(ctx, halo_commit_kernel, dim3(1), dim3(1), 0, st, d);
    }
    cudaEventRecord(ctx->ev[6], st);
    if (do_records) { launch_pdl(ctx, assemble_kernel, dim3(d.F), dim3(128), 0, st, d); }
    cudaEventRecord(ctx->ev[7], st);
    if (int rc = launch_status(ctx, "entity stages")) return rc;
    CU(cudaGetLastError());
    return E2SEC_OK;
}

static int check_device_error(e2sec_ctx *ctx)
{
    int err = 0;
    CU(cudaMemcpyAsync(&err, ctx->error.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemsetAsync(ctx->error.p, 0, sizeof(int), ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (err == E2SEC_ERR_CAPACITY) return fail(ctx, err, "stored points exceed the frame capacity");
    if (err == E2SEC_ERR_STATE) return fail(ctx, err, "radius selection lost its rank (internal error)");
    if (err == E2SEC_ERR_IO) return fail(ctx, err, "a compressed cloud is malformed, or a label file does not hold label_w x label_h non-negative integers");
    if (err) return fail(ctx, err, "device-side error");
    return E2SEC_OK;
}

// uint8 label images (e2sec_process_batch_labels8): widened to int32 on the device, 4 labels per thread
__global__ void labels_widen_kernel(const uchar4 *src, int4 *dst, size_t n4, const unsigned char *src1, int *dst1, size_t n)
{
    pdl_enter();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n4) { const uchar4 v = src[i]; dst[i] = make_int4(v.x, v.y, v.z, v.w); }
    if (i < n - 4 * n4) dst1[4 * n4 + i] = src1[4 * n4 + i];
}

static int process_batch_impl(e2sec_ctx *ctx, const float *xyz, const void *labels, int label_bytes, int32_t n_frames,
                              int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                              const int32_t *frame_ids, int on_device, double *records, bool wait)
{
    if (!ctx) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (!ctx->configured) return fail(ctx, E2SEC_ERR_STATE, "e2sec_configure has not been called");
    if (!xyz || !labels || !frame_ids) return fail(ctx, E2SEC_ERR_ARG, "null input");
    if (n_frames < 0 || n_frames > ctx->maxF) return fail(ctx, E2SEC_ERR_CAPACITY, "n_frames exceeds max_frames");
    if (n_frames == 0) return E2SEC_OK;
    const int N = ctx->N;
    const long long lsz = (long long)label_w * label_h;
    const bool upscale = (lsz != N);
    if (lsz <= 0 || lsz > N) return fail(ctx, E2SEC_ERR_ARG, "label image must be non-empty and not larger than the cloud");
    if (upscale) {
        if ((long long)width * height != N || label_w > width || label_h > height)
            return fail(ctx, E2SEC_ERR_ARG, "label upscaling needs width*height == n_points and label dims <= cloud dims");
        if (upload_maps(ctx, label_w, label_h, width, height)) return E2SEC_ERR_CUDA;
    }
    cudaStream_t st = ctx->stream;
    Dev d = ctx->d;
    d.F = n_frames;
    d.cap = N;
    d.S = (N + E2_SEG - 1) / E2_SEG;
    d.label_size = (int)lsz;
    d.rowmap = upscale ? (const int *)ctx->rowmap.p : nullptr;
    d.colmap = upscale ? (const int *)ctx->colmap.p : nullptr;
    d.width = upscale ? width : N;
    d.label_w = label_w;
    int stage_slot = -1;
    cudaEventRecord(ctx->ev[0], st);
    const size_t n_lab = (size_t)n_frames * (size_t)lsz;
    if (label_bytes == 1 && ensure(ctx, ctx->lab8_stage, (size_t)2 * ctx->maxF * N)) return E2SEC_ERR_CUDA;
    if (label_bytes == -1) {
        // e2sec_process_batch_pcd: the FILES travel (compressed clouds, label text); the copy stream uploads them and decodes
        // them into the staging slot while the main stream still works on the previous batch
        const size_t slack = 4096;                  // the decoder's input ring reads whole aligned 1 KB blocks around a stream
        const size_t fb = (ctx->pcd.files_bytes + 15) / 16 * 16 + 2 * slack, lb = (ctx->pcd.ltext_bytes + 15) / 16 * 16 + 64;
        if (ctx->pcd_stage.bytes < 2 * fb && ensure(ctx, ctx->pcd_stage, 2 * fb + fb / 2)) return E2SEC_ERR_CUDA;
        if (ctx->ltext_stage.bytes < 2 * lb && ensure(ctx, ctx->ltext_stage, 2 * lb + lb / 2)) return E2SEC_ERR_CUDA;
        if (ensure(ctx, ctx->pcd_frames, (size_t)2 * ctx->maxF * sizeof(LzfFrame)) || ensure(ctx, ctx->pcd_status, (size_t)4 * ctx->maxF * sizeof(int)) ||
            ensure(ctx, ctx->ltext_off, (size_t)2 * (ctx->maxF + 1) * sizeof(long long)))
            return E2SEC_ERR_CUDA;
        stage_slot = (int)(ctx->stage_idx++ & 1u);
        float *xs = (float *)ctx->xyz_stage.p + (size_t)stage_slot * ctx->maxF * N * 3;
        int *ls = (int *)ctx->lab_stage.p + (size_t)stage_slot * ctx->maxF * N;
        unsigned char *fs = (unsigned char *)ctx->pcd_stage.p + (size_t)stage_slot * (ctx->pcd_stage.bytes / 2 / 16 * 16) + slack;
        unsigned char *ts = (unsigned char *)ctx->ltext_stage.p + (size_t)stage_slot * (ctx->ltext_stage.bytes / 2 / 16 * 16);
        LzfFrame *fj = (LzfFrame *)ctx->pcd_frames.p + (size_t)stage_slot * ctx->maxF;
        int *stt = (int *)ctx->pcd_status.p + (size_t)stage_slot * 2 * ctx->maxF;      // [clouds][label images]
        long long *lo = (long long *)ctx->ltext_off.p + (size_t)stage_slot * (ctx->maxF + 1);
        cudaStream_t cs = ctx->copy_stream;
        if (ctx->slot_used[stage_slot]) CU(cudaStreamWaitEvent(cs, ctx->slot_free[stage_slot], 0));
        CU(cudaMemcpyAsync(fs, ctx->pcd.files, ctx->pcd.files_bytes, cudaMemcpyHostToDevice, cs));
        CU(cudaMemcpyAsync(ts, ctx->pcd.ltext, ctx->pcd.ltext_bytes, cudaMemcpyHostToDevice, cs));
        CU(cudaMemcpyAsync(fj, ctx->pcd.frames.data(), n_frames * sizeof(LzfFrame), cudaMemcpyHostToDevice, cs));
        CU(cudaMemcpyAsync(lo, ctx->pcd.loff.data(), (n_frames + 1) * sizeof(long long), cudaMemcpyHostToDevice, cs));
        lzf_decode_kernel<<<n_frames, 32, 0, cs>>>(fs, fj, xs, N, stt, (int *)ctx->error.p); LAUNCHED();
        label_text_kernel<<<n_frames, E2_LT_THREADS, 0, cs>>>(ts, lo, ls, (int)lsz, stt + ctx->maxF, (int *)ctx->error.p); LAUNCHED();
        CU(cudaGetLastError());
        CU(cudaEventRecord(ctx->copy_done[stage_slot], cs));
        CU(cudaStreamWaitEvent(st, ctx->copy_done[stage_slot], 0));
        d.xyz = xs;
        d.labels = ls;
    } else if (on_device && label_bytes == 4) {
        d.xyz = xyz;
        d.labels = (const int *)labels;
    } else {
        // slot s is refilled on the copy stream as soon as the kernels that read it two batches ago are done, so the
        // upload of this batch overlaps the kernels of the previous one when batches are queued asynchronously
        stage_slot = (int)(ctx->stage_idx++ & 1u);
        float *xs = (float *)ctx->xyz_stage.p + (size_t)stage_slot * ctx->maxF * N * 3;
        int *ls = (int *)ctx->lab_stage.p + (size_t)stage_slot * ctx->maxF * N;
        unsigned char *l8 = label_bytes == 1 ? (unsigned char *)ctx->lab8_stage.p + (size_t)stage_slot * ctx->maxF * N : nullptr;
        const unsigned char *src8 = (const unsigned char *)labels;
        d.xyz = xyz;
        if (!on_device) {
            if (ctx->slot_used[stage_slot]) CU(cudaStreamWaitEvent(ctx->copy_stream, ctx->slot_free[stage_slot], 0));
            CU(cudaMemcpyAsync(xs, xyz, (size_t)n_frames * N * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->copy_stream));
            if (label_bytes == 1) CU(cudaMemcpyAsync(l8, labels, n_lab, cudaMemcpyHostToDevice, ctx->copy_stream));
            else CU(cudaMemcpyAsync(ls, labels, n_lab * sizeof(int), cudaMemcpyHostToDevice, ctx->copy_stream));
            CU(cudaEventRecord(ctx->copy_done[stage_slot], ctx->copy_stream));
            CU(cudaStreamWaitEvent(st, ctx->copy_done[stage_slot], 0));
            d.xyz = xs;
            src8 = l8;
        }
        if (label_bytes == 1) {
            const size_t n4 = ((uintptr_t)src8 % 4 == 0 && (uintptr_t)ls % 16 == 0) ? n_lab / 4 : 0, work = n4 > n_lab - 4 * n4 ? n4 : n_lab - 4 * n4;
            launch_pdl(ctx, labels_widen_kernel, dim3((unsigned)((work + 255) / 256)), dim3(256), 0, st, (const uchar4 *)src8,
                       (int4 *)ls, n4, src8, ls, n_lab);
        }
        d.labels = ls;
    }
    CU(cudaMemcpyAsync(ctx->frame_ids.p, frame_ids, n_frames * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(ctx->flags.p, 0, n_frames * sizeof(int), st));      // the device error flag is sticky until checked
    // single-pass cluster kernel when the bulk copies can be issued (16-byte aligned frames), else the two-pass kernels
    const bool fused = !ctx->force_two_pass && (N % 4 == 0) && ((uintptr_t)d.xyz % 16 == 0) && ((uintptr_t)d.labels % 16 == 0) &&
                       (upscale || d.label_size % 4 == 0) && N <= (1 << 26);
    d.ent_box = nullptr;
    if (fused && !ctx->seg_cluster) {
        // unit stream: persistent warps pull 2048-point units, then a per-frame scan and one warp per unit to emit
        d.ent_box = (unsigned *)ctx->ent_box.p;
        Dev du = d;
        du.S = (N + E2_SS_UNIT - 1) / E2_SS_UNIT;
        unsigned *next = (unsigned *)ctx->seg_next.p;
        const long long units = (long long)n_frames * du.S;
        const int ctas = (int)std::min<long long>((units + E2_SS_WARPS - 1) / E2_SS_WARPS, (long long)ctx->sm_count * E2_SS_CTAS);
        cudaEventRecord(ctx->ev[8], st);
        if (upscale) launch_pdl(ctx, seg_stream_kernel<true>, dim3(ctas), dim3(E2_SS_THREADS), sizeof(SegStreamSmem), st, du, next);
        else launch_pdl(ctx, seg_stream_kernel<false>, dim3(ctas), dim3(E2_SS_THREADS), sizeof(SegStreamSmem), st, du, next);
        cudaEventRecord(ctx->ev[9], st);
        ctx->ev_stream = true;
        launch_pdl(ctx, seg_prefix_kernel, dim3(n_frames), dim3(256), 0, st, du);
        launch_pdl(ctx, seg_emit_kernel, dim3((du.S + E2_SS_WARPS - 1) / E2_SS_WARPS, n_frames), dim3(E2_SS_THREADS), 0, st, du, next);
    } else if (fused) {
        const int tpf = (N + E2_SF_TILE - 1) / E2_SF_TILE;
        int C = ctx->sf_cluster;
        while (C > 1 && C > tpf) C >>= 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)n_frames * C);
        cfg.blockDim = dim3(E2_SF_THREADS);
        cfg.dynamicSmemBytes = sizeof(SegFusedSmem);
        cfg.stream = st;
        cudaLaunchAttribute attr[2];
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[1].val.programmaticStreamSerializationAllowed = 1;
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = C;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = ctx->pdl ? 2 : 1;
        if (upscale) CU(cudaLaunchKernelEx(&cfg, seg_fused_kernel<true>, d));
        else CU(cudaLaunchKernelEx(&cfg, seg_fused_kernel<false>, d));
        LAUNCHED();
    } else {
        const dim3 sgrid((d.S + E2_SEG_WARPS - 1) / E2_SEG_WARPS, n_frames);
        launch_pdl(ctx, seg_count_kernel, dim3(sgrid), dim3(32 * E2_SEG_WARPS), 0, st, d);
        launch_pdl(ctx, seg_scan_kernel, dim3(n_frames), dim3(E2SEC_MAX_ENTITIES), 0, st, d);
        launch_pdl(ctx, seg_scatter_kernel, dim3(sgrid), dim3(32 * E2_SEG_WARPS), 0, st, d);
    }
    if (int lrc = launch_status(ctx, "segmentation")) return lrc;
    if (stage_slot >= 0) { CU(cudaEventRecord(ctx->slot_free[stage_slot], st)); ctx->slot_used[stage_slot] = true; }   // inputs consumed
    cudaEventRecord(ctx->ev[1], st);
    int rc = run_entity_stages(ctx, d, true, true, true, true);
    if (rc) return rc;
    if (records)
        CU(cudaMemcpyAsync(records, ctx->records.p, (size_t)n_frames * d.rec_doubles * sizeof(double), cudaMemcpyDefault, st));
    return wait ? check_device_error(ctx) : E2SEC_OK;
}

extern "C" int e2sec_process_batch(e2sec_ctx *ctx, const float *xyz, const int32_t *labels, int32_t n_frames,
                                   int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                   const int32_t *frame_ids, int on_device, double *records)
{
    return process_batch_impl(ctx, xyz, labels, 4, n_frames, label_w, label_h, width, height, frame_ids, on_device, records, true);
}

extern "C" int e2sec_process_batch_async(e2sec_ctx *ctx, const float *xyz, const int32_t *labels, int32_t n_frames,
                                         int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                         const int32_t *frame_ids, int on_device, double *records)
{
    return process_batch_impl(ctx, xyz, labels, 4, n_frames, label_w, label_h, width, height, frame_ids, on_device, records, false);
}

extern "C" int e2sec_process_batch_labels8(e2sec_ctx *ctx, const float *xyz, const uint8_t *labels, int32_t n_frames,
                                           int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                           const int32_t *frame_ids, int on_device, double *records)
{
    return process_batch_impl(ctx, xyz, labels, 1, n_frames, label_w, label_h, width, height, frame_ids, on_device, records, true);
}

extern "C" int e2sec_process_batch_labels8_async(e2sec_ctx *ctx, const float *xyz, const uint8_t *labels, int32_t n_frames,
                                                 int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                                 const int32_t *frame_ids, int on_device, double *records)
{
    return process_batch_impl(ctx, xyz, labels, 1, n_frames, label_w, label_h, width, height, frame_ids, on_device, records, false);
}

// e2sec_process_batch_pcd: headers are read here (a few hundred bytes per file), everything else happens on the device
static int process_batch_pcd(e2sec_ctx *ctx, const uint8_t *pcd_files, const int64_t *pcd_offsets, const uint8_t *label_files,
                             const int64_t *label_offsets, int32_t n_frames, int32_t label_w, int32_t label_h, int32_t width,
                             int32_t height, const int32_t *frame_ids, double *records, bool wait)
{
    if (!ctx) return E2SEC_ERR_ARG;
    if (!ctx->configured) return fail(ctx, E2SEC_ERR_STATE, "e2sec_configure has not been called");
    if (!pcd_files || !pcd_offsets || !label_files || !label_offsets || !frame_ids) return fail(ctx, E2SEC_ERR_ARG, "null input");
    if (n_frames < 0 || n_frames > ctx->maxF) return fail(ctx, E2SEC_ERR_CAPACITY, "n_frames exceeds max_frames");
    if (n_frames == 0) return E2SEC_OK;
    if ((long long)width * height != ctx->N) return fail(ctx, E2SEC_ERR_ARG, "width * height must equal n_points of e2sec_configure");
    ctx->pcd.frames.resize(n_frames);
    ctx->pcd.loff.assign(label_offsets, label_offsets + n_frames + 1);
    for (int f = 0; f < n_frames; ++f) {
        const int64_t a = pcd_offsets[f], b = pcd_offsets[f + 1];
        if (a < pcd_offsets[0] || b < a) return fail(ctx, E2SEC_ERR_ARG, "pcd_offsets must be non-decreasing");
        long long payload = 0, points = 0;
        unsigned comp = 0;
        int w = 0, h = 0;
        const int rc = e2sec_pcd_locate(pcd_files + a, (size_t)(b - a), &payload, &comp, &points, &w, &h);
        if (rc == E2SEC_ERR_ARG)
            return fail(ctx, rc, "cloud " + std::to_string(f) + " is not a binary_compressed PCD with float32 x, y, z as its first fields "
                                 "(decode it with e2sec_read_pcd_xyz and use e2sec_process_batch)");
        if (rc) return fail(ctx, rc, "cloud " + std::to_string(f) + ": malformed PCD header");
        if (points != ctx->N) return fail(ctx, E2SEC_ERR_ARG, "cloud " + std::to_string(f) + " holds " + std::to_string(points) + " points, the context was configured for " + std::to_string(ctx->N));
        ctx->pcd.frames[f].payload = (long long)(a - pcd_offsets[0]) + payload;
        ctx->pcd.frames[f].comp = comp;
        ctx->pcd.frames[f].need = (unsigned)(12ll * points);
        if (label_offsets[f + 1] < label_offsets[f]) return fail(ctx, E2SEC_ERR_ARG, "label_offsets must be non-decreasing");
    }
    for (auto &v : ctx->pcd.loff) v -= label_offsets[0];
    ctx->pcd.files = pcd_files + pcd_offsets[0];
    ctx->pcd.files_bytes = (size_t)(pcd_offsets[n_frames] - pcd_offsets[0]);
    ctx->pcd.ltext = label_files + label_offsets[0];
    ctx->pcd.ltext_bytes = (size_t)(label_offsets[n_frames] - label_offsets[0]);
    return process_batch_impl(ctx, (const float *)pcd_files, label_files, -1, n_frames, label_w, label_h, width, height, frame_ids, 0,
                              records, wait);
}

extern "C" int e2sec_process_batch_pcd(e2sec_ctx *ctx, const uint8_t *pcd_files, const int64_t *pcd_offsets, const uint8_t *label_files,
                                       const int64_t *label_offsets, int32_t n_frames, int32_t label_w, int32_t label_h,
                                       int32_t width, int32_t height, const int32_t *frame_ids, double *records)
{
    return process_batch_pcd(ctx, pcd_files, pcd_offsets, label_files, label_offsets, n_frames, label_w, label_h, width, height,
                             frame_ids, records, true);
}

extern "C" int e2sec_process_batch_pcd_async(e2sec_ctx *ctx, const uint8_t *pcd_files, const int64_t *pcd_offsets,
                                             const uint8_t *label_files, const int64_t *label_offsets, int32_t n_frames,
                                             int32_t label_w, int32_t label_h, int32_t width, int32_t height,
                                             const int32_t *frame_ids, double *records)
{
    return process_batch_pcd(ctx, pcd_files, pcd_offsets, label_files, label_offsets, n_frames, label_w, label_h, width, height,
                             frame_ids, records, false);
}

// the staged inputs of the batch queued last (device memory): what the decoders produced, for callers that want to look
extern "C" int e2sec_staged_inputs_dev(e2sec_ctx *ctx, const float **xyz_dev, const int32_t **labels_dev)
{
    if (!ctx || !ctx->configured) return E2SEC_ERR_ARG;
    const int slot = (int)((ctx->stage_idx - 1u) & 1u);
    if (xyz_dev) *xyz_dev = (const float *)ctx->xyz_stage.p + (size_t)slot * ctx->maxF * ctx->N * 3;
    if (labels_dev) *labels_dev = (const int32_t *)ctx->lab_stage.p + (size_t)slot * ctx->maxF * ctx->N;
    return E2SEC_OK;
}

extern "C" int e2sec_copy_staged_inputs(e2sec_ctx *ctx, float *xyz, int32_t *labels, int32_t n_frames, int32_t label_size)
{
    if (!ctx || !ctx->configured || n_frames < 0 || n_frames > ctx->maxF || label_size < 0 || label_size > ctx->N) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    const float *x;
    const int32_t *l;
    e2sec_staged_inputs_dev(ctx, &x, &l);
    CU(cudaStreamSynchronize(ctx->stream));
    if (xyz) CU(cudaMemcpy(xyz, x, (size_t)n_frames * ctx->N * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    if (labels) CU(cudaMemcpy(labels, l, (size_t)n_frames * label_size * sizeof(int32_t), cudaMemcpyDeviceToHost));
    return E2SEC_OK;
}

// The clouds behind a record (for callers that track rigid models, icp=True :1677-1738): entity `entity` of frame `frame` of
// the batch processed last.  which = 0: every valid point of the label, transformed (pcd[i], :1516); 1: the filtered cloud
// (filtered_pcd_voxel[i], :1519-1546: the seed point first -- it counts twice, pyx:48 -- then the kept points in raster order).
extern "C" int e2sec_entity_cloud(e2sec_ctx *ctx, int32_t frame, int32_t entity, int32_t which, double *out, int32_t capacity, int32_t *n_out)
{
    if (!ctx || !ctx->configured || !n_out || frame < 0 || frame >= ctx->maxF || (which != 0 && which != 1)) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    const Dev &d = ctx->d;
    const int K = ctx->d.K, e = entity + d.ent_shift;
    if (entity < 0 || e >= K) return fail(ctx, E2SEC_ERR_ARG, "e2sec_entity_cloud: entity out of range");
    CU(cudaStreamSynchronize(ctx->stream));
    int off = -1, raw = 0, nl = 0, seed = -1;
    const size_t fe = (size_t)frame * K + e;
    CU(cudaMemcpy(&off, d.ent_off + fe, sizeof(int), cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(&raw, d.ent_raw + fe, sizeof(int), cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(&nl, d.n_list + fe, sizeof(int), cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(&seed, d.seed_pos + fe, sizeof(int), cudaMemcpyDeviceToHost));
    if (off < 0) { *n_out = 0; return E2SEC_OK; }                 // index 0 / the ground label: never stored
    const int n = which == 0 ? raw : nl + (seed >= 0 ? 1 : 0);
    *n_out = n;
    if (!out || n == 0) return E2SEC_OK;
    if (n > capacity) return fail(ctx, E2SEC_ERR_CAPACITY, "e2sec_entity_cloud: capacity too small");
    const size_t pb = (size_t)frame * ctx->N + off;
    std::vector<double> x(raw), y(raw), z(raw);
    CU(cudaMemcpy(x.data(), d.px + pb, sizeof(double) * raw, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(y.data(), d.py + pb, sizeof(double) * raw, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(z.data(), d.pz + pb, sizeof(double) * raw, cudaMemcpyDeviceToHost));
    if (which == 0) {
        for (int i = 0; i < raw; ++i) { out[3 * (size_t)i] = x[i]; out[3 * (size_t)i + 1] = y[i]; out[3 * (size_t)i + 2] = z[i]; }
        return E2SEC_OK;
    }
    std::vector<unsigned> fl(nl > 0 ? nl : 1);
    if (nl) CU(cudaMemcpy(fl.data(), d.flist + pb, sizeof(unsigned) * nl, cudaMemcpyDeviceToHost));
    size_t o = 0;
    if (seed >= 0) { out[0] = x[seed]; out[1] = y[seed]; out[2] = z[seed]; o = 1; }
    for (int i = 0; i < nl; ++i, ++o) { const unsigned q = fl[i]; out[3 * o] = x[q]; out[3 * o + 1] = y[q]; out[3 * o + 2] = z[q]; }
    return E2SEC_OK;
}

extern "C" int e2sec_copy_records(e2sec_ctx *ctx, double *dst, int32_t first, int32_t count)
{
    if (!ctx || !dst || first < 0 || count < 0) return E2SEC_ERR_ARG;
    if (!ctx->configured || first + count > ctx->maxF) return fail(ctx, E2SEC_ERR_ARG, "e2sec_copy_records: range outside the batch");
    cudaSetDevice(ctx->device);
    const size_t w = (size_t)ctx->d.rec_doubles * sizeof(double);
    if (count) CU(cudaMemcpyAsync(dst, (const char *)ctx->records.p + (size_t)first * w, (size_t)count * w, cudaMemcpyDefault, ctx->stream));
    return E2SEC_OK;
}

extern "C" int e2sec_stage_ms(e2sec_ctx *ctx, float out[8])
{
    if (!ctx || !out) return E2SEC_ERR_ARG;
    CU(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < 8; ++i) out[i] = 0.f;
    for (int i = 0; i < 7; ++i)
        if (cudaEventElapsedTime(&out[i], ctx->ev[i], ctx->ev[i + 1]) != cudaSuccess) out[i] = -1.f;
    // out[7]: the streaming segmentation kernel alone (seg_stream_kernel), the HBM-bound kernel of the path
    if (ctx->ev_stream && cudaEventElapsedTime(&out[7], ctx->ev[8], ctx->ev[9]) != cudaSuccess) out[7] = -1.f;
    cudaGetLastError();
    return E2SEC_OK;
}

// ---- operators -----------------------------------------------------------------------------------
// A scratch "batch" of one frame holding `ne` clouds as entities 1..ne (entity 0 stays empty).
static int load_clouds(e2sec_ctx *ctx, Dev &d, int ne, const double *const *pts, const int *ns)
{
    size_t total = 0;
    for (int i = 0; i < ne; ++i) total += (size_t)ns[i];
    const int K = ne + 1;
    if (total > 0x3fffffff) return fail(ctx, E2SEC_ERR_CAPACITY, "cloud too large");
    if (!ctx->cap_total || total > ctx->cap_total || K > ctx->K) {
        // grow the scratch without disturbing a configured sequence's shape
        const int keepF = ctx->maxF, keepN = ctx->N, keepK = ctx->K;
        size_t want = total > 65536 ? total : 65536;
        int rc = allocate(ctx, 1, (int)want, K > keepK ? K : keepK);
        if (rc) return rc;
        if (ctx->configured) {
            rc = allocate(ctx, keepF, keepN, K > keepK ? K : keepK);
            if (rc) return rc;
            ctx->K = keepK;
        }
    }
    if (ctx->halo_info.p) CU(cudaMemsetAsync(ctx->halo_info.p, 0, 4 * sizeof(int), ctx->stream));   // the scratch overwrites the halo
    d = ctx->d;
    d.radius_mode = E2SEC_RADIUS_EXACT;       // operators work at the reference's own seams: exact values
    d.F = 1;
    d.K = K;
    d.cap = (int)ctx->cap_total;
    d.N = d.cap;
    d.S = 0;
    d.hand_idx = 1;
    d.ent_shift = 0;
    d.ground_idx = -1;
    d.n_ground = 0;
    d.rec_doubles = (int)e2sec_record_doubles(K);
    std::vector<int> raw(K, 0), off(K, -1);
    std::vector<double> soa;
    size_t o = 0;
    for (int i = 0; i < ne; ++i) {
        const int n = ns[i];
        raw[i + 1] = n;
        off[i + 1] = (int)o;
        soa.resize((size_t)3 * n);
        for (int j = 0; j < n; ++j)
            for (int a = 0; a < 3; ++a) soa[(size_t)a * n + j] = pts[i][3 * (size_t)j + a];
        if (n) {
            CU(cudaMemcpyAsync(d.px + o, soa.data(), sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaMemcpyAsync(d.py + o, soa.data() + n, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaMemcpyAsync(d.pz + o, soa.data() + 2 * (size_t)n, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
        }
        o += n;
    }
    int zero = 0;
    CU(cudaMemcpyAsync(d.ent_raw, raw.data(), K * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d.ent_off, off.data(), K * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync((void *)d.frame_ids, &zero, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemsetAsync(d.error, 0, sizeof(int), ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return E2SEC_OK;
}

extern "C" double e2sec_distance(const double p1[3], const double p2[3])
{
    const double dx = p1[0] - p2[0], dy = p1[1] - p2[1], dz = p1[2] - p2[2];
    return std::sqrt((dx * dx + dy * dy) + dz * dz);
}

extern "C" int e2sec_axis_median(e2sec_ctx *ctx, const double *pts, int32_t n, double median[3])
{
    if (!ctx || !pts || !median || n < 1) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    Dev d;
    const double *pp[1] = {pts};
    int nn[1] = {n};
    // the kernel skips clouds of <= 3 points (the reference never filters them): pad the call by repetition
    std::vector<double> rep;
    if (n <= 3) {
        // np.median of 1..3 values on the host (trivial, not a data path)
        for (int a = 0; a < 3; ++a) {
            double v[3];
            for (int i = 0; i < n; ++i) v[i] = pts[3 * i + a];
            for (int i = 0; i < n; ++i)
                for (int j = i + 1; j < n; ++j)
                    if (v[j] < v[i]) { double t = v[i]; v[i] = v[j]; v[j] = t; }
            median[a] = (n & 1) ? v[n / 2] : (v[n / 2 - 1] + v[n / 2]) / 2.0;
        }
        return E2SEC_OK;
    }
    int rc = load_clouds(ctx, d, 1, pp, nn);
    if (rc) return rc;
    axis_median_kernel<<<dim3(3, d.K), E2_SEL_THREADS, 0, ctx->stream>>>(d); LAUNCHED();
    CU(cudaMemcpyAsync(median, d.med + 3, 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return E2SEC_OK;
}

static int run_radius_only(e2sec_ctx *ctx, const Dev &d)
{
    cudaStream_t st = ctx->stream;
    const int FK = d.F * d.K;
    PairPlan plan{(unsigned *)ctx->plan_off.p, (unsigned *)ctx->plan_total.p, (unsigned *)ctx->plan_split.p};
    u64 *below_acc = (u64 *)ctx->below_acc.p, *min_above = (u64 *)ctx->min_above.p, *single = (u64 *)ctx->single_count.p;
    const int grid = ctx->sm_count * E2_PAIR_CTAS;
    radius_init_kernel<<<FK, 256, 0, st>>>(d, plan, below_acc, min_above, single, grid); LAUNCHED();      // + the plan of the first pass
    const size_t smem = E2_PAIR_SMEM_HIST;
    for (int round = 0; round < 12; ++round) {     // an operator call may take as many rounds as it needs
        unsigned total = 0;
        CU(cudaMemcpyAsync(&total, plan.total, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (!total) break;
        pair_pass_kernel<false><<<grid, E2_PAIR_THREADS, smem, st>>>(d, plan, below_acc, min_above); LAUNCHED();
        radius_select_kernel<<<(FK + 7) / 8, 256, 0, st>>>(d, plan, below_acc, single, 0, grid); LAUNCHED();
    }
    radius_plan_kernel<<<1, 1024, 0, st>>>(d, plan, 1, ctx->sm_count * 8); LAUNCHED();
    pair_pass_kernel<true><<<ctx->sm_count * 8, E2_PAIR_THREADS, E2_PAIR_SMEM_COLLECT, st>>>(d, plan, below_acc, min_above); LAUNCHED();
    radius_final_kernel<<<FK, E2_SEL_THREADS, 0, st>>>(d, min_above, single); LAUNCHED();
    CU(cudaGetLastError());
    return E2SEC_OK;
}

extern "C" int e2sec_pairwise_median(e2sec_ctx *ctx, const double *pts, int32_t n, double *median)
{
    if (!ctx || !pts || !median || n < 2) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (n <= 3) {       // 1 or 3 pairs: host arithmetic in pdist's expression order
        double v[3];
        int m = 0;
        for (int i = 0; i < n; ++i)
            for (int j = i + 1; j < n; ++j) v[m++] = e2sec_distance(pts + 3 * i, pts + 3 * j);
        for (int i = 0; i < m; ++i)
            for (int j = i + 1; j < m; ++j)
                if (v[j] < v[i]) { double t = v[i]; v[i] = v[j]; v[j] = t; }
        *median = (m & 1) ? v[m / 2] : (v[m / 2 - 1] + v[m / 2]) / 2.0;
        return E2SEC_OK;
    }
    Dev d;
    const double *pp[1] = {pts};
    int nn[1] = {n};
    int rc = load_clouds(ctx, d, 1, pp, nn);
    if (rc) return rc;
    rc = run_radius_only(ctx, d);
    if (rc) return rc;
    CU(cudaMemcpyAsync(median, d.radius + 1, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    return check_device_error(ctx);
}

extern "C" int e2sec_region_filter(e2sec_ctx *ctx, const double *pcd, int32_t n, const double center[3],
                                   double median_distance, int32_t mode, uint32_t seed, double *out, int32_t *n_out)
{
    if (!ctx || !pcd || !center || !out || !n_out || n < 1) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (n == 2) {       // pyx:35-42: a two-point cloud returns np.zeros((1, 3))
        out[0] = out[1] = out[2] = 0.0;
        *n_out = 1;
        return E2SEC_OK;
    }
    Dev d;
    const double *pp[1] = {pcd};
    // the kernel's n <= 3 shortcut belongs to _process (:1520), not to the operator: pad tiny clouds by running them
    // through the same code with the guard satisfied is impossible, so clouds of 1 or 3 points go through as 4+ only
    if (n <= 3) return fail(ctx, E2SEC_ERR_ARG, "region_filter operator needs at least 4 points (the reference path never calls it with fewer, :1520)");
    int nn[1] = {n};
    int rc = load_clouds(ctx, d, 1, pp, nn);
    if (rc) return rc;
    d.filter_mode = mode;
    d.filter_seed = seed;       // frame id 0, entity 1 -> seed + 7919; undo so that `seed` is the seed used
    d.filter_seed = seed - 7919u;
    CU(cudaMemcpyAsync(d.med + 3, center, 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d.radius + 1, &median_distance, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    region_filter_kernel<<<d.K, E2_FILTER_THREADS, 0, ctx->stream>>>(d); LAUNCHED();
    int m = 0, sp = -1;
    CU(cudaMemcpyAsync(&m, d.n_list + 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(&sp, d.seed_pos + 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    std::vector<unsigned> fl(m > 0 ? m : 1);
    if (m) CU(cudaMemcpyAsync(fl.data(), d.flist, sizeof(unsigned) * m, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < 3; ++a) out[a] = pcd[3 * (size_t)sp + a];      // dicti[0] = first_point
    for (int i = 0; i < m; ++i)
        for (int a = 0; a < 3; ++a) out[3 * (size_t)(i + 1) + a] = pcd[3 * (size_t)fl[i] + a];
    *n_out = m + 1;
    return check_device_error(ctx);
}

extern "C" int e2sec_min_distance(e2sec_ctx *ctx, const double *a, int32_t na, const double *b, int32_t nb,
                                  double limit, double *dist)
{
    if (!ctx || !a || !b || !dist || na < 0 || nb < 0) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (na == 0 || nb == 0) { *dist = INFINITY; return E2SEC_OK; }
    Dev d;
    const double *pp[2] = {a, b};
    int nn[2] = {na, nb};
    int rc = load_clouds(ctx, d, 2, pp, nn);
    if (rc) return rc;
    d.thresh = (limit > 0.0) ? limit : 1e150;
    d.filter_mode = E2SEC_FILTER_CLOSURE;
    // both clouds pass unfiltered: flist = identity
    std::vector<unsigned> id((size_t)(na > nb ? na : nb));
    for (size_t i = 0; i < id.size(); ++i) id[i] = (unsigned)i;
    int nl[3] = {0, na, nb}, sp[3] = {-1, -1, -1};
    CU(cudaMemcpyAsync(d.flist, id.data(), sizeof(unsigned) * na, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d.flist + na, id.data(), sizeof(unsigned) * nb, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d.n_list, nl, sizeof(nl), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d.seed_pos, sp, sizeof(sp), cudaMemcpyHostToDevice, ctx->stream));
    entity_stats_kernel<<<d.K, E2_STATS_THREADS, 0, ctx->stream>>>(d); LAUNCHED();
    pair_distance_kernel<<<dim3(d.K * (d.K - 1) / 2 + d.K, 1), E2_PD_THREADS, 0, ctx->stream>>>(d); LAUNCHED();   // in-frame slots only
    u64 key = 0;
    CU(cudaMemcpyAsync(&key, d.pair_key + (size_t)1 * (d.K + 1) + 2, sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    double d2;
    memcpy(&d2, &key, sizeof(double));
    *dist = std::sqrt(d2);
    return check_device_error(ctx);
}

// ---- eSEC_analyser.getDissimilarityMatrix (:782-841) / similarity_manipulations (:843-864) -----------------------------
extern "C" int e2sec_dissimilarity(e2sec_ctx *ctx, const char *codes, const int64_t *offsets, const int32_t *n_cols,
                                   int32_t n_tables, int32_t rows, double divisor, double *D, uint32_t *counts)
{
    if (!ctx) return E2SEC_ERR_ARG;
    if (!codes || !offsets || !n_cols || !D) return fail(ctx, E2SEC_ERR_ARG, "e2sec_dissimilarity: NULL argument");
    if (n_tables < 1 || rows < 3 || rows % 3 || !(divisor > 0.0))
        return fail(ctx, E2SEC_ERR_ARG, "e2sec_dissimilarity: need n_tables >= 1, rows a positive multiple of 3, divisor > 0");
    const int M = n_tables, opal = rows / 3;
    int Lmax = 0;
    int64_t end = 0;
    for (int m = 0; m < M; ++m) {
        // a table without columns has no last column to repeat (the reference raises KeyError at :932 / :940)
        if (n_cols[m] < 1) return fail(ctx, E2SEC_ERR_ARG, "e2sec_dissimilarity: table " + std::to_string(m) + " has no columns");
        if (offsets[m] < 0) return fail(ctx, E2SEC_ERR_ARG, "e2sec_dissimilarity: negative offset");
        Lmax = std::max(Lmax, (int)n_cols[m]);
        end = std::max<int64_t>(end, offsets[m] + (int64_t)rows * n_cols[m] * E2_DS_CODE);
    }
    const size_t E = (size_t)Lmax * opal;
    const size_t smem = 3 * E * sizeof(u64);
    if (smem > 200 * 1024) return fail(ctx, E2SEC_ERR_CAPACITY, "e2sec_dissimilarity: a table has more than 200 KB of codes");
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    int rc = 0;
    rc |= ensure(ctx, ctx->ds_codes, (size_t)end);
    rc |= ensure(ctx, ctx->ds_off, sizeof(long long) * M);
    rc |= ensure(ctx, ctx->ds_ncols, sizeof(int) * M);
    rc |= ensure(ctx, ctx->ds_packed, sizeof(u64) * 3 * E * M);
    rc |= ensure(ctx, ctx->ds_counts, sizeof(unsigned) * 3 * M * M);
    rc |= ensure(ctx, ctx->ds_D, sizeof(double) * M * M);
    if (rc) return E2SEC_ERR_CUDA;
    static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
    CU(cudaMemcpyAsync(ctx->ds_codes.p, codes, (size_t)end, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->ds_off.p, offsets, sizeof(long long) * M, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->ds_ncols.p, n_cols, sizeof(int) * M, cudaMemcpyHostToDevice, st));
    DissimDev d;
    d.codes = (const unsigned char *)ctx->ds_codes.p;
    d.off = (const long long *)ctx->ds_off.p;
    d.ncols = (const int *)ctx->ds_ncols.p;
    d.M = M; d.rows = rows; d.opal = opal; d.Lmax = Lmax; d.E = (int)E;
    d.packed = (u64 *)ctx->ds_packed.p;
    d.counts = (unsigned *)ctx->ds_counts.p;
    d.D = (double *)ctx->ds_D.p;
    d.divisor = divisor;
    CU(cudaFuncSetAttribute(dissim_pairs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CU(cudaEventRecord(ctx->ds_ev[0], st));
    dissim_pack_kernel<<<M, E2_DS_THREADS, 0, st>>>(d); LAUNCHED();
    dissim_pairs_kernel<<<dim3((M + E2_DS_JTILE - 1) / E2_DS_JTILE, M), E2_DS_THREADS, smem, st>>>(d); LAUNCHED();
    CU(cudaEventRecord(ctx->ds_ev[1], st));
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(D, d.D, sizeof(double) * M * M, cudaMemcpyDeviceToHost, st));
    if (counts) CU(cudaMemcpyAsync(counts, d.counts, sizeof(unsigned) * 3 * M * M, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    ctx->ds_timed = true;
    return E2SEC_OK;
}

extern "C" int e2sec_dissimilarity_ms(e2sec_ctx *ctx, float *ms)
{
    if (!ctx || !ms) return E2SEC_ERR_ARG;
    if (!ctx->ds_timed) return fail(ctx, E2SEC_ERR_STATE, "e2sec_dissimilarity_ms: no call to time yet");
    CU(cudaEventElapsedTime(ms, ctx->ds_ev[0], ctx->ds_ev[1]));
    return E2SEC_OK;
}

// ---- phase 2 on the device (phase2.cuh) ------------------------------------------------------------------------------------
static_assert(sizeof(Phase2State) <= E2SEC_PHASE2_STATE_BYTES, "Phase2State outgrew its ABI size");
static_assert(E2_P2_COL == E2SEC_PHASE2_COLUMN_BYTES, "packed column width");

struct e2sec_phase2 {
    e2sec_ctx *ctx = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ready = nullptr, done = nullptr;
    Phase2Dev a;
    DevBuf state, cols, col_frame, n_cols, mailbox;
    void *peer_mailbox = nullptr;               // the next rank's mailbox, opened from its IPC handle ...
    bool peer_is_ipc = false;                   // ... or an object of this very process (e2sec_phase2_connect_local)
};

static void phase2_initial_state(Phase2State &s)
{
    memset(&s, 0, sizeof(s));
    for (int k = 0; k < 3; ++k) s.sel.oindex[k] = -1;
    s.prev_frame_id = -1;
}

extern "C" int e2sec_phase2_create(e2sec_ctx *ctx, int32_t max_columns, e2sec_phase2 **out)
{
    if (!ctx || !out) return E2SEC_ERR_ARG;
    *out = nullptr;
    if (!ctx->configured) return fail(ctx, E2SEC_ERR_STATE, "e2sec_phase2_create: e2sec_configure has not been called");
    if (max_columns < 1) return fail(ctx, E2SEC_ERR_ARG, "e2sec_phase2_create: max_columns must be positive");
    cudaSetDevice(ctx->device);
    const e2sec_params &p = ctx->params;
    e2sec_phase2 *q = new e2sec_phase2();
    q->ctx = ctx;
    Phase2Dev &a = q->a;
    memset(&a, 0, sizeof(a));
    a.K = p.n_entities;
    a.hand_idx = a.ground_idx = a.support_idx = -1;
    for (int e = 0; e < a.K; ++e) {
        if (p.entity_label[e] == p.hand_label) a.hand_idx = e;
        if (p.entity_label[e] == p.ground_label) a.ground_idx = e;
        if (p.support_hand >= 0 && p.entity_label[e] == p.support_hand) a.support_idx = e;
    }
    if (p.support_hand >= 0 && a.support_idx < 0) { delete q; return fail(ctx, E2SEC_ERR_ARG, "support_hand is not in total_unique_labels"); }
    a.relations = p.relations ? p.relations : 3;
    a.thresh = p.thresh;
    a.rec_doubles = (int)e2sec_record_doubles(a.K);
    a.ground = rel_ent_undefined();
    a.ground.defined = 1;
    a.ground.n = p.n_ground;
    for (int i = 0; i < p.n_ground; ++i)
        for (int k = 0; k < 3; ++k) {
            const double v = ctx->ground[3 * (size_t)i + k];
            if (i == 0 || v < a.ground.lo[k]) a.ground.lo[k] = v;
            if (i == 0 || v > a.ground.hi[k]) a.ground.hi[k] = v;
        }
    for (int k = 0; k < 3; ++k) a.ground.center[k] = (a.ground.lo[k] + a.ground.hi[k]) * 0.5;
    a.max_columns = max_columns;
    bool ok = cudaStreamCreateWithFlags(&q->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&q->ready, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&q->done, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && !ensure(ctx, q->state, sizeof(Phase2State)) && !ensure(ctx, q->cols, (size_t)max_columns * E2_P2_COL) &&
         !ensure(ctx, q->col_frame, (size_t)max_columns * sizeof(int)) && !ensure(ctx, q->n_cols, sizeof(int)) &&
         !ensure(ctx, q->mailbox, sizeof(Phase2Mailbox)) && cudaMemset(q->mailbox.p, 0, sizeof(Phase2Mailbox)) == cudaSuccess;
    if (!ok) { e2sec_phase2_destroy(q); return fail(ctx, E2SEC_ERR_CUDA, "e2sec_phase2_create: cannot allocate"); }
    a.state = (Phase2State *)q->state.p;
    a.cols = (unsigned char *)q->cols.p;
    a.col_frame = (int *)q->col_frame.p;
    a.n_cols = (int *)q->n_cols.p;
    a.inbox = (Phase2Mailbox *)q->mailbox.p;
    a.timeout_ns = 20000000000ull;
    *out = q;
    return e2sec_phase2_reset(q);
}

extern "C" void e2sec_phase2_destroy(e2sec_phase2 *q)
{
    if (!q) return;
    cudaSetDevice(q->ctx->device);
    if (q->stream) cudaStreamSynchronize(q->stream);
    if (q->peer_mailbox && q->peer_is_ipc) cudaIpcCloseMemHandle(q->peer_mailbox);
    for (DevBuf *b : {&q->state, &q->cols, &q->col_frame, &q->n_cols, &q->mailbox})
        if (b->p) cudaFree(b->p);
    if (q->ready) cudaEventDestroy(q->ready);
    if (q->done) cudaEventDestroy(q->done);
    if (q->stream) cudaStreamDestroy(q->stream);
    delete q;
}

extern "C" int e2sec_phase2_reset(e2sec_phase2 *q)
{
    if (!q) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    Phase2State s;
    phase2_initial_state(s);
    CU(cudaMemcpyAsync(q->state.p, &s, sizeof(s), cudaMemcpyHostToDevice, q->stream));
    CU(cudaMemsetAsync(q->n_cols.p, 0, sizeof(int), q->stream));
    CU(cudaStreamSynchronize(q->stream));       // `s` lives on this stack frame
    return E2SEC_OK;
}

static_assert(sizeof(cudaIpcMemHandle_t) <= E2SEC_IPC_HANDLE_BYTES, "IPC handle size");

extern "C" int e2sec_phase2_mailbox_handle(e2sec_phase2 *q, void *handle)
{
    if (!q || !handle) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, q->mailbox.p));
    memset(handle, 0, E2SEC_IPC_HANDLE_BYTES);
    memcpy(handle, &h, sizeof(h));
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_connect(e2sec_phase2 *q, const void *next_handle)
{
    if (!q || !next_handle) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    if (q->peer_mailbox && q->peer_is_ipc) cudaIpcCloseMemHandle(q->peer_mailbox);
    q->peer_mailbox = nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, next_handle, sizeof(h));
    CU(cudaIpcOpenMemHandle(&q->peer_mailbox, h, cudaIpcMemLazyEnablePeerAccess));
    q->peer_is_ipc = true;
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_connect_local(e2sec_phase2 *q, e2sec_phase2 *next)
{
    if (!q || !next) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    if (q->peer_mailbox && q->peer_is_ipc) cudaIpcCloseMemHandle(q->peer_mailbox);
    q->peer_mailbox = nullptr;
    if (next->ctx->device != ctx->device) {
        int can = 0;
        CU(cudaDeviceCanAccessPeer(&can, ctx->device, next->ctx->device));
        if (!can) return fail(ctx, E2SEC_ERR_CUDA, "e2sec_phase2_connect_local: no peer access between the two devices");
        const cudaError_t e = cudaDeviceEnablePeerAccess(next->ctx->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU(e);
        cudaGetLastError();
    }
    q->peer_mailbox = next->mailbox.p;
    q->peer_is_ipc = false;
    return E2SEC_OK;
}

static int phase2_launch(e2sec_phase2 *q, e2sec_ctx *src, const double *records_dev, int32_t n_frames, uint32_t wait_ticket,
                         uint32_t post_ticket);

extern "C" int e2sec_phase2_run(e2sec_phase2 *q, e2sec_ctx *src, const double *records_dev, int32_t n_frames)
{
    return phase2_launch(q, src, records_dev, n_frames, 0u, 0u);
}

extern "C" int e2sec_phase2_run_ring(e2sec_phase2 *q, e2sec_ctx *src, const double *records_dev, int32_t n_frames,
                                     uint32_t wait_ticket, uint32_t post_ticket)
{
    if (!q) return E2SEC_ERR_ARG;
    if (post_ticket && !q->peer_mailbox) return fail(q->ctx, E2SEC_ERR_STATE, "e2sec_phase2_run_ring: e2sec_phase2_connect has not been called");
    return phase2_launch(q, src, records_dev, n_frames, wait_ticket, post_ticket);
}

static int phase2_launch(e2sec_phase2 *q, e2sec_ctx *src, const double *records_dev, int32_t n_frames, uint32_t wait_ticket,
                         uint32_t post_ticket)
{
    if (!q || !records_dev || n_frames < 0) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    if (n_frames == 0 && !wait_ticket && !post_ticket) return E2SEC_OK;
    if (src) {                                   // the records are produced on src's stream
        CU(cudaEventRecord(q->ready, src->stream));
        CU(cudaStreamWaitEvent(q->stream, q->ready, 0));
    }
    Phase2Dev a = q->a;
    a.records = records_dev;
    a.n_frames = n_frames;
    a.wait_ticket = wait_ticket;
    a.post_ticket = post_ticket;
    a.outbox = (Phase2Mailbox *)q->peer_mailbox;
    phase2_kernel<<<1, E2_P2_THREADS, 0, q->stream>>>(a);
    ++ctx->launches;
    CU(cudaGetLastError());
    if (src) {                                   // and may be overwritten by src's next batch only once they have been read
        CU(cudaEventRecord(q->done, q->stream));
        CU(cudaStreamWaitEvent(src->stream, q->done, 0));
    }
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_state_get(e2sec_phase2 *q, void *dst)
{
    if (!q || !dst) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    CU(cudaMemcpyAsync(dst, q->state.p, sizeof(Phase2State), cudaMemcpyDefault, q->stream));
    CU(cudaStreamSynchronize(q->stream));
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_state_set(e2sec_phase2 *q, const void *src)
{
    if (!q || !src) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    CU(cudaMemcpyAsync(q->state.p, src, sizeof(Phase2State), cudaMemcpyDefault, q->stream));
    CU(cudaStreamSynchronize(q->stream));
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_sync(e2sec_phase2 *q, int32_t *n_columns)
{
    if (!q) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    Phase2State s;
    int n = 0;
    CU(cudaMemcpyAsync(&s, q->state.p, sizeof(s), cudaMemcpyDeviceToHost, q->stream));
    CU(cudaMemcpyAsync(&n, q->n_cols.p, sizeof(int), cudaMemcpyDeviceToHost, q->stream));
    CU(cudaStreamSynchronize(q->stream));
    if (n_columns) *n_columns = n;
    if (s.error == E2SEC_ERR_CAPACITY) return fail(ctx, E2SEC_ERR_CAPACITY, "phase 2: more columns than max_columns");
    if (s.error) return fail(ctx, E2SEC_ERR_STATE, "phase 2: a record was measured against another previous kept frame than the one "
                                                   "phase 2 saw last (feed frames in order; a shard starts with its halo frame)");
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_fetch(e2sec_phase2 *q, uint8_t *columns, int32_t *frames, int32_t capacity, int32_t *n_columns)
{
    if (!q || !n_columns) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    int n = 0;
    int rc = e2sec_phase2_sync(q, &n);
    if (rc) return rc;
    *n_columns = n;
    if (n > capacity) return fail(ctx, E2SEC_ERR_CAPACITY, "e2sec_phase2_fetch: capacity too small");
    if (n && columns) CU(cudaMemcpyAsync(columns, q->cols.p, (size_t)n * E2_P2_COL, cudaMemcpyDeviceToHost, q->stream));
    if (n && frames) CU(cudaMemcpyAsync(frames, q->col_frame.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, q->stream));
    CU(cudaStreamSynchronize(q->stream));
    return E2SEC_OK;
}

// everything a gather between GPUs needs, in ONE fixed-size device (or host) buffer: [int32 n][12 bytes unused]
// [capacity x E2SEC_PHASE2_COLUMN_BYTES codes][capacity x int32 frame ids]; waits for the object's stream
extern "C" int e2sec_phase2_export(e2sec_phase2 *q, void *dst, int32_t capacity)
{
    if (!q || !dst || capacity < 1) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    const int cap = capacity < q->a.max_columns ? capacity : q->a.max_columns;
    char *o = (char *)dst;
    CU(cudaMemcpyAsync(o, q->n_cols.p, sizeof(int), cudaMemcpyDefault, q->stream));
    CU(cudaMemcpyAsync(o + 16, q->cols.p, (size_t)cap * E2_P2_COL, cudaMemcpyDefault, q->stream));
    CU(cudaMemcpyAsync(o + 16 + (size_t)capacity * E2_P2_COL, q->col_frame.p, (size_t)cap * sizeof(int), cudaMemcpyDefault, q->stream));
    CU(cudaStreamSynchronize(q->stream));
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_columns_dev(e2sec_phase2 *q, const uint8_t **columns_dev, const int32_t **frames_dev, const int32_t **n_dev)
{
    if (!q) return E2SEC_ERR_ARG;
    if (columns_dev) *columns_dev = (const uint8_t *)q->cols.p;
    if (frames_dev) *frames_dev = (const int32_t *)q->col_frame.p;
    if (n_dev) *n_dev = (const int32_t *)q->n_cols.p;
    return E2SEC_OK;
}

extern "C" int e2sec_phase2_bindings(e2sec_phase2 *q, int32_t labels[3], int32_t absent[3])
{
    if (!q) return E2SEC_ERR_ARG;
    e2sec_ctx *ctx = q->ctx;
    cudaSetDevice(ctx->device);
    Phase2State s;
    CU(cudaMemcpyAsync(&s, q->state.p, sizeof(s), cudaMemcpyDeviceToHost, q->stream));
    CU(cudaStreamSynchronize(q->stream));
    for (int k = 0; k < 3; ++k) {
        if (labels) labels[k] = s.sel.count[k] ? s.sel.oindex[k] : -1;
        if (absent) absent[k] = s.absent[k] ? 1 : 0;
    }
    return E2SEC_OK;
}

extern "C" void *e2sec_phase2_stream(e2sec_phase2 *q) { return q ? (void *)q->stream : nullptr; }

#ifdef E2_CELL_PROFILE
// debugging aid (not in e2sec.h): clock ticks per phase of cell_filter_kernel summed over CTAs; reset != 0 clears them
extern "C" int e2sec_debug_cell_profile(unsigned long long out[16], int reset)
{
    cudaDeviceSynchronize();
    if (out && cudaMemcpyFromSymbol(out, g_cell_prof, sizeof(unsigned long long) * 16) != cudaSuccess) return E2SEC_ERR_CUDA;
    if (reset) { unsigned long long z[16] = {0}; if (cudaMemcpyToSymbol(g_cell_prof, z, sizeof(z)) != cudaSuccess) return E2SEC_ERR_CUDA; }
    return E2SEC_OK;
}
#endif

#ifdef E2_LZ_PROFILE
// debugging aid (not in e2sec.h): clock ticks per stage of lzf_decode_kernel summed over warps; reset != 0 clears them
extern "C" int e2sec_debug_lzf_profile(unsigned long long out[16], int reset)
{
    cudaDeviceSynchronize();
    if (out && cudaMemcpyFromSymbol(out, g_lz_prof, sizeof(unsigned long long) * 16) != cudaSuccess) return E2SEC_ERR_CUDA;
    if (reset) { unsigned long long z[16] = {0}; if (cudaMemcpyToSymbol(g_lz_prof, z, sizeof(z)) != cudaSuccess) return E2SEC_ERR_CUDA; }
    return E2SEC_OK;
}
#endif

// ---- frame-0 setup geometry (frame0.cuh) ---------------------------------------------------------------------------------
// small scratch device buffers of the operators below
struct F0Buf {
    void *p = nullptr;
    ~F0Buf() { if (p) cudaFree(p); }
    int get(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16) == cudaSuccess ? 0 : 1; }
};

extern "C" int e2sec_plane_inlier_counts(e2sec_ctx *ctx, const double *pts, int32_t n, const double *planes, int32_t m,
                                         double threshold, int32_t *counts)
{
    if (!ctx || !pts || !planes || !counts || n < 1 || m < 1 || m > E2_F0_MAXPLANES) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    F0Buf dp, dpl, dc;
    if (dp.get(sizeof(double) * 3 * n) || dpl.get(sizeof(double) * 4 * m) || dc.get(sizeof(int) * m)) return fail(ctx, E2SEC_ERR_CUDA, "cudaMalloc failed");
    CU(cudaMemcpyAsync(dp.p, pts, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(dpl.p, planes, sizeof(double) * 4 * m, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(dc.p, 0, sizeof(int) * m, st));
    const int grid = std::min((n + E2_F0_THREADS - 1) / E2_F0_THREADS, ctx->sm_count * 8);
    plane_count_kernel<<<grid, E2_F0_THREADS, 0, st>>>((const double *)dp.p, n, (const double *)dpl.p, m, threshold, (int *)dc.p); LAUNCHED();
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(counts, dc.p, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return E2SEC_OK;
}

extern "C" int e2sec_plane_inlier_mask(e2sec_ctx *ctx, const double *pts, int32_t n, const double plane[4], double threshold, uint8_t *mask)
{
    if (!ctx || !pts || !plane || !mask || n < 1) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    F0Buf dp, dm;
    if (dp.get(sizeof(double) * 3 * n) || dm.get(n)) return fail(ctx, E2SEC_ERR_CUDA, "cudaMalloc failed");
    CU(cudaMemcpyAsync(dp.p, pts, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    plane_mask_kernel<<<(n + 255) / 256, 256, 0, st>>>((const double *)dp.p, n, plane[0], plane[1], plane[2], plane[3], threshold, (unsigned char *)dm.p); LAUNCHED();
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(mask, dm.p, n, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return E2SEC_OK;
}

extern "C" int e2sec_voxel_down_sample(e2sec_ctx *ctx, const double *pts, int32_t n, double voxel, double *out, int32_t *n_out)
{
    if (!ctx || !pts || !out || !n_out || n < 1 || !(voxel > 0.0)) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    double lo[3] = {pts[0], pts[1], pts[2]}, hi[3] = {pts[0], pts[1], pts[2]};
    for (int i = 1; i < n; ++i)
        for (int a = 0; a < 3; ++a) { const double v = pts[3 * (size_t)i + a]; if (v < lo[a]) lo[a] = v; if (v > hi[a]) hi[a] = v; }
    double org[3];
    long long dim[3];
    for (int a = 0; a < 3; ++a) {
        if (!std::isfinite(lo[a]) || !std::isfinite(hi[a])) return fail(ctx, E2SEC_ERR_ARG, "e2sec_voxel_down_sample: non-finite coordinates");
        org[a] = lo[a] - voxel * 0.5;                         // Open3D: voxel_min_bound = min_bound - voxel_size * 0.5
        dim[a] = (long long)std::floor((hi[a] - org[a]) / voxel) + 2;
    }
    if ((double)dim[0] * (double)dim[1] * (double)dim[2] > 9e18) return fail(ctx, E2SEC_ERR_CAPACITY, "e2sec_voxel_down_sample: voxel grid too large");
    F0Buf dp, dk, di, ds, dout;
    if (dp.get(sizeof(double) * 3 * n) || dk.get(sizeof(unsigned long long) * n) || di.get(sizeof(int) * n) || ds.get(sizeof(int) * (n + 1)) ||
        dout.get(sizeof(double) * 3 * n))
        return fail(ctx, E2SEC_ERR_CUDA, "cudaMalloc failed");
    CU(cudaMemcpyAsync(dp.p, pts, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    const int grid = (n + 255) / 256;
    voxel_key_kernel<<<grid, 256, 0, st>>>((const double *)dp.p, n, org[0], org[1], org[2], voxel, dim[1], dim[2], (unsigned long long *)dk.p, (int *)di.p); LAUNCHED();
    thrust::device_ptr<unsigned long long> k((unsigned long long *)dk.p);
    thrust::device_ptr<int> ix((int *)di.p), start((int *)ds.p);
    thrust::stable_sort_by_key(thrust::cuda::par.on(st), k, k + n, ix);
    voxel_start_kernel<<<grid, 256, 0, st>>>((const unsigned long long *)dk.p, n, (int *)ds.p); LAUNCHED();
    int last_flag = 0, last_excl = 0;
    CU(cudaMemcpyAsync(&last_flag, (int *)ds.p + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
    thrust::exclusive_scan(thrust::cuda::par.on(st), start, start + n, start);
    CU(cudaMemcpyAsync(&last_excl, (int *)ds.p + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
    voxel_mean_kernel<<<grid, 256, 0, st>>>((const double *)dp.p, n, (const unsigned long long *)dk.p, (const int *)di.p, (double *)dout.p, (int *)ds.p); LAUNCHED();
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    const int m = last_excl + last_flag;
    CU(cudaMemcpy(out, dout.p, sizeof(double) * 3 * m, cudaMemcpyDeviceToHost));
    *n_out = m;
    return E2SEC_OK;
}

extern "C" int e2sec_icp_correspondences(e2sec_ctx *ctx, const double *source, int32_t n_source, const double *target, int32_t n_target,
                                         const double transform[16], double max_distance, double stats[17])
{
    if (!ctx || !source || !target || !transform || !stats || n_source < 1 || n_target < 1 || !(max_distance > 0.0)) return E2SEC_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    F0Buf dsrc, dtgt, dnn, dd2, dmv, dst;
    if (dsrc.get(sizeof(double) * 3 * n_source) || dtgt.get(sizeof(double) * 3 * n_target) || dnn.get(sizeof(int) * n_source) ||
        dd2.get(sizeof(double) * n_source) || dmv.get(sizeof(double) * 3 * n_source) || dst.get(sizeof(double) * 17))
        return fail(ctx, E2SEC_ERR_CUDA, "cudaMalloc failed");
    CU(cudaMemcpyAsync(dsrc.p, source, sizeof(double) * 3 * n_source, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(dtgt.p, target, sizeof(double) * 3 * n_target, cudaMemcpyHostToDevice, st));
    IcpParams p;
    memcpy(p.T, transform, sizeof(p.T));
    p.max_d2 = max_distance * max_distance;
    const long long threads = (long long)n_source * 32;
    icp_nn_kernel<<<(unsigned)((threads + E2_F0_THREADS - 1) / E2_F0_THREADS), E2_F0_THREADS, 0, st>>>((const double *)dsrc.p, n_source, (const double *)dtgt.p, n_target, p,
                                                                                                     (int *)dnn.p, (double *)dd2.p, (double *)dmv.p); LAUNCHED();
    icp_accumulate_kernel<<<1, 1024, 0, st>>>((const double *)dmv.p, (const double *)dtgt.p, (const int *)dnn.p, (const double *)dd2.p, n_source, (double *)dst.p); LAUNCHED();
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(stats, dst.p, sizeof(double) * 17, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return E2SEC_OK;
}

// ---- the correspondence step without allocations (include/e2sec.h: e2sec_icp_correspondences_static) --------------------
// The scratch of icp_nn_kernel / icp_accumulate_kernel in a file-scope device array: per source point the moved point
// (3 doubles), the nearest distance (double) and the nearest index (int), planes one after the other.  Every GPU that
// loads the library holds the 2.25 MiB until the process exits; all calls on a device share it.
#define E2_STATIC_ICP_BYTES ((size_t)E2SEC_STATIC_ICP_MAX_SOURCE * (3 * sizeof(double) + sizeof(double) + sizeof(int)))
static __device__ __align__(256) unsigned char g_icp_static_scratch[E2_STATIC_ICP_BYTES];

// In host code `g_icp_static_scratch` is only the symbol's host-side stand-in: its device address is asked for once per
// device, with that device current, and kept here.
#define E2_STATIC_MAX_DEVICES 64
static unsigned char *g_icp_static_addr[E2_STATIC_MAX_DEVICES];
static std::mutex g_icp_static_mutex;

static int icp_static_scratch(e2sec_ctx *ctx, unsigned char **out)
{
    if (ctx->device < 0 || ctx->device >= E2_STATIC_MAX_DEVICES) return fail(ctx, E2SEC_ERR_ARG, "e2sec_icp_correspondences_static: device index out of range");
    std::lock_guard<std::mutex> lock(g_icp_static_mutex);
    if (!g_icp_static_addr[ctx->device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_icp_static_scratch);          // the current device's copy (cudaSetDevice by the caller)
        if (e != cudaSuccess || !p) return fail(ctx, E2SEC_ERR_CUDA, std::string("cudaGetSymbolAddress(g_icp_static_scratch): ") + cudaGetErrorString(e));
        g_icp_static_addr[ctx->device] = (unsigned char *)p;
    }
    *out = g_icp_static_addr[ctx->device];
    return E2SEC_OK;
}

extern "C" int e2sec_icp_correspondences_static(e2sec_ctx *ctx, const double *source_dev, int32_t n_source, const double *target_dev,
                                                int32_t n_target, const double transform[16], double max_distance, double *stats_dev)
{
    if (!ctx || !source_dev || !target_dev || !transform || !stats_dev || n_source < 1 || n_target < 1 || !(max_distance > 0.0)) return E2SEC_ERR_ARG;
    if (n_source > E2SEC_STATIC_ICP_MAX_SOURCE)
        return fail(ctx, E2SEC_ERR_CAPACITY, "e2sec_icp_correspondences_static: n_source " + std::to_string(n_source) + " exceeds the static scratch (E2SEC_STATIC_ICP_MAX_SOURCE = " +
                                                 std::to_string(E2SEC_STATIC_ICP_MAX_SOURCE) + " points); use e2sec_icp_correspondences");
    CU(cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    if (int rc = icp_static_scratch(ctx, &base)) return rc;
    double *moved = reinterpret_cast<double *>(base);
    double *nn_d2 = moved + 3 * (size_t)E2SEC_STATIC_ICP_MAX_SOURCE;
    int *nn = reinterpret_cast<int *>(nn_d2 + (size_t)E2SEC_STATIC_ICP_MAX_SOURCE);
    cudaStream_t st = ctx->stream;
    IcpParams p;
    memcpy(p.T, transform, sizeof(p.T));
    p.max_d2 = max_distance * max_distance;
    // nothing in the scratch has to start at zero: icp_nn_kernel writes moved / nn_d2 / nn for every i < n_source and
    // icp_accumulate_kernel reads exactly those; stats_dev[0..16] are all written.
    const long long threads = (long long)n_source * 32;
    icp_nn_kernel<<<(unsigned)((threads + E2_F0_THREADS - 1) / E2_F0_THREADS), E2_F0_THREADS, 0, st>>>(source_dev, n_source, target_dev, n_target, p, nn, nn_d2, moved); LAUNCHED();
    icp_accumulate_kernel<<<1, 1024, 0, st>>>(moved, target_dev, nn, nn_d2, n_source, stats_dev); LAUNCHED();
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    return E2SEC_OK;
}
