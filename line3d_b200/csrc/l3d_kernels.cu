// line3d_b200/csrc/l3d_kernels.cu -- sm_100a kernels and C ABI (include/l3d_b200.h) of the
// Line3D matching-and-scoring path.  Written from scratch for B200; the reference's
// cudawrapper.cu is cited only for the semantics each stage must reproduce.
//
// Everything stays on the device.  Pair tests (phase A) run per batch of views, the rest (phase B) per group of
// views of one dependency level (see VArgs); DESIGN.md section 4 has the scheduling.
//
//   hoist      per pair-tested (view, neighbour), per segment: epipolar lines of the endpoints (FP64 -> FP32)
//   pre-test   every (src, tgt) pair: conservative tests on the four epipolar residuals
//              q_m^T F p_n  -> 1 candidate bit per pair                   [FP32 issue bound]
//   exact      candidates only: the reference's arithmetic, bit for bit   (l3d_device.cuh)
//   assemble   count -> scan -> fill: the view's match list in the reference's canonical
//              (src, cam, tgt) order, new raw matches merged with the reversed records
//              that earlier views delivered -- no sort, no host round trip
//   score      matches of a segment are bucketed by depth; every match looks for supporters in the
//              buckets its depth window touches: a 2-compare depth gate that is a superset of the
//              reference's 3-D gate, exact evaluation of the survivors through a per-warp queue
//   reduce     best match per segment, median depth, filter (score > 1), and the
//              reversed hand-over lists for later views, regrouped on the device
//
// Tensor cores are not used: nothing here is a dense contraction.
#include "../../include/l3d_b200.h"
#include "l3d_device.cuh"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

using namespace l3d;

// ---- device code and small utilities, one file per stage (all in this translation unit) ----
#include "l3d_common.cuh"
#include "l3d_pairtests.cuh"
#include "l3d_scoring.cuh"
#include "l3d_groups.cuh"
#include "l3d_multigpu.cuh"
#include "l3d_collinearity.cuh"
#include "l3d_postprocess.cuh"
#include "l3d_geometry.cuh"

// ---- host side: state, phase A, phase B, schedule ----
#include "l3d_context.h"
#include "l3d_phase_a.h"
#include "l3d_phase_b.h"
#include "l3d_schedule.h"

// ==========================================================================================
// C ABI
// ==========================================================================================
extern "C" {

int l3d_create(int device, l3d_ctx** out)
{
    if (!out) { g_create_error = "l3d_create: null output pointer"; return 1; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) { g_create_error = std::string("l3d_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path"; return 1; }
    if (device < 0 || device >= n) { g_create_error = "l3d_create: device index out of range"; return 1; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_error = std::string("l3d_create: cudaSetDevice: ") + cudaGetErrorString(e); return 1; }
    l3d_ctx* ctx = new l3d_ctx();
    ctx->device = device;
    { int sm = 0; if (cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sm > 0) ctx->n_sm = sm; }
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);        // the per-view chain is latency-bound: it goes first
    if ((e = cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, prio_hi)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); delete ctx; return 1; }
    {   // keep stream-ordered allocations cached across synchronisations (hand-over buffers are allocated per view)
        cudaMemPool_t pool; unsigned long long keep = ~0ull;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    for (int k = 0; k < 2; ++k) cudaEventCreate(&ctx->ev[k]);
    for (int q = 0; q < 2; ++q) for (int k = 0; k < 6; ++k) cudaEventCreate(&ctx->grp[q].ev[k]);
    for (int q = 0; q < 2; ++q) for (int k = 0; k < 2; ++k) cudaEventCreateWithFlags(&ctx->grp[q].ev_x[k], cudaEventDisableTiming);
    if ((e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); cudaStreamDestroy(ctx->stream); delete ctx; return 1; }
    if ((e = cudaStreamCreateWithPriority(&ctx->fin_stream, cudaStreamNonBlocking, prio_hi)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); l3d_destroy(ctx); return 1; }
    for (int q = 0; q < 2; ++q) {
        PhaseASet& set = ctx->pa[q];
        if ((e = cudaStreamCreateWithPriority(&set.stream, cudaStreamNonBlocking, prio_lo)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); l3d_destroy(ctx); return 1; }
        for (int k = 0; k < 3; ++k) cudaEventCreate(&set.pa_ev[k]);
        for (int k = 0; k < 5; ++k) cudaEventCreate(&set.pa_kev[k]);
    }
    for (int k = 0; k < 2; ++k) cudaEventCreate(&ctx->mark_ev[k]);
    *out = ctx;
    return 0;
}

int l3d_comm_unique_id(void* id)
{
    std::string why;
    NcclApi* api = nccl_api(&why);
    if (!api || !id) { g_create_error = api ? "l3d_comm_unique_id: null argument" : "l3d_comm_unique_id: " + why; return 1; }
    ncclUniqueId u;
    static_assert(sizeof(ncclUniqueId) == L3D_COMM_ID_BYTES, "ncclUniqueId size");
    const ncclResult_t r = api->GetUniqueId(&u);
    if (r != ncclSuccess) { g_create_error = std::string("ncclGetUniqueId: ") + api->GetErrorString(r); return 1; }
    memcpy(id, &u, sizeof u);
    return 0;
}

int l3d_comm_init(l3d_ctx* ctx, int rank, int world, const void* id, int result_root)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (world < 1 || world > MR_MAX_RANKS || rank < 0 || rank >= world || (world > 1 && !id) || result_root >= world) { ctx->err = "l3d_comm_init: bad arguments (1 <= world <= 16)"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        l3d_reset_matching(ctx);
        if (l3d_comm_destroy(ctx)) return 1;
        if (world == 1) return 0;
        std::string why;
        ctx->nccl = nccl_api(&why);
        if (!ctx->nccl) { ctx->err = "l3d_comm_init: " + why; return 1; }
        ncclUniqueId u; memcpy(&u, id, sizeof u);
        L3D_NCCL(ctx->nccl->CommInitRank(&ctx->comm, world, u, rank));
        ctx->rank = rank; ctx->world = world; ctx->result_root = result_root;
        return 0;
    });
}

int l3d_comm_destroy(l3d_ctx* ctx)
{
    if (!ctx) return 1;
    if (ctx->comm) {
        cudaSetDevice(ctx->device);
        if (!ctx->comm_broken) { drain_pending(ctx); cudaStreamSynchronize(ctx->stream); }
        mr_release_window(ctx);
        if (ctx->comm_broken) ctx->nccl->CommAbort(ctx->comm); else ctx->nccl->CommDestroy(ctx->comm);
        ctx->comm = nullptr;
    }
    ctx->rank = 0; ctx->world = 1; ctx->result_root = -1; ctx->comm_broken = false;
    return 0;
}

int l3d_reset_matching(l3d_ctx* ctx)
{
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    drain_pending(ctx);
    for (size_t v = 0; v < ctx->delivered.size(); ++v) ctx->delivered[v].clear();
    ctx->deliver_bufs.clear();
    ctx->pool.put_all();                                 // hand-over buffers never taken, or not yet given back
    ctx->level.clear();
    ctx->plan.clear(); ctx->order.clear(); ctx->results.clear(); ctx->call_next = 0; ctx->plan_started = false; ctx->batch_begin = ctx->batch_end = 0; ctx->fly_begin = ctx->fly_end = 0;
    ctx->arena_chunk = 0; ctx->arena_used = 0;           // retained results are given up, their memory is reused
    cudaStreamSynchronize(ctx->stream);
    return 0;
}

void l3d_destroy(l3d_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    l3d_reset_matching(ctx);
    l3d_comm_destroy(ctx);
    ctx->d_segs.release(); ctx->d_viewd.release(); ctx->pa[0].release(); ctx->pa[1].release();
    ctx->d_tile_sums.release(); ctx->d_total.release();
    ctx->grp[0].release(); ctx->grp[1].release(); ctx->pool.release();
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->fin_stream) cudaStreamDestroy(ctx->fin_stream);
    for (size_t k = 0; k < ctx->arena.size(); ++k) cudaFreeHost(ctx->arena[k].first);
    for (int k = 0; k < 2; ++k) if (ctx->mark_ev[k]) cudaEventDestroy(ctx->mark_ev[k]);
    ctx->h_filtered.release(); ctx->h_ints.release();
    ctx->d_csegs.release(); ctx->d_cbits.release(); ctx->d_ccnt.release(); ctx->d_coff.release(); ctx->d_cpairs.release(); ctx->h_cpairs.release();
    ctx->h_cands.release(); ctx->d_cands.release(); ctx->d_hscore.release(); ctx->d_ew.release(); ctx->d_ecnt.release(); ctx->d_eoff.release(); ctx->d_kept.release(); ctx->h_kept.release(); ctx->d_ekeys.release(); ctx->d_eorder.release(); ctx->h_eorder.release();
    ctx->d_fit_start.release(); ctx->d_fit_count.release(); ctx->d_fit_member.release(); ctx->d_fit_cam.release(); ctx->d_fit_camopen.release(); ctx->d_fit_open.release(); ctx->d_fit_dbl.release(); ctx->d_fit_order.release(); ctx->d_fit_dist.release();
    ctx->d_geo_cams.release(); ctx->d_geo_start.release(); ctx->d_geo_other.release(); ctx->d_geo_ids.release(); ctx->d_geo_count.release(); ctx->d_geo_sim.release(); ctx->d_geo_sims.release(); ctx->d_geo_pairs.release(); ctx->d_geo_F.release();
    ctx->d_hyp.release(); ctx->d_hin.release(); ctx->d_hout.release(); ctx->d_pairs.release(); ctx->d_aff.release(); ctx->d_medians.release(); ctx->d_flush.release(); ctx->d_rdd.release(); ctx->h_rdd.release();
    for (int k = 0; k < 2; ++k) if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* l3d_last_error(const l3d_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }
uint64_t l3d_kernel_launches(const l3d_ctx* ctx) { return ctx ? ctx->launches : 0; }

int l3d_collinearity(l3d_ctx* ctx, const float* segments, int n, float collin_s, const l3d_collin_pair** pairs, size_t* count)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (!pairs || !count || (n > 0 && !segments)) { ctx->err = "l3d_collinearity: null argument"; return 1; }
        *pairs = nullptr; *count = 0;
        if (n <= 1) return 0;
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int wpr = (n + 31) / 32;
        L3D_CUDA(ctx->d_csegs.ensure(size_t(n))); L3D_CUDA(ctx->d_cbits.ensure(size_t(n) * wpr));
        L3D_CUDA(ctx->d_ccnt.ensure(size_t(n) + 1)); L3D_CUDA(ctx->d_coff.ensure(size_t(n) + 2));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_csegs.p, segments, sizeof(float4) * size_t(n), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemsetAsync(ctx->d_ccnt.p, 0, sizeof(int) * (size_t(n) + 1), st));
        const float s2 = collin_s * collin_s;                                       // cudawrapper.cu:850
        k_collin_count<<<dim3((wpr + 63) / 64, n), 64, 0, st>>>(ctx->d_csegs.p, n, s2, ctx->d_ccnt.p, ctx->d_cbits.p, wpr);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        if (scan_exclusive(ctx, ctx->d_ccnt.p, ctx->d_coff.p, n)) return 1;
        int total = 0;
        if (fetch_int(ctx, ctx->d_coff.p + n, &total)) return 1;
        if (total == 0) return 0;
        L3D_CUDA(ctx->d_cpairs.ensure(size_t(total))); L3D_CUDA(ctx->h_cpairs.ensure(size_t(total)));
        k_collin_fill<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_csegs.p, n, s2, ctx->d_coff.p, ctx->d_cbits.p, wpr, ctx->d_cpairs.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(ctx->h_cpairs.p, ctx->d_cpairs.p, sizeof(l3d_collin_pair) * size_t(total), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        *pairs = ctx->h_cpairs.p; *count = size_t(total);
        return 0;
    });
}

int l3d_collinearity_prealloc(const float* d_segments, int n, float collin_s, l3d_collin_pair* d_pairs, int capacity,
                              int* d_count, void* cuda_stream, char* err, size_t err_len)
{
    // No cudaMalloc, cudaFree, cudaMallocAsync or any other allocating call may appear in this function.
    auto fail = [&](int code, const char* msg) { if (err && err_len) snprintf(err, err_len, "l3d_collinearity_prealloc: %s", msg); return code; };
    if (n < 0 || capacity < 0 || (n > 0 && !d_segments) || (capacity > 0 && !d_pairs) || !d_count) return fail(1, "null or negative argument");
    if (n > PREALLOC_MAX_SEGMENTS) {
        char b[160]; snprintf(b, sizeof b, "%d segments exceed the built-in limit of %d (use l3d_collinearity, which allocates)", n, PREALLOC_MAX_SEGMENTS);
        return fail(2, b);
    }
    // the kernels read the segments as float4 and write 12-byte records and an int: a misaligned pointer would raise a
    // sticky device fault that takes the caller's whole CUDA context with it, so it is refused here
    if ((reinterpret_cast<uintptr_t>(d_segments) & 15u) || (reinterpret_cast<uintptr_t>(d_pairs) & 3u) || (reinterpret_cast<uintptr_t>(d_count) & 3u))
        return fail(1, "d_segments must be 16-byte aligned, d_pairs and d_count 4-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // Set-up, once per device and with that device current: the array's name is only a host-side handle for the
    // symbol, the kernels need its address in device memory.  The per-device cache is a table of atomics: two threads
    // that make their first call at the same time both look the address up and store the same value.
    static std::atomic<unsigned int*> scratch_of_device[64];
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) return fail(3, e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    if (st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread) {      // the work must go to the device whose array was looked up
        int sdev = -1;
        e = cudaStreamGetDevice(st, &sdev);
        if (e != cudaSuccess) return fail(7, cudaGetErrorString(e));
        if (sdev != dev) return fail(7, "cuda_stream belongs to another device than the current one (cudaSetDevice to the stream's device first)");
    }
    unsigned int* scratch = scratch_of_device[dev].load(std::memory_order_acquire);
    if (!scratch) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_prealloc_scratch);
        if (e != cudaSuccess || !p) return fail(4, e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned null");
        scratch = static_cast<unsigned int*>(p);
        scratch_of_device[dev].store(scratch, std::memory_order_release);
    }
    unsigned* bits = scratch;
    int* cnt = reinterpret_cast<int*>(scratch + PREALLOC_CNT_OFF);
    int* off = reinterpret_cast<int*>(scratch + PREALLOC_OFF_OFF);
    if (n <= 1) {
        e = cudaMemsetAsync(d_count, 0, sizeof(int), st);
        return e == cudaSuccess ? 0 : fail(5, cudaGetErrorString(e));
    }
    const int wpr = (n + 31) / 32;
    const float s2 = collin_s * collin_s;                                       // cudawrapper.cu:850
    // the counts are accumulated with atomicAdd and must start from zero; the bitmask is fully rewritten
    e = cudaMemsetAsync(cnt, 0, sizeof(int) * (size_t(n) + 1), st);
    if (e != cudaSuccess) return fail(5, cudaGetErrorString(e));
    k_collin_count<<<dim3((wpr + 63) / 64, n), 64, 0, st>>>(reinterpret_cast<const float4*>(d_segments), n, s2, cnt, bits, wpr);
    k_scan_small<<<1, SCAN_T, 0, st>>>(cnt, off, n);
    k_collin_fill_capped<<<(n * 32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(d_segments), n, s2, off, bits, wpr, d_pairs, capacity, d_count);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(6, cudaGetErrorString(e));
    return 0;
}

int l3d_set_views(l3d_ctx* ctx, int n_views, const l3d_view* views)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n_views <= 0 || !views) { ctx->err = "l3d_set_views: no views"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        l3d_reset_matching(ctx);
        ctx->V = n_views;
        ctx->views.assign(views, views + n_views);
        ctx->seg_off.assign(size_t(n_views) + 1, 0);
        ctx->max_seg = 0;
        for (int v = 0; v < n_views; ++v) {
            if (views[v].n_segments < 0 || (views[v].n_segments > 0 && !views[v].segments && !(ctx->world > 1 && ctx->rank != 0))) { ctx->err = "l3d_set_views: bad segment array"; return 1; }
            if (views[v].n_segments > 65535) { ctx->err = "l3d_set_views: more than 65535 segments in a view"; return 1; }
            ctx->seg_off[v + 1] = ctx->seg_off[v] + views[v].n_segments;
            ctx->max_seg = std::max(ctx->max_seg, views[v].n_segments);
        }
        ctx->total_segs = ctx->seg_off[n_views];
        L3D_CUDA(ctx->d_segs.ensure(size_t(std::max(ctx->total_segs, 1))));
        L3D_CUDA(ctx->d_viewd.ensure(size_t(n_views)));
        std::vector<ViewD> vd(n_views);
        for (int v = 0; v < n_views; ++v) {
            // (several GPUs on one scene: a rank other than 0 may leave the segments out -- pointer NULL with n_segments
            // set -- and gets them from rank 0 in the broadcast below)
            if (views[v].n_segments && views[v].segments)
                L3D_CUDA(cudaMemcpyAsync(ctx->d_segs.p + ctx->seg_off[v], views[v].segments, sizeof(float4) * size_t(views[v].n_segments), cudaMemcpyHostToDevice, ctx->stream));
            memcpy(vd[v].RtKinv, views[v].RtKinv_d, sizeof vd[v].RtKinv); memcpy(vd[v].C, views[v].C_d, sizeof vd[v].C);
            vd[v].k_lower = views[v].k_lower; vd[v].k_upper = views[v].k_upper; vd[v].seg_off = ctx->seg_off[v]; vd[v].n_seg = views[v].n_segments;
            ctx->views[v].segments = nullptr;
        }
        L3D_CUDA(cudaMemcpyAsync(ctx->d_viewd.p, vd.data(), sizeof(ViewD) * size_t(n_views), cudaMemcpyHostToDevice, ctx->stream));
        // several GPUs on one scene: the segment table is replicated from rank 0 over NVLink (the camera tables are a few
        // hundred bytes per view and every rank has computed them from the same input)
        if (ctx->world > 1 && ctx->total_segs > 0)
            L3D_NCCL(ctx->nccl->Broadcast(ctx->d_segs.p, ctx->d_segs.p, size_t(ctx->total_segs) * sizeof(float4), ncclChar, 0, ctx->comm, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->delivered.assign(size_t(n_views), std::vector<Delivered>());
        ctx->n_hyp = 0;
        return 0;
    });
}

int l3d_deliver(l3d_ctx* ctx, int to_view, int from_view, size_t n, const l3d_match* records)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (to_view < 0 || to_view >= ctx->V || from_view < 0 || from_view >= ctx->V || to_view == from_view) { ctx->err = "l3d_deliver: view index out of range"; return 1; }
        if (n > 0 && !records) { ctx->err = "l3d_deliver: null argument"; return 1; }
        if (n > size_t(0x7fffffff)) { ctx->err = "l3d_deliver: too many records"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        drain_pending(ctx);
        const int n_to = ctx->views[to_view].n_segments, n_from = ctx->views[from_view].n_segments;
        // grouped by the receiving view's segment (counting sort), as k_deliver_scatter leaves them
        std::vector<int> start(size_t(n_to) + 2, 0), tgt(std::max<size_t>(n, 1)); std::vector<float4> dep(std::max<size_t>(n, 1));
        for (size_t k = 0; k < n; ++k) {
            if (records[k].src_seg < 0 || records[k].src_seg >= n_to || records[k].tgt_seg < 0 || records[k].tgt_seg >= n_from) { ctx->err = "l3d_deliver: segment index out of range"; return 1; }
            ++start[size_t(records[k].src_seg) + 1];
        }
        int max_group = 0;
        for (int q = 0; q < n_to; ++q) { max_group = std::max(max_group, start[size_t(q) + 1]); start[size_t(q) + 1] += start[q]; }
        start[size_t(n_to) + 1] = start[n_to];
        { std::vector<int> cur(start.begin(), start.begin() + n_to);
          for (size_t k = 0; k < n; ++k) { const int pos = cur[records[k].src_seg]++; tgt[pos] = records[k].tgt_seg; dep[pos] = make_float4(records[k].depths[0], records[k].depths[1], records[k].depths[2], records[k].depths[3]); } }
        DeliverBuf buf; buf.refs = 1; buf.counts.assign(1, int(n)); buf.maxg.assign(1, max_group);
        L3D_CUDA(ctx->pool.get(sizeof(int) * (size_t(n_to) + 2), reinterpret_cast<void**>(&buf.start)));
        L3D_CUDA(ctx->pool.get(sizeof(int) * std::max<size_t>(n, 1), reinterpret_cast<void**>(&buf.tgt)));
        L3D_CUDA(ctx->pool.get(sizeof(float4) * std::max<size_t>(n, 1), reinterpret_cast<void**>(&buf.dep)));
        L3D_CUDA(cudaMemcpyAsync(buf.start, start.data(), sizeof(int) * (size_t(n_to) + 2), cudaMemcpyHostToDevice, ctx->stream));
        if (n) {
            L3D_CUDA(cudaMemcpyAsync(buf.tgt, tgt.data(), sizeof(int) * n, cudaMemcpyHostToDevice, ctx->stream));
            L3D_CUDA(cudaMemcpyAsync(buf.dep, dep.data(), sizeof(float4) * n, cudaMemcpyHostToDevice, ctx->stream));
        }
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));          // the host vectors go away
        Delivered dl; dl.from_view = from_view; dl.count = int(n); dl.count_all = (long long)n; dl.max_group = max_group; dl.t_start = buf.start; dl.t_tgt = buf.tgt; dl.t_dep = buf.dep;
        dl.owner = int(ctx->deliver_bufs.size()); dl.slot = 0;
        ctx->deliver_bufs.push_back(buf);
        ctx->delivered[to_view].push_back(dl);
        return 0;
    });
}

int l3d_plan_matching(l3d_ctx* ctx, int n_views, const l3d_view_plan* plans)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        cudaSetDevice(ctx->device);
        drain_pending(ctx);
        ctx->plan.clear(); ctx->order.clear(); ctx->results.clear(); ctx->call_next = 0; ctx->plan_started = false; ctx->batch_begin = ctx->batch_end = 0; ctx->fly_begin = ctx->fly_end = 0;
        if (n_views <= 0) return 0;
        if (!plans) { ctx->err = "l3d_plan_matching: null argument"; return 1; }
        for (int k = 0; k < n_views; ++k) {
            const l3d_view_plan& p = plans[k];
            if (p.view < 0 || p.view >= ctx->V || p.n_neighbors <= 0 || !p.neighbors) { ctx->plan.clear(); ctx->err = "l3d_plan_matching: bad entry"; return 1; }
            PlannedView pv; pv.view = p.view; pv.nb.assign(p.neighbors, p.neighbors + p.n_neighbors);
            for (int c = 0; c < p.n_neighbors; ++c) {
                if (p.neighbors[c].view < 0 || p.neighbors[c].view >= ctx->V || p.neighbors[c].view == p.view) { ctx->plan.clear(); ctx->err = "l3d_plan_matching: bad neighbour"; return 1; }
                // (several GPUs on one scene: the rows of the rank with most of them -- the batches must be cut the same way on every rank)
                if (p.neighbors[c].to_be_matched) pv.pair_tests += uint64_t((ctx->views[p.view].n_segments + ctx->world - 1) / ctx->world) * uint64_t(ctx->views[p.neighbors[c].view].n_segments);
            }
            ctx->plan.push_back(pv);
        }
        // Execution order.  View b needs view a first only if a hands records over to b (a comes earlier and marks b
        // "deliver"), so the schedule is a dependency graph, usually of a few independent chains (views along a street;
        // C2: 200 views, longest chain 50).  The library works through it level by level -- the views of one level do not
        // depend on each other and run side by side -- and hands the results out in the order of the calls.
        {
            const size_t n = ctx->plan.size();
            std::vector<int> at(size_t(ctx->V), -1), level(n, 1);
            for (size_t k = 0; k < n; ++k) at[ctx->plan[k].view] = int(k);
            for (size_t k = 0; k < n; ++k)
                for (size_t c = 0; c < ctx->plan[k].nb.size(); ++c) {
                    const l3d_neighbor& nbr = ctx->plan[k].nb[c];
                    const int to = at[nbr.view];
                    if (nbr.deliver && to > int(k)) level[to] = std::max(level[to], level[k] + 1);
                }
            ctx->level = level;
            ctx->order.resize(n);
            for (size_t k = 0; k < n; ++k) ctx->order[k] = k;
            std::stable_sort(ctx->order.begin(), ctx->order.end(), [&level](size_t a, size_t b) { return level[a] < level[b]; });
            l3d_ctx::Stashed empty; memset(&empty, 0, sizeof empty);
            ctx->results.assign(n, empty);
        }
        // Batches: the pair tests of batch i + 1 run beside the per-view work of batch i, so the schedule is cut into
        // about eight of them (not smaller than 2.5e8 pair tests: a batch has fixed costs), none above the memory budget.
        uint64_t total = 0;
        for (size_t k = 0; k < ctx->plan.size(); ++k) total += ctx->plan[k].pair_tests;
        {   // tuning knobs, read once per schedule
            const char* e;
            ctx->bg_blocks = (e = getenv("L3D_BG_BLOCKS")) ? std::max(0, std::min(atoi(e), 16)) : 0;
            ctx->p1_rows = (e = getenv("L3D_P1_ROWS")) ? std::max(128, atoi(e) / 128 * 128) : 0;
            ctx->no_overlap = (e = getenv("L3D_NO_OVERLAP")) != nullptr && atoi(e) != 0;
            ctx->batch_div = (e = getenv("L3D_BATCH_DIV")) ? std::max(1, std::min(atoi(e), 256)) : 8;
            // (one scene on several GPUs: a rank's share of a view is small and every group costs an exchange, so larger groups)
            ctx->group_views = (e = getenv("L3D_GROUP_VIEWS")) ? std::max(1, std::min(atoi(e), VG_MAX)) : (ctx->world > 1 ? 16 : 8);
        }
        ctx->batch_pairs = ctx->batch_override ? ctx->batch_override : std::min<uint64_t>(ctx->pair_budget, std::max<uint64_t>(total / uint64_t(ctx->batch_div), 250000000ull));
        if (ctx->stage_events && !ctx->batch_override) ctx->batch_pairs = ctx->pair_budget;      // per-stage timing: nothing runs beside anything else
        return 0;
    });
}

int l3d_debug_set_batch_pairs(l3d_ctx* ctx, uint64_t pairs)
{
    if (!ctx) return 1;
    ctx->batch_override = pairs;
    return 0;
}

int l3d_debug_set_pretest(l3d_ctx* ctx, int mode)
{
    if (!ctx || mode < 0 || mode > 1) return 1;
    ctx->pretest_mode = mode;
    return 0;
}

int l3d_debug_set_score_block(l3d_ctx* ctx, int hypotheses)
{
    if (!ctx || !(hypotheses == 0 || hypotheses == 32 || hypotheses == 64 || hypotheses == 128)) return 1;
    ctx->score_block = hypotheses;
    return 0;
}

int l3d_debug_set_candidate_limit(l3d_ctx* ctx, uint64_t limit)
{
    if (!ctx) return 1;
    ctx->cand_limit = (limit == 0 || limit > 0x7fff0000ull) ? 0x7fff0000ull : limit;
    return 0;
}

int l3d_debug_chain_times(l3d_ctx* ctx, int on, double out[24])
{
    if (!ctx) return 1;
    if (out) for (int k = 0; k < 24; ++k) { out[k] = ctx->chain_ms[k]; ctx->chain_ms[k] = 0.0; }
    ctx->chain_prof = on != 0;
    return 0;
}

int l3d_debug_host_times(l3d_ctx* ctx, double out[8])
{
    if (!ctx || !out) return 1;
    for (int k = 0; k < 8; ++k) { out[k] = ctx->host_s[k]; ctx->host_s[k] = 0.0; }
    return 0;
}

int l3d_retain_results(l3d_ctx* ctx, int on)
{
    if (!ctx) return 1;
    ctx->retain = on != 0;
    return 0;
}

int l3d_match_view(l3d_ctx* ctx, int view, int n_neighbors, const l3d_neighbor* neighbors, float sigma_p, float sigma_a, l3d_match_result* result)
{
    const int rc = guarded(ctx, [&]() { return match_view_impl(ctx, view, n_neighbors, neighbors, sigma_p, sigma_a, result, true); });
    if (rc && ctx && ctx->world > 1) ctx->comm_broken = true;
    return rc;
}
int l3d_match_view_resident(l3d_ctx* ctx, int view, int n_neighbors, const l3d_neighbor* neighbors, float sigma_p, float sigma_a, l3d_match_result* result)
{
    const int rc = guarded(ctx, [&]() { return match_view_impl(ctx, view, n_neighbors, neighbors, sigma_p, sigma_a, result, false); });
    if (rc && ctx && ctx->world > 1) ctx->comm_broken = true;
    return rc;
}

// ---- the neighbour graph and the fundamental matrices (l3d_geometry.cuh) ----
static const char* geometry_args_ok(int n_views, const l3d_camera* cams, const int32_t* start, const int32_t* other, const float* sim, const int32_t* count, const int32_t* out)
{
    if (n_views < 0 || (n_views > 0 && (!cams || !start || !count))) return "null argument";
    if (n_views > 0 && start[0] != 0) return "start[0] must be 0";
    for (int v = 0; v < n_views; ++v) if (start[v + 1] < start[v]) return "start must not decrease";
    const int nnz = n_views > 0 ? start[n_views] : 0;
    if (nnz > 0 && (!other || !sim || !out)) return "null argument";
    for (int v = 0; v < n_views; ++v)
        for (int q = start[v]; q < start[v + 1]; ++q) {
            if (other[q] < 0 || other[q] >= n_views || other[q] == v) return "co-visible view index out of range";
            if (q > start[v] && other[q - 1] >= other[q]) return "co-visible views must be in ascending order";
        }
    return nullptr;
}

int l3d_visual_neighbors(l3d_ctx* ctx, int n_views, const l3d_camera* cams, const int32_t* start, const int32_t* other, const float* similarity,
                         float min_baseline, int k, int32_t* count, int32_t* neighbors)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (const char* bad = geometry_args_ok(n_views, cams, start, other, similarity, count, neighbors)) { ctx->err = std::string("l3d_visual_neighbors: ") + bad; return 1; }
        if (n_views == 0) return 0;
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const size_t nnz = size_t(start[n_views]);
        L3D_CUDA(ctx->d_geo_cams.ensure(size_t(n_views))); L3D_CUDA(ctx->d_geo_start.ensure(size_t(n_views) + 1)); L3D_CUDA(ctx->d_geo_count.ensure(size_t(n_views)));
        L3D_CUDA(ctx->d_geo_other.ensure(std::max<size_t>(nnz, 1))); L3D_CUDA(ctx->d_geo_ids.ensure(std::max<size_t>(nnz, 1)));
        L3D_CUDA(ctx->d_geo_sim.ensure(std::max<size_t>(nnz, 1))); L3D_CUDA(ctx->d_geo_sims.ensure(std::max<size_t>(nnz, 1)));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_geo_cams.p, cams, sizeof(l3d_camera) * size_t(n_views), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_geo_start.p, start, sizeof(int) * (size_t(n_views) + 1), cudaMemcpyHostToDevice, st));
        if (nnz) {
            L3D_CUDA(cudaMemcpyAsync(ctx->d_geo_other.p, other, sizeof(int) * nnz, cudaMemcpyHostToDevice, st));
            L3D_CUDA(cudaMemcpyAsync(ctx->d_geo_sim.p, similarity, sizeof(float) * nnz, cudaMemcpyHostToDevice, st));
        }
        k_visual_neighbors<<<(n_views + 63) / 64, 64, 0, st>>>(ctx->d_geo_cams.p, n_views, ctx->d_geo_start.p, ctx->d_geo_other.p, ctx->d_geo_sim.p, min_baseline, k,
                                                            ctx->d_geo_count.p, ctx->d_geo_ids.p, ctx->d_geo_sims.p);
        L3D_CUDA(cudaGetLastError());
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(count, ctx->d_geo_count.p, sizeof(int) * size_t(n_views), cudaMemcpyDeviceToHost, st));
        if (nnz) L3D_CUDA(cudaMemcpyAsync(neighbors, ctx->d_geo_ids.p, sizeof(int) * nnz, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        return 0;
    });
}

int l3d_fundamentals(l3d_ctx* ctx, int n_views, const l3d_camera* cams, size_t n_pairs, const int32_t* pairs, double* F)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (n_pairs == 0) return 0;
        if (n_views <= 0 || !cams || !pairs || !F) { ctx->err = "l3d_fundamentals: null argument"; return 1; }
        for (size_t p = 0; p < 2 * n_pairs; ++p) if (pairs[p] < 0 || pairs[p] >= n_views) { ctx->err = "l3d_fundamentals: view index out of range"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        L3D_CUDA(ctx->d_geo_cams.ensure(size_t(n_views))); L3D_CUDA(ctx->d_geo_pairs.ensure(n_pairs)); L3D_CUDA(ctx->d_geo_F.ensure(9 * n_pairs));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_geo_cams.p, cams, sizeof(l3d_camera) * size_t(n_views), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_geo_pairs.p, pairs, sizeof(int2) * n_pairs, cudaMemcpyHostToDevice, st));
        k_fundamentals<<<unsigned((n_pairs + 127) / 128), 128, 0, st>>>(ctx->d_geo_cams.p, ctx->d_geo_pairs.p, n_pairs, ctx->d_geo_F.p);
        L3D_CUDA(cudaGetLastError());
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(F, ctx->d_geo_F.p, sizeof(double) * 9 * n_pairs, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        return 0;
    });
}

static const char* clusters_ok(const l3d_similarity* tr, int n_clusters, const int32_t* start, const void* a, const uint32_t* cam, const int32_t* count, const double* seg)
{
    if (n_clusters < 0 || !tr || (n_clusters > 0 && (!start || !count))) return "null argument";
    if (n_clusters > 0 && start[0] != 0) return "start[0] must be 0";
    for (int c = 0; c < n_clusters; ++c) if (start[c + 1] < start[c]) return "start must not decrease";
    if (n_clusters > 0 && start[n_clusters] > 0 && (!a || !cam || !seg)) return "null argument";
    for (int c = 0; c < n_clusters; ++c)
        for (int q = start[c] + 1; q < start[c + 1]; ++q) if (cam[q - 1] > cam[q]) return "the members of a cluster must be in (camera, segment) order";
    return nullptr;
}

int l3d_fit_lines(l3d_ctx* ctx, const l3d_similarity* back, int n_clusters, const int32_t* start, const int32_t* member, const uint32_t* member_camera,
                  int32_t* seg_count, double* segments)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (const char* bad = clusters_ok(back, n_clusters, start, member, member_camera, seg_count, segments)) { ctx->err = std::string("l3d_fit_lines: ") + bad; return 1; }
        if (n_clusters == 0) return 0;
        const size_t total = size_t(start[n_clusters]);
        for (size_t q = 0; q < total; ++q) if (member[q] < 0 || size_t(member[q]) >= ctx->n_hyp) { ctx->err = "l3d_fit_lines: hypothesis index out of range (indices refer to the last l3d_unproject)"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const size_t t1 = std::max<size_t>(total, 1);
        L3D_CUDA(ctx->d_fit_start.ensure(size_t(n_clusters) + 1)); L3D_CUDA(ctx->d_fit_count.ensure(size_t(n_clusters)));
        L3D_CUDA(ctx->d_fit_member.ensure(t1)); L3D_CUDA(ctx->d_fit_cam.ensure(t1)); L3D_CUDA(ctx->d_fit_camopen.ensure(t1)); L3D_CUDA(ctx->d_fit_open.ensure(t1));
        L3D_CUDA(ctx->d_fit_dbl.ensure(4 * 6 * t1)); L3D_CUDA(ctx->d_fit_order.ensure(2 * t1)); L3D_CUDA(ctx->d_fit_dist.ensure(2 * t1));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_fit_start.p, start, sizeof(int) * (size_t(n_clusters) + 1), cudaMemcpyHostToDevice, st));
        if (total) {
            L3D_CUDA(cudaMemcpyAsync(ctx->d_fit_member.p, member, sizeof(int) * total, cudaMemcpyHostToDevice, st));
            L3D_CUDA(cudaMemcpyAsync(ctx->d_fit_cam.p, member_camera, sizeof(unsigned) * total, cudaMemcpyHostToDevice, st));
        }
        double* pts = ctx->d_fit_dbl.p; double* A = pts + 6 * t1; double* G = A + 6 * t1; double* seg = G + 6 * t1;
        k_fit_lines<<<n_clusters, FIT_T, 0, st>>>(*back, n_clusters, ctx->d_fit_start.p, ctx->d_fit_member.p, ctx->d_fit_cam.p, ctx->d_hyp.p, pts, A, G,
                                                        ctx->d_fit_order.p, ctx->d_fit_dist.p, ctx->d_fit_open.p, ctx->d_fit_camopen.p, ctx->d_fit_count.p, seg);
        L3D_CUDA(cudaGetLastError());
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(seg_count, ctx->d_fit_count.p, sizeof(int) * size_t(n_clusters), cudaMemcpyDeviceToHost, st));
        if (total) L3D_CUDA(cudaMemcpyAsync(segments, seg, sizeof(double) * 6 * total, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        return 0;
    });
}

int l3d_debug_fit_lines_host(const l3d_similarity* back, int n_clusters, const int32_t* start, const double* points, const uint32_t* member_camera,
                             int32_t* seg_count, double* segments)
{
    try {
        if (clusters_ok(back, n_clusters, start, points, member_camera, seg_count, segments)) return 1;
        for (int c = 0; c < n_clusters; ++c) {
            const int b = start[c], m = start[c + 1] - b;
            std::vector<double> A(size_t(m) * 6 + 1), G(size_t(m) * 6 + 1); std::vector<int> order(size_t(m) * 2 + 1), cam_open(size_t(m) + 1);
            std::vector<float> dist(size_t(m) * 2 + 1); std::vector<unsigned char> open_line(size_t(m) + 1);
            seg_count[c] = geo_fit_cluster(*back, m, points + size_t(b) * 6, member_camera + b, A.data(), G.data(), order.data(), dist.data(), open_line.data(), cam_open.data(),
                                           segments + size_t(b) * 6);
        }
        return 0;
    } catch (...) { return 1; }
}

int l3d_debug_visual_neighbors_host(int n_views, const l3d_camera* cams, const int32_t* start, const int32_t* other, const float* similarity,
                                    float min_baseline, int k, int32_t* count, int32_t* neighbors)
{
    try {
        if (geometry_args_ok(n_views, cams, start, other, similarity, count, neighbors)) return 1;
        std::vector<float> sims(n_views > 0 ? size_t(start[n_views]) + 1 : 1);
        for (int v = 0; v < n_views; ++v)
            count[v] = geo_select_neighbors(cams, v, start[v + 1] - start[v], other + start[v], similarity + start[v], min_baseline, k, neighbors + start[v], sims.data() + start[v]);
        return 0;
    } catch (...) { return 1; }
}

int l3d_debug_fundamentals_host(int n_views, const l3d_camera* cams, size_t n_pairs, const int32_t* pairs, double* F)
{
    if (n_pairs == 0) return 0;
    if (n_views <= 0 || !cams || !pairs || !F) return 1;
    for (size_t p = 0; p < n_pairs; ++p) {
        if (pairs[2 * p] < 0 || pairs[2 * p] >= n_views || pairs[2 * p + 1] < 0 || pairs[2 * p + 1] >= n_views) return 1;
        geo_fundamental(cams[pairs[2 * p]], cams[pairs[2 * p + 1]], F + 9 * p);
    }
    return 0;
}

int l3d_unproject(l3d_ctx* ctx, size_t n, const l3d_hypothesis_in* in, l3d_segment3d* out)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n == 0) { ctx->n_hyp = 0; return 0; }
        if (!in || !out) { ctx->err = "l3d_unproject: null argument"; return 1; }
        for (size_t k = 0; k < n; ++k)
            if (in[k].view < 0 || in[k].view >= ctx->V || in[k].seg < 0 || in[k].seg >= ctx->views[in[k].view].n_segments) { ctx->err = "l3d_unproject: view/segment out of range"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        L3D_CUDA(ctx->d_hin.ensure(n)); L3D_CUDA(ctx->d_hout.ensure(n)); L3D_CUDA(ctx->d_hyp.ensure(n));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_hin.p, in, sizeof(l3d_hypothesis_in) * n, cudaMemcpyHostToDevice, ctx->stream));
        k_unproject<<<unsigned((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_segs.p, ctx->d_viewd.p, ctx->d_hin.p, n, ctx->d_hyp.p, ctx->d_hout.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(out, ctx->d_hout.p, sizeof(l3d_segment3d) * n, cudaMemcpyDeviceToHost, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->n_hyp = n;
        return 0;
    });
}

int l3d_affinity(l3d_ctx* ctx, const float* median_depth, size_t n_pairs, const l3d_pair* pairs, float sigma_a, float* out)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n_pairs == 0) return 0;
        if (!median_depth || !pairs || !out) { ctx->err = "l3d_affinity: null argument"; return 1; }
        if (ctx->n_hyp == 0) { ctx->err = "l3d_affinity: no hypotheses resident (call l3d_unproject first)"; return 1; }
        for (size_t k = 0; k < n_pairs; ++k)
            if (pairs[k].a < 0 || size_t(pairs[k].a) >= ctx->n_hyp || pairs[k].b < 0 || size_t(pairs[k].b) >= ctx->n_hyp) { ctx->err = "l3d_affinity: pair index out of range"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        L3D_CUDA(ctx->d_pairs.ensure(n_pairs)); L3D_CUDA(ctx->d_aff.ensure(n_pairs)); L3D_CUDA(ctx->d_medians.ensure(size_t(ctx->V)));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_medians.p, median_depth, sizeof(float) * size_t(ctx->V), cudaMemcpyHostToDevice, ctx->stream));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_pairs.p, pairs, sizeof(l3d_pair) * n_pairs, cudaMemcpyHostToDevice, ctx->stream));
        const float two_log = 2.0f * logf(0.01f);                                   // view.cc:376
        k_affinity<<<unsigned((n_pairs + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_hyp.p, ctx->d_viewd.p, ctx->d_medians.p, ctx->d_pairs.p, n_pairs, sigma_a, two_log, ctx->d_aff.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(out, ctx->d_aff.p, sizeof(float) * n_pairs, cudaMemcpyDeviceToHost, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        return 0;
    });
}

int l3d_candidate_buffer(l3d_ctx* ctx, size_t n, l3d_candidate** buf)
{
    if (!ctx || !buf) return 1;
    return guarded(ctx, [&]() -> int {
        L3D_CUDA(cudaSetDevice(ctx->device));
        L3D_CUDA(ctx->h_cands.ensure(std::max<size_t>(n, 1)));
        *buf = ctx->h_cands.p;
        return 0;
    });
}

int l3d_affinity_edges(l3d_ctx* ctx, const float* median_depth, const float* hyp_score, size_t n, const l3d_candidate* cands, float sigma_a,
                       const l3d_kept_edge** kept, size_t* n_kept, const uint32_t** order)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (!kept || !n_kept) { ctx->err = "l3d_affinity_edges: null argument"; return 1; }
        *kept = nullptr; *n_kept = 0;
        if (order) *order = nullptr;
        if (n == 0) return 0;
        if (!median_depth || !hyp_score || !cands) { ctx->err = "l3d_affinity_edges: null argument"; return 1; }
        if (ctx->n_hyp == 0) { ctx->err = "l3d_affinity_edges: no hypotheses resident (call l3d_unproject first)"; return 1; }
        if (n > size_t(0x7fffffff) - AE_T) { ctx->err = "l3d_affinity_edges: too many candidates"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int blocks = int((n + AE_T - 1) / AE_T);
        L3D_CUDA(ctx->d_cands.ensure(n)); L3D_CUDA(ctx->d_ew.ensure(n)); L3D_CUDA(ctx->d_hscore.ensure(ctx->n_hyp)); L3D_CUDA(ctx->d_medians.ensure(size_t(ctx->V)));
        L3D_CUDA(ctx->d_ecnt.ensure(size_t(blocks) + 1)); L3D_CUDA(ctx->d_eoff.ensure(size_t(blocks) + 2));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_medians.p, median_depth, sizeof(float) * size_t(ctx->V), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_hscore.p, hyp_score, sizeof(float) * ctx->n_hyp, cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_cands.p, cands, sizeof(l3d_candidate) * n, cudaMemcpyHostToDevice, st));
        const float two_log = 2.0f * logf(0.01f);                                   // view.cc:376
        k_affinity_edges<<<blocks, AE_T, 0, st>>>(ctx->d_hyp.p, ctx->d_viewd.p, ctx->d_medians.p, ctx->d_hscore.p, ctx->d_cands.p, n, sigma_a, two_log, ctx->d_ew.p, ctx->d_ecnt.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        if (scan_exclusive(ctx, ctx->d_ecnt.p, ctx->d_eoff.p, blocks)) return 1;
        int total = 0;
        if (fetch_int(ctx, ctx->d_eoff.p + blocks, &total)) return 1;
        if (total == 0) return 0;
        L3D_CUDA(ctx->d_kept.ensure(size_t(total))); L3D_CUDA(ctx->h_kept.ensure(size_t(total)));
        k_affinity_compact<<<blocks, AE_T, 0, st>>>(ctx->d_ew.p, n, ctx->d_eoff.p, ctx->d_kept.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(ctx->h_kept.p, ctx->d_kept.p, sizeof(l3d_kept_edge) * size_t(total), cudaMemcpyDeviceToHost, st));
        if (order) {                                        // the clustering's edge order, while the records travel
            int n_pad = BS_TILE; while (n_pad < total) n_pad <<= 1;
            L3D_CUDA(ctx->d_ekeys.ensure(size_t(n_pad))); L3D_CUDA(ctx->d_eorder.ensure(size_t(total))); L3D_CUDA(ctx->h_eorder.ensure(size_t(total)));
            k_edge_keys<<<(n_pad + 255) / 256, 256, 0, st>>>(ctx->d_kept.p, total, n_pad, ctx->d_ekeys.p);
            ctx->launches += 1;
            for (int k = 2; k <= n_pad; k <<= 1) {
                int j = k >> 1;
                for (; j >= BS_TILE; j >>= 1) { k_bitonic_global<<<n_pad / 256, 256, 0, st>>>(ctx->d_ekeys.p, j, k); ctx->launches += 1; }
                k_bitonic_shared<<<n_pad / BS_TILE, BS_TILE / 2, 0, st>>>(ctx->d_ekeys.p, j, k);
                ctx->launches += 1;
            }
            k_edge_order<<<(total + 255) / 256, 256, 0, st>>>(ctx->d_ekeys.p, total, ctx->d_eorder.p);
            ctx->launches += 1;
            L3D_CUDA(cudaGetLastError());
            L3D_CUDA(cudaMemcpyAsync(ctx->h_eorder.p, ctx->d_eorder.p, sizeof(unsigned) * size_t(total), cudaMemcpyDeviceToHost, st));
        }
        L3D_CUDA(cudaStreamSynchronize(st));
        *kept = ctx->h_kept.p; *n_kept = size_t(total);
        if (order) *order = ctx->h_eorder.p;
        return 0;
    });
}

int l3d_diffuse(l3d_ctx* ctx, int n_nodes, size_t nnz, const l3d_edge* edges, l3d_edge* out)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (nnz == 0 || n_nodes <= 0) return 0;
        if (!edges || !out) { ctx->err = "l3d_diffuse: null argument"; return 1; }
        if (nnz > size_t(0x7ffffff0)) { ctx->err = "l3d_diffuse: more than 2^31 entries"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int N = int(nnz);
        for (int k = 0; k < N; ++k)
            if (edges[k].i < 0 || edges[k].i >= n_nodes || edges[k].j < 0 || edges[k].j >= n_nodes) { ctx->err = "l3d_diffuse: node id out of range"; return 1; }
        // W column-sorted (sparsematrix.cc:81-85), P row-sorted (sparsematrix.cc:158-161); the comparators order by the
        // integer ids only, stable, so equal keys (none for a valid input) keep their list order.  Two stable counting
        // passes per order (secondary key first) instead of comparison sorts: O(nnz + nodes).
        auto order_by = [&](const std::vector<int>& in, bool by_col_then_row, std::vector<int>& outv) {
            std::vector<int> tmp(N), cnt(size_t(n_nodes) + 1);
            for (int pass = 0; pass < 2; ++pass) {
                const bool use_i = by_col_then_row ? (pass == 0) : (pass == 1);          // secondary key first
                const std::vector<int>& src = pass == 0 ? in : tmp;
                std::vector<int>& dst = pass == 0 ? tmp : outv;
                std::fill(cnt.begin(), cnt.end(), 0);
                for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[src[k]]; ++cnt[size_t(use_i ? e.i : e.j) + 1]; }
                for (int d = 0; d < n_nodes; ++d) cnt[size_t(d) + 1] += cnt[d];
                for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[src[k]]; dst[cnt[use_i ? e.i : e.j]++] = src[k]; }
            }
        };
        std::vector<int> ident(N), wi(N), pi(N);
        for (int k = 0; k < N; ++k) ident[k] = k;
        order_by(ident, true, wi);                                                  // W: by (col, row)
        order_by(wi, false, pi);                                                    // P is built from W's stored order: by (row, col); pi holds input indices
        // pinned staging, one buffer: prow pcol wcol tpos (N ints each), rstart cstart (n_nodes each), pval wval (N floats each)
        const size_t n_int = size_t(4) * N + size_t(2) * n_nodes, n_all = n_int + size_t(2) * N;
        L3D_CUDA(ctx->h_rdd.ensure(n_all)); L3D_CUDA(ctx->d_rdd.ensure(n_all + size_t(N)));
        int* prow = ctx->h_rdd.p; int* pcol = prow + N; int* wcol = pcol + N; int* tpos = wcol + N; int* rstart = tpos + N; int* cstart = rstart + n_nodes;
        float* pval = reinterpret_cast<float*>(cstart + n_nodes); float* wval = pval + N;
        for (int d = 0; d < n_nodes; ++d) { rstart[d] = -1; cstart[d] = -1; }
        for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[wi[k]]; wcol[k] = e.j; wval[k] = e.w; if (cstart[e.j] < 0) cstart[e.j] = k; }
        for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[pi[k]]; prow[k] = e.i; pcol[k] = e.j; pval[k] = e.w; if (rstart[e.i] < 0) rstart[e.i] = k; }
        // cudawrapper.cu:809-827: first entry (r, c) of row r = col_y with col == row_y (the transposed entry's place);
        // the rows are sorted by column, so it is found by bisection; and the algorithmic traffic of one iteration:
        // per entry the lock-step walk reads min(row length, column length) pairs of 16 bytes from each operand
        double walk = 0.0;
        {
            std::vector<int> rlen(size_t(n_nodes), 0), clen(size_t(n_nodes), 0);
            for (int k = 0; k < N; ++k) { ++rlen[prow[k]]; ++clen[wcol[k]]; }
            const size_t nt = std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), 16));
            std::vector<double> part(nt, 0.0);
            auto work = [&](size_t t) {
                const int y0 = int(size_t(N) * t / nt), y1 = int(size_t(N) * (t + 1) / nt);
                for (int y = y0; y < y1; ++y) {
                    const int r = pcol[y], c = prow[y];
                    int lo = rstart[r], hi = lo < 0 ? -1 : lo + rlen[r];
                    int found = -1;
                    while (lo >= 0 && lo < hi) { const int mid = (lo + hi) >> 1; if (pcol[mid] < c) lo = mid + 1; else hi = mid; }
                    if (lo >= 0 && lo < rstart[r] + rlen[r] && prow[lo] == r && pcol[lo] == c) found = lo;
                    tpos[y] = found;
                    part[t] += double(std::min(rlen[r], clen[c]));
                }
            };
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t);
            work(0);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
            for (size_t t = 0; t < nt; ++t) walk += part[t];
        }
        int* d_prow = ctx->d_rdd.p; int* d_pcol = d_prow + N; int* d_wcol = d_pcol + N; int* d_tpos = d_wcol + N; int* d_rstart = d_tpos + N; int* d_cstart = d_rstart + n_nodes;
        float* d_pa = reinterpret_cast<float*>(d_cstart + n_nodes); float* d_wval = d_pa + N; float* d_pb = d_wval + N;
        L3D_CUDA(cudaMemcpyAsync(ctx->d_rdd.p, ctx->h_rdd.p, sizeof(int) * n_all, cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(d_pb, d_pa, sizeof(float) * N, cudaMemcpyDeviceToDevice, st));
        const int rb = (n_nodes + 255) / 256, eb = (N + 255) / 256;
        k_rdd_normalise<<<rb, 256, 0, st>>>(d_prow, d_pa, d_rstart, n_nodes, N);
        ctx->launches += 1;
        float* P = d_pa; float* Pn = d_pb;
        float step_ms = 0.f;
        for (int it = 0; it < 10; ++it) {                                           // L3D_RDD_MAX_ITER, cudawrapper.h:35
            if (it == 1) L3D_CUDA(cudaEventRecord(ctx->ev[0], st));                 // the step kernel's own duration: iteration 2 (warm)
            k_rdd_step<<<eb, 256, 0, st>>>(d_prow, d_pcol, P, d_wcol, d_wval, d_rstart, d_cstart, d_tpos, Pn, N);
            if (it == 1) L3D_CUDA(cudaEventRecord(ctx->ev[1], st));
            std::swap(P, Pn);
            ctx->launches += 1;
            if (it < 9) { k_rdd_normalise<<<rb, 256, 0, st>>>(d_prow, P, d_rstart, n_nodes, N); ctx->launches += 1; }
        }
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(pval, P, sizeof(float) * N, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        cudaEventElapsedTime(&step_ms, ctx->ev[0], ctx->ev[1]);
        ctx->rdd_step_ms = step_ms; ctx->rdd_step_bytes = 32.0 * walk + 16.0 * double(N); ctx->rdd_nnz = N;
        for (int k = 0; k < N; ++k) { out[k].i = prow[k]; out[k].j = pcol[k]; out[k].w = pval[k]; }
        return 0;
    });
}

int l3d_diffuse_stats(l3d_ctx* ctx, double out[3])
{
    if (!ctx || !out) return 1;
    out[0] = ctx->rdd_step_ms; out[1] = ctx->rdd_step_bytes; out[2] = double(ctx->rdd_nnz);
    return 0;
}

int l3d_debug_math(l3d_ctx* ctx, int which, int n, const float* in, float* out)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n <= 0) return 0;
        L3D_CUDA(cudaSetDevice(ctx->device));
        float *d_in, *d_out;
        L3D_CUDA(cudaMalloc(&d_in, sizeof(float) * n)); L3D_CUDA(cudaMalloc(&d_out, sizeof(float) * n));
        L3D_CUDA(cudaMemcpyAsync(d_in, in, sizeof(float) * n, cudaMemcpyHostToDevice, ctx->stream));
        k_debug_math<<<(n + 255) / 256, 256, 0, ctx->stream>>>(which, d_in, d_out, n);
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(out, d_out, sizeof(float) * n, cudaMemcpyDeviceToHost, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        cudaFree(d_in); cudaFree(d_out);
        return 0;
    });
}

int l3d_last_timings(l3d_ctx* ctx, float ms[6])
{
    if (!ctx || !ms) return 1;
    for (int k = 0; k < 6; ++k) ms[k] = ctx->last_ms[k];
    return 0;
}

int l3d_set_stage_timing(l3d_ctx* ctx, int on)
{
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    drain_pending(ctx);
    ctx->stage_events = on != 0;
    return 0;
}

int l3d_mark(l3d_ctx* ctx, int which)
{
    if (!ctx || which < 0 || which > 1) return 1;
    L3D_CUDA(cudaSetDevice(ctx->device));
    if (which == 1) {            // the closing mark also comes after the side stream's last result records
        L3D_CUDA(cudaEventRecord(ctx->ev[1], ctx->fin_stream));
        L3D_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev[1], 0));
    }
    L3D_CUDA(cudaEventRecord(ctx->mark_ev[which], ctx->stream));
    return 0;
}

int l3d_marks_elapsed(l3d_ctx* ctx, float* ms)
{
    if (!ctx || !ms) return 1;
    L3D_CUDA(cudaSetDevice(ctx->device));
    L3D_CUDA(cudaEventSynchronize(ctx->mark_ev[1]));
    L3D_CUDA(cudaEventElapsedTime(ms, ctx->mark_ev[0], ctx->mark_ev[1]));
    return 0;
}

int l3d_last_kernel_timings(l3d_ctx* ctx, float ms[3])
{
    if (!ctx || !ms) return 1;
    ms[0] = ctx->pa_kernel_ms[0]; ms[1] = ctx->pa_kernel_ms[1]; ms[2] = ctx->pa_kernel_ms[2];
    return 0;
}

int l3d_flush_l2(l3d_ctx* ctx)
{
    if (!ctx) return 1;
    L3D_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = size_t(256) << 20;
    L3D_CUDA(ctx->d_flush.ensure(bytes));
    L3D_CUDA(cudaMemsetAsync(ctx->d_flush.p, ctx->flush_value++ & 0xff, bytes, ctx->stream));
    L3D_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// which = 0: scalar FFMA chains, 1: packed FFMA2 chains.  Best of five launches.
static int measure_fp32(l3d_ctx* ctx, int which, float* tflops)
{
    L3D_CUDA(cudaSetDevice(ctx->device));
    cudaDeviceProp prop; L3D_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 13;
    float* d = nullptr;
    L3D_CUDA(cudaMalloc(&d, sizeof(float) * size_t(blocks) * threads));
    const float b = 1.0000001f, c = 1e-7f;
    float best = 0.f;
    for (int r = 0; r < 6; ++r) {                       // the first launch warms up
        L3D_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
        if (which == 0) k_fma_probe<<<blocks, threads, 0, ctx->stream>>>(d, iters, b, c);
        else k_fma2_probe<<<blocks, threads, 0, ctx->stream>>>(d, iters, b, c);
        L3D_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        float ms = 0.f; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        const double fl = double(blocks) * threads * double(iters) * 128.0 * 2.0;      // 128 FMAs per thread and loop trip in both variants
        if (r > 0) best = std::max(best, float(fl / (double(ms) * 1e-3) / 1e12));
    }
    ctx->launches += 6;
    cudaFree(d);
    *tflops = best;
    return 0;
}

int l3d_measure_fp32_peak(l3d_ctx* ctx, float* tflops)
{
    if (!ctx || !tflops) return 1;
    return measure_fp32(ctx, 0, tflops);
}

int l3d_measure_fp32x2_peak(l3d_ctx* ctx, float* tflops)
{
    if (!ctx || !tflops) return 1;
    return measure_fp32(ctx, 1, tflops);
}

}  // extern "C"
