// line3d_b200/csrc/l3d_phase_a.h -- host side of phase A: launch helpers and run_phase_a (one batch of pair tests: hoist, pre-tests, exact test, bounds read back once).
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

template <typename K, typename... Args>
inline cudaError_t pdl_launch(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

template <typename C>
int scan_exclusive(C* ctx, const int* in, int* out /* n + 1 */, int n)
{
    if (n <= SCAN_SMALL) {
        L3D_CUDA(pdl_launch(k_scan_small, dim3(1), dim3(SCAN_T), 0, ctx->stream, in, out, n));
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        return 0;
    }
    const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (tiles > SCAN_TILE) { ctx->err = "scan: input too long"; return 1; }
    L3D_CUDA(ctx->d_tile_sums.ensure(size_t(std::max(tiles, 1))));
    L3D_CUDA(ctx->d_total.ensure(1));
    L3D_CUDA(pdl_launch(k_scan_tiles, dim3(std::max(tiles, 1)), dim3(SCAN_T), 0, ctx->stream, in, out, ctx->d_tile_sums.p, n));
    L3D_CUDA(pdl_launch(k_scan_sums, dim3(1), dim3(SCAN_T), 0, ctx->stream, ctx->d_tile_sums.p, tiles, ctx->d_total.p));
    L3D_CUDA(pdl_launch(k_scan_add, dim3(std::max(tiles, 1)), dim3(SCAN_T), 0, ctx->stream, out, ctx->d_tile_sums.p, n, ctx->d_total.p));
    ctx->launches += 3;
    L3D_CUDA(cudaGetLastError());
    return 0;
}

// rows of a view this process holds: all of them, or every world-th segment starting at its rank
inline int local_rows(const l3d_ctx* ctx, int n_segments) { return ctx->world <= 1 ? n_segments : (n_segments > ctx->rank ? (n_segments - ctx->rank + ctx->world - 1) / ctx->world : 0); }

int fetch_int(l3d_ctx* ctx, const int* dptr, int* out)
{
    L3D_CUDA(ctx->h_ints.ensure(16));
    L3D_CUDA(cudaMemcpyAsync(ctx->h_ints.p, dptr, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    L3D_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = ctx->h_ints.p[0];
    return 0;
}

// Phase A for a set of (source view, neighbour) jobs: hoist, pre-test, scan, exact test.
// Results stay on the device: d_pa_off (candidate offsets per global row = staging ranges),
// d_pa_raw (raw matches per global row), d_stage_col / d_stage_dep.
int run_phase_a(l3d_ctx* cx, PhaseASet* ctx, const std::vector<JobHost>& jh)   // ctx: the set being filled (the error macro writes ctx->err)
{
    L3D_CUDA(cudaSetDevice(cx->device));
    cudaStream_t st = ctx->stream;
    const int n_jobs = int(jh.size());
    // Pair tests that run beside the per-view chain must leave it room: the chain is a sequence of short dependent
    // kernels whose blocks can only start when registers and shared memory are free, and kernels that fill every SM
    // make each of them wait for blocks to retire.  In the background the large kernels therefore ask for unused
    // dynamic shared memory, which caps them at 3-4 blocks per SM (about half of each SM stays free); alone they run
    // at full occupancy.
    size_t bg_p1 = 0, bg_p2 = 0, bg_ex = 0, bg_ef = 0;
    if (ctx->background && cx->bg_blocks > 0) {
        // what leaves room for bg_blocks blocks per SM: 228 KB of shared memory, 1 KB reserved per block
        auto cap = [&](const void* fn, size_t* dyn) -> cudaError_t {
            cudaFuncAttributes a; cudaError_t e = cudaFuncGetAttributes(&a, fn); if (e != cudaSuccess) return e;
            const long long per = (228ll * 1024) / (cx->bg_blocks + 1) + 1024;      // one block more would not fit
            long long d = per - 1024 - (long long)a.sharedSizeBytes;
            d = std::max(0ll, std::min(d, 227ll * 1024 - (long long)a.sharedSizeBytes));
            *dyn = size_t(d);
            return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(d));
        };
        L3D_CUDA(cap(reinterpret_cast<const void*>(&k_pretest<4, 6>), &bg_p1)); L3D_CUDA(cap(reinterpret_cast<const void*>(&k_pretest2), &bg_p2));
        L3D_CUDA(cap(reinterpret_cast<const void*>(&k_expand), &bg_ex)); L3D_CUDA(cap(reinterpret_cast<const void*>(&k_exact_flat), &bg_ef));
    }
    ctx->jobs_host.assign(size_t(n_jobs), JobDev());
    ctx->job_cands.assign(size_t(n_jobs), 0);
    ctx->job_rowmax.assign(size_t(n_jobs), 0);
    ctx->job_raw.assign(size_t(n_jobs), 0); ctx->job_raw_rowmax.assign(size_t(n_jobs), 0);
    ctx->pa_kev_set[0] = ctx->pa_kev_set[1] = false;
    if (n_jobs == 0) return 0;
    long long sh_tot = 0, th_tot = 0, words = 0, rows = 0, blocks = 0, l1_words = 0, l1_blocks = 0;
    const int p1_rows = cx->p1_rows > 0 ? cx->p1_rows : P1_ROWS;
    int max_n = 1;
    for (int j = 0; j < n_jobs; ++j) {
        const l3d_view& sv = cx->views[jh[j].src_view]; const l3d_view& tv = cx->views[jh[j].tgt_view];
        JobDev& d = ctx->jobs_host[j];
        memcpy(d.F, jh[j].F, sizeof d.F); memcpy(d.RtKinv_s, sv.RtKinv, sizeof d.RtKinv_s); memcpy(d.RtKinv_t, tv.RtKinv, sizeof d.RtKinv_t);
        memcpy(d.C_s, sv.C, sizeof d.C_s); memcpy(d.C_t, tv.C, sizeof d.C_t);
        d.src_off = cx->seg_off[jh[j].src_view]; d.n_src = local_rows(cx, sv.n_segments); d.src_first = cx->world > 1 ? cx->rank : 0; d.src_stride = cx->world > 1 ? cx->world : 1; d.tgt_off = cx->seg_off[jh[j].tgt_view]; d.n_tgt = tv.n_segments;
        d.sh_off = int(sh_tot); d.th_off = int(th_tot); d.cand_off = words; d.wpr = (d.n_tgt + 31) / 32;
        d.row_off = int(rows); d.blk_off = int(blocks); d.col_blocks = std::max((d.n_tgt + PT_COLS - 1) / PT_COLS, 1);
        sh_tot += d.n_src; th_tot += d.n_tgt; words += (long long)d.n_src * d.wpr; rows += d.n_src;
        blocks += (long long)((d.n_src + PT_ROWS - 1) / PT_ROWS) * d.col_blocks;
        d.l1_off = l1_words; d.l1_blk_off = int(l1_blocks);
        l1_words += (long long)((d.n_src + 31) / 32) * d.n_tgt;
        l1_blocks += (long long)((d.n_src + p1_rows - 1) / p1_rows) * ((d.n_tgt + P1_COLS - 1) / P1_COLS);
        max_n = std::max(max_n, std::max(d.n_src, d.n_tgt));
        if (sh_tot > 0x7fffffff || th_tot > 0x7fffffff || rows > 0x7ffffff0 || blocks > 0x7ffffff0 || l1_blocks > 0x7ffffff0) { ctx->err = "phase A: batch too large"; return 1; }
    }
    L3D_CUDA(ctx->d_jobs.ensure(size_t(n_jobs)));
    L3D_CUDA(cudaMemcpyAsync(ctx->d_jobs.p, ctx->jobs_host.data(), sizeof(JobDev) * size_t(n_jobs), cudaMemcpyHostToDevice, st));
    L3D_CUDA(ctx->d_sh.ensure(size_t(std::max<long long>(sh_tot, 1)))); L3D_CUDA(ctx->d_sh1.ensure(size_t(std::max<long long>(sh_tot, 1)))); L3D_CUDA(ctx->d_th.ensure(size_t(std::max<long long>(th_tot, 1))));
    L3D_CUDA(ctx->d_cand.ensure(size_t(std::max<long long>(words, 1)))); L3D_CUDA(ctx->d_row_job.ensure(size_t(rows) + 1));
    L3D_CUDA(ctx->d_l1.ensure(size_t(std::max<long long>(l1_words, 1))));
    L3D_CUDA(ctx->d_pa_cnt.ensure(size_t(rows) + 1)); L3D_CUDA(ctx->d_pa_off.ensure(size_t(rows) + 2)); L3D_CUDA(ctx->d_pa_raw.ensure(size_t(rows) + 1));
    L3D_CUDA(cudaMemsetAsync(ctx->d_pa_cnt.p, 0, sizeof(int) * (size_t(rows) + 1), st));
    L3D_CUDA(cudaEventRecord(ctx->pa_ev[0], st));
    if (rows > 0 && blocks > 0) {
        k_hoist<<<dim3((max_n + 127) / 128, n_jobs), 128, 0, st>>>(cx->d_segs.p, ctx->d_jobs.p, ctx->d_sh.p, ctx->d_sh1.p, ctx->d_th.p, ctx->d_row_job.p, cx->pretest_mode);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[0], st));
        k_pretest<4, 6><<<unsigned(l1_blocks), P1_BLOCK, bg_p1, st>>>(ctx->d_sh1.p, ctx->d_th.p, ctx->d_jobs.p, n_jobs, ctx->d_l1.p, p1_rows);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[1], st));
        k_pretest2<<<unsigned(blocks), PT_BLOCK, bg_p2, st>>>(ctx->d_sh.p, ctx->d_th.p, ctx->d_jobs.p, n_jobs, ctx->d_l1.p, ctx->d_cand.p, ctx->d_pa_cnt.p);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[4], st));
        ctx->pa_kev_set[0] = true;
        ctx->launches += 3;
        L3D_CUDA(cudaGetLastError());
    }
    if (scan_exclusive(ctx, ctx->d_pa_cnt.p, ctx->d_pa_off.p, int(rows))) return 1;
    L3D_CUDA(cudaEventRecord(ctx->pa_ev[1], st));
    // candidate totals per job (and overall) come back in one small copy
    L3D_CUDA(ctx->d_pa_idx.ensure(size_t(n_jobs) + 1)); L3D_CUDA(ctx->d_pa_tot.ensure(size_t(n_jobs) + 1)); L3D_CUDA(ctx->h_res.ensure(2 * size_t(n_jobs) + 64));
    {
        std::vector<int> idx(size_t(n_jobs) + 1);
        for (int j = 0; j < n_jobs; ++j) idx[j] = ctx->jobs_host[j].row_off;
        idx[n_jobs] = int(rows);
        L3D_CUDA(cudaMemcpyAsync(ctx->d_pa_idx.p, idx.data(), sizeof(int) * idx.size(), cudaMemcpyHostToDevice, st));
        k_gather_int<<<(n_jobs + 256) / 256, 256, 0, st>>>(ctx->d_pa_off.p, ctx->d_pa_idx.p, n_jobs + 1, ctx->d_pa_tot.p);
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(ctx->h_res.p, ctx->d_pa_tot.p, sizeof(int) * (size_t(n_jobs) + 1), cudaMemcpyDeviceToHost, st));
        // largest candidate count of a single row, per job
        std::vector<int> lens(static_cast<size_t>(n_jobs));
        for (int j = 0; j < n_jobs; ++j) lens[j] = ctx->jobs_host[j].n_src;
        L3D_CUDA(ctx->d_rm_begin.ensure(size_t(n_jobs))); L3D_CUDA(ctx->d_rm_len.ensure(size_t(n_jobs))); L3D_CUDA(ctx->d_rm_out.ensure(size_t(n_jobs)));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_rm_begin.p, idx.data(), sizeof(int) * size_t(n_jobs), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_rm_len.p, lens.data(), sizeof(int) * size_t(n_jobs), cudaMemcpyHostToDevice, st));
        k_range_max<<<n_jobs, 256, 0, st>>>(ctx->d_pa_cnt.p, ctx->d_rm_begin.p, ctx->d_rm_len.p, 0, ctx->d_rm_out.p);
        L3D_CUDA(ctx->d_job_sum64.ensure(size_t(n_jobs))); L3D_CUDA(ctx->h_job_sum64.ensure(size_t(n_jobs)));
        k_range_sum64<<<n_jobs, 256, 0, st>>>(ctx->d_pa_cnt.p, ctx->d_rm_begin.p, ctx->d_rm_len.p, ctx->d_job_sum64.p);
        ctx->launches += 2;
        L3D_CUDA(cudaMemcpyAsync(ctx->h_res.p + n_jobs + 1, ctx->d_rm_out.p, sizeof(int) * size_t(n_jobs), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->h_job_sum64.p, ctx->d_job_sum64.p, sizeof(unsigned long long) * size_t(n_jobs), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));        // idx / lens are stack vectors: must be consumed before they go away
    }
    {   // Candidate offsets are 32-bit: a batch whose candidates do not fit (a candidate rate far above the usual 2 %:
        // degenerate epipolar geometry, or the pre-test switched off) is reported so that the caller can cut it smaller.
        unsigned long long all = 0ull;
        for (int j = 0; j < n_jobs; ++j) all += ctx->h_job_sum64.p[j];
        if (all > cx->cand_limit) {
            char b[200]; snprintf(b, sizeof b, "pair tests: %llu candidates in one batch exceed the limit of %llu (32-bit candidate index)", all, (unsigned long long)cx->cand_limit);
            ctx->err = b;
            return 2;
        }
    }
    for (int j = 0; j < n_jobs; ++j) ctx->job_cands[j] = uint64_t(ctx->h_res.p[j + 1] - ctx->h_res.p[j]);
    for (int j = 0; j < n_jobs; ++j) ctx->job_rowmax[j] = ctx->h_res.p[n_jobs + 1 + j];
    const int n_cand = ctx->h_res.p[n_jobs];
    L3D_CUDA(ctx->d_stage_col.ensure(size_t(std::max(n_cand, 1)))); L3D_CUDA(ctx->d_stage_dep.ensure(size_t(std::max(n_cand, 1))));
    L3D_CUDA(ctx->d_stage_row.ensure(size_t(std::max(n_cand, 1)))); L3D_CUDA(ctx->d_stage_ok.ensure(size_t(std::max(n_cand, 1))));
    L3D_CUDA(cudaMemsetAsync(ctx->d_pa_raw.p, 0, sizeof(int) * (size_t(rows) + 1), st));
    if (rows > 0 && n_cand > 0) {
        k_expand<<<unsigned((rows * 32 + 255) / 256), 256, bg_ex, st>>>(ctx->d_jobs.p, ctx->d_row_job.p, int(rows), ctx->d_cand.p, ctx->d_pa_off.p, ctx->d_stage_col.p, ctx->d_stage_row.p);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[2], st));
        k_exact_flat<<<unsigned((n_cand + EF_T - 1) / EF_T), EF_T, bg_ef, st>>>(cx->d_segs.p, ctx->d_jobs.p, ctx->d_row_job.p, n_cand, ctx->d_stage_col.p, ctx->d_stage_row.p,
                                                                          ctx->d_stage_ok.p, ctx->d_stage_dep.p, ctx->d_pa_raw.p);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[3], st));
        ctx->pa_kev_set[1] = true;
        ctx->launches += 2;
        L3D_CUDA(cudaGetLastError());
    }
    L3D_CUDA(cudaEventRecord(ctx->pa_ev[2], st));
    ctx->pa_ms_pending = true;
    // raw matches per job and per row: tight bounds for the per-view work (the candidate counts are 2.5x larger).
    // Waiting for them here costs one idle gap per batch, not per view.
    if (rows > 0 && n_cand > 0) {
        L3D_CUDA(ctx->d_pa_idx.ensure(2 * size_t(n_jobs) + 2));
        k_range_sum_max<<<n_jobs, 256, 0, st>>>(ctx->d_pa_raw.p, ctx->d_rm_begin.p, ctx->d_rm_len.p, ctx->d_pa_idx.p, ctx->d_pa_idx.p + n_jobs);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(ctx->h_res.p, ctx->d_pa_idx.p, sizeof(int) * 2 * size_t(n_jobs), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        for (int j = 0; j < n_jobs; ++j) { ctx->job_raw[j] = uint64_t(ctx->h_res.p[j]); ctx->job_raw_rowmax[j] = ctx->h_res.p[n_jobs + j]; }
    }
    return 0;
}

bool same_neighbors(const std::vector<l3d_neighbor>& a, int n, const l3d_neighbor* b)
{
    if (int(a.size()) != n) return false;
    for (int c = 0; c < n; ++c)
        if (a[c].view != b[c].view || (a[c].to_be_matched != 0) != (b[c].to_be_matched != 0) || memcmp(a[c].F, b[c].F, sizeof a[c].F) != 0) return false;
    return true;
}

}  // namespace
