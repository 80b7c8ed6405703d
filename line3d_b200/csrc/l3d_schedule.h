// line3d_b200/csrc/l3d_schedule.h -- the announced schedule: batches, the helper thread, groups in flight, l3d_match_view's implementation, and the exception guard of the C ABI.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

int join_worker(l3d_ctx* ctx);

// Collects the views that were put on the GPU ahead of their calls and are not going to be asked for (and the helper thread).
void drain_pending(l3d_ctx* ctx)
{
    join_worker(ctx);
    while (ctx->grp_begin < ctx->grp_end) { ViewGroup& gr = ctx->grp[ctx->grp_begin % 2]; finish_group(ctx, gr, false, nullptr); ctx->fly_begin += size_t(gr.count); ++ctx->grp_begin; }
    ctx->fly_begin = ctx->fly_end = 0; ctx->grp_begin = ctx->grp_end = 0;
}

// The batch of the announced schedule that starts at position `from` of the execution order: its job list, and
// where it ends.
size_t plan_batch(l3d_ctx* ctx, size_t from, std::vector<JobHost>& jh)
{
    size_t e = from; uint64_t pairs = 0;
    while (e < ctx->order.size() && (e == from || pairs + ctx->plan[ctx->order[e]].pair_tests <= ctx->batch_pairs)) {
        PlannedView& pv = ctx->plan[ctx->order[e]];
        if (e > from && jh.size() + pv.nb.size() > 65535) break;        // one grid row per job in k_hoist
        pv.first_job = int(jh.size());
        for (size_t c = 0; c < pv.nb.size(); ++c)
            if (pv.nb[c].to_be_matched) { JobHost j; j.src_view = pv.view; j.tgt_view = pv.nb[c].view; memcpy(j.F, pv.nb[c].F, sizeof j.F); jh.push_back(j); }
        pairs += pv.pair_tests; ++e;
    }
    return e;
}

// Pair tests of the batch starting at `from`, on the helper thread and the second stream, into the set that is not
// in use.  The per-view work of a batch keeps only a fraction of the SMs busy (short dependent kernels); the pair
// tests of the next batch fill the rest.
void start_worker(l3d_ctx* ctx, size_t from)
{
    PhaseASet* set = &ctx->pa[ctx->cur ^ 1];
    std::vector<JobHost> jh;
    set->begin = from; set->end = plan_batch(ctx, from, jh);
    ctx->pa_worker_rc = 0;
    set->background = true;
    ctx->pa_worker = std::thread([ctx, set, jh]() {
        try { ctx->pa_worker_rc = run_phase_a(ctx, set, jh); }
        catch (const std::exception& e) { set->err = std::string("pair tests (helper thread): ") + e.what(); ctx->pa_worker_rc = 1; }
        catch (...) { set->err = "pair tests (helper thread): unknown exception"; ctx->pa_worker_rc = 1; }
    });
    ctx->pa_worker_active = true;
}

int join_worker(l3d_ctx* ctx)
{
    if (!ctx->pa_worker_active) return 0;
    ctx->pa_worker.join();
    ctx->pa_worker_active = false;
    return ctx->pa_worker_rc;
}

// More views of the schedule onto the GPU, in execution order, as far as slots and resident pair tests allow; starts
// the next batch when the current one has been collected completely.
int top_up(l3d_ctx* ctx)
{
    if (ctx->fly_end >= ctx->batch_end) {
        if (ctx->fly_begin < ctx->fly_end || ctx->fly_end >= ctx->order.size()) return 0;
        bool have = false;
        if (ctx->pa_worker_active) {                             // its pair tests were started while the previous batch was scored
            PhaseASet& nx = ctx->pa[ctx->cur ^ 1];
            const auto tj = std::chrono::steady_clock::now();
            const int rc = join_worker(ctx);
            ctx->host_s[4] += std::chrono::duration<double>(std::chrono::steady_clock::now() - tj).count();
            if (rc == 0) {
                if (nx.begin != ctx->fly_end) { ctx->err = "l3d_match_view: internal error (batch out of step)"; return 1; }
                ctx->cur ^= 1; have = true;
            } else if (rc != 2) { ctx->err = nx.err; return 1; }
            // rc == 2: more candidates than a batch can index -- the batch is redone below, cut smaller
        }
        while (!have) {
            std::vector<JobHost> jh;
            PhaseASet& set = ctx->pa[ctx->cur];
            set.begin = ctx->fly_end; set.end = plan_batch(ctx, ctx->fly_end, jh);
            set.background = false;
            const auto tf = std::chrono::steady_clock::now();
            const int rc = run_phase_a(ctx, &set, jh);
            ctx->host_s[5] += std::chrono::duration<double>(std::chrono::steady_clock::now() - tf).count();
            if (rc == 2 && set.end - set.begin > 1 && ctx->world <= 1) {             // halve the batch and try again (one GPU only: ranks must cut alike)
                uint64_t pairs = 0;
                for (size_t e = set.begin; e < set.end; ++e) pairs += ctx->plan[ctx->order[e]].pair_tests;
                ctx->batch_pairs = std::max<uint64_t>(pairs / 2, 1);
                continue;
            }
            if (rc) { ctx->err = set.err + (rc == 2 ? " (a single view: not supported)" : ""); return 1; }
            have = true;
        }
        PhaseASet& set = ctx->pa[ctx->cur];
        const auto tt = std::chrono::steady_clock::now();
        if (mr_job_totals(ctx, set)) return 1;
        ctx->host_s[6] += std::chrono::duration<double>(std::chrono::steady_clock::now() - tt).count();
        ctx->batch_begin = set.begin; ctx->batch_end = set.end;
        ctx->launches += set.launches; set.launches = 0;
        if (ctx->batch_end < ctx->order.size() && !ctx->stage_events && !ctx->no_overlap) start_worker(ctx, ctx->batch_end);      // the following batch, meanwhile (not when stages are timed: nothing may run beside anything)
    }
    // the next groups: consecutive views of one dependency level, two groups in flight
    while (ctx->fly_end < ctx->batch_end && ctx->grp_end - ctx->grp_begin < 2) {
        GroupView gv[VG_MAX]; int count = 0;
        const int lvl = ctx->level[ctx->order[ctx->fly_end]];
        while (count < ctx->group_views && ctx->fly_end + size_t(count) < ctx->batch_end && ctx->level[ctx->order[ctx->fly_end + size_t(count)]] == lvl) {
            PlannedView& pv = ctx->plan[ctx->order[ctx->fly_end + size_t(count)]];
            GroupView v = {pv.view, int(pv.nb.size()), pv.nb.data(), pv.first_job, ctx->order[ctx->fly_end + size_t(count)]};
            gv[count++] = v;
        }
        ViewGroup& gr = ctx->grp[ctx->grp_end % 2];
        gr.pos0 = ctx->fly_end;
        if (ctx->world > 1) {
            // the record windows must hold this group; growing them is a collective step with nothing in flight
            // (every rank computes the same bound at the same point of the same schedule)
            const size_t need = mr_group_need(ctx, gv, count);
            if (need > ctx->win_cap) {
                if (ctx->grp_begin < ctx->grp_end) return 0;          // collect what is in flight first; the caller's loop comes back
                if (mr_ensure_window(ctx, need)) return 1;
            }
        }
        const auto t0 = std::chrono::steady_clock::now();
        if (enqueue_group(ctx, gr, gv, count, ctx->run_sigma_p, ctx->run_sigma_a)) return 1;
        ctx->host_s[0] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); ctx->host_s[3] += 1.0;
        ctx->fly_end += size_t(count); ++ctx->grp_end;
    }
    return 0;
}

int match_view_impl(l3d_ctx* ctx, int view, int n_nb, const l3d_neighbor* nb, float sigma_p, float sigma_a, l3d_match_result* res, bool copy_back)
{
    if (!ctx) return 1;
    if (n_nb <= 0 || !nb || !res) { ctx->err = "l3d_match_view: no neighbours / null argument"; return 1; }
    if (view < 0 || view >= ctx->V) { ctx->err = "l3d_match_view: view index out of range"; return 1; }
    L3D_CUDA(cudaSetDevice(ctx->device));
    memset(res, 0, sizeof *res);
    const bool on_plan = ctx->call_next < ctx->plan.size() && ctx->plan[ctx->call_next].view == view && same_neighbors(ctx->plan[ctx->call_next].nb, n_nb, nb);
    if (ctx->plan_started && !(on_plan && ctx->run_sigma_p == sigma_p && ctx->run_sigma_a == sigma_a && ctx->run_copy_back == copy_back)) {
        ctx->err = "l3d_match_view: the call does not follow the schedule announced with l3d_plan_matching (same views, neighbours, sigmas and entry point, in order); call l3d_reset_matching to start over";
        return 1;
    }
    if (on_plan) {
        // ---- the next view of the announced schedule: the library works through the schedule in its own order
        // (by dependency level) and ahead of the caller; wait until this view's result has been collected
        if (!ctx->plan_started) { ctx->plan_started = true; ctx->run_sigma_p = sigma_p; ctx->run_sigma_a = sigma_a; ctx->run_copy_back = copy_back; }
        ctx->want_records = ctx->run_copy_back;
        const size_t p = ctx->call_next;
        while (!ctx->results[p].ready) {
            if (top_up(ctx)) { drain_pending(ctx); return 1; }
            if (ctx->grp_begin >= ctx->grp_end) { ctx->err = "l3d_match_view: internal error (schedule exhausted)"; return 1; }
            ViewGroup& gr = ctx->grp[ctx->grp_begin % 2];
            const int rc = finish_group(ctx, gr, ctx->run_copy_back, nullptr);
            ctx->fly_begin += size_t(gr.count); ++ctx->grp_begin;
            if (rc) { drain_pending(ctx); return 1; }
        }
        if (ctx->copies_pending) { L3D_CUDA(cudaStreamSynchronize(ctx->copy_stream)); ctx->copies_pending = false; }
        *res = ctx->results[p].res;
        memcpy(ctx->last_ms, ctx->results[p].ms, sizeof ctx->last_ms); memcpy(ctx->pa_kernel_ms, ctx->results[p].kms, sizeof ctx->pa_kernel_ms);
        ++ctx->call_next;
        if (top_up(ctx)) { drain_pending(ctx); return 1; }       // keep the GPU busy while the caller works
        return 0;
    }
    // ---- a call off the schedule (before it started) voids it: this view alone -------------------
    drain_pending(ctx);
    ctx->plan.clear(); ctx->order.clear(); ctx->results.clear(); ctx->call_next = 0; ctx->batch_begin = ctx->batch_end = 0; ctx->fly_begin = ctx->fly_end = 0;
    std::vector<JobHost> jh;
    for (int c = 0; c < n_nb; ++c)
        if (nb[c].to_be_matched) { JobHost j; j.src_view = view; j.tgt_view = nb[c].view; memcpy(j.F, nb[c].F, sizeof j.F); jh.push_back(j); }
    PhaseASet& set = ctx->pa[ctx->cur];
    set.background = false;
    ctx->want_records = copy_back;
    if (run_phase_a(ctx, &set, jh)) { ctx->err = set.err; return 1; }
    if (mr_job_totals(ctx, set)) return 1;
    ctx->launches += set.launches; set.launches = 0;
    GroupView one = {view, n_nb, nb, 0, 0};
    if (ctx->world > 1 && mr_ensure_window(ctx, mr_group_need(ctx, &one, 1))) return 1;
    if (enqueue_group(ctx, ctx->grp[0], &one, 1, sigma_p, sigma_a)) return 1;
    if (finish_group(ctx, ctx->grp[0], copy_back, res)) return 1;
    if (ctx->copies_pending) { L3D_CUDA(cudaStreamSynchronize(ctx->copy_stream)); ctx->copies_pending = false; }
    return 0;
}

// Nothing throws across the C ABI: the host-side containers can (out of memory).
template <typename F>
int guarded(l3d_ctx* ctx, F f)
{
    try { return f(); }
    catch (const std::exception& e) { if (ctx) ctx->err = std::string("exception: ") + e.what(); return 1; }
    catch (...) { if (ctx) ctx->err = "unknown exception"; return 1; }
}

}  // namespace
