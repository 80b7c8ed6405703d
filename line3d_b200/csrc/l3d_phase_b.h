// line3d_b200/csrc/l3d_phase_b.h -- host side of phase B: enqueue_group / finish_group (one launch of every per-view kernel per group of views), the multi-GPU record windows.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

// Phase B of a group of views (all of one dependency level, see VArgs), split in two so that the next group can be
// put on the stream before the host waits for this one: enqueue_group does the host's bookkeeping for every view,
// uploads the group's tables and arguments once and launches every kernel once (sizes come from upper bounds, the
// real counts stay on the device); finish_group waits for the group's result records and fetches the views' output.
struct GroupView { int view; int n_nb; const l3d_neighbor* nb; int first_job; size_t plan_index; };

// Several GPUs on one scene: how many records the views of a group can put out at most, all ranks together (what the
// record window has to hold) -- the same bound enqueue_group computes, without touching anything.
size_t mr_group_need(l3d_ctx* ctx, const GroupView* gv, int count)
{
    PhaseASet& pa = ctx->pa[ctx->cur];
    size_t need = MR_HEADER_BYTES;
    for (int g = 0; g < count; ++g) {
        if (gv[g].view < 0 || gv[g].view >= ctx->V) continue;
        int n_slots = 0; size_t f_ub = 0;
        for (int c = 0; c < gv[g].n_nb; ++c) {
            const l3d_neighbor& n = gv[g].nb[c];
            if (n.to_be_matched) { const size_t j = size_t(gv[g].first_job + n_slots++); if (j < pa.job_raw_all.size()) f_ub += size_t(pa.job_raw_all[j]); }
            else {
                const std::vector<Delivered>& dl = ctx->delivered[gv[g].view];
                for (size_t k = 0; k < dl.size(); ++k) if (dl[k].from_view == n.view) { f_ub += size_t(dl[k].count_all); break; }
            }
        }
        need += (size_t(ctx->views[gv[g].view].n_segments) * 8 + 255) & ~size_t(255);
        need += 2 * size_t(ctx->world) * std::max<size_t>(f_ub, 1) * sizeof(l3d_match);
    }
    return need;
}

int enqueue_group(l3d_ctx* ctx, ViewGroup& gr, const GroupView* gv, int count, float sigma_p, float sigma_a)
{
    cudaStream_t st = ctx->stream;
    PhaseASet& pa = ctx->pa[ctx->cur];
    const bool mr = ctx->world > 1;
    gr.count = count;
    L3D_CUDA(gr.h_args.ensure(VG_MAX)); L3D_CUDA(gr.d_args.ensure(VG_MAX));
    if (mr) {
        L3D_CUDA(gr.h_allrec.ensure(size_t(VG_MAX) * MR_MAX_RANKS));
        // the records of the group two groups ago may still be on their way to the host out of this group's buffers
        if (gr.ev_x[1]) L3D_CUDA(cudaStreamWaitEvent(st, gr.ev_x[1], 0));
    }
    int res_stride = 8;
    for (int g = 0; g < count; ++g) { int nd = 0; for (int c = 0; c < gv[g].n_nb; ++c) nd += gv[g].nb[c].deliver ? 1 : 0; res_stride = std::max(res_stride, 8 + nd); }
    gr.res_stride = res_stride;
    L3D_CUDA(gr.d_vres.ensure(size_t(VG_MAX) * res_stride)); L3D_CUDA(gr.h_vres.ensure(size_t(VG_MAX) * res_stride));
    size_t cams_total = 0;
    for (int g = 0; g < count; ++g) {
        if (gv[g].view < 0 || gv[g].view >= ctx->V) { ctx->err = "l3d_match_view: view index out of range"; return 1; }
        for (int c = 0; c < gv[g].n_nb; ++c) {
            if (gv[g].nb[c].view < 0 || gv[g].nb[c].view >= ctx->V || gv[g].nb[c].view == gv[g].view) { ctx->err = "l3d_match_view: bad neighbour view index"; return 1; }
            if (c > 0 && gv[g].nb[c - 1].view >= gv[g].nb[c].view) { ctx->err = "l3d_match_view: neighbours must be in ascending view order"; return 1; }
        }
        gr.cam_off[g] = cams_total; cams_total += size_t(gv[g].n_nb);
    }
    gr.cam_off[count] = cams_total;
    L3D_CUDA(gr.h_cams.ensure(cams_total)); L3D_CUDA(gr.d_cams.ensure(cams_total));

    // ---- per view: tables, bounds, buffers, hand-over records, kernel arguments ------------------------------
    size_t zero_total = 0;
    int mx_src = 1, mx_cams = 1, mx_items = 1, mx_cta = 1, mn_sc_h = SC_T; long mx_warps = 1; size_t mx_M = 0, mx_smem = 0;
    for (int g = 0; g < count; ++g) mn_sc_h = std::min(mn_sc_h, gv[g].n_nb <= 192 ? 128 : (gv[g].n_nb <= 384 ? 64 : 32));
    if (ctx->score_block) mn_sc_h = std::min(mn_sc_h, ctx->score_block);
    bool any_scored = false, any_window = false, any_quadratic = false, any_deliver = false;
    struct Z { size_t items, iters, best, hist, cursor, end; } z[VG_MAX];
    size_t hist_total = 0, window_need = MR_HEADER_BYTES;
    const int slot_q = int(&gr - ctx->grp);
    memset(&gr.lay, 0, sizeof gr.lay);
    for (int g = 0; g < count; ++g) {
        const int view = gv[g].view, n_nb = gv[g].n_nb; const l3d_neighbor* nb = gv[g].nb; const int first_job = gv[g].first_job;
        ViewSlot& sl = gr.slot[g];
        ViewJob& job = sl.job;
        sl.plan_index = gv[g].plan_index;
        const l3d_view& sv = ctx->views[view];
        const int n_full = sv.n_segments;
        const int n_src = local_rows(ctx, n_full);         // the rows of the view this process works on
        CamTable* hc = gr.h_cams.p + gr.cam_off[g];
        job.view = view; job.nb.assign(nb, nb + n_nb); job.sigma_p = sigma_p; job.sigma_a = sigma_a;
        job.consumed.clear(); job.owner = -1;
        int n_slots = 0, n_deliver = 0, max_deliver_seg = 1;
        uint64_t pair_tests = 0, cand_ub = 0, raw_ub = 0, raw_ub_all = 0, delivered_ub = 0, delivered_ub_all = 0;      // raw_ub_all: over all ranks' rows
        size_t seg_ub = 0;                               // upper bound on the matches of one source segment
        std::vector<int> consumed;                       // indices into ctx->delivered[view]
        for (int c = 0; c < n_nb; ++c) {
            const l3d_neighbor& n = nb[c];
            const l3d_view& tv = ctx->views[n.view];
            CamTable& ct = hc[c];
            memcpy(ct.F, n.F, sizeof ct.F); memcpy(ct.RtKinv, tv.RtKinv, sizeof ct.RtKinv);
            memcpy(ct.P, tv.P, sizeof ct.P); memcpy(ct.C, tv.C, sizeof ct.C);
            ct.seg_off = ctx->seg_off[n.view]; ct.n_seg = tv.n_segments;
            ct.t_start = nullptr; ct.t_tgt = nullptr; ct.t_dep = nullptr; ct.slot = -1; ct.deliver = -1; ct.row_off = 0;
            if (n.to_be_matched) {
                const int j = first_job + n_slots;
                ct.mode = 0; ct.slot = n_slots++; ct.row_off = pa.jobs_host[j].row_off;
                cand_ub += pa.job_cands[j]; raw_ub += pa.job_raw[j]; raw_ub_all += pa.job_raw_all[j]; seg_ub += size_t(pa.job_raw_rowmax[j]);
                pair_tests += uint64_t(n_src) * uint64_t(tv.n_segments);
            } else {
                ct.mode = 2;
                std::vector<Delivered>& dl = ctx->delivered[view];
                for (size_t k = 0; k < dl.size(); ++k)
                    if (dl[k].from_view == n.view) {
                        ct.mode = 1; ct.t_start = dl[k].t_start; ct.t_tgt = dl[k].t_tgt; ct.t_dep = dl[k].t_dep;
                        delivered_ub += uint64_t(dl[k].count); delivered_ub_all += uint64_t(dl[k].count_all); seg_ub += size_t(dl[k].max_group);
                        consumed.push_back(int(k)); job.consumed.push_back(std::make_pair(dl[k].owner, dl[k].slot));
                        break;
                    }
            }
            if (n.deliver) { ct.deliver = n_deliver++; max_deliver_seg = std::max(max_deliver_seg, tv.n_segments); }
        }
        job.pair_tests = pair_tests; job.cand_ub = cand_ub;
        const size_t M_ub = size_t(raw_ub + delivered_ub);          // upper bound on the number of matches to score
        const size_t F_ub_all = size_t(raw_ub_all + (mr ? delivered_ub_all : delivered_ub));  // ... and on the output records of any rank (the same number on every rank)
        if (mr) {                                                    // the view's places in the group's window
            gr.lay.samp_off[g] = (long long)window_need; window_need += (size_t(n_full) * 8 + 255) & ~size_t(255);
            gr.lay.cap[g] = (long long)std::max<size_t>(F_ub_all, 1);
            gr.lay.rec_off[g] = (long long)window_need; window_need += size_t(ctx->world) * size_t(gr.lay.cap[g]) * sizeof(l3d_match);
            gr.lay.full_off[g] = (long long)window_need; window_need += size_t(ctx->world) * size_t(gr.lay.cap[g]) * sizeof(l3d_match);
        }
        const bool scored = n_slots > 0;                             // cudawrapper.cu:877-878
        const bool delivering = n_deliver > 0 && scored;
        const int stride = max_deliver_seg + 1;                      // one spare key per slot keeps t_start[n_seg] addressable
        const int n_keys = delivering ? n_deliver * stride : 0;
        const bool windowed = sv.spatial_k > 0.0f;
        const int SORT_MAX = 1 << 20;                                // the window kernel's index has 20 bits
        const bool quadratic = scored && (!windowed || seg_ub > size_t(SORT_MAX));
        // hypotheses per scoring block: the per-(camera, hypothesis) maxima must fit in shared memory whatever the number
        // of neighbours (k = -1: every co-visible view)
        const int sc_h = mn_sc_h;                        // (one launch covers the group: the same for all its views)
        const int max_items = int(M_ub / size_t(sc_h)) + n_src + 1;
        const int n_cta = int(std::max<size_t>((M_ub + RM_T - 1) / RM_T, 1));
        const size_t sm_bytes = size_t(n_nb) * size_t(sc_h) * sizeof(int);
        if (sm_bytes > 200 * 1024) { ctx->err = "l3d_match_view: more than 1600 neighbours in one view"; return 1; }
        // buffers of this view
        L3D_CUDA(sl.d_cnt.ensure(size_t(n_src) * n_nb + 1)); L3D_CUDA(sl.d_off.ensure(size_t(n_src) * n_nb + 2));
        L3D_CUDA(sl.d_mids.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_mdep.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_score.ensure(std::max<size_t>(M_ub, 1)));
        L3D_CUDA(sl.d_filtered.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_cta.ensure(size_t(n_cta) + 1)); L3D_CUDA(sl.d_F.ensure(4));
        if (!mr) L3D_CUDA(sl.d_dsamples.ensure(size_t(n_src) * 2 + 2));
        L3D_CUDA(sl.d_items.ensure(size_t(2 * max_items)));
        if (scored && windowed) {
            L3D_CUDA(sl.d_sk.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_sd2.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_spl.ensure(std::max<size_t>(M_ub, 1)));
            L3D_CUDA(sl.d_boff.ensure(size_t(n_src) * (BK_B + 1))); L3D_CUDA(sl.d_brange.ensure(size_t(std::max(n_src, 1))));
        }
        // what has to start from zero: one region per view in the group's zero buffer, then the hand-over group sizes of
        // all views (one block: with several GPUs on one scene they are summed over the ranks in one all-reduce)
        z[g].items = zero_total; z[g].iters = z[g].items + 4; z[g].best = z[g].iters + 4;
        z[g].cursor = z[g].best + 2 * size_t(n_src) + 2; z[g].end = (z[g].cursor + size_t(n_keys) + 2 + 3) & ~size_t(3);
        zero_total = z[g].end;
        z[g].hist = hist_total; hist_total += (size_t(n_keys) + 2 + 3) & ~size_t(3);
        // hand-over buffers
        DeliverBuf buf;
        if (delivering) {
            buf.refs = n_deliver;
            L3D_CUDA(ctx->pool.get(sizeof(int) * (size_t(n_keys) + 2), reinterpret_cast<void**>(&buf.start)));
            // (several GPUs on one scene: every rank holds the complete hand-over lists, all ranks' records)
            L3D_CUDA(ctx->pool.get(sizeof(int) * std::max<size_t>(mr ? F_ub_all : M_ub, 1), reinterpret_cast<void**>(&buf.tgt)));
            L3D_CUDA(ctx->pool.get(sizeof(float4) * std::max<size_t>(mr ? F_ub_all : M_ub, 1), reinterpret_cast<void**>(&buf.dep)));
        }
        job.M_ub = M_ub; job.scored = scored; job.delivering = delivering; job.n_deliver = n_deliver;
        // kernel arguments (pointers into the zero buffer are filled in below, when it has its size)
        VArgs& va = gr.h_args.p[g];
        memset(&va, 0, sizeof va);
        va.cams = gr.d_cams.p + gr.cam_off[g]; va.n_cams = n_nb; va.n_src = n_src;
        memcpy(va.src.RtKinv, sv.RtKinv, sizeof va.src.RtKinv); memcpy(va.src.C, sv.C, sizeof va.src.C);
        va.src.spatial_k = sv.spatial_k; va.src.cmax = std::max(std::fabs(sv.C[0]), std::max(std::fabs(sv.C[1]), std::fabs(sv.C[2])));
        va.src.seg_off = ctx->seg_off[view]; va.src.n_seg = n_src;
        va.src.row_first = mr ? ctx->rank : 0; va.src.row_stride = mr ? ctx->world : 1;
        va.mr = mr ? 1 : 0; va.n_seg_full = n_full;
        va.sp.sigma_p = sigma_p; va.sp.sigma_a = sigma_a;
        va.raw_cnt = pa.d_pa_raw.p; va.cand_off = pa.d_pa_off.p; va.stage_col = pa.d_stage_col.p; va.stage_dep = pa.d_stage_dep.p; va.stage_ok = pa.d_stage_ok.p;
        va.cnt = sl.d_cnt.p; va.off = sl.d_off.p; va.mids = sl.d_mids.p; va.mdep = sl.d_mdep.p; va.score = sl.d_score.p;
        va.sk = sl.d_sk.p; va.sd2 = sl.d_sd2.p; va.spl = sl.d_spl.p; va.boff = sl.d_boff.p; va.brange = sl.d_brange.p; va.items = sl.d_items.p; va.max_items = max_items;
        va.sc_h = sc_h;
        va.scored = scored ? 1 : 0; va.windowed = windowed ? 1 : 0; va.quadratic = quadratic ? 1 : 0; va.m_hi = SORT_MAX;
        va.cta_kept = sl.d_cta.p; va.n_cta = n_cta; va.Fout = sl.d_F.p; va.filtered = sl.d_filtered.p; va.dsamples = sl.d_dsamples.p;
        va.delivering = delivering ? 1 : 0; va.stride = stride; va.n_keys = n_keys; va.n_deliver = n_deliver;
        va.dl_start = buf.start; va.dl_tgt = buf.tgt; va.dl_dep = buf.dep;
        va.vres = gr.d_vres.p + size_t(g) * res_stride;
        // the hand-over records exist from now on (stream order); until this view is collected their sizes are bounds
        if (delivering) {
            job.owner = int(ctx->deliver_bufs.size());
            buf.counts.assign(size_t(n_deliver), 0); buf.maxg.assign(size_t(n_deliver), 0);
            for (int c = 0; c < n_nb; ++c) {
                const int d = hc[c].deliver;
                if (d < 0) continue;
                const uint64_t ub = std::min<uint64_t>(hc[c].slot >= 0 ? pa.job_raw_all[first_job + hc[c].slot] : uint64_t(mr ? F_ub_all : M_ub), 0x7fffffffull);
                Delivered dl; dl.from_view = view; dl.owner = job.owner; dl.slot = d;
                dl.count = int(ub); dl.max_group = int(std::min<uint64_t>(ub, uint64_t(n_src)));      // a source segment matches a target segment once
                dl.count_all = (long long)ub;
                if (mr) dl.max_group = int(std::min<uint64_t>(ub, uint64_t(n_full)));
                dl.t_start = buf.start + size_t(d) * stride; dl.t_tgt = buf.tgt; dl.t_dep = buf.dep;
                buf.counts[d] = dl.count; buf.maxg[d] = dl.max_group;
                ctx->delivered[nb[c].view].push_back(dl);
            }
            ctx->deliver_bufs.push_back(buf);
        }
        // the consumed deliveries leave the list; their memory is given back when this view has been collected
        {
            std::vector<Delivered>& dl = ctx->delivered[view];
            for (size_t k = 0; k < consumed.size(); ++k) dl[consumed[k]].from_view = -1;
            std::vector<Delivered> keep;
            for (size_t k = 0; k < dl.size(); ++k) if (dl[k].from_view >= 0) keep.push_back(dl[k]);
            dl.swap(keep);
        }
        job.active = true;
        mx_src = std::max(mx_src, n_src); mx_cams = std::max(mx_cams, n_nb); mx_warps = std::max(mx_warps, long(n_src) * n_nb);
        mx_M = std::max(mx_M, M_ub); mx_cta = std::max(mx_cta, n_cta);
        if (scored) { any_scored = true; mx_items = std::max(mx_items, max_items); mx_smem = std::max(mx_smem, sm_bytes); }
        any_window |= scored && windowed; any_quadratic |= quadratic; any_deliver |= delivering;
    }
    gr.hist_off = zero_total; gr.hist_total = hist_total;
    zero_total += hist_total;
    const size_t routed_off = zero_total;                     // several GPUs: records routed per (view, destination rank), and a fault word
    zero_total += size_t(VG_MAX) * MR_MAX_RANKS + 4;
    L3D_CUDA(gr.d_zero.ensure(zero_total + 4));
    if (mr && window_need > ctx->win_cap) { ctx->err = "l3d_match_view: internal error (window too small)"; return 1; }
    for (int g = 0; g < count; ++g) {
        VArgs& va = gr.h_args.p[g];
        va.nitems = gr.d_zero.p + z[g].items; va.iters = reinterpret_cast<unsigned long long*>(gr.d_zero.p + z[g].iters);
        va.best = reinterpret_cast<unsigned long long*>(gr.d_zero.p + z[g].best); va.hist = gr.d_zero.p + gr.hist_off + z[g].hist; va.cursor = gr.d_zero.p + z[g].cursor;
        if (mr) va.dsamples = reinterpret_cast<float*>(ctx->win[slot_q] + gr.lay.samp_off[g]);
    }
    if (mx_smem > ctx->score_smem) {
        L3D_CUDA(cudaFuncSetAttribute(k_score, cudaFuncAttributeMaxDynamicSharedMemorySize, int(mx_smem)));
        L3D_CUDA(cudaFuncSetAttribute(k_score_window, cudaFuncAttributeMaxDynamicSharedMemorySize, int(mx_smem)));
        ctx->score_smem = mx_smem;
    }

    // ---- one upload, one memset, one launch of every kernel for the whole group ------------------------------
    const unsigned G = unsigned(count);
    const VArgs* A = gr.d_args.p;
    L3D_CUDA(cudaMemcpyAsync(gr.d_cams.p, gr.h_cams.p, sizeof(CamTable) * cams_total, cudaMemcpyHostToDevice, st));
    L3D_CUDA(cudaMemcpyAsync(gr.d_args.p, gr.h_args.p, sizeof(VArgs) * size_t(count), cudaMemcpyHostToDevice, st));
    L3D_CUDA(cudaMemsetAsync(gr.d_zero.p, 0, sizeof(int) * zero_total, st));
    gr.n_kev = 0;
    // (per-kernel events, only when asked for: l3d_debug_chain_times)
#define L3D_KEV(slot) do { if (ctx->chain_prof && ctx->stage_events && gr.n_kev < 24) { if (!gr.kev[gr.n_kev]) cudaEventCreate(&gr.kev[gr.n_kev]); cudaEventRecord(gr.kev[gr.n_kev], st); gr.kev_slot[gr.n_kev] = (slot); ++gr.n_kev; } } while (0)
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[2], st));
    L3D_KEV(0);
    L3D_CUDA(pdl_launch(k_gather_counts, dim3((mx_src + 255) / 256, mx_cams, G), dim3(256), 0, st, A));
    L3D_KEV(1);
    L3D_CUDA(pdl_launch(k_scan_views, dim3(1, G), dim3(SCAN_T), 0, st, A, 0));
    L3D_KEV(2);
    L3D_CUDA(pdl_launch(k_fill, dim3(unsigned((mx_warps * 32 + 255) / 256), G), dim3(256), 0, st, A, ctx->d_segs.p));
    ctx->launches += 3;
    L3D_KEV(3);
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[3], st));
    if (any_window) {
        L3D_CUDA(pdl_launch(k_bucket_segments, dim3(mx_src, G), dim3(BK_T), 0, st, A));
        L3D_KEV(4);
        L3D_CUDA(pdl_launch(k_score_window, dim3(mx_items, G), dim3(mn_sc_h), mx_smem, st, A, ctx->d_segs.p));
        ctx->launches += 2;
    }
    if (any_quadratic) {
        L3D_CUDA(pdl_launch(k_score_items, dim3((mx_src + 255) / 256, G), dim3(256), 0, st, A));
        L3D_CUDA(pdl_launch(k_score, dim3(mx_items, G), dim3(mn_sc_h), mx_smem, st, A, ctx->d_segs.p));
        ctx->launches += 2;
    }
    L3D_KEV(5);
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[4], st));
    if (any_scored) { L3D_CUDA(pdl_launch(k_reduce_matches, dim3(mx_cta, G), dim3(RM_T), 0, st, A)); ctx->launches += 1; }
    L3D_KEV(6);
    L3D_CUDA(pdl_launch(k_filter_compact, dim3(mx_cta, G), dim3(RM_T), 0, st, A));
    ctx->launches += 1;
    L3D_KEV(7);
    if (mr) {
        // Every rank stores its records, depth samples and counts into all ranks' windows (regions fixed by shared
        // bounds), a one-word all-reduce tells that everybody's stores have landed, and the hand-over lists and the
        // medians are made from the own window.
        const unsigned pb = unsigned(std::min<size_t>(std::max<size_t>((mx_M + 255) / 256, 1), size_t(4 * ctx->n_sm)));
        const bool collect = ctx->want_records;
        const int root = !collect ? -1 : (ctx->result_root < 0 ? -2 : ctx->result_root);
        int* routed = gr.d_zero.p + routed_off; int* fault = routed + VG_MAX * MR_MAX_RANKS;
        const unsigned seq = ++ctx->win_seq[slot_q];
        const unsigned char* window = ctx->win[slot_q];
        L3D_CUDA(pdl_launch(k_mr_push, dim3(pb, G), dim3(256), 0, st, A, gr.lay, ctx->peers[slot_q], ctx->rank, ctx->world, root, routed));
        L3D_KEV(8);
        L3D_CUDA(pdl_launch(k_mr_signal, dim3(1), dim3(VG_MAX * MR_MAX_RANKS), 0, st, A, int(count), ctx->peers[slot_q], ctx->rank, ctx->world, (const int*)routed, seq));
        L3D_CUDA(pdl_launch(k_mr_wait, dim3(1), dim3(32), 0, st, window, ctx->world, seq, fault));
        ctx->launches += 3;
        L3D_KEV(9);
        gr.d_fault = fault;
        if (any_deliver) {
            const unsigned hb = unsigned(std::min<size_t>(std::max<size_t>(pb, 1), size_t(4 * ctx->n_sm)));
            L3D_CUDA(pdl_launch(k_mr_hist, dim3(hb, G), dim3(256), 0, st, A, window, gr.lay, ctx->world));
            L3D_KEV(10);
            L3D_CUDA(pdl_launch(k_scan_views, dim3(1, G), dim3(SCAN_T), 0, st, A, 1));
            L3D_KEV(11);
            L3D_CUDA(pdl_launch(k_mr_deliver_scatter, dim3(hb, G), dim3(256), 0, st, A, window, gr.lay, ctx->world));
            ctx->launches += 3;
            L3D_KEV(12);
        }
        L3D_CUDA(pdl_launch(k_finish_view, dim3(1, G), dim3(MS_T), 0, st, A));
        ctx->launches += 1;
        L3D_KEV(13);
        if (collect && (ctx->result_root < 0 || ctx->result_root == ctx->rank)) {
            // for the copy to the host: the ranks' regions of each view put end to end.  On this stream, so that the
            // window is only ever read in stream order (the other ranks write the group after next into it as soon as
            // this rank has joined the next group's all-reduce).
            size_t bound = 0;
            for (int g = 0; g < count; ++g) bound += size_t(gr.lay.cap[g]);
            L3D_CUDA(gr.d_gather.ensure(bound));
            k_mr_gather<<<dim3(std::min(2 * pb, unsigned(4 * ctx->n_sm)), G), 256, 0, st>>>(window, gr.lay, ctx->world, gr.d_gather.p);
            ctx->launches += 1;
            L3D_KEV(14);
        }
        L3D_CUDA(cudaEventRecord(gr.ev_x[0], st));      // the gathered records are complete
        L3D_CUDA(cudaGetLastError());
        if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[5], st));
        L3D_CUDA(cudaMemcpyAsync(gr.h_vres.p, gr.d_vres.p, sizeof(int) * size_t(count) * res_stride, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaMemcpyAsync(gr.h_allrec.p, window, sizeof(RankRec) * VG_MAX * size_t(ctx->world), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaMemcpyAsync(&gr.h_allrec.p[size_t(VG_MAX) * MR_MAX_RANKS - 1].pad, gr.d_fault, sizeof(int), cudaMemcpyDeviceToHost, st));      // (a spare word of the pinned table)
        L3D_CUDA(cudaEventRecord(gr.ev[0], st));
        return 0;
    }
    if (any_deliver) {
        L3D_CUDA(pdl_launch(k_scan_views, dim3(1, G), dim3(SCAN_T), 0, st, A, 1));
        L3D_CUDA(pdl_launch(k_deliver_scatter, dim3(unsigned((std::max<size_t>(mx_M, 1) + 255) / 256), G), dim3(256), 0, st, A));
        ctx->launches += 2;
        L3D_KEV(12);
    }
    // The last kernel (medians, result records) and the download of the records are needed by the host only: the next
    // group does not wait for them.  They go to a side stream (this group's buffers are not touched again before the
    // host has collected it), except when stages are being timed.
    cudaStream_t fs = st;
    if (!ctx->stage_events) {
        fs = ctx->fin_stream;
        L3D_CUDA(cudaEventRecord(gr.ev[1], st));
        L3D_CUDA(cudaStreamWaitEvent(fs, gr.ev[1], 0));
    }
    L3D_CUDA(pdl_launch(k_finish_view, dim3(1, G), dim3(MS_T), 0, fs, A));
    ctx->launches += 1;
    L3D_KEV(13);
    L3D_CUDA(cudaGetLastError());
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[5], st));
    L3D_CUDA(cudaMemcpyAsync(gr.h_vres.p, gr.d_vres.p, sizeof(int) * size_t(count) * res_stride, cudaMemcpyDeviceToHost, fs));
    L3D_CUDA(cudaEventRecord(gr.ev[0], fs));
    return 0;
}

// n records of pinned host memory that stay where they are until l3d_reset_matching
int arena_take(l3d_ctx* ctx, size_t n, l3d_match** out)
{
    while (ctx->arena_chunk < ctx->arena.size() && ctx->arena_used + n > ctx->arena[ctx->arena_chunk].second) { ++ctx->arena_chunk; ctx->arena_used = 0; }
    if (ctx->arena_chunk == ctx->arena.size()) {
        const size_t cap = std::max<size_t>(n, size_t(4) << 20);          // 128 MiB chunks
        l3d_match* p = nullptr;
        L3D_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&p), cap * sizeof(l3d_match), cudaHostAllocDefault));
        ctx->arena.push_back(std::make_pair(p, cap)); ctx->arena_used = 0;
    }
    *out = ctx->arena[ctx->arena_chunk].first + ctx->arena_used;
    ctx->arena_used += n;
    return 0;
}

// Several GPUs on one scene: the record windows.  Every rank allocates two (one per group in flight), exports them
// (CUDA IPC) and maps every other rank's; k_mr_push then stores a group's records into all of them.  Called by all
// ranks at the same point with the same size, with no group in flight.
int mr_ensure_window(l3d_ctx* ctx, size_t need)
{
    if (need <= ctx->win_cap) return 0;
    cudaStream_t st = ctx->stream;
    const int W = ctx->world;
    L3D_CUDA(ctx->d_bar.ensure(4)); L3D_CUDA(ctx->d_ipc.ensure(size_t(2) * sizeof(cudaIpcMemHandle_t) * W)); L3D_CUDA(ctx->h_ipc.ensure(size_t(2) * sizeof(cudaIpcMemHandle_t) * W));
    L3D_CUDA(cudaStreamSynchronize(st));
    for (int q = 0; q < 2; ++q)
        for (int r = 0; r < W; ++r) { if (r != ctx->rank && ctx->peers[q].base[r]) cudaIpcCloseMemHandle(ctx->peers[q].base[r]); ctx->peers[q].base[r] = nullptr; }
    // nobody frees a window that another rank still has mapped
    L3D_CUDA(cudaMemsetAsync(ctx->d_bar.p, 0, sizeof(int), st));
    L3D_NCCL(ctx->nccl->AllReduce(ctx->d_bar.p, ctx->d_bar.p, 1, ncclInt, ncclSum, ctx->comm, st));
    L3D_CUDA(cudaStreamSynchronize(st));
    const size_t cap = std::max<size_t>(need + need / 4, size_t(64) << 20);      // bytes; at least 64 MiB
    cudaIpcMemHandle_t mine[2];
    for (int q = 0; q < 2; ++q) {
        if (ctx->win[q]) cudaFree(ctx->win[q]);
        ctx->win[q] = nullptr;
        L3D_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->win[q]), cap));
        L3D_CUDA(cudaMemsetAsync(ctx->win[q], 0, MR_HEADER_BYTES, st));          // the arrival flags start below every sequence number
        L3D_CUDA(cudaIpcGetMemHandle(&mine[q], ctx->win[q]));
    }
    ctx->win_cap = 0;
    unsigned char* d_send = ctx->d_ipc.p; unsigned char* d_recv = ctx->d_ipc.p;      // in place: each rank's part at its own offset
    L3D_CUDA(cudaMemcpyAsync(d_send + size_t(ctx->rank) * sizeof mine, mine, sizeof mine, cudaMemcpyHostToDevice, st));
    L3D_NCCL(ctx->nccl->AllGather(d_send + size_t(ctx->rank) * sizeof mine, d_recv, sizeof mine, ncclChar, ctx->comm, st));
    L3D_CUDA(cudaMemcpyAsync(ctx->h_ipc.p, d_recv, sizeof mine * size_t(W), cudaMemcpyDeviceToHost, st));
    L3D_CUDA(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r)
        for (int q = 0; q < 2; ++q) {
            if (r == ctx->rank) { ctx->peers[q].base[r] = ctx->win[q]; continue; }
            cudaIpcMemHandle_t h; memcpy(&h, ctx->h_ipc.p + (size_t(r) * 2 + q) * sizeof h, sizeof h);
            void* ptr = nullptr;
            L3D_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
            ctx->peers[q].base[r] = static_cast<unsigned char*>(ptr);
        }
    ctx->win_cap = cap;
    return 0;
}

void mr_release_window(l3d_ctx* ctx)
{
    for (int q = 0; q < 2; ++q) {
        for (int r = 0; r < MR_MAX_RANKS; ++r) { if (r != ctx->rank && ctx->peers[q].base[r]) cudaIpcCloseMemHandle(ctx->peers[q].base[r]); ctx->peers[q].base[r] = nullptr; }
    }
    if (ctx->comm && ctx->d_bar.p && !ctx->comm_broken) {        // nobody frees a window that another rank still has mapped
        cudaMemsetAsync(ctx->d_bar.p, 0, sizeof(int), ctx->stream);
        ctx->nccl->AllReduce(ctx->d_bar.p, ctx->d_bar.p, 1, ncclInt, ncclSum, ctx->comm, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
    }
    for (int q = 0; q < 2; ++q) { if (ctx->win[q]) cudaFree(ctx->win[q]); ctx->win[q] = nullptr; }
    ctx->win_cap = 0;
}

// Raw matches per job summed over the ranks: the bounds every rank sizes the views' buffers with must cover what the
// other ranks may hand over.  One small all-reduce per batch of pair tests, on the main stream.
int mr_job_totals(l3d_ctx* ctx, PhaseASet& pa)
{
    const size_t n = pa.job_raw.size();
    pa.job_raw_all = pa.job_raw;
    if (ctx->world <= 1 || n == 0) return 0;
    L3D_CUDA(ctx->d_tot.ensure(n)); L3D_CUDA(ctx->h_tot.ensure(n));
    for (size_t j = 0; j < n; ++j) ctx->h_tot.p[j] = pa.job_raw[j];
    L3D_CUDA(cudaMemcpyAsync(ctx->d_tot.p, ctx->h_tot.p, sizeof(unsigned long long) * n, cudaMemcpyHostToDevice, ctx->stream));
    L3D_NCCL(ctx->nccl->AllReduce(ctx->d_tot.p, ctx->d_tot.p, n, ncclUint64, ncclSum, ctx->comm, ctx->stream));
    L3D_CUDA(cudaMemcpyAsync(ctx->h_tot.p, ctx->d_tot.p, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost, ctx->stream));
    L3D_CUDA(cudaStreamSynchronize(ctx->stream));
    for (size_t j = 0; j < n; ++j) pa.job_raw_all[j] = ctx->h_tot.p[j];
    return 0;
}

// Waits for a group, turns its result records into l3d_match_results (results[plan index], or *single for a view
// matched outside a schedule) and fetches the output records of all its views with one synchronise.
int finish_group(l3d_ctx* ctx, ViewGroup& gr, bool copy_back, l3d_match_result* single)
{
    const auto t_in = std::chrono::steady_clock::now();
    L3D_CUDA(cudaEventSynchronize(gr.ev[0]));
    const auto t_waited = std::chrono::steady_clock::now();
    ctx->host_s[1] += std::chrono::duration<double>(t_waited - t_in).count();
    struct Tail { l3d_ctx* c; std::chrono::steady_clock::time_point t; ~Tail() { c->host_s[2] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t).count(); } } tail = {ctx, t_waited};
    PhaseASet& pa = ctx->pa[ctx->cur];
    float ms_a[2] = {0.f, 0.f};
    if (pa.pa_ms_pending) {                            // phase A of a whole batch is booked on the first view collected after it
        cudaEventElapsedTime(&ms_a[0], pa.pa_ev[0], pa.pa_ev[1]); cudaEventElapsedTime(&ms_a[1], pa.pa_ev[1], pa.pa_ev[2]);
        for (int k = 0; k < 2; ++k) { float ms = 0.f; if (pa.pa_kev_set[k]) cudaEventElapsedTime(&ms, pa.pa_kev[2 * k], pa.pa_kev[2 * k + 1]); ctx->pa_kernel_ms[k] = ms; }
        { float ms = 0.f; if (pa.pa_kev_set[0]) cudaEventElapsedTime(&ms, pa.pa_kev[1], pa.pa_kev[4]); ctx->pa_kernel_ms[2] = ms; }
        pa.pa_ms_pending = false;
    } else ctx->pa_kernel_ms[0] = ctx->pa_kernel_ms[1] = ctx->pa_kernel_ms[2] = 0.f;
    float ms_b[3] = {0.f, 0.f, 0.f};
    if (ctx->stage_events) for (int k = 0; k < 3; ++k) cudaEventElapsedTime(&ms_b[k], gr.ev[k + 2], gr.ev[k + 3]);
    for (int k = 1; k < gr.n_kev; ++k) { float ms = 0.f; cudaEventElapsedTime(&ms, gr.kev[k - 1], gr.kev[k]); ctx->chain_ms[gr.kev_slot[k]] += double(ms); }
    gr.n_kev = 0;
    const bool mr = ctx->world > 1;
    int view_off[VG_MAX], view_cnt[VG_MAX];              // several GPUs: where each view's records are in this group's window
    if (mr) {
        long long total = 0; int max_cnt = 0;
        for (int v = 0; v < gr.count; ++v) {
            view_off[v] = int(total);
            for (int r = 0; r < ctx->world; ++r) {
                const int f = gr.h_allrec.p[size_t(r) * VG_MAX + v].F;
                if (f < 0 || (long long)f > gr.lay.cap[v]) { ctx->err = "l3d_match_view: internal error (record window overrun)"; return 1; }
                total += f; max_cnt = std::max(max_cnt, f);
            }
            view_cnt[v] = int(total - view_off[v]);
        }
        if (total > 0x7fffffffll) { ctx->err = "l3d_match_view: too many output records in one group of views"; return 1; }
        if (gr.h_allrec.p[size_t(VG_MAX) * MR_MAX_RANKS - 1].pad != 0) { ctx->err = "l3d_match_view: a rank did not deliver its records of a group of views within ten seconds"; return 1; }
        copy_back = copy_back && (ctx->result_root < 0 || ctx->result_root == ctx->rank);      // who gets the records on the host
        (void)max_cnt;
        if (copy_back && total > 0) L3D_CUDA(cudaStreamWaitEvent(ctx->copy_stream, gr.ev_x[0], 0));
    }
    bool copies = false;
    for (int g = 0; g < gr.count; ++g) {
        ViewSlot& sl = gr.slot[g];
        ViewJob& job = sl.job;
        job.active = false;
        const int* hr = gr.h_vres.p + size_t(g) * gr.res_stride;
        l3d_match_result* res = single ? single : &ctx->results[sl.plan_index].res;
        memset(res, 0, sizeof *res);
        res->pair_tests = job.pair_tests; res->n_candidates = job.cand_ub; res->scored = job.scored ? 1 : 0;
        int M = hr[0], F = (job.M_ub > 0) ? hr[1] : 0;
        float median = 1.0f;                              // line3D.cc:811
        if (mr) {                                          // counters: this rank's rows; records and median: the whole view
            long long m_all = 0;
            for (int r = 0; r < ctx->world; ++r) m_all += gr.h_allrec.p[size_t(r) * VG_MAX + g].M;
            const RankRec& mine = gr.h_allrec.p[size_t(ctx->rank) * VG_MAX + g];
            M = mine.M; F = view_cnt[g];
            if (job.scored && m_all > 0) memcpy(&median, &hr[2], sizeof(float));
            if (job.scored) { res->score_iters = mine.iters_all; res->score_iters_other_cam = mine.iters_all - mine.iters_same; }
        } else if (job.scored && M > 0) {
            memcpy(&median, &hr[2], sizeof(float));
            const unsigned long long all = (unsigned long long)(unsigned)hr[4] | ((unsigned long long)(unsigned)hr[5] << 32);
            const unsigned long long same = (unsigned long long)(unsigned)hr[6] | ((unsigned long long)(unsigned)hr[7] << 32);
            res->score_iters = all; res->score_iters_other_cam = all - same;
        }
        if (job.delivering) {                              // exact sizes of what this view handed over
            DeliverBuf& b = ctx->deliver_bufs[job.owner];
            for (int d = 0; d < job.n_deliver; ++d) { b.counts[d] = hr[8 + d]; b.maxg[d] = std::min(b.maxg[d], hr[8 + d]); }
            for (size_t c = 0; c < job.nb.size(); ++c) {
                if (!job.nb[c].deliver) continue;
                std::vector<Delivered>& dl = ctx->delivered[job.nb[c].view];
                for (size_t k = 0; k < dl.size(); ++k)
                    if (dl[k].owner == job.owner) { dl[k].count = b.counts[dl[k].slot]; dl[k].max_group = b.maxg[dl[k].slot]; }
            }
        }
        for (size_t k = 0; k < job.consumed.size(); ++k) {
            DeliverBuf& b = ctx->deliver_bufs[job.consumed[k].first];
            res->n_delivered += uint64_t(b.counts[job.consumed[k].second]);
            if (--b.refs == 0) {                           // the last receiver has been collected: nothing on the GPU reads it any more
                ctx->pool.put(b.start); ctx->pool.put(b.tgt); ctx->pool.put(b.dep);
                b.start = nullptr; b.tgt = nullptr; b.dep = nullptr;
            }
        }
        res->n_scored = uint64_t(M);
        res->n_raw = uint64_t(M) - res->n_delivered;
        res->median_depth = median;
        res->n_filtered = uint64_t(F);
        l3d_match* dst = nullptr;
        if (copy_back && F > 0) {                          // on its own stream: the next group is already running
            if (ctx->retain || !single) { if (arena_take(ctx, size_t(F), &dst)) return 1; }      // results of a schedule are collected ahead of their calls
            else { L3D_CUDA(ctx->h_filtered.ensure(size_t(F))); dst = ctx->h_filtered.p; }
            const l3d_match* from = mr ? gr.d_gather.p + view_off[g] : sl.d_filtered.p;
            L3D_CUDA(cudaMemcpyAsync(dst, from, sizeof(l3d_match) * size_t(F), cudaMemcpyDeviceToHost, ctx->copy_stream));
            copies = true;
        }
        res->filtered = dst;
        if (!single) {
            l3d_ctx::Stashed& stash = ctx->results[sl.plan_index];
            memset(stash.ms, 0, sizeof stash.ms); stash.kms[0] = stash.kms[1] = stash.kms[2] = 0.f;
            if (g == 0) {                                  // the group's device times are booked on its first view
                stash.ms[0] = ms_a[0]; stash.ms[1] = ms_a[1]; stash.ms[2] = ms_b[0]; stash.ms[3] = ms_b[1]; stash.ms[4] = ms_b[2];
                stash.ms[5] = ms_a[0] + ms_a[1] + ms_b[0] + ms_b[1] + ms_b[2];
                stash.kms[0] = ctx->pa_kernel_ms[0]; stash.kms[1] = ctx->pa_kernel_ms[1]; stash.kms[2] = ctx->pa_kernel_ms[2];
            }
            stash.ready = true;
        } else {
            ctx->last_ms[0] = ms_a[0]; ctx->last_ms[1] = ms_a[1]; ctx->last_ms[2] = ms_b[0]; ctx->last_ms[3] = ms_b[1]; ctx->last_ms[4] = ms_b[2];
            ctx->last_ms[5] = ms_a[0] + ms_a[1] + ms_b[0] + ms_b[1] + ms_b[2];
        }
    }
    if (mr) {
        // the next group can go onto the GPU before these copies have arrived: they are waited for when a result is
        // handed to the caller (copies_pending), and before this group's buffers are used again (ev_x[1])
        if (copies) { L3D_CUDA(cudaEventRecord(gr.ev_x[1], ctx->copy_stream)); ctx->copies_pending = true; }
    } else if (copies) L3D_CUDA(cudaStreamSynchronize(ctx->copy_stream));
    return 0;
}

}  // namespace
