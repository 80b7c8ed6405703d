// line3d_b200/csrc/l3d_context.h -- host-side state: hand-over bookkeeping, view slots and groups, the two sets of pair-test buffers, and the context behind l3d_ctx.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

struct Delivered {            // records a source view handed to a later view
    int from_view = -1;
    int count = 0;            // records (a bound until the sending view has been collected; several GPUs: this rank's share)
    long long count_all = 0;  // several GPUs: a bound on all ranks' records together, the same number on every rank (it fixes the window layout)
    int* t_start = nullptr;   // n_seg(to) + 1, points into the owning buffer
    int* t_tgt = nullptr;
    float4* t_dep = nullptr;
    int owner = -1;           // index into ctx->deliver_bufs
    int max_group = 0;        // most records delivered to a single segment
    int slot = 0;             // position among the owner's deliveries
};
struct DeliverBuf { int* start = nullptr; int* tgt = nullptr; float4* dep = nullptr; int refs = 0; std::vector<int> counts, maxg; };   // counts/maxg: per slot, bounds until the owner is collected

struct PlannedView {           // one entry of l3d_plan_matching
    int view = -1;
    std::vector<l3d_neighbor> nb;
    uint64_t pair_tests = 0;     // of this process; with several GPUs on one scene: of the rank with most rows (the same number on every rank)
    int first_job = -1;         // valid while the view's batch is resident
    uint64_t n_candidates = 0;
};
struct JobHost { int src_view, tgt_view; float F[9]; };
struct ViewJob {                 // phase B of one view: on the stream, not yet collected
    bool active = false;
    int view = -1, parity = 0;
    std::vector<l3d_neighbor> nb;
    float sigma_p = 0.f, sigma_a = 0.f;
    uint64_t pair_tests = 0, cand_ub = 0;
    size_t M_ub = 0;
    bool scored = false, delivering = false;
    int n_deliver = 0, owner = -1;
    std::vector<std::pair<int, int> > consumed;      // (deliver buffer, slot) of every delivery the view took in
};
// The per-view work runs on groups of views (see VArgs): a slot holds one view's buffers, a group up to VG_MAX
// slots plus what the group's views share (one upload of tables and arguments, one memset, one result download).
// Two groups exist: the host prepares and launches the next one while the GPU works on the current one.
struct ViewSlot {
    ViewJob job;
    size_t plan_index = 0;
    DBuf<int> d_cnt, d_off, d_cta, d_F, d_boff;
    DBuf<int4> d_mids; DBuf<float4> d_mdep; DBuf<float> d_score, d_dsamples, d_sk, d_sd2; DBuf<unsigned> d_spl;
    DBuf<int2> d_items; DBuf<uint4> d_brange; DBuf<l3d_match> d_filtered;
    void release()
    {
        d_cnt.release(); d_off.release(); d_cta.release(); d_F.release(); d_boff.release(); d_mids.release(); d_mdep.release(); d_score.release();
        d_dsamples.release(); d_sk.release(); d_sd2.release(); d_spl.release(); d_items.release(); d_brange.release(); d_filtered.release();
    }
};
struct ViewGroup {
    int count = 0; size_t pos0 = 0;      // positions [pos0, pos0 + count) of the execution order
    ViewSlot slot[VG_MAX];
    HBuf<VArgs> h_args; DBuf<VArgs> d_args;
    HBuf<CamTable> h_cams; DBuf<CamTable> d_cams; size_t cam_off[VG_MAX + 1];
    DBuf<int> d_zero; DBuf<int> d_vres; HBuf<int> h_vres; int res_stride = 8;     // result ints per view: 8 scalars + one count per hand-over slot
    // several GPUs on one scene: depth samples of the group's views (one buffer: one all-reduce), the ranks' records,
    // the packed records of this rank, all ranks' packed records, all ranks' records view by view
    size_t hist_off = 0, hist_total = 0;
    HBuf<RankRec> h_allrec; MrLayout lay; DBuf<l3d_match> d_gather; int* d_fault = nullptr;
    cudaEvent_t ev_x[2] = {nullptr, nullptr};      // [0] exchange done on the main stream, [1] this group's records copied to the host
    cudaEvent_t kev[24] = {}; int kev_slot[24] = {}; int n_kev = 0;      // per-kernel events of the chain (l3d_debug_chain_times)
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // [0] results on the host, [2..5] stage boundaries
    void release()
    {
        for (int k = 0; k < VG_MAX; ++k) slot[k].release();
        h_args.release(); d_args.release(); h_cams.release(); d_cams.release(); d_zero.release(); d_vres.release(); h_vres.release();
        h_allrec.release(); d_gather.release();
        for (int k = 0; k < 2; ++k) if (ev_x[k]) { cudaEventDestroy(ev_x[k]); ev_x[k] = nullptr; }
        for (int k = 0; k < 24; ++k) if (kev[k]) { cudaEventDestroy(kev[k]); kev[k] = nullptr; }
        for (int k = 0; k < 6; ++k) if (ev[k]) { cudaEventDestroy(ev[k]); ev[k] = nullptr; }
    }
};
// Device memory for the hand-over buffers, recycled on the host: a buffer goes back when the last receiving view has
// been collected, so nothing on the GPU can still be reading it (stream-ordered allocation across streams cost 10 us
// per call here).
struct DevPool {
    struct Block { void* p; size_t cap; bool free; };
    std::vector<Block> blocks;
    cudaError_t get(size_t bytes, void** out)
    {
        size_t best = blocks.size();
        for (size_t k = 0; k < blocks.size(); ++k)
            if (blocks[k].free && blocks[k].cap >= bytes && blocks[k].cap <= 4 * std::max<size_t>(bytes, 4096) && (best == blocks.size() || blocks[k].cap < blocks[best].cap)) best = k;
        if (best == blocks.size()) {
            size_t cap = 4096; while (cap < bytes) cap <<= 1;
            void* p = nullptr;
            cudaError_t e = cudaMalloc(&p, cap);
            if (e != cudaSuccess) return e;
            Block b = {p, cap, true}; blocks.push_back(b);
        }
        blocks[best].free = false; *out = blocks[best].p;
        return cudaSuccess;
    }
    void put(void* p) { for (size_t k = 0; k < blocks.size(); ++k) if (blocks[k].p == p) { blocks[k].free = true; return; } }
    void put_all() { for (size_t k = 0; k < blocks.size(); ++k) blocks[k].free = true; }
    void release() { for (size_t k = 0; k < blocks.size(); ++k) cudaFree(blocks[k].p); blocks.clear(); }
};

// Everything one batch of pair tests (phase A) produces and needs: two of these exist, so that a helper thread can
// run the next batch's pair tests on a second stream while the per-view work (phase B) of the current batch, which
// leaves most of the GPU idle, goes on.
struct PhaseASet {
    cudaStream_t stream = nullptr;
    std::string err;
    uint64_t launches = 0;
    size_t begin = 0, end = 0;                 // plan entries covered
    DBuf<SrcHoist> d_sh; DBuf<SrcHoist1> d_sh1; DBuf<TgtHoist> d_th; DBuf<int> d_row_job;
    DBuf<unsigned> d_cand, d_l1; DBuf<int> d_tile_sums, d_total;
    DBuf<int> d_stage_col, d_stage_row; DBuf<float4> d_stage_dep; DBuf<unsigned char> d_stage_ok;
    DBuf<JobDev> d_jobs; DBuf<int> d_pa_cnt, d_pa_off, d_pa_raw, d_pa_idx, d_pa_tot;
    std::vector<JobDev> jobs_host; std::vector<uint64_t> job_cands, job_raw, job_raw_all /* summed over the ranks (several GPUs on one scene) */; std::vector<int> job_rowmax, job_raw_rowmax; DBuf<int> d_rm_begin, d_rm_len, d_rm_out;
    DBuf<unsigned long long> d_job_sum64; HBuf<unsigned long long> h_job_sum64;
    bool pa_ms_pending = false;
    bool background = false;      // this batch's pair tests run beside the per-view chain of the batch before (helper thread)
    cudaEvent_t pa_ev[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t pa_kev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // [0][1] around k_pretest, [1][4] around k_pretest2, [2][3] around k_exact_flat
    bool pa_kev_set[2] = {false, false};
    HBuf<int> h_res;
    void release()
    {
        d_sh.release(); d_sh1.release(); d_th.release(); d_row_job.release(); d_cand.release(); d_l1.release(); d_tile_sums.release(); d_total.release();
        d_stage_col.release(); d_stage_row.release(); d_stage_dep.release(); d_stage_ok.release(); d_jobs.release();
        d_pa_cnt.release(); d_pa_off.release(); d_pa_raw.release(); d_pa_idx.release(); d_pa_tot.release();
        d_rm_begin.release(); d_rm_len.release(); d_rm_out.release(); h_res.release(); d_job_sum64.release(); h_job_sum64.release();
        for (int k = 0; k < 3; ++k) if (pa_ev[k]) { cudaEventDestroy(pa_ev[k]); pa_ev[k] = nullptr; }
        for (int k = 0; k < 5; ++k) if (pa_kev[k]) { cudaEventDestroy(pa_kev[k]); pa_kev[k] = nullptr; }
        if (stream) { cudaStreamDestroy(stream); stream = nullptr; }
    }
};

struct l3d_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    // several GPUs on one scene (l3d_comm_init)
    NcclApi* nccl = nullptr; ncclComm_t comm = nullptr; int rank = 0, world = 1, result_root = -1;
    bool comm_broken = false;          // a call failed while several ranks were at work: no collective step is attempted any more (the others may be gone)
    bool copies_pending = false, want_records = true;      // want_records: the current calls copy the output records to the host
    bool chain_prof = false; double chain_ms[24] = {};      // device ms per kernel of the per-view chain, summed over the groups (with stage timing)
    // the clusters' 3-D lines (l3d_fit_lines): CSR of the clusters, and per member scratch (points, centred products, sweep state)
    DBuf<int> d_fit_start, d_fit_count, d_fit_member, d_fit_camopen, d_fit_order; DBuf<unsigned> d_fit_cam; DBuf<unsigned char> d_fit_open; DBuf<double> d_fit_dbl; DBuf<float> d_fit_dist;
    // the neighbour graph and the fundamental matrices (l3d_visual_neighbors, l3d_fundamentals)
    DBuf<l3d_camera> d_geo_cams; DBuf<int> d_geo_start, d_geo_other, d_geo_ids, d_geo_count; DBuf<float> d_geo_sim, d_geo_sims; DBuf<int2> d_geo_pairs; DBuf<double> d_geo_F;
    double host_s[8] = {0, 0, 0, 0, 0, 0, 0, 0};      // host seconds in enqueue_group / finish_group (waiting) / finish_group (rest) / number of groups / waiting for the helper thread / pair tests in the foreground / totals over the ranks (l3d_debug_host_times)
    // how the pair tests of the next batch share the GPU with the per-view chain (read from the environment by l3d_plan_matching)
    int bg_blocks = 0;                              // L3D_BG_BLOCKS: blocks per SM the large pair-test kernels may take in the background (0: no cap)
    int p1_rows = 0;                                // L3D_P1_ROWS: source rows per block of k_pretest (0: P1_ROWS)
    bool no_overlap = false;                        // L3D_NO_OVERLAP: pair tests never run beside the chain
    int batch_div = 8;                              // L3D_BATCH_DIV: the schedule is cut into about this many batches
    int group_views = 8;                            // L3D_GROUP_VIEWS: views per group of the per-view chain (at most VG_MAX)
    // the record windows (one per group in flight): every rank's is mapped into every process, the ranks store their
    // output records straight into each other's
    unsigned char* win[2] = {nullptr, nullptr}; size_t win_cap = 0 /* bytes */; MrPeers peers[2]; unsigned win_seq[2] = {0u, 0u};
    DBuf<int> d_bar; DBuf<unsigned long long> d_tot; HBuf<unsigned long long> h_tot; DBuf<unsigned char> d_ipc; HBuf<unsigned char> h_ipc;
    uint64_t launches = 0;
    // views
    int V = 0;
    std::vector<l3d_view> views;          // host copies (segments pointer not kept)
    std::vector<int> seg_off;
    int total_segs = 0, max_seg = 0;
    DBuf<float4> d_segs;
    DBuf<ViewD> d_viewd;
    std::vector<std::vector<Delivered> > delivered;   // per destination view
    std::vector<DeliverBuf> deliver_bufs;
    // scratch
    DBuf<int> d_tile_sums, d_total;
    size_t score_smem = 48 * 1024; int n_sm = 148;
    // phase A (pair tests): the batch the per-view work reads (pa[cur]) and the one a helper thread prepares meanwhile
    PhaseASet pa[2]; int cur = 0;
    std::thread pa_worker; bool pa_worker_active = false; int pa_worker_rc = 0;
    std::vector<PlannedView> plan;                  // as announced (the order of the caller's calls)
    std::vector<size_t> order;                      // plan entries in the order they are executed: by dependency level
    struct Stashed { l3d_match_result res; bool ready; float ms[6], kms[3]; };
    std::vector<Stashed> results;                   // per plan entry, until the caller asks for it
    size_t call_next = 0; bool plan_started = false, run_copy_back = true;
    size_t batch_begin = 0, batch_end = 0;          // positions in `order` whose pair tests are resident
    uint64_t pair_budget = 24000000000ull;          // most pair tests of one batch (3 GB of candidate bits)
    uint64_t batch_pairs = 24000000000ull;          // pair tests per batch of the announced schedule (l3d_plan_matching)
    uint64_t batch_override = 0;                    // l3d_debug_set_batch_pairs
    int pretest_mode = 0;                           // l3d_debug_set_pretest
    int score_block = 0;                            // l3d_debug_set_score_block
    uint64_t cand_limit = 0x7fff0000ull;            // most candidates of one batch (32-bit candidate index); l3d_debug_set_candidate_limit
    float pa_kernel_ms[3] = {0.f, 0.f, 0.f};
    // views in flight: positions [fly_begin, fly_end) of `order` are on the GPU, as the groups [grp_begin, grp_end)
    ViewGroup grp[2]; size_t grp_begin = 0, grp_end = 0;
    size_t fly_begin = 0, fly_end = 0; std::vector<int> level;
    float run_sigma_p = 0.f, run_sigma_a = 0.f;
    DevPool pool;
    cudaStream_t copy_stream = nullptr, fin_stream = nullptr;
    bool stage_events = false;         // per-stage CUDA events inside a view (l3d_set_stage_timing): they separate kernels that otherwise overlap
    cudaEvent_t mark_ev[2] = {nullptr, nullptr};
    // retained results (l3d_retain_results): pinned chunks that are never moved, handed out front to back
    bool retain = false; std::vector<std::pair<l3d_match*, size_t> > arena; size_t arena_chunk = 0, arena_used = 0;
    HBuf<l3d_match> h_filtered; HBuf<int> h_ints;
    // collinearity
    DBuf<float4> d_csegs; DBuf<unsigned> d_cbits; DBuf<int> d_ccnt, d_coff; DBuf<l3d_collin_pair> d_cpairs; HBuf<l3d_collin_pair> h_cpairs;
    // hypotheses
    DBuf<Hyp> d_hyp; size_t n_hyp = 0; DBuf<l3d_hypothesis_in> d_hin; DBuf<l3d_segment3d> d_hout;
    DBuf<l3d_pair> d_pairs; DBuf<float> d_aff, d_medians;
    HBuf<l3d_candidate> h_cands; DBuf<l3d_candidate> d_cands; DBuf<float> d_hscore, d_ew; DBuf<int> d_ecnt, d_eoff; DBuf<l3d_kept_edge> d_kept; HBuf<l3d_kept_edge> h_kept; DBuf<unsigned long long> d_ekeys; DBuf<unsigned> d_eorder; HBuf<unsigned> h_eorder;
    DBuf<unsigned char> d_flush; int flush_value = 0;
    // diffusion: one device and one pinned buffer for all its arrays; the step kernel's duration and algorithmic bytes of the last call
    DBuf<int> d_rdd; HBuf<int> h_rdd; float rdd_step_ms = 0.f; double rdd_step_bytes = 0.0; int rdd_nnz = 0;
    // timing
    cudaEvent_t ev[2] = {nullptr, nullptr};          // around the FP32 peak probe
    float last_ms[6] = {0, 0, 0, 0, 0, 0};
};
