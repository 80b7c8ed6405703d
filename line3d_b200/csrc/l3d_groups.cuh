// line3d_b200/csrc/l3d_groups.cuh -- phase B kernels: one launch per group of views (VArgs), thin wrappers around l3d_scoring.cuh.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

// ------------------------------------------------------------------------------------------
// The per-view kernels run on GROUPS of views.  Which view needs which other view first (the hand-over of filtered
// matches) is a dependency graph, usually a few independent chains (views along a street; C2: 200 views, longest
// chain 50), and the views of one dependency level are independent.  One launch of each kernel below covers all views
// of a level: blockIdx.y (z for k_gather_counts) picks the view, whose buffers and sizes come from a VArgs record, and
// the grid's x extent is the largest any view of the group needs.  Per level, not per view: the host's launch work
// (it was the bottleneck at ~95 us per view) and the GPU's chain of short dependent kernels.
// ------------------------------------------------------------------------------------------
struct VArgs {
    // tables of the view
    const CamTable* cams; int n_cams, n_src; SrcTable src; ScoreParams sp;
    // pair-test results of the view's batch
    const int* raw_cnt; const int* cand_off; const int* stage_col; const float4* stage_dep; const unsigned char* stage_ok;
    // match list
    int* cnt; int* off; int4* mids; float4* mdep; float* score;
    // scoring
    float* sk; float* sd2; unsigned* spl; int* boff; uint4* brange; int2* items; int max_items; int scored, windowed, quadratic, m_hi;
    // zeroed per view
    int* nitems; unsigned long long* iters; unsigned long long* best; int* hist; int* cursor;
    // reduce, filter
    int* cta_kept; int n_cta; int* Fout; l3d_match* filtered; float* dsamples;
    // hand-over
    int delivering, stride, n_keys, n_deliver; int* dl_start; int* dl_tgt; float4* dl_dep;
    int* vres;
    // several GPUs on one scene (l3d_comm_init): this process holds the rows row_first + r * row_stride of the view
    int mr, n_seg_full;
    int sc_h;          // hypotheses per block of the scoring kernels (their shared memory holds cameras x sc_h maxima)
};

constexpr int VG_MAX = 32;         // most views in one group (l3d_ctx::group_views is how many are used)
// What one rank reports about one view of a group (all-gathered): its output records, its matches, its iteration counts.
struct RankRec { int F, M; unsigned long long iters_all, iters_same; int routed; int pad; };      // routed: records sent to the rank that holds this table
constexpr int MR_MAX_RANKS = 16;
// The group's window of every rank, as mapped into this process (CUDA IPC): [rank] -> base of that rank's window.
// A window holds, for one group of views: the ranks' records of each view (RankRec[rank][VG_MAX]), the depth samples of
// every segment of each view, and the output records of each view -- one region per rank, sized by the bound all ranks
// share, so that a rank knows where its records go in everybody's window without knowing the other ranks' counts.
struct MrPeers { unsigned char* base[MR_MAX_RANKS]; };
struct MrLayout { long long samp_off[VG_MAX], rec_off[VG_MAX], full_off[VG_MAX], cap[VG_MAX]; };      // per view of the group: bytes, bytes, bytes, records per rank
constexpr size_t MR_COUNTS_BYTES = size_t(MR_MAX_RANKS) * VG_MAX * 32;      // the RankRec table at the start of a window ...
constexpr size_t MR_FLAGS_OFF = MR_COUNTS_BYTES;                       // ... then one arrival flag per sending rank (unsigned) ...
constexpr size_t MR_HEADER_BYTES = MR_COUNTS_BYTES + 256;             // ... and the views' regions
#define L3D_VIEW(dim) L3D_PDL_PROLOGUE(); const VArgs& va = all[blockIdx.dim]

__global__ void k_gather_counts(const VArgs* __restrict__ all)
{
    L3D_VIEW(z);
    if (int(blockIdx.y) >= va.n_cams) return;
    b_gather_counts(va.cams, va.n_cams, va.n_src, va.src, va.raw_cnt, va.cnt);
}
// which = 0: count matrix -> offsets; 1: hand-over group sizes -> group starts.  One block per view.
__global__ void __launch_bounds__(SCAN_T) k_scan_views(const VArgs* __restrict__ all, int which)
{
    L3D_VIEW(y);
    if (which == 0) block_scan_global(va.cnt, va.off, va.n_src * va.n_cams);
    else if (va.delivering) block_scan_global(va.hist, va.dl_start, va.n_keys);
}
__global__ void k_fill(const VArgs* __restrict__ all, const float4* __restrict__ segs)
{
    L3D_VIEW(y);
    (void)segs;
    b_fill(va.cams, va.n_cams, va.n_src, va.src, va.off, va.cand_off, va.stage_col, va.stage_dep, va.stage_ok, va.mids, va.mdep);
}
__global__ void __launch_bounds__(BK_T) k_bucket_segments(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.scored || !va.windowed || int(blockIdx.x) >= va.n_src) return;
    b_bucket_segments(va.off, va.n_cams, va.mids, va.mdep, va.sk, va.sd2, va.spl, va.boff, va.brange, va.m_hi, va.sc_h, va.nitems, va.items);
}
__global__ void __launch_bounds__(SC_T, 12) k_score_window(const VArgs* __restrict__ all, const float4* __restrict__ segs)
{
    L3D_VIEW(y);
    if (!va.scored || !va.windowed) return;
    b_score_window(segs, va.cams, va.n_cams, va.src, va.off, va.mids, va.mdep, va.sk, va.sd2, va.spl, va.boff, va.brange, va.items, va.nitems, va.sp, va.score);
}
// the quadratic path: views without the 3-D gate, and segments too long for the window kernel's index
__global__ void k_score_items(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.scored || !va.quadratic) return;
    b_score_items(va.off, va.n_cams, va.n_src, va.windowed ? 0 : 0x7fffffff, va.windowed ? va.m_hi : 0x7fffffff, 1, va.nitems + 1, va.items + va.max_items, va.sc_h);
}
__global__ void __launch_bounds__(SC_T) k_score(const VArgs* __restrict__ all, const float4* __restrict__ segs)
{
    L3D_VIEW(y);
    if (!va.scored || !va.quadratic) return;
    b_score(segs, va.cams, va.n_cams, va.src, va.off, va.mids, va.mdep, va.items + va.max_items, va.nitems + 1, va.sp, va.score);
}
__global__ void __launch_bounds__(RM_T) k_reduce_matches(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.scored || int(blockIdx.x) >= va.n_cta) return;
    b_reduce_matches(va.off + va.n_src * va.n_cams, va.off, va.n_cams, va.score, va.mids, va.best, va.cta_kept, va.iters);
}
__global__ void __launch_bounds__(RM_T) k_filter_compact(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (int(blockIdx.x) >= va.n_cta) return;
    if (va.scored) b_filter_compact(va.off + va.n_src * va.n_cams, va.score, va.mids, va.mdep, va.cta_kept, va.filtered, va.Fout, va.cams, va.stride, (va.delivering && !va.mr) ? va.hist : nullptr, va.n_cta);
    else {          // no neighbour was pair-tested (cudawrapper.cu:877-878): all records, unscored; RM_T / 256 chunks of 256 per block
        b_copy_unscored_at(va.mids, va.mdep, va.off + va.n_src * va.n_cams, va.filtered, blockIdx.x * RM_T + threadIdx.x);
    }
}
__global__ void k_deliver_scatter(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.delivering) return;
    b_deliver_scatter(va.filtered, va.Fout, va.cams, va.stride, va.dl_start, va.cursor, va.dl_tgt, va.dl_dep);
}
__global__ void __launch_bounds__(MS_T) k_finish_view(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    const int* Mptr = va.off + va.n_src * va.n_cams;
    b_finish_view(va.dsamples, va.scored ? 2 * (va.mr ? va.n_seg_full : va.n_src) : 0, va.best, va.mdep, Mptr, va.scored ? va.Fout : Mptr, va.scored ? va.iters : nullptr,
                  va.delivering ? va.dl_start : nullptr, va.stride, va.delivering ? va.n_deliver : 0, va.vres, va.mr != 0);
}

}  // namespace
