// line3d_b200/csrc/l3d_multigpu.cuh -- one scene on several GPUs: record windows in peer memory, routed push, arrival flags, hand-over lists from routed records.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

// ------------------------------------------------------------------------------------------
// Several GPUs on ONE scene (l3d_comm_init).  The segments of every view are dealt out to the ranks in turn (segment
// s belongs to rank s mod ranks), so each rank pair-tests, assembles and scores its rows of every view, and the views
// are still processed level by level.  What a later view needs from an earlier one -- the reversed records, grouped by
// ITS segments -- is put together from all ranks' output records of the group:
//   k_mr_push      after the view's records have been filtered: every record that a later view will consume goes to
//                  the rank that holds the target segment (peer store over NVLink into that rank's window, into this
//                  rank's region of the view, appended through a counter); the depth samples of the segments this rank
//                  owns go to every rank; and, if a rank collects the records for the host, all records go to it as well
//   k_mr_signal    this rank's counts of each view into every rank's window, then -- after a system-wide fence -- its
//                  arrival flag
//   k_mr_wait      spins until every rank's flag for this group has arrived (no collective call, no host)
//   k_mr_hist, k_scan_views, k_mr_deliver_scatter   the hand-over lists for the segments this rank holds, from the
//                  records all ranks routed to it
//   k_finish_view  the median, from all ranks' samples in the own window
// The regions of a window are fixed by bounds all ranks share, so nothing has to be known of the others to store into it.
// ------------------------------------------------------------------------------------------
__global__ void k_mr_push(const VArgs* __restrict__ all, MrLayout lay, MrPeers peers, int rank, int W, int root, int* __restrict__ routed /* [VG_MAX][MR_MAX_RANKS], zeroed */)
{
    L3D_PDL_PROLOGUE();
    const int v = blockIdx.y;
    const VArgs& va = all[v];
    const int* Mptr = va.off + va.n_src * va.n_cams;
    const int F = va.scored ? *va.Fout : *Mptr;
    const int lane = threadIdx.x & 31;
    // records a later view consumes: to the owner of the target segment
    if (va.delivering) {
        for (int k0 = blockIdx.x * blockDim.x; k0 < F; k0 += gridDim.x * blockDim.x) {
            const int k = k0 + threadIdx.x;
            int dest = -1; uint4 lo, hi;
            if (k < F) {
                const uint4* src = reinterpret_cast<const uint4*>(va.filtered + k);
                lo = src[0]; hi = src[1];
                const int cam = int(lo.y), tgt = int(lo.z);                 // l3d_match: src_seg, cam, tgt_seg, confidence | depths
                if (va.cams[cam].deliver >= 0) dest = tgt % W;
            }
            // one counter bump per destination and warp
            const unsigned active = __ballot_sync(0xffffffffu, dest >= 0);
            if (dest >= 0) {
                const unsigned peers_same = __match_any_sync(active, dest);
                const int leader = __ffs(peers_same) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&routed[v * MR_MAX_RANKS + dest], __popc(peers_same));
                base = __shfl_sync(peers_same, base, leader);
                const int pos = base + __popc(peers_same & ((1u << lane) - 1u));
                uint4* dst = reinterpret_cast<uint4*>(peers.base[dest] + lay.rec_off[v]) + size_t(2) * (size_t(rank) * size_t(lay.cap[v]) + size_t(pos));
                dst[0] = lo; dst[1] = hi;
            }
        }
    }
    // all records, in order, to the rank that hands them to its host
    for (int r = (root == -2 ? 0 : root); r >= 0 && r < W; r = (root == -2 ? r + 1 : W)) {      // root: one rank, -2: every rank, -1: nobody
        const uint4* src = reinterpret_cast<const uint4*>(va.filtered);
        uint4* dst = reinterpret_cast<uint4*>(peers.base[r] + lay.full_off[v]) + size_t(2) * size_t(rank) * size_t(lay.cap[v]);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * F; i += gridDim.x * blockDim.x) dst[i] = src[i];
    }
    // depth samples of the segments this rank owns: both depths of the segment's best match if its score exceeds 0.5,
    // infinity otherwise (cudawrapper.cu:1090-1110); every segment is written by exactly one rank, to every rank
    if (va.scored) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < va.n_src * W; i += gridDim.x * blockDim.x) {
            const int row = i / W, r = i - row * W;
            const unsigned long long key = va.best[row];
            float2 ab = make_float2(INFINITY, INFINITY);
            if (__uint_as_float(unsigned(key >> 32)) > 0.5f) { const float4 dp = va.mdep[0xffffffffu - unsigned(key & 0xffffffffull)]; ab = make_float2(dp.x, dp.y); }
            reinterpret_cast<float2*>(peers.base[r] + lay.samp_off[v])[row_segment(va.src, row)] = ab;
        }
    }
}
// one thread per (view, destination rank): this rank's records of the group's views, then the flags
__global__ void k_mr_signal(const VArgs* __restrict__ all, int n_views, MrPeers peers, int rank, int W, const int* __restrict__ routed, unsigned seq)
{
    L3D_PDL_PROLOGUE();
    const int t = threadIdx.x;
    if (t < n_views * W) {
        const int v = t / W, r = t - v * W;
        const VArgs& va = all[v];
        const int* Mptr = va.off + va.n_src * va.n_cams;
        RankRec rec; rec.M = *Mptr; rec.F = va.scored ? *va.Fout : *Mptr;
        rec.iters_all = va.scored ? va.iters[0] : 0ull; rec.iters_same = va.scored ? va.iters[1] : 0ull;
        rec.routed = routed[v * MR_MAX_RANKS + r]; rec.pad = 0;
        reinterpret_cast<RankRec*>(peers.base[r])[rank * VG_MAX + v] = rec;
    }
    __threadfence_system();                 // the push kernel's stores (complete: it has ended) and the table, before the flags
    __syncthreads();
    if (t < W) {
        __threadfence_system();
        volatile unsigned* flag = reinterpret_cast<volatile unsigned*>(peers.base[t] + MR_FLAGS_OFF) + rank;
        *flag = seq;
    }
}
// waits for the flags of all ranks; gives up after about ten seconds and says so in *fault (the host reports it)
__global__ void k_mr_wait(const unsigned char* __restrict__ window, int W, unsigned seq, int* __restrict__ fault)
{
    // The next kernel's blocks are let onto the SMs only when the flags have arrived (at the end of this kernel): while
    // this rank waits for the slowest one they would sit there idle, in the way of the next batch's pair tests.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int r = threadIdx.x;
    if (r < W) {
        const volatile unsigned* flag = reinterpret_cast<const volatile unsigned*>(window + MR_FLAGS_OFF) + r;
        const long long t0 = clock64();
        while (int(*flag - seq) < 0) {
            if (clock64() - t0 > 20000000000ll) { atomicExch(fault, 1 + r); break; }
            __nanosleep(200);
        }
    }
    __threadfence_system();
}
// sizes of the hand-over groups, from the records all ranks routed here
__global__ void k_mr_hist(const VArgs* __restrict__ all, const unsigned char* __restrict__ window, MrLayout lay, int W)
{
    L3D_PDL_PROLOGUE();
    const int v = blockIdx.y;
    const VArgs& va = all[v];
    if (!va.delivering) return;
    const RankRec* rec = reinterpret_cast<const RankRec*>(window);
    for (int r = 0; r < W; ++r) {
        const int n = rec[r * VG_MAX + v].routed;
        const l3d_match* f = reinterpret_cast<const l3d_match*>(window + lay.rec_off[v]) + size_t(r) * size_t(lay.cap[v]);
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
            const int d = va.cams[f[k].cam].deliver;
            if (d >= 0) atomicAdd(&va.hist[d * va.stride + f[k].tgt_seg], 1);
        }
    }
}
__global__ void k_mr_deliver_scatter(const VArgs* __restrict__ all, const unsigned char* __restrict__ window, MrLayout lay, int W)
{
    L3D_PDL_PROLOGUE();
    const int v = blockIdx.y;
    const VArgs& va = all[v];
    if (!va.delivering) return;
    const RankRec* rec = reinterpret_cast<const RankRec*>(window);
    for (int r = 0; r < W; ++r) {
        const int n = rec[r * VG_MAX + v].routed;
        const l3d_match* f = reinterpret_cast<const l3d_match*>(window + lay.rec_off[v]) + size_t(r) * size_t(lay.cap[v]);
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
            const l3d_match m = f[k];
            const int d = va.cams[m.cam].deliver;
            if (d < 0) continue;
            const int key = d * va.stride + m.tgt_seg;
            const int pos = va.dl_start[key] + atomicAdd(&va.cursor[key], 1);
            va.dl_tgt[pos] = m.src_seg;
            va.dl_dep[pos] = make_float4(m.depths[2], m.depths[3], m.depths[0], m.depths[1]);     // line3D.cc:852-855
        }
    }
}
// the ranks' records of one group, view by view and rank by rank, into one contiguous buffer (for the copy to the host)
__global__ void k_mr_gather(const unsigned char* __restrict__ window, MrLayout lay, int W, l3d_match* __restrict__ out)
{
    const int v = blockIdx.y;
    const RankRec* rec = reinterpret_cast<const RankRec*>(window);
    long long off = 0;
    for (int vv = 0; vv < v; ++vv) for (int r = 0; r < W; ++r) off += rec[r * VG_MAX + vv].F;
    for (int r = 0; r < W; ++r) {
        const int n = rec[r * VG_MAX + v].F;
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const l3d_match*>(window + lay.full_off[v]) + size_t(r) * size_t(lay.cap[v]));
        uint4* dst = reinterpret_cast<uint4*>(out + off);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * n; i += gridDim.x * blockDim.x) dst[i] = src[i];
        off += n;
    }
}

}  // namespace
