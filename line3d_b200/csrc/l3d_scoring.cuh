// line3d_b200/csrc/l3d_scoring.cuh -- phase B device code: scoring (quadratic and bucketed window), per-segment reduction, filter, median, hand-over lists.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

// ------------------------------------------------------------------------------------------
// scoring (cudawrapper.cu:614-714).  Block = (chunk of SC_T matches of one source segment);
// every thread owns one match y and walks the segment's other matches camera group by
// camera group.  The reference's 3-D gate  |P-Q| > spatial_k |C-P|  (cudawrapper.cu:388-401)
// compares two points on the same unit ray, so it can only pass when the depths differ by
// at most spatial_k * depth (+ rounding): that 2-compare test on (d1, d2) pairs staged in
// shared memory rejects ~97-99.8 % of the iterations; the rest are queued per warp and
// evaluated with the reference's exact arithmetic, 32 at a time.
// ------------------------------------------------------------------------------------------
constexpr int SC_T = 128;
constexpr int SC_WARPS = SC_T / 32;
constexpr int SC_TILE = 2048;      // matches staged per smem tile
constexpr int SC_QCAP = 128;       // per-warp queue of gate survivors

struct ScoreParams { float sigma_p, sigma_a; };

// work items: (source segment, chunk of SC_T matches); order is irrelevant to the result
__device__ __forceinline__ void b_score_items(const int* __restrict__ off, int n_cams, int n_src, int m_lo, int m_hi, int outside, int* __restrict__ counter, int2* __restrict__ items, int chunk)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n_src) return;
    int m = off[(row + 1) * n_cams] - off[row * n_cams];
    const bool inside = m > m_lo && m <= m_hi;
    if (inside == (outside != 0)) m = 0;     // this list takes the segments inside (m_lo, m_hi], or those outside it
    const int n = (m + chunk - 1) / chunk;
    if (n == 0) return;
    const int base = atomicAdd(counter, n);
    for (int k = 0; k < n; ++k) items[base + k] = make_int2(row, k);
}

// exact evaluation of queued (hypothesis y, other match i) pairs: cudawrapper.cu:662-705
__device__ __forceinline__ void score_drain(int count, const unsigned* q, int lane, const float2* s_yd, int seg_start,
                                            const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                            V3 ray1, V3 ray2, V3 C, float spatial_k, const CamTable* __restrict__ cams,
                                            const float4* __restrict__ segs, float sigma_p, float sigma_a, int* s_max, int stride)
{
    if (lane < count) {
        const unsigned e = q[lane];
        const int yl = int(e >> 24), idx = int(e & 0xffffffu);
        const float2 dy = s_yd[yl];
        const int4 id = m_ids[seg_start + idx];
        const float4 di = m_dep[seg_start + idx];
        const CamTable& ct = cams[id.y];
        V3 P1 = xadd3(C, xscale(dy.x, ray1)), P2 = xadd3(C, xscale(dy.y, ray2));          // cudawrapper.cu:644-645
        V3 Q1 = xadd3(C, xscale(di.x, ray1)), Q2 = xadd3(C, xscale(di.y, ray2));          // cudawrapper.cu:671-672
        V3 pr1 = xproject(ct.P, P1), pr2 = xproject(ct.P, P2);
        if (pr1.z == 1.0f && pr2.z == 1.0f) {
            float conf = xhypothesis_confidence(pr1, pr2, P1, P2, Q1, Q2, C, segs[ct.seg_off + id.z], sigma_p, sigma_a, spatial_k);
            if (conf > 0.5f) atomicMax(&s_max[id.y * stride + yl], __float_as_int(conf));
        }
    }
}

__device__ __forceinline__ void b_score(const float4* __restrict__ segs, const CamTable* __restrict__ cams, int n_cams, SrcTable src,
                                                const int* __restrict__ off, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                const int2* __restrict__ items, const int* __restrict__ n_items,
                                                ScoreParams sp, float* __restrict__ score)
{
    extern __shared__ int s_max[];                 // [n_cams][H] best confidence per (camera, hypothesis), as int bits; H = blockDim.x hypotheses per block
    const int H = blockDim.x;
    __shared__ float2 s_d[SC_TILE];
    __shared__ float2 s_yd[SC_T];
    __shared__ unsigned s_q[SC_WARPS][SC_QCAP];
    if (int(blockIdx.x) >= *n_items) return;
    const int2 item = items[blockIdx.x];
    const int row = item.x;
    const int seg_start = off[row * n_cams], seg_end = off[(row + 1) * n_cams];
    const int M = seg_end - seg_start;
    const int y0 = item.y * H;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y = y0 + tid;
    const bool live = y < M;
    int my_cam = -1; float2 my_d = make_float2(0.f, 0.f);
    if (live) { my_cam = m_ids[seg_start + y].y; float4 d = m_dep[seg_start + y]; my_d = make_float2(d.x, d.y); }
    s_yd[tid] = my_d;
    for (int c = 0; c < n_cams; ++c) s_max[c * H + tid] = 0;
    // per-segment invariants (cudawrapper.cu:638-645)
    const float4 sg = segs[src.seg_off + row_segment(src, row)];
    const V3 p1 = mk3(sg.x, sg.y, 1.0f), p2 = mk3(sg.z, sg.w, 1.0f);
    const V3 ray1 = xnormalize(xmat3(src.RtKinv, p1, false)), ray2 = xnormalize(xmat3(src.RtKinv, p2, false));
    const V3 C = mk3(src.C[0], src.C[1], src.C[2]);
    // superset gate:  |dy - di| <= thr  must hold for the exact gate to be able to pass
    const float k = src.spatial_k;
    const float thr1 = (k * fabsf(my_d.x) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.x) + src.cmax)) * 1.00001f;
    const float thr2 = (k * fabsf(my_d.y) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.y) + src.cmax)) * 1.00001f;
    const bool gate_on = k > 0.0f;
    unsigned* q = s_q[warp];
    int qn = 0;
    // this thread skips its own camera's range [skip0, skip1) (cudawrapper.cu:674-675)
    int skip0 = 0, skip1 = 0;
    if (live) { skip0 = off[row * n_cams + my_cam] - seg_start; skip1 = off[row * n_cams + my_cam + 1] - seg_start; }
    for (int t0 = 0; t0 < M; t0 += SC_TILE) {
        const int tn = min(SC_TILE, M - t0);
        __syncthreads();
        for (int i = tid; i < tn; i += H) { float4 d = m_dep[seg_start + t0 + i]; s_d[i] = make_float2(d.x, d.y); }
        __syncthreads();
        for (int i0 = 0; i0 < tn; i0 += 32) {
            unsigned hits = 0;
            if (live) {
                const int lim = min(32, tn - i0);
                if (gate_on) {
                    unsigned mk = 1u;
#pragma unroll 8
                    for (int b = 0; b < 32; ++b, mk <<= 1) {
                        const float2 d = s_d[i0 + b];           // beyond tn: stale but masked below
                        bool pass = (fabsf(my_d.x - d.x) <= thr1) && (fabsf(my_d.y - d.y) <= thr2);
                        hits |= pass ? mk : 0u;
                    }
                    if (lim < 32) hits &= (1u << lim) - 1u;
                } else {
                    hits = (lim >= 32) ? 0xffffffffu : ((1u << lim) - 1u);
                }
                // drop the entries of this thread's own camera
                const int g = t0 + i0;
                if (g < skip1 && g + 32 > skip0) {
                    const int lo = max(skip0 - g, 0), hi = min(skip1 - g, 32);
                    const unsigned m = ((hi >= 32) ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
                    hits &= ~m;
                }
            }
            while (__any_sync(0xffffffffu, hits != 0u)) {
                int n = __popc(hits), ex = n;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { int v = __shfl_up_sync(0xffffffffu, ex, o); if (lane >= o) ex += v; }
                ex -= n;
                int pos = qn + ex, pushed = 0;
                while (hits && pos < SC_QCAP) { int b = __ffs(hits) - 1; hits &= hits - 1; q[pos++] = (unsigned(tid) << 24) | unsigned(t0 + i0 + b); ++pushed; }
                int tot = pushed;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                qn += tot;
                __syncwarp();
                while (qn >= 32) {
                    score_drain(32, q + (qn - 32), lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, k, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
                    qn -= 32;
                    __syncwarp();
                }
            }
        }
    }
    if (qn > 0) score_drain(qn, q, lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, k, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
    __syncwarp();
    // sum of the per-camera best confidences in camera order (cudawrapper.cu:677-687, 709)
    float total = 0.0f;
    for (int c = 0; c < n_cams; ++c) total = xa(total, __int_as_float(s_max[c * H + tid]));
    if (live) score[seg_start + y] = total;
}

// ------------------------------------------------------------------------------------------
// scoring, sorted-window variant.  The superset gate of k_score is |d1_y - d1_i| <= thr1(y) and
// |d2_y - d2_i| <= thr2(y).  If the segment's matches are sorted by d1, the partners of y that can
// pass the first condition are one contiguous window, found by binary search: the quadratic walk
// over all (y, i) pairs becomes O(M log M) + the window contents.  The set of pairs handed to the
// exact evaluation is the same as in k_score (the same two fabsf tests are applied inside the
// window), so the scores are bit-identical; only the order of evaluation changes, and the result
// is a per-camera maximum, which does not depend on it.
// One CTA per source segment with lo < M <= MP matches; the segment is sorted once in shared
// memory (bitonic, key d1, payload d2 + camera + original index).
// ------------------------------------------------------------------------------------------
// Grouping the matches of a segment by depth, for the windowed scoring.  A hypothesis can only be supported by matches
// whose first depth lies within a small relative window of its own (the 3-D gate, see k_score_window), so each
// segment's matches are distributed over BK_B buckets by d1 -- a counting sort: three barriers per segment whatever
// its length, where a comparison sort of the longest segments took 66 barrier-separated steps and set the kernel's
// duration.  The bucket of a depth is linear in the bit pattern of the float between the segment's smallest and
// largest d1 (positive floats order like unsigned integers, and the pattern is a piecewise-linear log2, which suits
// a window that is relative): monotonic, so the buckets a window touches are the buckets of its two ends.  The
// order inside a bucket is whatever the atomics give; the scores do not depend on it (per-camera maxima).
// Output: (d1, d2, camera << 20 | index) at the segment's own offsets in bucket order, the bucket boundaries, and
// (bits of min d1, bits of max d1, scale) per segment.
constexpr int BK_T = 128;
constexpr int BK_B = 512;
__device__ __forceinline__ int depth_bucket(unsigned bits, unsigned bits_min, float scale)
{
    return min(BK_B - 1, int(float(bits - bits_min) * scale));
}

__device__ __forceinline__ void b_bucket_segments(const int* __restrict__ off, int n_cams, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                          float* __restrict__ g_k, float* __restrict__ g_d2, unsigned* __restrict__ g_pl,
                                                          int* __restrict__ g_boff /* [n_src][BK_B + 1] */, uint4* __restrict__ g_brange,
                                                          int m_hi, int chunk, int* __restrict__ n_items, int2* __restrict__ items)
{
    __shared__ int s_cnt[BK_B], s_start[BK_B + 1];
    __shared__ unsigned s_lo[BK_T / 32], s_hi[BK_T / 32];
    __shared__ int s_item0;
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg_start = off[row * n_cams], M = off[(row + 1) * n_cams] - seg_start;
    if (M <= 0 || M > m_hi) return;
    // the scoring kernel's work items of this segment: (segment, chunk of `chunk` hypotheses); their order is irrelevant
    const int n_chunks = (M + chunk - 1) / chunk;
    if (tid == 0) s_item0 = atomicAdd(n_items, n_chunks);
    unsigned lo = 0xffffffffu, hi = 0u;
    for (int i = tid; i < M; i += BK_T) { const unsigned b = __float_as_uint(m_dep[seg_start + i].x); lo = min(lo, b); hi = max(hi, b); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
    if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; }
    for (int b = tid; b < BK_B; b += BK_T) s_cnt[b] = 0;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < BK_T / 32; ++w) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
    const float scale = float(BK_B) / (float(hi - lo) + 1.0f);
    for (int i = tid; i < M; i += BK_T) atomicAdd(&s_cnt[depth_bucket(__float_as_uint(m_dep[seg_start + i].x), lo, scale)], 1);
    __syncthreads();
    if (warp == 0) {                               // exclusive scan of the BK_B counts: 4 per lane
        int c[BK_B / 32], sum = 0;
#pragma unroll
        for (int r = 0; r < BK_B / 32; ++r) { c[r] = s_cnt[lane * (BK_B / 32) + r]; sum += c[r]; }
        int ex = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, ex, o); if (lane >= o) ex += t; }
        ex -= sum;
#pragma unroll
        for (int r = 0; r < BK_B / 32; ++r) { s_start[lane * (BK_B / 32) + r] = ex; ex += c[r]; }
        if (lane == 31) s_start[BK_B] = ex;
    }
    __syncthreads();
    int* boff = g_boff + size_t(row) * (BK_B + 1);
    for (int b = tid; b < BK_B; b += BK_T) boff[b] = s_start[b];
    if (tid == 0) { boff[BK_B] = s_start[BK_B]; g_brange[row] = make_uint4(lo, hi, __float_as_uint(scale), 0u); }
    for (int c = tid; c < n_chunks; c += BK_T) items[s_item0 + c] = make_int2(row, c);
    for (int b = tid; b < BK_B; b += BK_T) s_cnt[b] = 0;      // now the fill cursors
    __syncthreads();
    for (int i = tid; i < M; i += BK_T) {
        const float4 d = m_dep[seg_start + i];
        const int bk = depth_bucket(__float_as_uint(d.x), lo, scale);
        const int pos = seg_start + s_start[bk] + atomicAdd(&s_cnt[bk], 1);
        g_k[pos] = d.x; g_d2[pos] = d.y;
        g_pl[pos] = (unsigned(m_ids[seg_start + i].y) << 20) | unsigned(i);
    }
}

// work items (segment, chunk of SC_T hypotheses) as in k_score; the partners of a hypothesis are looked for in the
// depth buckets its window touches
__device__ __forceinline__ void b_score_window(const float4* __restrict__ segs, const CamTable* __restrict__ cams, int n_cams, SrcTable src,
                                                       const int* __restrict__ off, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                       const float* __restrict__ g_k, const float* __restrict__ g_d2, const unsigned* __restrict__ g_pl,
                                                       const int* __restrict__ g_boff, const uint4* __restrict__ g_brange,
                                                       const int2* __restrict__ items, const int* __restrict__ n_items,
                                                       ScoreParams sp, float* __restrict__ score)
{
    extern __shared__ int s_max[];                 // [n_cams][H], H = blockDim.x hypotheses per block
    const int H = blockDim.x;
    __shared__ float2 s_yd[SC_T];
    __shared__ unsigned s_q[SC_WARPS][SC_QCAP];
    if (int(blockIdx.x) >= *n_items) return;
    const int2 item = items[blockIdx.x];
    const int row = item.x;
    const int seg_start = off[row * n_cams], M = off[(row + 1) * n_cams] - seg_start;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y = item.y * H + tid;
    const bool live = y < M;
    // thread y takes the y-th match in BUCKET order: the lanes of a warp then hold neighbouring depths, their windows
    // overlap (coalesced loads) and are about equally long (little divergence in the scan below)
    int my_cam = -1, my_idx = 0; float2 my_d = make_float2(0.f, 0.f);
    if (live) { const unsigned pl = g_pl[seg_start + y]; my_cam = int(pl >> 20); my_idx = int(pl & 0xfffffu); my_d = make_float2(g_k[seg_start + y], g_d2[seg_start + y]); }
    s_yd[tid] = my_d;
    for (int c = 0; c < n_cams; ++c) s_max[c * H + tid] = 0;
    const float4 sg = segs[src.seg_off + row_segment(src, row)];
    const V3 p1 = mk3(sg.x, sg.y, 1.0f), p2 = mk3(sg.z, sg.w, 1.0f);
    const V3 ray1 = xnormalize(xmat3(src.RtKinv, p1, false)), ray2 = xnormalize(xmat3(src.RtKinv, p2, false));
    const V3 C = mk3(src.C[0], src.C[1], src.C[2]);
    const float kk = src.spatial_k;
    const float thr1 = (kk * fabsf(my_d.x) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.x) + src.cmax)) * 1.00001f;
    const float thr2 = (kk * fabsf(my_d.y) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.y) + src.cmax)) * 1.00001f;
    const float* K = g_k + seg_start; const float* D2 = g_d2 + seg_start; const unsigned* PL = g_pl + seg_start;
    int j = 0, jend = 0;
    if (live) {        // slightly widened window; the exact fabsf tests follow
        const float lo = my_d.x - thr1 * 1.001f - 1e-30f, hi = my_d.x + thr1 * 1.001f + 1e-30f;
        const uint4 br = g_brange[row];
        const float dmin = __uint_as_float(br.x), dmax = __uint_as_float(br.y), scale = __uint_as_float(br.z);
        if (hi >= dmin && lo <= dmax) {
            const int* boff = g_boff + size_t(row) * (BK_B + 1);
            j = boff[depth_bucket(__float_as_uint(fmaxf(lo, dmin)), br.x, scale)];
            jend = boff[depth_bucket(__float_as_uint(fminf(hi, dmax)), br.x, scale) + 1];
        }
    }
    __syncwarp();
    unsigned* q = s_q[warp];
    int qn = 0;
    while (__any_sync(0xffffffffu, j < jend)) {
        bool hit = false; unsigned pl = 0;
        if (j < jend) {
            hit = (fabsf(my_d.x - K[j]) <= thr1) && (fabsf(my_d.y - D2[j]) <= thr2);
            if (hit) { pl = PL[j]; hit = int(pl >> 20) != my_cam; }
            ++j;
        }
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (m) {
            if (hit) q[qn + __popc(m & ((1u << lane) - 1u))] = (unsigned(tid) << 24) | (pl & 0xfffffu);
            qn += __popc(m);
            __syncwarp();
            if (qn >= 32) {
                score_drain(32, q + (qn - 32), lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, kk, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
                qn -= 32;
                __syncwarp();
            }
        }
    }
    if (qn > 0) score_drain(qn, q, lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, kk, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
    __syncwarp();
    if (live) {        // sum of the per-camera best confidences in camera order (cudawrapper.cu:677-687, 709)
        float total = 0.0f;
        for (int c = 0; c < n_cams; ++c) total = xa(total, __int_as_float(s_max[c * H + tid]));
        score[seg_start + my_idx] = total;
    }
}

// ------------------------------------------------------------------------------------------
// per-segment reduction (cudawrapper.cu:1028-1064, 1090-1110): best match of every segment (first maximum) and
// the matches with score > 1.  One thread per match, all loads independent (a warp per segment walking its matches
// was a chain of dependent loads as long as the longest segment): the best match is a 64-bit atomicMax on
// (score bits, ~match index) -- scores are positive, so their bits order like the values, and among equal scores the
// smaller index wins as in the reference's strict '>' scan --, the kept matches are counted per CTA for the ordered
// compaction in k_filter_compact, and the two iteration counters are summed on the way.
// ------------------------------------------------------------------------------------------
constexpr int RM_T = 1024;
__device__ __forceinline__ void b_reduce_matches(const int* __restrict__ Mptr, const int* __restrict__ off, int n_cams, const float* __restrict__ score,
                                                         const int4* __restrict__ m_ids, unsigned long long* __restrict__ best /* [n_src], zeroed */,
                                                         int* __restrict__ cta_kept, unsigned long long* __restrict__ iters /* [0] all pairs, [1] same-camera pairs */)
{
    __shared__ int s_k[RM_T / 32];
    __shared__ unsigned long long s_a[RM_T / 32], s_b[RM_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = blockIdx.x * RM_T + tid, M = *Mptr;
    int keep = 0; unsigned long long all = 0ull, same = 0ull;
    if (k < M) {
        const float c = score[k]; const int4 id = m_ids[k];
        keep = c > 1.0f;
        const int* o = off + id.w * n_cams;
        all = (unsigned long long)(o[n_cams] - o[0]);                 // sum over the matches of their segment's size = sum of squares
        same = (unsigned long long)(o[id.y + 1] - o[id.y]);
        if (c > 0.0f) {
            const unsigned long long key = ((unsigned long long)__float_as_uint(c) << 32) | (unsigned long long)(0xffffffffu - unsigned(k));
            if (key > best[id.w]) atomicMax(&best[id.w], key);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { keep += __shfl_xor_sync(0xffffffffu, keep, o); all += __shfl_xor_sync(0xffffffffu, all, o); same += __shfl_xor_sync(0xffffffffu, same, o); }
    if (lane == 0) { s_k[warp] = keep; s_a[warp] = all; s_b[warp] = same; }
    __syncthreads();
    if (warp == 0) {
        keep = s_k[lane]; all = s_a[lane]; same = s_b[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { keep += __shfl_xor_sync(0xffffffffu, keep, o); all += __shfl_xor_sync(0xffffffffu, all, o); same += __shfl_xor_sync(0xffffffffu, same, o); }
        if (lane == 0) { cta_kept[blockIdx.x] = keep; if (all) atomicAdd(iters, all); if (same) atomicAdd(iters + 1, same); }
    }
}

// median of the finite samples: sorted[count / 2], -1 if none (cudawrapper.cu:1066-1076).
// Radix select in one block: the samples are positive floats or +inf (no sample), so their bit patterns order like
// unsigned integers; four passes of an 8-bit histogram over the samples that still match the prefix found so far
// narrow the wanted rank down to one value.
constexpr int MS_T = 1024;
__device__ float block_median_select(const float* __restrict__ dsamples, int n)
{
    __shared__ int s_hist[256];
    __shared__ int s_cnt[MS_T / 32];
    __shared__ unsigned s_prefix; __shared__ int s_rank;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int finite = 0;
    for (int i = tid; i < n; i += MS_T) finite += (dsamples[i] < INFINITY);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) finite += __shfl_xor_sync(0xffffffffu, finite, o);
    if (lane == 0) s_cnt[warp] = finite;
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    int count = 0;
#pragma unroll
    for (int w = 0; w < MS_T / 32; ++w) count += s_cnt[w];
    if (count == 0) return -1.0f;                   // (the same for every thread of the block)
    if (tid == 0) { s_prefix = 0u; s_rank = count / 2; }
    unsigned prefix = 0u, mask = 0u;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = tid; i < n; i += MS_T) {       // the samples are re-read every pass (16 KB, L1/L2)
            const float x = dsamples[i];
            const unsigned v = __float_as_uint(x);
            if (x < INFINITY && (v & mask) == prefix) atomicAdd(&s_hist[(v >> shift) & 255u], 1);
        }
        __syncthreads();
        if (warp == 0) {                            // the bin in which the cumulative count passes the wanted rank
            int c[8], sum = 0;
#pragma unroll
            for (int r = 0; r < 8; ++r) { c[r] = s_hist[lane * 8 + r]; sum += c[r]; }
            int inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
            const int rank = s_rank;
            int ex = inc - sum;
            if (ex <= rank && rank < inc) {
#pragma unroll
                for (int r = 0; r < 8; ++r) { if (rank < ex + c[r]) { s_prefix = prefix | (unsigned(lane * 8 + r) << shift); s_rank = rank - ex; break; } ex += c[r]; }
            }
        }
        __syncthreads();
        prefix = s_prefix; mask |= 255u << shift;
        if (tid < 256) s_hist[tid] = 0;
        __syncthreads();
    }
    return __uint_as_float(prefix);
}

// The last kernel of a view: the median depth, and all the small results in one buffer so that one copy brings
// them to the host: out[0] = matches, out[1] = output records, out[2] = median depth (float bits), out[4..7] = two
// 64-bit iteration counters, out[8 + d] = records handed to delivery slot d.  One block.
__device__ __forceinline__ void b_finish_view(float* __restrict__ dsamples, int n_samples, const unsigned long long* __restrict__ best, const float4* __restrict__ m_dep,
                                                      const int* __restrict__ M, const int* __restrict__ F,
                                                      const unsigned long long* __restrict__ iters, const int* __restrict__ start, int stride, int n_deliver,
                                                      int* __restrict__ out, bool samples_ready)
{
    // depth samples: both depths of every segment's best match if its score exceeds 0.5 (cudawrapper.cu:1090-1110)
    // (with several GPUs on one scene every rank's k_mr_push has written them into this rank's window already)
    for (int row = threadIdx.x; !samples_ready && 2 * row < n_samples; row += MS_T) {
        const unsigned long long key = best[row];
        float a = INFINITY, b = INFINITY;
        if (__uint_as_float(unsigned(key >> 32)) > 0.5f) { const float4 d = m_dep[0xffffffffu - unsigned(key & 0xffffffffull)]; a = d.x; b = d.y; }
        dsamples[2 * row] = a; dsamples[2 * row + 1] = b;
    }
    __syncthreads();
    const float median = (n_samples > 0) ? block_median_select(dsamples, n_samples) : 0.0f;
    if (threadIdx.x == 0) {
        out[0] = *M; out[1] = *F; out[2] = __float_as_int(median); out[3] = 0;
        const unsigned long long a = iters ? iters[0] : 0ull, b = iters ? iters[1] : 0ull;
        out[4] = int(a & 0xffffffffull); out[5] = int(a >> 32); out[6] = int(b & 0xffffffffull); out[7] = int(b >> 32);
    }
    for (int d = threadIdx.x; d < n_deliver; d += blockDim.x) out[8 + d] = start[(d + 1) * stride] - start[d * stride];
}

// filtered records (score > 1) in match order = the reference's order: an ordered compaction of the whole match list.
// CTA b keeps cta_kept[b] records (k_reduce_matches), so it starts at the sum of the counts before it; inside the CTA
// a ballot/prefix gives every kept match its place.  Also counts the records per hand-over group (destination
// view's segment), the first step of the regrouping below.
__device__ __forceinline__ void b_filter_compact(const int* __restrict__ Mptr, const float* __restrict__ score, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                         const int* __restrict__ cta_kept, l3d_match* __restrict__ out, int* __restrict__ Fout,
                                                         const CamTable* __restrict__ cams, int stride, int* __restrict__ hist /* or null */, int n_cta)
{
    __shared__ int s_w[RM_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = blockIdx.x * RM_T + tid, M = *Mptr;
    int part = 0;
    for (int i = tid; i < int(blockIdx.x); i += RM_T) part += cta_kept[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_w[warp] = part;
    __syncthreads();
    int base = 0;
#pragma unroll
    for (int w = 0; w < RM_T / 32; ++w) base += s_w[w];
    __syncthreads();
    const float c = (k < M) ? score[k] : 0.0f;
    const bool keep = c > 1.0f;
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_w[warp] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < RM_T / 32; ++w) { const int n = s_w[w]; if (w < warp) before += n; total += n; }
    if (keep) {
        const int r = base + before + __popc(m & ((1u << lane) - 1u));
        const int4 id = m_ids[k]; const float4 d = m_dep[k];
        l3d_match rec; rec.src_seg = id.x; rec.cam = id.y; rec.tgt_seg = id.z; rec.confidence = xd(c, 2.0f);
        rec.depths[0] = d.x; rec.depths[1] = d.y; rec.depths[2] = d.z; rec.depths[3] = d.w;
        out[r] = rec;
        if (hist) { const int dl = cams[id.y].deliver; if (dl >= 0) atomicAdd(&hist[dl * stride + id.z], 1); }     // hand-over group sizes
    }
    if (int(blockIdx.x) == n_cta - 1 && tid == 0) *Fout = base + total;
}
// unscored variant (no neighbour to pair-test): all records, confidence 0
__device__ __forceinline__ void b_copy_unscored_at(const int4* __restrict__ m_ids, const float4* __restrict__ m_dep, const int* __restrict__ Mptr, l3d_match* __restrict__ out, int k)
{
    if (k >= *Mptr) return;
    int4 id = m_ids[k]; float4 d = m_dep[k];
    l3d_match rec; rec.src_seg = id.x; rec.cam = id.y; rec.tgt_seg = id.z; rec.confidence = 0.0f;
    rec.depths[0] = d.x; rec.depths[1] = d.y; rec.depths[2] = d.z; rec.depths[3] = d.w;
    out[k] = rec;
}

// ------------------------------------------------------------------------------------------
// hand-over (line3D.cc:838-858): the filtered records towards each neighbour that asked for
// delivery, reversed and regrouped by the neighbour's segment, target ids ascending
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void b_deliver_scatter(const l3d_match* __restrict__ f, const int* __restrict__ Fptr, const CamTable* __restrict__ cams, int stride,
                                  const int* __restrict__ start, int* __restrict__ cursor, int* __restrict__ t_tgt, float4* __restrict__ t_dep)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= *Fptr) return;
    const l3d_match m = f[k];
    const int d = cams[m.cam].deliver;
    if (d < 0) return;
    const int key = d * stride + m.tgt_seg;
    const int pos = start[key] + atomicAdd(&cursor[key], 1);
    t_tgt[pos] = m.src_seg;
    t_dep[pos] = make_float4(m.depths[2], m.depths[3], m.depths[0], m.depths[1]);     // line3D.cc:852-855
}

}  // namespace
