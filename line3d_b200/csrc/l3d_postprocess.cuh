// line3d_b200/csrc/l3d_postprocess.cuh -- hypotheses and affinities in FP64, edge order (bitonic), diffusion, and the probes used by tests and bench.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

// ------------------------------------------------------------------------------------------
// hypotheses (view.cc:302-342) and affinity (line3D.cc:1600-1691), FP64
// ------------------------------------------------------------------------------------------
struct ViewD { double RtKinv[9]; double C[3]; float k_lower, k_upper; int seg_off; int n_seg; };
struct Hyp { double P1[3], P2[3], dir[3]; float d1, d2; int view; int pad; };

__device__ __forceinline__ void d_normalize(double* v)
{
    double n = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(v[0], v[0]), __dmul_rn(v[1], v[1])), __dmul_rn(v[2], v[2])));
    v[0] = __ddiv_rn(v[0], n); v[1] = __ddiv_rn(v[1], n); v[2] = __ddiv_rn(v[2], n);
}
__global__ void k_unproject(const float4* __restrict__ segs, const ViewD* __restrict__ views, const l3d_hypothesis_in* __restrict__ in, size_t n,
                            Hyp* __restrict__ hyp, l3d_segment3d* __restrict__ out)
{
    const size_t k = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const l3d_hypothesis_in h = in[k];
    const ViewD& v = views[h.view];
    const float4 s = segs[v.seg_off + h.seg];
    const double p[2][3] = {{double(s.x), double(s.y), 1.0}, {double(s.z), double(s.w), 1.0}};
    const double dep[2] = {double(h.depth1), double(h.depth2)};
    Hyp o; double P[2][3];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        double r[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < 3; ++c) acc = __dadd_rn(acc, __dmul_rn(v.RtKinv[a * 3 + c], p[e][c]));
            r[a] = acc;
        }
        d_normalize(r);
#pragma unroll
        for (int a = 0; a < 3; ++a) P[e][a] = __dadd_rn(v.C[a], __dmul_rn(r[a], dep[e]));
    }
    double d[3] = {__dsub_rn(P[1][0], P[0][0]), __dsub_rn(P[1][1], P[0][1]), __dsub_rn(P[1][2], P[0][2])};
    d_normalize(d);
    l3d_segment3d so;
#pragma unroll
    for (int a = 0; a < 3; ++a) { o.P1[a] = P[0][a]; o.P2[a] = P[1][a]; o.dir[a] = d[a]; so.P1[a] = P[0][a]; so.P2[a] = P[1][a]; so.dir[a] = d[a]; }
    o.d1 = h.depth1; o.d2 = h.depth2; o.view = h.view; o.pad = 0;
    hyp[k] = o; out[k] = so;
}

// line3D.cc:1684-1691: |P1 + (dir (X - P1)^T) dir - X|, the outer product evaluated first
__device__ __forceinline__ float d_point_line(const Hyp& s, const double* X)
{
    double v[3] = {__dsub_rn(X[0], s.P1[0]), __dsub_rn(X[1], s.P1[1]), __dsub_rn(X[2], s.P1[2])};
    double e[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < 3; ++j) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(s.dir[i], v[j]), s.dir[j]));
        e[i] = __dsub_rn(__dadd_rn(s.P1[i], acc), X[i]);
    }
    return float(sqrt(__dadd_rn(__dadd_rn(__dmul_rn(e[0], e[0]), __dmul_rn(e[1], e[1])), __dmul_rn(e[2], e[2]))));
}
// view.cc:353-377
__device__ __forceinline__ float d_sim_term(float d, float depth, float k_lower, float k_upper, float median, float neg2log)
{
    float lo = depth < median ? xm(k_lower, depth) : xm(k_lower, median);
    float up = depth < median ? xm(k_upper, depth) : xm(k_upper, median);
    if (d < lo) return 1.0f;
    float diff = xs(up, lo);
    float sig = xd(xm(-diff, diff), neg2log);                      // -(d2-d1)^2 / (2 logf(0.01f))
    float t = xs(d, lo);
    return xexpf(xd(xm(-t, t), xm(2.0f, sig)));
}
__global__ void k_affinity(const Hyp* __restrict__ hyp, const ViewD* __restrict__ views, const float* __restrict__ median,
                           const l3d_pair* __restrict__ pairs, size_t n, float sigma_a, float two_log001, float* __restrict__ out)
{
    const size_t k = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const Hyp a = hyp[pairs[k].a], b = hyp[pairs[k].b];
    const ViewD& va = views[a.view]; const ViewD& vb = views[b.view];
    const float ma = median[a.view], mb = median[b.view];
    float s1 = d_sim_term(d_point_line(b, a.P1), a.d1, va.k_lower, va.k_upper, ma, two_log001);
    float s2 = d_sim_term(d_point_line(b, a.P2), a.d2, va.k_lower, va.k_upper, ma, two_log001);
    float s3 = d_sim_term(d_point_line(a, b.P1), b.d1, vb.k_lower, vb.k_upper, mb, two_log001);
    float s4 = d_sim_term(d_point_line(a, b.P2), b.d2, vb.k_lower, vb.k_upper, mb, two_log001);
    float w_d = fminf(fminf(s1, s2), fminf(s3, s4));
    double dt = __dadd_rn(__dadd_rn(__dmul_rn(a.dir[0], b.dir[0]), __dmul_rn(a.dir[1], b.dir[1])), __dmul_rn(a.dir[2], b.dir[2]));
    float angle = float(acos(fmax(fmin(dt, 1.0), -1.0)) / 3.14159265358979323846 * 180.0);
    if (angle > 90.0f) angle = xs(180.0f, angle);
    float w_a = xexpf(xd(xm(-angle, angle), xm(xm(2.0f, sigma_a), sigma_a)));
    float sim = fminf(w_d, w_a);
    out[k] = (sim <= 0.01f) ? 0.0f : sim;
}

// Candidate pairs -> affinity edges on the device (line3D.cc:1014-1016, 1085-1087, 1163-1165): the similarity of the two
// hypotheses, the edge weight  mult * 0.5 * (score_a + score_b) * sim  in the host code's order of operations, the
// threshold of the candidate's kind, and an ordered compaction of the survivors -- only they travel back to the host.
constexpr int AE_T = 256;
__device__ __forceinline__ float affinity_of(const Hyp& a, const Hyp& b, const ViewD* __restrict__ views, const float* __restrict__ median, float sigma_a, float two_log001)
{
    const ViewD& va = views[a.view]; const ViewD& vb = views[b.view];
    const float ma = median[a.view], mb = median[b.view];
    float s1 = d_sim_term(d_point_line(b, a.P1), a.d1, va.k_lower, va.k_upper, ma, two_log001);
    float s2 = d_sim_term(d_point_line(b, a.P2), a.d2, va.k_lower, va.k_upper, ma, two_log001);
    float s3 = d_sim_term(d_point_line(a, b.P1), b.d1, vb.k_lower, vb.k_upper, mb, two_log001);
    float s4 = d_sim_term(d_point_line(a, b.P2), b.d2, vb.k_lower, vb.k_upper, mb, two_log001);
    float w_d = fminf(fminf(s1, s2), fminf(s3, s4));
    double dt = __dadd_rn(__dadd_rn(__dmul_rn(a.dir[0], b.dir[0]), __dmul_rn(a.dir[1], b.dir[1])), __dmul_rn(a.dir[2], b.dir[2]));
    float angle = float(acos(fmax(fmin(dt, 1.0), -1.0)) / 3.14159265358979323846 * 180.0);
    if (angle > 90.0f) angle = xs(180.0f, angle);
    float w_a = xexpf(xd(xm(-angle, angle), xm(xm(2.0f, sigma_a), sigma_a)));
    float sim = fminf(w_d, w_a);
    return (sim <= 0.01f) ? 0.0f : sim;
}
__global__ void __launch_bounds__(AE_T) k_affinity_edges(const Hyp* __restrict__ hyp, const ViewD* __restrict__ views, const float* __restrict__ median, const float* __restrict__ score,
                                                         const l3d_candidate* __restrict__ cands, size_t n, float sigma_a, float two_log001,
                                                         float* __restrict__ w_out, int* __restrict__ cta_kept)
{
    __shared__ int s_n[AE_T / 32];
    const size_t k = size_t(blockIdx.x) * AE_T + threadIdx.x;
    float w = 0.0f; bool keep = false;
    if (k < n) {
        const l3d_candidate c = cands[k];
        const float sim = affinity_of(hyp[c.a], hyp[c.b], views, median, sigma_a, two_log001);
        w = xm(xm(xm(c.mult, 0.5f), xa(score[c.a], score[c.b])), sim);
        keep = w > c.thr;
        w_out[k] = keep ? w : 0.0f;
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if ((threadIdx.x & 31) == 0) s_n[threadIdx.x >> 5] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int q = 0; q < AE_T / 32; ++q) t += s_n[q]; cta_kept[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(AE_T) k_affinity_compact(const float* __restrict__ w_in, size_t n, const int* __restrict__ cta_off, l3d_kept_edge* __restrict__ out)
{
    __shared__ int s_n[AE_T / 32];
    const size_t k = size_t(blockIdx.x) * AE_T + threadIdx.x;
    const float w = (k < n) ? w_in[k] : 0.0f;
    const bool keep = w != 0.0f;                       // (a kept weight exceeds a positive threshold)
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_n[warp] = __popc(m);
    __syncthreads();
    int before = 0;
    for (int q = 0; q < warp; ++q) before += s_n[q];
    if (keep) { l3d_kept_edge e; e.index = unsigned(k); e.w = w; out[cta_off[blockIdx.x] + before + __popc(m & ((1u << lane) - 1u))] = e; }
}

// The order in which the clustering takes the edges (clustering.cc:12: a stable sort by ascending weight): the kept
// candidates sorted by the 64-bit key (weight bits, position) -- weights are positive, so their bit patterns order like
// the values, and the position makes every key unique, so any sort gives the stable order.  Bitonic network over a
// power of two of keys (padding: all ones); strides below the tile size run in shared memory.
constexpr int BS_TILE = 2048;
__global__ void k_edge_keys(const l3d_kept_edge* __restrict__ kept, int n, int n_pad, unsigned long long* __restrict__ keys)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pad) return;
    keys[k] = (k < n) ? (((unsigned long long)__float_as_uint(kept[k].w) << 32) | (unsigned long long)unsigned(k)) : ~0ull;
}
__global__ void k_bitonic_global(unsigned long long* __restrict__ keys, int j, int k)
{
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned p = i ^ unsigned(j);
    if (p > i) {
        const unsigned long long a = keys[i], b = keys[p];
        const bool up = (i & unsigned(k)) == 0;
        if ((a > b) == up) { keys[i] = b; keys[p] = a; }
    }
}
__global__ void __launch_bounds__(BS_TILE / 2) k_bitonic_shared(unsigned long long* __restrict__ keys, int j_first, int k)
{
    __shared__ unsigned long long s[BS_TILE];
    const unsigned base = blockIdx.x * BS_TILE;
    s[threadIdx.x] = keys[base + threadIdx.x]; s[threadIdx.x + BS_TILE / 2] = keys[base + threadIdx.x + BS_TILE / 2];
    __syncthreads();
    for (int j = j_first; j > 0; j >>= 1) {
        const unsigned t = threadIdx.x;
        const unsigned i = ((t & ~unsigned(j - 1)) << 1) | (t & unsigned(j - 1));      // the lower index of the t-th pair at stride j
        const unsigned p = i | unsigned(j);
        const bool up = ((base + i) & unsigned(k)) == 0;
        const unsigned long long a = s[i], b = s[p];
        if ((a > b) == up) { s[i] = b; s[p] = a; }
        __syncthreads();
    }
    keys[base + threadIdx.x] = s[threadIdx.x]; keys[base + threadIdx.x + BS_TILE / 2] = s[threadIdx.x + BS_TILE / 2];
}
__global__ void k_edge_order(const unsigned long long* __restrict__ keys, int n, unsigned* __restrict__ order)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) order[k] = unsigned(keys[k] & 0xffffffffull);
}

// ------------------------------------------------------------------------------------------
// diffusion (cudawrapper.cu:717-829, 1131-1191)
// ------------------------------------------------------------------------------------------
// P: row-sorted entries (row, col, val); W: column-sorted.  rstart/cstart: first entry of a
// row of P / a column of W.  tpos[y] = position in P of the entry (col_y, row_y).
__global__ void k_rdd_normalise(const int* __restrict__ prow, float* __restrict__ pval, const int* __restrict__ rstart, int n, int nnz)
{
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= n) return;
    const int s = rstart[y];
    if (s < 0) return;
    float sum = 0.0f;
    int i = s;
    while (i < nnz && prow[i] == y) { sum = xa(sum, pval[i]); ++i; }
    if (sum < L3D_EPS_F) sum = L3D_EPS_F;
    for (int k = s; k < i; ++k) pval[k] = xd(pval[k], sum);
}
__global__ void k_rdd_step(const int* __restrict__ prow, const int* __restrict__ pcol, const float* __restrict__ pval,
                           const int* __restrict__ wcol, const float* __restrict__ wval,
                           const int* __restrict__ rstart, const int* __restrict__ cstart, const int* __restrict__ tpos,
                           float* __restrict__ pnext, int nnz)
{
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= nnz) return;
    const int r = pcol[y], c = prow[y];                     // transposed (cudawrapper.cu:778-780)
    float mul = 0.0f;
    int sp = rstart[r], sw = cstart[c];
    while (sp >= 0 && sw >= 0 && sp < nnz && sw < nnz) {    // lockstep walk (cudawrapper.cu:786-800)
        if (prow[sp] != r || wcol[sw] != c) break;
        mul = xa(mul, xm(pval[sp], wval[sw]));
        ++sp; ++sw;
    }
    mul = xm(mul, pval[y]);
    if (mul < L3D_EPS_F) mul = L3D_EPS_F;
    if (tpos[y] >= 0) pnext[tpos[y]] = mul;
}

// libm-compatible float functions, exposed for tests
__global__ void k_debug_math(int which, const float* __restrict__ in, float* __restrict__ out, int n)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    out[k] = which == 0 ? xexpf(in[k]) : xacosf(in[k]);
}

// FP32 throughput probes: register-only chains, operands in registers (not immediates), 16 independent chains per
// thread, 8 steps unrolled: 128 FFMA (64 FFMA2) per loop trip against 3 loop instructions.
__global__ void k_fma_probe(float* out, int iters, float b, float c)
{
    float a[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = threadIdx.x * 1e-3f + float(k);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int k = 0; k < 16; ++k) a[k] = fmaf(a[k], b, c);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fma2_probe(float* out, int iters, float b, float c)
{
    f32x2 a[8];
    const f32x2 B = pk2(b, b * 1.0000001f), C = pk2(c, c * 1.5f);
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = pk2(threadIdx.x * 1e-3f + float(k), threadIdx.x * 2e-3f + float(k));
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fma2(a[k], B, C);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { unsigned lo, hi; up2(a[k], lo, hi); s += __uint_as_float(lo) + __uint_as_float(hi); }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace
