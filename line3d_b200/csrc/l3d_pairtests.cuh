// line3d_b200/csrc/l3d_pairtests.cuh -- phase A: hoisting, the two conservative pre-tests, candidate expansion and the exact pair test with triangulation.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

// ------------------------------------------------------------------------------------------
// hoist: per (pair-tested neighbour, segment) epipolar data for the pre-test
//   src side:  A_n = F p_n (n = 1,2) and w'_n = |(A_n)_xy|
//   tgt side:  q1, q2 and w_m = |(F^T q_m)_xy|
// evaluated in FP64 from the FP32 tables, rounded once.
// ------------------------------------------------------------------------------------------
struct SrcHoist { float4 a; float4 b; };   // a = (A1x, A1y, A1z, W1)   b = (A2x, A2y, A2z, W2),  A_n = F p_n / rho,  W_n = 2 eps |(F p_n)_xy|
struct TgtHoist { float4 q; float2 w; float2 pad; };   // w_m = 2 eps kappa |(F^T q_m)_xy|
struct SrcHoist1 { float4 p; float4 m; };  // level 1: p = (A_1 + A_2) / rho, m = (A_1 - A_2) / rho  (sum and difference of the residuals at the two endpoints)

// one pair-tested (source view, neighbour) combination; phase A runs over an array of these
struct JobDev {
    float F[9], RtKinv_s[9], RtKinv_t[9], C_s[3], C_t[3];
    int src_off, n_src, tgt_off, n_tgt;      // segment ranges in the global segment array (n_src: rows of this job)
    int src_first, src_stride;               // row r of the job is segment src_off + src_first + r * src_stride of the view
    int sh_off, th_off;                      // first entry in the hoist arrays
    long long cand_off;                      // first word in the candidate bitmask
    int wpr;                                 // words per row
    int row_off;                             // first global row (= source segment of this job) in the per-row arrays
    int blk_off, col_blocks;                 // level-2 blocks of this job: blk_off + rb * col_blocks + cb
    long long l1_off;                        // first word of the job's level-1 bits: [row group][target column]
    int l1_blk_off;                          // level-1 blocks of this job start here
};
__device__ __forceinline__ int job_src_segment(const JobDev& jb, int row) { return jb.src_off + jb.src_first + row * jb.src_stride; }

__device__ __forceinline__ int find_job_by_block(const JobDev* __restrict__ jobs, int n_jobs, int blk)
{
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (jobs[mid].blk_off <= blk) lo = mid; else hi = mid - 1; }
    return lo;
}

#define L3D_PRETEST_EPS_PX 0.25
#define L3D_PRETEST_LO 0.09f
#define L3D_PRETEST_HI 0.28f
#define L3D_PRETEST_MAXABS 1e6f     // beyond this the float tests could overflow: the pre-test is switched off for the segment

// How far (in units of the 0.25 px allowance) the reference's own FP32 arithmetic can move the supporting line of the
// segment (x1, y1)-(x2, y2) or an intersection with it, relative to coordinates around 2 kpx and long segments, for
// which the allowance was validated:
//   * cross(p1, p2).z = x1 y2 - y1 x2 is a difference of two rounded products (cudawrapper.cu:546-547): its error of up
//     to 1.5 ulp(|x y|) moves the line by that much divided by the segment's length -- 0.06 px for a 12 px segment at
//     2 kpx, but 1.5 px for a 3 px one, and 16 times more at 8 kpx;
//   * the intersections are cross products of such lines: their error grows with the square of the coordinates.
// The factor is >= 1; margins are multiplied by it (source side: the residuals are divided by it instead).
__device__ __forceinline__ double pretest_noise_factor(double x1, double y1, double x2, double y2)
{
    const double c = fmax(fmax(fabs(x1), fabs(y1)), fmax(fabs(x2), fabs(y2)));
    const double g = fmax(1.0, (c / 2048.0) * (c / 2048.0));
    const double m = fmax(fmax(fabs(x1 * y2), fabs(y1 * x2)), 1.0);
    const double ulp = scalbn(1.0, ilogb(m) - 23);
    const double len = fmax(hypot(x1 - x2, y1 - y2), 1e-3);
    return g * (1.0 + (3.0 * ulp / len) / L3D_PRETEST_EPS_PX);
}

// mode 0: normal; 1: the pre-test is switched off for every segment (all pairs go to the exact test; l3d_debug_set_pretest)
__global__ void k_hoist(const float4* __restrict__ segs, const JobDev* __restrict__ jobs, SrcHoist* __restrict__ sh, SrcHoist1* __restrict__ sh1, TgtHoist* __restrict__ th, int* __restrict__ row_job, int mode)
{
    const JobDev& jb = jobs[blockIdx.y];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= jb.n_src && idx >= jb.n_tgt) return;
    // The tests are homogeneous in F: it is scaled by a power of two (exact) so that its largest entry lies in [1, 2)
    // and squares of residuals stay far from the ends of the float range.
    double F[9], fmaxabs = 0.0;
#pragma unroll
    for (int k = 0; k < 9; ++k) { F[k] = double(jb.F[k]); fmaxabs = fmax(fmaxabs, fabs(F[k])); }
    const bool f_ok = fmaxabs > 0.0 && fmaxabs < 1e300;      // (a NaN fails both comparisons)
    if (f_ok) {
        const double sc = scalbn(1.0, -ilogb(fmaxabs));
#pragma unroll
        for (int k = 0; k < 9; ++k) F[k] *= sc;
    }
    const float INF = __int_as_float(0x7f800000);
    if (idx < jb.n_src) {
        float4 s = segs[job_src_segment(jb, idx)];
        double p[2][2] = {{double(s.x), double(s.y)}, {double(s.z), double(s.w)}};
        const double rho = pretest_noise_factor(p[0][0], p[0][1], p[1][0], p[1][1]);
        float o[2][4]; double A[2][3];
        bool ok = f_ok && mode == 0;
#pragma unroll
        for (int n = 0; n < 2; ++n) {
            double ax = F[0] * p[n][0] + F[1] * p[n][1] + F[2];
            double ay = F[3] * p[n][0] + F[4] * p[n][1] + F[5];
            double az = F[6] * p[n][0] + F[7] * p[n][1] + F[8];
            A[n][0] = ax / rho; A[n][1] = ay / rho; A[n][2] = az / rho;
            o[n][0] = float(A[n][0]); o[n][1] = float(A[n][1]); o[n][2] = float(A[n][2]); o[n][3] = float(2.0 * L3D_PRETEST_EPS_PX * sqrt(ax * ax + ay * ay));
            ok = ok && fabsf(o[n][0]) < L3D_PRETEST_MAXABS && fabsf(o[n][1]) < L3D_PRETEST_MAXABS && fabsf(o[n][2]) < L3D_PRETEST_MAXABS && o[n][3] < L3D_PRETEST_MAXABS;
        }
        SrcHoist1 h1;
        h1.p = make_float4(float(A[0][0] + A[1][0]), float(A[0][1] + A[1][1]), float(A[0][2] + A[1][2]), 0.f);
        h1.m = make_float4(float(A[0][0] - A[1][0]), float(A[0][1] - A[1][1]), float(A[0][2] - A[1][2]), 0.f);
        // Values the float tests cannot handle switch the pre-test off for this segment: residuals of +1 at one endpoint
        // and -1 at the other ("the epipolar line crosses the segment") and infinite margins keep every pair at both levels.
        if (!ok) {
            o[0][0] = o[0][1] = o[1][0] = o[1][1] = 0.f; o[0][2] = 1.f; o[1][2] = -1.f; o[0][3] = o[1][3] = INF;
            h1.p = make_float4(0.f, 0.f, 0.f, 0.f); h1.m = make_float4(0.f, 0.f, 2.f, 0.f);      // sum 0, difference 2
        }
        SrcHoist h; h.a = make_float4(o[0][0], o[0][1], o[0][2], o[0][3]); h.b = make_float4(o[1][0], o[1][1], o[1][2], o[1][3]);
        sh[jb.sh_off + idx] = h; sh1[jb.sh_off + idx] = h1;
        row_job[jb.row_off + idx] = blockIdx.y;          // row -> job, so that later kernels need no search
    }
    if (idx < jb.n_tgt) {
        float4 t = segs[jb.tgt_off + idx];
        double q[2][2] = {{double(t.x), double(t.y)}, {double(t.z), double(t.w)}};
        const double kappa = pretest_noise_factor(q[0][0], q[0][1], q[1][0], q[1][1]);
        float w[2];
        bool ok = f_ok && mode == 0 && fabsf(t.x) < L3D_PRETEST_MAXABS && fabsf(t.y) < L3D_PRETEST_MAXABS && fabsf(t.z) < L3D_PRETEST_MAXABS && fabsf(t.w) < L3D_PRETEST_MAXABS;
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            double ex = F[0] * q[m][0] + F[3] * q[m][1] + F[6];
            double ey = F[1] * q[m][0] + F[4] * q[m][1] + F[7];
            w[m] = float(2.0 * L3D_PRETEST_EPS_PX * kappa * sqrt(ex * ex + ey * ey));
            ok = ok && w[m] > 1e-12f && w[m] < L3D_PRETEST_MAXABS;      // (w / 2)^2 must be a normal float
        }
        if (!ok) { t = make_float4(0.f, 0.f, 0.f, 0.f); w[0] = w[1] = INF; }
        TgtHoist h; h.q = t; h.w = make_float2(w[0], w[1]); h.pad = make_float2(0.f, 0.f);
        th[jb.th_off + idx] = h;
    }
}

// ------------------------------------------------------------------------------------------
// pre-test.  For the pair (src s = p1p2, tgt t = q1q2) let E[m][n] = q_m^T F p_n.  The
// intersection of the src line with the epipolar line of q_m sits at parameter
//     tau_m = E[m][1] / (E[m][1] - E[m][2])        (0 at p1, 1 at p2)
// and the intersection of the tgt line with the epipolar line of p_n at
//     tau'_n = E[1][n] / (E[1][n] - E[2][n]).
// The reference's overlap value (cudawrapper.cu:209-252) is the 1-D intersection over
// union of [0,1] with [tau_1,tau_2] (resp. tau'), or 0.  A pair is dropped here only if,
// with every tau moved by an error bound worth L3D_PRETEST_EPS_PX pixels of intersection
// noise, an overlap still cannot reach the reference's thresholds minus a margin; all
// others go to the exact test.  The result is a superset of the reference's accepted set.
// ------------------------------------------------------------------------------------------

__device__ __forceinline__ float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ void side_bounds(float e1a, float e1b, float e2a, float e2b, float w1, float w2, float& I_hi, float& U_lo)
{
    // endpoint m: tau = ema / (ema - emb)
    float d1 = e1a - e1b, d2 = e2a - e2b;
    // approximate reciprocals (one MUFU each, 1 ulp): the thresholds below keep a 10 % margin; a zero or denormal
    // denominator gives inf, then NaN/inf bounds, and those comparisons keep the pair
    float r1 = fast_rcp(d1), r2 = fast_rcp(d2);
    float t1 = e1a * r1, t2 = e2a * r2;
    float err = 0.5f * (w1 * fabsf(r1) * (1.0f + 2.0f * fabsf(t1)) + w2 * fabsf(r2) * (1.0f + 2.0f * fabsf(t2)));
    float lo = fminf(t1, t2), hi = fmaxf(t1, t2);
    I_hi = fminf(hi, 1.0f) - fmaxf(lo, 0.0f) + err;
    U_lo = fmaxf(hi, 1.0f) - fminf(lo, 0.0f) - err;
}

// Packed FP32 pairs (sm_100: FFMA2 / FADD2 work on 64-bit register pairs, one issue slot for two operations; an operand
// given as pk2(x, x) is encoded as a scalar that is broadcast to both halves).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void up2(f32x2 v, unsigned& lo, unsigned& hi) { asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// Two levels, two kernels.  Level 1 (k_pretest) is a sign test on the four residuals: the src segment lies entirely on
// one side of the epipolar line of q_m when E[m][1] and E[m][2] have the same sign; if that holds for both target
// endpoints, by more than the allowance (L3D_PRETEST_EPS_PX pixels), and both intersections fall beyond the same end of
// the segment (|E[m][1]| - |E[m][2]| has the same sign for m = 1, 2), the overlap is zero whatever the rounding.  About
// 96 % of all pairs end here.  It works on the sum and the difference of the two residuals of a line,
//     S_m = E[m][1] + E[m][2] = q_m . (A_1 + A_2),      D_m = E[m][1] - E[m][2] = q_m . (A_1 - A_2),
// which the hoisted source vectors give directly (packed over m: 2 FFMA2 each).  |S_m| - |D_m| is twice the smaller
// magnitude when the residuals share a sign and negative when they do not, so
//     x_m = |S_m| - |D_m| - w_m                   sign set: not on one side by more than the allowance
// and, when they share a sign, |E[m][1]| - |E[m][2]| = sign(S_m) D_m, so the sign of (S_1 D_1)(S_2 D_2) tells whether
// the two intersections lie beyond different ends.  The verdict is the sign bit of x_1 | x_2 | (S_1 D_1)(S_2 D_2):
// 4 FFMA2 + 2 FADD + 1 FADD2 + 1 FMUL2 + 1 FMUL on the FMA pipe (15 of its cycles per warp and 32 pairs), one LOP3 and one
// funnel shift on the ALU pipe (2 cycles each, and on this part they are cycles the FMA pipe loses), no branch.
// Only the source side is tested: on the frozen scenes it rejects 95.8 % of all pairs and the mirrored test on the
// target side would add 0.4 % for as many instructions again -- level 2 covers both sides anyway.
// Its output is one bit per pair, 32 source rows of one target column per word (what a lane holds after 32 rows),
// stored [row group][column]: a warp's store is one coalesced line.
// Level 2 (k_pretest2) reads those words, compacts the ~4 % survivors through a per-warp queue, applies the interval
// test on full warps and writes the candidate bitmask [row][column / 32] and the candidates per row.
constexpr int P1_COLS = 256;        // target segments per block: 4 warps x 32 lanes x 2 columns per lane
constexpr int P1_BLOCK = 128;
constexpr int P1_CHUNK = 128;       // source rows staged in shared memory at a time
constexpr int P1_ROWS = 512;        // source rows per block

__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// QX = (q_1.x, q_2.x), QY = (q_1.y, q_2.y), nW = (-w_1, -w_2); Ap, Am: the hoisted sum and difference vectors of the row
__device__ __forceinline__ unsigned pretest_level1(f32x2 QX, f32x2 QY, f32x2 nW, const float4 Ap, const float4 Am)
{
    const f32x2 S = fma2(QX, pk2(Ap.x, Ap.x), fma2(QY, pk2(Ap.y, Ap.y), pk2(Ap.z, Ap.z)));      // (S_1, S_2)
    const f32x2 D = fma2(QX, pk2(Am.x, Am.x), fma2(QY, pk2(Am.y, Am.y), pk2(Am.z, Am.z)));      // (D_1, D_2)
    unsigned sl, sh_, dl, dh;
    up2(S, sl, sh_); up2(D, dl, dh);
    const float y1 = fabsf(__uint_as_float(sl)) - fabsf(__uint_as_float(dl)), y2 = fabsf(__uint_as_float(sh_)) - fabsf(__uint_as_float(dh));
    const f32x2 X = add2(pk2(y1, y2), nW);
    const f32x2 P = mul2(S, D);
    unsigned xl, xh, pl, ph;
    up2(X, xl, xh); up2(P, pl, ph);
    const float pp = __uint_as_float(pl) * __uint_as_float(ph);
    return xl | xh | __float_as_uint(pp);      // sign bit clear: surely no overlap on the src side
}

template <int UNROLL, int MINB>
__global__ void __launch_bounds__(P1_BLOCK, MINB) k_pretest(const SrcHoist1* __restrict__ sh, const TgtHoist* __restrict__ th,
                                                            const JobDev* __restrict__ jobs, int n_jobs, unsigned* __restrict__ l1, int rows_per_block)
{
    __shared__ float4 s_a[P1_CHUNK], s_b[P1_CHUNK];
    __shared__ int s_job;
    if (threadIdx.x == 0) {      // the job this block belongs to: blocks are numbered job by job (l1 blocks: l1_blk_off)
        int lo = 0, hi = n_jobs - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (jobs[mid].l1_blk_off <= int(blockIdx.x)) lo = mid; else hi = mid - 1; }
        s_job = lo;
    }
    __syncthreads();
    const JobDev& jb = jobs[s_job];
    const int local = blockIdx.x - jb.l1_blk_off;
    const int col_blocks = (jb.n_tgt + P1_COLS - 1) / P1_COLS;
    const int rb = local / col_blocks, cb = local - rb * col_blocks;
    const int row_begin = rb * rows_per_block, row_end = min(jb.n_src, row_begin + rows_per_block);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = cb * P1_COLS + warp * 64 + lane, c1 = c0 + 32;
    f32x2 QX0 = 0, QY0 = 0, NT0 = 0, QX1 = 0, QY1 = 0, NT1 = 0;
    if (c0 < jb.n_tgt) { const TgtHoist h = th[jb.th_off + c0]; QX0 = pk2(h.q.x, h.q.z); QY0 = pk2(h.q.y, h.q.w); NT0 = pk2(-h.w.x, -h.w.y); }
    if (c1 < jb.n_tgt) { const TgtHoist h = th[jb.th_off + c1]; QX1 = pk2(h.q.x, h.q.z); QY1 = pk2(h.q.y, h.q.w); NT1 = pk2(-h.w.x, -h.w.y); }
    unsigned* out = l1 + jb.l1_off;
    // the rows of the next chunk travel through registers while the current chunk is computed
    float4 na = make_float4(0.f, 0.f, 0.f, 0.f), nb = na;
    if (row_begin + tid < row_end) { const SrcHoist1 h = sh[jb.sh_off + row_begin + tid]; na = h.p; nb = h.m; }
    for (int r0 = row_begin; r0 < row_end; r0 += P1_CHUNK) {
        __syncthreads();                                   // the previous chunk has been read by every warp
        s_a[tid] = na; s_b[tid] = nb;
        __syncthreads();
        na = make_float4(0.f, 0.f, 0.f, 0.f); nb = na;
        if (r0 + P1_CHUNK + tid < row_end) { const SrcHoist1 h = sh[jb.sh_off + r0 + P1_CHUNK + tid]; na = h.p; nb = h.m; }
        const int groups = min(P1_CHUNK, row_end - r0 + 31) >> 5;
        for (int g = 0; g < groups; ++g) {
            unsigned acc0 = 0u, acc1 = 0u;
#pragma unroll UNROLL
            for (int rr = 0; rr < 32; ++rr) {
                const float4 A1 = s_a[g * 32 + rr], A2 = s_b[g * 32 + rr];          // broadcast
                acc0 = __funnelshift_l(pretest_level1(QX0, QY0, NT0, A1, A2), acc0, 1);      // row rr of the group ends at bit 31 - rr
                acc1 = __funnelshift_l(pretest_level1(QX1, QY1, NT1, A1, A2), acc1, 1);
            }
            const int left = row_end - (r0 + g * 32);      // rows beyond the job's last (zero hoists) are masked out
            if (left < 32) { const unsigned m = ~((1u << (32 - left)) - 1u); acc0 &= m; acc1 &= m; }
            const size_t word = size_t((r0 >> 5) + g) * size_t(jb.n_tgt);
            if (c0 < jb.n_tgt) out[word + c0] = acc0;
            if (c1 < jb.n_tgt) out[word + c1] = acc1;
        }
    }
}

// level 2 for one pair: false = an overlap surely misses a threshold
__device__ __forceinline__ bool pretest_level2(const float4 A1, const float4 A2, const float4 qq, const float2 ww)
{
    const float e11 = fmaf(qq.x, A1.x, fmaf(qq.y, A1.y, A1.z));
    const float e12 = fmaf(qq.x, A2.x, fmaf(qq.y, A2.y, A2.z));
    const float e21 = fmaf(qq.z, A1.x, fmaf(qq.w, A1.y, A1.z));
    const float e22 = fmaf(qq.z, A2.x, fmaf(qq.w, A2.y, A2.z));
    float Ia, Ua, Ib, Ub;
    side_bounds(e11, e12, e21, e22, ww.x, ww.y, Ia, Ua);      // along the src segment
    side_bounds(e11, e21, e12, e22, A1.w, A2.w, Ib, Ub);      // along the tgt segment
    // keep unless an overlap surely misses a threshold; NaN/inf comparisons keep
    const bool drop = (Ia < L3D_PRETEST_LO * Ua) || (Ib < L3D_PRETEST_LO * Ub) ||
                      ((Ia < L3D_PRETEST_HI * Ua) && (Ib < L3D_PRETEST_HI * Ub));
    return !drop;
}

constexpr int PT_ROWS = 128;        // level 2: source segments per block tile (4 words of level-1 bits per lane)
constexpr int PT_COLS = 256;        // target segments per block tile: 8 warps x 32 lanes
constexpr int PT_BLOCK = 256;
constexpr int PT_Q = 320;           // per-warp queue.  A word of 32 rows leaves about 43 survivors in a warp; a word
                                    // that would not fit (degenerate geometry) is evaluated without the queue

__global__ void __launch_bounds__(PT_BLOCK) k_pretest2(const SrcHoist* __restrict__ sh, const TgtHoist* __restrict__ th,
                                                       const JobDev* __restrict__ jobs, int n_jobs, const unsigned* __restrict__ l1,
                                                       unsigned* __restrict__ cand, int* __restrict__ cand_cnt)
{
    __shared__ float4 s_a[PT_ROWS], s_b[PT_ROWS];
    __shared__ float4 s_q[PT_COLS];
    __shared__ float2 s_w[PT_COLS];
    __shared__ unsigned s_bits[PT_ROWS][PT_COLS / 32];
    __shared__ unsigned short s_queue[PT_BLOCK / 32][PT_Q];
    __shared__ int s_job;
    if (threadIdx.x == 0) s_job = find_job_by_block(jobs, n_jobs, blockIdx.x);
    __syncthreads();
    const JobDev& jb = jobs[s_job];
    const int local = blockIdx.x - jb.blk_off;
    const int rb = local / jb.col_blocks, cb = local - rb * jb.col_blocks;
    const int col0 = cb * PT_COLS, row0 = rb * PT_ROWS;
    const int ncol = min(PT_COLS, jb.n_tgt - col0), nrow = min(PT_ROWS, jb.n_src - row0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the level-1 words of this lane's column: row groups row0 / 32 .. of the job
    unsigned words[PT_ROWS / 32];
    {
        const unsigned* in = l1 + jb.l1_off + size_t(row0 >> 5) * size_t(jb.n_tgt) + col0 + tid;
#pragma unroll
        for (int g = 0; g < PT_ROWS / 32; ++g) words[g] = (tid < ncol && g * 32 < nrow) ? in[size_t(g) * size_t(jb.n_tgt)] : 0u;
    }
    {
        if (tid < ncol) { TgtHoist h = th[jb.th_off + col0 + tid]; s_q[tid] = h.q; s_w[tid] = h.w; }
        else { s_q[tid] = make_float4(0.f, 0.f, 0.f, 0.f); s_w[tid] = make_float2(0.f, 0.f); }
        if (tid < PT_ROWS) {
            if (tid < nrow) { SrcHoist h = sh[jb.sh_off + row0 + tid]; s_a[tid] = h.a; s_b[tid] = h.b; }
            else { s_a[tid] = make_float4(0.f, 0.f, 0.f, 0.f); s_b[tid] = make_float4(0.f, 0.f, 0.f, 0.f); }
        }
        for (int k = tid; k < PT_ROWS * (PT_COLS / 32); k += PT_BLOCK) (&s_bits[0][0])[k] = 0u;
    }
    __syncthreads();
    unsigned short* queue = s_queue[warp];
    int qn = 0;
    auto level2 = [&](int first, int count) {
        if (lane < count) {
            const unsigned e = queue[first + lane];
            const int r = int(e >> 5), c = warp * 32 + int(e & 31u);
            if (pretest_level2(s_a[r], s_b[r], s_q[c], s_w[c])) atomicOr(&s_bits[r][warp], 1u << (e & 31u));
        }
    };
#pragma unroll
    for (int g = 0; g < PT_ROWS / 32; ++g) {
        unsigned acc = words[g];                             // bit 31 - rr = row g * 32 + rr of the tile
        const int n = __popc(acc);
        int pre = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
        const int total = __shfl_sync(0xffffffffu, pre, 31);
        if (total == 0) continue;
        if (qn + total <= PT_Q) {
            pre += qn - n;
            while (acc) { const int bit = 31 - __clz(acc); acc &= ~(1u << bit); queue[pre++] = (unsigned short)(((g * 32 + 31 - bit) << 5) | lane); }
            qn += total;
            __syncwarp();
            while (qn >= 32) { level2(qn - 32, 32); qn -= 32; }
            __syncwarp();
        } else {            // more survivors than the queue holds: every lane walks its own bits
            while (acc) {
                const int bit = 31 - __clz(acc); acc &= ~(1u << bit);
                const int r = g * 32 + 31 - bit;
                if (pretest_level2(s_a[r], s_b[r], s_q[tid], s_w[tid])) atomicOr(&s_bits[r][warp], 1u << lane);
            }
        }
    }
    if (qn > 0) level2(0, qn);
    __syncthreads();
    // write the tile's bitmask rows and count the candidates per row
    const int wpt = PT_COLS / 32;                        // 8 words per row
    for (int k = tid; k < nrow * wpt; k += PT_BLOCK) {
        const int r = k / wpt, w = k - r * wpt;
        if (col0 + w * 32 < jb.n_tgt) cand[jb.cand_off + (long long)(row0 + r) * jb.wpr + (col0 >> 5) + w] = s_bits[r][w];
    }
    if (tid < nrow) {
        int total = 0;
#pragma unroll
        for (int w = 0; w < wpt; ++w) total += __popc(s_bits[tid][w]);
        if (total) atomicAdd(&cand_cnt[jb.row_off + row0 + tid], total);
    }
}

// ------------------------------------------------------------------------------------------
// exact test.  Candidates of a row rarely fill whole warps and most of them fail the overlap test, so a
// row-per-warp kernel (the first version, see profiles/r1_v1_kernels.md) idled more than half of its lanes.
// Here every candidate is its own thread:
//   k_expand      warp per row: the row's candidate columns, in order, into a flat list (+ the row of each)
//   k_exact_flat  thread per candidate: overlap test; the block compacts the survivors through shared memory
//                 and its first threads triangulate them (idle warps cost no issue slots); results are written
//                 at the candidate's own index: ok flag, 4 depths; raw matches per row are counted with atomics
// k_fill then copies the ok candidates of a row, in order, to their final position.
// ------------------------------------------------------------------------------------------
__global__ void k_expand(const JobDev* __restrict__ jobs, const int* __restrict__ row_job, int n_rows, const unsigned* __restrict__ cand, const int* __restrict__ cand_off,
                         int* __restrict__ c_col, int* __restrict__ c_row)
{
    const int grow = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (grow >= n_rows) return;
    int o = cand_off[grow];
    if (cand_off[grow + 1] == o) return;
    const JobDev& jb = jobs[row_job[grow]];
    const unsigned* words = cand + jb.cand_off + (long long)(grow - jb.row_off) * jb.wpr;
    const int n_words = (jb.n_tgt + 31) >> 5;
    for (int w0 = 0; w0 < n_words; w0 += 32) {
        unsigned bits = (w0 + lane < n_words) ? words[w0 + lane] : 0u;
        int n = __popc(bits), ex = n;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, ex, s); if (lane >= s) ex += y; }
        const int tot = __shfl_sync(0xffffffffu, ex, 31);
        int pos = o + ex - n;
        while (bits) { int b = __ffs(bits) - 1; bits &= bits - 1; c_col[pos] = (w0 + lane) * 32 + b; c_row[pos] = grow; ++pos; }
        o += tot;
    }
}

constexpr int EF_T = 256;
__global__ void __launch_bounds__(EF_T, 6) k_exact_flat(const float4* __restrict__ segs, const JobDev* __restrict__ jobs, const int* __restrict__ row_job, int n_cand,
                                                     const int* __restrict__ c_col, const int* __restrict__ c_row,
                                                     unsigned char* __restrict__ c_ok, float4* __restrict__ c_dep, int* __restrict__ raw_cnt)
{
    __shared__ float s_is[8][EF_T];
    __shared__ int s_who[EF_T];               // candidate (block-local, low 8 bits) and job of each overlap survivor
    __shared__ int s_warp_n[EF_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x * EF_T + tid;
    bool ok1 = false; PairIsect is; int job = 0;
    if (c < n_cand) {
        const int grow = c_row[c];
        job = row_job[grow];
        const JobDev& jb = jobs[job];
        ok1 = xpair_overlap(segs[job_src_segment(jb, grow - jb.row_off)], segs[jb.tgt_off + c_col[c]], jb.F, is);
        c_ok[c] = 0;
    }
    // block-wide ordered compaction of the survivors
    const unsigned m = __ballot_sync(0xffffffffu, ok1);
    if (lane == 0) s_warp_n[warp] = __popc(m);
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < EF_T / 32; ++w) { const int n = s_warp_n[w]; if (w < warp) base += n; total += n; }
    if (ok1) {
        const int p = base + __popc(m & ((1u << lane) - 1u));
        s_who[p] = tid | (job << 8);
        s_is[0][p] = is.ax; s_is[1][p] = is.ay; s_is[2][p] = is.bx; s_is[3][p] = is.by;
        s_is[4][p] = is.cx; s_is[5][p] = is.cy; s_is[6][p] = is.dx; s_is[7][p] = is.dy;
    }
    __syncthreads();
    if (tid < total) {
        const int cc = blockIdx.x * EF_T + (s_who[tid] & (EF_T - 1));
        const int grow = c_row[cc];
        const JobDev& jb = jobs[s_who[tid] >> 8];
        PairIsect i2;
        i2.ax = s_is[0][tid]; i2.ay = s_is[1][tid]; i2.bx = s_is[2][tid]; i2.by = s_is[3][tid];
        i2.cx = s_is[4][tid]; i2.cy = s_is[5][tid]; i2.dx = s_is[6][tid]; i2.dy = s_is[7][tid];
        float4 dep;
        const bool ok = xpair_depths(segs[job_src_segment(jb, grow - jb.row_off)], segs[jb.tgt_off + c_col[cc]], i2, jb.RtKinv_s, jb.RtKinv_t,
                                     mk3(jb.C_s[0], jb.C_s[1], jb.C_s[2]), mk3(jb.C_t[0], jb.C_t[1], jb.C_t[2]), dep);
        if (ok) { c_ok[cc] = 1; c_dep[cc] = dep; atomicAdd(&raw_cnt[grow], 1); }
    }
}

// out_sum[b], out_max[b] over v[begin[b] .. begin[b] + len[b]); one block per b.  After the exact test: raw matches
// per job and the most raw matches of a single row, which bound every per-view buffer and grid.
__global__ void k_range_sum_max(const int* __restrict__ v, const int* __restrict__ begin, const int* __restrict__ len, int* __restrict__ out_sum, int* __restrict__ out_max)
{
    __shared__ int sh_s[32], sh_m[32];
    const int b0 = begin[blockIdx.x], n = len[blockIdx.x];
    int s = 0, m = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { const int x = v[b0 + i]; s += x; m = max(m, x); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); m = max(m, __shfl_xor_sync(0xffffffffu, m, o)); }
    if ((threadIdx.x & 31) == 0) { sh_s[threadIdx.x >> 5] = s; sh_m[threadIdx.x >> 5] = m; }
    __syncthreads();
    if (threadIdx.x < 32) {
        s = (threadIdx.x < (blockDim.x >> 5)) ? sh_s[threadIdx.x] : 0; m = (threadIdx.x < (blockDim.x >> 5)) ? sh_m[threadIdx.x] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); m = max(m, __shfl_xor_sync(0xffffffffu, m, o)); }
        if (threadIdx.x == 0) { out_sum[blockIdx.x] = s; out_max[blockIdx.x] = m; }
    }
}

// out[b] = max over i in [begin[b], begin[b] + len[b]) of (diff ? v[i + 1] - v[i] : v[i]); one block per b.
// Used for upper bounds on the number of matches of a single segment (which scoring kernels are needed).
__global__ void k_range_max(const int* __restrict__ v, const int* __restrict__ begin, const int* __restrict__ len, int diff, int* __restrict__ out)
{
    __shared__ int sh[32];
    const int b0 = begin[blockIdx.x], n = len[blockIdx.x];
    int m = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) m = max(m, diff ? v[b0 + i + 1] - v[b0 + i] : v[b0 + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (threadIdx.x == 0) out[blockIdx.x] = m;
    }
}

// out[b] = sum over v[begin[b] .. begin[b] + len[b]) in 64 bits (the 32-bit scan of the same values may have wrapped)
__global__ void k_range_sum64(const int* __restrict__ v, const int* __restrict__ begin, const int* __restrict__ len, unsigned long long* __restrict__ out)
{
    __shared__ unsigned long long sh[32];
    const int b0 = begin[blockIdx.x], n = len[blockIdx.x];
    unsigned long long s = 0ull;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += (unsigned long long)(unsigned)v[b0 + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) out[blockIdx.x] = s;
    }
}

// dst[k] = src[idx[k]]
__global__ void k_gather_int(const int* __restrict__ src, const int* __restrict__ idx, int n, int* __restrict__ dst)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) dst[k] = src[idx[k]];
}

// counts of the delivered records per (row, cam) for the cams in mode 1
__device__ __forceinline__ void b_gather_counts(const CamTable* __restrict__ cams, int n_cams, int n_src, const SrcTable& src, const int* __restrict__ raw_cnt, int* __restrict__ cnt)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    const int cam = blockIdx.y;
    if (row >= n_src) return;
    const CamTable& ct = cams[cam];
    int c = 0;
    if (ct.mode == 0) c = raw_cnt[ct.row_off + row];
    else if (ct.mode == 1) { const int sg = row_segment(src, row); c = ct.t_start ? (ct.t_start[sg + 1] - ct.t_start[sg]) : 0; }
    cnt[row * n_cams + cam] = c;
}

// fill: one warp per (row, cam) copies the row's records to their final place
__device__ __forceinline__ void b_fill(const CamTable* __restrict__ cams, int n_cams, int n_src, const SrcTable& src,
                       const int* __restrict__ off, const int* __restrict__ cand_off,
                       const int* __restrict__ stage_col, const float4* __restrict__ stage_dep, const unsigned char* __restrict__ stage_ok,
                       int4* __restrict__ m_ids, float4* __restrict__ m_dep)
{
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (gw >= n_src * n_cams) return;
    const int row = gw / n_cams, cam = gw - row * n_cams;
    const int o = off[gw], n = off[gw + 1] - o;
    if (n == 0) return;
    const CamTable& ct = cams[cam];
    const int sg = row_segment(src, row);            // match ids carry the segment (x) and this process's row (w)
    if (ct.mode == 0) {
        const int b = cand_off[ct.row_off + row], e = cand_off[ct.row_off + row + 1];
        int w = o;
        for (int k0 = b; k0 < e; k0 += 32) {
            const int k = k0 + lane;
            const bool ok = k < e && stage_ok[k] != 0;
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (ok) { const int r = w + __popc(m & ((1u << lane) - 1u)); m_ids[r] = make_int4(sg, cam, stage_col[k], row); m_dep[r] = stage_dep[k]; }
            w += __popc(m);
        }
    } else if (ct.mode == 1) {
        // delivered records of one segment arrive in the order of the sender's atomics; they go out in ascending
        // target id (ids are unique within a group and groups are tiny: each record counts the smaller ones)
        const int b = ct.t_start[sg];
        for (int k = lane; k < n; k += 32) {
            const int me = ct.t_tgt[b + k];
            int rank = 0;
            for (int i = 0; i < n; ++i) rank += (ct.t_tgt[b + i] < me);
            m_ids[o + rank] = make_int4(sg, cam, me, row); m_dep[o + rank] = ct.t_dep[b + k];
        }
    }
}

}  // namespace
