// line3d_b200/csrc/l3d_device.cuh -- device arithmetic of the Line3D path, "exact tier".
//
// The decisions of the path (accept / reject a segment pair, gate a hypothesis, keep a
// match) are threshold tests on FP32 values.  To reproduce the reference's decisions and
// values bit for bit, the functions here evaluate the reference's formulas in the
// reference's source order with individually rounded IEEE operations (__fmul_rn,
// __fadd_rn, __fdiv_rn, __fsqrt_rn: never contracted into FMAs, never approximated).
// They run only on the few percent of pairs that survive the cheap conservative
// pre-tests in l3d_kernels.cu, so their cost does not set the throughput.
// Citations are into /root/reference.
#pragma once
#include <cuda_runtime.h>

namespace l3d {

#define L3D_EPS_F 1e-12f           // cudawrapper.h:43
#define L3D_OVERLAP_LO 0.10f       // cudawrapper.h:45
#define L3D_OVERLAP_HI 0.30f       // cudawrapper.h:46
#define L3D_COLLIN_T 0.50f         // cudawrapper.h:44

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ float xm(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xa(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xs(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float xd(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float xsqrt(float a) { return __fsqrt_rn(a); }

// helper_math.h:1248-1251, 1420-1423, 1291-1294, 1304-1318 (+ host rsqrtf, :62-65)
__device__ __forceinline__ float xdot(V3 a, V3 b) { return xa(xa(xm(a.x, b.x), xm(a.y, b.y)), xm(a.z, b.z)); }
__device__ __forceinline__ V3 xcross(V3 a, V3 b)
{
    return mk3(xs(xm(a.y, b.z), xm(a.z, b.y)), xs(xm(a.z, b.x), xm(a.x, b.z)), xs(xm(a.x, b.y), xm(a.y, b.x)));
}
__device__ __forceinline__ V3 xsub3(V3 a, V3 b) { return mk3(xs(a.x, b.x), xs(a.y, b.y), xs(a.z, b.z)); }
__device__ __forceinline__ V3 xadd3(V3 a, V3 b) { return mk3(xa(a.x, b.x), xa(a.y, b.y), xa(a.z, b.z)); }
__device__ __forceinline__ V3 xscale(float s, V3 a) { return mk3(xm(s, a.x), xm(s, a.y), xm(s, a.z)); }
__device__ __forceinline__ float xlength(V3 v) { return xsqrt(xdot(v, v)); }
__device__ __forceinline__ V3 xnormalize(V3 v) { float inv = xd(1.0f, xsqrt(xdot(v, v))); return mk3(xm(v.x, inv), xm(v.y, inv), xm(v.z, inv)); }

// row-major 3x3 times vector, accumulator from 0, column index ascending
// (cudawrapper.cu:144-163 D_epipolar_line, 270-303 D_get_ray_src/_tgt)
__device__ __forceinline__ V3 xmat3(const float* M, V3 p, bool transpose)
{
    float v[3] = {p.x, p.y, p.z}, l[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        float acc = 0.0f;
#pragma unroll
        for (int c = 0; c < 3; ++c) acc = xa(acc, xm(transpose ? M[c * 3 + r] : M[r * 3 + c], v[c]));
        l[r] = acc;
    }
    return mk3(l[0], l[1], l[2]);
}
// cudawrapper.cu:255-267
__device__ __forceinline__ V3 xhom(V3 p)
{
    if (fabsf(p.z) > L3D_EPS_F) return mk3(xd(p.x, p.z), xd(p.y, p.z), 1.0f);
    return mk3(0.0f, 0.0f, 0.0f);
}
// cudawrapper.cu:95-99
__device__ __forceinline__ float xlen2d(V3 a, V3 b) { float vx = xs(a.x, b.x), vy = xs(a.y, b.y); return xsqrt(xa(xm(vx, vx), xm(vy, vy))); }
// cudawrapper.cu:135-141
__device__ __forceinline__ bool xon_segment(V3 p1, V3 p2, V3 q)
{
    float ax = xs(p1.x, q.x), ay = xs(p1.y, q.y), bx = xs(p2.x, q.x), by = xs(p2.y, q.y);
    return xa(xm(ax, bx), xm(ay, by)) < L3D_EPS_F;
}
// cudawrapper.cu:209-252
__device__ inline float xoverlap(V3 s1, V3 s2, V3 q1, V3 q2)
{
    float len_src = xlen2d(s1, s2), len_tgt = xlen2d(q1, q2);
    if (len_src < 1.0f || len_tgt < 1.0f) return 0.0f;
    bool q1_in = xon_segment(s1, s2, q1), q2_in = xon_segment(s1, s2, q2);
    if (q1_in && q2_in) return xd(len_tgt, len_src);
    bool s1_in = xon_segment(q1, q2, s1), s2_in = xon_segment(q1, q2, s2);
    if (s1_in && s2_in) return xd(len_src, len_tgt);
    if (q1_in) {
        float len1 = xlen2d(s2, q2), len2 = xlen2d(s1, q2);
        if (s1_in && len1 > L3D_EPS_F) return xd(xlen2d(q1, s1), len1);
        else if (len2 > L3D_EPS_F) return xd(xlen2d(q1, s2), len2);
    } else if (q2_in) {
        float len1 = xlen2d(s1, q1), len2 = xlen2d(s2, q1);
        if (s2_in && len1 > L3D_EPS_F) return xd(xlen2d(q2, s2), len1);
        else if (len2 > L3D_EPS_F) return xd(xlen2d(q2, s1), len2);
    }
    return 0.0f;
}
// cudawrapper.cu:306-335
__device__ inline float xtri_depth(V3 p1, V3 p2, V3 C1, V3 C2, const float* RtKinv1, const float* RtKinv2, bool for_src)
{
    V3 ray1 = xnormalize(xmat3(RtKinv1, p1, false));
    V3 ray2 = xnormalize(xmat3(RtKinv2, p2, false));
    V3 w0 = xsub3(C1, C2);
    float a = xdot(ray1, ray1), b = xdot(ray1, ray2), c = xdot(ray2, ray2), d = xdot(ray1, w0), e = xdot(ray2, w0);
    float denom = xs(xm(a, c), xm(b, b));
    if (fabsf(denom) > L3D_EPS_F)
        return for_src ? xd(xs(xm(b, e), xm(c, d)), denom) : xd(xs(xm(a, e), xm(b, d)), denom);
    return -1.0f;
}

// cudawrapper.cu:546-588: the overlap part of the pair test.  On acceptance the four intersection
// points (x, y of l2_p1, l2_p2, l1_q1, l1_q2) are returned for the triangulation.
struct PairIsect { float ax, ay, bx, by, cx, cy, dx, dy; };
__device__ inline bool xpair_overlap(float4 s, float4 t, const float* F, PairIsect& is)
{
    V3 p1 = mk3(s.x, s.y, 1.0f), p2 = mk3(s.z, s.w, 1.0f), line1 = xcross(p1, p2);
    V3 q1 = mk3(t.x, t.y, 1.0f), q2 = mk3(t.z, t.w, 1.0f), line2 = xcross(q1, q2);
    V3 epi_p1 = xmat3(F, p1, false), epi_p2 = xmat3(F, p2, false);
    V3 epi_q1 = xmat3(F, q1, true), epi_q2 = xmat3(F, q2, true);
    V3 l2_p1 = xhom(xcross(line2, epi_p1)), l2_p2 = xhom(xcross(line2, epi_p2));
    V3 l1_q1 = xhom(xcross(line1, epi_q1)), l1_q2 = xhom(xcross(line1, epi_q2));
    if (l2_p1.z == 0.0f || l2_p2.z == 0.0f || l1_q1.z == 0.0f || l1_q2.z == 0.0f) return false;
    float o1 = xoverlap(p1, p2, l1_q1, l1_q2), o2 = xoverlap(q1, q2, l2_p1, l2_p2);
    if (!(fminf(o1, o2) > L3D_OVERLAP_LO && fmaxf(o1, o2) > L3D_OVERLAP_HI)) return false;
    is.ax = l2_p1.x; is.ay = l2_p1.y; is.bx = l2_p2.x; is.by = l2_p2.y;
    is.cx = l1_q1.x; is.cy = l1_q1.y; is.dx = l1_q2.x; is.dy = l1_q2.y;
    return true;
}
// cudawrapper.cu:590-601 + :931: the four depths; true when all are > 0
__device__ inline bool xpair_depths(float4 s, float4 t, const PairIsect& is, const float* RtKinv_s, const float* RtKinv_t, V3 C_s, V3 C_t, float4& depths)
{
    V3 p1 = mk3(s.x, s.y, 1.0f), p2 = mk3(s.z, s.w, 1.0f), q1 = mk3(t.x, t.y, 1.0f), q2 = mk3(t.z, t.w, 1.0f);
    V3 l2_p1 = mk3(is.ax, is.ay, 1.0f), l2_p2 = mk3(is.bx, is.by, 1.0f), l1_q1 = mk3(is.cx, is.cy, 1.0f), l1_q2 = mk3(is.dx, is.dy, 1.0f);
    depths.x = xtri_depth(p1, l2_p1, C_s, C_t, RtKinv_s, RtKinv_t, true);
    depths.y = xtri_depth(p2, l2_p2, C_s, C_t, RtKinv_s, RtKinv_t, true);
    depths.z = xtri_depth(l1_q1, q1, C_s, C_t, RtKinv_s, RtKinv_t, false);
    depths.w = xtri_depth(l1_q2, q2, C_s, C_t, RtKinv_s, RtKinv_t, false);
    return depths.x > 0.0f && depths.y > 0.0f && depths.z > 0.0f && depths.w > 0.0f;
}
// cudawrapper.cu:538-611.  Returns true when all four depths are > 0 (cudawrapper.cu:931).
__device__ inline bool xpair_test(float4 s, float4 t, const float* F, const float* RtKinv_s, const float* RtKinv_t,
                                  V3 C_s, V3 C_t, float4& depths)
{
    PairIsect is;
    if (!xpair_overlap(s, t, F, is)) return false;
    return xpair_depths(s, t, is, RtKinv_s, RtKinv_t, C_s, C_t, depths);
}

// cudawrapper.cu:58-61
__device__ __forceinline__ float xdist_p2l(V3 line, V3 p)
{
    return fabsf(xd(xa(xa(xm(line.x, p.x), xm(line.y, p.y)), line.z), xsqrt(xa(xm(line.x, line.x), xm(line.y, line.y)))));
}
// cudawrapper.cu:355-377; P row-major 3x4
__device__ __forceinline__ V3 xproject(const float* P, V3 X)
{
    float v[4] = {X.x, X.y, X.z, 1.0f}, p[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        float acc = 0.0f;
#pragma unroll
        for (int c = 0; c < 4; ++c) acc = xa(acc, xm(P[r * 4 + c], v[c]));
        p[r] = acc;
    }
    if (fabsf(p[2]) > L3D_EPS_F) return mk3(xd(p[0], p[2]), xd(p[1], p[2]), 1.0f);
    return mk3(0.0f, 0.0f, 0.0f);
}

// expf / acosf exactly as the C library behind the reference's host build evaluates them
// (glibc 2.39).  The oracle and the CPU build of the reference call libm; to agree with
// them bit for bit the device runs the same two algorithms:
//   * expf: the table-driven double-precision routine glibc adopted from ARM's optimized
//     routines (N = 32 table of 2^(i/32), cubic polynomial, result rounded once to float);
//     T[i] = bits(2^(i/32)) - (i << 47), regenerated here from that definition;
//   * acosf: the fdlibm single-precision routine (rational approximation R(z) = p(z)/q(z)
//     on three ranges), every operation individually rounded in float.
// tests/test_device_math.py checks both against libm on random and edge inputs.
__device__ const unsigned long long L3D_EXP2F_TAB[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};
__device__ __forceinline__ float xexpf(float x)
{
    if (x < -103.972084f) return 0.0f;                       // underflow to zero, as libm
    const double InvLn2N = 0x1.71547652b82fep+0 * 32.0;
    const double C0 = 0x1.c6af84b912394p-5 / 32.0 / 32.0 / 32.0, C1 = 0x1.ebfce50fac4f3p-3 / 32.0 / 32.0, C2 = 0x1.62e42ff0c52d6p-1 / 32.0;
    const double SHIFT = 0x1.8p+52;
    double z = __dmul_rn(InvLn2N, double(x));
    double kd = __dadd_rn(z, SHIFT);
    unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd = __dsub_rn(kd, SHIFT);
    double r = __dsub_rn(z, kd);
    unsigned long long t = L3D_EXP2F_TAB[ki & 31ull] + (ki << 47);
    double s = __longlong_as_double((long long)t);
    double zz = __dadd_rn(__dmul_rn(C0, r), C1);
    double r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(C2, r), 1.0);
    y = __dadd_rn(__dmul_rn(zz, r2), y);
    y = __dmul_rn(y, s);
    return __double2float_rn(y);
}
__device__ __forceinline__ float xacos_pq(float z)
{
    const float pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f, pS3 = -4.0055535734e-02f,
                pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f;
    const float qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f, qS3 = -6.8828397989e-01f, qS4 = 7.7038154006e-02f;
    float p = xm(z, xa(pS0, xm(z, xa(pS1, xm(z, xa(pS2, xm(z, xa(pS3, xm(z, xa(pS4, xm(z, pS5)))))))))));
    float q = xa(1.0f, xm(z, xa(qS1, xm(z, xa(qS2, xm(z, xa(qS3, xm(z, qS4))))))));
    return xd(p, q);
}
__device__ inline float xacosf(float x)
{
    const float pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f, pi = 3.1415925026e+00f;
    const unsigned ix = __float_as_uint(x) & 0x7fffffffu;
    if (ix == 0x3f800000u) return (x > 0.0f) ? 0.0f : xa(pi, xm(2.0f, pio2_lo));
    if (ix > 0x3f800000u) return __int_as_float(0x7fc00000);
    if (ix < 0x3f000000u) {                                   // |x| < 0.5
        if (ix <= 0x32800000u) return xa(pio2_hi, pio2_lo);
        float r = xacos_pq(xm(x, x));
        return xs(pio2_hi, xs(x, xs(pio2_lo, xm(x, r))));
    }
    if (x < 0.0f) {                                           // x < -0.5
        float z = xm(xa(1.0f, x), 0.5f);
        float r = xacos_pq(z), s = xsqrt(z);
        float w = xs(xm(r, s), pio2_lo);
        return xs(pi, xm(2.0f, xa(s, w)));
    }
    float z = xm(xs(1.0f, x), 0.5f);                          // x > 0.5
    float s = xsqrt(z);
    float df = __uint_as_float(__float_as_uint(s) & 0xfffff000u);
    float c = xd(xs(z, xm(df, df)), xa(s, df));
    float r = xacos_pq(z);
    float w = xa(xm(r, s), c);
    return xm(2.0f, xa(df, w));
}

// cudawrapper.cu:118-130
__device__ inline float xangle_deg_3d(V3 P1, V3 P2, V3 Q1, V3 Q2)
{
    V3 v1 = xnormalize(xsub3(P1, P2)), v2 = xnormalize(xsub3(Q1, Q2));
    float d = xdot(v1, v2);
    float cl = fmaxf(fminf(d, 1.0f), -1.0f);
    float angle = float(double(xacosf(cl)) / 3.1415926535897931e+0 * 180.0);
    if (angle > 90.0f) angle = xs(180.0f, angle);
    return angle;
}
// cudawrapper.cu:380-427
__device__ inline float xhypothesis_confidence(V3 p1, V3 p2, V3 P1, V3 P2, V3 Q1, V3 Q2, V3 C, float4 tgt,
                                               float sigma_p, float sigma_a, float spatial_k)
{
    if (spatial_k > 0.0f) {
        float depth1 = xlength(xsub3(C, P1)), depth2 = xlength(xsub3(C, P2));
        float unc1 = xm(spatial_k, depth1), unc2 = xm(spatial_k, depth2);
        float dist1 = xlength(xsub3(P1, Q1)), dist2 = xlength(xsub3(P2, Q2));
        if (dist1 > unc1 || dist2 > unc2) return 0.0f;
    }
    V3 line1 = xcross(p1, p2);
    V3 q1 = mk3(tgt.x, tgt.y, 1.0f), q2 = mk3(tgt.z, tgt.w, 1.0f), line2 = xcross(q1, q2);
    float d1 = fmaxf(xdist_p2l(line2, p1), xdist_p2l(line2, p2));
    float d2 = fmaxf(xdist_p2l(line1, q1), xdist_p2l(line1, q2));
    float dist = fmaxf(d1, d2);
    float sigma_sqr_a = xm(sigma_a, sigma_a), sigma_sqr_d = xm(sigma_p, sigma_p);
    // callers only use the value when it exceeds 0.5 (cudawrapper.cu:699); exp(-0.7) < 0.4966,
    // so an exponent below -0.7 can never get there: the angle and the exponentials are skipped
    float arg_d = xd(xm(-dist, dist), xm(2.0f, sigma_sqr_d));
    if (arg_d < -0.7f) return 0.0f;
    float angle = xangle_deg_3d(P1, P2, Q1, Q2);
    float arg_a = xd(xm(-angle, angle), xm(2.0f, sigma_sqr_a));
    if (arg_a < -0.7f) return 0.0f;
    return fminf(xexpf(arg_d), xexpf(arg_a));
}

// cudawrapper.cu:476-535 for one unordered pair
__device__ inline float xcollinearity(float4 a, float4 b, float sigma_sqr)
{
    V3 p1 = mk3(a.x, a.y, 1.0f), p2 = mk3(a.z, a.w, 1.0f), line1 = xcross(p1, p2);
    V3 q1 = mk3(b.x, b.y, 1.0f), q2 = mk3(b.z, b.w, 1.0f), line2 = xcross(q1, q2);
    float d1 = fmaxf(xdist_p2l(line2, p1), xdist_p2l(line2, p2));
    float d2 = fmaxf(xdist_p2l(line1, q1), xdist_p2l(line1, q2));
    float d = fmaxf(d1, d2);
    float arg = xd(xm(-d, d), xm(2.0f, sigma_sqr));
    if (arg < -0.7f) return 0.0f;                 // exp(-0.7) < 0.4966 < L3D_COLLIN_T: cannot pass
    float aff = xexpf(arg);
    if (!(aff > L3D_COLLIN_T)) return 0.0f;
    float pos1 = xa(xm(xs(q1.x, p1.x), xs(q2.x, p1.x)), xm(xs(q1.y, p1.y), xs(q2.y, p1.y)));
    float pos2 = xa(xm(xs(q1.x, p2.x), xs(q2.x, p2.x)), xm(xs(q1.y, p2.y), xs(q2.y, p2.y)));
    float pos3 = xa(xm(xs(p1.x, q1.x), xs(p2.x, q1.x)), xm(xs(p1.y, q1.y), xs(p2.y, q1.y)));
    float pos4 = xa(xm(xs(p1.x, q2.x), xs(p2.x, q2.x)), xm(xs(p1.y, q2.y), xs(p2.y, q2.y)));
    if (pos1 > -L3D_EPS_F && pos2 > -L3D_EPS_F && pos3 > -L3D_EPS_F && pos4 > -L3D_EPS_F) return aff;
    return 0.0f;
}

}  // namespace l3d
