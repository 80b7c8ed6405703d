// line3d_b200/csrc/l3d_geometry.cuh -- the preparation in front of phase A: which views a view is matched with
// (line3D.cc:476-549) and the fundamental matrix of a view pair (line3D.cc:1949-1993), FP64.
//
// One source, two targets.  The kernels (l3d_visual_neighbors, l3d_fundamentals) run these functions on the device with
// individually rounded operations (__dmul_rn ...: never contracted into FMAs); compiled for the host the same functions
// use plain IEEE double arithmetic, which is what the reference's host code does (Eigen products accumulate from zero
// with ascending inner index; the 3x3 inverse is by cofactors times the reciprocal determinant).  The host build is
// only reachable through l3d_debug_*_host and exists so that the arithmetic can be pinned against the reference's
// numbers without a GPU (tests/test_geometry_host.py); the product calls the kernels.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

#ifdef __CUDA_ARCH__
#define GEO_MUL(a, b) __dmul_rn((a), (b))
#define GEO_ADD(a, b) __dadd_rn((a), (b))
#define GEO_SUB(a, b) __dsub_rn((a), (b))
#define GEO_DIV(a, b) __ddiv_rn((a), (b))
#define GEO_SQRT(a) __dsqrt_rn((a))
#else
#define GEO_MUL(a, b) ((a) * (b))
#define GEO_ADD(a, b) ((a) + (b))
#define GEO_SUB(a, b) ((a) - (b))
#define GEO_DIV(a, b) ((a) / (b))
#define GEO_SQRT(a) std::sqrt((a))
#endif
#define GEO_HD __host__ __device__ inline

// r = a b, row-major 3x3: every entry a sum that starts at zero and takes the products in ascending inner index
GEO_HD void geo_mul33(const double* a, const double* b, double* r)
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s = GEO_ADD(s, GEO_MUL(a[i * 3 + k], b[k * 3 + j]));
            r[i * 3 + j] = s;
        }
}
GEO_HD void geo_mul31(const double* a, const double* v, double* r)
{
    for (int i = 0; i < 3; ++i) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s = GEO_ADD(s, GEO_MUL(a[i * 3 + k], v[k]));
        r[i] = s;
    }
}
GEO_HD void geo_transpose33(const double* a, double* r)
{
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r[i * 3 + j] = a[j * 3 + i];
}
// cofactors times 1 / det, the determinant expanded along the first row
GEO_HD void geo_inverse33(const double* m, double* r)
{
#define M_(i, j) m[(i) * 3 + (j)]
    const double c00 = GEO_SUB(GEO_MUL(M_(1, 1), M_(2, 2)), GEO_MUL(M_(1, 2), M_(2, 1)));
    const double c01 = GEO_SUB(GEO_MUL(M_(1, 2), M_(2, 0)), GEO_MUL(M_(1, 0), M_(2, 2)));
    const double c02 = GEO_SUB(GEO_MUL(M_(1, 0), M_(2, 1)), GEO_MUL(M_(1, 1), M_(2, 0)));
    const double det = GEO_ADD(GEO_ADD(GEO_MUL(M_(0, 0), c00), GEO_MUL(M_(0, 1), c01)), GEO_MUL(M_(0, 2), c02));
    const double id = GEO_DIV(1.0, det);
    r[0] = GEO_MUL(c00, id); r[3] = GEO_MUL(c01, id); r[6] = GEO_MUL(c02, id);
    r[1] = GEO_MUL(GEO_SUB(GEO_MUL(M_(0, 2), M_(2, 1)), GEO_MUL(M_(0, 1), M_(2, 2))), id);
    r[4] = GEO_MUL(GEO_SUB(GEO_MUL(M_(0, 0), M_(2, 2)), GEO_MUL(M_(0, 2), M_(2, 0))), id);
    r[7] = GEO_MUL(GEO_SUB(GEO_MUL(M_(0, 1), M_(2, 0)), GEO_MUL(M_(0, 0), M_(2, 1))), id);
    r[2] = GEO_MUL(GEO_SUB(GEO_MUL(M_(0, 1), M_(1, 2)), GEO_MUL(M_(0, 2), M_(1, 1))), id);
    r[5] = GEO_MUL(GEO_SUB(GEO_MUL(M_(0, 2), M_(1, 0)), GEO_MUL(M_(0, 0), M_(1, 2))), id);
    r[8] = GEO_MUL(GEO_SUB(GEO_MUL(M_(0, 0), M_(1, 1)), GEO_MUL(M_(0, 1), M_(1, 0))), id);
#undef M_
}

// line3D.cc:1968-1993: R = R2 R1^T, t = t2 - R t1, F = K2^-T [t]x R K1^-1 (1 = source view, 2 = neighbour)
GEO_HD void geo_fundamental(const l3d_camera& v1, const l3d_camera& v2, double* F)
{
    double R1t[9], R[9], Rt1[3], t[3], T[9], E[9], K2t[9], K2ti[9], K1i[9], A[9];
    geo_transpose33(v1.R, R1t);
    geo_mul33(v2.R, R1t, R);
    geo_mul31(R, v1.t, Rt1);
    for (int i = 0; i < 3; ++i) t[i] = GEO_SUB(v2.t[i], Rt1[i]);
    T[0] = 0.0;   T[1] = -t[2]; T[2] = t[1];
    T[3] = t[2];  T[4] = 0.0;   T[5] = -t[0];
    T[6] = -t[1]; T[7] = t[0];  T[8] = 0.0;
    geo_mul33(T, R, E);
    geo_transpose33(v2.K, K2t);
    geo_inverse33(K2t, K2ti);
    geo_inverse33(v1.K, K1i);
    geo_mul33(K2ti, E, A);
    geo_mul33(A, K1i, F);
}

// distance of two camera centres as the host compares it with the minimum baseline: the FP64 norm, rounded to float
GEO_HD float geo_baseline(const double* a, const double* b)
{
    double s = 0.0;
    for (int i = 0; i < 3; ++i) { const double d = GEO_SUB(a[i], b[i]); s = GEO_ADD(s, GEO_MUL(d, d)); }
    return float(GEO_SQRT(s));
}

// line3D.cc:498-546 for one view.  other/similarity: the m co-visible views in ascending order.  A view is kept if it is
// farther than min_baseline from `self` and from every view kept before it; the kept ones are ranked by similarity
// (descending, stable), cut to k (k <= 0: all) and returned in ascending order.  ids/sims: scratch and output, m entries.
GEO_HD int geo_select_neighbors(const l3d_camera* cams, int self, int m, const int* other, const float* similarity,
                                float min_baseline, int k, int* ids, float* sims)
{
    int n = 0;
    for (int q = 0; q < m; ++q) {
        const int o = other[q];
        if (!(geo_baseline(cams[self].C, cams[o].C) > min_baseline)) continue;
        bool ok = true;
        for (int e = 0; e < n && ok; ++e) if (geo_baseline(cams[ids[e]].C, cams[o].C) <= min_baseline) ok = false;
        if (ok) { ids[n] = o; sims[n] = similarity[q]; ++n; }
    }
    // stable insertion sort by similarity, descending
    for (int i = 1; i < n; ++i) {
        const int id = ids[i]; const float s = sims[i];
        int j = i - 1;
        while (j >= 0 && sims[j] < s) { ids[j + 1] = ids[j]; sims[j + 1] = sims[j]; --j; }
        ids[j + 1] = id; sims[j + 1] = s;
    }
    if (k > 0 && n > k) n = k;
    for (int i = 1; i < n; ++i) {                    // ascending ids
        const int id = ids[i]; const float s = sims[i];
        int j = i - 1;
        while (j >= 0 && ids[j] > id) { ids[j + 1] = ids[j]; sims[j + 1] = sims[j]; --j; }
        ids[j + 1] = id; sims[j + 1] = s;
    }
    return n;
}

__global__ void k_fundamentals(const l3d_camera* __restrict__ cams, const int2* __restrict__ pairs, size_t n, double* __restrict__ F)
{
    const size_t p = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= n) return;
    double f[9];
    geo_fundamental(cams[pairs[p].x], cams[pairs[p].y], f);
    for (int i = 0; i < 9; ++i) F[p * 9 + i] = f[i];
}
// one thread per view: the selection is sequential in the co-visible views (each is tested against those kept so far)
__global__ void k_visual_neighbors(const l3d_camera* __restrict__ cams, int n_views, const int* __restrict__ start, const int* __restrict__ other,
                                   const float* __restrict__ similarity, float min_baseline, int k, int* __restrict__ count, int* __restrict__ ids, float* __restrict__ sims)
{
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n_views) return;
    const int b = start[v];
    count[v] = geo_select_neighbors(cams, v, start[v + 1] - b, other + b, similarity + b, min_baseline, k, ids + b, sims + b);
}

}  // namespace
