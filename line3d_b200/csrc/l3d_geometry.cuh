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

// ------------------------------------------------------------------------------------------
// 3-D line of a cluster (line3D.cc:1306-1368, 1392-1451, 1479-1597).  The cluster's hypotheses, in
// (camera, segment) order, are taken back to world coordinates; the line goes through their centroid along the
// dominant direction of their scatter matrix; the end points are swept along it and every stretch seen by at least
// three cameras becomes a 3-D segment.  Everything in FP64 in the host code's order of operations -- the scatter matrix
// as the product g (I - 1/n) g^T it is written as, entry by entry, and the one-sided Jacobi SVD of the 3x3 result.
// ------------------------------------------------------------------------------------------
GEO_HD double geo_dot3(const double* a, const double* b)
{
    double s = 0.0;
    for (int i = 0; i < 3; ++i) s = GEO_ADD(s, GEO_MUL(a[i], b[i]));
    return s;
}
// line3D.cc:1782-1786: R^-1 (P / scale - t)
GEO_HD void geo_inverse_transform(const l3d_similarity& tr, const double* P, double* out)
{
    double v[3];
    for (int i = 0; i < 3; ++i) v[i] = GEO_ADD(GEO_MUL(P[i], tr.scale_inv), tr.tneg[i]);
    geo_mul31(tr.Rinv, v, out);
}
// one-sided Jacobi on the columns of a 3x3 matrix (row-major a); returns the left singular vector of the largest
// singular value (the first one among equals)
GEO_HD void geo_dominant_direction(const double* a_in, double* dir)
{
    double A[9];
    for (int i = 0; i < 9; ++i) A[i] = a_in[i];
#define A_(i, j) A[(i) * 3 + (j)]
    for (int sweep = 0; sweep < 60; ++sweep) {
        bool rotated = false;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double alpha = 0.0, beta = 0.0, gamma = 0.0;
                for (int i = 0; i < 3; ++i) {
                    alpha = GEO_ADD(alpha, GEO_MUL(A_(i, p), A_(i, p))); beta = GEO_ADD(beta, GEO_MUL(A_(i, q), A_(i, q))); gamma = GEO_ADD(gamma, GEO_MUL(A_(i, p), A_(i, q)));
                }
                if (gamma == 0.0 || fabs(gamma) <= GEO_MUL(1e-15, GEO_SQRT(GEO_MUL(alpha, beta)))) continue;
                rotated = true;
                const double zeta = GEO_DIV(GEO_SUB(beta, alpha), GEO_MUL(2.0, gamma));
                const double t = GEO_DIV(zeta >= 0 ? 1.0 : -1.0, GEO_ADD(fabs(zeta), GEO_SQRT(GEO_ADD(1.0, GEO_MUL(zeta, zeta)))));
                const double c = GEO_DIV(1.0, GEO_SQRT(GEO_ADD(1.0, GEO_MUL(t, t)))), sn = GEO_MUL(c, t);
                for (int i = 0; i < 3; ++i) {
                    const double ap = A_(i, p), aq = A_(i, q);
                    A_(i, p) = GEO_SUB(GEO_MUL(c, ap), GEO_MUL(sn, aq)); A_(i, q) = GEO_ADD(GEO_MUL(sn, ap), GEO_MUL(c, aq));
                }
            }
        if (!rotated) break;
    }
    double sv[3];
    for (int j = 0; j < 3; ++j) { double s = 0.0; for (int i = 0; i < 3; ++i) s = GEO_ADD(s, GEO_MUL(A_(i, j), A_(i, j))); sv[j] = GEO_SQRT(s); }
    // the singular values in descending order (stable), then the first maximum of that list: the first column with the largest value
    int best = 0;
    for (int j = 1; j < 3; ++j) if (sv[j] > sv[best]) best = j;
    // its rank among the columns decides the unit vector a zero singular value gets (all three are zero then: rank 0)
    for (int i = 0; i < 3; ++i) dir[i] = sv[best] > 0.0 ? GEO_DIV(A_(i, best), sv[best]) : (i == 0 ? 1.0 : 0.0);
#undef A_
}

// One cluster in steps, so that the device can spread the independent entries of a step over a block while the host
// build runs the same functions one after the other.  Points: 2m x 3, P1 and P2 of the m hypotheses alternating.
GEO_HD void geo_fit_transform(const l3d_similarity& tr, const double* pts_in, double* A, int k)
{
    geo_inverse_transform(tr, pts_in + size_t(k) * 3, A + size_t(k) * 3);
}
GEO_HD void geo_fit_centroid(const double* A, int n, double* P)          // sequential: the sum takes the points in order
{
    double szd = 0.0;
    P[0] = P[1] = P[2] = 0.0;
    for (int k = 0; k < n; k += 2) {
        for (int i = 0; i < 3; ++i) P[i] = GEO_ADD(P[i], A[size_t(k) * 3 + i]);
        for (int i = 0; i < 3; ++i) P[i] = GEO_ADD(P[i], A[size_t(k + 1) * 3 + i]);
        szd = GEO_ADD(szd, 2.0);
    }
    for (int i = 0; i < 3; ++i) P[i] = GEO_DIV(P[i], szd);
}
// entry (i, j) of g Cm, g = the points as columns, Cm = I - (1/n) 1 1^T
GEO_HD double geo_fit_centered(const double* A, int n, int j, int i)
{
    const double inv_n = GEO_MUL(1.0, GEO_DIV(1.0, double(n)));
    const double c_diag = GEO_SUB(1.0, inv_n), c_off = GEO_SUB(0.0, inv_n);
    double s = 0.0;
    for (int k = 0; k < n; ++k) s = GEO_ADD(s, GEO_MUL(A[size_t(k) * 3 + i], k == j ? c_diag : c_off));
    return s;
}
// entry (i, l) of (g Cm) g^T
GEO_HD double geo_fit_scatter(const double* G, const double* A, int n, int i, int l)
{
    double s = 0.0;
    for (int j = 0; j < n; ++j) s = GEO_ADD(s, GEO_MUL(G[size_t(j) * 3 + i], A[size_t(j) * 3 + l]));
    return s;
}
// direction of the line and its end towards -dir: the projection of an end point that lies farthest that way (sequential)
GEO_HD void geo_fit_line(const double* scat, const double* A, const double* P, int n, double* min_point)
{
    double dir[3];
    geo_dominant_direction(scat, dir);
    { const double nn = GEO_SQRT(geo_dot3(dir, dir)); for (int i = 0; i < 3; ++i) dir[i] = GEO_DIV(dir[i], nn); }
    double min_len = 0.0;
    min_point[0] = min_point[1] = min_point[2] = 0.0;
    for (int k = 0; k < n; ++k) {
        double d[3], pr[3], e[3];
        for (int i = 0; i < 3; ++i) d[i] = GEO_SUB(A[size_t(k) * 3 + i], P[i]);
        const double nrm = GEO_SQRT(geo_dot3(dir, dir));
        const double q = GEO_DIV(geo_dot3(dir, d), GEO_MUL(nrm, nrm));
        for (int i = 0; i < 3; ++i) pr[i] = GEO_ADD(P[i], GEO_MUL(dir[i], q));
        for (int i = 0; i < 3; ++i) e[i] = GEO_SUB(P[i], pr[i]);
        const double loc = geo_dot3(dir, e);
        if (loc <= min_len) { min_len = loc; for (int i = 0; i < 3; ++i) min_point[i] = pr[i]; }
    }
}
GEO_HD float geo_fit_dist(const double* A, const double* min_point, int k)
{
    double d[3];
    for (int i = 0; i < 3; ++i) d[i] = GEO_SUB(A[size_t(k) * 3 + i], min_point[i]);
    return float(GEO_SQRT(geo_dot3(d, d)));
}
// where point e stands in the stable ascending order by distance
GEO_HD int geo_fit_rank(const float* dist, int n, int e)
{
    const float d = dist[e];
    int r = 0;
    for (int k = 0; k < n; ++k) r += (dist[k] < d || (dist[k] == d && k < e)) ? 1 : 0;
    return r;
}
// sweep (sequential): a hypothesis is open between its two end points; a stretch with at least three cameras open is a segment
GEO_HD int geo_fit_sweep(const double* A, const int* order, const unsigned* cam, int m, unsigned char* open_line, int* cam_open, double* seg_out)
{
    const int n = 2 * m;
    for (int k = 0; k < m; ++k) { open_line[k] = 0; cam_open[k] = 0; }
    int cams_open = 0, n_seg = 0; bool opened = false; double start[3] = {0.0, 0.0, 0.0};
    for (int r = 0; r < n; ++r) {
        const int pt = order[r], k = pt >> 1;
        // the counter of a camera is kept at the first hypothesis of that camera (members are in camera order)
        int c0 = k; while (c0 > 0 && cam[c0 - 1] == cam[k]) --c0;
        if (!open_line[k]) { open_line[k] = 1; if (cam_open[c0]++ == 0) ++cams_open; }
        else { open_line[k] = 0; if (--cam_open[c0] == 0) --cams_open; }
        if (opened && cams_open < 3) {
            for (int i = 0; i < 3; ++i) { seg_out[size_t(n_seg) * 6 + i] = start[i]; seg_out[size_t(n_seg) * 6 + 3 + i] = A[size_t(pt) * 3 + i]; }
            ++n_seg; opened = false;
        } else if (!opened && cams_open >= 3) { for (int i = 0; i < 3; ++i) start[i] = A[size_t(pt) * 3 + i]; opened = true; }
    }
    return n_seg;
}

// The whole cluster, one step after the other (host build; the kernel below runs the same steps on a block).
// cam: the camera of each hypothesis (members in (camera, segment) order).  Scratch: A (2m x 3), G (2m x 3), order (2m),
// dist (2m), open_line (m), cam_open (m).  Output: up to m segments of 6 doubles; returns how many.
GEO_HD int geo_fit_cluster(const l3d_similarity& tr, int m, const double* pts_in, const unsigned* cam, double* A, double* G, int* order, float* dist,
                           unsigned char* open_line, int* cam_open, double* seg_out)
{
    const int n = 2 * m;
    if (m <= 0) return 0;
    double P[3], scat[9], min_point[3];
    for (int k = 0; k < n; ++k) geo_fit_transform(tr, pts_in, A, k);
    geo_fit_centroid(A, n, P);
    for (int j = 0; j < n; ++j) for (int i = 0; i < 3; ++i) G[size_t(j) * 3 + i] = geo_fit_centered(A, n, j, i);
    for (int i = 0; i < 3; ++i) for (int l = 0; l < 3; ++l) scat[i * 3 + l] = geo_fit_scatter(G, A, n, i, l);
    geo_fit_line(scat, A, P, n, min_point);
    for (int k = 0; k < n; ++k) dist[k] = geo_fit_dist(A, min_point, k);
    for (int k = 0; k < n; ++k) order[geo_fit_rank(dist, n, k)] = k;
    return geo_fit_sweep(A, order, cam, m, open_line, cam_open, seg_out);
}

// one block per cluster; the points of a hypothesis come from the resident array of l3d_unproject
constexpr int FIT_T = 128;
__global__ void __launch_bounds__(FIT_T) k_fit_lines(l3d_similarity tr, int n_clusters, const int* __restrict__ start, const int* __restrict__ member,
                            const unsigned* __restrict__ cam, const Hyp* __restrict__ hyp, double* __restrict__ pts_all, double* __restrict__ A_all, double* __restrict__ G_all,
                            int* __restrict__ order_all, float* __restrict__ dist_all, unsigned char* __restrict__ open_line, int* __restrict__ cam_open,
                            int* __restrict__ seg_count, double* __restrict__ seg_out)
{
    const int c = blockIdx.x;
    if (c >= n_clusters) return;
    const int b = start[c], m = start[c + 1] - b, n = 2 * m, t = threadIdx.x;
    if (m <= 0) { if (t == 0) seg_count[c] = 0; return; }
    double* pts = pts_all + size_t(b) * 6; double* A = A_all + size_t(b) * 6; double* G = G_all + size_t(b) * 6;
    int* order = order_all + size_t(b) * 2; float* dist = dist_all + size_t(b) * 2;
    __shared__ double sP[3], sScat[9], sMin[3];
    for (int k = t; k < m; k += FIT_T) {
        const Hyp& h = hyp[member[b + k]];
        for (int i = 0; i < 3; ++i) { pts[size_t(2 * k) * 3 + i] = h.P1[i]; pts[size_t(2 * k + 1) * 3 + i] = h.P2[i]; }
        geo_fit_transform(tr, pts, A, 2 * k); geo_fit_transform(tr, pts, A, 2 * k + 1);
    }
    __syncthreads();
    if (t == 0) geo_fit_centroid(A, n, sP);
    for (int e = t; e < 3 * n; e += FIT_T) G[e] = geo_fit_centered(A, n, e / 3, e % 3);
    __syncthreads();
    if (t < 9) sScat[t] = geo_fit_scatter(G, A, n, t / 3, t % 3);
    __syncthreads();
    if (t == 0) geo_fit_line(sScat, A, sP, n, sMin);
    __syncthreads();
    for (int k = t; k < n; k += FIT_T) dist[k] = geo_fit_dist(A, sMin, k);
    __syncthreads();
    for (int k = t; k < n; k += FIT_T) order[geo_fit_rank(dist, n, k)] = k;
    __syncthreads();
    if (t == 0) seg_count[c] = geo_fit_sweep(A, order, cam + b, m, open_line + b, cam_open + b, seg_out + size_t(b) * 6);
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
