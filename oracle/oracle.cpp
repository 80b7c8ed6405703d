// oracle/oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see oracle.h).
//
// CPU restatement of the reference's matching-and-scoring path.  Every function cites the
// reference lines it follows (paths are relative to /root/reference).  FP32 arithmetic is
// written in the reference's source order and compiled with -ffp-contract=off, so that it
// is bit-identical to the reference's own sources built for the CPU (oracle/_ref); libm's
// expf/acosf/logf/acos are used exactly where the reference's host build would call them.
// Third-party arithmetic (Eigen 3x3 inverse / products / JacobiSVD) comes from
// compat/eigen3 because Eigen itself is absent from the reference tree and this image.
#include "oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <list>
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../compat/eigen3/Eigen/Eigen"

namespace {

// ---------------------------------------------------------------------------------------
// float3 arithmetic -- helper_math.h:1248-1251 (dot), 1291-1294 (length), 1304-1318
// (normalize = v * rsqrtf(dot), host rsqrtf = 1/sqrtf: helper_math.h:62-65), 1420-1423 (cross)
// ---------------------------------------------------------------------------------------
struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };
inline f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
inline f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
inline f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(f3 v) { return sqrtf(dot(v, v)); }
inline f3 normalize(f3 v) { float inv = 1.0f / sqrtf(dot(v, v)); return v * inv; }

const float EPS_G = 1e-12;            // cudawrapper.h:43
const float COLLIN_AFF_T = 0.50f;     // cudawrapper.h:44
const float MIN_OVERLAP_LOWER = 0.10f;// cudawrapper.h:45
const float MIN_OVERLAP_UPPER = 0.30f;// cudawrapper.h:46
const int RDD_ITER = 10;              // cudawrapper.h:35
const float MIN_AFFINITY = 0.25f;     // commons.h:64
const float COLLINEARITY_S = 2.0f;    // commons.h:48

// cudawrapper.cu:58-61
inline float dist_p2l_2d(f3 line, f3 p) { return fabsf((line.x * p.x + line.y * p.y + line.z) / sqrtf(line.x * line.x + line.y * line.y)); }
// cudawrapper.cu:95-99
inline float seg_len_2d(f3 p1, f3 p2) { f3 v = p1 - p2; return sqrtf(v.x * v.x + v.y * v.y); }
// cudawrapper.cu:135-141
inline bool point_on_segment(f3 p1, f3 p2, f3 q)
{
    float ax = p1.x - q.x, ay = p1.y - q.y, bx = p2.x - q.x, by = p2.y - q.y;
    return (ax * bx + ay * by) < EPS_G;
}
// cudawrapper.cu:118-130 -- acos(float) is the float overload; the division by the double
// constant CUDART_PI promotes the degree conversion to double
inline float angle_deg_3d(f3 P1, f3 P2, f3 Q1, f3 Q2)
{
    f3 v1 = normalize(P1 - P2), v2 = normalize(Q1 - Q2);
    float d = dot(v1, v2);
    float cl = fmaxf(fminf(d, 1.0f), -1.0f);
    float angle = float(double(acosf(cl)) / 3.1415926535897931e+0 * double(180.0f));
    if (angle > 90.0f) angle = 180.0f - angle;
    return angle;
}
// M[r][c] row-major 3x3 times p, accumulator from 0, c ascending (cudawrapper.cu:144-163, 270-303)
inline f3 mat3_mul(const float* M, f3 p, bool transpose)
{
    float v[3] = {p.x, p.y, p.z}, l[3] = {0.0f, 0.0f, 0.0f};
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c)
            l[r] += (transpose ? M[c * 3 + r] : M[r * 3 + c]) * v[c];
    return mk3(l[0], l[1], l[2]);
}
// cudawrapper.cu:255-267
inline f3 normalize_hom(f3 p)
{
    if (fabsf(p.z) > EPS_G) { p.x /= p.z; p.y /= p.z; p.z /= p.z; p.z = 1; return p; }
    return mk3(0, 0, 0);
}
// cudawrapper.cu:209-252
float segment_overlap(f3 s1, f3 s2, f3 q1, f3 q2)
{
    float len_src = seg_len_2d(s1, s2), len_tgt = seg_len_2d(q1, q2);
    if (len_src < 1.0f || len_tgt < 1.0f) return 0.0f;
    if (point_on_segment(s1, s2, q1) && point_on_segment(s1, s2, q2)) return len_tgt / len_src;
    else if (point_on_segment(q1, q2, s1) && point_on_segment(q1, q2, s2)) return len_src / len_tgt;
    else if (point_on_segment(s1, s2, q1)) {
        float len1 = seg_len_2d(s2, q2), len2 = seg_len_2d(s1, q2);
        if (point_on_segment(q1, q2, s1) && len1 > EPS_G) return seg_len_2d(q1, s1) / len1;
        else if (len2 > EPS_G) return seg_len_2d(q1, s2) / len2;
    } else if (point_on_segment(s1, s2, q2)) {
        float len1 = seg_len_2d(s1, q1), len2 = seg_len_2d(s2, q1);
        if (point_on_segment(q1, q2, s2) && len1 > EPS_G) return seg_len_2d(q2, s2) / len1;
        else if (len2 > EPS_G) return seg_len_2d(q2, s1) / len2;
    }
    return 0.0f;
}
// cudawrapper.cu:306-335
float triangulation_depth(f3 p1, f3 p2, f3 C1, f3 C2, const float* RtKinv1, const float* RtKinv2, bool for_src)
{
    f3 ray1 = normalize(mat3_mul(RtKinv1, p1, false));
    f3 ray2 = normalize(mat3_mul(RtKinv2, p2, false));
    f3 w0 = C1 - C2;
    float a = dot(ray1, ray1), b = dot(ray1, ray2), c = dot(ray2, ray2), d = dot(ray1, w0), e = dot(ray2, w0);
    float denom = a * c - b * b;
    if (fabsf(denom) > EPS_G) return for_src ? (b * e - c * d) / denom : (a * e - b * d) / denom;
    return -1.0f;
}

// cudawrapper.cu:538-611; ov receives the two overlap values (-1 when an intersection is invalid)
bool pair_test(f4 s, f4 t, const float* F, const float* RtKinv_s, const float* RtKinv_t, f3 C_s, f3 C_t, f4& result, float* ov)
{
    result.x = result.y = result.z = result.w = 0.0f;
    f3 p1 = mk3(s.x, s.y, 1.0f), p2 = mk3(s.z, s.w, 1.0f), line1 = cross(p1, p2);
    f3 q1 = mk3(t.x, t.y, 1.0f), q2 = mk3(t.z, t.w, 1.0f), line2 = cross(q1, q2);
    f3 epi_p1 = mat3_mul(F, p1, false), epi_p2 = mat3_mul(F, p2, false);
    f3 epi_q1 = mat3_mul(F, q1, true), epi_q2 = mat3_mul(F, q2, true);
    f3 l2_p1 = normalize_hom(cross(line2, epi_p1)), l2_p2 = normalize_hom(cross(line2, epi_p2));
    f3 l1_q1 = normalize_hom(cross(line1, epi_q1)), l1_q2 = normalize_hom(cross(line1, epi_q2));
    if (int(l2_p1.z) == 0 || int(l2_p2.z) == 0 || int(l1_q1.z) == 0 || int(l1_q2.z) == 0) { if (ov) ov[0] = ov[1] = -1.0f; return false; }
    float o1 = segment_overlap(p1, p2, l1_q1, l1_q2), o2 = segment_overlap(q1, q2, l2_p1, l2_p2);
    if (ov) { ov[0] = o1; ov[1] = o2; }
    if (fminf(o1, o2) > MIN_OVERLAP_LOWER && fmaxf(o1, o2) > MIN_OVERLAP_UPPER) {
        result.x = triangulation_depth(p1, l2_p1, C_s, C_t, RtKinv_s, RtKinv_t, true);
        result.y = triangulation_depth(p2, l2_p2, C_s, C_t, RtKinv_s, RtKinv_t, true);
        result.z = triangulation_depth(l1_q1, q1, C_s, C_t, RtKinv_s, RtKinv_t, false);
        result.w = triangulation_depth(l1_q2, q2, C_s, C_t, RtKinv_s, RtKinv_t, false);
        return true;
    }
    return false;
}

// cudawrapper.cu:355-377; P row-major 3x4
inline f3 project_point(const float* P, f3 X)
{
    float v[4] = {X.x, X.y, X.z, 1.0f}, p[3] = {0.0f, 0.0f, 0.0f};
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) p[r] += P[r * 4 + c] * v[c];
    if (fabsf(p[2]) > EPS_G) return mk3(p[0] / p[2], p[1] / p[2], 1.0f);
    return mk3(0, 0, 0);
}
// cudawrapper.cu:380-427; *gate_passed reports whether the 3-D distance gate let the pair through
float hypothesis_confidence(f3 p1, f3 p2, f3 P1, f3 P2, f3 Q1, f3 Q2, f3 C, f4 tgt, float sigma_p, float sigma_a, float spatial_k, bool* gate_passed)
{
    if (spatial_k > 0.0f) {
        float depth1 = length(C - P1), depth2 = length(C - P2);
        float unc1 = spatial_k * depth1, unc2 = spatial_k * depth2;
        float dist1 = length(P1 - Q1), dist2 = length(P2 - Q2);
        if (dist1 > unc1 || dist2 > unc2) { *gate_passed = false; return 0.0f; }
    }
    *gate_passed = true;
    f3 line1 = cross(p1, p2);
    f3 q1 = mk3(tgt.x, tgt.y, 1.0f), q2 = mk3(tgt.z, tgt.w, 1.0f), line2 = cross(q1, q2);
    float d1 = fmaxf(dist_p2l_2d(line2, p1), dist_p2l_2d(line2, p2));
    float d2 = fmaxf(dist_p2l_2d(line1, q1), dist_p2l_2d(line1, q2));
    float dist = fmaxf(d1, d2);
    float angle = angle_deg_3d(P1, P2, Q1, Q2);
    float sigma_sqr_a = sigma_a * sigma_a, sigma_sqr_d = sigma_p * sigma_p;
    float d = expf(-dist * dist / (2.0f * sigma_sqr_d));
    return fminf(d, expf(-angle * angle / (2.0f * sigma_sqr_a)));
}

struct Counters { double pair_tests = 0, accepted = 0, raw = 0, score_iters = 0, gate_pass = 0; };

// cudawrapper.cu:614-714, one "thread" y
float verify_match(long y, const f4* data, const f4* depths, const int* seg_start, const int* seg_count,
                   const f4* segs_src, const float* RtKinv_src, f3 C_src, const f4* tgt_segs, const int* cam_offsets,
                   const float* projections, float sigma_p, float sigma_a, float spatial_k, double* iters, double* gates)
{
    int srcID = int(data[y].x), camID = int(data[y].y);
    f4 sg = segs_src[srcID];
    f3 p1 = mk3(sg.x, sg.y, 1.0f), p2 = mk3(sg.z, sg.w, 1.0f);
    // D_unproject_point_src, cudawrapper.cu:338-344
    f3 ray1 = normalize(mat3_mul(RtKinv_src, p1, false)), ray2 = normalize(mat3_mul(RtKinv_src, p2, false));
    f3 P1 = C_src + depths[y].x * ray1, P2 = C_src + depths[y].y * ray2;
    int start = seg_start[srcID], end = start + seg_count[srcID];
    float confidence = 0.0f, current_confidence = 0.0f;
    int current_cam = -1;
    double it = 0, gp = 0;
    for (int i = start; i < end; ++i) {
        if (i == y) continue;
        int camID2 = int(data[i].y), tgtID2 = int(data[i].z);
        f3 Q1 = C_src + depths[i].x * ray1, Q2 = C_src + depths[i].y * ray2;
        if (camID2 == camID) continue;
        if (camID2 != current_cam) {
            if (current_cam != -1) confidence += current_confidence;
            current_confidence = 0.0f; current_cam = camID2;
        }
        it += 1;
        f3 proj1 = project_point(projections + camID2 * 12, P1), proj2 = project_point(projections + camID2 * 12, P2);
        if (int(proj1.z) == 1 && int(proj2.z) == 1) {
            bool g = false;
            float conf = hypothesis_confidence(proj1, proj2, P1, P2, Q1, Q2, C_src, tgt_segs[tgtID2 + cam_offsets[camID2]], sigma_p, sigma_a, spatial_k, &g);
            if (g) gp += 1;
            if (conf > 0.5f && conf > current_confidence) current_confidence = conf;
        }
    }
    confidence += current_confidence;
    if (iters) *iters += it;
    if (gates) *gates += gp;
    return confidence;
}

// sparsematrix.h:37-66
struct MatchingPair { unsigned segID1, camID2, segID2; f4 depths; float confidence; };
// sparsematrix.h:68-79
bool sort_pairs(const MatchingPair& a, const MatchingPair& b)
{
    if (a.segID1 < b.segID1) return true;
    if (a.segID1 == b.segID1 && a.camID2 < b.camID2) return true;
    if (a.segID1 == b.segID1 && a.camID2 == b.camID2 && a.segID2 < b.segID2) return true;
    return false;
}

struct NeighbourTables {                 // what line3D.cc:716-800 packs for one source view
    std::vector<float> F, RtKinv, P, C;  // per local camera: 9, 9, 12, 3 floats
    std::vector<int> offset, count;      // line3D.cc:779
    std::vector<f4> tgt_segs;            // line3D.cc:770-776
    std::map<unsigned, unsigned> local2global;
};

// cudawrapper.cu:858-1128.  `matches` holds the pre-existing (localised) records on entry
// and the filtered, globalised records on return; median_depth is only written when the
// function gets as far as the reference does.
void compute_pairwise_matches(const std::vector<f4>& segs_src, const float* RtKinv_src, f3 C_src, NeighbourTables& nt,
                              const std::list<unsigned>& toBeMatched, std::list<MatchingPair>& matches,
                              float sigma_p, float sigma_a, float spatial_k, float& median_depth, int threads, Counters& cnt)
{
    if (toBeMatched.empty()) return;                                            // :877-878
    const int height = int(segs_src.size());
    for (std::list<unsigned>::const_iterator it = toBeMatched.begin(); it != toBeMatched.end(); ++it) {   // :900-945
        const unsigned cam = *it;
        const int off = nt.offset[cam], width = nt.count[cam];
        std::vector<std::vector<std::pair<int, f4> > > rows(height);
        double acc = 0;
        #pragma omp parallel for schedule(dynamic, 8) num_threads(threads) reduction(+:acc)
        for (int i = 0; i < height; ++i)
            for (int j = 0; j < width; ++j) {
                f4 d;
                if (pair_test(segs_src[i], nt.tgt_segs[off + j], &nt.F[cam * 9], RtKinv_src, &nt.RtKinv[cam * 9], C_src,
                              mk3(nt.C[cam * 3], nt.C[cam * 3 + 1], nt.C[cam * 3 + 2]), d, 0)) acc += 1;
                if (d.x > 0.0f && d.y > 0.0f && d.z > 0.0f && d.w > 0.0f) rows[i].push_back(std::make_pair(j, d));   // :931
            }
        cnt.pair_tests += double(height) * width; cnt.accepted += acc;
        for (int i = 0; i < height; ++i)
            for (size_t k = 0; k < rows[i].size(); ++k) {
                MatchingPair mp; mp.segID1 = unsigned(i); mp.segID2 = unsigned(rows[i][k].first); mp.camID2 = cam;
                mp.depths = rows[i][k].second; mp.confidence = 0.0f;
                matches.push_back(mp); cnt.raw += 1;
            }
    }
    matches.sort(sort_pairs);                                                   // :951 (stable)
    if (matches.empty()) return;                                                // :955-956
    const long M = long(matches.size());
    std::vector<f4> data(M), depths(M);
    std::vector<int> seg_start(height, -1), seg_count(height, -1);              // :960-961
    {
        long pos = 0;
        for (std::list<MatchingPair>::iterator it = matches.begin(); it != matches.end(); ++it, ++pos) {   // :963-1003
            const int s = int(it->segID1);
            if (seg_start[s] < 0) { seg_start[s] = int(pos); seg_count[s] = 0; }
            seg_count[s] += 1;
            data[pos].x = float(s); data[pos].y = float(it->camID2); data[pos].z = float(it->segID2); data[pos].w = 0.0f;
            depths[pos] = it->depths;
        }
    }
    std::vector<float> conf(M);
    double iters = 0, gates = 0;
    #pragma omp parallel for schedule(dynamic, 64) num_threads(threads) reduction(+:iters, gates)
    for (long y = 0; y < M; ++y)                                                 // :1013-1019
        conf[y] = verify_match(y, data.data(), depths.data(), seg_start.data(), seg_count.data(), segs_src.data(), RtKinv_src, C_src,
                               nt.tgt_segs.data(), nt.offset.data(), nt.P.data(), sigma_p, sigma_a, spatial_k, &iters, &gates);
    cnt.score_iters += iters; cnt.gate_pass += gates;
    matches.clear();                                                            // :1022
    std::vector<float> dsamples;
    const float conf_t = 1.00f;
    for (int i = 0; i < height; ++i) {                                          // :1028-1064
        int start = seg_start[i], end = start + seg_count[i];
        if (start < 0) continue;
        float max_conf = 0.0f, ds1 = 0.0f, ds2 = 0.0f;
        for (int k = start; k < end; ++k)
            if (conf[k] > max_conf) { max_conf = conf[k]; ds1 = depths[k].x; ds2 = depths[k].y; }
        if (max_conf > conf_t / 2.0f) { dsamples.push_back(ds1); dsamples.push_back(ds2); }
    }
    median_depth = -1.0f;                                                       // :1066-1076
    if (!dsamples.empty()) { std::sort(dsamples.begin(), dsamples.end()); median_depth = dsamples[dsamples.size() / 2]; }
    for (long i = 0; i < M; ++i)                                                 // :1090-1110
        if (conf[i] > conf_t) {
            MatchingPair mp; mp.segID1 = unsigned(data[i].x); mp.camID2 = nt.local2global[unsigned(data[i].y)];
            mp.segID2 = unsigned(data[i].z); mp.depths = depths[i]; mp.confidence = conf[i] / 2.0f;
            matches.push_back(mp);
        }
}

// cudawrapper.cu:476-535 for one (x < y) pair
inline float collinearity_pair(f4 a, f4 b, float sigma_sqr)
{
    f3 p1 = mk3(a.x, a.y, 1.0f), p2 = mk3(a.z, a.w, 1.0f), line1 = cross(p1, p2);
    f3 q1 = mk3(b.x, b.y, 1.0f), q2 = mk3(b.z, b.w, 1.0f), line2 = cross(q1, q2);
    float d1 = fmaxf(dist_p2l_2d(line2, p1), dist_p2l_2d(line2, p2));
    float d2 = fmaxf(dist_p2l_2d(line1, q1), dist_p2l_2d(line1, q2));
    float d = fmaxf(d1, d2);
    float aff = expf(-d * d / (2.0f * sigma_sqr));
    float result = 0.0f;
    if (aff > COLLIN_AFF_T) {
        float pos1 = (q1.x - p1.x) * (q2.x - p1.x) + (q1.y - p1.y) * (q2.y - p1.y);
        float pos2 = (q1.x - p2.x) * (q2.x - p2.x) + (q1.y - p2.y) * (q2.y - p2.y);
        float pos3 = (p1.x - q1.x) * (p2.x - q1.x) + (p1.y - q1.y) * (p2.y - q1.y);
        float pos4 = (p1.x - q2.x) * (p2.x - q2.x) + (p1.y - q2.y) * (p2.y - q2.y);
        if (pos1 > -EPS_G && pos2 > -EPS_G && pos3 > -EPS_G && pos4 > -EPS_G) result = aff;
    }
    return result;
}

// ---------------------------------------------------------------------------------------
// sparse matrix + diffusion
// ---------------------------------------------------------------------------------------
struct Edge { int i, j; float w; };                                             // clustering.h:57-61
bool edges_by_col(const Edge& a, const Edge& b) { return a.j < b.j || (a.j == b.j && a.i < b.i); }   // clustering.h:95-103
bool edges_by_row(const Edge& a, const Edge& b) { return a.i < b.i || (a.i == b.i && a.j < b.j); }   // clustering.h:105-113
struct Sparse { std::vector<f4> e; std::vector<int> start; bool row_sorted; };

// sparsematrix.cc:63-117 (default: column-sorted, normalisation factor 1)
Sparse sparse_from_edges(std::list<Edge> A, int n, bool by_row)
{
    Sparse S; S.row_sorted = by_row;
    if (by_row) A.sort(edges_by_row); else A.sort(edges_by_col);
    S.e.resize(A.size()); S.start.assign(n, -1);
    int pos = 0, cur = -1;
    for (std::list<Edge>::iterator it = A.begin(); it != A.end(); ++it, ++pos) {
        f4 d; d.x = float(it->i); d.y = float(it->j); d.z = it->w / 1.0f; d.w = 0.0f;
        S.e[pos] = d;
        int rc = by_row ? it->i : it->j;
        if (cur != rc) { S.start[rc] = pos; cur = rc; }
    }
    return S;
}
// sparsematrix.cc:120-191 with change_sorting = true
Sparse sparse_resorted(const Sparse& M, int n)
{
    std::list<Edge> A;
    for (size_t k = 0; k < M.e.size(); ++k) { Edge e; e.i = int(M.e[k].x); e.j = int(M.e[k].y); e.w = M.e[k].z; A.push_back(e); }
    return sparse_from_edges(A, n, !M.row_sorted);
}
// cudawrapper.cu:717-762
void row_normalise(Sparse& P, int n)
{
    const int nnz = int(P.e.size());
    for (int y = 0; y < n; ++y) {
        int start = P.start[y];
        float sum = 0.0f;
        int i = start;
        while (i >= 0 && i < nnz) { if (int(P.e[i].x) != y) break; sum += P.e[i].z; ++i; }
        if (sum < EPS_G) sum = EPS_G;
        i = start;
        while (i >= 0 && i < nnz) { if (int(P.e[i].x) != y) break; P.e[i].z /= sum; ++i; }
    }
}
// cudawrapper.cu:765-829
void diffusion_step(const Sparse& P, const Sparse& W, Sparse& Pp)
{
    const int nnz = int(P.e.size());
    for (int y = 0; y < nnz; ++y) {
        f4 data = P.e[y];
        int r = int(data.y), c = int(data.x);
        float mul = 0.0f;
        int sp = P.start[r], sw = W.start[c];
        while (sp >= 0 && sw >= 0 && sp < nnz && sw < nnz) {
            f4 d1 = P.e[sp], d2 = W.e[sw];
            if (int(d1.x) != r || int(d2.y) != c) break;
            mul += (d1.z * d2.z);
            ++sp; ++sw;
        }
        mul *= data.z;
        if (mul < EPS_G) mul = EPS_G;
        int s = Pp.start[r];
        bool found = false;
        while (s >= 0 && s < nnz && !found) {
            if (int(Pp.e[s].x) != r) break;
            if (int(Pp.e[s].y) == c) { Pp.e[s].z = mul; found = true; }
            ++s;
        }
    }
}
// cudawrapper.cu:1131-1191
Sparse rdd(const Sparse& W, int n)
{
    Sparse P = sparse_resorted(W, n);
    Sparse Pp = P;
    row_normalise(P, n);
    for (int i = 0; i < RDD_ITER; ++i) {
        diffusion_step(P, W, Pp);
        std::swap(P, Pp);
        if (i < RDD_ITER - 1) row_normalise(P, n);
    }
    return P;
}

// ---------------------------------------------------------------------------------------
// host side: views, neighbours, hypotheses, affinities, clustering
// ---------------------------------------------------------------------------------------
typedef Eigen::Matrix3d M3; typedef Eigen::Vector3d V3;

struct Seg2D {                                                                  // commons.h:81-99
    unsigned cam, seg;
    Seg2D(unsigned c = 0, unsigned s = 0) : cam(c), seg(s) {}
    bool operator<(const Seg2D& o) const { return cam < o.cam || (cam == o.cam && seg < o.seg); }
};
struct Seg3D { V3 P1, P2, dir; float depth_p1, depth_p2; unsigned cam, seg; };  // commons.h:69-78
struct Hypothesis { Seg3D s3; Seg2D src, tgt; float score; bool valid; };       // commons.h:102-160 (fields used)

struct View {
    unsigned id, width, height;
    M3 K, Kinv, R, Rt, RtKinv; V3 t, C, pp; Eigen::MatrixXd P;
    float unc_upper_px, unc_lower_px, k_upper, k_lower, median_depth;
    std::vector<f4> segs;
    std::map<unsigned, std::map<unsigned, float> > collin;
    bool file_exists; std::list<MatchingPair> file;                             // stands in for matches_<id>_raw.bin

    // view.cc:90-121 / 124-147
    double shifted_distance(float dist_px)
    {
        V3 n = RtKinv * pp; n.normalize();
        V3 Pp = C + n;
        V3 sh(pp.x() + dist_px, pp.y(), 1.0);
        V3 d = RtKinv * sh; d.normalize();
        double tt = (Pp.dot(n) - n.dot(C)) / (n.dot(d));
        V3 Q = C + tt * d;
        return (Pp - Q).norm();
    }
    void define_uncertainty() { k_upper = float(shifted_distance(unc_upper_px)); k_lower = float(shifted_distance(unc_lower_px)); }
    float specific_k(float dist_px) { return float(shifted_distance(dist_px)); }
    // view.cc:6-66
    void init()
    {
        pp(0) = float(width) / 2.0f; pp(1) = float(height) / 2.0f; pp(2) = 1.0;
        Kinv = K.inverse(); Rt = R.transpose(); RtKinv = Rt * Kinv; C = Rt * (-1.0 * t);
        Eigen::MatrixXd Rpt(3, 4); Rpt.block<3, 3>(0, 0) = R; Rpt.block<3, 1>(0, 3) = t;
        P = K * Rpt;
        median_depth = 1.0f; file_exists = false;
        define_uncertainty();
    }
    // view.cc:227-261
    void transform(Eigen::Matrix4d& Qinv, double scale)
    {
        t *= scale;
        Eigen::MatrixXd Rtm(3, 4); Rtm.block<3, 3>(0, 0) = R; Rtm.block<3, 1>(0, 3) = t;
        Rtm *= Qinv;
        R = Rtm.block<3, 3>(0, 0); t = Rtm.block<3, 1>(0, 3);
        Rt = R.transpose(); P = K * Rtm; RtKinv = Rt * Kinv; C = Rt * (-1.0 * t);
        define_uncertainty();
    }
    // view.cc:353-377
    float lower_unc(float depth) { return depth < median_depth ? k_lower * depth : k_lower * median_depth; }
    float upper_unc(float depth) { return depth < median_depth ? k_upper * depth : k_upper * median_depth; }
    float sigma_sq(float depth) { float d1 = lower_unc(depth), d2 = upper_unc(depth); return -(d2 - d1) * (d2 - d1) / (2.0f * logf(0.01f)); }
    // view.cc:302-342
    Seg3D unproject(unsigned sid, float d1, float d2)
    {
        Seg3D s;
        V3 p1(segs[sid].x, segs[sid].y, 1.0), p2(segs[sid].z, segs[sid].w, 1.0);
        V3 r1 = RtKinv * p1; r1.normalize();
        V3 r2 = RtKinv * p2; r2.normalize();
        s.P1 = C + r1 * d1; s.P2 = C + r2 * d2;
        s.dir = s.P2 - s.P1; s.dir.normalize();
        s.cam = id; s.seg = sid; s.depth_p1 = d1; s.depth_p2 = d2;
        return s;
    }
    // view.cc:162-197 (the file is a list kept in memory here)
    void add_matches(std::list<MatchingPair>& m, bool remove_old, bool only_best)
    {
        if (only_best) {
            std::map<unsigned, std::list<MatchingPair> > best;
            for (std::list<MatchingPair>::iterator it = m.begin(); it != m.end(); ++it) best[it->segID1].push_back(*it);
            m.clear();
            for (std::map<unsigned, std::list<MatchingPair> >::iterator it = best.begin(); it != best.end(); ++it) {
                it->second.sort([](const MatchingPair& a, const MatchingPair& b) { return a.confidence > b.confidence; });
                m.push_back(it->second.front());
            }
        }
        if (file_exists && !remove_old) file.insert(file.end(), m.begin(), m.end());
        else { file = m; file_exists = true; }
    }
};

struct Pipeline {
    int neighbors; float unc_upper, unc_lower, sigma_p, sigma_a, min_baseline; bool collin, diffusion; int threads;
    std::map<unsigned, View*> views;
    std::map<unsigned, unsigned> num_wps;
    std::map<unsigned, std::map<unsigned, unsigned> > common_wps;
    std::map<unsigned, std::map<unsigned, float> > view_sim;
    std::map<unsigned, std::map<unsigned, bool> > wp2views, vn, matched;
    std::map<unsigned, std::map<unsigned, M3> > fund;
    std::map<Seg2D, Hypothesis> best_match;
    std::map<Seg2D, std::map<Seg2D, bool> > potential;
    double tr_scale; M3 tr_R; V3 tr_t; double tr_scale_inv; M3 tr_Rinv; V3 tr_tneg; Eigen::Matrix4d Qinv;
    Counters cnt;
    double t_match = 0, t_diff = 0, t_cluster = 0;
    // records for the dumps
    std::vector<int32_t> match_log; std::vector<float> match_median; std::vector<double> filtered, edges_pre, edges_cluster;
    int cluster_nodes = 0;
    struct FinalLine { std::list<std::pair<V3, V3> > s3; std::list<Seg2D> s2; };
    std::list<FinalLine> result;

    ~Pipeline() { for (std::map<unsigned, View*>::iterator it = views.begin(); it != views.end(); ++it) delete it->second; }

    // line3D.cc:1874-1935
    void process_worldpoints(unsigned vid, const std::vector<uint32_t>& wps)
    {
        num_wps[vid] = 0;
        for (size_t k = 0; k < wps.size(); ++k) {
            std::map<unsigned, bool>& seen = wp2views[wps[k]];
            if (seen.size() == 2) {
                std::map<unsigned, bool>::iterator w = seen.begin();
                unsigned v1 = w->first; ++w; unsigned v2 = w->first;
                if (common_wps[v1].find(v2) == common_wps[v1].end()) { common_wps[v1][v2] = 1; common_wps[v2][v1] = 1; }
                else { ++common_wps[v1][v2]; ++common_wps[v2][v1]; }
                ++num_wps[v1]; ++num_wps[v2];
            }
            if (seen.size() >= 2) {
                for (std::map<unsigned, bool>::iterator w = seen.begin(); w != seen.end(); ++w) {
                    if (common_wps[w->first].find(vid) == common_wps[w->first].end()) { common_wps[w->first][vid] = 1; common_wps[vid][w->first] = 1; }
                    else { ++common_wps[w->first][vid]; ++common_wps[vid][w->first]; }
                }
                ++num_wps[vid];
            }
            seen[vid] = true;
        }
    }

    // line3D.cc:95-217 with the segments supplied directly (cache path, :160-168) and
    // segments.h:60-102 for the collinearity map
    void add_image(unsigned id, unsigned w, unsigned h, const double* K, const double* R, const double* t,
                   const std::vector<float>& segs, const std::vector<uint32_t>& wps)
    {
        if (views.find(id) != views.end() || wps.empty() || w == 0 || h == 0) return;   // :111-127
        View* v = new View();
        v->id = id; v->width = w; v->height = h;
        for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) { v->K(r, c) = K[r * 3 + c]; v->R(r, c) = R[r * 3 + c]; } v->t(r) = t[r]; }
        v->unc_upper_px = unc_upper; v->unc_lower_px = unc_lower;
        const int n = int(segs.size() / 4);
        v->segs.resize(n);
        for (int i = 0; i < n; ++i) { v->segs[i].x = segs[i * 4]; v->segs[i].y = segs[i * 4 + 1]; v->segs[i].z = segs[i * 4 + 2]; v->segs[i].w = segs[i * 4 + 3]; }
        if (collin) {
            const float s2 = COLLINEARITY_S * COLLINEARITY_S;                  // cudawrapper.cu:850
            std::vector<std::vector<std::pair<int, float> > > rows(n);
            #pragma omp parallel for schedule(dynamic, 8) num_threads(threads)
            for (int i = 0; i < n; ++i)
                for (int j = i + 1; j < n; ++j) { float r = collinearity_pair(v->segs[i], v->segs[j], s2); if (r > 0.0f) rows[i].push_back(std::make_pair(j, r)); }
            for (int i = 0; i < n; ++i)
                for (size_t k = 0; k < rows[i].size(); ++k) { v->collin[i][rows[i][k].first] = rows[i][k].second; v->collin[rows[i][k].first][i] = rows[i][k].second; }   // segments.h:82-98
        }
        v->init();
        views[id] = v;
        process_worldpoints(id, wps);
    }

    // line3D.cc:476-549
    void find_visual_neighbors()
    {
        vn.clear();
        for (std::map<unsigned, std::map<unsigned, unsigned> >::iterator it = common_wps.begin(); it != common_wps.end(); ++it) {
            if (view_sim.find(it->first) != view_sim.end()) continue;
            for (std::map<unsigned, unsigned>::iterator n = it->second.begin(); n != it->second.end(); ++n) {
                float sim = 2.0f * float(n->second) / (float(num_wps[it->first] + num_wps[n->first]));
                if (sim > 1e-12) view_sim[it->first][n->first] = sim;
            }
        }
        for (std::map<unsigned, std::map<unsigned, float> >::iterator s = view_sim.begin(); s != view_sim.end(); ++s) {
            std::list<std::pair<unsigned, float> > cand;
            for (std::map<unsigned, float>::iterator n = s->second.begin(); n != s->second.end(); ++n) {
                if (views.find(n->first) == views.end()) continue;
                if (!(baseline(s->first, n->first) > min_baseline)) continue;
                bool ok = true;
                for (std::list<std::pair<unsigned, float> >::iterator e = cand.begin(); e != cand.end() && ok; ++e)
                    if (baseline(e->first, n->first) <= min_baseline) ok = false;
                if (ok) cand.push_back(std::make_pair(n->first, n->second));
            }
            cand.sort([](const std::pair<unsigned, float>& a, const std::pair<unsigned, float>& b) { return a.second > b.second; });   // commons.h:177-181
            if (neighbors > 0 && int(cand.size()) > neighbors) cand.resize(neighbors);
            for (std::list<std::pair<unsigned, float> >::iterator e = cand.begin(); e != cand.end(); ++e) vn[s->first][e->first] = true;
        }
    }
    float baseline(unsigned a, unsigned b) { return float((views[a]->C - views[b]->C).norm()); }    // view.cc:446-449

    // line3D.cc:552-617, 1694-1779
    void transform_geometry()
    {
        fund.clear();
        double size = double(views.size());
        V3 m(0.0, 0.0, 0.0);
        std::vector<V3> in_pts;
        for (std::map<unsigned, View*>::iterator it = views.begin(); it != views.end(); ++it) { m += it->second->C; in_pts.push_back(it->second->C); }
        m /= size;
        double q = 0.0;
        for (std::map<unsigned, View*>::iterator it = views.begin(); it != views.end(); ++it) q += (it->second->C - m).norm();
        q /= size;
        q = sqrtf(2.0) / q;
        Eigen::MatrixXd T(4, 4);
        T(0,0) = q; T(0,1) = 0; T(0,2) = 0; T(0,3) = -q * m(0);
        T(1,0) = 0; T(1,1) = q; T(1,2) = 0; T(1,3) = -q * m(1);
        T(2,0) = 0; T(2,1) = 0; T(2,2) = q; T(2,3) = -q * m(2);
        T(3,0) = 0; T(3,1) = 0; T(3,2) = 0; T(3,3) = 1.0;
        std::vector<V3> out_pts; V3 cog_out(0.0, 0.0, 0.0);
        for (size_t i = 0; i < in_pts.size(); ++i) {
            Eigen::Vector4d tv = T * Eigen::Vector4d(in_pts[i](0), in_pts[i](1), in_pts[i](2), 1.0);
            V3 t3(tv(0), tv(1), tv(2)); cog_out += t3; out_pts.push_back(t3);
        }
        cog_out /= size;
        // findSimilarityTransform, line3D.cc:1694-1718
        const size_t n = in_pts.size();
        double ssum = 0.0;
        for (size_t i = 0; i < n; ++i) ssum += (out_pts[i] - cog_out).norm() / (in_pts[i] - m).norm();
        tr_scale = ssum / double(n);
        V3 cog_in = m; cog_in *= tr_scale;
        for (size_t i = 0; i < n; ++i) in_pts[i] *= tr_scale;
        // euclideanTransformation, line3D.cc:1721-1755
        for (size_t i = 0; i < n; ++i) { in_pts[i] -= cog_in; out_pts[i] -= cog_out; }
        M3 H; H.setZero();
        for (size_t i = 0; i < n; ++i) { M3 outer = out_pts[i] * in_pts[i].transpose(); H += outer; }
        Eigen::JacobiSVD<Eigen::MatrixXd> svd(H, Eigen::ComputeThinU | Eigen::ComputeThinV);
        M3 Vt = svd.matrixV().transpose(), U = svd.matrixU();
        tr_R = U * Vt;
        if (tr_R.determinant() < 0) { Vt.row(2) *= -1; tr_R = U * Vt; }
        tr_t = cog_out - tr_R * cog_in;
        tr_t /= tr_scale;
        // applyTransformation, line3D.cc:1758-1779
        Eigen::Matrix4d Q;
        Q << tr_R.row(0), tr_t.x() * tr_scale, tr_R.row(1), tr_t.y() * tr_scale, tr_R.row(2), tr_t.z() * tr_scale, 0.0, 0.0, 0.0, 1.0;
        Qinv = Q.inverse();
        tr_scale_inv = 1.0 / tr_scale; tr_Rinv = tr_R.transpose(); tr_tneg = -tr_t;
        for (std::map<unsigned, View*>::iterator it = views.begin(); it != views.end(); ++it) it->second->transform(Qinv, tr_scale);
    }
    V3 inverse_transform(V3 P) { return tr_Rinv * (P * tr_scale_inv + tr_tneg); }               // line3D.cc:1782-1786

    // line3D.cc:1968-1993 / 1949-1965
    M3 fundamental(unsigned a, unsigned b)
    {
        View* v1 = views[a]; View* v2 = views[b];
        M3 R = v2->R * v1->R.transpose();
        V3 t = v2->t - R * v1->t;
        M3 T;
        T(0,0) = 0.0; T(0,1) = -t.z(); T(0,2) = t.y();
        T(1,0) = t.z(); T(1,1) = 0.0; T(1,2) = -t.x();
        T(2,0) = -t.y(); T(2,1) = t.x(); T(2,2) = 0.0;
        M3 E = T * R;
        return v2->K.transpose().inverse() * E * v1->K.inverse();
    }
    void compute_fundamentals(unsigned vid)
    {
        for (std::map<unsigned, bool>::iterator it = vn[vid].begin(); it != vn[vid].end(); ++it)
            if (fund[vid].find(it->first) == fund[vid].end()) { M3 F = fundamental(vid, it->first); M3 Ft = F.transpose(); fund[vid][it->first] = F; fund[it->first][vid] = Ft; }
    }

    // line3D.cc:698-885
    void perform_matching(unsigned vid)
    {
        if (vn[vid].empty()) return;
        View* v = views[vid];
        std::map<unsigned, unsigned> g2l;
        NeighbourTables nt;
        std::list<unsigned> toBeMatched;
        unsigned local = 0; int total = 0;
        for (std::map<unsigned, bool>::iterator it = vn[vid].begin(); it != vn[vid].end(); ++it, ++local) {
            g2l[it->first] = local; nt.local2global[local] = it->first;
            if (matched[vid].find(it->first) == matched[vid].end()) toBeMatched.push_back(local);
            View* u = views[it->first];
            M3 F = fund[vid][it->first];
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { nt.F.push_back(float(F(r, c))); }
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { nt.RtKinv.push_back(float(u->RtKinv(r, c))); }
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) nt.P.push_back(float(u->P(r, c)));
            for (int r = 0; r < 3; ++r) nt.C.push_back(float(u->C(r)));
            nt.offset.push_back(total); nt.count.push_back(int(u->segs.size())); total += int(u->segs.size());
            nt.tgt_segs.insert(nt.tgt_segs.end(), u->segs.begin(), u->segs.end());
        }
        float RtKinv_src[9];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) RtKinv_src[r * 3 + c] = float(v->RtKinv(r, c));
        f3 C_src = mk3(float(v->C.x()), float(v->C.y()), float(v->C.z()));
        // view.cc:200-224
        std::list<MatchingPair> matches;
        if (v->file_exists)
            for (std::list<MatchingPair>::iterator it = v->file.begin(); it != v->file.end(); ++it)
                if (g2l.find(it->camID2) != g2l.end()) { MatchingPair mp = *it; mp.camID2 = g2l[mp.camID2]; matches.push_back(mp); }
        const size_t n_pre = matches.size();
        float median = 1.0f;                                                     // :811
        auto t0 = std::chrono::steady_clock::now();
        compute_pairwise_matches(v->segs, RtKinv_src, C_src, nt, toBeMatched, matches, sigma_p, sigma_a,
                                 v->specific_k(2.0f * sigma_p), median, threads, cnt);
        t_match += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        match_log.push_back(int32_t(vid)); match_log.push_back(int32_t(toBeMatched.size())); match_log.push_back(int32_t(n_pre)); match_log.push_back(int32_t(matches.size()));
        match_median.push_back(median);
        for (std::list<MatchingPair>::iterator it = matches.begin(); it != matches.end(); ++it) {
            const double row[9] = {double(vid), double(it->segID1), double(it->camID2), double(it->segID2), it->depths.x, it->depths.y, it->depths.z, it->depths.w, it->confidence};
            filtered.insert(filtered.end(), row, row + 9);
        }
        v->median_depth = median;                                               // :835
        std::map<unsigned, std::list<MatchingPair> > other;                     // :838-866
        for (std::list<MatchingPair>::iterator it = matches.begin(); it != matches.end(); ++it) {
            unsigned cam = it->camID2;
            if (vn[cam].find(vid) != vn[cam].end() && matched[cam].find(vid) == matched[cam].end()) {
                MatchingPair r; r.segID1 = it->segID2; r.camID2 = vid; r.segID2 = it->segID1; r.confidence = 0.0f;
                r.depths.x = it->depths.z; r.depths.y = it->depths.w; r.depths.z = it->depths.x; r.depths.w = it->depths.y;
                other[cam].push_back(r);
            }
            Seg2D ref(vid, it->segID1), tgt(it->camID2, it->segID2);
            potential[ref][tgt] = true; potential[tgt][ref] = true;
        }
        for (std::map<unsigned, std::list<MatchingPair> >::iterator o = other.begin(); o != other.end(); ++o)
            if (views.find(o->first) != views.end()) views[o->first]->add_matches(o->second, false, false);   // :868-872
        for (std::map<unsigned, bool>::iterator it = vn[vid].begin(); it != vn[vid].end(); ++it) {           // :875-881
            matched[vid][it->first] = true;
            if (vn[it->first].find(vid) != vn[it->first].end()) matched[it->first][vid] = true;
        }
        v->add_matches(matches, true, true);                                    // :884
    }

    // line3D.cc:899-965
    void greedy_selection()
    {
        best_match.clear();
        for (std::map<unsigned, View*>::iterator it = views.begin(); it != views.end(); ++it) {
            if (!it->second->file_exists) continue;
            std::map<Seg2D, std::list<MatchingPair> > per_seg;
            for (std::list<MatchingPair>::iterator m = it->second->file.begin(); m != it->second->file.end(); ++m) per_seg[Seg2D(it->first, m->segID1)].push_back(*m);
            for (std::map<Seg2D, std::list<MatchingPair> >::iterator s = per_seg.begin(); s != per_seg.end(); ++s) {
                s->second.sort([](const MatchingPair& a, const MatchingPair& b) { return a.confidence > b.confidence; });
                MatchingPair mp = s->second.front();
                mp.confidence = fminf(mp.confidence, 1.0f);
                Hypothesis h; h.src = s->first; h.tgt = Seg2D(mp.camID2, mp.segID2); h.valid = true; h.score = mp.confidence;
                h.s3 = views[s->first.cam]->unproject(s->first.seg, mp.depths.x, mp.depths.y);
                best_match[s->first] = h;
            }
        }
    }

    // line3D.cc:1684-1691
    static float dist_point_line(const Seg3D& s, const V3& X)
    {
        V3 proj = s.P1 + (s.dir * ((X - s.P1).transpose()) * s.dir);
        return float((proj - X).norm());
    }
    // line3D.cc:1600-1681
    static float similarity(const Seg3D& a, View* va, const Seg3D& b, View* vb, float sigma_a)
    {
        float d1 = dist_point_line(b, a.P1), d2 = dist_point_line(b, a.P2);
        float min_d1 = va->lower_unc(a.depth_p1), min_d2 = va->lower_unc(a.depth_p2);
        float s1 = va->sigma_sq(a.depth_p1), s2 = va->sigma_sq(a.depth_p2);
        float sim1 = (d1 < min_d1) ? 1.0f : expf(-(d1 - min_d1) * (d1 - min_d1) / (2.0f * s1));
        float sim2 = (d2 < min_d2) ? 1.0f : expf(-(d2 - min_d2) * (d2 - min_d2) / (2.0f * s2));
        float w12 = fminf(sim1, sim2);
        float d3 = dist_point_line(a, b.P1), d4 = dist_point_line(a, b.P2);
        float min_d3 = vb->lower_unc(b.depth_p1), min_d4 = vb->lower_unc(b.depth_p2);
        float s3 = vb->sigma_sq(b.depth_p1), s4 = vb->sigma_sq(b.depth_p2);
        float sim3 = (d3 < min_d3) ? 1.0f : expf(-(d3 - min_d3) * (d3 - min_d3) / (2.0f * s3));
        float sim4 = (d4 < min_d4) ? 1.0f : expf(-(d4 - min_d4) * (d4 - min_d4) / (2.0f * s4));
        float w_d = fminf(w12, fminf(sim3, sim4));
        float angle = float(acos(fmax(fmin(a.dir.dot(b.dir), 1.0), -1.0)) / M_PI * double(180.0f));
        if (angle > 90.0f) angle = 180.0f - angle;
        float w_a = expf(-angle * angle / (2.0f * sigma_a * sigma_a));
        float sim = fminf(w_d, w_a);
        return (sim <= 0.01f) ? 0.0f : sim;
    }

    // line3D.cc:968-1252
    void cluster_segments()
    {
        result.clear();
        std::list<Edge> A;
        unsigned next_id = 0;
        std::map<Seg2D, unsigned> g2l; std::map<unsigned, Seg2D> l2g;
        std::map<Seg2D, std::map<Seg2D, bool> > used;
        auto id_of = [&](const Seg2D& s) -> unsigned {
            std::map<Seg2D, unsigned>::iterator f = g2l.find(s);
            if (f != g2l.end()) return f->second;
            unsigned id = next_id++; g2l[s] = id; l2g[id] = s; return id;
        };
        auto add_edge = [&](std::list<Edge>& out, const Seg2D& a, const Seg2D& b, float w) {
            unsigned ia = id_of(a); unsigned ib = id_of(b);            // src id first, then target id (:1018-1050)
            Edge e; e.i = int(ia); e.j = int(ib); e.w = w; out.push_back(e);
            e.i = int(ib); e.j = int(ia); out.push_back(e);
        };
        for (std::map<Seg2D, Hypothesis>::iterator it = best_match.begin(); it != best_match.end(); ++it) {
            const Seg2D src = it->first;
            Hypothesis C = it->second;
            if (!C.valid) continue;
            std::list<Edge> local;
            std::map<Seg2D, bool>& corr = potential[src];
            for (std::map<Seg2D, bool>::iterator c = corr.begin(); c != corr.end(); ++c) {
                Seg2D tgt = c->first;
                if (used[src].find(tgt) != used[src].end()) continue;
                used[src][tgt] = true; used[tgt][src] = true;
                std::map<Seg2D, Hypothesis>::iterator bt = best_match.find(tgt);
                if (bt == best_match.end()) continue;
                if (bt->second.valid) {
                    float w = 0.5f * (C.score + bt->second.score) * similarity(C.s3, views[C.s3.cam], bt->second.s3, views[bt->second.s3.cam], sigma_a);
                    if (w > MIN_AFFINITY) add_edge(local, src, tgt, w);
                }
                std::map<unsigned, std::map<unsigned, float> >& cm = views[tgt.cam]->collin;
                std::map<unsigned, std::map<unsigned, float> >::iterator cr = cm.find(tgt.seg);
                if (cr == cm.end()) continue;
                for (std::map<unsigned, float>::iterator tc = cr->second.begin(); tc != cr->second.end(); ++tc) {
                    Seg2D tgtc(tgt.cam, tc->first);
                    if (used[src].find(tgtc) != used[src].end()) continue;
                    used[src][tgtc] = true; used[tgtc][src] = true;
                    std::map<Seg2D, Hypothesis>::iterator b3 = best_match.find(tgtc);
                    if (b3 == best_match.end() || !b3->second.valid) continue;
                    float w = 0.5f * (C.score + b3->second.score) * similarity(C.s3, views[C.s3.cam], b3->second.s3, views[b3->second.s3.cam], sigma_a);
                    if (w > 0.01f) add_edge(local, src, tgtc, w);
                }
            }
            std::map<unsigned, std::map<unsigned, float> >& cm = views[src.cam]->collin;
            std::map<unsigned, std::map<unsigned, float> >::iterator cr = cm.find(src.seg);
            if (cr != cm.end())
                for (std::map<unsigned, float>::iterator ci = cr->second.begin(); ci != cr->second.end(); ++ci) {
                    Seg2D tgt(src.cam, ci->first);
                    float collin_w = ci->second;
                    if (used[src].find(tgt) != used[src].end()) continue;
                    used[src][tgt] = true; used[tgt][src] = true;
                    std::map<Seg2D, Hypothesis>::iterator bt = best_match.find(tgt);
                    if (bt == best_match.end() || !bt->second.valid) continue;
                    float w = collin_w * 0.5f * (C.score + bt->second.score) * similarity(C.s3, views[C.s3.cam], bt->second.s3, views[bt->second.s3.cam], sigma_a);
                    if (w > 0.01f) add_edge(local, src, tgt, w);
                }
            A.splice(A.end(), local);
        }
        if (A.empty()) return;
        const int n_nodes = int(l2g.size());
        if (diffusion) {                                                        // line3D.cc:1255-1303
            auto t0 = std::chrono::steady_clock::now();
            Sparse W = sparse_from_edges(A, n_nodes, false);
            for (size_t k = 0; k < W.e.size(); ++k) { edges_pre.push_back(W.e[k].x); edges_pre.push_back(W.e[k].y); edges_pre.push_back(W.e[k].z); }
            Sparse Pm = rdd(W, n_nodes);
            A.clear();
            std::map<int, std::map<int, float> > entries;
            for (size_t k = 0; k < Pm.e.size(); ++k) {
                int s1 = int(Pm.e[k].x), s2 = int(Pm.e[k].y); float w12 = Pm.e[k].z, w21 = w12;
                if (entries[s2].find(s1) != entries[s2].end()) w21 = entries[s2][s1];
                float w = fminf(w12, w21);
                entries[s1][s2] = w; entries[s2][s1] = w;
            }
            for (std::map<int, std::map<int, float> >::iterator a = entries.begin(); a != entries.end(); ++a)
                for (std::map<int, float>::iterator b = a->second.begin(); b != a->second.end(); ++b) { Edge e; e.i = a->first; e.j = b->first; e.w = b->second; A.push_back(e); }
            t_diff += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
        cluster_nodes = n_nodes;
        for (std::list<Edge>::iterator e = A.begin(); e != A.end(); ++e) { edges_cluster.push_back(e->i); edges_cluster.push_back(e->j); edges_cluster.push_back(e->w); }
        auto t0 = std::chrono::steady_clock::now();
        // clustering.cc:6-47 with universe.h:59-115
        std::list<Edge> sorted = A;
        sorted.sort([](const Edge& a, const Edge& b) { return a.w < b.w; });
        std::vector<int> parent(n_nodes), rank(n_nodes, 0), sz(n_nodes, 1);
        std::vector<float> thr(n_nodes, 1.0f);
        for (int i = 0; i < n_nodes; ++i) parent[i] = i;
        auto find = [&](int x) { int y = x; while (y != parent[y]) y = parent[y]; parent[x] = y; return y; };
        for (std::list<Edge>::iterator e = sorted.begin(); e != sorted.end(); ++e) {
            int a = find(e->i), b = find(e->j);
            if (a != b && e->w <= thr[a] && e->w <= thr[b]) {
                if (rank[a] > rank[b]) { parent[b] = a; sz[a] += sz[b]; }
                else { parent[a] = b; sz[b] += sz[a]; if (rank[a] == rank[b]) rank[b]++; }
                a = find(a);
                thr[a] = e->w + 1.0f / float(sz[a]);
            }
        }
        t_cluster += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        // line3D.cc:1306-1368
        std::map<unsigned, std::list<Seg2D> > c2s; std::map<unsigned, std::map<unsigned, bool> > c2c;
        for (std::map<unsigned, Seg2D>::iterator it = l2g.begin(); it != l2g.end(); ++it) { unsigned cl = unsigned(find(int(it->first))); c2s[cl].push_back(it->second); c2c[cl][it->second.cam] = true; }
        for (std::map<unsigned, std::list<Seg2D> >::iterator c = c2s.begin(); c != c2s.end(); ++c) {
            if (c2c[c->first].size() < 4) continue;
            std::map<Seg2D, std::pair<V3, V3> > t3;                              // :1371-1389
            for (std::list<Seg2D>::iterator s = c->second.begin(); s != c->second.end(); ++s) {
                std::map<Seg2D, Hypothesis>::iterator b = best_match.find(*s);
                if (b != best_match.end()) t3[*s] = std::make_pair(inverse_transform(b->second.s3.P1), inverse_transform(b->second.s3.P2));
            }
            FinalLine fl;
            align(t3, fl);
            if (!fl.s3.empty()) result.push_back(fl);
        }
    }

    // line3D.cc:1392-1451, 1454-1476, 1479-1597
    void align(std::map<Seg2D, std::pair<V3, V3> >& t3, FinalLine& fl)
    {
        if (t3.empty()) return;
        V3 P(0, 0, 0);
        int n = int(t3.size()) * 2;
        Eigen::MatrixXd g(3, n);
        int j = 0; double sz = 0.0;
        for (std::map<Seg2D, std::pair<V3, V3> >::iterator it = t3.begin(); it != t3.end(); ++it) {
            P += it->second.first; P += it->second.second;
            for (int r = 0; r < 3; ++r) g(r, j) = it->second.first(r);
            ++j;
            for (int r = 0; r < 3; ++r) g(r, j) = it->second.second(r);
            ++j;
            sz += 2.0;
        }
        P /= sz;
        Eigen::MatrixXd Cm = Eigen::MatrixXd::Identity(n, n) - (1.0 / double(n)) * Eigen::MatrixXd::Constant(n, n, 1.0);
        Eigen::MatrixXd Scat = g * Cm * g.transpose();
        Eigen::JacobiSVD<Eigen::MatrixXd> svd(Scat, Eigen::ComputeThinU);
        Eigen::MatrixXd U = svd.matrixU(); Eigen::VectorXd S = svd.singularValues();
        int mp; S.maxCoeff(&mp);
        V3 dir(U(0, mp), U(1, mp), U(2, mp)); dir.normalize();
        V3 L1 = P;
        struct Pt { V3 P; unsigned seg3, cam; float dist; };
        std::list<Pt> pts;
        V3 min_point(0.0, 0.0, 0.0), max_point(0.0, 0.0, 0.0);
        double min_len = 0.0, max_len = 0.0;
        unsigned sid = 0;
        for (std::map<Seg2D, std::pair<V3, V3> >::iterator it = t3.begin(); it != t3.end(); ++it, ++sid) {
            V3 A1 = it->second.first, A2 = it->second.second;
            V3 pr1 = L1 + (dir.dot(A1 - L1) / (dir.norm() * dir.norm())) * dir;
            V3 pr2 = L1 + (dir.dot(A2 - L1) / (dir.norm() * dir.norm())) * dir;
            double loc1 = dir.dot(L1 - pr1);
            if (loc1 <= min_len) { min_len = loc1; min_point = pr1; }
            if (loc1 >= max_len) { max_len = loc1; max_point = pr1; }
            double loc2 = dir.dot(L1 - pr2);
            if (loc2 <= min_len) { min_len = loc2; min_point = pr2; }
            if (loc2 >= max_len) { max_len = loc2; max_point = pr2; }
            Pt a; a.P = A1; a.seg3 = sid; a.cam = it->first.cam; a.dist = 0; Pt b = a; b.P = A2;
            pts.push_back(a); pts.push_back(b);
        }
        for (std::list<Pt>::iterator p = pts.begin(); p != pts.end(); ++p) p->dist = float((p->P - min_point).norm());
        pts.sort([](const Pt& a, const Pt& b) { return a.dist < b.dist; });
        std::map<unsigned, unsigned> open; std::map<unsigned, bool> open_lines;
        bool opened = false; V3 start(0, 0, 0);
        for (std::list<Pt>::iterator p = pts.begin(); p != pts.end(); ++p) {
            if (open_lines.find(p->seg3) == open_lines.end()) { open_lines[p->seg3] = true; if (open.find(p->cam) == open.end()) open[p->cam] = 1; else ++open[p->cam]; }
            else { open_lines.erase(p->seg3); --open[p->cam]; if (open[p->cam] == 0) open.erase(p->cam); }
            if (opened && open.size() < 3) { fl.s3.push_back(std::make_pair(start, p->P)); opened = false; }
            else if (!opened && open.size() >= 3) { start = p->P; opened = true; }
        }
        for (std::map<Seg2D, std::pair<V3, V3> >::iterator it = t3.begin(); it != t3.end(); ++it) fl.s2.push_back(it->first);
    }

    // Not in the reference: multi-process sharding by connected component of the neighbour graph (DESIGN.md
    // section 5).  Components are dealt out whole, most expensive first, to the least loaded rank.
    int shard_rank = 0, shard_world = 1;
    std::set<unsigned> shard_views()
    {
        std::map<unsigned, unsigned> root;                                   // union-find over view ids
        std::function<unsigned(unsigned)> find = [&](unsigned x) { while (root[x] != x) x = root[x] = root[root[x]]; return x; };
        for (auto& a : vn) { if (!root.count(a.first)) root[a.first] = a.first; for (auto& b : a.second) if (!root.count(b.first)) root[b.first] = b.first; }
        for (auto& a : vn) for (auto& b : a.second) { unsigned x = find(a.first), y = find(b.first); if (x != y) root[std::max(x, y)] = std::min(x, y); }
        std::map<unsigned, double> cost; std::map<unsigned, std::vector<unsigned> > members;   // keyed by smallest view id of the component
        for (auto& r : root) members[find(r.first)].push_back(r.first);
        for (auto& a : vn) for (auto& b : a.second) cost[find(a.first)] += double(views[a.first]->segs.size()) * double(views[b.first]->segs.size());
        std::vector<unsigned> order;
        for (auto& m : members) order.push_back(m.first);
        std::stable_sort(order.begin(), order.end(), [&](unsigned a, unsigned b) { return cost[a] > cost[b]; });
        std::vector<double> load(size_t(shard_world), 0.0);
        std::set<unsigned> mine;
        for (unsigned c : order) {
            size_t best = 0;
            for (size_t r = 1; r < load.size(); ++r) if (load[r] < load[best]) best = r;
            load[best] += cost[c];
            if (int(best) == shard_rank) mine.insert(members[c].begin(), members[c].end());
        }
        return mine;
    }

    // line3D.cc:345-374, 620-648, 888-896
    void compute_3d_model()
    {
        if (views.size() < 4) return;
        matched.clear(); potential.clear(); result.clear();
        find_visual_neighbors();
        const std::set<unsigned> mine = shard_views();
        transform_geometry();
        for (std::map<unsigned, std::map<unsigned, bool> >::iterator it = vn.begin(); it != vn.end(); ++it) {
            if (it->second.empty() || !mine.count(it->first)) continue;
            compute_fundamentals(it->first);
            perform_matching(it->first);
        }
        greedy_selection();
        cluster_segments();
    }
};

// ---------------------------------------------------------------------------------------
// scene file + npy dumps (same layout as oracle/ref_harness.cpp)
// ---------------------------------------------------------------------------------------
struct SceneView { uint32_t id, width, height; double K[9], R[9], t[3]; std::vector<float> segs; std::vector<uint32_t> wps; };
bool read_scene(const char* path, std::vector<SceneView>& views)
{
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    char magic[8]; f.read(magic, 8);
    if (std::memcmp(magic, "L3DSCN01", 8) != 0) return false;
    uint32_t nv, rsv; f.read(reinterpret_cast<char*>(&nv), 4); f.read(reinterpret_cast<char*>(&rsv), 4);
    views.resize(nv);
    for (uint32_t i = 0; i < nv; ++i) {
        SceneView& v = views[i];
        uint32_t hdr[5]; f.read(reinterpret_cast<char*>(hdr), 20);
        v.id = hdr[0]; v.width = hdr[1]; v.height = hdr[2];
        f.read(reinterpret_cast<char*>(v.K), 72); f.read(reinterpret_cast<char*>(v.R), 72); f.read(reinterpret_cast<char*>(v.t), 24);
        v.segs.resize(size_t(hdr[3]) * 4); if (hdr[3]) f.read(reinterpret_cast<char*>(v.segs.data()), std::streamsize(v.segs.size() * 4));
        v.wps.resize(hdr[4]); if (hdr[4]) f.read(reinterpret_cast<char*>(v.wps.data()), std::streamsize(v.wps.size() * 4));
    }
    return bool(f);
}
void write_npy(const std::string& path, const char* descr, size_t elem, size_t rows, size_t cols, const void* data)
{
    std::ostringstream h;
    h << "{'descr': '" << descr << "', 'fortran_order': False, 'shape': (" << rows << ", " << cols << "), }";
    std::string hs = h.str();
    size_t total = 10 + hs.size() + 1, pad = (64 - total % 64) % 64;
    hs += std::string(pad, ' '); hs += "\n";
    std::ofstream f(path.c_str(), std::ios::binary);
    const unsigned char magic[8] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0};
    f.write(reinterpret_cast<const char*>(magic), 8);
    uint16_t hl = uint16_t(hs.size()); f.write(reinterpret_cast<const char*>(&hl), 2);
    f.write(hs.data(), std::streamsize(hs.size()));
    if (rows > 0 && cols > 0) f.write(static_cast<const char*>(data), std::streamsize(rows * cols * elem));
}
void npy_f64(const std::string& p, const std::vector<double>& v, size_t cols) { write_npy(p, "<f8", 8, v.size() / cols, cols, v.data()); }
void npy_i32(const std::string& p, const std::vector<int32_t>& v, size_t cols) { write_npy(p, "<i4", 4, v.size() / cols, cols, v.data()); }
void npy_f32(const std::string& p, const std::vector<float>& v, size_t cols) { write_npy(p, "<f4", 4, v.size() / cols, cols, v.data()); }

}  // namespace

static int g_shard_rank = 0, g_shard_world = 1;

extern "C" {

void oracle_set_shard(int rank, int world) { g_shard_world = world < 1 ? 1 : world; g_shard_rank = (rank < 0 || rank >= g_shard_world) ? 0 : rank; }

int oracle_run_scene(const char* scene_path, int neighbors, float unc_upper, float unc_lower, float sigma_p, float sigma_a,
                     float min_baseline, int collinearity, int diffusion, const char* dump_dir, int threads, double* timings, double* counters)
{
    std::vector<SceneView> sv;
    if (!read_scene(scene_path, sv)) return -1;
    Pipeline pl;
    pl.neighbors = neighbors;
    // line3D.cc:18-28
    pl.unc_upper = fabsf(unc_upper); pl.unc_lower = fabsf(unc_lower);
    if (pl.unc_lower < 1.0f) pl.unc_lower = 1.0f;
    if (pl.unc_upper <= pl.unc_lower) pl.unc_upper = pl.unc_lower + 1.0f;
    pl.sigma_p = sigma_p; pl.sigma_a = sigma_a; pl.min_baseline = min_baseline; pl.collin = collinearity != 0; pl.diffusion = diffusion != 0;
    pl.threads = threads < 1 ? 1 : threads;
    pl.shard_rank = g_shard_rank; pl.shard_world = g_shard_world;
    auto t0 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < sv.size(); ++i) pl.add_image(sv[i].id, sv[i].width, sv[i].height, sv[i].K, sv[i].R, sv[i].t, sv[i].segs, sv[i].wps);
    auto t1 = std::chrono::steady_clock::now();
    pl.compute_3d_model();
    auto t2 = std::chrono::steady_clock::now();
    if (timings) {
        timings[0] = std::chrono::duration<double>(t1 - t0).count(); timings[1] = std::chrono::duration<double>(t2 - t1).count();
        timings[2] = pl.t_match; timings[3] = pl.t_diff; timings[4] = pl.t_cluster;
    }
    if (counters) { counters[0] = pl.cnt.pair_tests; counters[1] = pl.cnt.accepted; counters[2] = pl.cnt.raw; counters[3] = pl.cnt.score_iters; counters[4] = pl.cnt.gate_pass; }
    if (dump_dir && dump_dir[0]) {
        const std::string d = std::string(dump_dir) + "/";
        std::vector<int32_t> vn;
        for (auto& a : pl.vn) for (auto& b : a.second) { vn.push_back(int32_t(a.first)); vn.push_back(int32_t(b.first)); }
        npy_i32(d + "visual_neighbors.npy", vn, 2);
        std::vector<double> vw;
        for (auto& a : pl.views) {
            View* v = a.second;
            vw.push_back(a.first); vw.push_back(v->median_depth); vw.push_back(v->k_upper); vw.push_back(v->k_lower); vw.push_back(v->specific_k(2.0f * sigma_p));
            for (int i = 0; i < 3; ++i) vw.push_back(v->C(i));
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(v->RtKinv(r, c));
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) vw.push_back(v->P(r, c));
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(v->K(r, c));
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(v->R(r, c));
            for (int i = 0; i < 3; ++i) vw.push_back(v->t(i));
        }
        npy_f64(d + "views.npy", vw, 50);
        std::vector<double> fs;
        for (auto& a : pl.fund) for (auto& b : a.second) { fs.push_back(a.first); fs.push_back(b.first); for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) fs.push_back(b.second(r, c)); }
        npy_f64(d + "fundamentals.npy", fs, 11);
        std::vector<double> co;
        for (auto& a : pl.views) for (auto& r : a.second->collin) for (auto& c : r.second) { co.push_back(a.first); co.push_back(r.first); co.push_back(c.first); co.push_back(c.second); }
        npy_f64(d + "collinearity.npy", co, 4);
        npy_i32(d + "match_log.npy", pl.match_log, 4);
        npy_f32(d + "match_median.npy", pl.match_median, 1);
        npy_f64(d + "filtered_matches.npy", pl.filtered, 9);
        std::vector<double> bm;
        for (auto& a : pl.best_match) {
            const Hypothesis& h = a.second;
            bm.push_back(a.first.cam); bm.push_back(a.first.seg); bm.push_back(h.tgt.cam); bm.push_back(h.tgt.seg); bm.push_back(h.score); bm.push_back(h.valid ? 1.0 : 0.0);
            for (int i = 0; i < 3; ++i) bm.push_back(h.s3.P1(i));
            for (int i = 0; i < 3; ++i) bm.push_back(h.s3.P2(i));
            for (int i = 0; i < 3; ++i) bm.push_back(h.s3.dir(i));
            bm.push_back(h.s3.depth_p1); bm.push_back(h.s3.depth_p2);
        }
        npy_f64(d + "best_match.npy", bm, 17);
        std::vector<int32_t> pc;
        for (auto& a : pl.potential) for (auto& b : a.second) { pc.push_back(int32_t(a.first.cam)); pc.push_back(int32_t(a.first.seg)); pc.push_back(int32_t(b.first.cam)); pc.push_back(int32_t(b.first.seg)); }
        npy_i32(d + "potential.npy", pc, 4);
        npy_f64(d + "edges_pre_diffusion.npy", pl.edges_pre, 3);
        npy_f64(d + "edges_cluster.npy", pl.edges_cluster, 3);
        std::vector<int32_t> nn(1, pl.cluster_nodes); npy_i32(d + "cluster_nodes.npy", nn, 1);
        std::vector<double> s3; std::vector<int32_t> s2; int li = 0;
        for (auto it = pl.result.begin(); it != pl.result.end(); ++it, ++li) {
            for (auto& p : it->s3) { s3.push_back(li); for (int i = 0; i < 3; ++i) s3.push_back(p.first(i)); for (int i = 0; i < 3; ++i) s3.push_back(p.second(i)); }
            for (auto& q : it->s2) { s2.push_back(li); s2.push_back(int32_t(q.cam)); s2.push_back(int32_t(q.seg)); }
        }
        npy_f64(d + "lines_seg3d.npy", s3, 7);
        npy_i32(d + "lines_seg2d.npy", s2, 3);
        std::vector<double> tr; tr.push_back(pl.tr_scale);
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) tr.push_back(pl.tr_R(r, c));
        for (int i = 0; i < 3; ++i) tr.push_back(pl.tr_t(i));
        npy_f64(d + "transform.npy", tr, 13);
    }
    return int(pl.result.size());
}

int oracle_pair_test(const float* s, const float* t, const float* F, const float* RtKinv_src, const float* RtKinv_tgt, const float* C_src, const float* C_tgt, float* out)
{
    f4 a = {s[0], s[1], s[2], s[3]}, b = {t[0], t[1], t[2], t[3]}, d; float ov[2];
    bool acc = pair_test(a, b, F, RtKinv_src, RtKinv_tgt, mk3(C_src[0], C_src[1], C_src[2]), mk3(C_tgt[0], C_tgt[1], C_tgt[2]), d, ov);
    out[0] = d.x; out[1] = d.y; out[2] = d.z; out[3] = d.w; out[4] = ov[0]; out[5] = ov[1];
    return acc ? 1 : 0;
}

long oracle_match_pair(const float* segs_src, int n_src, const float* segs_tgt, int n_tgt, const float* F, const float* RtKinv_src, const float* RtKinv_tgt,
                       const float* C_src, const float* C_tgt, int* out_ij, float* out_depths, long cap, long* band)
{
    long n = 0, nb = 0;
    f3 Cs = mk3(C_src[0], C_src[1], C_src[2]), Ct = mk3(C_tgt[0], C_tgt[1], C_tgt[2]);
    for (int i = 0; i < n_src; ++i)
        for (int j = 0; j < n_tgt; ++j) {
            f4 a = {segs_src[i * 4], segs_src[i * 4 + 1], segs_src[i * 4 + 2], segs_src[i * 4 + 3]};
            f4 b = {segs_tgt[j * 4], segs_tgt[j * 4 + 1], segs_tgt[j * 4 + 2], segs_tgt[j * 4 + 3]};
            f4 d; float ov[2];
            pair_test(a, b, F, RtKinv_src, RtKinv_tgt, Cs, Ct, d, ov);
            if (ov[0] >= 0.0f) {
                float lo = fminf(ov[0], ov[1]), hi = fmaxf(ov[0], ov[1]);
                if (fabsf(lo - MIN_OVERLAP_LOWER) < 1e-5f || fabsf(hi - MIN_OVERLAP_UPPER) < 1e-5f) ++nb;
            }
            if (d.x > 0.0f && d.y > 0.0f && d.z > 0.0f && d.w > 0.0f) {
                if (n < cap) { out_ij[2 * n] = i; out_ij[2 * n + 1] = j; out_depths[4 * n] = d.x; out_depths[4 * n + 1] = d.y; out_depths[4 * n + 2] = d.z; out_depths[4 * n + 3] = d.w; }
                ++n;
            }
        }
    if (band) *band = nb;
    return n;
}

void oracle_score_matches(const float* data, const float* depths, long M, const float* segs_src, int n_src, const float* RtKinv_src, const float* C_src,
                          const float* tgt_segs, const int* cam_offsets, const float* projections, int n_cams, float sigma_p, float sigma_a, float spatial_k, float* score)
{
    (void)n_cams;
    std::vector<int> start(n_src, -1), count(n_src, -1);
    const f4* dt = reinterpret_cast<const f4*>(data);
    for (long k = 0; k < M; ++k) { int s = int(dt[k].x); if (start[s] < 0) { start[s] = int(k); count[s] = 0; } count[s] += 1; }
    #pragma omp parallel for schedule(dynamic, 64)
    for (long y = 0; y < M; ++y)
        score[y] = verify_match(y, dt, reinterpret_cast<const f4*>(depths), start.data(), count.data(), reinterpret_cast<const f4*>(segs_src), RtKinv_src,
                                mk3(C_src[0], C_src[1], C_src[2]), reinterpret_cast<const f4*>(tgt_segs), cam_offsets, projections, sigma_p, sigma_a, spatial_k, 0, 0);
}

void oracle_collinearity(const float* segs, int n, float collin_s, float* out)
{
    const f4* s = reinterpret_cast<const f4*>(segs);
    for (int y = 0; y < n; ++y)
        for (int x = 0; x < n; ++x) {
            if (x == y) out[size_t(y) * n + x] = 0.0f;
            else if (x < y) { float r = collinearity_pair(s[x], s[y], collin_s * collin_s); out[size_t(y) * n + x] = r; out[size_t(x) * n + y] = r; }
        }
}

void oracle_diffusion(const int* ij, const float* w, int nnz, int n, float* out)
{
    std::list<Edge> A;
    for (int k = 0; k < nnz; ++k) { Edge e; e.i = ij[2 * k]; e.j = ij[2 * k + 1]; e.w = w[k]; A.push_back(e); }
    Sparse W = sparse_from_edges(A, n, false);
    Sparse P = rdd(W, n);
    for (size_t k = 0; k < P.e.size(); ++k) { out[3 * k] = P.e[k].x; out[3 * k + 1] = P.e[k].y; out[3 * k + 2] = P.e[k].z; }
}

float oracle_similarity(const double* h1, const double* v1, const double* h2, const double* v2, float sigma_a)
{
    Seg3D a, b; View va, vb;
    for (int i = 0; i < 3; ++i) { a.P1(i) = h1[i]; a.P2(i) = h1[3 + i]; a.dir(i) = h1[6 + i]; b.P1(i) = h2[i]; b.P2(i) = h2[3 + i]; b.dir(i) = h2[6 + i]; }
    a.depth_p1 = float(h1[9]); a.depth_p2 = float(h1[10]); b.depth_p1 = float(h2[9]); b.depth_p2 = float(h2[10]);
    va.k_lower = float(v1[0]); va.k_upper = float(v1[1]); va.median_depth = float(v1[2]);
    vb.k_lower = float(v2[0]); vb.k_upper = float(v2[1]); vb.median_depth = float(v2[2]);
    return Pipeline::similarity(a, &va, b, &vb, sigma_a);
}

}  // extern "C"
