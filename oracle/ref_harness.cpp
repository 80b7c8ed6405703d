// oracle/ref_harness.cpp  --  TEST INFRASTRUCTURE, not product code.
//
// Drives the REFERENCE's own code (compiled from /root/reference by oracle/Makefile into
// oracle/_ref/libline3d_ref.so) through its public API on a synthetic scene file and
// dumps what it computed as .npy arrays, so that oracle/oracle.cpp (our restatement) and
// the CUDA path can be pinned against the reference itself.  Everything in here is new
// glue: it calls Line3D::addImage / compute3Dmodel / getResult (line3D.h:58-101) and
// observes three operator-boundary functions whose link names the Makefile renames with
// -D (compute_pairwise_matches, replicator_dynamics_diffusion: cudawrapper.h:49-75;
// performClustering: clustering.h:121) so that the wrappers below can record their
// inputs/outputs and forward to the untouched originals.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <sstream>
#include <string>
#include <vector>

// the dumps need a handful of private members of L3D::Line3D (views_, best_match_, ...)
#define private public
#include "line3D.h"
#undef private

namespace L3D {
// the originals, renamed at compile time of cudawrapper.cu / clustering.cc
void compute_pairwise_matches_real(L3D::DataArray<float>*, L3D::DataArray<float>*, L3D::DataArray<float4>*,
                                   L3D::DataArray<float>*, L3D::DataArray<float>*, const float3,
                                   L3D::DataArray<float>*, L3D::DataArray<float>*, L3D::DataArray<int2>*,
                                   std::list<unsigned int>&, std::list<L3D::L3DMatchingPair>&,
                                   std::map<unsigned int, unsigned int>&, const unsigned int, const unsigned int,
                                   const float, const float, const float, const float, const float, float&,
                                   const bool, const std::string);
void replicator_dynamics_diffusion_real(L3D::SparseMatrix*& W, const bool verbose, const std::string prefix);
CLUniverse* performClustering_real(std::list<CLEdge> edges, int numNodes, float c);
}

// --------------------------------------------------------------------------------------
// npy writer
// --------------------------------------------------------------------------------------
static void write_npy(const std::string& path, const char* descr, size_t elem, const std::vector<size_t>& shape, const void* data)
{
    std::ostringstream h;
    h << "{'descr': '" << descr << "', 'fortran_order': False, 'shape': (";
    size_t n = 1;
    for (size_t i = 0; i < shape.size(); ++i) { h << shape[i] << (shape.size() == 1 || i + 1 < shape.size() ? "," : ""); n *= shape[i]; }
    h << "), }";
    std::string hs = h.str();
    size_t total = 10 + hs.size() + 1;
    size_t pad = (64 - total % 64) % 64;
    hs += std::string(pad, ' ');
    hs += "\n";
    std::ofstream f(path.c_str(), std::ios::binary);
    const unsigned char magic[8] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0};
    f.write(reinterpret_cast<const char*>(magic), 8);
    uint16_t hl = uint16_t(hs.size());
    f.write(reinterpret_cast<const char*>(&hl), 2);
    f.write(hs.data(), std::streamsize(hs.size()));
    if (n) f.write(static_cast<const char*>(data), std::streamsize(n * elem));
}
static void npy_f64(const std::string& p, const std::vector<double>& v, size_t cols)
{ write_npy(p, "<f8", 8, {cols ? v.size() / cols : 0, cols}, v.data()); }
static void npy_i32(const std::string& p, const std::vector<int32_t>& v, size_t cols)
{ write_npy(p, "<i4", 4, {cols ? v.size() / cols : 0, cols}, v.data()); }
static void npy_f32(const std::string& p, const std::vector<float>& v, size_t cols)
{ write_npy(p, "<f4", 4, {cols ? v.size() / cols : 0, cols}, v.data()); }

// --------------------------------------------------------------------------------------
// recording hooks
// --------------------------------------------------------------------------------------
struct Recorder {
    bool on = false;
    std::string dir;
    double t_match = 0.0, t_diffusion = 0.0, t_cluster = 0.0;
    std::vector<int32_t> view_log;       // per performMatching call: vID, #toBeMatched, #pre-existing, #filtered
    std::vector<float> median_log;       // per call: median depth returned
    std::vector<double> filtered;        // rows: vID, seg1, cam2(global), seg2, dx,dy,dz,dw, conf
    std::vector<double> edges_pre;       // W handed to diffusion (row, col, val) in stored order
    std::vector<double> edges_cluster;   // edge list handed to clustering (i, j, w) in list order
    int cluster_nodes = 0;
};
static Recorder g_rec;

namespace L3D {
void compute_pairwise_matches(L3D::DataArray<float>* segments_src, L3D::DataArray<float>* RtKinv_src,
                              L3D::DataArray<float4>* segments_tgt, L3D::DataArray<float>* RtKinv_tgt,
                              L3D::DataArray<float>* camCenters_tgt, const float3 camCenter_src,
                              L3D::DataArray<float>* fundamentals, L3D::DataArray<float>* projections,
                              L3D::DataArray<int2>* offsets, std::list<unsigned int>& toBeMatched,
                              std::list<L3D::L3DMatchingPair>& matches, std::map<unsigned int, unsigned int>& local2global,
                              const unsigned int maxSegments, const unsigned int vID,
                              const float uncertainty_k_upper, const float uncertainty_k_lower,
                              const float sigma_p, const float sigma_a, const float spatial_k, float& median_depth,
                              const bool verbose, const std::string prefix)
{
    const size_t n_pre = matches.size();
    auto t0 = std::chrono::steady_clock::now();
    compute_pairwise_matches_real(segments_src, RtKinv_src, segments_tgt, RtKinv_tgt, camCenters_tgt, camCenter_src,
                                  fundamentals, projections, offsets, toBeMatched, matches, local2global, maxSegments, vID,
                                  uncertainty_k_upper, uncertainty_k_lower, sigma_p, sigma_a, spatial_k, median_depth,
                                  verbose, prefix);
    g_rec.t_match += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (!g_rec.on) return;
    g_rec.view_log.push_back(int32_t(vID));
    g_rec.view_log.push_back(int32_t(toBeMatched.size()));
    g_rec.view_log.push_back(int32_t(n_pre));
    g_rec.view_log.push_back(int32_t(matches.size()));
    g_rec.median_log.push_back(median_depth);
    for (std::list<L3D::L3DMatchingPair>::iterator it = matches.begin(); it != matches.end(); ++it) {
        const double row[9] = {double(vID), double(it->segID1_), double(it->camID2_), double(it->segID2_),
                               it->depths_.x, it->depths_.y, it->depths_.z, it->depths_.w, it->confidence_};
        g_rec.filtered.insert(g_rec.filtered.end(), row, row + 9);
    }
}

void replicator_dynamics_diffusion(L3D::SparseMatrix*& W, const bool verbose, const std::string prefix)
{
    if (g_rec.on)
        for (unsigned int i = 0; i < W->entries()->width(); ++i) {
            float4 e = W->entries()->dataCPU(i, 0)[0];
            g_rec.edges_pre.push_back(e.x); g_rec.edges_pre.push_back(e.y); g_rec.edges_pre.push_back(e.z);
        }
    auto t0 = std::chrono::steady_clock::now();
    replicator_dynamics_diffusion_real(W, verbose, prefix);
    g_rec.t_diffusion += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

CLUniverse* performClustering(std::list<CLEdge> edges, int numNodes, float c)
{
    if (g_rec.on) {
        g_rec.cluster_nodes = numNodes;
        for (std::list<CLEdge>::iterator it = edges.begin(); it != edges.end(); ++it) {
            g_rec.edges_cluster.push_back(it->i_); g_rec.edges_cluster.push_back(it->j_); g_rec.edges_cluster.push_back(it->w_);
        }
    }
    auto t0 = std::chrono::steady_clock::now();
    CLUniverse* u = performClustering_real(edges, numNodes, c);
    g_rec.t_cluster += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return u;
}
}  // namespace L3D

// --------------------------------------------------------------------------------------
// scene file (line3d_b200/scenes.py)
// --------------------------------------------------------------------------------------
struct SceneView {
    uint32_t id, width, height;
    double K[9], R[9], t[3];
    std::vector<float> segs;
    std::vector<uint32_t> wps;
};
static bool read_scene(const char* path, std::vector<SceneView>& views)
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

static Eigen::Matrix3d mat3(const double* m)
{
    Eigen::Matrix3d M;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) M(r, c) = m[r * 3 + c];
    return M;
}

extern "C" {

// Runs the whole reference pipeline on a scene.  Segments enter the way the reference's
// own cache does: an L3DSegments object (built by the reference's constructor, which also
// runs its collinearity kernel) is serialised under the cache file name addImage looks
// for, and addImage is called with loadAndStoreSegments = true (line3D.cc:143-168).
// timings[0..4] = addImage total, compute3Dmodel total, compute_pairwise_matches total,
// diffusion, clustering (seconds).  Returns the number of final 3-D lines, or <0 on error.
int ref_run_scene(const char* scene_path, const char* data_dir, int neighbors, float unc_upper, float unc_lower,
                  float sigma_p, float sigma_a, float min_baseline, int collinearity, int diffusion, int verbose,
                  const char* dump_dir, double* timings)
{
    std::vector<SceneView> views;
    if (!read_scene(scene_path, views)) return -1;
    g_rec = Recorder();
    g_rec.on = (dump_dir != 0 && dump_dir[0] != 0);
    if (g_rec.on) g_rec.dir = dump_dir;

    std::streambuf* old_cout = 0;
    std::ostringstream sink;
    if (!verbose) old_cout = std::cout.rdbuf(sink.rdbuf());

    L3D::Line3D* l3d = new L3D::Line3D(data_dir, neighbors, unc_upper, unc_lower, sigma_p, sigma_a, min_baseline,
                                       collinearity != 0, verbose != 0);
    auto t0 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < views.size(); ++i) {
        SceneView& v = views[i];
        std::list<float4> segs;
        for (size_t s = 0; s + 3 < v.segs.size(); s += 4) segs.push_back(make_float4(v.segs[s], v.segs[s + 1], v.segs[s + 2], v.segs[s + 3]));
        {
            L3D::L3DSegments cache(segs, collinearity != 0);
            std::stringstream str;
            str << data_dir << "/segments_" << v.id << "_" << v.width << "x" << v.height << (collinearity ? "_coll1.bin" : "_coll0.bin");
            L3D::serializeToFile(str.str(), cache);
        }
        cv::Mat image(int(v.height), int(v.width), CV_8UC3);
        std::list<unsigned int> wps(v.wps.begin(), v.wps.end());
        l3d->addImage(v.id, image, mat3(v.K), mat3(v.R), Eigen::Vector3d(v.t[0], v.t[1], v.t[2]), wps,
                      std::max(v.width, v.height), true);
    }
    auto t1 = std::chrono::steady_clock::now();
    l3d->compute3Dmodel(diffusion != 0);
    auto t2 = std::chrono::steady_clock::now();
    if (timings) {
        timings[0] = std::chrono::duration<double>(t1 - t0).count();
        timings[1] = std::chrono::duration<double>(t2 - t1).count();
        timings[2] = g_rec.t_match; timings[3] = g_rec.t_diffusion; timings[4] = g_rec.t_cluster;
    }
    std::list<L3D::L3DFinalLine3D> result;
    l3d->getResult(result);
    const int n_lines = int(result.size());

    if (g_rec.on) {
        const std::string d = g_rec.dir + "/";
        // visual neighbours, per-view scalars and matrices (after scene normalisation)
        std::vector<int32_t> vn;
        for (auto& a : l3d->visual_neighbors_) for (auto& b : a.second) { vn.push_back(int32_t(a.first)); vn.push_back(int32_t(b.first)); }
        npy_i32(d + "visual_neighbors.npy", vn, 2);
        std::vector<double> vw;   // id, median, k_upper, k_lower, spatial_k(2*sigma_p), C(3), RtKinv(9), P(12), K(9), R(9), t(3)
        for (auto& a : l3d->views_) {
            L3D::L3DView* v = a.second;
            vw.push_back(a.first); vw.push_back(v->median_depth()); vw.push_back(v->uncertainty_k_upper()); vw.push_back(v->uncertainty_k_lower());
            vw.push_back(v->specificSpatialUncertaintyK(2.0f * sigma_p));
            Eigen::Vector3d C = v->C(); for (int i = 0; i < 3; ++i) vw.push_back(C(i));
            Eigen::Matrix3d M = v->RtKinv(); for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(M(r, c));
            Eigen::MatrixXd P = v->P(); for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) vw.push_back(P(r, c));
            M = v->K(); for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(M(r, c));
            M = v->R(); for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(M(r, c));
            Eigen::Vector3d t = v->t(); for (int i = 0; i < 3; ++i) vw.push_back(t(i));
        }
        npy_f64(d + "views.npy", vw, 50);
        // fundamentals
        std::vector<double> fs;
        for (auto& a : l3d->fundamentals_) for (auto& b : a.second) {
            fs.push_back(a.first); fs.push_back(b.first);
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) fs.push_back(b.second(r, c));
        }
        npy_f64(d + "fundamentals.npy", fs, 11);
        // collinearity maps
        std::vector<double> co;
        for (auto& a : l3d->views_) {
            std::map<unsigned int, std::map<unsigned int, float> >* m = a.second->seg_collinearities();
            if (!m) continue;
            for (auto& r : *m) for (auto& c : r.second) { co.push_back(a.first); co.push_back(r.first); co.push_back(c.first); co.push_back(c.second); }
        }
        npy_f64(d + "collinearity.npy", co, 4);
        // matching log
        npy_i32(d + "match_log.npy", g_rec.view_log, 4);
        npy_f32(d + "match_median.npy", g_rec.median_log, 1);
        npy_f64(d + "filtered_matches.npy", g_rec.filtered, 9);
        // best hypotheses (map order)
        std::vector<double> bm;
        for (auto& a : l3d->best_match_) {
            L3D::L3DCorrespondenceRRW c = a.second;
            L3D::L3DSegment3D s = c.src_seg3D();
            bm.push_back(a.first.camID()); bm.push_back(a.first.segID());
            bm.push_back(c.tgt().camID()); bm.push_back(c.tgt().segID());
            bm.push_back(c.score()); bm.push_back(c.valid() ? 1.0 : 0.0);
            for (int i = 0; i < 3; ++i) bm.push_back(s.P1_(i));
            for (int i = 0; i < 3; ++i) bm.push_back(s.P2_(i));
            for (int i = 0; i < 3; ++i) bm.push_back(s.dir_(i));
            bm.push_back(s.depth_p1_); bm.push_back(s.depth_p2_);
        }
        npy_f64(d + "best_match.npy", bm, 17);
        std::vector<int32_t> pc;
        for (auto& a : l3d->potential_correspondences_) for (auto& b : a.second) {
            pc.push_back(int32_t(a.first.camID())); pc.push_back(int32_t(a.first.segID()));
            pc.push_back(int32_t(b.first.camID())); pc.push_back(int32_t(b.first.segID()));
        }
        npy_i32(d + "potential.npy", pc, 4);
        npy_f64(d + "edges_pre_diffusion.npy", g_rec.edges_pre, 3);
        npy_f64(d + "edges_cluster.npy", g_rec.edges_cluster, 3);
        std::vector<int32_t> nn(1, g_rec.cluster_nodes); npy_i32(d + "cluster_nodes.npy", nn, 1);
        // final lines
        std::vector<double> s3; std::vector<int32_t> s2;
        int li = 0;
        for (auto it = result.begin(); it != result.end(); ++it, ++li) {
            L3D::L3DFinalLine3D fl = *it;
            for (auto& p : *fl.segments3D()) {
                s3.push_back(li);
                for (int i = 0; i < 3; ++i) s3.push_back(p.first(i));
                for (int i = 0; i < 3; ++i) s3.push_back(p.second(i));
            }
            for (auto& q : *fl.segments2D()) { s2.push_back(li); s2.push_back(int32_t(q.camID())); s2.push_back(int32_t(q.segID())); }
        }
        npy_f64(d + "lines_seg3d.npy", s3, 7);
        npy_i32(d + "lines_seg2d.npy", s2, 3);
        // transform
        std::vector<double> tr;
        tr.push_back(l3d->transf_scale_);
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) tr.push_back(l3d->transf_R_(r, c));
        for (int i = 0; i < 3; ++i) tr.push_back(l3d->transf_t_(i));
        npy_f64(d + "transform.npy", tr, 13);
    }
    delete l3d;
    if (old_cout) std::cout.rdbuf(old_cout);
    g_rec.on = false;
    return n_lines;
}

// Direct call of the reference's collinearity operator (cudawrapper.cu:833-855) on n
// segments; out is a dense row-major n*n float array as the reference downloads it.
int ref_collinearity(const float* segs, int n, float collin_s, float* out)
{
    L3D::DataArray<float>* s = new L3D::DataArray<float>(4, n);
    for (int i = 0; i < n; ++i) for (int c = 0; c < 4; ++c) s->dataCPU(c, i)[0] = segs[i * 4 + c];
    L3D::DataArray<float>* rel = new L3D::DataArray<float>(n, n, true);
    s->upload();
    L3D::compute_collinearity(s, rel, collin_s);
    rel->download();
    for (int y = 0; y < n; ++y) for (int x = 0; x < n; ++x) out[size_t(y) * n + x] = rel->dataCPU(x, y)[0];
    delete rel; delete s;
    return 0;
}

// Direct call of the reference's diffusion (cudawrapper.cu:1131-1191) on an edge list
// (i, j, w) with n nodes, through the reference's SparseMatrix (sparsematrix.cc:63-117).
// out receives nnz rows (row, col, val) in the returned matrix's stored order.
int ref_diffusion(const int* ij, const float* w, int nnz, int n, float* out)
{
    std::list<L3D::CLEdge> A;
    for (int k = 0; k < nnz; ++k) { L3D::CLEdge e; e.i_ = ij[2 * k]; e.j_ = ij[2 * k + 1]; e.w_ = w[k]; A.push_back(e); }
    L3D::SparseMatrix* W = new L3D::SparseMatrix(A, n);
    L3D::replicator_dynamics_diffusion_real(W, false, "");
    W->download();
    for (unsigned int i = 0; i < W->entries()->width(); ++i) {
        float4 e = W->entries()->dataCPU(i, 0)[0];
        out[3 * i] = e.x; out[3 * i + 1] = e.y; out[3 * i + 2] = e.z;
    }
    delete W;
    return 0;
}

}  // extern "C"
