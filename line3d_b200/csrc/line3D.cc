// line3d_b200/csrc/line3D.cc -- host side of the Line3D API over the C ABI (l3d_b200.h).
//
// Fresh implementation of the reference's orchestration (line3D.cc, view.cc, segments.h,
// clustering.cc of /root/reference; cited per function).  Containers are flat (vectors,
// sorted arrays, hash sets) instead of the reference's nested std::map / std::list, and
// views exchange matches through device memory instead of files; the observable behaviour
// -- messages, guards, iteration orders, tie-breaking -- is the reference's.
#include "line3D.h"

#include "../../include/l3d_b200.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <set>
#include <sstream>
#include <stdexcept>
#include <thread>

namespace L3D
{
namespace
{
typedef Eigen::Matrix3d M3;
typedef Eigen::Vector3d V3;

// Worker threads for the host stages: up to 16, and a fair share of the cores when several processes of one job sit on
// the machine (torchrun sets LOCAL_WORLD_SIZE; L3D_HOST_THREADS overrides).
inline size_t host_threads()
{
    static const size_t n = [] {
        if (const char* e = getenv("L3D_HOST_THREADS")) { const long v = atol(e); if (v > 0) return size_t(v); }
        size_t hw = std::max<size_t>(1, std::thread::hardware_concurrency()), share = 1;
        if (const char* e = getenv("LOCAL_WORLD_SIZE")) { const long v = atol(e); if (v > 0) share = size_t(v); }
        return std::max<size_t>(1, std::min<size_t>(hw / share, 16));
    }();
    return n;
}
inline double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
inline uint64_t seg_key(unsigned cam, unsigned seg) { return (uint64_t(cam) << 32) | uint64_t(seg); }

struct CollinEntry { unsigned other; float w; };

// one image (view.h:40-153)
struct ViewData
{
    unsigned id, width, height;
    M3 K, Kinv, R, Rt, RtKinv, R0; V3 t, C, pp, t0; Eigen::MatrixXd P;
    float unc_upper_px, unc_lower_px, k_upper, k_lower, median_depth;
    std::vector<float4> segs;
    std::vector<std::vector<CollinEntry> > collin;     // per segment, ascending partner id (segments.h:82-98)
    int index;                                         // position in the l3d_set_views array

    // view.cc:90-147: lateral 3-D displacement at depth 1 of a pixel dist_px away from the principal point
    double shifted(float dist_px) const
    {
        V3 n = RtKinv * pp; n.normalize();
        V3 Pl = C + n;
        V3 sh(pp.x() + dist_px, pp.y(), 1.0);
        V3 d = RtKinv * sh; d.normalize();
        double tt = (Pl.dot(n) - n.dot(C)) / (n.dot(d));
        V3 Q = C + tt * d;
        return (Pl - Q).norm();
    }
    void defineUncertainty() { k_upper = float(shifted(unc_upper_px)); k_lower = float(shifted(unc_lower_px)); }
    // view.cc:16-34
    void init()
    {
        pp(0) = float(width) / 2.0f; pp(1) = float(height) / 2.0f; pp(2) = 1.0;
        Kinv = K.inverse(); Rt = R.transpose(); RtKinv = Rt * Kinv; C = Rt * (-1.0 * t);
        Eigen::MatrixXd Rpt(3, 4); Rpt.block<3, 3>(0, 0) = R; Rpt.block<3, 1>(0, 3) = t;
        P = K * Rpt;
        median_depth = 1.0f;
        defineUncertainty();
    }
    // view.cc:227-261
    void transform(Eigen::Matrix4d& Qinv, double scale)
    {
        t *= scale;
        Eigen::MatrixXd Rtm(3, 4); Rtm.block<3, 3>(0, 0) = R; Rtm.block<3, 1>(0, 3) = t;
        Rtm *= Qinv;
        R = Rtm.block<3, 3>(0, 0); t = Rtm.block<3, 1>(0, 3);
        Rt = R.transpose(); P = K * Rtm; RtKinv = Rt * Kinv; C = Rt * (-1.0 * t);
        defineUncertainty();
    }
};

struct BestMatch { bool has; unsigned tgt_cam, tgt_seg; float conf, d1, d2; };
struct Hypothesis { unsigned cam, seg, tgt_cam, tgt_seg; float score; V3 P1, P2, dir; float d1, d2; };
struct Edge { int i, j; float w; };

// tiny .npy writer for debugDump (same files as oracle/ref_harness.cpp writes)
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

std::string cacheName(const std::string& dir, unsigned id, unsigned w, unsigned h, bool collin)
{
    std::stringstream s;
    s << dir << "/segments_" << id << "_" << w << "x" << h << (collin ? "_coll1.bin" : "_coll0.bin");   // line3D.cc:144-150
    return s.str();
}
}  // namespace

struct Line3D::Impl
{
    bool verbose, computation, use_collinearity, measure_only, match_only, flush_l2, keep_debug, stage_timing;
    bool split; int split_rank, split_world;          // several GPUs on one scene (setSplit)
    std::string prefix, separator, data_directory;
    int matching_neighbors; float min_baseline, unc_upper, unc_lower, sigma_p, sigma_a;
    int shard_rank, shard_world; std::set<unsigned> my_views;
    l3d_ctx* ctx;

    std::map<unsigned, ViewData*> views;
    std::map<unsigned, unsigned> num_wps;
    std::map<unsigned, std::map<unsigned, unsigned> > common_wps;
    std::map<unsigned, std::map<unsigned, float> > view_sim;
    std::map<unsigned, std::map<unsigned, bool> > wp2views;
    std::map<unsigned, std::vector<unsigned> > vn;                 // visual neighbours, ascending
    std::map<unsigned, std::map<unsigned, M3> > fund;

    std::map<unsigned, std::vector<BestMatch> > best;              // per view, per segment
    struct PC { unsigned seg, tcam, tseg; };                          // (this view, seg) <-> (tcam, tseg), line3D.cc:861-865
    std::vector<std::vector<PC> > pc_fwd, pc_rev, pc;                // per view index: as produced / as received / merged, sorted, unique
    std::vector<ViewData*> view_at;                                  // view index -> view
    std::vector<int> id2idx;                                         // view id -> view index or -1
    std::vector<Hypothesis> hyps;                                  // (cam, seg) ascending
    std::vector<std::vector<int> > hyp_of;                            // [view index][segment] -> hypothesis or -1

    std::list<L3DFinalLine3D> result;
    double tr_scale, tr_scale_inv; M3 tr_R, tr_Rinv; V3 tr_t, tr_tneg; Eigen::Matrix4d Qinv;

    // dumps / stats
    std::vector<int32_t> match_log; std::vector<float> match_median; std::vector<double> filtered_dump, edges_pre, edges_cluster;
    int cluster_nodes;
    L3DRunStats stats;

    Impl() : match_only(false), split(false), split_rank(0), split_world(1), shard_rank(0), shard_world(1), ctx(0), cluster_nodes(0) { memset(&stats, 0, sizeof stats); }

    void fail(const char* what)
    {
        std::string msg = std::string(what) + ": " + l3d_last_error(ctx);
        std::cerr << prefix << msg << std::endl;
        throw std::runtime_error(msg);
    }

    // line3D.cc:1874-1935
    void processWorldpoints(unsigned vid, std::list<unsigned>& wps)
    {
        num_wps[vid] = 0;
        for (std::list<unsigned>::iterator it = wps.begin(); it != wps.end(); ++it) {
            std::map<unsigned, bool>& seen = wp2views[*it];
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

    // segments.h:60-102 with the dense relation replaced by the compacted pair list
    void buildCollinearity(ViewData* v)
    {
        v->collin.assign(v->segs.size(), std::vector<CollinEntry>());
        if (!use_collinearity || v->segs.size() < 2) return;
        const l3d_collin_pair* pairs = 0; size_t n = 0;
        if (l3d_collinearity(ctx, reinterpret_cast<const float*>(v->segs.data()), int(v->segs.size()), L3D_DEF_COLLINEARITY_S, &pairs, &n)) fail("l3d_collinearity");
        for (size_t k = 0; k < n; ++k) {
            CollinEntry a = {unsigned(pairs[k].j), pairs[k].value}, b = {unsigned(pairs[k].i), pairs[k].value};
            v->collin[pairs[k].i].push_back(a); v->collin[pairs[k].j].push_back(b);
        }
        for (size_t s = 0; s < v->collin.size(); ++s)
            std::sort(v->collin[s].begin(), v->collin[s].end(), [](const CollinEntry& x, const CollinEntry& y) { return x.other < y.other; });
    }

    // common part of addImage / addImage_fixed_sim (line3D.cc:95-217, 220-342)
    ViewData* addView(unsigned imageID, const cv::Mat& image, const M3& K, const M3& R, const V3& t, int maxImgWidth, bool loadAndStore)
    {
        unsigned new_width = image.cols, new_height = image.rows;
        if (maxImgWidth > 0 && std::max(image.rows, image.cols) > maxImgWidth) {                     // line3D.cc:134-139
            float scale = float(maxImgWidth) / fmax(image.rows, image.cols);
            new_width = round(float(image.cols) * scale); new_height = round(float(image.rows) * scale);
        }
        std::cout << prefix << "adding image [" << imageID << "] [" << new_width << "x" << new_height << "]" << std::endl;
        const std::string file = cacheName(data_directory, imageID, new_width, new_height, use_collinearity);
        boost::filesystem::path fpath(file);
        if (boost::filesystem::exists(fpath) && !loadAndStore) boost::filesystem::remove(fpath);     // line3D.cc:154-157
        std::vector<float4> segs;
        if (boost::filesystem::exists(fpath) && loadAndStore) {
            if (verbose) std::cout << prefix << "segments data found" << std::endl;
            if (!readSegmentCache(file, segs)) { std::cerr << prefix << "segment cache unreadable: " << file << std::endl; return 0; }
        } else {
            if (verbose) std::cout << prefix << "performing line segment detection..." << std::endl;
            // line3D.cc:174-190.  2-D line detection (OpenCV LSD) is upstream of this library and not
            // available on this image: segments have to arrive through the cache.
            std::cerr << prefix << "no line segment detector in this build: provide " << file << " (Line3D::writeSegmentCache)" << std::endl;
            return 0;
        }
        if (verbose) std::cout << prefix << "#segments: " << segs.size() << " (final)" << std::endl;
        ViewData* v = new ViewData();
        v->id = imageID; v->width = image.cols; v->height = image.rows;                              // line3D.cc:202-207
        v->K = K; v->R = R; v->t = t; v->R0 = R; v->t0 = t; v->unc_upper_px = unc_upper; v->unc_lower_px = unc_lower;
        v->segs.swap(segs); v->index = -1;
        buildCollinearity(v);
        v->init();
        views[imageID] = v;
        if (verbose) {
            std::cout << prefix << "minimum uncertainty in depth=1: " << v->k_lower << std::endl;
            std::cout << prefix << "maximum uncertainty in depth=1: " << v->k_upper << std::endl;
        }
        return v;
    }

    static bool readSegmentCache(const std::string& file, std::vector<float4>& segs)
    {
        std::ifstream f(file.c_str(), std::ios::binary);
        if (!f) return false;
        char magic[8]; f.read(magic, 8);
        if (memcmp(magic, "L3DSEG01", 8) != 0) return false;
        uint32_t n = 0, flags = 0; f.read(reinterpret_cast<char*>(&n), 4); f.read(reinterpret_cast<char*>(&flags), 4);
        segs.resize(n);
        if (n) f.read(reinterpret_cast<char*>(segs.data()), std::streamsize(size_t(n) * sizeof(float4)));
        return bool(f);
    }

    bool guards(unsigned imageID, const cv::Mat& image, bool linked, const char* what)
    {
        if (computation) { std::cerr << "reconstruction already performed! cannot add more images (try reset first)" << std::endl; return false; }   // line3D.cc:101-105
        if (views.size() == 0) std::cout << prefix << ">>> LOADING DATA <<<" << std::endl;
        if (views.find(imageID) != views.end()) { std::cerr << prefix << "imageID already in use!" << std::endl; return false; }                   // :111-115
        if (!linked) { std::cerr << prefix << "unlinked images cannot be added! (" << what << ")" << std::endl; return false; }                      // :116-120
        if (image.rows == 0 || image.cols == 0) { std::cerr << prefix << "image is empty!" << std::endl; return false; }                           // :123-127
        return true;
    }

    float baseline(unsigned a, unsigned b) { return float((views[a]->C - views[b]->C).norm()); }     // view.cc:446-449

    // line3D.cc:476-549
    void findVisualNeighbors()
    {
        vn.clear();
        std::cout << prefix << separator << std::endl;
        std::cout << prefix << ">>> FINDING VISUAL NEIGHBORS <<<" << std::endl;
        for (std::map<unsigned, std::map<unsigned, unsigned> >::iterator it = common_wps.begin(); it != common_wps.end(); ++it) {
            if (view_sim.find(it->first) != view_sim.end()) continue;
            std::cout << prefix << "computing similarities for image [" << it->first << "]" << std::endl;
            for (std::map<unsigned, unsigned>::iterator n = it->second.begin(); n != it->second.end(); ++n) {
                float sim = 2.0f * float(n->second) / (float(num_wps[it->first] + num_wps[n->first]));
                if (sim > L3D_EPS) view_sim[it->first][n->first] = sim;
            }
        }
        // The selection itself (baseline against the view and against the neighbours accepted so far, ranking by
        // similarity, cut to k; line3D.cc:498-546) runs on the device for all views at once: l3d_visual_neighbors gets
        // every view's co-visible views in ascending order with their similarities, in CSR form.
        std::vector<std::map<unsigned, std::map<unsigned, float> >::iterator> todo;
        for (std::map<unsigned, std::map<unsigned, float> >::iterator s = view_sim.begin(); s != view_sim.end(); ++s) { vn[s->first]; todo.push_back(s); }
        std::vector<size_t> found(todo.size(), 0);
        {
            std::map<unsigned, int> idx_of; std::vector<unsigned> id_of; std::vector<l3d_camera> cams;
            for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) {
                idx_of[it->first] = int(id_of.size()); id_of.push_back(it->first);
                cams.push_back(camera_of(it->second));
            }
            const int nv = int(id_of.size());
            std::vector<int32_t> start(size_t(nv) + 1, 0), other, count(static_cast<size_t>(nv), 0); std::vector<float> sim;
            std::vector<int> todo_of(static_cast<size_t>(nv), -1);
            for (size_t q = 0; q < todo.size(); ++q) { std::map<unsigned, int>::iterator f = idx_of.find(todo[q]->first); if (f != idx_of.end()) todo_of[f->second] = int(q); }
            for (int v = 0; v < nv; ++v) {
                if (todo_of[v] >= 0)
                    for (std::map<unsigned, float>::iterator n = todo[todo_of[v]]->second.begin(); n != todo[todo_of[v]]->second.end(); ++n) {
                        std::map<unsigned, int>::iterator f = idx_of.find(n->first);
                        if (f == idx_of.end() || f->second == v) continue;          // a view that was never added has no camera
                        other.push_back(f->second); sim.push_back(n->second);
                    }
                start[size_t(v) + 1] = int32_t(other.size());
            }
            std::vector<int32_t> out(std::max<size_t>(other.size(), 1));
            if (l3d_visual_neighbors(ctx, nv, cams.data(), start.data(), other.data(), sim.data(), min_baseline, matching_neighbors, count.data(), out.data())) fail("l3d_visual_neighbors");
            for (int v = 0; v < nv; ++v) {
                if (todo_of[v] < 0) continue;
                std::vector<unsigned>& lst = vn.find(id_of[v])->second;
                for (int e = 0; e < count[v]; ++e) lst.push_back(id_of[out[size_t(start[v]) + e]]);      // ascending index = ascending id
                found[todo_of[v]] = size_t(count[v]);
            }
        }
        for (size_t q = 0; q < todo.size(); ++q) {
            std::cout << prefix << "setting VNs for image [" << todo[q]->first << "]" << std::endl;
            if (verbose) std::cout << prefix << found[q] << " visual neighbors found" << std::endl;
        }
    }

    // Deal the connected components of the (undirected) neighbour graph out to the ranks.
    void assignShard()
    {
        my_views.clear();
        if (split) {                // every rank works on every view (its share of the view's segments)
            for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) my_views.insert(it->first);
            return;
        }
        std::map<unsigned, std::set<unsigned> > adj;
        for (std::map<unsigned, std::vector<unsigned> >::iterator it = vn.begin(); it != vn.end(); ++it)
            for (size_t k = 0; k < it->second.size(); ++k) { adj[it->first].insert(it->second[k]); adj[it->second[k]].insert(it->first); }
        std::map<unsigned, int> comp; std::vector<double> cost; std::vector<std::vector<unsigned> > members;
        for (std::map<unsigned, std::set<unsigned> >::iterator it = adj.begin(); it != adj.end(); ++it) {
            if (comp.count(it->first)) continue;
            const int c = int(cost.size()); cost.push_back(0.0); members.push_back(std::vector<unsigned>());
            std::vector<unsigned> stack(1, it->first); comp[it->first] = c;
            while (!stack.empty()) {
                const unsigned v = stack.back(); stack.pop_back();
                members[c].push_back(v);
                std::map<unsigned, std::vector<unsigned> >::iterator nv = vn.find(v);
                if (nv != vn.end()) for (size_t k = 0; k < nv->second.size(); ++k) cost[c] += double(views[v]->segs.size()) * double(views[nv->second[k]]->segs.size());
                for (std::set<unsigned>::iterator n = adj[v].begin(); n != adj[v].end(); ++n) if (!comp.count(*n)) { comp[*n] = c; stack.push_back(*n); }
            }
        }
        std::vector<int> order(cost.size());
        for (size_t c = 0; c < order.size(); ++c) order[c] = int(c);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
        std::vector<double> load(size_t(std::max(shard_world, 1)), 0.0);
        for (size_t k = 0; k < order.size(); ++k) {
            size_t best = 0;
            for (size_t r = 1; r < load.size(); ++r) if (load[r] < load[best]) best = r;
            load[best] += cost[order[k]];
            if (int(best) == shard_rank) my_views.insert(members[order[k]].begin(), members[order[k]].end());
        }
    }

    // line3D.cc:552-617, 1694-1779
    void transformGeometry()
    {
        std::cout << prefix << separator << std::endl;
        std::cout << prefix << ">>> TRANSFORMING SCENE GEOMETRY <<<" << std::endl;
        fund.clear();
        const double size = double(views.size());
        V3 m(0.0, 0.0, 0.0);
        std::vector<V3> in_pts;
        for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) { m += it->second->C; in_pts.push_back(it->second->C); }
        m /= size;
        double q = 0.0;
        for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) q += (it->second->C - m).norm();
        q /= size;
        q = sqrtf(2.0) / q;
        Eigen::MatrixXd T(4, 4);
        T(0,0) = q; T(0,1) = 0.0; T(0,2) = 0.0; T(0,3) = -q * m(0);
        T(1,0) = 0.0; T(1,1) = q; T(1,2) = 0.0; T(1,3) = -q * m(1);
        T(2,0) = 0.0; T(2,1) = 0.0; T(2,2) = q; T(2,3) = -q * m(2);
        T(3,0) = 0.0; T(3,1) = 0.0; T(3,2) = 0.0; T(3,3) = 1.0;
        std::vector<V3> out_pts; V3 cog_out(0.0, 0.0, 0.0);
        for (size_t i = 0; i < in_pts.size(); ++i) {
            Eigen::Vector4d tv = T * Eigen::Vector4d(in_pts[i](0), in_pts[i](1), in_pts[i](2), 1.0);
            V3 t3(tv(0), tv(1), tv(2)); cog_out += t3; out_pts.push_back(t3);
        }
        cog_out /= size;
        if (verbose) {
            std::cout << prefix << "COG_in   = [" << m.x() << " " << m.y() << " " << m.z() << "]" << std::endl;
            std::cout << prefix << "COG_out  = [" << cog_out.x() << " " << cog_out.y() << " " << cog_out.z() << "]" << std::endl;
            std::cout << prefix << "variance = " << q << std::endl;
        }
        std::cout << prefix << "computing similarity transform..." << std::endl;
        const size_t n = in_pts.size();
        double ssum = 0.0;
        for (size_t i = 0; i < n; ++i) ssum += (out_pts[i] - cog_out).norm() / (in_pts[i] - m).norm();
        tr_scale = ssum / double(n);
        V3 cog_in = m; cog_in *= tr_scale;
        for (size_t i = 0; i < n; ++i) in_pts[i] *= tr_scale;
        for (size_t i = 0; i < n; ++i) { in_pts[i] -= cog_in; out_pts[i] -= cog_out; }
        M3 H; H.setZero();
        for (size_t i = 0; i < n; ++i) { M3 outer = out_pts[i] * in_pts[i].transpose(); H += outer; }
        Eigen::JacobiSVD<Eigen::MatrixXd> svd(H, Eigen::ComputeThinU | Eigen::ComputeThinV);
        M3 Vt = svd.matrixV().transpose(), U = svd.matrixU();
        tr_R = U * Vt;
        if (tr_R.determinant() < 0) { Vt.row(2) *= -1; tr_R = U * Vt; }
        tr_t = cog_out - tr_R * cog_in;
        tr_t /= tr_scale;
        Eigen::Matrix4d Q;
        Q << tr_R.row(0), tr_t.x() * tr_scale, tr_R.row(1), tr_t.y() * tr_scale, tr_R.row(2), tr_t.z() * tr_scale, 0.0, 0.0, 0.0, 1.0;
        Qinv = Q.inverse();
        tr_scale_inv = 1.0 / tr_scale; tr_Rinv = tr_R.transpose(); tr_tneg = -tr_t;
        for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) it->second->transform(Qinv, tr_scale);
    }
    V3 inverseTransform(const V3& P) { return tr_Rinv * (P * tr_scale_inv + tr_tneg); }              // line3D.cc:1782-1786

    // line3D.cc:1968-1993
    static l3d_camera camera_of(const ViewData* v)
    {
        l3d_camera c;
        for (int r = 0; r < 3; ++r) { for (int k = 0; k < 3; ++k) { c.K[r * 3 + k] = v->K(r, k); c.R[r * 3 + k] = v->R(r, k); } c.t[r] = v->t(r); c.C[r] = v->C(r); }
        return c;
    }

    // line3D.cc:1949-1993 for every pair the schedule needs and that is not known yet, in one call: the pair a view
    // meets first is computed, the reverse pair gets the transpose (:1959-1962)
    void prepareFundamentals()
    {
        std::vector<l3d_camera> cams; std::map<unsigned, int> idx_of;
        for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) { idx_of[it->first] = int(cams.size()); cams.push_back(camera_of(it->second)); }
        std::vector<int32_t> pairs; std::vector<std::pair<unsigned, unsigned> > ids; std::set<std::pair<unsigned, unsigned> > fresh;
        for (std::map<unsigned, std::vector<unsigned> >::iterator it = vn.begin(); it != vn.end(); ++it) {
            const unsigned vid = it->first;
            if (it->second.empty() || !my_views.count(vid)) continue;
            for (size_t c = 0; c < it->second.size(); ++c) {
                const unsigned n = it->second[c];
                if (fund[vid].find(n) != fund[vid].end() || fresh.count(std::make_pair(vid, n))) continue;
                fresh.insert(std::make_pair(vid, n)); fresh.insert(std::make_pair(n, vid));
                ids.push_back(std::make_pair(vid, n)); pairs.push_back(idx_of[vid]); pairs.push_back(idx_of[n]);
            }
        }
        if (ids.empty()) return;
        std::vector<double> F(9 * ids.size());
        if (l3d_fundamentals(ctx, int(cams.size()), cams.data(), ids.size(), pairs.data(), F.data())) fail("l3d_fundamentals");
        for (size_t p = 0; p < ids.size(); ++p) {
            M3 Fm;
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Fm(r, c) = F[9 * p + size_t(r) * 3 + c];
            M3 Ft = Fm.transpose();
            fund[ids[p].first][ids[p].second] = Fm; fund[ids[p].second][ids[p].first] = Ft;
        }
    }

    void uploadViews()
    {
        std::vector<l3d_view> lv(views.size());
        int idx = 0;
        view_at.assign(views.size(), (ViewData*)0);
        id2idx.assign(views.empty() ? 0 : size_t(views.rbegin()->first) + 1, -1);
        pc_fwd.assign(views.size(), std::vector<PC>()); pc_rev.assign(views.size(), std::vector<PC>()); pc.assign(views.size(), std::vector<PC>());
        for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it, ++idx) {
            ViewData* v = it->second; v->index = idx; view_at[idx] = v; id2idx[it->first] = idx;
            l3d_view& o = lv[idx];
            memset(&o, 0, sizeof o);
            // a rank only needs the segments of its own connected components (every neighbour of one of its views is in them)
            const bool mine = shard_world <= 1 || my_views.count(it->first) != 0;
            o.id = int(v->id); o.n_segments = mine ? int(v->segs.size()) : 0; o.segments = reinterpret_cast<const float*>(v->segs.data());
            if (split && split_rank != 0) o.segments = 0;      // replicated from rank 0 over NCCL (l3d_set_views)
            for (int r = 0; r < 3; ++r) {
                for (int c = 0; c < 3; ++c) { o.RtKinv[r * 3 + c] = float(v->RtKinv(r, c)); o.RtKinv_d[r * 3 + c] = v->RtKinv(r, c); }   // line3D.cc:746, 790
                for (int c = 0; c < 4; ++c) o.P[r * 4 + c] = float(v->P(r, c));                                                            // line3D.cc:756
                o.C[r] = float(v->C(r)); o.C_d[r] = v->C(r);                                                                               // line3D.cc:761-763, 801-803
            }
            o.spatial_k = float(v->shifted(2.0f * sigma_p));                                                                               // line3D.cc:820
            o.k_lower = v->k_lower; o.k_upper = v->k_upper;
        }
        if (l3d_set_views(ctx, int(lv.size()), lv.data())) fail("l3d_set_views");
        for (size_t k = 0; k < lv.size(); ++k) stats.h2d_bytes += (lv[k].segments ? size_t(lv[k].n_segments) * 16 : 0) + sizeof(l3d_view);
    }

    // line3D.cc:620-648 + 698-885
    void matchViews()
    {
        std::cout << prefix << separator << std::endl;
        std::cout << prefix << ">>> MATCHING IMAGES <<<" << std::endl;
        // t_matching is the wall time of this whole function: the library runs a view ahead of the host, so the time
        // spent inside the calls alone (t_matching_calls) would leave out the host's own share of the loop
        const double t_match0 = now_s();
        if (l3d_set_stage_timing(ctx, stage_timing ? 1 : 0) || l3d_mark(ctx, 0)) fail("l3d_set_stage_timing");
        if (getenv("L3D_PROFILE")) l3d_debug_chain_times(ctx, stage_timing ? 1 : 0, 0);
        // Which neighbours a view pair-tests and to which it hands its matches depends only on the
        // neighbour graph and the view order (line3D.cc:732-736, 844-845, 875-881): work the whole
        // schedule out first and announce it, so the library can batch the pair tests.
        prepareFundamentals();
        std::map<unsigned, std::vector<l3d_neighbor> > schedule;
        {
            // done[x]: the views x has been paired with so far -- a std::set per view in the reference (matched_); here a
            // sorted vector per view id, and the neighbour lists reachable by view id, so that the 10^4..10^5 membership
            // tests of a large scene need no tree walks (the schedule is part of every step's matching time)
            const size_t n_ids = id2idx.size();
            std::vector<std::vector<unsigned> > done(n_ids);
            std::vector<const std::vector<unsigned>*> vn_of(n_ids, (const std::vector<unsigned>*)0);
            for (std::map<unsigned, std::vector<unsigned> >::iterator it = vn.begin(); it != vn.end(); ++it) if (it->first < n_ids) vn_of[it->first] = &it->second;
            auto has = [](const std::vector<unsigned>& s, unsigned x) { return std::binary_search(s.begin(), s.end(), x); };
            auto put = [](std::vector<unsigned>& s, unsigned x) { std::vector<unsigned>::iterator p = std::lower_bound(s.begin(), s.end(), x); if (p == s.end() || *p != x) s.insert(p, x); };
            auto nb_of = [&](unsigned a, unsigned b) { return a < n_ids && vn_of[a] && std::binary_search(vn_of[a]->begin(), vn_of[a]->end(), b); };      // b in VN(a)
            for (std::map<unsigned, std::vector<unsigned> >::iterator it = vn.begin(); it != vn.end(); ++it) {
                const unsigned vid = it->first;
                if (it->second.empty() || !my_views.count(vid)) continue;
                if (vid >= n_ids) fail("matching: a view with neighbours that was never added");
                std::vector<l3d_neighbor>& nb = schedule[vid];
                nb.resize(it->second.size());
                std::map<unsigned, M3>& fv = fund[vid];
                for (size_t c = 0; c < it->second.size(); ++c) {
                    const unsigned n = it->second[c];
                    std::map<unsigned, M3>::iterator f = fv.find(n);
                    if (f == fv.end() || n >= n_ids || id2idx[n] < 0) fail("fundamental matrix missing");                                  // (prepareFundamentals, :1949-1965)
                    const M3& F = f->second;
                    nb[c].view = id2idx[n];
                    for (int r = 0; r < 3; ++r) for (int cc = 0; cc < 3; ++cc) nb[c].F[r * 3 + cc] = float(F(r, cc));                    // :745
                    nb[c].to_be_matched = has(done[vid], n) ? 0 : 1;                                                                     // :732-736
                    nb[c].deliver = (nb_of(n, vid) && !has(done[n], vid)) ? 1 : 0;                                                       // :844-845
                }
                for (size_t c = 0; c < it->second.size(); ++c) {                                                                         // :875-881
                    const unsigned n = it->second[c];
                    put(done[vid], n);
                    if (nb_of(n, vid)) put(done[n], vid);
                }
            }
            std::vector<l3d_view_plan> plans;
            for (std::map<unsigned, std::vector<l3d_neighbor> >::iterator it = schedule.begin(); it != schedule.end(); ++it) {
                l3d_view_plan p; p.view = views[it->first]->index; p.n_neighbors = int(it->second.size()); p.neighbors = it->second.data();
                plans.push_back(p);
            }
            view_records.clear();
            if (l3d_retain_results(ctx, measure_only ? 0 : 1)) fail("l3d_retain_results");
            if (l3d_plan_matching(ctx, int(plans.size()), plans.data())) fail("l3d_plan_matching");
        }
        for (std::map<unsigned, std::vector<unsigned> >::iterator it = vn.begin(); it != vn.end(); ++it) {
            const unsigned vid = it->first;
            std::cout << prefix << "matching image [" << vid << "] with " << it->second.size() << " VNs\n";      // (no flush per view: the loop runs beside the GPU)
            if (it->second.empty() || !my_views.count(vid)) continue;
            ViewData* v = views[vid];
            std::vector<l3d_neighbor>& nb = schedule[vid];
            l3d_match_result res;
            const double t0 = now_s();
            int rc = measure_only ? l3d_match_view_resident(ctx, v->index, int(nb.size()), nb.data(), sigma_p, sigma_a, &res)
                                  : l3d_match_view(ctx, v->index, int(nb.size()), nb.data(), sigma_p, sigma_a, &res);
            if (rc) fail("l3d_match_view");
            stats.t_matching_calls += now_s() - t0;
            float ms[6]; l3d_last_timings(ctx, ms);
            for (int k = 0; k < 5; ++k) stats.t_stage[k] += double(ms[k]) * 1e-3;
            if (ms[0] > 0.f) { float km[3]; l3d_last_kernel_timings(ctx, km); stats.t_k_pretest += double(km[0]) * 1e-3; stats.t_k_exact += double(km[1]) * 1e-3; stats.t_k_pretest2 += double(km[2]) * 1e-3; }
            stats.score_iters_other_cam += res.score_iters_other_cam;
            stats.h2d_bytes += nb.size() * 160; stats.d2h_bytes += ((measure_only || !res.filtered) ? 0 : res.n_filtered * sizeof(l3d_match)) + 64;
            stats.pair_tests += res.pair_tests; stats.candidates += res.n_candidates; stats.raw_matches += res.n_raw;
            stats.scored_matches += res.n_scored; stats.filtered_matches += res.n_filtered; stats.score_iters += res.score_iters;
            if (verbose) {
                std::cout << prefix << "existing matches:  " << res.n_delivered << std::endl;
                std::cout << prefix << "#raw_matches:          " << res.n_scored << std::endl;
                std::cout << prefix << "#filtered_matches (2): " << res.n_filtered << std::endl;
            }
            v->median_depth = res.median_depth;                                                                                      // :835
            size_t n_tbm = 0; for (size_t c = 0; c < nb.size(); ++c) n_tbm += nb[c].to_be_matched ? 1 : 0;
            match_log.push_back(int32_t(vid)); match_log.push_back(int32_t(n_tbm)); match_log.push_back(int32_t(res.n_delivered)); match_log.push_back(int32_t(res.n_filtered));
            match_median.push_back(res.median_depth);
            // the records stay in the library's pinned memory (l3d_retain_results) and are filed after the loop, on
            // all cores, while nothing else waits for them: the GPU is not held up by the host's bookkeeping
            if (!measure_only && !match_only && (res.filtered || res.n_filtered == 0)) { ViewRecords vr = {vid, v->index, res.scored != 0, res.filtered, res.n_filtered, &it->second}; view_records.push_back(vr); }
            else if (!measure_only && !match_only) { std::vector<BestMatch>& bm = best[vid]; bm.assign(v->segs.size(), BestMatch()); for (size_t s = 0; s < bm.size(); ++s) bm[s].has = false; }      // (the later stages do not run in the measuring modes)
        }
        { float span = 0.f; if (l3d_mark(ctx, 1) || l3d_marks_elapsed(ctx, &span)) fail("l3d_mark"); stats.t_matching_device += double(span) * 1e-3; }   // CUDA events around the whole loop
        stats.t_matching += now_s() - t_match0;
        if (getenv("L3D_PROFILE")) {
            double hs[8]; l3d_debug_host_times(ctx, hs);
            if (stage_timing) {
                double cm[24]; l3d_debug_chain_times(ctx, 0, cm);
                static const char* nm[15] = {"", "gather_counts", "scan", "fill", "bucket", "score", "reduce", "filter", "push", "wait", "hist", "scan2", "scatter", "finish", "gather"};
                std::cerr << "[prof] chain ms:";
                for (int k = 1; k < 15; ++k) std::cerr << " " << nm[k] << " " << cm[k];
                std::cerr << std::endl;
            }
            std::cerr << "[prof] matching " << (now_s() - t_match0) * 1e3 << " ms: host enqueue " << hs[0] * 1e3 << " ms, waiting " << hs[1] * 1e3 << " ms, collecting " << hs[2] * 1e3 << " ms, helper " << hs[4] * 1e3 << " ms, foreground pair tests " << hs[5] * 1e3 << " ms, totals " << hs[6] * 1e3 << " ms, " << hs[3] << " groups" << std::endl;
        }
    }

    // view.cc:162-224 (a view takes in its matches: best match per segment) and the potential-correspondence
    // lists of line3D.cc:861-871, from the records the matching loop left in place.  Pass 1 (per source view):
    // best matches, the view's own list, and its records grouped by target camera.  Pass 2 (per target view):
    // the received records in ascending source-view order, as a sequential loop would have appended them.
    struct ViewRecords { unsigned vid; int index; bool scored; const l3d_match* recs; uint64_t n; const std::vector<unsigned>* cams; };
    std::vector<ViewRecords> view_records;
    void fileRecords()
    {
        const size_t nr = view_records.size();
        for (size_t k = 0; k < nr; ++k) best[view_records[k].vid];                    // create the map nodes before the threads run
        std::vector<std::vector<PC> > grouped(nr);                                    // reversed records of a source view, by local camera
        std::vector<std::vector<size_t> > group_off(nr);
        std::vector<std::vector<std::pair<size_t, size_t> > > sources(view_at.size());   // target view index -> (source record set, local camera)
        for (size_t k = 0; k < nr; ++k) {
            const ViewRecords& vr = view_records[k];
            for (size_t c = 0; c < vr.cams->size(); ++c) {
                const unsigned cam = vr.scored ? (*vr.cams)[c] : unsigned(c);         // cudawrapper.cu:877-878, line3D.cc:843 -- reproduced
                if (cam < id2idx.size() && id2idx[cam] >= 0) sources[id2idx[cam]].push_back(std::make_pair(k, c));
            }
        }
        auto pass1 = [&](size_t t, size_t nt) {
            for (size_t k = t; k < nr; k += nt) {
                const ViewRecords& vr = view_records[k];
                std::vector<BestMatch>& bm = best.find(vr.vid)->second;
                bm.assign(view_at[vr.index]->segs.size(), BestMatch());
                for (size_t s = 0; s < bm.size(); ++s) bm[s].has = false;
                const size_t nc = vr.cams->size();
                std::vector<size_t>& off = group_off[k]; off.assign(nc + 1, 0);
                std::vector<PC>& fwd = pc_fwd[vr.index]; fwd.resize(vr.n);
                for (uint64_t i = 0; i < vr.n; ++i) {
                    const l3d_match& m = vr.recs[i];
                    const unsigned cam = vr.scored ? (*vr.cams)[m.cam] : unsigned(m.cam);
                    PC f = {unsigned(m.src_seg), cam, unsigned(m.tgt_seg)};                                                         // :861-865
                    fwd[i] = f;
                    ++off[size_t(m.cam) + 1];
                    BestMatch& b = bm[m.src_seg];                                                                                   // view.cc:165-183: first maximum
                    if (!b.has || m.confidence > b.conf) { b.has = true; b.tgt_cam = cam; b.tgt_seg = unsigned(m.tgt_seg); b.conf = m.confidence; b.d1 = m.depths[0]; b.d2 = m.depths[1]; }
                }
                for (size_t c = 0; c < nc; ++c) off[c + 1] += off[c];
                std::vector<size_t> cur(off.begin(), off.end() - 1);
                std::vector<PC>& g = grouped[k]; g.resize(vr.n);
                for (uint64_t i = 0; i < vr.n; ++i) {
                    const l3d_match& m = vr.recs[i];
                    PC r = {unsigned(m.tgt_seg), vr.vid, unsigned(m.src_seg)};
                    g[cur[m.cam]++] = r;
                }
            }
        };
        auto pass2 = [&](size_t t, size_t nt) {
            for (size_t w = t; w < sources.size(); w += nt) {
                size_t total = 0;
                for (size_t q = 0; q < sources[w].size(); ++q) { const std::vector<size_t>& off = group_off[sources[w][q].first]; total += off[sources[w][q].second + 1] - off[sources[w][q].second]; }
                std::vector<PC>& rev = pc_rev[w]; rev.reserve(total);
                for (size_t q = 0; q < sources[w].size(); ++q) {
                    const size_t k = sources[w][q].first, c = sources[w][q].second;
                    rev.insert(rev.end(), grouped[k].begin() + group_off[k][c], grouped[k].begin() + group_off[k][c + 1]);
                }
            }
        };
        const size_t nt = std::max<size_t>(1, host_threads());
        for (int pass = 1; pass <= 2; ++pass) {
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) { if (pass == 1) pool.emplace_back(pass1, t, nt); else pool.emplace_back(pass2, t, nt); }
            if (pass == 1) pass1(0, nt); else pass2(0, nt);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
        }
        if (keep_debug)
            for (size_t k = 0; k < nr; ++k) {
                const ViewRecords& vr = view_records[k];
                for (uint64_t i = 0; i < vr.n; ++i) {
                    const l3d_match& m = vr.recs[i];
                    const unsigned cam = vr.scored ? (*vr.cams)[m.cam] : unsigned(m.cam);
                    const double row[9] = {double(vr.vid), double(m.src_seg), double(cam), double(m.tgt_seg), m.depths[0], m.depths[1], m.depths[2], m.depths[3], m.confidence};
                    filtered_dump.insert(filtered_dump.end(), row, row + 9);
                }
            }
        view_records.clear();
    }

    // line3D.cc:899-965 + view.cc:302-342 on the device
    void buildHypotheses()
    {
        std::cout << prefix << separator << std::endl;
        std::cout << prefix << ">>> OPTIMIZING CORRESPONDENCES (greedy) <<<" << std::endl;
        fileRecords();
        hyps.clear(); hyp_of.assign(views.size(), std::vector<int>());
        for (size_t k = 0; k < view_at.size(); ++k) hyp_of[k].assign(view_at[k]->segs.size(), -1);
        std::vector<l3d_hypothesis_in> in;
        for (std::map<unsigned, std::vector<BestMatch> >::iterator it = best.begin(); it != best.end(); ++it)
            for (size_t s = 0; s < it->second.size(); ++s) {
                const BestMatch& b = it->second[s];
                if (!b.has) continue;
                Hypothesis h; h.cam = it->first; h.seg = unsigned(s); h.tgt_cam = b.tgt_cam; h.tgt_seg = b.tgt_seg;
                h.score = fminf(b.conf, 1.0f); h.d1 = b.d1; h.d2 = b.d2;                                                            // :937
                hyp_of[views[it->first]->index][s] = int(hyps.size());
                hyps.push_back(h);
                l3d_hypothesis_in hi; hi.view = views[it->first]->index; hi.seg = int(s); hi.depth1 = b.d1; hi.depth2 = b.d2;
                in.push_back(hi);
            }
        std::vector<l3d_segment3d> out(in.size());
        if (!in.empty() && l3d_unproject(ctx, in.size(), in.data(), out.data())) fail("l3d_unproject");
        for (size_t k = 0; k < hyps.size(); ++k)
            for (int a = 0; a < 3; ++a) { hyps[k].P1(a) = out[k].P1[a]; hyps[k].P2(a) = out[k].P2[a]; hyps[k].dir(a) = out[k].dir[a]; }
        stats.hypotheses = hyps.size();
        if (verbose) std::cout << prefix << "#clusterable_segments:  " << hyps.size() << std::endl;
    }

    // line3D.cc:968-1252: candidate enumeration in the reference's order, affinities on the device
    void clusterSegments(bool perform_diffusion)
    {
        std::cout << prefix << separator << std::endl;
        std::cout << prefix << ">>> CLUSTERING 2D SEGMENTS (global) <<<" << std::endl;
        result.clear();
        std::cout << prefix << "computing affinity matrix..." << std::endl;
        const double t0 = now_s();
        const bool prof = getenv("L3D_PROFILE") != 0; double tp = now_s();
        // potential correspondences per view: the records a view produced are already in (seg, cam, seg) order,
        // the ones it received are sorted here; both are merged and de-duplicated (a std::map in the reference)
        {
            const size_t nv = view_at.size();
            auto work = [&](size_t v0, size_t step) {
                auto less = [](const PC& a, const PC& b) { return a.seg != b.seg ? a.seg < b.seg : (a.tcam != b.tcam ? a.tcam < b.tcam : a.tseg < b.tseg); };
                auto same = [](const PC& a, const PC& b) { return a.seg == b.seg && a.tcam == b.tcam && a.tseg == b.tseg; };
                for (size_t v = v0; v < nv; v += step) {
                    std::vector<PC>& f = pc_fwd[v]; std::vector<PC>& r = pc_rev[v]; std::vector<PC>& o = pc[v];
                    if (!std::is_sorted(f.begin(), f.end(), less)) std::sort(f.begin(), f.end(), less);
                    std::sort(r.begin(), r.end(), less);
                    o.resize(f.size() + r.size());
                    std::merge(f.begin(), f.end(), r.begin(), r.end(), o.begin(), less);
                    o.erase(std::unique(o.begin(), o.end(), same), o.end());
                    std::vector<PC>().swap(f); std::vector<PC>().swap(r);
                }
            };
            const size_t nt = std::max<size_t>(1, host_threads());
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t, nt);
            work(0, nt);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
        }
        if (prof) { std::cerr << "[prof] pc merge " << now_s() - tp << std::endl; tp = now_s(); }
        typedef l3d_candidate Cand;                   // { a, b, mult, thr }
        Cand* cands = 0; size_t n_cands = 0;          // the views' lists end to end, in the library's pinned memory
        // dense numbering of every segment, for the "used" pair set
        std::vector<uint64_t> view_base(view_at.size() + 1, 0);
        for (size_t k = 0; k < view_at.size(); ++k) view_base[k + 1] = view_base[k] + view_at[k]->segs.size();
        const uint64_t total = view_base[view_at.size()];
        // a key outside every view can only come from the unscored-view quirk; such a target has no
        // hypothesis and no collinear partners, so it can never produce an edge: it is skipped
        auto dense = [&](unsigned cam, unsigned seg, int& vidx, uint64_t& out) -> bool {
            if (cam >= id2idx.size() || id2idx[cam] < 0) return false;
            vidx = id2idx[cam];
            if (seg >= view_at[vidx]->segs.size()) return false;
            out = view_base[vidx] + seg; return true;
        };
        // Set of unordered segment pairs already looked at (line3D.cc:980, 1001-1005): a pair is dealt with at its first
        // encounter, whether or not it passes.  Pairs only ever join a segment of the view whose turn it is (the source)
        // with a segment of a neighbouring view or of the same view, so the set is one bit matrix per view pair, and the
        // matrix of views (lo, hi) is written during lo's turns and again during hi's turns, by nobody else.  That makes
        // the views' turns independent but for one thing -- what lo marked before hi starts:
        //   phase 1  every view v, in parallel: its turns restricted to targets in LATER views w; the pairs it marks
        //            there (on matrices that start empty) are kept as a list T(v, w);
        //   phase 2  every view v, in parallel: all its turns in order with private matrices -- those of earlier views
        //            w start from T(w, v), the others empty -- emitting the candidates.
        // The per-view candidate lists, concatenated in view order, are the sequential loop's list.
        const int nviews = int(view_at.size());
        std::vector<size_t> h_begin(size_t(nviews) + 1, 0);
        {   // hypotheses are ordered by (camera id, segment): one contiguous range per view, in view order
            std::vector<size_t> cnt(size_t(nviews), 0), first(size_t(nviews), hyps.size());
            for (size_t hidx = 0; hidx < hyps.size(); ++hidx) {
                int svi; uint64_t d; if (!dense(hyps[hidx].cam, hyps[hidx].seg, svi, d)) continue;
                if (cnt[svi]++ == 0) first[svi] = hidx;
            }
            for (int v = 0; v < nviews; ++v) h_begin[v] = first[v];
            h_begin[nviews] = hyps.size();
            for (int v = nviews - 1; v >= 0; --v) if (cnt[v] == 0) h_begin[v] = h_begin[v + 1];
        }
        std::vector<std::map<int, std::vector<uint32_t> > > touched(static_cast<size_t>(nviews));      // T(v, w), keys (seg of v) << 16 | seg of w
        std::vector<std::vector<Cand> > cands_of(static_cast<size_t>(nviews));
        auto run_pass = [&](int v, bool emit, std::vector<std::vector<uint64_t> >& spare) {      // spare: this thread's retired matrices (no fresh pages per pass)
            std::vector<std::vector<uint64_t> > bm(static_cast<size_t>(nviews)); std::vector<size_t> pitch(size_t(nviews), 0);
            auto mark = [&](unsigned sseg, int tv, unsigned tseg) -> bool {   // true if the pair was new
                const int lo = std::min(v, tv), hi = std::max(v, tv);
                std::vector<uint64_t>& w = bm[tv];
                if (w.empty()) {
                    if (!spare.empty()) { w.swap(spare.back()); spare.pop_back(); }
                    pitch[tv] = (view_at[hi]->segs.size() + 63) / 64;
                    w.assign(std::max<size_t>(view_at[lo]->segs.size() * pitch[tv], 1), 0);
                    if (emit && tv < v) {                             // what the earlier view marked during its own turns
                        std::map<int, std::vector<uint32_t> >::const_iterator f = touched[tv].find(v);
                        if (f != touched[tv].end())
                            for (size_t k = 0; k < f->second.size(); ++k) { const uint32_t key = f->second[k]; w[size_t(key >> 16) * pitch[tv] + ((key & 0xffffu) >> 6)] |= uint64_t(1) << (key & 63u); }
                    }
                }
                unsigned r, c;                                          // row: segment of the lower view, column: of the higher
                if (v < tv) { r = sseg; c = tseg; } else if (v > tv) { r = tseg; c = sseg; } else { r = std::min(sseg, tseg); c = std::max(sseg, tseg); }
                uint64_t& word = w[size_t(r) * pitch[tv] + (c >> 6)];
                const uint64_t bit = uint64_t(1) << (c & 63);
                if (word & bit) return false;
                word |= bit;
                if (!emit) touched[v][tv].push_back((uint32_t(sseg) << 16) | uint32_t(tseg));
                return true;
            };
            std::vector<Cand>& out = cands_of[v];
            const std::vector<PC>& list = pc[v];
            size_t cur = 0;
            for (size_t hidx = h_begin[v]; hidx < h_begin[v + 1]; ++hidx) {
                const Hypothesis& C = hyps[hidx];
                int svi; uint64_t dsrc; if (!dense(C.cam, C.seg, svi, dsrc) || svi != v) continue;
                while (cur < list.size() && list[cur].seg < C.seg) ++cur;
                for (size_t k = cur; k < list.size() && list[k].seg == C.seg; ++k) {                                                // :996-1138
                    const unsigned tcam = list[k].tcam, tseg = list[k].tseg;
                    int tvi; uint64_t dt; if (!dense(tcam, tseg, tvi, dt)) continue;
                    if (!emit && tvi <= v) continue;
                    if (!mark(C.seg, tvi, tseg)) continue;
                    const int ht = hyp_of[tvi][tseg];
                    if (ht < 0) continue;                                                                                           // :1007
                    if (emit) { Cand c = {int(hidx), ht, 1.0f, L3D_MIN_AFFINITY}; out.push_back(c); }                               // :1014-1016
                    const std::vector<CollinEntry>& cl = view_at[tvi]->collin[tseg];                                                // :1065-1135
                    for (size_t e = 0; e < cl.size(); ++e) {
                        if (!mark(C.seg, tvi, cl[e].other)) continue;
                        const int hc = hyp_of[tvi][cl[e].other];
                        if (hc < 0) continue;
                        if (emit) { Cand c3 = {int(hidx), hc, 1.0f, 0.01f}; out.push_back(c3); }                                    // :1085-1087
                    }
                }
                if (!emit) continue;
                const std::vector<CollinEntry>& cs = view_at[v]->collin[C.seg];                                                     // :1141-1214
                for (size_t e = 0; e < cs.size(); ++e) {
                    if (!mark(C.seg, v, cs[e].other)) continue;
                    const int hc = hyp_of[v][cs[e].other];
                    if (hc < 0) continue;
                    Cand c2 = {int(hidx), hc, cs[e].w, 0.01f}; out.push_back(c2);                                                   // :1163-1165
                }
            }
            for (int t = 0; t < nviews; ++t) if (!bm[t].empty()) { spare.push_back(std::vector<uint64_t>()); spare.back().swap(bm[t]); }
        };
        for (int phase = 1; phase <= 2; ++phase) {
            const size_t nt = std::max<size_t>(1, std::min<size_t>(host_threads(), size_t(nviews)));
            auto work = [&](size_t t0) { std::vector<std::vector<uint64_t> > spare; for (int v = int(t0); v < nviews; v += int(nt)) run_pass(v, phase == 2, spare); };
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t);
            work(0);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
        }
        {   // the views' lists, end to end (copied in parallel: 10^7..10^8 records)
            std::vector<size_t> at(size_t(nviews) + 1, 0);
            for (int v = 0; v < nviews; ++v) at[v + 1] = at[v] + cands_of[v].size();
            n_cands = at[nviews];
            if (l3d_candidate_buffer(ctx, n_cands, &cands)) fail("l3d_candidate_buffer");
            const size_t nt = std::max<size_t>(1, std::min<size_t>(host_threads(), size_t(nviews)));
            auto work = [&](size_t t0) {
                for (int v = int(t0); v < nviews; v += int(nt)) { if (!cands_of[v].empty()) memcpy(&cands[at[v]], cands_of[v].data(), sizeof(Cand) * cands_of[v].size()); std::vector<Cand>().swap(cands_of[v]); }
            };
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t);
            work(0);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
        }
        touched.clear();
        if (prof) { std::cerr << "[prof] enumerate " << now_s() - tp << " cands " << n_cands << std::endl; tp = now_s(); }
        // similarities, edge weights and thresholds on the device (line3D.cc:1600-1691, 1014-1016, 1085-1087, 1163-1165):
        // only the accepted candidates come back
        std::vector<float> medians(views.size()), scores(hyps.size());
        for (std::map<unsigned, ViewData*>::iterator it = views.begin(); it != views.end(); ++it) medians[it->second->index] = it->second->median_depth;
        for (size_t k = 0; k < hyps.size(); ++k) scores[k] = hyps[k].score;
        const l3d_kept_edge* kept = 0; size_t n_kept = 0; const uint32_t* kept_order = 0;
        if (n_cands && l3d_affinity_edges(ctx, medians.data(), scores.data(), n_cands, cands, sigma_a, &kept, &n_kept, perform_diffusion ? (const uint32_t**)0 : &kept_order)) fail("l3d_affinity_edges");
        stats.affinity_pairs = n_cands;
        if (prof) { std::cerr << "[prof] l3d_affinity_edges " << now_s() - tp << " kept " << n_kept << std::endl; tp = now_s(); }
        // edges and node ids in encounter order (line3D.cc:1018-1060)
        std::vector<Edge> A;
        std::vector<int> node_of(hyps.size(), -1); std::vector<int> hyp_of_node;
        A.reserve(2 * n_kept);
        for (size_t q = 0; q < n_kept; ++q) {
            const Cand& c = cands[kept[q].index]; const float w = kept[q].w;
            if (node_of[c.a] < 0) { node_of[c.a] = int(hyp_of_node.size()); hyp_of_node.push_back(c.a); }
            if (node_of[c.b] < 0) { node_of[c.b] = int(hyp_of_node.size()); hyp_of_node.push_back(c.b); }
            Edge e = {node_of[c.a], node_of[c.b], w}; A.push_back(e);
            Edge r = {node_of[c.b], node_of[c.a], w}; A.push_back(r);
        }
        if (prof) { std::cerr << "[prof] edges " << now_s() - tp << " A " << A.size() << std::endl; }
        stats.t_affinity += now_s() - t0;
        const int n_nodes = int(hyp_of_node.size());
        if (verbose) {
            std::cout << prefix << "A: #num_entries = " << A.size() << std::endl;
            std::cout << prefix << "A: #num_rows    = " << n_nodes << std::endl;
        }
        if (A.empty()) return;                                                                                                      // :1232-1233
        if (perform_diffusion) {                                                                                                    // :1255-1303
            std::cout << prefix << "replicator dynamics diffusion..." << std::endl;
            const double td = now_s();
            std::vector<l3d_edge> in(A.size()), out(A.size());
            for (size_t k = 0; k < A.size(); ++k) { in[k].i = A[k].i; in[k].j = A[k].j; in[k].w = A[k].w; }
            if (keep_debug) {   // W as the reference stores it (column-sorted) for the dump
                std::vector<l3d_edge> w(in);
                std::stable_sort(w.begin(), w.end(), [](const l3d_edge& a, const l3d_edge& b) { return a.j < b.j || (a.j == b.j && a.i < b.i); });
                for (size_t k = 0; k < w.size(); ++k) { edges_pre.push_back(w[k].i); edges_pre.push_back(w[k].j); edges_pre.push_back(w[k].w); }
            }
            if (l3d_diffuse(ctx, n_nodes, in.size(), in.data(), out.data())) fail("l3d_diffuse");
            { double ds[3]; if (!l3d_diffuse_stats(ctx, ds)) { stats.rdd_step_ms = ds[0]; stats.rdd_step_bytes = ds[1]; stats.rdd_entries = ds[2]; } }
            // symmetrise with min(w_ij, w_ji), then (i, j) order (line3D.cc:1267-1299).  The reference walks the result in
            // stored order (sorted by row, then column) and keeps, per unordered pair, the smaller of the two directions
            // in a map of maps, which it then reads out in (i, j) order -- the stored order again.  Here: every entry
            // looks its transposed entry up by bisection in that entry's row.
            {
                const size_t n = out.size();
                std::vector<size_t> row_begin(size_t(n_nodes) + 1, 0);
                for (size_t k = 0; k < n; ++k) ++row_begin[size_t(out[k].i) + 1];
                for (int r = 0; r < n_nodes; ++r) row_begin[size_t(r) + 1] += row_begin[r];
                A.resize(n);
                auto work = [&](size_t t0, size_t nt) {
                    for (size_t k = n * t0 / nt; k < n * (t0 + 1) / nt; ++k) {
                        const int s1 = out[k].i, s2 = out[k].j;
                        size_t lo = row_begin[s2], hi = row_begin[size_t(s2) + 1];
                        while (lo < hi) { const size_t mid = (lo + hi) >> 1; if (out[mid].j < s1) lo = mid + 1; else hi = mid; }
                        float w = out[k].w;
                        if (lo < row_begin[size_t(s2) + 1] && out[lo].j == s1) w = fminf(w, out[lo].w);
                        Edge e = {s1, s2, w}; A[k] = e;
                    }
                };
                const size_t nt = std::max<size_t>(1, host_threads());
                std::vector<std::thread> pool;
                for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t, nt);
                work(0, nt);
                for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
            }
            stats.t_diffusion += now_s() - td;
        }
        stats.edges = A.size();
        cluster_nodes = n_nodes;
        if (keep_debug) for (size_t k = 0; k < A.size(); ++k) { edges_cluster.push_back(A[k].i); edges_cluster.push_back(A[k].j); edges_cluster.push_back(A[k].w); }
        std::cout << prefix << "graph clustering..." << std::endl;
        const double tc = now_s();
        // Felzenszwalb-Huttenlocher as the reference runs it (clustering.cc:6-47, universe.h:59-115):
        // edges stably sorted by ascending weight, threshold w + c / size with c = 1
        std::vector<Edge> sorted;
        if (kept_order && A.size() == 2 * n_kept) {
            // sorted on the device (l3d_affinity_edges): the two directions of a candidate carry the same weight and
            // follow each other in A, so they stay together, in order, in the stable order by weight
            sorted.resize(A.size());
            for (size_t t = 0; t < n_kept; ++t) { sorted[2 * t] = A[2 * size_t(kept_order[t])]; sorted[2 * t + 1] = A[2 * size_t(kept_order[t]) + 1]; }
        } else {
            sorted = A;
            // stable sort by weight: positive floats order like their bit patterns, so four stable 8-bit counting passes
            // do it (a comparison sort of 10^6..10^7 edges was most of this stage); anything else goes the general way
            bool positive = true;
            for (size_t k = 0; k < sorted.size() && positive; ++k) positive = sorted[k].w > 0.0f;
            if (positive && sorted.size() > 256) {
                std::vector<Edge> tmp(sorted.size());
                for (int shift = 0; shift < 32; shift += 8) {
                    size_t count[257] = {0};
                    for (size_t k = 0; k < sorted.size(); ++k) { uint32_t b; memcpy(&b, &sorted[k].w, 4); ++count[((b >> shift) & 255u) + 1]; }
                    for (int d = 0; d < 256; ++d) count[d + 1] += count[d];
                    for (size_t k = 0; k < sorted.size(); ++k) { uint32_t b; memcpy(&b, &sorted[k].w, 4); tmp[count[(b >> shift) & 255u]++] = sorted[k]; }
                    sorted.swap(tmp);
                }
            } else std::stable_sort(sorted.begin(), sorted.end(), [](const Edge& a, const Edge& b) { return a.w < b.w; });
        }
        std::vector<int> parent(n_nodes), rank(n_nodes, 0), sz(n_nodes, 1);
        std::vector<float> thr(n_nodes, 1.0f);
        for (int i = 0; i < n_nodes; ++i) parent[i] = i;
        auto find = [&](int x) { int y = x; while (y != parent[y]) y = parent[y]; parent[x] = y; return y; };
        for (size_t k = 0; k < sorted.size(); ++k) {
            int a = find(sorted[k].i), b = find(sorted[k].j);
            if (a == b) continue;
            if (sorted[k].w <= thr[a] && sorted[k].w <= thr[b]) {
                if (rank[a] > rank[b]) { parent[b] = a; sz[a] += sz[b]; }
                else { parent[a] = b; sz[b] += sz[a]; if (rank[a] == rank[b]) rank[b]++; }
                a = find(a);
                thr[a] = sorted[k].w + 1.0f / float(sz[a]);
            }
        }
        stats.t_clustering += now_s() - tc;
        // line3D.cc:1306-1368
        const double tl = now_s();
        // the clusters as ranges of one array, in ascending order of their root node (the reference's std::map order),
        // members in node order; no per-node map or set operations
        std::vector<int> root(static_cast<size_t>(n_nodes));
        for (int n = 0; n < n_nodes; ++n) root[n] = find(n);
        std::vector<size_t> cbeg(size_t(n_nodes) + 1, 0);
        for (int n = 0; n < n_nodes; ++n) ++cbeg[size_t(root[n]) + 1];
        for (int r = 0; r < n_nodes; ++r) cbeg[size_t(r) + 1] += cbeg[r];
        std::vector<int> members(static_cast<size_t>(n_nodes));
        {
            std::vector<size_t> cur(cbeg.begin(), cbeg.end() - 1);
            for (int n = 0; n < n_nodes; ++n) members[cur[root[n]]++] = hyp_of_node[n];
        }
        std::vector<int> todo;                                   // roots of the non-empty clusters
        for (int r = 0; r < n_nodes; ++r) if (cbeg[size_t(r) + 1] > cbeg[r]) todo.push_back(r);
        if (verbose) std::cout << prefix << "#clusters_total:  " << todo.size() << std::endl;
        // Only clusters seen by at least four cameras are fitted (line3D.cc:1334).  The host puts each one's hypotheses in
        // (camera, segment) order (the reference's key order; equal keys cannot occur), on all cores; the fits themselves
        // run on the device, all clusters in one call (l3d_fit_lines, on the hypotheses l3d_unproject left there).
        unsigned valid = 0;
        {
            std::vector<std::vector<int> > sorted(todo.size());
            auto work = [&](size_t t0, size_t step) {
                std::vector<unsigned> seen;
                for (size_t k = t0; k < todo.size(); k += step) {
                    const int* mem = members.data() + cbeg[todo[k]]; const size_t m = cbeg[size_t(todo[k]) + 1] - cbeg[todo[k]];
                    if (m < 4) continue;
                    seen.clear();
                    for (size_t q = 0; q < m; ++q) seen.push_back(hyps[mem[q]].cam);
                    std::sort(seen.begin(), seen.end());
                    if (std::unique(seen.begin(), seen.end()) - seen.begin() < 4) continue;
                    sorted[k].assign(mem, mem + m);
                    std::sort(sorted[k].begin(), sorted[k].end(), [&](int a, int b) { return seg_key(hyps[a].cam, hyps[a].seg) < seg_key(hyps[b].cam, hyps[b].seg); });
                }
            };
            const size_t nt = std::max<size_t>(1, std::min<size_t>(host_threads(), todo.size()));
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t, nt);
            work(0, nt);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
            std::vector<size_t> fitted;                         // indices into todo
            std::vector<int32_t> start(1, 0), member; std::vector<uint32_t> member_cam;
            for (size_t k = 0; k < todo.size(); ++k) {
                if (sorted[k].empty()) continue;
                fitted.push_back(k);
                for (size_t q = 0; q < sorted[k].size(); ++q) { member.push_back(sorted[k][q]); member_cam.push_back(hyps[sorted[k][q]].cam); }
                start.push_back(int32_t(member.size()));
            }
            std::vector<int32_t> count(fitted.size(), 0); std::vector<double> segs(6 * std::max<size_t>(member.size(), 1));
            l3d_similarity back;
            back.scale_inv = tr_scale_inv;
            for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) back.Rinv[r * 3 + c] = tr_Rinv(r, c); back.tneg[r] = tr_tneg(r); }
            if (!fitted.empty() && l3d_fit_lines(ctx, &back, int(fitted.size()), start.data(), member.data(), member_cam.data(), count.data(), segs.data())) fail("l3d_fit_lines");
            for (size_t f = 0; f < fitted.size(); ++f) {
                if (count[f] <= 0) continue;
                std::list<std::pair<V3, V3> > s3; std::list<L3DSegment2D> s2;
                const double* sg = segs.data() + 6 * size_t(start[f]);
                for (int e = 0; e < count[f]; ++e) s3.push_back(std::make_pair(V3(sg[6 * e], sg[6 * e + 1], sg[6 * e + 2]), V3(sg[6 * e + 3], sg[6 * e + 4], sg[6 * e + 5])));
                const std::vector<int>& ord = sorted[fitted[f]];
                for (size_t q = 0; q < ord.size(); ++q) s2.push_back(L3DSegment2D(hyps[ord[q]].cam, hyps[ord[q]].seg));
                result.push_back(L3DFinalLine3D(s2, s3)); ++valid;
            }
        }
        if (verbose) std::cout << prefix << "#clusters_valid:  " << valid << std::endl;
        stats.t_lines += now_s() - tl;
        std::cout << prefix << result.size() << " 3D lines found!" << std::endl;
    }
};

// ------------------------------------------------------------------------------------------
// public API
// ------------------------------------------------------------------------------------------
Line3D::Line3D(const std::string data_directory, const int matchingNeighbors, const float uncertainty_t_upper_2D,
               const float uncertainty_t_lower_2D, const float sigma_p, const float sigma_a, const float min_baseline,
               bool useCollinearity, bool verbose)
{
    impl_ = new Impl();
    Impl& d = *impl_;
    d.verbose = verbose; d.prefix = "[L3D] "; d.separator = "----------------------------------------";
    d.data_directory = data_directory; d.matching_neighbors = matchingNeighbors;
    d.unc_upper = fabs(uncertainty_t_upper_2D); d.unc_lower = fabs(uncertainty_t_lower_2D);              // line3D.cc:18-28
    d.computation = false; d.use_collinearity = useCollinearity; d.min_baseline = min_baseline; d.measure_only = false; d.match_only = false; d.flush_l2 = false; d.keep_debug = false; d.stage_timing = false;
    if (d.unc_lower < 1.0f) d.unc_lower = 1.0f;
    if (d.unc_upper <= d.unc_lower) d.unc_upper = d.unc_lower + 1.0f;
    d.sigma_a = sigma_a; d.sigma_p = sigma_p;
    boost::filesystem::path dir(data_directory);
    boost::filesystem::create_directory(dir);
    d.tr_scale = 1.0; d.tr_scale_inv = 1.0; d.tr_R = M3::Identity(); d.tr_Rinv = M3::Identity(); d.tr_t = V3(0.0, 0.0, 0.0); d.tr_tneg = V3(0.0, 0.0, 0.0);
    int device = 0;
    if (const char* e = getenv("L3D_DEVICE")) device = atoi(e);
    if (l3d_create(device, &d.ctx)) {
        std::string msg = std::string("Line3D: cannot open the CUDA device: ") + l3d_last_error(0);
        std::cerr << d.prefix << msg << std::endl;
        delete impl_; impl_ = 0;
        throw std::runtime_error(msg);
    }
    std::cout << d.prefix << "Line3D - http://www.icg.tugraz.at/ - AerialVisionGroup" << std::endl;
    std::cout << d.prefix << "(c) 2015, Manuel Hofer" << std::endl;
    std::cout << d.prefix << "(B200 matching/scoring path: line3d-b200)" << std::endl;
    std::cout << d.prefix << d.separator << std::endl;
}

Line3D::~Line3D()
{
    if (!impl_) return;
    std::cout << impl_->prefix << impl_->separator << std::endl;
    std::cout << impl_->prefix << ">>> EXITING SYSTEM <<<" << std::endl;
    std::cout << impl_->prefix << "cleaning up..." << std::endl;
    reset();
    l3d_destroy(impl_->ctx);
    delete impl_;
}

void Line3D::reset()                                                             // line3D.cc:62-92
{
    Impl& d = *impl_;
    d.num_wps.clear(); d.common_wps.clear(); d.view_sim.clear(); d.wp2views.clear(); d.vn.clear(); d.fund.clear();
    d.best.clear(); d.pc_fwd.clear(); d.pc_rev.clear(); d.pc.clear(); d.hyps.clear(); d.hyp_of.clear(); d.result.clear();
    for (std::map<unsigned, ViewData*>::iterator it = d.views.begin(); it != d.views.end(); ++it) delete it->second;
    d.views.clear();
    d.computation = false;
    if (d.ctx) l3d_reset_matching(d.ctx);
}

void Line3D::addImage(const unsigned int imageID, const cv::Mat image, const Eigen::Matrix3d K, const Eigen::Matrix3d R,
                      const Eigen::Vector3d t, std::list<unsigned int>& worldpointIDs, const int maxImgWidth, const bool loadAndStoreSegments)
{
    Impl& d = *impl_;
    if (!d.guards(imageID, image, worldpointIDs.size() != 0, "no worldpoints")) return;
    if (!d.addView(imageID, image, K, R, t, maxImgWidth, loadAndStoreSegments)) return;
    if (d.verbose) std::cout << d.prefix << "updating view neighborhoods using " << worldpointIDs.size() << " wps" << std::endl;
    d.processWorldpoints(imageID, worldpointIDs);                                 // line3D.cc:216
}

void Line3D::addImage_fixed_sim(const unsigned int imageID, const cv::Mat image, const Eigen::Matrix3d K, const Eigen::Matrix3d R,
                                const Eigen::Vector3d t, std::map<unsigned int, float>& viewSimilarity, const int maxImgWidth,
                                const bool loadAndStoreSegments)
{
    Impl& d = *impl_;
    if (!d.guards(imageID, image, viewSimilarity.size() != 0, "no view similarities")) return;
    if (!d.addView(imageID, image, K, R, t, maxImgWidth, loadAndStoreSegments)) return;
    for (std::map<unsigned, float>::iterator it = viewSimilarity.begin(); it != viewSimilarity.end(); ++it)   // line3D.cc:1938-1946
        if (it->second > 0.01f) d.view_sim[imageID][it->first] = it->second;
}

void Line3D::compute3Dmodel(bool perform_diffusion)                              // line3D.cc:345-374
{
    Impl& d = *impl_;
    if (d.views.size() < 4) { std::cerr << d.prefix << "not enough images! can't compute 3D model..." << std::endl; return; }
    d.computation = true;
    d.best.clear(); d.result.clear();
    d.match_log.clear(); d.match_median.clear(); d.filtered_dump.clear(); d.edges_pre.clear(); d.edges_cluster.clear(); d.cluster_nodes = 0;
    memset(&d.stats, 0, sizeof d.stats);
    const uint64_t launches0 = l3d_kernel_launches(d.ctx);
    const double t0 = now_s();
    d.findVisualNeighbors();
    d.assignShard();
    d.transformGeometry();
    d.stats.t_neighbors = now_s() - t0;
    double t1 = now_s();
    d.uploadViews();
    d.stats.t_upload = now_s() - t1;
    if (d.flush_l2) l3d_flush_l2(d.ctx);
    d.matchViews();
    // (several GPUs on one scene: the records are gathered on rank 0, which alone goes on to cluster them)
    if (!d.measure_only && !d.match_only && (!d.split || d.split_rank == 0)) {
        t1 = now_s();
        d.buildHypotheses();
        d.stats.t_hypotheses = now_s() - t1;
        d.clusterSegments(perform_diffusion);
    }
    d.stats.t_total = now_s() - t0;
    d.stats.kernel_launches = l3d_kernel_launches(d.ctx) - launches0;
}

void Line3D::getResult(std::list<L3D::L3DFinalLine3D>& result) { result.clear(); result = impl_->result; }   // line3D.cc:377-381
unsigned int Line3D::numCameras() { return impl_->views.size(); }
L3DRunStats Line3D::runStats() const { return impl_->stats; }
void Line3D::setMeasureOnly(bool on) { impl_->measure_only = on; }
void Line3D::setMatchOnly(bool on) { impl_->match_only = on; }
void Line3D::setFlushL2(bool on) { impl_->flush_l2 = on; }
void Line3D::setStageTiming(bool on) { impl_->stage_timing = on; }
void Line3D::setKeepDebug(bool on) { impl_->keep_debug = on; }
void Line3D::resetGeometry()
{
    for (std::map<unsigned, ViewData*>::iterator it = impl_->views.begin(); it != impl_->views.end(); ++it) { it->second->R = it->second->R0; it->second->t = it->second->t0; it->second->init(); }
}
void Line3D::setSplit(int rank, int world, const unsigned char* nccl_id)
{
    Impl& d = *impl_;
    if (l3d_comm_init(d.ctx, rank, world, nccl_id, 0)) d.fail("l3d_comm_init");
    d.split = world > 1; d.split_rank = d.split ? rank : 0; d.split_world = d.split ? world : 1;
    d.shard_rank = 0; d.shard_world = 1;
}
void Line3D::setShard(int rank, int world) { impl_->shard_world = world < 1 ? 1 : world; impl_->shard_rank = (rank < 0 || rank >= impl_->shard_world) ? 0 : rank; }

float4 Line3D::getSegment2D(L3D::L3DSegment2D& seg2D)                            // line3D.cc:2004-2013, view.cc:275-288
{
    Impl& d = *impl_;
    float4 zero; zero.x = zero.y = zero.z = zero.w = 0.0f;
    std::map<unsigned, ViewData*>::iterator it = d.views.find(seg2D.camID());
    if (it == d.views.end()) { std::cerr << d.prefix << "no view with ID " << seg2D.camID() << "!" << std::endl; return zero; }
    if (seg2D.segID() < it->second->segs.size()) return it->second->segs[seg2D.segID()];
    return zero;
}

void Line3D::save3DLinesAsSTL(std::list<L3D::L3DFinalLine3D>& result, std::string filename)   // line3D.cc:384-430
{
    std::ofstream file; file.open(filename.c_str());
    file << "solid lineModel" << std::endl;
    for (std::list<L3DFinalLine3D>::iterator it = result.begin(); it != result.end(); ++it) {
        std::list<std::pair<V3, V3> >* s3 = it->segments3D();
        for (std::list<std::pair<V3, V3> >::iterator s = s3->begin(); s != s3->end(); ++s) {
            char a[6][50];
            sprintf(a[0], "%e", s->first.x()); sprintf(a[1], "%e", s->first.y()); sprintf(a[2], "%e", s->first.z());
            sprintf(a[3], "%e", s->second.x()); sprintf(a[4], "%e", s->second.y()); sprintf(a[5], "%e", s->second.z());
            file << " facet normal 1.0e+000 0.0e+000 0.0e+000" << std::endl;
            file << "  outer loop" << std::endl;
            file << "   vertex " << a[0] << " " << a[1] << " " << a[2] << std::endl;
            file << "   vertex " << a[3] << " " << a[4] << " " << a[5] << std::endl;
            file << "   vertex " << a[0] << " " << a[1] << " " << a[2] << std::endl;
            file << "  endloop" << std::endl;
            file << " endfacet" << std::endl;
        }
    }
    file << "endsolid lineModel" << std::endl;
    file.close();
}

void Line3D::save3DLinesAsTXT(std::list<L3D::L3DFinalLine3D>& result, std::string filename)   // line3D.cc:433-473
{
    std::ofstream file; file.open(filename.c_str());
    for (std::list<L3DFinalLine3D>::iterator it = result.begin(); it != result.end(); ++it) {
        std::list<std::pair<V3, V3> >* s3 = it->segments3D();
        if (s3->size() == 0) continue;
        file << s3->size() << " ";
        for (std::list<std::pair<V3, V3> >::iterator s = s3->begin(); s != s3->end(); ++s) {
            file << s->first.x() << " " << s->first.y() << " " << s->first.z() << " ";
            file << s->second.x() << " " << s->second.y() << " " << s->second.z() << " ";
        }
        std::list<L3DSegment2D>* s2 = it->segments2D();
        file << s2->size() << " ";
        for (std::list<L3DSegment2D>::iterator q = s2->begin(); q != s2->end(); ++q) {
            file << q->camID() << " " << q->segID() << " ";
            float4 c = getSegment2D(*q);
            file << c.x << " " << c.y << " " << c.z << " " << c.w << " ";
        }
        file << std::endl;
    }
    file.close();
}

bool Line3D::writeSegmentCache(const std::string& data_directory, unsigned int imageID, unsigned int width, unsigned int height,
                               bool collinearity_flag, const std::vector<float4>& segments)
{
    std::ofstream f(cacheName(data_directory, imageID, width, height, collinearity_flag).c_str(), std::ios::binary);
    if (!f) return false;
    f.write("L3DSEG01", 8);
    uint32_t n = uint32_t(segments.size()), flags = 0;
    f.write(reinterpret_cast<const char*>(&n), 4); f.write(reinterpret_cast<const char*>(&flags), 4);
    if (n) f.write(reinterpret_cast<const char*>(segments.data()), std::streamsize(size_t(n) * sizeof(float4)));
    return bool(f);
}

void Line3D::debugDump(const std::string& directory)
{
    Impl& d = *impl_;
    const std::string p = directory + "/";
    std::vector<int32_t> vn;
    for (std::map<unsigned, std::vector<unsigned> >::iterator a = d.vn.begin(); a != d.vn.end(); ++a)
        for (size_t k = 0; k < a->second.size(); ++k) { vn.push_back(int32_t(a->first)); vn.push_back(int32_t(a->second[k])); }
    npy_i32(p + "visual_neighbors.npy", vn, 2);
    std::vector<double> vw;
    for (std::map<unsigned, ViewData*>::iterator a = d.views.begin(); a != d.views.end(); ++a) {
        ViewData* v = a->second;
        vw.push_back(a->first); vw.push_back(v->median_depth); vw.push_back(v->k_upper); vw.push_back(v->k_lower); vw.push_back(float(v->shifted(2.0f * d.sigma_p)));
        for (int i = 0; i < 3; ++i) vw.push_back(v->C(i));
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(v->RtKinv(r, c));
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) vw.push_back(v->P(r, c));
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(v->K(r, c));
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) vw.push_back(v->R(r, c));
        for (int i = 0; i < 3; ++i) vw.push_back(v->t(i));
    }
    npy_f64(p + "views.npy", vw, 50);
    std::vector<double> fs;
    for (std::map<unsigned, std::map<unsigned, M3> >::iterator a = d.fund.begin(); a != d.fund.end(); ++a)
        for (std::map<unsigned, M3>::iterator b = a->second.begin(); b != a->second.end(); ++b) {
            fs.push_back(a->first); fs.push_back(b->first);
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) fs.push_back(b->second(r, c));
        }
    npy_f64(p + "fundamentals.npy", fs, 11);
    std::vector<double> co;
    for (std::map<unsigned, ViewData*>::iterator a = d.views.begin(); a != d.views.end(); ++a)
        for (size_t s = 0; s < a->second->collin.size(); ++s)
            for (size_t k = 0; k < a->second->collin[s].size(); ++k) { co.push_back(a->first); co.push_back(double(s)); co.push_back(a->second->collin[s][k].other); co.push_back(a->second->collin[s][k].w); }
    npy_f64(p + "collinearity.npy", co, 4);
    npy_i32(p + "match_log.npy", d.match_log, 4);
    npy_f32(p + "match_median.npy", d.match_median, 1);
    npy_f64(p + "filtered_matches.npy", d.filtered_dump, 9);
    std::vector<double> bm;
    for (size_t k = 0; k < d.hyps.size(); ++k) {
        const Hypothesis& h = d.hyps[k];
        bm.push_back(h.cam); bm.push_back(h.seg); bm.push_back(h.tgt_cam); bm.push_back(h.tgt_seg); bm.push_back(h.score); bm.push_back(1.0);
        for (int i = 0; i < 3; ++i) bm.push_back(h.P1(i));
        for (int i = 0; i < 3; ++i) bm.push_back(h.P2(i));
        for (int i = 0; i < 3; ++i) bm.push_back(h.dir(i));
        bm.push_back(h.d1); bm.push_back(h.d2);
    }
    npy_f64(p + "best_match.npy", bm, 17);
    std::vector<int32_t> pc;
    for (size_t v = 0; v < d.pc.size(); ++v)
        for (size_t k = 0; k < d.pc[v].size(); ++k) {
            pc.push_back(int32_t(d.view_at[v]->id)); pc.push_back(int32_t(d.pc[v][k].seg));
            pc.push_back(int32_t(d.pc[v][k].tcam)); pc.push_back(int32_t(d.pc[v][k].tseg));
        }
    npy_i32(p + "potential.npy", pc, 4);
    npy_f64(p + "edges_pre_diffusion.npy", d.edges_pre, 3);
    npy_f64(p + "edges_cluster.npy", d.edges_cluster, 3);
    std::vector<int32_t> nn(1, d.cluster_nodes); npy_i32(p + "cluster_nodes.npy", nn, 1);
    std::vector<double> s3; std::vector<int32_t> s2; int li = 0;
    for (std::list<L3DFinalLine3D>::iterator it = d.result.begin(); it != d.result.end(); ++it, ++li) {
        for (std::list<std::pair<V3, V3> >::iterator s = it->segments3D()->begin(); s != it->segments3D()->end(); ++s) {
            s3.push_back(li);
            for (int i = 0; i < 3; ++i) s3.push_back(s->first(i));
            for (int i = 0; i < 3; ++i) s3.push_back(s->second(i));
        }
        for (std::list<L3DSegment2D>::iterator q = it->segments2D()->begin(); q != it->segments2D()->end(); ++q) { s2.push_back(li); s2.push_back(int32_t(q->camID())); s2.push_back(int32_t(q->segID())); }
    }
    npy_f64(p + "lines_seg3d.npy", s3, 7);
    npy_i32(p + "lines_seg2d.npy", s2, 3);
    std::vector<double> tr; tr.push_back(d.tr_scale);
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) tr.push_back(d.tr_R(r, c));
    for (int i = 0; i < 3; ++i) tr.push_back(d.tr_t(i));
    npy_f64(p + "transform.npy", tr, 13);
}

}  // namespace L3D
