// line3d_b200/csrc/host_harness.cpp -- C entry points that drive the Line3D C++ API
// (line3D.h) on a synthetic scene file (line3d_b200/scenes.py), for tests and bench.py.
// Segments enter through the segment cache, the way the reference lets precomputed
// segments in (line3D.cc:143-168).
#include "line3D.h"

#include <chrono>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <vector>

namespace {
struct SceneView { uint32_t id, width, height; double K[9], R[9], t[3]; std::vector<float4> segs; std::vector<uint32_t> wps; };
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
        v.segs.resize(hdr[3]); if (hdr[3]) f.read(reinterpret_cast<char*>(v.segs.data()), std::streamsize(size_t(hdr[3]) * 16));
        v.wps.resize(hdr[4]); if (hdr[4]) f.read(reinterpret_cast<char*>(v.wps.data()), std::streamsize(size_t(hdr[4]) * 4));
    }
    return bool(f);
}
Eigen::Matrix3d mat3(const double* m) { Eigen::Matrix3d M; for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) M(r, c) = m[r * 3 + c]; return M; }
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}

extern "C" {

// stats_out: 20 doubles -- pair_tests, candidates, raw, scored, filtered, score_iters, hypotheses, affinity_pairs,
// edges, kernel_launches, t_neighbors, t_upload, t_matching, t_matching_device, t_hypotheses, t_affinity, t_diffusion,
// t_clustering, t_lines, t_total.  timings_out: addImage total, compute3Dmodel total (seconds, host clock).
// repeat > 1 runs compute3Dmodel several times (the reference allows that, line3D.cc:356-358) and reports the last.
int l3dhost_run_scene(const char* scene_path, const char* data_dir, int neighbors, float unc_upper, float unc_lower,
                      float sigma_p, float sigma_a, float min_baseline, int collinearity, int diffusion, int verbose,
                      int measure_only, int repeat, const char* dump_dir, double* timings_out, double* stats_out)
{
    std::vector<SceneView> sv;
    if (!read_scene(scene_path, sv)) return -1;
    std::streambuf* old_cout = 0; std::ostringstream sink;
    if (!verbose) old_cout = std::cout.rdbuf(sink.rdbuf());
    int n_lines = -2;
    try {
        L3D::Line3D l3d(data_dir, neighbors, unc_upper, unc_lower, sigma_p, sigma_a, min_baseline, collinearity != 0, verbose != 0);
        l3d.setMeasureOnly(measure_only != 0);
        const double t0 = now_s();
        for (size_t i = 0; i < sv.size(); ++i) {
            SceneView& v = sv[i];
            L3D::Line3D::writeSegmentCache(data_dir, v.id, v.width, v.height, collinearity != 0, v.segs);
            cv::Mat image(int(v.height), int(v.width), CV_8UC3);
            std::list<unsigned int> wps(v.wps.begin(), v.wps.end());
            l3d.addImage(v.id, image, mat3(v.K), mat3(v.R), Eigen::Vector3d(v.t[0], v.t[1], v.t[2]), wps, int(std::max(v.width, v.height)), true);
        }
        const double t1 = now_s();
        double t2 = t1, t3 = t1;
        for (int r = 0; r < (repeat < 1 ? 1 : repeat); ++r) { t2 = now_s(); l3d.compute3Dmodel(diffusion != 0); t3 = now_s(); }
        if (timings_out) { timings_out[0] = t1 - t0; timings_out[1] = t3 - t2; }
        if (stats_out) {
            L3D::L3DRunStats s = l3d.runStats();
            const double v[20] = {double(s.pair_tests), double(s.candidates), double(s.raw_matches), double(s.scored_matches), double(s.filtered_matches),
                                  double(s.score_iters), double(s.hypotheses), double(s.affinity_pairs), double(s.edges), double(s.kernel_launches),
                                  s.t_neighbors, s.t_upload, s.t_matching, s.t_matching_device, s.t_hypotheses, s.t_affinity, s.t_diffusion,
                                  s.t_clustering, s.t_lines, s.t_total};
            std::memcpy(stats_out, v, sizeof v);
        }
        if (dump_dir && dump_dir[0]) l3d.debugDump(dump_dir);
        std::list<L3D::L3DFinalLine3D> result;
        l3d.getResult(result);
        n_lines = int(result.size());
    } catch (const std::exception& e) {
        if (old_cout) std::cout.rdbuf(old_cout);
        std::cerr << "l3dhost_run_scene: " << e.what() << std::endl;
        return -3;
    }
    if (old_cout) std::cout.rdbuf(old_cout);
    return n_lines;
}

}  // extern "C"
