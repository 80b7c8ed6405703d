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

struct Session {
    L3D::Line3D* l3d = nullptr;
    bool verbose = false;
    double t_add = 0.0;
};
struct Quiet {      // silences std::cout for the scope unless verbose
    std::streambuf* old = nullptr; std::ostringstream sink;
    explicit Quiet(bool verbose) { if (!verbose) old = std::cout.rdbuf(sink.rdbuf()); }
    ~Quiet() { if (old) std::cout.rdbuf(old); }
};
}

extern "C" {

// Line3D(...) + addImage for every view of the scene.  Returns an opaque session or NULL.
void* l3dhost_open(const char* scene_path, const char* data_dir, int neighbors, float unc_upper, float unc_lower,
                   float sigma_p, float sigma_a, float min_baseline, int collinearity, int verbose)
{
    std::vector<SceneView> sv;
    if (!read_scene(scene_path, sv)) { std::cerr << "l3dhost_open: cannot read " << scene_path << std::endl; return nullptr; }
    Quiet q(verbose != 0);
    Session* s = new Session();
    s->verbose = verbose != 0;
    try {
        s->l3d = new L3D::Line3D(data_dir, neighbors, unc_upper, unc_lower, sigma_p, sigma_a, min_baseline, collinearity != 0, verbose != 0);
        const double t0 = now_s();
        for (size_t i = 0; i < sv.size(); ++i) {
            SceneView& v = sv[i];
            L3D::Line3D::writeSegmentCache(data_dir, v.id, v.width, v.height, collinearity != 0, v.segs);
            cv::Mat image(int(v.height), int(v.width), CV_8UC3);
            std::list<unsigned int> wps(v.wps.begin(), v.wps.end());
            s->l3d->addImage(v.id, image, mat3(v.K), mat3(v.R), Eigen::Vector3d(v.t[0], v.t[1], v.t[2]), wps, int(std::max(v.width, v.height)), true);
        }
        s->t_add = now_s() - t0;
    } catch (const std::exception& e) {
        std::cerr << "l3dhost_open: " << e.what() << std::endl;
        delete s->l3d; delete s;
        return nullptr;
    }
    return s;
}

// compute3Dmodel + getResult.  stats_out: 40 doubles (see line3d_b200/host.py STAT_NAMES).
// Returns the number of 3-D lines or < 0.
int l3dhost_compute(void* session, int diffusion, int measure_only, int flush_l2, int keep_debug, int stage_timing, double* wall_out, double* stats_out)
{
    Session* s = static_cast<Session*>(session);
    if (!s || !s->l3d) return -1;
    Quiet q(s->verbose);
    try {
        s->l3d->setMeasureOnly(measure_only == 1);
        s->l3d->setMatchOnly(measure_only == 2);
        s->l3d->setFlushL2(flush_l2 != 0);
        s->l3d->setKeepDebug(keep_debug != 0);
        s->l3d->setStageTiming(stage_timing != 0);
        const double t0 = now_s();
        s->l3d->compute3Dmodel(diffusion != 0);
        const double t1 = now_s();
        if (wall_out) { wall_out[0] = s->t_add; wall_out[1] = t1 - t0; }
        if (stats_out) {
            L3D::L3DRunStats st = s->l3d->runStats();
            const double v[40] = {double(st.pair_tests), double(st.candidates), double(st.raw_matches), double(st.scored_matches), double(st.filtered_matches),
                                  double(st.score_iters), double(st.score_iters_other_cam), double(st.hypotheses), double(st.affinity_pairs), double(st.edges),
                                  double(st.kernel_launches), st.t_neighbors, st.t_upload, st.t_matching, st.t_matching_device, st.t_hypotheses, st.t_affinity,
                                  st.t_diffusion, st.t_clustering, st.t_lines, st.t_total, st.t_stage[0], st.t_stage[1], st.t_stage[2], st.t_stage[3], st.t_stage[4],
                                  double(st.h2d_bytes), double(st.d2h_bytes), st.t_matching_calls, st.t_k_pretest, st.t_k_exact, st.t_k_pretest2,
                                  st.rdd_step_ms, st.rdd_step_bytes, st.rdd_entries, 0.0, 0.0, 0.0, 0.0, 0.0};
            std::memcpy(stats_out, v, sizeof v);
        }
        std::list<L3D::L3DFinalLine3D> result;
        s->l3d->getResult(result);
        return int(result.size());
    } catch (const std::exception& e) {
        std::cerr << "l3dhost_compute: " << e.what() << std::endl;
        return -3;
    }
}

int l3dhost_reset_geometry(void* session)
{
    Session* s = static_cast<Session*>(session);
    if (!s || !s->l3d) return -1;
    s->l3d->resetGeometry();
    return 0;
}

int l3dhost_set_shard(void* session, int rank, int world)
{
    Session* s = static_cast<Session*>(session);
    if (!s || !s->l3d) return -1;
    s->l3d->setShard(rank, world);
    return 0;
}

int l3dhost_set_split(void* session, int rank, int world, const unsigned char* nccl_id)
{
    Session* s = static_cast<Session*>(session);
    if (!s || !s->l3d) return -1;
    try { s->l3d->setSplit(rank, world, nccl_id); } catch (const std::exception& e) { std::cerr << "l3dhost_set_split: " << e.what() << std::endl; return -2; }
    return 0;
}

int l3dhost_dump(void* session, const char* dump_dir)
{
    Session* s = static_cast<Session*>(session);
    if (!s || !s->l3d || !dump_dir) return -1;
    s->l3d->debugDump(dump_dir);
    return 0;
}

void l3dhost_close(void* session)
{
    Session* s = static_cast<Session*>(session);
    if (!s) return;
    { Quiet q(s->verbose); delete s->l3d; }
    delete s;
}

}  // extern "C"
