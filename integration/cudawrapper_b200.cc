// integration/cudawrapper_b200.cc -- the reference's operator boundary on top of libl3d_b200.so.
//
// Line3D's host code (line3D.cc, view.cc, segments.h, sparsematrix.cc) calls three functions declared in
// cudawrapper.h:49-75 and defined in cudawrapper.cu, its only CUDA translation unit.  This file defines the same three
// functions -- same names, same signatures, same in/out conventions -- over the C ABI of include/l3d_b200.h, so that
// a maintainer can drop cudawrapper.cu from the build, add this file and link libl3d_b200.so: the reference's own,
// unmodified host code then runs on the sm_100a kernels.  (cudawrapper.cu itself no longer compiles with CUDA >= 12:
// it uses texture references.)
//
//     compute_collinearity            cudawrapper.cu:833-855    -> l3d_collinearity
//     compute_pairwise_matches        cudawrapper.cu:858-1128   -> l3d_set_views + l3d_deliver + l3d_match_view
//     replicator_dynamics_diffusion   cudawrapper.cu:1131-1191  -> l3d_diffuse
//
// It needs the reference's own headers (dataArray.h, sparsematrix.h, clustering.h) on the include path; nothing of
// them is copied here.  tests/test_gpu_operator_binding.py builds the reference's host sources against it
// (oracle/Makefile, target refhost) and compares the result with the reference's golden output.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <iostream>
#include <list>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "dataArray.h"
#include "sparsematrix.h"
#include "clustering.h"

#include "l3d_b200.h"

namespace {

l3d_ctx* context()
{
    static l3d_ctx* ctx = 0;
    if (!ctx) {
        int device = 0;
        cudaGetDevice(&device);
        if (l3d_create(device, &ctx)) { std::cerr << "[L3D] line3d-b200: " << l3d_last_error(0) << std::endl; throw std::runtime_error(l3d_last_error(0)); }
    }
    return ctx;
}
void check(int rc, const char* what)
{
    if (rc) { std::string m = std::string(what) + ": " + l3d_last_error(context()); std::cerr << "[L3D] " << m << std::endl; throw std::runtime_error(m); }
}

}  // namespace

namespace L3D
{
    // cudawrapper.cu:833-855.  `relation` is a dense N x N device array the caller downloads and scans
    // (segments.h:73-101); the library returns the compacted upper triangle, which is expanded here.
    void compute_collinearity(L3D::DataArray<float>* segments, L3D::DataArray<float>* relation, const float collin_s)
    {
        const unsigned int n = segments->height();
        std::vector<float> segs(size_t(n) * 4);
        for (unsigned int i = 0; i < n; ++i) for (unsigned int c = 0; c < 4; ++c) segs[size_t(i) * 4 + c] = segments->dataCPU(c, i)[0];
        const l3d_collin_pair* pairs = 0; size_t count = 0;
        check(l3d_collinearity(context(), segs.data(), int(n), collin_s, &pairs, &count), "l3d_collinearity");
        std::vector<float> dense(size_t(n) * n, 0.0f);
        for (size_t k = 0; k < count; ++k) {
            dense[size_t(pairs[k].j) * n + size_t(pairs[k].i)] = pairs[k].value;          // the kernel writes both halves (cudawrapper.cu:531-532)
            dense[size_t(pairs[k].i) * n + size_t(pairs[k].j)] = pairs[k].value;
        }
        cudaMemcpy2D(relation->dataGPU(), relation->pitchGPU(), dense.data(), size_t(n) * sizeof(float), size_t(n) * sizeof(float), n, cudaMemcpyHostToDevice);
    }

    // cudawrapper.cu:858-1128.  The packed tables of line3D.cc:708-803 become one small scene: view 0 is the source
    // image, views 1..k its neighbours in local-id order; the matches already filed for the source (the in-list, local
    // camera ids, view.cc:200-224) are handed to the library per neighbour; the out-list is the library's filtered
    // records with the camera ids made global again (cudawrapper.cu:1089-1110).
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
        (void)maxSegments;
        if (toBeMatched.size() == 0) return;                                       // cudawrapper.cu:877-878
        l3d_ctx* ctx = context();
        const unsigned int k = offsets->width(), n_src = segments_src->height();
        std::vector<std::vector<float> > segs(size_t(k) + 1);
        std::vector<l3d_view> views(size_t(k) + 1);
        memset(views.data(), 0, sizeof(l3d_view) * views.size());
        segs[0].resize(size_t(n_src) * 4);
        for (unsigned int i = 0; i < n_src; ++i) for (unsigned int c = 0; c < 4; ++c) segs[0][size_t(i) * 4 + c] = segments_src->dataCPU(c, i)[0];
        views[0].id = int(vID); views[0].n_segments = int(n_src); views[0].segments = segs[0].data();
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { views[0].RtKinv[r * 3 + c] = RtKinv_src->dataCPU(c, r)[0]; views[0].RtKinv_d[r * 3 + c] = views[0].RtKinv[r * 3 + c]; }
        views[0].C[0] = camCenter_src.x; views[0].C[1] = camCenter_src.y; views[0].C[2] = camCenter_src.z;
        for (int c = 0; c < 3; ++c) views[0].C_d[c] = views[0].C[c];
        views[0].spatial_k = spatial_k; views[0].k_lower = uncertainty_k_lower; views[0].k_upper = uncertainty_k_upper;
        std::vector<l3d_neighbor> nb(k);
        for (unsigned int loc = 0; loc < k; ++loc) {
            const int2 oc = offsets->dataCPU(loc, 0)[0];
            l3d_view& v = views[size_t(loc) + 1];
            segs[size_t(loc) + 1].resize(size_t(oc.y) * 4);
            for (int i = 0; i < oc.y; ++i) { const float4 s = segments_tgt->dataCPU(oc.x + i, 0)[0]; float* d = &segs[size_t(loc) + 1][size_t(i) * 4]; d[0] = s.x; d[1] = s.y; d[2] = s.z; d[3] = s.w; }
            v.id = int(local2global[loc]); v.n_segments = oc.y; v.segments = segs[size_t(loc) + 1].data();
            for (int r = 0; r < 3; ++r) {
                for (int c = 0; c < 3; ++c) { v.RtKinv[r * 3 + c] = RtKinv_tgt->dataCPU(c, loc * 3 + r)[0]; v.RtKinv_d[r * 3 + c] = v.RtKinv[r * 3 + c]; nb[loc].F[r * 3 + c] = fundamentals->dataCPU(c, loc * 3 + r)[0]; }
                for (int c = 0; c < 4; ++c) v.P[r * 4 + c] = projections->dataCPU(c, loc * 3 + r)[0];
                v.C[r] = camCenters_tgt->dataCPU(r, loc)[0]; v.C_d[r] = v.C[r];
            }
            nb[loc].view = int(loc) + 1; nb[loc].to_be_matched = 0; nb[loc].deliver = 0;
        }
        for (std::list<unsigned int>::iterator it = toBeMatched.begin(); it != toBeMatched.end(); ++it) if (*it < k) nb[*it].to_be_matched = 1;
        check(l3d_set_views(ctx, int(views.size()), views.data()), "l3d_set_views");
        // the matches earlier views filed for this one, per neighbour
        std::vector<std::vector<l3d_match> > pre(k);
        for (std::list<L3D::L3DMatchingPair>::iterator it = matches.begin(); it != matches.end(); ++it) {
            if (it->camID2_ >= k) continue;
            l3d_match m; m.src_seg = int(it->segID1_); m.cam = int(it->camID2_); m.tgt_seg = int(it->segID2_); m.confidence = it->confidence_;
            m.depths[0] = it->depths_.x; m.depths[1] = it->depths_.y; m.depths[2] = it->depths_.z; m.depths[3] = it->depths_.w;
            pre[it->camID2_].push_back(m);
        }
        for (unsigned int loc = 0; loc < k; ++loc)
            if (!pre[loc].empty() && !nb[loc].to_be_matched) check(l3d_deliver(ctx, 0, int(loc) + 1, pre[loc].size(), pre[loc].data()), "l3d_deliver");
        l3d_match_result res;
        check(l3d_match_view(ctx, 0, int(k), nb.data(), sigma_p, sigma_a, &res), "l3d_match_view");
        if (verbose) std::cout << prefix << "#raw_matches:          " << res.n_scored << std::endl;
        if (res.n_scored == 0) return;                                              // cudawrapper.cu:955-956
        matches.clear();                                                            // cudawrapper.cu:1022
        median_depth = res.median_depth;
        for (uint64_t i = 0; i < res.n_filtered; ++i) {                             // cudawrapper.cu:1089-1110
            const l3d_match& m = res.filtered[i];
            L3D::L3DMatchingPair mp;
            mp.segID1_ = unsigned(m.src_seg); mp.camID2_ = local2global[unsigned(m.cam)]; mp.segID2_ = unsigned(m.tgt_seg);
            mp.depths_ = make_float4(m.depths[0], m.depths[1], m.depths[2], m.depths[3]);
            mp.confidence_ = m.confidence; mp.active_ = true;
            matches.push_back(mp);
        }
        if (verbose) std::cout << prefix << "#filtered_matches (2): " << matches.size() << std::endl;
    }

    // cudawrapper.cu:1131-1191: W comes in column-sorted, is deleted, and comes back as the row-sorted result.
    void replicator_dynamics_diffusion(L3D::SparseMatrix*& W, const bool verbose, const std::string prefix)
    {
        const unsigned int nnz = W->num_entries(), n = W->num_rows_cols();
        if (nnz == 0 || n == 0) return;
        if (verbose) std::cout << prefix << "diffusion on " << nnz << " entries" << std::endl;
        std::vector<l3d_edge> in(nnz), out(nnz);
        for (unsigned int i = 0; i < nnz; ++i) { const float4 e = W->entries()->dataCPU(i, 0)[0]; in[i].i = int(e.x); in[i].j = int(e.y); in[i].w = e.z; }
        check(l3d_diffuse(context(), int(n), nnz, in.data(), out.data()), "l3d_diffuse");
        std::list<float4> entries;
        for (unsigned int i = 0; i < nnz; ++i) entries.push_back(make_float4(float(out[i].i), float(out[i].j), out[i].w, 0.0f));
        delete W;                                                                   // cudawrapper.cu:1186-1188
        W = new L3D::SparseMatrix(entries, n, 1.0f, true, true);
    }
}
