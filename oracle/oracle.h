/* oracle/oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * C interface of the CPU oracle: our own restatement of the reference's (manhofer/Line3D)
 * matching-and-scoring path, written from the reference's sources with file:line
 * citations at every function (oracle.cpp).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / reference arm may load it; the product (line3d_b200/) never
 * does and fails loudly when its CUDA library is missing.
 *
 * Pinning: the reference ships no tests or golden vectors, so the
 * oracle is pinned against the reference ITSELF: oracle/Makefile compiles the reference's
 * own sources for the CPU (oracle/_ref/libline3d_ref.so), tests/test_oracle_vs_ref.py
 * compares the two on seeded scenes, and tests/golden/ holds outputs of that reference
 * build (made by tests/golden/make_golden.py) which the oracle must reproduce even where
 * /root/reference is absent.
 */
#ifndef L3D_ORACLE_H_
#define L3D_ORACLE_H_
#ifdef __cplusplus
extern "C" {
#endif

/* Whole pipeline (addImage ... compute3Dmodel ... getResult) on a scene file written by
 * line3d_b200/scenes.py; writes the same .npy dumps as oracle/ref_harness.cpp into
 * dump_dir (may be NULL/""), returns the number of final 3-D lines (<0: error).
 * timings[0..4]: addImage, compute3Dmodel, matching+scoring, diffusion, clustering (s).
 * counters[0..3]: pair tests, overlap-accepted pairs, raw matches, scoring inner
 * iterations on a different camera (work accounting for BASELINE.md). */
int oracle_run_scene(const char* scene_path, int neighbors, float unc_upper, float unc_lower,
                     float sigma_p, float sigma_a, float min_baseline, int collinearity, int diffusion,
                     const char* dump_dir, int threads, double* timings, double* counters);

/* Multi-process sharding for the next oracle_run_scene calls: only the views of this rank's connected
 * components of the neighbour graph are matched (not in the reference; mirrors Line3D::setShard). */
void oracle_set_shard(int rank, int world);

/* One pair test (cudawrapper.cu:538-611).  seg = p1x p1y p2x p2y; F, RtKinv row-major 3x3.
 * out[0..3] = depths (0 if rejected); out[4], out[5] = the two overlap values, or -1 if an
 * intersection was invalid.  Returns 1 if the overlap test accepted the pair. */
int oracle_pair_test(const float* seg_src, const float* seg_tgt, const float* F,
                     const float* RtKinv_src, const float* RtKinv_tgt,
                     const float* C_src, const float* C_tgt, float* out);

/* Raw matching of one source view against one target view (cudawrapper.cu:901-945): every
 * (i, j) whose four depths are > 0, in (i, j) order.  out_ij: int[2*cap], out_depths:
 * float[4*cap].  Returns the number found (may exceed cap; only cap are written).
 * band: if non-NULL, counts pairs whose overlap values lie within 1e-5 of 0.10 / 0.30. */
long oracle_match_pair(const float* segs_src, int n_src, const float* segs_tgt, int n_tgt,
                       const float* F, const float* RtKinv_src, const float* RtKinv_tgt,
                       const float* C_src, const float* C_tgt,
                       int* out_ij, float* out_depths, long cap, long* band);

/* Scoring of one view's sorted match list (cudawrapper.cu:614-714, 958-1019).
 * data: float[4*M] (src, camLocal, tgt, 0) sorted by (src, cam, tgt); depths: float[4*M];
 * cam tables indexed by local camera id; tgt_segs: all neighbours' segments concatenated,
 * cam_offsets[cam] = first index.  Writes score[M]. */
void oracle_score_matches(const float* data, const float* depths, long M,
                          const float* segs_src, int n_src, const float* RtKinv_src, const float* C_src,
                          const float* tgt_segs, const int* cam_offsets, const float* projections,
                          int n_cams, float sigma_p, float sigma_a, float spatial_k, float* score);

/* Collinearity relation (cudawrapper.cu:476-535): dense row-major n*n. */
void oracle_collinearity(const float* segs, int n, float collin_s, float* out);

/* Diffusion (sparsematrix.cc:63-117, cudawrapper.cu:1131-1191): edges (i, j, w) in list
 * order, n nodes; out: nnz rows (row, col, val) in the returned matrix's stored order. */
void oracle_diffusion(const int* ij, const float* w, int nnz, int n, float* out);

/* 3-D affinity of two hypotheses (line3D.cc:1600-1681).  hyp = P1(3) P2(3) dir(3) as
 * doubles, depth1 depth2 as doubles; view = k_lower, k_upper, median_depth. */
float oracle_similarity(const double* hyp1, const double* view1, const double* hyp2, const double* view2, float sigma_a);

#ifdef __cplusplus
}
#endif
#endif
