/* include/l3d_b200.h -- C ABI of libl3d_b200.so: the Blackwell (sm_100a) implementation of
 * Line3D's data-parallel matching-and-scoring path.
 *
 * This is the drop-in boundary.  It replaces the three C++ operator functions the
 * reference's host code calls into its CUDA translation unit (cudawrapper.h:49-75):
 *
 *     compute_collinearity(...)            -> l3d_collinearity
 *     compute_pairwise_matches(...)        -> l3d_set_views + l3d_match_view
 *     replicator_dynamics_diffusion(...)   -> l3d_diffuse
 *
 * and moves onto the GPU the two pieces of the path that the reference runs in host
 * double precision (view.cc:302-342 unprojectSegment, line3D.cc:1600-1691
 * similarity_coll3D)                       -> l3d_unproject, l3d_affinity.
 *
 * Plain C: POD structs, pointers and sizes only; no STL, Eigen or torch types.  Every
 * call returns 0 on success, non-zero on failure with a message from l3d_last_error().
 * A context is bound to one CUDA device and one host thread; it owns all device memory,
 * nothing is global.  Host input buffers are caller-owned and only read during the call.
 * There is no CPU fallback: without a usable CUDA device l3d_create fails.
 */
#ifndef L3D_B200_H_
#define L3D_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct l3d_ctx l3d_ctx;

/* ---- lifetime ------------------------------------------------------------------------ */
int l3d_create(int device, l3d_ctx** out);
void l3d_destroy(l3d_ctx* ctx);
const char* l3d_last_error(const l3d_ctx* ctx);   /* ctx may be NULL: last l3d_create error */

/* ---- collinearity: replaces compute_collinearity (cudawrapper.h:49-51,
 *      cudawrapper.cu:476-535, 833-855) and the dense N*N download + triangle scan of
 *      segments.h:73-101.  segments: n x (p1x p1y p2x p2y).  On return *pairs points to
 *      *count records with i < j and value > 0, sorted by (i, j); the buffer belongs to
 *      the context and stays valid until the next l3d_collinearity call or l3d_destroy. */
typedef struct { int32_t i, j; float value; } l3d_collin_pair;
int l3d_collinearity(l3d_ctx* ctx, const float* segments, int n, float collin_s,
                     const l3d_collin_pair** pairs, size_t* count);

/* ---- collinearity for callers that must not allocate.  No counterpart in the reference.
 *
 *      Who needs it: a host program that hands out its whole GPU memory budget at start-up
 *      and forbids (or cannot survive) later device allocations -- a long-running
 *      reconstruction service sharing the card with other components, or a frame loop with
 *      a latency bound, for which cudaMalloc/cudaFree are unacceptable because they
 *      synchronise the device for an unbounded time and may fail once the card is full.
 *      l3d_collinearity grows context-owned buffers with cudaMalloc; this variant allocates
 *      nothing, ever: its working memory is a fixed array compiled into the library
 *      (about 2.1 MB of device memory on every GPU that loads the library, held until the
 *      process exits), and inputs and outputs are the caller's own DEVICE buffers.
 *
 *      d_segments  n x 4 floats on the device, 16-byte aligned (it is read as float4; a
 *                  misaligned pointer is refused with return code 1)
 *      d_pairs     room for `capacity` records on the device (4-byte aligned)
 *      d_count     one int on the device: receives the number of collinear pairs found.  If
 *                  it exceeds `capacity`, only the first `capacity` records (in (i, j) order)
 *                  were written.
 *      cuda_stream a cudaStream_t (NULL = default stream); the work is enqueued, not waited for.
 *                  The stream's device must be the CURRENT device (return code 7 otherwise):
 *                  the built-in array is looked up on the current device.
 *      err/err_len optional buffer for a message when the return value is non-zero.
 *
 *      Limit: n <= 4096 segments per image; larger inputs are refused with return code 2
 *      (the reference's own detector cap is 3000, commons.h:44).
 *      The first call on each device looks up the device address of the built-in array
 *      (cudaGetSymbolAddress) with that device current; make that first call during
 *      start-up, before any stream capture is begun (the look-up result is cached per
 *      device in an atomic table: concurrent first calls are safe, both store the same
 *      address).  All calls on a device share the one
 *      array and it keeps whatever the previous call left in it (the part that has to start
 *      from zero is cleared by each call); two calls whose GPU work overlaps in time -- on
 *      different streams or from different threads -- race on it and must be serialised by
 *      the caller.  Results are identical to l3d_collinearity. */
int l3d_collinearity_prealloc(const float* d_segments, int n, float collin_s, l3d_collin_pair* d_pairs, int capacity,
                              int* d_count, void* cuda_stream, char* err, size_t err_len);

/* ---- scene tables: what line3D.cc:738-803 packs per view, uploaded once for all views.
 *      Matrices are row-major.  The float tables are the reference's double values cast
 *      to float (line3D.cc:745-763, 790); the double ones feed l3d_unproject. */
typedef struct {
    int32_t id;              /* caller's view id (reported back, not interpreted)        */
    int32_t n_segments;
    const float* segments;   /* n_segments x 4                                           */
    float RtKinv[9];         /* R^T K^-1                    view.cc:26                   */
    float C[3];              /* camera centre               view.cc:27                   */
    float P[12];             /* K [R|t]                     view.cc:31-34                */
    float spatial_k;         /* specificSpatialUncertaintyK(2 sigma_p), line3D.cc:820    */
    float k_lower, k_upper;  /* view.cc:90-121                                           */
    double RtKinv_d[9];
    double C_d[3];
} l3d_view;
int l3d_set_views(l3d_ctx* ctx, int n_views, const l3d_view* views);

/* ---- matching + scoring of one view: replaces compute_pairwise_matches
 *      (cudawrapper.h:54-71, cudawrapper.cu:858-1128) together with the exchange of
 *      reversed matches between views that the reference routes through files
 *      (line3D.cc:838-872, view.cc:162-224).
 *
 *      neighbors[] lists the view's visual neighbours in ascending view id; the position
 *      in the array is the reference's local camera id (line3D.cc:723-730).
 *        to_be_matched  = 1: pair-test this neighbour now (line3D.cc:732-736);
 *                         0: it was processed earlier as a mutual neighbour; the records
 *                            it delivered to this view are consumed instead.
 *        deliver        = 1: hand this view's filtered matches, reversed, to that
 *                            neighbour for when it is processed (line3D.cc:844-858).
 *      Views must be processed in the reference's order (ascending id, line3D.cc:626).
 *      If no neighbour has to_be_matched set the reference does not score at all
 *      (cudawrapper.cu:877-878): the delivered records come back unscored (scored = 0,
 *      confidence 0, median_depth 1.0), sorted by (src, cam, tgt).
 *
 *      Results: records with score > 1 (stored score = score / 2, cudawrapper.cu:1096-1098)
 *      in (src, cam, tgt) order, cam = position in neighbors[].  The record buffer is
 *      context-owned pinned host memory, valid until the next l3d_match_view. */
typedef struct {
    int32_t view;            /* index into the l3d_set_views array                       */
    float F[9];              /* fundamental matrix src -> this neighbour, line3D.cc:739   */
    int32_t to_be_matched;
    int32_t deliver;
} l3d_neighbor;
typedef struct {
    int32_t src_seg;
    int32_t cam;             /* position in neighbors[]                                   */
    int32_t tgt_seg;
    float confidence;
    float depths[4];         /* src p1, src p2, tgt q1, tgt q2 (cudawrapper.cu:594-601)   */
} l3d_match;
typedef struct {
    uint64_t pair_tests;     /* (src seg, tgt seg) evaluations done for this view         */
    uint64_t n_candidates;   /* pairs the conservative pre-test passed to the exact test  */
    uint64_t n_raw;          /* new raw matches (all four depths > 0)                     */
    uint64_t n_delivered;    /* records consumed from earlier mutual neighbours           */
    uint64_t n_scored;       /* n_raw + n_delivered                                       */
    uint64_t n_filtered;
    uint64_t score_iters;    /* sum over src segments of (matches of the segment)^2       */
    uint64_t score_iters_other_cam; /* ... minus the same-camera pairs: the scoring work unit  */
    int32_t scored;
    float median_depth;      /* cudawrapper.cu:1066-1076                                  */
    const l3d_match* filtered;
} l3d_match_result;
/* Limits: at most 1600 neighbours per view (the scoring kernel keeps one maximum per (camera, hypothesis) in shared
 * memory); a batch of pair tests holds at most 65 535 (view, neighbour) jobs and 2^31 - 65 536 candidates (batches are
 * cut accordingly, see l3d_debug_set_candidate_limit); segments, matches and records of one view are counted with
 * 32 bits.  Anything beyond fails with a message in l3d_last_error, never silently. */
int l3d_match_view(l3d_ctx* ctx, int view, int n_neighbors, const l3d_neighbor* neighbors,
                   float sigma_p, float sigma_a, l3d_match_result* result);
/* Hands records to a view from outside: what the reference's host does when it loads the matches earlier views have
 * filed for a view and passes them into compute_pairwise_matches through its in/out list (view.cc:200-224,
 * line3D.cc:806).  `records` are in the RECEIVING view's frame: src_seg = its segment, tgt_seg = the segment of
 * from_view, depths as l3d_match lists them (cam and confidence are ignored).  The next l3d_match_view of to_view that
 * lists from_view with to_be_matched = 0 consumes them like records delivered inside the library.  Used by the
 * operator-level binding (integration/cudawrapper_b200.cc). */
int l3d_deliver(l3d_ctx* ctx, int to_view, int from_view, size_t n, const l3d_match* records);
/* Optional: with on != 0, l3d_match_result::filtered of every following l3d_match_view call stays valid until
 * l3d_reset_matching / l3d_set_views / l3d_destroy instead of until the next call (the records of all views are kept
 * in pinned host memory, 32 bytes each, reused after a reset).  A caller that files the records into its own
 * structures (view.cc:162-224 in the reference) can then do so after the loop, off the GPU's critical path.
 * No counterpart in the reference. */
int l3d_retain_results(l3d_ctx* ctx, int on);
/* Optional: announce the whole schedule (the views in the order they will be passed to
 * l3d_match_view, each with its neighbour array) before the first l3d_match_view call.
 * Which neighbours get pair-tested is a static function of the neighbour graph and the view
 * order (line3D.cc:732-736, 875-881), so the pair tests of many views can then run as a few
 * large launches ahead of the per-view scoring, instead of one small launch per neighbour as
 * in cudawrapper.cu:900-945, and the scoring of the next announced view is put on the GPU before the
 * host waits for the current one.  Results are identical with or without a plan.  Because of that run-ahead the
 * announced views must then be asked for in order, with the announced neighbours and the same sigmas throughout:
 * a call that deviates before the schedule has started voids the schedule, one that deviates while a view is
 * already running ahead fails (l3d_last_error says so; l3d_reset_matching starts over). */
/* Tuning, read from the environment once per schedule (defaults are what bench.py measures): L3D_BATCH_DIV (about how
 * many batches of pair tests a schedule is cut into, 8), L3D_GROUP_VIEWS (views of one dependency level per launch of
 * the per-view kernels: 8, with several GPUs on one scene 16, at most 32), L3D_NO_OVERLAP=1 (pair tests of the next
 * batch do not run beside the per-view work), L3D_BG_BLOCKS (blocks per SM the pair-test kernels may take while they
 * run beside it, 0 = no cap), L3D_P1_ROWS (source rows per block of the level-1 pre-test, 512). */
typedef struct { int32_t view; int32_t n_neighbors; const l3d_neighbor* neighbors; } l3d_view_plan;
int l3d_plan_matching(l3d_ctx* ctx, int n_views, const l3d_view_plan* plans);
/* Tests only: pair tests per batch of the next announced schedules (0 = automatic: about an eighth of the schedule,
 * at least 2.5e8, at most 2.4e10).  Small values force many batches on small scenes. */
int l3d_debug_set_batch_pairs(l3d_ctx* ctx, uint64_t pairs);
/* Tests only: mode 1 switches the conservative pre-tests off (every pair of a computed view pair goes to the exact
 * test, as in cudawrapper.cu:538-611); mode 0 (default) switches them back on.  Results must be identical. */
int l3d_debug_set_pretest(l3d_ctx* ctx, int mode);
/* Tests only: candidates (pairs that survive the pre-tests) are indexed with 32 bits per batch of pair tests.  A batch
 * with more than 2147418112 of them -- a candidate rate far above the usual 2 %, e.g. degenerate epipolar geometry --
 * is cut in two and redone; if a single view alone exceeds the limit l3d_match_view fails with a message.  This call
 * lowers the limit so that a test can reach that path with a small scene (0 = back to the default). */
int l3d_debug_set_candidate_limit(l3d_ctx* ctx, uint64_t limit);
/* Tests only: the scoring kernels hold one maximum per (camera, hypothesis) in shared memory and therefore take 128
 * hypotheses per block for up to 192 neighbours, 64 up to 384 and 32 up to 1600 (more are refused).  This call forces
 * 32 or 64 (0 = automatic) so that a small scene exercises the blocks a view with hundreds of neighbours gets. */
int l3d_debug_set_score_block(l3d_ctx* ctx, int hypotheses);
/* Forget all delivered-but-unconsumed records and any plan (a new compute3Dmodel run, line3D.cc:356). */
int l3d_reset_matching(l3d_ctx* ctx);

/* ---- several GPUs on ONE scene.  No counterpart in the reference (single implicit device, cudawrapper.cu).
 *      One process and one context per GPU of one node.  After l3d_comm_init the contexts of the `world` ranks work
 *      on the same scene together: the segments of every view are dealt out to the ranks in turn (segment s of a view
 *      belongs to rank s mod world), each rank pair-tests, assembles and scores its rows of every view, and the views
 *      are still processed in the reference's order of dependence.  What travels between the GPUs: over NCCL (from the
 *      process's libnccl.so.2, opened at run time) the segment table from rank 0 at l3d_set_views (a rank other than 0
 *      may pass segments = NULL), the raw-match totals per batch of pair tests (the bounds all ranks size their buffers
 *      with) and the handshake when the exchange windows are (re)allocated; per group of independent views, by stores
 *      into CUDA-IPC-mapped windows of the other ranks (NVLink) and an arrival flag per rank, every output record that
 *      a later view consumes (to the rank that holds the target segment), the depth samples for the median and the
 *      ranks' counts -- and all output records to the rank(s) that hand them to the host.
 *      Every rank must make the same calls in the same order with the same arguments (l3d_set_views, l3d_plan_matching,
 *      l3d_match_view ...); a rank that fails leaves the others waiting (ten seconds inside a group, then they fail too).
 *      Results: n_filtered, median_depth and `filtered` describe the whole view -- `filtered` holds the records of ALL
 *      ranks (grouped by rank, each group in (src, cam, tgt) order) on rank result_root, or on every rank if
 *      result_root < 0, and is NULL elsewhere; pair_tests, n_candidates, n_raw, n_delivered, n_scored and the score
 *      iteration counters count this rank's rows (their sums over the ranks are the single-GPU values).
 *      l3d_comm_unique_id: rank 0 obtains an id (ncclGetUniqueId) and hands it to the other ranks by any means. */
#define L3D_COMM_ID_BYTES 128
int l3d_comm_unique_id(void* id /* L3D_COMM_ID_BYTES bytes */);
int l3d_comm_init(l3d_ctx* ctx, int rank, int world, const void* id, int result_root);
int l3d_comm_destroy(l3d_ctx* ctx);

/* Device-only variant used for measurement: same work, but nothing is copied back except
 * the counters; filtered is NULL. */
int l3d_match_view_resident(l3d_ctx* ctx, int view, int n_neighbors, const l3d_neighbor* neighbors,
                            float sigma_p, float sigma_a, l3d_match_result* result);

/* ---- in front of the pair tests: the neighbour graph and the fundamental matrices.
 *      l3d_camera is a view's calibration as the host holds it (FP64, row-major; C = -R^T t as the host computed it).
 *      Both calls evaluate the host code's formulas in the host code's order of operations with individually rounded
 *      FP64 operations, so the results are the host code's, bit for bit.
 *
 *      l3d_visual_neighbors replaces the selection loop of Line3D::findVisualNeighbors (line3D.cc:498-546): for every
 *      view the co-visible views (ascending view index) and their similarities (line3D.cc:483-495) are passed in CSR
 *      form; a co-visible view is kept if its camera centre is farther than min_baseline from the view's and from that
 *      of every view kept before it, the kept ones are ranked by similarity (descending, stable), cut to k (k <= 0:
 *      all) and returned in ascending order: count[v] of them at neighbors[start[v]...].
 *
 *      l3d_fundamentals replaces Line3D::fundamental (line3D.cc:1949-1993) for a list of (source, neighbour) index
 *      pairs: R = R2 R1^T, t = t2 - R t1, F = K2^-T [t]x R K1^-1; 9 doubles per pair, row-major. */
typedef struct { double K[9]; double R[9]; double t[3]; double C[3]; } l3d_camera;
int l3d_visual_neighbors(l3d_ctx* ctx, int n_views, const l3d_camera* cameras, const int32_t* start /* n_views + 1 */,
                         const int32_t* other, const float* similarity, float min_baseline, int k,
                         int32_t* count /* n_views */, int32_t* neighbors /* start[n_views] */);
int l3d_fundamentals(l3d_ctx* ctx, int n_views, const l3d_camera* cameras, size_t n_pairs, const int32_t* pairs /* 2 per pair */,
                     double* F /* 9 per pair */);
/* ---- after the clustering: the 3-D line of every cluster (line3D.cc:1306-1368, 1392-1451, 1479-1597).
 *      A cluster is a list of hypotheses (indices into the array of the last l3d_unproject, which is still resident), in
 *      (camera, segment) order, with the camera id of each; clusters are given in CSR form.  Per cluster: the end
 *      points go back to world coordinates (`back`, line3D.cc:1782-1786), the line runs through their centroid along
 *      the dominant direction of their scatter matrix, and a sweep along it emits every stretch over which at least
 *      three cameras see the line.  seg_count[c] segments of 6 doubles (start, end) are written at segments +
 *      6 * start[c]; a cluster of m hypotheses yields at most m.  FP64, the host code's order of operations. */
typedef struct { double scale_inv; double Rinv[9]; double tneg[3]; } l3d_similarity;      /* P -> Rinv (P * scale_inv + tneg) */
int l3d_fit_lines(l3d_ctx* ctx, const l3d_similarity* back, int n_clusters, const int32_t* start /* n_clusters + 1 */,
                  const int32_t* member, const uint32_t* member_camera, int32_t* seg_count /* n_clusters */, double* segments /* 6 * start[n_clusters] */);
/* Tests only: the same source compiled for the host (no GPU needed), so that the arithmetic can be checked against the
 * reference's numbers in the CPU test suite.  Not used by the product. */
int l3d_debug_fit_lines_host(const l3d_similarity* back, int n_clusters, const int32_t* start, const double* points /* 6 per member: P1, P2 */,
                             const uint32_t* member_camera, int32_t* seg_count, double* segments);
int l3d_debug_visual_neighbors_host(int n_views, const l3d_camera* cameras, const int32_t* start, const int32_t* other,
                                    const float* similarity, float min_baseline, int k, int32_t* count, int32_t* neighbors);
int l3d_debug_fundamentals_host(int n_views, const l3d_camera* cameras, size_t n_pairs, const int32_t* pairs, double* F);

/* ---- hypotheses: replaces L3DView::unprojectSegment (view.cc:302-342), FP64 on device.
 *      The n hypotheses stay resident for l3d_affinity. */
typedef struct { int32_t view; int32_t seg; float depth1, depth2; } l3d_hypothesis_in;
typedef struct { double P1[3], P2[3], dir[3]; } l3d_segment3d;
int l3d_unproject(l3d_ctx* ctx, size_t n, const l3d_hypothesis_in* in, l3d_segment3d* out);

/* ---- pairwise 3-D affinity: replaces Line3D::similarity_coll3D (line3D.cc:1600-1691)
 *      for a list of candidate pairs (indices into the l3d_unproject array), in any order.
 *      median_depth: one value per view of l3d_set_views (view.cc:353-377 cap).
 *      out[k] = the reference's sim in [0,1] (0 where sim <= 0.01). */
typedef struct { int32_t a, b; } l3d_pair;
int l3d_affinity(l3d_ctx* ctx, const float* median_depth, size_t n_pairs, const l3d_pair* pairs,
                 float sigma_a, float* out);

/* ---- candidate pairs -> affinity edges, thresholded on the device: the similarity of l3d_affinity followed by the
 *      reference's edge weight and acceptance test (line3D.cc:1014-1016, 1085-1087, 1163-1165)
 *          w = mult * 0.5 * (score[a] + score[b]) * sim,     kept if w > thr
 *      (mult: the collinearity weight for pairs inside one image, 1 otherwise; thr: 0.25 for a direct correspondence,
 *      0.01 for the others).  Only the kept candidates come back, in candidate order, as (index into cands, w) -- for
 *      the urban scene 2.4e6 of 2.2e7.  l3d_candidate_buffer hands out pinned host memory (context-owned, valid until
 *      the next call) to build the candidate list in; any host pointer is accepted by l3d_affinity_edges, pinned
 *      memory just travels faster.  hyp_score: one value per hypothesis of the last l3d_unproject.
 *      order (optional, may be NULL): receives the kept candidates' positions in the order the clustering takes its
 *      edges -- a stable sort by ascending weight (clustering.cc:12) -- sorted on the device: order[t] is the index
 *      into kept[] of the t-th lightest edge (equal weights in candidate order).
 *      The result buffers belong to the context and stay valid until the next l3d_affinity_edges call. */
typedef struct { int32_t a, b; float mult, thr; } l3d_candidate;
typedef struct { uint32_t index; float w; } l3d_kept_edge;
int l3d_candidate_buffer(l3d_ctx* ctx, size_t n, l3d_candidate** buf);
int l3d_affinity_edges(l3d_ctx* ctx, const float* median_depth, const float* hyp_score, size_t n, const l3d_candidate* cands,
                       float sigma_a, const l3d_kept_edge** kept, size_t* n_kept, const uint32_t** order);

/* ---- diffusion: replaces SparseMatrix construction (sparsematrix.cc:63-191) and
 *      replicator_dynamics_diffusion (cudawrapper.h:74, cudawrapper.cu:1131-1191).
 *      edges: nnz directed entries (i, j, w), unique (i, j), structurally symmetric, every
 *      node 0..n_nodes-1 present.  out: the nnz entries of the result in the reference's
 *      stored order (sorted by row, then column). */
typedef struct { int32_t i, j; float w; } l3d_edge;
int l3d_diffuse(l3d_ctx* ctx, int n_nodes, size_t nnz, const l3d_edge* edges, l3d_edge* out);
/* Measurement: of the last l3d_diffuse call, [0] the duration in ms of one step kernel (the second of the ten
 * iterations, CUDA events), [1] the algorithmic bytes one iteration moves -- per entry 16 bytes plus 32 bytes per
 * lock-step pair, min(row length, column length) of them --, [2] the number of entries. */
int l3d_diffuse_stats(l3d_ctx* ctx, double out[3]);

/* ---- measurement helpers -------------------------------------------------------------- */
/* Device time in ms of the kernels of the last l3d_match_view[_resident] call, by stage:
 * [0] pre-test, [1] exact test, [2] assembly, [3] scoring, [4] reduce/filter/deliver,
 * [5] their sum.  Timed with CUDA events on the context's stream.  [0] and [1] are reported on the first call
 * after a batch of pair tests ran; [2]..[4] only with l3d_set_stage_timing(ctx, 1), zero otherwise: the events
 * between the stages keep consecutive kernels from overlapping (programmatic dependent launch) and cost about
 * 13 % of a step, so they are off unless asked for. */
int l3d_last_timings(l3d_ctx* ctx, float ms[6]);
/* Duration in ms of the three large launches of the last batch of pair tests alone: [0] the level-1 pre-test kernel
 * (k_pretest), [1] the exact-test kernel, [2] the level-2 pre-test kernel (k_pretest2); valid after the
 * l3d_last_timings call that reported a non-zero ms[0]. */
int l3d_last_kernel_timings(l3d_ctx* ctx, float ms[3]);
int l3d_set_stage_timing(l3d_ctx* ctx, int on);
/* Two CUDA events on the context's stream for timing any span of calls on the device: l3d_mark(ctx, 0) before,
 * l3d_mark(ctx, 1) after, l3d_marks_elapsed waits for the second and returns the milliseconds between them. */
int l3d_mark(l3d_ctx* ctx, int which);
int l3d_marks_elapsed(l3d_ctx* ctx, float* ms);
/* Overwrites a 256 MiB scratch buffer so that nothing of the previous step is left in L2. */
int l3d_flush_l2(l3d_ctx* ctx);
/* FP32 FFMA throughput of the device in TFLOP/s (register-only FMA chains, CUDA events):
 * the measured denominator for the FP32-pipe roofline. */
int l3d_measure_fp32_peak(l3d_ctx* ctx, float* tflops);
/* The same with packed FFMA2 instructions (two FP32 FMAs per issue slot on sm_100). */
int l3d_measure_fp32x2_peak(l3d_ctx* ctx, float* tflops);
/* The device's float exp (which = 0) / acos (which = 1) as used inside the kernels, on an
 * array: they are written to agree bit for bit with glibc's expf / acosf, which the
 * reference's host build calls; tests compare them against libm. */
int l3d_debug_math(l3d_ctx* ctx, int which, int n, const float* in, float* out);
/* Host seconds this context has spent since the last call of this function [0] putting groups of views on the GPU,
 * [1] waiting for their results, [2] collecting them; [3] = number of groups.  For finding out whether the host or
 * the GPU limits the per-view chain.  [4] waiting for the helper thread's pair tests of the next batch, [5] pair tests
 * run in the foreground (first batch), [6] the per-batch totals over the ranks (several GPUs); [7] unused. */
int l3d_debug_host_times(l3d_ctx* ctx, double out[8]);
/* With on != 0 and l3d_set_stage_timing, a CUDA event follows every kernel of the per-view chain; out (may be NULL)
 * receives the device ms since the last call, summed over the groups: [1] gather_counts [2] scan [3] fill [4] bucket
 * [5] score [6] reduce [7] filter [8] push [9] signal + waiting for the other ranks' flags [10] hist [11] scan [12] scatter [13] finish [14] gather. */
int l3d_debug_chain_times(l3d_ctx* ctx, int on, double out[24]);
/* Number of kernels this library has launched on the context since creation. */
uint64_t l3d_kernel_launches(const l3d_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* L3D_B200_H_ */
