// line3d_b200/csrc/line3D.h -- the Line3D public C++ API, unchanged in signature
// (reference: line3D.h:58-101, commons.h:42-66, 81-99, 215-238), implemented from scratch
// on top of the C ABI in include/l3d_b200.h.  main_bundler.cpp and main_vsfm.cpp of the
// reference compile against this header as they are.
//
// What is behind it: neighbour selection, scene normalisation, the per-view packing and
// reverse-match bookkeeping, candidate enumeration, graph clustering and 3-D line fitting
// stay on the host (fresh code in line3D.cc); epipolar matching, triangulation, scoring,
// hypothesis unprojection, pairwise affinity, diffusion and collinearity run on the GPU.
// There is no CPU path for those: without a CUDA device construction fails loudly.
#ifndef L3D_B200_LINE3D_H_
#define L3D_B200_LINE3D_H_

#include <list>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include <vector_types.h>   // float4 (CUDA toolkit)

#include "opencv/cv.h"
#include "eigen3/Eigen/Eigen"
#include "boost/filesystem.hpp"

// defaults, commons.h:42-66
#define L3D_DEF_MAX_IMG_WIDTH 1920
#define L3D_DEF_MIN_LINE_LENGTH_F 0.005f
#define L3D_DEF_MAX_NUM_SEGMENTS 3000
#define L3D_DEF_LOAD_AND_STORE_SEGMENTS true
#define L3D_DEF_COLLINEARITY_S 2.0f
#define L3D_DEF_COLLINEARITY_FOR_CLUSTERING true
#define L3D_DEF_MATCHING_NEIGHBORS 10
#define L3D_DEF_UNCERTAINTY_UPPER_T 5.0f
#define L3D_DEF_UNCERTAINTY_LOWER_T 1.0f
#define L3D_DEF_MIN_BASELINE_T 0.25f
#define L3D_DEF_SIGMA_P 3.5f
#define L3D_DEF_SIGMA_A 10.0f
#define L3D_DEF_PERFORM_RDD false
#define L3D_MIN_AFFINITY 0.25f
#define L3D_EPS 1e-12

namespace L3D
{
    // 2D segment id (sortable), commons.h:81-99
    class L3DSegment2D
    {
    public:
        L3DSegment2D() : camID_(0), segID_(0) {}
        L3DSegment2D(unsigned int camID, unsigned int segID) : camID_(camID), segID_(segID) {}
        inline unsigned int camID() const { return camID_; }
        inline unsigned int segID() const { return segID_; }
        inline bool operator==(const L3DSegment2D& rhs) const { return camID_ == rhs.camID_ && segID_ == rhs.segID_; }
        inline bool operator<(const L3DSegment2D& rhs) const { return camID_ < rhs.camID_ || (camID_ == rhs.camID_ && segID_ < rhs.segID_); }
        inline bool operator!=(const L3DSegment2D& rhs) const { return !((*this) == rhs); }
    private:
        unsigned int camID_, segID_;
    };

    // clustered 3D line, commons.h:215-238
    class L3DFinalLine3D
    {
    public:
        L3DFinalLine3D(std::list<L3D::L3DSegment2D> segments2D,
                       std::list<std::pair<Eigen::Vector3d, Eigen::Vector3d> > segments3D)
        {
            segments3D_ = segments3D;
            segments2D_ = segments2D;
        }
        std::list<std::pair<Eigen::Vector3d, Eigen::Vector3d> >* segments3D() { return &segments3D_; }
        std::list<L3D::L3DSegment2D>* segments2D() { return &segments2D_; }
    private:
        std::list<std::pair<Eigen::Vector3d, Eigen::Vector3d> > segments3D_;
        std::list<L3D::L3DSegment2D> segments2D_;
    };

    // Counters and timers of the last compute3Dmodel() (extension; not in the reference).
    struct L3DRunStats
    {
        unsigned long long pair_tests, candidates, raw_matches, scored_matches, filtered_matches, score_iters, score_iters_other_cam;
        unsigned long long hypotheses, affinity_pairs, edges, kernel_launches;
        double t_neighbors, t_upload, t_matching, t_matching_device, t_hypotheses, t_affinity, t_diffusion, t_clustering, t_lines, t_total;
        double t_stage[5];   // device seconds: pre-test, exact test, assembly, scoring, reduce/filter/deliver
        unsigned long long h2d_bytes, d2h_bytes;
        double t_k_pretest, t_k_exact;   // device seconds of the level-1 pre-test and exact-test kernels alone
        double t_k_pretest2;             // ... and of the level-2 pre-test kernel
        double rdd_step_ms, rdd_step_bytes, rdd_entries;   // diffusion: one step kernel's duration, its algorithmic bytes, the matrix entries
        double t_matching_calls;   // part of t_matching spent inside l3d_match_view (the rest is the host's bookkeeping between the calls)
    };

    class Line3D
    {
    public:
        // line3D.h:61-66 -- note: upper bound before lower bound
        Line3D(const std::string data_directory, const int matchingNeighbors = L3D_DEF_MATCHING_NEIGHBORS,
               const float uncertainty_t_upper_2D = L3D_DEF_UNCERTAINTY_UPPER_T,
               const float uncertainty_t_lower_2D = L3D_DEF_UNCERTAINTY_LOWER_T,
               const float sigma_p = L3D_DEF_SIGMA_P, const float sigma_a = L3D_DEF_SIGMA_A,
               const float min_baseline = L3D_DEF_MIN_BASELINE_T,
               bool useCollinearity = L3D_DEF_COLLINEARITY_FOR_CLUSTERING, bool verbose = false);
        ~Line3D();

        // line3D.h:70-74
        void addImage(const unsigned int imageID, const cv::Mat image,
                      const Eigen::Matrix3d K, const Eigen::Matrix3d R,
                      const Eigen::Vector3d t, std::list<unsigned int>& worldpointIDs,
                      const int maxImgWidth = L3D_DEF_MAX_IMG_WIDTH,
                      const bool loadAndStoreSegments = L3D_DEF_LOAD_AND_STORE_SEGMENTS);
        // line3D.h:76-80
        void addImage_fixed_sim(const unsigned int imageID, const cv::Mat image,
                                const Eigen::Matrix3d K, const Eigen::Matrix3d R,
                                const Eigen::Vector3d t, std::map<unsigned int, float>& viewSimilarity,
                                const int maxImgWidth = L3D_DEF_MAX_IMG_WIDTH,
                                const bool loadAndStoreSegments = L3D_DEF_LOAD_AND_STORE_SEGMENTS);

        void compute3Dmodel(bool perform_diffusion = L3D_DEF_PERFORM_RDD);          // line3D.h:83
        void getResult(std::list<L3D::L3DFinalLine3D>& result);                    // line3D.h:86
        float4 getSegment2D(L3D::L3DSegment2D& seg2D);                             // line3D.h:89
        void save3DLinesAsSTL(std::list<L3D::L3DFinalLine3D>& result, std::string filename);   // line3D.h:92
        void save3DLinesAsTXT(std::list<L3D::L3DFinalLine3D>& result, std::string filename);   // line3D.h:95
        unsigned int numCameras();                                                  // line3D.h:98
        void reset();                                                               // line3D.h:101

        // ---- extensions (not in the reference) -------------------------------------------
        // Segment cache in this library's own documented format: lets an
        // external detector feed segments through the reference's cache path
        // (line3D.cc:143-168).  Returns false if the file cannot be written.
        static bool writeSegmentCache(const std::string& data_directory, unsigned int imageID,
                                      unsigned int width, unsigned int height, bool collinearity_flag,
                                      const std::vector<float4>& segments);
        // Intermediate state of the last compute3Dmodel() as .npy files (test harness).
        void debugDump(const std::string& directory);
        L3DRunStats runStats() const;
        // run matching without copying per-view results to the host (measurement only;
        // the later stages are skipped)
        void setMeasureOnly(bool on);
        // flush the GPU's L2 before the matching phase of every compute3Dmodel() (measurement only)
        void setFlushL2(bool on);
        // Put every camera back to the pose it was added with.  compute3Dmodel() normalises the scene in place
        // (line3D.cc:364) and, as in the reference, a second call would normalise the normalised scene again (with
        // min_baseline then meaning something else); call this in between to repeat a run on identical input.
        void resetGeometry();
        // Multi-GPU: process only this rank's share of the views.  The neighbour graph's connected components are
        // independent of each other, so they are dealt out whole (largest first, to the least loaded rank); all ranks
        // must have added the same images.  rank 0 / world 1 = everything.
        void setShard(int rank, int world);
        // Multi-GPU, one connected scene: the `world` processes (one per GPU) have added the same images and make the
        // same calls; every view's segments are dealt out to the ranks in turn and the ranks exchange what later views
        // need over NCCL (include/l3d_b200.h, l3d_comm_init).  The matches are gathered on rank 0, which alone clusters
        // them: getResult() is empty on the other ranks.  nccl_id: L3D_COMM_ID_BYTES bytes from l3d_comm_unique_id on
        // rank 0, distributed by the caller.  world = 1 switches it off.
        void setSplit(int rank, int world, const unsigned char* nccl_id);
        // run matching with the records copied to the host, but stop there (measurement of the matching stage alone)
        void setMatchOnly(bool on);
        // keep the per-match records debugDump() writes (off by default: they are large)
        void setKeepDebug(bool on);
        void setStageTiming(bool on);   // per-stage device times in runStats().t_stage (slows matching by about 13 %)

    private:
        struct Impl;
        Impl* impl_;
        Line3D(const Line3D&);
        Line3D& operator=(const Line3D&);
    };
}

#endif  // L3D_B200_LINE3D_H_
