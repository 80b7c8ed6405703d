// line3d_b200/csrc/l3d_kernels.cu -- sm_100a kernels and C ABI (include/l3d_b200.h) of the
// Line3D matching-and-scoring path.  Written from scratch for B200; the reference's
// cudawrapper.cu is cited only for the semantics each stage must reproduce.
//
// Everything stays on the device.  Pair tests (phase A) run per batch of views, the rest (phase B) per group of
// views of one dependency level (see VArgs); DESIGN.md section 4 has the scheduling.
//
//   hoist      per pair-tested (view, neighbour), per segment: epipolar lines of the endpoints (FP64 -> FP32)
//   pre-test   every (src, tgt) pair: conservative tests on the four epipolar residuals
//              q_m^T F p_n  -> 1 candidate bit per pair                   [FP32 issue bound]
//   exact      candidates only: the reference's arithmetic, bit for bit   (l3d_device.cuh)
//   assemble   count -> scan -> fill: the view's match list in the reference's canonical
//              (src, cam, tgt) order, new raw matches merged with the reversed records
//              that earlier views delivered -- no sort, no host round trip
//   score      matches of a segment are bucketed by depth; every match looks for supporters in the
//              buckets its depth window touches: a 2-compare depth gate that is a superset of the
//              reference's 3-D gate, exact evaluation of the survivors through a per-warp queue
//   reduce     best match per segment, median depth, filter (score > 1), and the
//              reversed hand-over lists for later views, regrouped on the device
//
// Tensor cores are not used: nothing here is a dense contraction.
#include "../../include/l3d_b200.h"
#include "l3d_device.cuh"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

using namespace l3d;

// ------------------------------------------------------------------------------------------
// small utilities
// ------------------------------------------------------------------------------------------
namespace {

std::string g_create_error;

// NCCL is only needed when several GPUs work on one scene (l3d_comm_init): it is opened at run time, so that the
// single-GPU path has no dependency on it (and a process that already holds a libnccl -- torch ships its own -- shares it).
struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclCommAbort) CommAbort = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};
NcclApi* nccl_api(std::string* err)
{
    static NcclApi api; static bool tried = false; static std::string why;
    if (!tried) {
        tried = true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (int k = 0; k < 2 && !api.lib; ++k) api.lib = dlopen(names[k], RTLD_NOW | RTLD_GLOBAL);
        if (!api.lib) why = std::string("cannot open libnccl.so.2: ") + dlerror();
        else {
#define L3D_NCCL_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(api.lib, "nccl" #name)); if (!api.name) why = "libnccl has no nccl" #name;
            L3D_NCCL_SYM(GetUniqueId) L3D_NCCL_SYM(CommInitRank) L3D_NCCL_SYM(CommDestroy) L3D_NCCL_SYM(CommAbort) L3D_NCCL_SYM(AllGather) L3D_NCCL_SYM(AllReduce)
            L3D_NCCL_SYM(Broadcast) L3D_NCCL_SYM(GroupStart) L3D_NCCL_SYM(GroupEnd) L3D_NCCL_SYM(GetErrorString)
#undef L3D_NCCL_SYM
        }
    }
    if (!why.empty()) { if (err) *err = why; return nullptr; }
    return &api;
}
#define L3D_NCCL(call)                                                                              \
    do {                                                                                            \
        ncclResult_t r_ = (call);                                                                   \
        if (r_ != ncclSuccess) {                                                                    \
            char b_[512];                                                                           \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, ctx->nccl->GetErrorString(r_), __FILE__, __LINE__); \
            ctx->err = b_;                                                                          \
            return 1;                                                                               \
        }                                                                                           \
    } while (0)

#define L3D_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            char b_[512];                                                                           \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            ctx->err = b_;                                                                          \
            return 1;                                                                               \
        }                                                                                           \
    } while (0)

template<class T>
struct DBuf {                 // grow-only device buffer
    T* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        size_t want = std::max(n, cap + cap / 2);
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
template<class T>
struct HBuf {                 // grow-only pinned host buffer
    T* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        size_t want = std::max(n, cap + cap / 2);
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMallocHost(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// per (source view, neighbour) table on the device
struct CamTable {
    float F[9];
    float RtKinv[9];
    float P[12];
    float C[3];
    int seg_off;        // first segment of the neighbour in the global segment array
    int n_seg;
    int mode;           // 0 = pair-test now, 1 = consume delivered records, 2 = nothing
    int slot;           // index among the pair-tested neighbours (mode 0) or -1
    int row_off;        // mode 0: first global row of this (view, neighbour) job in the phase-A arrays
    int deliver;        // delivery slot (>= 0) or -1
    const int* t_start; // mode 1: delivered records grouped by this view's segment
    const int* t_tgt;
    const float4* t_dep;
};

struct SrcTable {       // the source view
    float RtKinv[9];
    float C[3];
    float spatial_k;
    float cmax;         // max |C_k|
    int seg_off, n_seg;
    int row_first, row_stride;   // row r of this process is segment row_first + r * row_stride of the view (one GPU: 0, 1;
                                 // several GPUs on one scene: rank, number of ranks -- the view's segments are dealt out in turn)
};
__device__ __forceinline__ int row_segment(const SrcTable& s, int row) { return s.row_first + row * s.row_stride; }

// ------------------------------------------------------------------------------------------
// exclusive scan (int), three phases, any length up to 4096 * 4096
// ------------------------------------------------------------------------------------------
constexpr int SCAN_T = 1024, SCAN_ITEMS = 4, SCAN_TILE = SCAN_T * SCAN_ITEMS;

__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* sh /* 32 ints */)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) sh[wid] = x;
    __syncthreads();
    if (wid == 0) {
        int s = (lane < (blockDim.x >> 5)) ? sh[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
        sh[lane] = s;
    }
    __syncthreads();
    int base = wid ? sh[wid - 1] : 0;
    if (total) *total = sh[(blockDim.x >> 5) - 1];
    __syncthreads();
    return base + x - v;
}

// Programmatic dependent launch: every kernel of the per-view chain starts with L3D_PDL_PROLOGUE and is launched with
// pdl_launch, so the blocks of kernel N+1 are set up on the SMs while the last blocks of kernel N still run and start
// the moment N has finished (griddepcontrol.wait returns when the preceding grid is complete and its writes are
// visible).  No kernel touches global memory before the prologue; launched the ordinary way the two instructions
// do nothing.
#define L3D_PDL_PROLOGUE() do { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); } while (0)

__global__ void __launch_bounds__(SCAN_T) k_scan_tiles(const int* __restrict__ in, int* __restrict__ out, int* __restrict__ tile_sums, int n)
{
    L3D_PDL_PROLOGUE();
    __shared__ int sh[32];
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
    int total;
    int ex = block_exclusive_scan(s, &total, sh);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { if (base + k < n) out[base + k] = ex; ex += v[k]; }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
// scans the tile sums in one block (<= 4096 tiles) and writes the grand total to out_total
__global__ void __launch_bounds__(SCAN_T) k_scan_sums(int* __restrict__ tile_sums, int n_tiles, int* __restrict__ out_total)
{
    L3D_PDL_PROLOGUE();
    __shared__ int sh[32];
    const int base = threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { v[k] = (base + k < n_tiles) ? tile_sums[base + k] : 0; s += v[k]; }
    int total;
    int ex = block_exclusive_scan(s, &total, sh);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { if (base + k < n_tiles) tile_sums[base + k] = ex; ex += v[k]; }
    if (threadIdx.x == 0) *out_total = total;
}
// Whole scan in one block for inputs up to SCAN_SMALL elements (all per-view scans): one launch instead of three.
// Every warp owns one contiguous chunk: pass 1 sums it (independent coalesced loads), one shuffle scan over the 32
// warp totals gives each warp its base, pass 2 re-reads the chunk (L1/L2) and scans it 32 elements at a time with
// a running carry.  One block-wide barrier in total; the tiled version this replaces had three per 4096 elements.
constexpr int SCAN_SMALL = 64 * 1024;
__device__ __forceinline__ void block_scan_global(const int* __restrict__ in, int* __restrict__ out /* n + 1 */, int n)      // blocks of SCAN_T threads
{
    __shared__ int s_tot[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int chunk = (((n + 31) >> 5) + 31) & ~31;              // elements per warp, a multiple of 32
    const int begin = w * chunk, end = min(n, begin + chunk);
    int s = 0;
#pragma unroll 4
    for (int i = begin + lane; i < end; i += 32) s += in[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) s_tot[w] = s;
    __syncthreads();
    int t = s_tot[lane], inc = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
    int carry = __shfl_sync(0xffffffffu, inc - t, w);            // exclusive prefix of this warp's total
    const int grand = __shfl_sync(0xffffffffu, inc, 31);
    if (threadIdx.x == 0) out[n] = grand;
    for (int i0 = begin; i0 < end; i0 += 32) {
        const int i = i0 + lane;
        const int v = (i < end) ? in[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (i < end) out[i] = carry + x - v;
        carry += __shfl_sync(0xffffffffu, x, 31);
    }
}
__global__ void __launch_bounds__(SCAN_T) k_scan_small(const int* __restrict__ in, int* __restrict__ out /* n + 1 */, int n)
{
    L3D_PDL_PROLOGUE();
    block_scan_global(in, out, n);
}
__global__ void __launch_bounds__(SCAN_T) k_scan_add(int* __restrict__ out, const int* __restrict__ tile_sums, int n, const int* __restrict__ total)
{
    L3D_PDL_PROLOGUE();
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    const int add = tile_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) if (base + k < n) out[base + k] += add;
    if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = *total;       // out has n + 1 entries
}

// ------------------------------------------------------------------------------------------
// hoist: per (pair-tested neighbour, segment) epipolar data for the pre-test
//   src side:  A_n = F p_n (n = 1,2) and w'_n = |(A_n)_xy|
//   tgt side:  q1, q2 and w_m = |(F^T q_m)_xy|
// evaluated in FP64 from the FP32 tables, rounded once.
// ------------------------------------------------------------------------------------------
struct SrcHoist { float4 a; float4 b; };   // a = (A1x, A1y, A1z, W1)   b = (A2x, A2y, A2z, W2),  A_n = F p_n / rho,  W_n = 2 eps |(F p_n)_xy|
struct TgtHoist { float4 q; float2 w; float2 pad; };   // w_m = 2 eps kappa |(F^T q_m)_xy|
struct SrcHoist1 { float4 p; float4 m; };  // level 1: p = (A_1 + A_2) / rho, m = (A_1 - A_2) / rho  (sum and difference of the residuals at the two endpoints)

// one pair-tested (source view, neighbour) combination; phase A runs over an array of these
struct JobDev {
    float F[9], RtKinv_s[9], RtKinv_t[9], C_s[3], C_t[3];
    int src_off, n_src, tgt_off, n_tgt;      // segment ranges in the global segment array (n_src: rows of this job)
    int src_first, src_stride;               // row r of the job is segment src_off + src_first + r * src_stride of the view
    int sh_off, th_off;                      // first entry in the hoist arrays
    long long cand_off;                      // first word in the candidate bitmask
    int wpr;                                 // words per row
    int row_off;                             // first global row (= source segment of this job) in the per-row arrays
    int blk_off, col_blocks;                 // level-2 blocks of this job: blk_off + rb * col_blocks + cb
    long long l1_off;                        // first word of the job's level-1 bits: [row group][target column]
    int l1_blk_off;                          // level-1 blocks of this job start here
};
__device__ __forceinline__ int job_src_segment(const JobDev& jb, int row) { return jb.src_off + jb.src_first + row * jb.src_stride; }

__device__ __forceinline__ int find_job_by_block(const JobDev* __restrict__ jobs, int n_jobs, int blk)
{
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (jobs[mid].blk_off <= blk) lo = mid; else hi = mid - 1; }
    return lo;
}

#define L3D_PRETEST_EPS_PX 0.25
#define L3D_PRETEST_LO 0.09f
#define L3D_PRETEST_HI 0.28f
#define L3D_PRETEST_MAXABS 1e6f     // beyond this the float tests could overflow: the pre-test is switched off for the segment

// How far (in units of the 0.25 px allowance) the reference's own FP32 arithmetic can move the supporting line of the
// segment (x1, y1)-(x2, y2) or an intersection with it, relative to coordinates around 2 kpx and long segments, for
// which the allowance was validated:
//   * cross(p1, p2).z = x1 y2 - y1 x2 is a difference of two rounded products (cudawrapper.cu:546-547): its error of up
//     to 1.5 ulp(|x y|) moves the line by that much divided by the segment's length -- 0.06 px for a 12 px segment at
//     2 kpx, but 1.5 px for a 3 px one, and 16 times more at 8 kpx;
//   * the intersections are cross products of such lines: their error grows with the square of the coordinates.
// The factor is >= 1; margins are multiplied by it (source side: the residuals are divided by it instead).
__device__ __forceinline__ double pretest_noise_factor(double x1, double y1, double x2, double y2)
{
    const double c = fmax(fmax(fabs(x1), fabs(y1)), fmax(fabs(x2), fabs(y2)));
    const double g = fmax(1.0, (c / 2048.0) * (c / 2048.0));
    const double m = fmax(fmax(fabs(x1 * y2), fabs(y1 * x2)), 1.0);
    const double ulp = scalbn(1.0, ilogb(m) - 23);
    const double len = fmax(hypot(x1 - x2, y1 - y2), 1e-3);
    return g * (1.0 + (3.0 * ulp / len) / L3D_PRETEST_EPS_PX);
}

// mode 0: normal; 1: the pre-test is switched off for every segment (all pairs go to the exact test; l3d_debug_set_pretest)
__global__ void k_hoist(const float4* __restrict__ segs, const JobDev* __restrict__ jobs, SrcHoist* __restrict__ sh, SrcHoist1* __restrict__ sh1, TgtHoist* __restrict__ th, int* __restrict__ row_job, int mode)
{
    const JobDev& jb = jobs[blockIdx.y];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= jb.n_src && idx >= jb.n_tgt) return;
    // The tests are homogeneous in F: it is scaled by a power of two (exact) so that its largest entry lies in [1, 2)
    // and squares of residuals stay far from the ends of the float range.
    double F[9], fmaxabs = 0.0;
#pragma unroll
    for (int k = 0; k < 9; ++k) { F[k] = double(jb.F[k]); fmaxabs = fmax(fmaxabs, fabs(F[k])); }
    const bool f_ok = fmaxabs > 0.0 && fmaxabs < 1e300;      // (a NaN fails both comparisons)
    if (f_ok) {
        const double sc = scalbn(1.0, -ilogb(fmaxabs));
#pragma unroll
        for (int k = 0; k < 9; ++k) F[k] *= sc;
    }
    const float INF = __int_as_float(0x7f800000);
    if (idx < jb.n_src) {
        float4 s = segs[job_src_segment(jb, idx)];
        double p[2][2] = {{double(s.x), double(s.y)}, {double(s.z), double(s.w)}};
        const double rho = pretest_noise_factor(p[0][0], p[0][1], p[1][0], p[1][1]);
        float o[2][4]; double A[2][3];
        bool ok = f_ok && mode == 0;
#pragma unroll
        for (int n = 0; n < 2; ++n) {
            double ax = F[0] * p[n][0] + F[1] * p[n][1] + F[2];
            double ay = F[3] * p[n][0] + F[4] * p[n][1] + F[5];
            double az = F[6] * p[n][0] + F[7] * p[n][1] + F[8];
            A[n][0] = ax / rho; A[n][1] = ay / rho; A[n][2] = az / rho;
            o[n][0] = float(A[n][0]); o[n][1] = float(A[n][1]); o[n][2] = float(A[n][2]); o[n][3] = float(2.0 * L3D_PRETEST_EPS_PX * sqrt(ax * ax + ay * ay));
            ok = ok && fabsf(o[n][0]) < L3D_PRETEST_MAXABS && fabsf(o[n][1]) < L3D_PRETEST_MAXABS && fabsf(o[n][2]) < L3D_PRETEST_MAXABS && o[n][3] < L3D_PRETEST_MAXABS;
        }
        SrcHoist1 h1;
        h1.p = make_float4(float(A[0][0] + A[1][0]), float(A[0][1] + A[1][1]), float(A[0][2] + A[1][2]), 0.f);
        h1.m = make_float4(float(A[0][0] - A[1][0]), float(A[0][1] - A[1][1]), float(A[0][2] - A[1][2]), 0.f);
        // Values the float tests cannot handle switch the pre-test off for this segment: residuals of +1 at one endpoint
        // and -1 at the other ("the epipolar line crosses the segment") and infinite margins keep every pair at both levels.
        if (!ok) {
            o[0][0] = o[0][1] = o[1][0] = o[1][1] = 0.f; o[0][2] = 1.f; o[1][2] = -1.f; o[0][3] = o[1][3] = INF;
            h1.p = make_float4(0.f, 0.f, 0.f, 0.f); h1.m = make_float4(0.f, 0.f, 2.f, 0.f);      // sum 0, difference 2
        }
        SrcHoist h; h.a = make_float4(o[0][0], o[0][1], o[0][2], o[0][3]); h.b = make_float4(o[1][0], o[1][1], o[1][2], o[1][3]);
        sh[jb.sh_off + idx] = h; sh1[jb.sh_off + idx] = h1;
        row_job[jb.row_off + idx] = blockIdx.y;          // row -> job, so that later kernels need no search
    }
    if (idx < jb.n_tgt) {
        float4 t = segs[jb.tgt_off + idx];
        double q[2][2] = {{double(t.x), double(t.y)}, {double(t.z), double(t.w)}};
        const double kappa = pretest_noise_factor(q[0][0], q[0][1], q[1][0], q[1][1]);
        float w[2];
        bool ok = f_ok && mode == 0 && fabsf(t.x) < L3D_PRETEST_MAXABS && fabsf(t.y) < L3D_PRETEST_MAXABS && fabsf(t.z) < L3D_PRETEST_MAXABS && fabsf(t.w) < L3D_PRETEST_MAXABS;
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            double ex = F[0] * q[m][0] + F[3] * q[m][1] + F[6];
            double ey = F[1] * q[m][0] + F[4] * q[m][1] + F[7];
            w[m] = float(2.0 * L3D_PRETEST_EPS_PX * kappa * sqrt(ex * ex + ey * ey));
            ok = ok && w[m] > 1e-12f && w[m] < L3D_PRETEST_MAXABS;      // (w / 2)^2 must be a normal float
        }
        if (!ok) { t = make_float4(0.f, 0.f, 0.f, 0.f); w[0] = w[1] = INF; }
        TgtHoist h; h.q = t; h.w = make_float2(w[0], w[1]); h.pad = make_float2(0.f, 0.f);
        th[jb.th_off + idx] = h;
    }
}

// ------------------------------------------------------------------------------------------
// pre-test.  For the pair (src s = p1p2, tgt t = q1q2) let E[m][n] = q_m^T F p_n.  The
// intersection of the src line with the epipolar line of q_m sits at parameter
//     tau_m = E[m][1] / (E[m][1] - E[m][2])        (0 at p1, 1 at p2)
// and the intersection of the tgt line with the epipolar line of p_n at
//     tau'_n = E[1][n] / (E[1][n] - E[2][n]).
// The reference's overlap value (cudawrapper.cu:209-252) is the 1-D intersection over
// union of [0,1] with [tau_1,tau_2] (resp. tau'), or 0.  A pair is dropped here only if,
// with every tau moved by an error bound worth L3D_PRETEST_EPS_PX pixels of intersection
// noise, an overlap still cannot reach the reference's thresholds minus a margin; all
// others go to the exact test.  The result is a superset of the reference's accepted set.
// ------------------------------------------------------------------------------------------

__device__ __forceinline__ float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ void side_bounds(float e1a, float e1b, float e2a, float e2b, float w1, float w2, float& I_hi, float& U_lo)
{
    // endpoint m: tau = ema / (ema - emb)
    float d1 = e1a - e1b, d2 = e2a - e2b;
    // approximate reciprocals (one MUFU each, 1 ulp): the thresholds below keep a 10 % margin; a zero or denormal
    // denominator gives inf, then NaN/inf bounds, and those comparisons keep the pair
    float r1 = fast_rcp(d1), r2 = fast_rcp(d2);
    float t1 = e1a * r1, t2 = e2a * r2;
    float err = 0.5f * (w1 * fabsf(r1) * (1.0f + 2.0f * fabsf(t1)) + w2 * fabsf(r2) * (1.0f + 2.0f * fabsf(t2)));
    float lo = fminf(t1, t2), hi = fmaxf(t1, t2);
    I_hi = fminf(hi, 1.0f) - fmaxf(lo, 0.0f) + err;
    U_lo = fmaxf(hi, 1.0f) - fminf(lo, 0.0f) - err;
}

// Packed FP32 pairs (sm_100: FFMA2 / FADD2 work on 64-bit register pairs, one issue slot for two operations; an operand
// given as pk2(x, x) is encoded as a scalar that is broadcast to both halves).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void up2(f32x2 v, unsigned& lo, unsigned& hi) { asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// Two levels, two kernels.  Level 1 (k_pretest) is a sign test on the four residuals: the src segment lies entirely on
// one side of the epipolar line of q_m when E[m][1] and E[m][2] have the same sign; if that holds for both target
// endpoints, by more than the allowance (L3D_PRETEST_EPS_PX pixels), and both intersections fall beyond the same end of
// the segment (|E[m][1]| - |E[m][2]| has the same sign for m = 1, 2), the overlap is zero whatever the rounding.  About
// 96 % of all pairs end here.  It works on the sum and the difference of the two residuals of a line,
//     S_m = E[m][1] + E[m][2] = q_m . (A_1 + A_2),      D_m = E[m][1] - E[m][2] = q_m . (A_1 - A_2),
// which the hoisted source vectors give directly (packed over m: 2 FFMA2 each).  |S_m| - |D_m| is twice the smaller
// magnitude when the residuals share a sign and negative when they do not, so
//     x_m = |S_m| - |D_m| - w_m                   sign set: not on one side by more than the allowance
// and, when they share a sign, |E[m][1]| - |E[m][2]| = sign(S_m) D_m, so the sign of (S_1 D_1)(S_2 D_2) tells whether
// the two intersections lie beyond different ends.  The verdict is the sign bit of x_1 | x_2 | (S_1 D_1)(S_2 D_2):
// 4 FFMA2 + 2 FADD + 1 FADD2 + 1 FMUL2 + 1 FMUL on the FMA pipe (15 of its cycles per warp and 32 pairs), one LOP3 and one
// funnel shift on the ALU pipe (2 cycles each, and on this part they are cycles the FMA pipe loses), no branch.
// Only the source side is tested: on the frozen scenes it rejects 95.8 % of all pairs and the mirrored test on the
// target side would add 0.4 % for as many instructions again -- level 2 covers both sides anyway.
// Its output is one bit per pair, 32 source rows of one target column per word (what a lane holds after 32 rows),
// stored [row group][column]: a warp's store is one coalesced line.
// Level 2 (k_pretest2) reads those words, compacts the ~4 % survivors through a per-warp queue, applies the interval
// test on full warps and writes the candidate bitmask [row][column / 32] and the candidates per row.
constexpr int P1_COLS = 256;        // target segments per block: 4 warps x 32 lanes x 2 columns per lane
constexpr int P1_BLOCK = 128;
constexpr int P1_CHUNK = 128;       // source rows staged in shared memory at a time
constexpr int P1_ROWS = 512;        // source rows per block

__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// QX = (q_1.x, q_2.x), QY = (q_1.y, q_2.y), nW = (-w_1, -w_2); Ap, Am: the hoisted sum and difference vectors of the row
__device__ __forceinline__ unsigned pretest_level1(f32x2 QX, f32x2 QY, f32x2 nW, const float4 Ap, const float4 Am)
{
    const f32x2 S = fma2(QX, pk2(Ap.x, Ap.x), fma2(QY, pk2(Ap.y, Ap.y), pk2(Ap.z, Ap.z)));      // (S_1, S_2)
    const f32x2 D = fma2(QX, pk2(Am.x, Am.x), fma2(QY, pk2(Am.y, Am.y), pk2(Am.z, Am.z)));      // (D_1, D_2)
    unsigned sl, sh_, dl, dh;
    up2(S, sl, sh_); up2(D, dl, dh);
    const float y1 = fabsf(__uint_as_float(sl)) - fabsf(__uint_as_float(dl)), y2 = fabsf(__uint_as_float(sh_)) - fabsf(__uint_as_float(dh));
    const f32x2 X = add2(pk2(y1, y2), nW);
    const f32x2 P = mul2(S, D);
    unsigned xl, xh, pl, ph;
    up2(X, xl, xh); up2(P, pl, ph);
    const float pp = __uint_as_float(pl) * __uint_as_float(ph);
    return xl | xh | __float_as_uint(pp);      // sign bit clear: surely no overlap on the src side
}

template <int UNROLL, int MINB>
__global__ void __launch_bounds__(P1_BLOCK, MINB) k_pretest(const SrcHoist1* __restrict__ sh, const TgtHoist* __restrict__ th,
                                                            const JobDev* __restrict__ jobs, int n_jobs, unsigned* __restrict__ l1, int rows_per_block)
{
    __shared__ float4 s_a[P1_CHUNK], s_b[P1_CHUNK];
    __shared__ int s_job;
    if (threadIdx.x == 0) {      // the job this block belongs to: blocks are numbered job by job (l1 blocks: l1_blk_off)
        int lo = 0, hi = n_jobs - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (jobs[mid].l1_blk_off <= int(blockIdx.x)) lo = mid; else hi = mid - 1; }
        s_job = lo;
    }
    __syncthreads();
    const JobDev& jb = jobs[s_job];
    const int local = blockIdx.x - jb.l1_blk_off;
    const int col_blocks = (jb.n_tgt + P1_COLS - 1) / P1_COLS;
    const int rb = local / col_blocks, cb = local - rb * col_blocks;
    const int row_begin = rb * rows_per_block, row_end = min(jb.n_src, row_begin + rows_per_block);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = cb * P1_COLS + warp * 64 + lane, c1 = c0 + 32;
    f32x2 QX0 = 0, QY0 = 0, NT0 = 0, QX1 = 0, QY1 = 0, NT1 = 0;
    if (c0 < jb.n_tgt) { const TgtHoist h = th[jb.th_off + c0]; QX0 = pk2(h.q.x, h.q.z); QY0 = pk2(h.q.y, h.q.w); NT0 = pk2(-h.w.x, -h.w.y); }
    if (c1 < jb.n_tgt) { const TgtHoist h = th[jb.th_off + c1]; QX1 = pk2(h.q.x, h.q.z); QY1 = pk2(h.q.y, h.q.w); NT1 = pk2(-h.w.x, -h.w.y); }
    unsigned* out = l1 + jb.l1_off;
    // the rows of the next chunk travel through registers while the current chunk is computed
    float4 na = make_float4(0.f, 0.f, 0.f, 0.f), nb = na;
    if (row_begin + tid < row_end) { const SrcHoist1 h = sh[jb.sh_off + row_begin + tid]; na = h.p; nb = h.m; }
    for (int r0 = row_begin; r0 < row_end; r0 += P1_CHUNK) {
        __syncthreads();                                   // the previous chunk has been read by every warp
        s_a[tid] = na; s_b[tid] = nb;
        __syncthreads();
        na = make_float4(0.f, 0.f, 0.f, 0.f); nb = na;
        if (r0 + P1_CHUNK + tid < row_end) { const SrcHoist1 h = sh[jb.sh_off + r0 + P1_CHUNK + tid]; na = h.p; nb = h.m; }
        const int groups = min(P1_CHUNK, row_end - r0 + 31) >> 5;
        for (int g = 0; g < groups; ++g) {
            unsigned acc0 = 0u, acc1 = 0u;
#pragma unroll UNROLL
            for (int rr = 0; rr < 32; ++rr) {
                const float4 A1 = s_a[g * 32 + rr], A2 = s_b[g * 32 + rr];          // broadcast
                acc0 = __funnelshift_l(pretest_level1(QX0, QY0, NT0, A1, A2), acc0, 1);      // row rr of the group ends at bit 31 - rr
                acc1 = __funnelshift_l(pretest_level1(QX1, QY1, NT1, A1, A2), acc1, 1);
            }
            const int left = row_end - (r0 + g * 32);      // rows beyond the job's last (zero hoists) are masked out
            if (left < 32) { const unsigned m = ~((1u << (32 - left)) - 1u); acc0 &= m; acc1 &= m; }
            const size_t word = size_t((r0 >> 5) + g) * size_t(jb.n_tgt);
            if (c0 < jb.n_tgt) out[word + c0] = acc0;
            if (c1 < jb.n_tgt) out[word + c1] = acc1;
        }
    }
}

// level 2 for one pair: false = an overlap surely misses a threshold
__device__ __forceinline__ bool pretest_level2(const float4 A1, const float4 A2, const float4 qq, const float2 ww)
{
    const float e11 = fmaf(qq.x, A1.x, fmaf(qq.y, A1.y, A1.z));
    const float e12 = fmaf(qq.x, A2.x, fmaf(qq.y, A2.y, A2.z));
    const float e21 = fmaf(qq.z, A1.x, fmaf(qq.w, A1.y, A1.z));
    const float e22 = fmaf(qq.z, A2.x, fmaf(qq.w, A2.y, A2.z));
    float Ia, Ua, Ib, Ub;
    side_bounds(e11, e12, e21, e22, ww.x, ww.y, Ia, Ua);      // along the src segment
    side_bounds(e11, e21, e12, e22, A1.w, A2.w, Ib, Ub);      // along the tgt segment
    // keep unless an overlap surely misses a threshold; NaN/inf comparisons keep
    const bool drop = (Ia < L3D_PRETEST_LO * Ua) || (Ib < L3D_PRETEST_LO * Ub) ||
                      ((Ia < L3D_PRETEST_HI * Ua) && (Ib < L3D_PRETEST_HI * Ub));
    return !drop;
}

constexpr int PT_ROWS = 128;        // level 2: source segments per block tile (4 words of level-1 bits per lane)
constexpr int PT_COLS = 256;        // target segments per block tile: 8 warps x 32 lanes
constexpr int PT_BLOCK = 256;
constexpr int PT_Q = 320;           // per-warp queue.  A word of 32 rows leaves about 43 survivors in a warp; a word
                                    // that would not fit (degenerate geometry) is evaluated without the queue

__global__ void __launch_bounds__(PT_BLOCK) k_pretest2(const SrcHoist* __restrict__ sh, const TgtHoist* __restrict__ th,
                                                       const JobDev* __restrict__ jobs, int n_jobs, const unsigned* __restrict__ l1,
                                                       unsigned* __restrict__ cand, int* __restrict__ cand_cnt)
{
    __shared__ float4 s_a[PT_ROWS], s_b[PT_ROWS];
    __shared__ float4 s_q[PT_COLS];
    __shared__ float2 s_w[PT_COLS];
    __shared__ unsigned s_bits[PT_ROWS][PT_COLS / 32];
    __shared__ unsigned short s_queue[PT_BLOCK / 32][PT_Q];
    __shared__ int s_job;
    if (threadIdx.x == 0) s_job = find_job_by_block(jobs, n_jobs, blockIdx.x);
    __syncthreads();
    const JobDev& jb = jobs[s_job];
    const int local = blockIdx.x - jb.blk_off;
    const int rb = local / jb.col_blocks, cb = local - rb * jb.col_blocks;
    const int col0 = cb * PT_COLS, row0 = rb * PT_ROWS;
    const int ncol = min(PT_COLS, jb.n_tgt - col0), nrow = min(PT_ROWS, jb.n_src - row0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the level-1 words of this lane's column: row groups row0 / 32 .. of the job
    unsigned words[PT_ROWS / 32];
    {
        const unsigned* in = l1 + jb.l1_off + size_t(row0 >> 5) * size_t(jb.n_tgt) + col0 + tid;
#pragma unroll
        for (int g = 0; g < PT_ROWS / 32; ++g) words[g] = (tid < ncol && g * 32 < nrow) ? in[size_t(g) * size_t(jb.n_tgt)] : 0u;
    }
    {
        if (tid < ncol) { TgtHoist h = th[jb.th_off + col0 + tid]; s_q[tid] = h.q; s_w[tid] = h.w; }
        else { s_q[tid] = make_float4(0.f, 0.f, 0.f, 0.f); s_w[tid] = make_float2(0.f, 0.f); }
        if (tid < PT_ROWS) {
            if (tid < nrow) { SrcHoist h = sh[jb.sh_off + row0 + tid]; s_a[tid] = h.a; s_b[tid] = h.b; }
            else { s_a[tid] = make_float4(0.f, 0.f, 0.f, 0.f); s_b[tid] = make_float4(0.f, 0.f, 0.f, 0.f); }
        }
        for (int k = tid; k < PT_ROWS * (PT_COLS / 32); k += PT_BLOCK) (&s_bits[0][0])[k] = 0u;
    }
    __syncthreads();
    unsigned short* queue = s_queue[warp];
    int qn = 0;
    auto level2 = [&](int first, int count) {
        if (lane < count) {
            const unsigned e = queue[first + lane];
            const int r = int(e >> 5), c = warp * 32 + int(e & 31u);
            if (pretest_level2(s_a[r], s_b[r], s_q[c], s_w[c])) atomicOr(&s_bits[r][warp], 1u << (e & 31u));
        }
    };
#pragma unroll
    for (int g = 0; g < PT_ROWS / 32; ++g) {
        unsigned acc = words[g];                             // bit 31 - rr = row g * 32 + rr of the tile
        const int n = __popc(acc);
        int pre = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
        const int total = __shfl_sync(0xffffffffu, pre, 31);
        if (total == 0) continue;
        if (qn + total <= PT_Q) {
            pre += qn - n;
            while (acc) { const int bit = 31 - __clz(acc); acc &= ~(1u << bit); queue[pre++] = (unsigned short)(((g * 32 + 31 - bit) << 5) | lane); }
            qn += total;
            __syncwarp();
            while (qn >= 32) { level2(qn - 32, 32); qn -= 32; }
            __syncwarp();
        } else {            // more survivors than the queue holds: every lane walks its own bits
            while (acc) {
                const int bit = 31 - __clz(acc); acc &= ~(1u << bit);
                const int r = g * 32 + 31 - bit;
                if (pretest_level2(s_a[r], s_b[r], s_q[tid], s_w[tid])) atomicOr(&s_bits[r][warp], 1u << lane);
            }
        }
    }
    if (qn > 0) level2(0, qn);
    __syncthreads();
    // write the tile's bitmask rows and count the candidates per row
    const int wpt = PT_COLS / 32;                        // 8 words per row
    for (int k = tid; k < nrow * wpt; k += PT_BLOCK) {
        const int r = k / wpt, w = k - r * wpt;
        if (col0 + w * 32 < jb.n_tgt) cand[jb.cand_off + (long long)(row0 + r) * jb.wpr + (col0 >> 5) + w] = s_bits[r][w];
    }
    if (tid < nrow) {
        int total = 0;
#pragma unroll
        for (int w = 0; w < wpt; ++w) total += __popc(s_bits[tid][w]);
        if (total) atomicAdd(&cand_cnt[jb.row_off + row0 + tid], total);
    }
}

// ------------------------------------------------------------------------------------------
// exact test.  Candidates of a row rarely fill whole warps and most of them fail the overlap test, so a
// row-per-warp kernel (the first version, see profiles/r1_v1_kernels.md) idled more than half of its lanes.
// Here every candidate is its own thread:
//   k_expand      warp per row: the row's candidate columns, in order, into a flat list (+ the row of each)
//   k_exact_flat  thread per candidate: overlap test; the block compacts the survivors through shared memory
//                 and its first threads triangulate them (idle warps cost no issue slots); results are written
//                 at the candidate's own index: ok flag, 4 depths; raw matches per row are counted with atomics
// k_fill then copies the ok candidates of a row, in order, to their final position.
// ------------------------------------------------------------------------------------------
__global__ void k_expand(const JobDev* __restrict__ jobs, const int* __restrict__ row_job, int n_rows, const unsigned* __restrict__ cand, const int* __restrict__ cand_off,
                         int* __restrict__ c_col, int* __restrict__ c_row)
{
    const int grow = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (grow >= n_rows) return;
    int o = cand_off[grow];
    if (cand_off[grow + 1] == o) return;
    const JobDev& jb = jobs[row_job[grow]];
    const unsigned* words = cand + jb.cand_off + (long long)(grow - jb.row_off) * jb.wpr;
    const int n_words = (jb.n_tgt + 31) >> 5;
    for (int w0 = 0; w0 < n_words; w0 += 32) {
        unsigned bits = (w0 + lane < n_words) ? words[w0 + lane] : 0u;
        int n = __popc(bits), ex = n;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, ex, s); if (lane >= s) ex += y; }
        const int tot = __shfl_sync(0xffffffffu, ex, 31);
        int pos = o + ex - n;
        while (bits) { int b = __ffs(bits) - 1; bits &= bits - 1; c_col[pos] = (w0 + lane) * 32 + b; c_row[pos] = grow; ++pos; }
        o += tot;
    }
}

constexpr int EF_T = 256;
__global__ void __launch_bounds__(EF_T, 6) k_exact_flat(const float4* __restrict__ segs, const JobDev* __restrict__ jobs, const int* __restrict__ row_job, int n_cand,
                                                     const int* __restrict__ c_col, const int* __restrict__ c_row,
                                                     unsigned char* __restrict__ c_ok, float4* __restrict__ c_dep, int* __restrict__ raw_cnt)
{
    __shared__ float s_is[8][EF_T];
    __shared__ int s_who[EF_T];               // candidate (block-local, low 8 bits) and job of each overlap survivor
    __shared__ int s_warp_n[EF_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x * EF_T + tid;
    bool ok1 = false; PairIsect is; int job = 0;
    if (c < n_cand) {
        const int grow = c_row[c];
        job = row_job[grow];
        const JobDev& jb = jobs[job];
        ok1 = xpair_overlap(segs[job_src_segment(jb, grow - jb.row_off)], segs[jb.tgt_off + c_col[c]], jb.F, is);
        c_ok[c] = 0;
    }
    // block-wide ordered compaction of the survivors
    const unsigned m = __ballot_sync(0xffffffffu, ok1);
    if (lane == 0) s_warp_n[warp] = __popc(m);
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < EF_T / 32; ++w) { const int n = s_warp_n[w]; if (w < warp) base += n; total += n; }
    if (ok1) {
        const int p = base + __popc(m & ((1u << lane) - 1u));
        s_who[p] = tid | (job << 8);
        s_is[0][p] = is.ax; s_is[1][p] = is.ay; s_is[2][p] = is.bx; s_is[3][p] = is.by;
        s_is[4][p] = is.cx; s_is[5][p] = is.cy; s_is[6][p] = is.dx; s_is[7][p] = is.dy;
    }
    __syncthreads();
    if (tid < total) {
        const int cc = blockIdx.x * EF_T + (s_who[tid] & (EF_T - 1));
        const int grow = c_row[cc];
        const JobDev& jb = jobs[s_who[tid] >> 8];
        PairIsect i2;
        i2.ax = s_is[0][tid]; i2.ay = s_is[1][tid]; i2.bx = s_is[2][tid]; i2.by = s_is[3][tid];
        i2.cx = s_is[4][tid]; i2.cy = s_is[5][tid]; i2.dx = s_is[6][tid]; i2.dy = s_is[7][tid];
        float4 dep;
        const bool ok = xpair_depths(segs[job_src_segment(jb, grow - jb.row_off)], segs[jb.tgt_off + c_col[cc]], i2, jb.RtKinv_s, jb.RtKinv_t,
                                     mk3(jb.C_s[0], jb.C_s[1], jb.C_s[2]), mk3(jb.C_t[0], jb.C_t[1], jb.C_t[2]), dep);
        if (ok) { c_ok[cc] = 1; c_dep[cc] = dep; atomicAdd(&raw_cnt[grow], 1); }
    }
}

// out_sum[b], out_max[b] over v[begin[b] .. begin[b] + len[b]); one block per b.  After the exact test: raw matches
// per job and the most raw matches of a single row, which bound every per-view buffer and grid.
__global__ void k_range_sum_max(const int* __restrict__ v, const int* __restrict__ begin, const int* __restrict__ len, int* __restrict__ out_sum, int* __restrict__ out_max)
{
    __shared__ int sh_s[32], sh_m[32];
    const int b0 = begin[blockIdx.x], n = len[blockIdx.x];
    int s = 0, m = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { const int x = v[b0 + i]; s += x; m = max(m, x); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); m = max(m, __shfl_xor_sync(0xffffffffu, m, o)); }
    if ((threadIdx.x & 31) == 0) { sh_s[threadIdx.x >> 5] = s; sh_m[threadIdx.x >> 5] = m; }
    __syncthreads();
    if (threadIdx.x < 32) {
        s = (threadIdx.x < (blockDim.x >> 5)) ? sh_s[threadIdx.x] : 0; m = (threadIdx.x < (blockDim.x >> 5)) ? sh_m[threadIdx.x] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); m = max(m, __shfl_xor_sync(0xffffffffu, m, o)); }
        if (threadIdx.x == 0) { out_sum[blockIdx.x] = s; out_max[blockIdx.x] = m; }
    }
}

// out[b] = max over i in [begin[b], begin[b] + len[b]) of (diff ? v[i + 1] - v[i] : v[i]); one block per b.
// Used for upper bounds on the number of matches of a single segment (which scoring kernels are needed).
__global__ void k_range_max(const int* __restrict__ v, const int* __restrict__ begin, const int* __restrict__ len, int diff, int* __restrict__ out)
{
    __shared__ int sh[32];
    const int b0 = begin[blockIdx.x], n = len[blockIdx.x];
    int m = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) m = max(m, diff ? v[b0 + i + 1] - v[b0 + i] : v[b0 + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (threadIdx.x == 0) out[blockIdx.x] = m;
    }
}

// out[b] = sum over v[begin[b] .. begin[b] + len[b]) in 64 bits (the 32-bit scan of the same values may have wrapped)
__global__ void k_range_sum64(const int* __restrict__ v, const int* __restrict__ begin, const int* __restrict__ len, unsigned long long* __restrict__ out)
{
    __shared__ unsigned long long sh[32];
    const int b0 = begin[blockIdx.x], n = len[blockIdx.x];
    unsigned long long s = 0ull;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += (unsigned long long)(unsigned)v[b0 + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) out[blockIdx.x] = s;
    }
}

// dst[k] = src[idx[k]]
__global__ void k_gather_int(const int* __restrict__ src, const int* __restrict__ idx, int n, int* __restrict__ dst)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) dst[k] = src[idx[k]];
}

// counts of the delivered records per (row, cam) for the cams in mode 1
__device__ __forceinline__ void b_gather_counts(const CamTable* __restrict__ cams, int n_cams, int n_src, const SrcTable& src, const int* __restrict__ raw_cnt, int* __restrict__ cnt)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    const int cam = blockIdx.y;
    if (row >= n_src) return;
    const CamTable& ct = cams[cam];
    int c = 0;
    if (ct.mode == 0) c = raw_cnt[ct.row_off + row];
    else if (ct.mode == 1) { const int sg = row_segment(src, row); c = ct.t_start ? (ct.t_start[sg + 1] - ct.t_start[sg]) : 0; }
    cnt[row * n_cams + cam] = c;
}

// fill: one warp per (row, cam) copies the row's records to their final place
__device__ __forceinline__ void b_fill(const CamTable* __restrict__ cams, int n_cams, int n_src, const SrcTable& src,
                       const int* __restrict__ off, const int* __restrict__ cand_off,
                       const int* __restrict__ stage_col, const float4* __restrict__ stage_dep, const unsigned char* __restrict__ stage_ok,
                       int4* __restrict__ m_ids, float4* __restrict__ m_dep)
{
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (gw >= n_src * n_cams) return;
    const int row = gw / n_cams, cam = gw - row * n_cams;
    const int o = off[gw], n = off[gw + 1] - o;
    if (n == 0) return;
    const CamTable& ct = cams[cam];
    const int sg = row_segment(src, row);            // match ids carry the segment (x) and this process's row (w)
    if (ct.mode == 0) {
        const int b = cand_off[ct.row_off + row], e = cand_off[ct.row_off + row + 1];
        int w = o;
        for (int k0 = b; k0 < e; k0 += 32) {
            const int k = k0 + lane;
            const bool ok = k < e && stage_ok[k] != 0;
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (ok) { const int r = w + __popc(m & ((1u << lane) - 1u)); m_ids[r] = make_int4(sg, cam, stage_col[k], row); m_dep[r] = stage_dep[k]; }
            w += __popc(m);
        }
    } else if (ct.mode == 1) {
        // delivered records of one segment arrive in the order of the sender's atomics; they go out in ascending
        // target id (ids are unique within a group and groups are tiny: each record counts the smaller ones)
        const int b = ct.t_start[sg];
        for (int k = lane; k < n; k += 32) {
            const int me = ct.t_tgt[b + k];
            int rank = 0;
            for (int i = 0; i < n; ++i) rank += (ct.t_tgt[b + i] < me);
            m_ids[o + rank] = make_int4(sg, cam, me, row); m_dep[o + rank] = ct.t_dep[b + k];
        }
    }
}

// ------------------------------------------------------------------------------------------
// scoring (cudawrapper.cu:614-714).  Block = (chunk of SC_T matches of one source segment);
// every thread owns one match y and walks the segment's other matches camera group by
// camera group.  The reference's 3-D gate  |P-Q| > spatial_k |C-P|  (cudawrapper.cu:388-401)
// compares two points on the same unit ray, so it can only pass when the depths differ by
// at most spatial_k * depth (+ rounding): that 2-compare test on (d1, d2) pairs staged in
// shared memory rejects ~97-99.8 % of the iterations; the rest are queued per warp and
// evaluated with the reference's exact arithmetic, 32 at a time.
// ------------------------------------------------------------------------------------------
constexpr int SC_T = 128;
constexpr int SC_WARPS = SC_T / 32;
constexpr int SC_TILE = 2048;      // matches staged per smem tile
constexpr int SC_QCAP = 128;       // per-warp queue of gate survivors

struct ScoreParams { float sigma_p, sigma_a; };

// work items: (source segment, chunk of SC_T matches); order is irrelevant to the result
__device__ __forceinline__ void b_score_items(const int* __restrict__ off, int n_cams, int n_src, int m_lo, int m_hi, int outside, int* __restrict__ counter, int2* __restrict__ items, int chunk)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n_src) return;
    int m = off[(row + 1) * n_cams] - off[row * n_cams];
    const bool inside = m > m_lo && m <= m_hi;
    if (inside == (outside != 0)) m = 0;     // this list takes the segments inside (m_lo, m_hi], or those outside it
    const int n = (m + chunk - 1) / chunk;
    if (n == 0) return;
    const int base = atomicAdd(counter, n);
    for (int k = 0; k < n; ++k) items[base + k] = make_int2(row, k);
}

// exact evaluation of queued (hypothesis y, other match i) pairs: cudawrapper.cu:662-705
__device__ __forceinline__ void score_drain(int count, const unsigned* q, int lane, const float2* s_yd, int seg_start,
                                            const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                            V3 ray1, V3 ray2, V3 C, float spatial_k, const CamTable* __restrict__ cams,
                                            const float4* __restrict__ segs, float sigma_p, float sigma_a, int* s_max, int stride)
{
    if (lane < count) {
        const unsigned e = q[lane];
        const int yl = int(e >> 24), idx = int(e & 0xffffffu);
        const float2 dy = s_yd[yl];
        const int4 id = m_ids[seg_start + idx];
        const float4 di = m_dep[seg_start + idx];
        const CamTable& ct = cams[id.y];
        V3 P1 = xadd3(C, xscale(dy.x, ray1)), P2 = xadd3(C, xscale(dy.y, ray2));          // cudawrapper.cu:644-645
        V3 Q1 = xadd3(C, xscale(di.x, ray1)), Q2 = xadd3(C, xscale(di.y, ray2));          // cudawrapper.cu:671-672
        V3 pr1 = xproject(ct.P, P1), pr2 = xproject(ct.P, P2);
        if (pr1.z == 1.0f && pr2.z == 1.0f) {
            float conf = xhypothesis_confidence(pr1, pr2, P1, P2, Q1, Q2, C, segs[ct.seg_off + id.z], sigma_p, sigma_a, spatial_k);
            if (conf > 0.5f) atomicMax(&s_max[id.y * stride + yl], __float_as_int(conf));
        }
    }
}

__device__ __forceinline__ void b_score(const float4* __restrict__ segs, const CamTable* __restrict__ cams, int n_cams, SrcTable src,
                                                const int* __restrict__ off, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                const int2* __restrict__ items, const int* __restrict__ n_items,
                                                ScoreParams sp, float* __restrict__ score)
{
    extern __shared__ int s_max[];                 // [n_cams][H] best confidence per (camera, hypothesis), as int bits; H = blockDim.x hypotheses per block
    const int H = blockDim.x;
    __shared__ float2 s_d[SC_TILE];
    __shared__ float2 s_yd[SC_T];
    __shared__ unsigned s_q[SC_WARPS][SC_QCAP];
    if (int(blockIdx.x) >= *n_items) return;
    const int2 item = items[blockIdx.x];
    const int row = item.x;
    const int seg_start = off[row * n_cams], seg_end = off[(row + 1) * n_cams];
    const int M = seg_end - seg_start;
    const int y0 = item.y * H;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y = y0 + tid;
    const bool live = y < M;
    int my_cam = -1; float2 my_d = make_float2(0.f, 0.f);
    if (live) { my_cam = m_ids[seg_start + y].y; float4 d = m_dep[seg_start + y]; my_d = make_float2(d.x, d.y); }
    s_yd[tid] = my_d;
    for (int c = 0; c < n_cams; ++c) s_max[c * H + tid] = 0;
    // per-segment invariants (cudawrapper.cu:638-645)
    const float4 sg = segs[src.seg_off + row_segment(src, row)];
    const V3 p1 = mk3(sg.x, sg.y, 1.0f), p2 = mk3(sg.z, sg.w, 1.0f);
    const V3 ray1 = xnormalize(xmat3(src.RtKinv, p1, false)), ray2 = xnormalize(xmat3(src.RtKinv, p2, false));
    const V3 C = mk3(src.C[0], src.C[1], src.C[2]);
    // superset gate:  |dy - di| <= thr  must hold for the exact gate to be able to pass
    const float k = src.spatial_k;
    const float thr1 = (k * fabsf(my_d.x) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.x) + src.cmax)) * 1.00001f;
    const float thr2 = (k * fabsf(my_d.y) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.y) + src.cmax)) * 1.00001f;
    const bool gate_on = k > 0.0f;
    unsigned* q = s_q[warp];
    int qn = 0;
    // this thread skips its own camera's range [skip0, skip1) (cudawrapper.cu:674-675)
    int skip0 = 0, skip1 = 0;
    if (live) { skip0 = off[row * n_cams + my_cam] - seg_start; skip1 = off[row * n_cams + my_cam + 1] - seg_start; }
    for (int t0 = 0; t0 < M; t0 += SC_TILE) {
        const int tn = min(SC_TILE, M - t0);
        __syncthreads();
        for (int i = tid; i < tn; i += H) { float4 d = m_dep[seg_start + t0 + i]; s_d[i] = make_float2(d.x, d.y); }
        __syncthreads();
        for (int i0 = 0; i0 < tn; i0 += 32) {
            unsigned hits = 0;
            if (live) {
                const int lim = min(32, tn - i0);
                if (gate_on) {
                    unsigned mk = 1u;
#pragma unroll 8
                    for (int b = 0; b < 32; ++b, mk <<= 1) {
                        const float2 d = s_d[i0 + b];           // beyond tn: stale but masked below
                        bool pass = (fabsf(my_d.x - d.x) <= thr1) && (fabsf(my_d.y - d.y) <= thr2);
                        hits |= pass ? mk : 0u;
                    }
                    if (lim < 32) hits &= (1u << lim) - 1u;
                } else {
                    hits = (lim >= 32) ? 0xffffffffu : ((1u << lim) - 1u);
                }
                // drop the entries of this thread's own camera
                const int g = t0 + i0;
                if (g < skip1 && g + 32 > skip0) {
                    const int lo = max(skip0 - g, 0), hi = min(skip1 - g, 32);
                    const unsigned m = ((hi >= 32) ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
                    hits &= ~m;
                }
            }
            while (__any_sync(0xffffffffu, hits != 0u)) {
                int n = __popc(hits), ex = n;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { int v = __shfl_up_sync(0xffffffffu, ex, o); if (lane >= o) ex += v; }
                ex -= n;
                int pos = qn + ex, pushed = 0;
                while (hits && pos < SC_QCAP) { int b = __ffs(hits) - 1; hits &= hits - 1; q[pos++] = (unsigned(tid) << 24) | unsigned(t0 + i0 + b); ++pushed; }
                int tot = pushed;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                qn += tot;
                __syncwarp();
                while (qn >= 32) {
                    score_drain(32, q + (qn - 32), lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, k, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
                    qn -= 32;
                    __syncwarp();
                }
            }
        }
    }
    if (qn > 0) score_drain(qn, q, lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, k, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
    __syncwarp();
    // sum of the per-camera best confidences in camera order (cudawrapper.cu:677-687, 709)
    float total = 0.0f;
    for (int c = 0; c < n_cams; ++c) total = xa(total, __int_as_float(s_max[c * H + tid]));
    if (live) score[seg_start + y] = total;
}

// ------------------------------------------------------------------------------------------
// scoring, sorted-window variant.  The superset gate of k_score is |d1_y - d1_i| <= thr1(y) and
// |d2_y - d2_i| <= thr2(y).  If the segment's matches are sorted by d1, the partners of y that can
// pass the first condition are one contiguous window, found by binary search: the quadratic walk
// over all (y, i) pairs becomes O(M log M) + the window contents.  The set of pairs handed to the
// exact evaluation is the same as in k_score (the same two fabsf tests are applied inside the
// window), so the scores are bit-identical; only the order of evaluation changes, and the result
// is a per-camera maximum, which does not depend on it.
// One CTA per source segment with lo < M <= MP matches; the segment is sorted once in shared
// memory (bitonic, key d1, payload d2 + camera + original index).
// ------------------------------------------------------------------------------------------
// Grouping the matches of a segment by depth, for the windowed scoring.  A hypothesis can only be supported by matches
// whose first depth lies within a small relative window of its own (the 3-D gate, see k_score_window), so each
// segment's matches are distributed over BK_B buckets by d1 -- a counting sort: three barriers per segment whatever
// its length, where a comparison sort of the longest segments took 66 barrier-separated steps and set the kernel's
// duration.  The bucket of a depth is linear in the bit pattern of the float between the segment's smallest and
// largest d1 (positive floats order like unsigned integers, and the pattern is a piecewise-linear log2, which suits
// a window that is relative): monotonic, so the buckets a window touches are the buckets of its two ends.  The
// order inside a bucket is whatever the atomics give; the scores do not depend on it (per-camera maxima).
// Output: (d1, d2, camera << 20 | index) at the segment's own offsets in bucket order, the bucket boundaries, and
// (bits of min d1, bits of max d1, scale) per segment.
constexpr int BK_T = 128;
constexpr int BK_B = 512;
__device__ __forceinline__ int depth_bucket(unsigned bits, unsigned bits_min, float scale)
{
    return min(BK_B - 1, int(float(bits - bits_min) * scale));
}

__device__ __forceinline__ void b_bucket_segments(const int* __restrict__ off, int n_cams, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                          float* __restrict__ g_k, float* __restrict__ g_d2, unsigned* __restrict__ g_pl,
                                                          int* __restrict__ g_boff /* [n_src][BK_B + 1] */, uint4* __restrict__ g_brange,
                                                          int m_hi, int chunk, int* __restrict__ n_items, int2* __restrict__ items)
{
    __shared__ int s_cnt[BK_B], s_start[BK_B + 1];
    __shared__ unsigned s_lo[BK_T / 32], s_hi[BK_T / 32];
    __shared__ int s_item0;
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg_start = off[row * n_cams], M = off[(row + 1) * n_cams] - seg_start;
    if (M <= 0 || M > m_hi) return;
    // the scoring kernel's work items of this segment: (segment, chunk of `chunk` hypotheses); their order is irrelevant
    const int n_chunks = (M + chunk - 1) / chunk;
    if (tid == 0) s_item0 = atomicAdd(n_items, n_chunks);
    unsigned lo = 0xffffffffu, hi = 0u;
    for (int i = tid; i < M; i += BK_T) { const unsigned b = __float_as_uint(m_dep[seg_start + i].x); lo = min(lo, b); hi = max(hi, b); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
    if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; }
    for (int b = tid; b < BK_B; b += BK_T) s_cnt[b] = 0;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < BK_T / 32; ++w) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
    const float scale = float(BK_B) / (float(hi - lo) + 1.0f);
    for (int i = tid; i < M; i += BK_T) atomicAdd(&s_cnt[depth_bucket(__float_as_uint(m_dep[seg_start + i].x), lo, scale)], 1);
    __syncthreads();
    if (warp == 0) {                               // exclusive scan of the BK_B counts: 4 per lane
        int c[BK_B / 32], sum = 0;
#pragma unroll
        for (int r = 0; r < BK_B / 32; ++r) { c[r] = s_cnt[lane * (BK_B / 32) + r]; sum += c[r]; }
        int ex = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, ex, o); if (lane >= o) ex += t; }
        ex -= sum;
#pragma unroll
        for (int r = 0; r < BK_B / 32; ++r) { s_start[lane * (BK_B / 32) + r] = ex; ex += c[r]; }
        if (lane == 31) s_start[BK_B] = ex;
    }
    __syncthreads();
    int* boff = g_boff + size_t(row) * (BK_B + 1);
    for (int b = tid; b < BK_B; b += BK_T) boff[b] = s_start[b];
    if (tid == 0) { boff[BK_B] = s_start[BK_B]; g_brange[row] = make_uint4(lo, hi, __float_as_uint(scale), 0u); }
    for (int c = tid; c < n_chunks; c += BK_T) items[s_item0 + c] = make_int2(row, c);
    for (int b = tid; b < BK_B; b += BK_T) s_cnt[b] = 0;      // now the fill cursors
    __syncthreads();
    for (int i = tid; i < M; i += BK_T) {
        const float4 d = m_dep[seg_start + i];
        const int bk = depth_bucket(__float_as_uint(d.x), lo, scale);
        const int pos = seg_start + s_start[bk] + atomicAdd(&s_cnt[bk], 1);
        g_k[pos] = d.x; g_d2[pos] = d.y;
        g_pl[pos] = (unsigned(m_ids[seg_start + i].y) << 20) | unsigned(i);
    }
}

// work items (segment, chunk of SC_T hypotheses) as in k_score; the partners of a hypothesis are looked for in the
// depth buckets its window touches
__device__ __forceinline__ void b_score_window(const float4* __restrict__ segs, const CamTable* __restrict__ cams, int n_cams, SrcTable src,
                                                       const int* __restrict__ off, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                       const float* __restrict__ g_k, const float* __restrict__ g_d2, const unsigned* __restrict__ g_pl,
                                                       const int* __restrict__ g_boff, const uint4* __restrict__ g_brange,
                                                       const int2* __restrict__ items, const int* __restrict__ n_items,
                                                       ScoreParams sp, float* __restrict__ score)
{
    extern __shared__ int s_max[];                 // [n_cams][H], H = blockDim.x hypotheses per block
    const int H = blockDim.x;
    __shared__ float2 s_yd[SC_T];
    __shared__ unsigned s_q[SC_WARPS][SC_QCAP];
    if (int(blockIdx.x) >= *n_items) return;
    const int2 item = items[blockIdx.x];
    const int row = item.x;
    const int seg_start = off[row * n_cams], M = off[(row + 1) * n_cams] - seg_start;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y = item.y * H + tid;
    const bool live = y < M;
    // thread y takes the y-th match in BUCKET order: the lanes of a warp then hold neighbouring depths, their windows
    // overlap (coalesced loads) and are about equally long (little divergence in the scan below)
    int my_cam = -1, my_idx = 0; float2 my_d = make_float2(0.f, 0.f);
    if (live) { const unsigned pl = g_pl[seg_start + y]; my_cam = int(pl >> 20); my_idx = int(pl & 0xfffffu); my_d = make_float2(g_k[seg_start + y], g_d2[seg_start + y]); }
    s_yd[tid] = my_d;
    for (int c = 0; c < n_cams; ++c) s_max[c * H + tid] = 0;
    const float4 sg = segs[src.seg_off + row_segment(src, row)];
    const V3 p1 = mk3(sg.x, sg.y, 1.0f), p2 = mk3(sg.z, sg.w, 1.0f);
    const V3 ray1 = xnormalize(xmat3(src.RtKinv, p1, false)), ray2 = xnormalize(xmat3(src.RtKinv, p2, false));
    const V3 C = mk3(src.C[0], src.C[1], src.C[2]);
    const float kk = src.spatial_k;
    const float thr1 = (kk * fabsf(my_d.x) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.x) + src.cmax)) * 1.00001f;
    const float thr2 = (kk * fabsf(my_d.y) * 1.00001f + 2e-6f * (2.0f * fabsf(my_d.y) + src.cmax)) * 1.00001f;
    const float* K = g_k + seg_start; const float* D2 = g_d2 + seg_start; const unsigned* PL = g_pl + seg_start;
    int j = 0, jend = 0;
    if (live) {        // slightly widened window; the exact fabsf tests follow
        const float lo = my_d.x - thr1 * 1.001f - 1e-30f, hi = my_d.x + thr1 * 1.001f + 1e-30f;
        const uint4 br = g_brange[row];
        const float dmin = __uint_as_float(br.x), dmax = __uint_as_float(br.y), scale = __uint_as_float(br.z);
        if (hi >= dmin && lo <= dmax) {
            const int* boff = g_boff + size_t(row) * (BK_B + 1);
            j = boff[depth_bucket(__float_as_uint(fmaxf(lo, dmin)), br.x, scale)];
            jend = boff[depth_bucket(__float_as_uint(fminf(hi, dmax)), br.x, scale) + 1];
        }
    }
    __syncwarp();
    unsigned* q = s_q[warp];
    int qn = 0;
    while (__any_sync(0xffffffffu, j < jend)) {
        bool hit = false; unsigned pl = 0;
        if (j < jend) {
            hit = (fabsf(my_d.x - K[j]) <= thr1) && (fabsf(my_d.y - D2[j]) <= thr2);
            if (hit) { pl = PL[j]; hit = int(pl >> 20) != my_cam; }
            ++j;
        }
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (m) {
            if (hit) q[qn + __popc(m & ((1u << lane) - 1u))] = (unsigned(tid) << 24) | (pl & 0xfffffu);
            qn += __popc(m);
            __syncwarp();
            if (qn >= 32) {
                score_drain(32, q + (qn - 32), lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, kk, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
                qn -= 32;
                __syncwarp();
            }
        }
    }
    if (qn > 0) score_drain(qn, q, lane, s_yd, seg_start, m_ids, m_dep, ray1, ray2, C, kk, cams, segs, sp.sigma_p, sp.sigma_a, s_max, H);
    __syncwarp();
    if (live) {        // sum of the per-camera best confidences in camera order (cudawrapper.cu:677-687, 709)
        float total = 0.0f;
        for (int c = 0; c < n_cams; ++c) total = xa(total, __int_as_float(s_max[c * H + tid]));
        score[seg_start + my_idx] = total;
    }
}

// ------------------------------------------------------------------------------------------
// per-segment reduction (cudawrapper.cu:1028-1064, 1090-1110): best match of every segment (first maximum) and
// the matches with score > 1.  One thread per match, all loads independent (a warp per segment walking its matches
// was a chain of dependent loads as long as the longest segment): the best match is a 64-bit atomicMax on
// (score bits, ~match index) -- scores are positive, so their bits order like the values, and among equal scores the
// smaller index wins as in the reference's strict '>' scan --, the kept matches are counted per CTA for the ordered
// compaction in k_filter_compact, and the two iteration counters are summed on the way.
// ------------------------------------------------------------------------------------------
constexpr int RM_T = 1024;
__device__ __forceinline__ void b_reduce_matches(const int* __restrict__ Mptr, const int* __restrict__ off, int n_cams, const float* __restrict__ score,
                                                         const int4* __restrict__ m_ids, unsigned long long* __restrict__ best /* [n_src], zeroed */,
                                                         int* __restrict__ cta_kept, unsigned long long* __restrict__ iters /* [0] all pairs, [1] same-camera pairs */)
{
    __shared__ int s_k[RM_T / 32];
    __shared__ unsigned long long s_a[RM_T / 32], s_b[RM_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = blockIdx.x * RM_T + tid, M = *Mptr;
    int keep = 0; unsigned long long all = 0ull, same = 0ull;
    if (k < M) {
        const float c = score[k]; const int4 id = m_ids[k];
        keep = c > 1.0f;
        const int* o = off + id.w * n_cams;
        all = (unsigned long long)(o[n_cams] - o[0]);                 // sum over the matches of their segment's size = sum of squares
        same = (unsigned long long)(o[id.y + 1] - o[id.y]);
        if (c > 0.0f) {
            const unsigned long long key = ((unsigned long long)__float_as_uint(c) << 32) | (unsigned long long)(0xffffffffu - unsigned(k));
            if (key > best[id.w]) atomicMax(&best[id.w], key);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { keep += __shfl_xor_sync(0xffffffffu, keep, o); all += __shfl_xor_sync(0xffffffffu, all, o); same += __shfl_xor_sync(0xffffffffu, same, o); }
    if (lane == 0) { s_k[warp] = keep; s_a[warp] = all; s_b[warp] = same; }
    __syncthreads();
    if (warp == 0) {
        keep = s_k[lane]; all = s_a[lane]; same = s_b[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { keep += __shfl_xor_sync(0xffffffffu, keep, o); all += __shfl_xor_sync(0xffffffffu, all, o); same += __shfl_xor_sync(0xffffffffu, same, o); }
        if (lane == 0) { cta_kept[blockIdx.x] = keep; if (all) atomicAdd(iters, all); if (same) atomicAdd(iters + 1, same); }
    }
}

// median of the finite samples: sorted[count / 2], -1 if none (cudawrapper.cu:1066-1076).
// Radix select in one block: the samples are positive floats or +inf (no sample), so their bit patterns order like
// unsigned integers; four passes of an 8-bit histogram over the samples that still match the prefix found so far
// narrow the wanted rank down to one value.
constexpr int MS_T = 1024;
__device__ float block_median_select(const float* __restrict__ dsamples, int n)
{
    __shared__ int s_hist[256];
    __shared__ int s_cnt[MS_T / 32];
    __shared__ unsigned s_prefix; __shared__ int s_rank;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int finite = 0;
    for (int i = tid; i < n; i += MS_T) finite += (dsamples[i] < INFINITY);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) finite += __shfl_xor_sync(0xffffffffu, finite, o);
    if (lane == 0) s_cnt[warp] = finite;
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    int count = 0;
#pragma unroll
    for (int w = 0; w < MS_T / 32; ++w) count += s_cnt[w];
    if (count == 0) return -1.0f;                   // (the same for every thread of the block)
    if (tid == 0) { s_prefix = 0u; s_rank = count / 2; }
    unsigned prefix = 0u, mask = 0u;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = tid; i < n; i += MS_T) {       // the samples are re-read every pass (16 KB, L1/L2)
            const float x = dsamples[i];
            const unsigned v = __float_as_uint(x);
            if (x < INFINITY && (v & mask) == prefix) atomicAdd(&s_hist[(v >> shift) & 255u], 1);
        }
        __syncthreads();
        if (warp == 0) {                            // the bin in which the cumulative count passes the wanted rank
            int c[8], sum = 0;
#pragma unroll
            for (int r = 0; r < 8; ++r) { c[r] = s_hist[lane * 8 + r]; sum += c[r]; }
            int inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
            const int rank = s_rank;
            int ex = inc - sum;
            if (ex <= rank && rank < inc) {
#pragma unroll
                for (int r = 0; r < 8; ++r) { if (rank < ex + c[r]) { s_prefix = prefix | (unsigned(lane * 8 + r) << shift); s_rank = rank - ex; break; } ex += c[r]; }
            }
        }
        __syncthreads();
        prefix = s_prefix; mask |= 255u << shift;
        if (tid < 256) s_hist[tid] = 0;
        __syncthreads();
    }
    return __uint_as_float(prefix);
}

// The last kernel of a view: the median depth, and all the small results in one buffer so that one copy brings
// them to the host: out[0] = matches, out[1] = output records, out[2] = median depth (float bits), out[4..7] = two
// 64-bit iteration counters, out[8 + d] = records handed to delivery slot d.  One block.
__device__ __forceinline__ void b_finish_view(float* __restrict__ dsamples, int n_samples, const unsigned long long* __restrict__ best, const float4* __restrict__ m_dep,
                                                      const int* __restrict__ M, const int* __restrict__ F,
                                                      const unsigned long long* __restrict__ iters, const int* __restrict__ start, int stride, int n_deliver,
                                                      int* __restrict__ out, bool samples_ready)
{
    // depth samples: both depths of every segment's best match if its score exceeds 0.5 (cudawrapper.cu:1090-1110)
    // (with several GPUs on one scene every rank's k_mr_push has written them into this rank's window already)
    for (int row = threadIdx.x; !samples_ready && 2 * row < n_samples; row += MS_T) {
        const unsigned long long key = best[row];
        float a = INFINITY, b = INFINITY;
        if (__uint_as_float(unsigned(key >> 32)) > 0.5f) { const float4 d = m_dep[0xffffffffu - unsigned(key & 0xffffffffull)]; a = d.x; b = d.y; }
        dsamples[2 * row] = a; dsamples[2 * row + 1] = b;
    }
    __syncthreads();
    const float median = (n_samples > 0) ? block_median_select(dsamples, n_samples) : 0.0f;
    if (threadIdx.x == 0) {
        out[0] = *M; out[1] = *F; out[2] = __float_as_int(median); out[3] = 0;
        const unsigned long long a = iters ? iters[0] : 0ull, b = iters ? iters[1] : 0ull;
        out[4] = int(a & 0xffffffffull); out[5] = int(a >> 32); out[6] = int(b & 0xffffffffull); out[7] = int(b >> 32);
    }
    for (int d = threadIdx.x; d < n_deliver; d += blockDim.x) out[8 + d] = start[(d + 1) * stride] - start[d * stride];
}

// filtered records (score > 1) in match order = the reference's order: an ordered compaction of the whole match list.
// CTA b keeps cta_kept[b] records (k_reduce_matches), so it starts at the sum of the counts before it; inside the CTA
// a ballot/prefix gives every kept match its place.  Also counts the records per hand-over group (destination
// view's segment), the first step of the regrouping below.
__device__ __forceinline__ void b_filter_compact(const int* __restrict__ Mptr, const float* __restrict__ score, const int4* __restrict__ m_ids, const float4* __restrict__ m_dep,
                                                         const int* __restrict__ cta_kept, l3d_match* __restrict__ out, int* __restrict__ Fout,
                                                         const CamTable* __restrict__ cams, int stride, int* __restrict__ hist /* or null */, int n_cta)
{
    __shared__ int s_w[RM_T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = blockIdx.x * RM_T + tid, M = *Mptr;
    int part = 0;
    for (int i = tid; i < int(blockIdx.x); i += RM_T) part += cta_kept[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_w[warp] = part;
    __syncthreads();
    int base = 0;
#pragma unroll
    for (int w = 0; w < RM_T / 32; ++w) base += s_w[w];
    __syncthreads();
    const float c = (k < M) ? score[k] : 0.0f;
    const bool keep = c > 1.0f;
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_w[warp] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < RM_T / 32; ++w) { const int n = s_w[w]; if (w < warp) before += n; total += n; }
    if (keep) {
        const int r = base + before + __popc(m & ((1u << lane) - 1u));
        const int4 id = m_ids[k]; const float4 d = m_dep[k];
        l3d_match rec; rec.src_seg = id.x; rec.cam = id.y; rec.tgt_seg = id.z; rec.confidence = xd(c, 2.0f);
        rec.depths[0] = d.x; rec.depths[1] = d.y; rec.depths[2] = d.z; rec.depths[3] = d.w;
        out[r] = rec;
        if (hist) { const int dl = cams[id.y].deliver; if (dl >= 0) atomicAdd(&hist[dl * stride + id.z], 1); }     // hand-over group sizes
    }
    if (int(blockIdx.x) == n_cta - 1 && tid == 0) *Fout = base + total;
}
// unscored variant (no neighbour to pair-test): all records, confidence 0
__device__ __forceinline__ void b_copy_unscored_at(const int4* __restrict__ m_ids, const float4* __restrict__ m_dep, const int* __restrict__ Mptr, l3d_match* __restrict__ out, int k)
{
    if (k >= *Mptr) return;
    int4 id = m_ids[k]; float4 d = m_dep[k];
    l3d_match rec; rec.src_seg = id.x; rec.cam = id.y; rec.tgt_seg = id.z; rec.confidence = 0.0f;
    rec.depths[0] = d.x; rec.depths[1] = d.y; rec.depths[2] = d.z; rec.depths[3] = d.w;
    out[k] = rec;
}

// ------------------------------------------------------------------------------------------
// hand-over (line3D.cc:838-858): the filtered records towards each neighbour that asked for
// delivery, reversed and regrouped by the neighbour's segment, target ids ascending
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void b_deliver_scatter(const l3d_match* __restrict__ f, const int* __restrict__ Fptr, const CamTable* __restrict__ cams, int stride,
                                  const int* __restrict__ start, int* __restrict__ cursor, int* __restrict__ t_tgt, float4* __restrict__ t_dep)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= *Fptr) return;
    const l3d_match m = f[k];
    const int d = cams[m.cam].deliver;
    if (d < 0) return;
    const int key = d * stride + m.tgt_seg;
    const int pos = start[key] + atomicAdd(&cursor[key], 1);
    t_tgt[pos] = m.src_seg;
    t_dep[pos] = make_float4(m.depths[2], m.depths[3], m.depths[0], m.depths[1]);     // line3D.cc:852-855
}
// ------------------------------------------------------------------------------------------
// The per-view kernels run on GROUPS of views.  Which view needs which other view first (the hand-over of filtered
// matches) is a dependency graph, usually a few independent chains (views along a street; C2: 200 views, longest
// chain 50), and the views of one dependency level are independent.  One launch of each kernel below covers all views
// of a level: blockIdx.y (z for k_gather_counts) picks the view, whose buffers and sizes come from a VArgs record, and
// the grid's x extent is the largest any view of the group needs.  Per level, not per view: the host's launch work
// (it was the bottleneck at ~95 us per view) and the GPU's chain of short dependent kernels.
// ------------------------------------------------------------------------------------------
struct VArgs {
    // tables of the view
    const CamTable* cams; int n_cams, n_src; SrcTable src; ScoreParams sp;
    // pair-test results of the view's batch
    const int* raw_cnt; const int* cand_off; const int* stage_col; const float4* stage_dep; const unsigned char* stage_ok;
    // match list
    int* cnt; int* off; int4* mids; float4* mdep; float* score;
    // scoring
    float* sk; float* sd2; unsigned* spl; int* boff; uint4* brange; int2* items; int max_items; int scored, windowed, quadratic, m_hi;
    // zeroed per view
    int* nitems; unsigned long long* iters; unsigned long long* best; int* hist; int* cursor;
    // reduce, filter
    int* cta_kept; int n_cta; int* Fout; l3d_match* filtered; float* dsamples;
    // hand-over
    int delivering, stride, n_keys, n_deliver; int* dl_start; int* dl_tgt; float4* dl_dep;
    int* vres;
    // several GPUs on one scene (l3d_comm_init): this process holds the rows row_first + r * row_stride of the view
    int mr, n_seg_full;
    int sc_h;          // hypotheses per block of the scoring kernels (their shared memory holds cameras x sc_h maxima)
};

constexpr int VG_MAX = 32;         // most views in one group (l3d_ctx::group_views is how many are used)
// What one rank reports about one view of a group (all-gathered): its output records, its matches, its iteration counts.
struct RankRec { int F, M; unsigned long long iters_all, iters_same; int routed; int pad; };      // routed: records sent to the rank that holds this table
constexpr int MR_MAX_RANKS = 16;
// The group's window of every rank, as mapped into this process (CUDA IPC): [rank] -> base of that rank's window.
// A window holds, for one group of views: the ranks' records of each view (RankRec[rank][VG_MAX]), the depth samples of
// every segment of each view, and the output records of each view -- one region per rank, sized by the bound all ranks
// share, so that a rank knows where its records go in everybody's window without knowing the other ranks' counts.
struct MrPeers { unsigned char* base[MR_MAX_RANKS]; };
struct MrLayout { long long samp_off[VG_MAX], rec_off[VG_MAX], full_off[VG_MAX], cap[VG_MAX]; };      // per view of the group: bytes, bytes, bytes, records per rank
constexpr size_t MR_COUNTS_BYTES = size_t(MR_MAX_RANKS) * VG_MAX * 32;      // the RankRec table at the start of a window ...
constexpr size_t MR_FLAGS_OFF = MR_COUNTS_BYTES;                       // ... then one arrival flag per sending rank (unsigned) ...
constexpr size_t MR_HEADER_BYTES = MR_COUNTS_BYTES + 256;             // ... and the views' regions
#define L3D_VIEW(dim) L3D_PDL_PROLOGUE(); const VArgs& va = all[blockIdx.dim]

__global__ void k_gather_counts(const VArgs* __restrict__ all)
{
    L3D_VIEW(z);
    if (int(blockIdx.y) >= va.n_cams) return;
    b_gather_counts(va.cams, va.n_cams, va.n_src, va.src, va.raw_cnt, va.cnt);
}
// which = 0: count matrix -> offsets; 1: hand-over group sizes -> group starts.  One block per view.
__global__ void __launch_bounds__(SCAN_T) k_scan_views(const VArgs* __restrict__ all, int which)
{
    L3D_VIEW(y);
    if (which == 0) block_scan_global(va.cnt, va.off, va.n_src * va.n_cams);
    else if (va.delivering) block_scan_global(va.hist, va.dl_start, va.n_keys);
}
__global__ void k_fill(const VArgs* __restrict__ all, const float4* __restrict__ segs)
{
    L3D_VIEW(y);
    (void)segs;
    b_fill(va.cams, va.n_cams, va.n_src, va.src, va.off, va.cand_off, va.stage_col, va.stage_dep, va.stage_ok, va.mids, va.mdep);
}
__global__ void __launch_bounds__(BK_T) k_bucket_segments(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.scored || !va.windowed || int(blockIdx.x) >= va.n_src) return;
    b_bucket_segments(va.off, va.n_cams, va.mids, va.mdep, va.sk, va.sd2, va.spl, va.boff, va.brange, va.m_hi, va.sc_h, va.nitems, va.items);
}
__global__ void __launch_bounds__(SC_T, 12) k_score_window(const VArgs* __restrict__ all, const float4* __restrict__ segs)
{
    L3D_VIEW(y);
    if (!va.scored || !va.windowed) return;
    b_score_window(segs, va.cams, va.n_cams, va.src, va.off, va.mids, va.mdep, va.sk, va.sd2, va.spl, va.boff, va.brange, va.items, va.nitems, va.sp, va.score);
}
// the quadratic path: views without the 3-D gate, and segments too long for the window kernel's index
__global__ void k_score_items(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.scored || !va.quadratic) return;
    b_score_items(va.off, va.n_cams, va.n_src, va.windowed ? 0 : 0x7fffffff, va.windowed ? va.m_hi : 0x7fffffff, 1, va.nitems + 1, va.items + va.max_items, va.sc_h);
}
__global__ void __launch_bounds__(SC_T) k_score(const VArgs* __restrict__ all, const float4* __restrict__ segs)
{
    L3D_VIEW(y);
    if (!va.scored || !va.quadratic) return;
    b_score(segs, va.cams, va.n_cams, va.src, va.off, va.mids, va.mdep, va.items + va.max_items, va.nitems + 1, va.sp, va.score);
}
__global__ void __launch_bounds__(RM_T) k_reduce_matches(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.scored || int(blockIdx.x) >= va.n_cta) return;
    b_reduce_matches(va.off + va.n_src * va.n_cams, va.off, va.n_cams, va.score, va.mids, va.best, va.cta_kept, va.iters);
}
__global__ void __launch_bounds__(RM_T) k_filter_compact(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (int(blockIdx.x) >= va.n_cta) return;
    if (va.scored) b_filter_compact(va.off + va.n_src * va.n_cams, va.score, va.mids, va.mdep, va.cta_kept, va.filtered, va.Fout, va.cams, va.stride, (va.delivering && !va.mr) ? va.hist : nullptr, va.n_cta);
    else {          // no neighbour was pair-tested (cudawrapper.cu:877-878): all records, unscored; RM_T / 256 chunks of 256 per block
        b_copy_unscored_at(va.mids, va.mdep, va.off + va.n_src * va.n_cams, va.filtered, blockIdx.x * RM_T + threadIdx.x);
    }
}
__global__ void k_deliver_scatter(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    if (!va.delivering) return;
    b_deliver_scatter(va.filtered, va.Fout, va.cams, va.stride, va.dl_start, va.cursor, va.dl_tgt, va.dl_dep);
}
__global__ void __launch_bounds__(MS_T) k_finish_view(const VArgs* __restrict__ all)
{
    L3D_VIEW(y);
    const int* Mptr = va.off + va.n_src * va.n_cams;
    b_finish_view(va.dsamples, va.scored ? 2 * (va.mr ? va.n_seg_full : va.n_src) : 0, va.best, va.mdep, Mptr, va.scored ? va.Fout : Mptr, va.scored ? va.iters : nullptr,
                  va.delivering ? va.dl_start : nullptr, va.stride, va.delivering ? va.n_deliver : 0, va.vres, va.mr != 0);
}

// ------------------------------------------------------------------------------------------
// Several GPUs on ONE scene (l3d_comm_init).  The segments of every view are dealt out to the ranks in turn (segment
// s belongs to rank s mod ranks), so each rank pair-tests, assembles and scores its rows of every view, and the views
// are still processed level by level.  What a later view needs from an earlier one -- the reversed records, grouped by
// ITS segments -- is put together from all ranks' output records of the group:
//   k_mr_push      after the view's records have been filtered: every record that a later view will consume goes to
//                  the rank that holds the target segment (peer store over NVLink into that rank's window, into this
//                  rank's region of the view, appended through a counter); the depth samples of the segments this rank
//                  owns go to every rank; and, if a rank collects the records for the host, all records go to it as well
//   k_mr_signal    this rank's counts of each view into every rank's window, then -- after a system-wide fence -- its
//                  arrival flag
//   k_mr_wait      spins until every rank's flag for this group has arrived (no collective call, no host)
//   k_mr_hist, k_scan_views, k_mr_deliver_scatter   the hand-over lists for the segments this rank holds, from the
//                  records all ranks routed to it
//   k_finish_view  the median, from all ranks' samples in the own window
// The regions of a window are fixed by bounds all ranks share, so nothing has to be known of the others to store into it.
// ------------------------------------------------------------------------------------------
__global__ void k_mr_push(const VArgs* __restrict__ all, MrLayout lay, MrPeers peers, int rank, int W, int root, int* __restrict__ routed /* [VG_MAX][MR_MAX_RANKS], zeroed */)
{
    L3D_PDL_PROLOGUE();
    const int v = blockIdx.y;
    const VArgs& va = all[v];
    const int* Mptr = va.off + va.n_src * va.n_cams;
    const int F = va.scored ? *va.Fout : *Mptr;
    const int lane = threadIdx.x & 31;
    // records a later view consumes: to the owner of the target segment
    if (va.delivering) {
        for (int k0 = blockIdx.x * blockDim.x; k0 < F; k0 += gridDim.x * blockDim.x) {
            const int k = k0 + threadIdx.x;
            int dest = -1; uint4 lo, hi;
            if (k < F) {
                const uint4* src = reinterpret_cast<const uint4*>(va.filtered + k);
                lo = src[0]; hi = src[1];
                const int cam = int(lo.y), tgt = int(lo.z);                 // l3d_match: src_seg, cam, tgt_seg, confidence | depths
                if (va.cams[cam].deliver >= 0) dest = tgt % W;
            }
            // one counter bump per destination and warp
            const unsigned active = __ballot_sync(0xffffffffu, dest >= 0);
            if (dest >= 0) {
                const unsigned peers_same = __match_any_sync(active, dest);
                const int leader = __ffs(peers_same) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&routed[v * MR_MAX_RANKS + dest], __popc(peers_same));
                base = __shfl_sync(peers_same, base, leader);
                const int pos = base + __popc(peers_same & ((1u << lane) - 1u));
                uint4* dst = reinterpret_cast<uint4*>(peers.base[dest] + lay.rec_off[v]) + size_t(2) * (size_t(rank) * size_t(lay.cap[v]) + size_t(pos));
                dst[0] = lo; dst[1] = hi;
            }
        }
    }
    // all records, in order, to the rank that hands them to its host
    for (int r = (root == -2 ? 0 : root); r >= 0 && r < W; r = (root == -2 ? r + 1 : W)) {      // root: one rank, -2: every rank, -1: nobody
        const uint4* src = reinterpret_cast<const uint4*>(va.filtered);
        uint4* dst = reinterpret_cast<uint4*>(peers.base[r] + lay.full_off[v]) + size_t(2) * size_t(rank) * size_t(lay.cap[v]);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * F; i += gridDim.x * blockDim.x) dst[i] = src[i];
    }
    // depth samples of the segments this rank owns: both depths of the segment's best match if its score exceeds 0.5,
    // infinity otherwise (cudawrapper.cu:1090-1110); every segment is written by exactly one rank, to every rank
    if (va.scored) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < va.n_src * W; i += gridDim.x * blockDim.x) {
            const int row = i / W, r = i - row * W;
            const unsigned long long key = va.best[row];
            float2 ab = make_float2(INFINITY, INFINITY);
            if (__uint_as_float(unsigned(key >> 32)) > 0.5f) { const float4 dp = va.mdep[0xffffffffu - unsigned(key & 0xffffffffull)]; ab = make_float2(dp.x, dp.y); }
            reinterpret_cast<float2*>(peers.base[r] + lay.samp_off[v])[row_segment(va.src, row)] = ab;
        }
    }
}
// one thread per (view, destination rank): this rank's records of the group's views, then the flags
__global__ void k_mr_signal(const VArgs* __restrict__ all, int n_views, MrPeers peers, int rank, int W, const int* __restrict__ routed, unsigned seq)
{
    L3D_PDL_PROLOGUE();
    const int t = threadIdx.x;
    if (t < n_views * W) {
        const int v = t / W, r = t - v * W;
        const VArgs& va = all[v];
        const int* Mptr = va.off + va.n_src * va.n_cams;
        RankRec rec; rec.M = *Mptr; rec.F = va.scored ? *va.Fout : *Mptr;
        rec.iters_all = va.scored ? va.iters[0] : 0ull; rec.iters_same = va.scored ? va.iters[1] : 0ull;
        rec.routed = routed[v * MR_MAX_RANKS + r]; rec.pad = 0;
        reinterpret_cast<RankRec*>(peers.base[r])[rank * VG_MAX + v] = rec;
    }
    __threadfence_system();                 // the push kernel's stores (complete: it has ended) and the table, before the flags
    __syncthreads();
    if (t < W) {
        __threadfence_system();
        volatile unsigned* flag = reinterpret_cast<volatile unsigned*>(peers.base[t] + MR_FLAGS_OFF) + rank;
        *flag = seq;
    }
}
// waits for the flags of all ranks; gives up after about ten seconds and says so in *fault (the host reports it)
__global__ void k_mr_wait(const unsigned char* __restrict__ window, int W, unsigned seq, int* __restrict__ fault)
{
    // The next kernel's blocks are let onto the SMs only when the flags have arrived (at the end of this kernel): while
    // this rank waits for the slowest one they would sit there idle, in the way of the next batch's pair tests.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int r = threadIdx.x;
    if (r < W) {
        const volatile unsigned* flag = reinterpret_cast<const volatile unsigned*>(window + MR_FLAGS_OFF) + r;
        const long long t0 = clock64();
        while (int(*flag - seq) < 0) {
            if (clock64() - t0 > 20000000000ll) { atomicExch(fault, 1 + r); break; }
            __nanosleep(200);
        }
    }
    __threadfence_system();
}
// sizes of the hand-over groups, from the records all ranks routed here
__global__ void k_mr_hist(const VArgs* __restrict__ all, const unsigned char* __restrict__ window, MrLayout lay, int W)
{
    L3D_PDL_PROLOGUE();
    const int v = blockIdx.y;
    const VArgs& va = all[v];
    if (!va.delivering) return;
    const RankRec* rec = reinterpret_cast<const RankRec*>(window);
    for (int r = 0; r < W; ++r) {
        const int n = rec[r * VG_MAX + v].routed;
        const l3d_match* f = reinterpret_cast<const l3d_match*>(window + lay.rec_off[v]) + size_t(r) * size_t(lay.cap[v]);
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
            const int d = va.cams[f[k].cam].deliver;
            if (d >= 0) atomicAdd(&va.hist[d * va.stride + f[k].tgt_seg], 1);
        }
    }
}
__global__ void k_mr_deliver_scatter(const VArgs* __restrict__ all, const unsigned char* __restrict__ window, MrLayout lay, int W)
{
    L3D_PDL_PROLOGUE();
    const int v = blockIdx.y;
    const VArgs& va = all[v];
    if (!va.delivering) return;
    const RankRec* rec = reinterpret_cast<const RankRec*>(window);
    for (int r = 0; r < W; ++r) {
        const int n = rec[r * VG_MAX + v].routed;
        const l3d_match* f = reinterpret_cast<const l3d_match*>(window + lay.rec_off[v]) + size_t(r) * size_t(lay.cap[v]);
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
            const l3d_match m = f[k];
            const int d = va.cams[m.cam].deliver;
            if (d < 0) continue;
            const int key = d * va.stride + m.tgt_seg;
            const int pos = va.dl_start[key] + atomicAdd(&va.cursor[key], 1);
            va.dl_tgt[pos] = m.src_seg;
            va.dl_dep[pos] = make_float4(m.depths[2], m.depths[3], m.depths[0], m.depths[1]);     // line3D.cc:852-855
        }
    }
}
// the ranks' records of one group, view by view and rank by rank, into one contiguous buffer (for the copy to the host)
__global__ void k_mr_gather(const unsigned char* __restrict__ window, MrLayout lay, int W, l3d_match* __restrict__ out)
{
    const int v = blockIdx.y;
    const RankRec* rec = reinterpret_cast<const RankRec*>(window);
    long long off = 0;
    for (int vv = 0; vv < v; ++vv) for (int r = 0; r < W; ++r) off += rec[r * VG_MAX + vv].F;
    for (int r = 0; r < W; ++r) {
        const int n = rec[r * VG_MAX + v].F;
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const l3d_match*>(window + lay.full_off[v]) + size_t(r) * size_t(lay.cap[v]));
        uint4* dst = reinterpret_cast<uint4*>(out + off);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * n; i += gridDim.x * blockDim.x) dst[i] = src[i];
        off += n;
    }
}

// ------------------------------------------------------------------------------------------
// collinearity (cudawrapper.cu:476-535): upper triangle, compacted on the device
// ------------------------------------------------------------------------------------------
// A pair can only pass if all four endpoint-to-line distances stay below sigma * sqrt(2 ln 2) (affinity > 0.5).  Two of
// them -- the other segment's endpoints against THIS row's line -- are tested first in plain FP32 with the line
// normalised once per row: 4 FMAs per pair, and all but about 1 % of the pairs end there.  The bound is widened by
// 0.5 px plus what the reference's own FP32 line through the row's endpoints can be off by (the cross product's
// rounding divided by the segment's length), so that no pair the exact evaluation accepts is dropped.
__global__ void k_collin_count(const float4* __restrict__ segs, int n, float sigma_sqr, int* __restrict__ cnt, unsigned* __restrict__ bits, int wpr)
{
    const int i = blockIdx.y;
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= wpr) return;
    const float4 a = segs[i];
    const float lx = a.y - a.w, ly = a.z - a.x, lz = a.x * a.w - a.y * a.z;          // p1 x p2
    const float len2 = lx * lx + ly * ly;
    const float inv = rsqrtf(fmaxf(len2, 1e-30f));
    const float nx = lx * inv, ny = ly * inv, nz = lz * inv;
    const float big = fmaxf(fmaxf(fabsf(a.x * a.w), fabsf(a.y * a.z)), 1.0f);
    const float ulp = __int_as_float((__float_as_int(big) & 0x7f800000) - (23 << 23));      // 2^(exponent - 23)
    float thr = 1.1775f * sqrtf(sigma_sqr) * 1.0001f + 0.5f + 2.0f * ulp * inv;
    if (!(len2 > 1e-12f) || !(thr < 1e30f)) thr = __int_as_float(0x7f800000);             // degenerate row: no pre-test
    unsigned m = 0;
    for (int b = 0; b < 32; ++b) {
        const int j = w * 32 + b;
        if (j > i && j < n) {
            const float4 q = segs[j];
            const float t1 = fabsf(fmaf(nx, q.x, fmaf(ny, q.y, nz))), t2 = fabsf(fmaf(nx, q.z, fmaf(ny, q.w, nz)));
            if (!(t1 > thr || t2 > thr) && xcollinearity(a, q, sigma_sqr) > 0.0f) m |= 1u << b;
        }
    }
    bits[size_t(i) * wpr + w] = m;
    if (m) atomicAdd(&cnt[i], __popc(m));
}
__global__ void k_collin_fill(const float4* __restrict__ segs, int n, float sigma_sqr, const int* __restrict__ off, const unsigned* __restrict__ bits,
                              int wpr, l3d_collin_pair* __restrict__ out)
{
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return;
    int o = off[i];
    if (off[i + 1] == o) return;
    const float4 a = segs[i];
    for (int w0 = 0; w0 < wpr; w0 += 32) {
        unsigned m = (w0 + lane < wpr) ? bits[size_t(i) * wpr + w0 + lane] : 0u;
        int c = __popc(m), ex = c;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, ex, s); if (lane >= s) ex += y; }
        int tot = __shfl_sync(0xffffffffu, ex, 31);
        int pos = o + ex - c;
        while (m) {
            int b = __ffs(m) - 1; m &= m - 1;
            int j = (w0 + lane) * 32 + b;
            l3d_collin_pair p; p.i = i; p.j = j; p.value = xcollinearity(a, segs[j], sigma_sqr);
            out[pos++] = p;
        }
        o += tot;
    }
}

// ------------------------------------------------------------------------------------------
// Collinearity without any allocation (l3d_collinearity_prealloc).
//
// Some programs that embed this library are not allowed to allocate device memory while
// they run: a reconstruction service that reserves its whole GPU budget at start-up and
// treats any later cudaMalloc as a fault, or a frame loop with a latency bound, where
// cudaMalloc / cudaFree would stall the device for an unpredictable time (both synchronise
// it) and can fail outright once the host application has filled the card.  For them the
// per-image collinearity pass gets a second entry point whose working memory (the pair
// bitmask, the per-row counts and their prefix sums) is this fixed array, part of the
// library image on every GPU that loads it, instead of buffers grown with cudaMalloc.
// The price is a hard upper limit on the number of segments per image.
// ------------------------------------------------------------------------------------------
constexpr int PREALLOC_MAX_SEGMENTS = 4096;
constexpr int PREALLOC_WPR = PREALLOC_MAX_SEGMENTS / 32;
constexpr size_t PREALLOC_BITS_WORDS = size_t(PREALLOC_MAX_SEGMENTS) * PREALLOC_WPR;          // 2 MiB of pair bits
constexpr size_t PREALLOC_CNT_OFF = PREALLOC_BITS_WORDS;                                      // n + 1 counts
constexpr size_t PREALLOC_OFF_OFF = PREALLOC_CNT_OFF + PREALLOC_MAX_SEGMENTS + 32;            // n + 2 offsets
constexpr size_t PREALLOC_WORDS = PREALLOC_OFF_OFF + PREALLOC_MAX_SEGMENTS + 32;
static __device__ __align__(256) unsigned int g_prealloc_scratch[PREALLOC_WORDS];

// like k_collin_fill, but never writes past `capacity` records and also reports the total
__global__ void k_collin_fill_capped(const float4* __restrict__ segs, int n, float sigma_sqr, const int* __restrict__ off, const unsigned* __restrict__ bits,
                                     int wpr, l3d_collin_pair* __restrict__ out, int capacity, int* __restrict__ total_out)
{
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i == 0 && lane == 0 && total_out) *total_out = off[n];
    if (i >= n) return;
    int o = off[i];
    if (off[i + 1] == o) return;
    const float4 a = segs[i];
    for (int w0 = 0; w0 < wpr; w0 += 32) {
        unsigned m = (w0 + lane < wpr) ? bits[size_t(i) * wpr + w0 + lane] : 0u;
        int c = __popc(m), ex = c;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, ex, s); if (lane >= s) ex += y; }
        int tot = __shfl_sync(0xffffffffu, ex, 31);
        int pos = o + ex - c;
        while (m) {
            int b = __ffs(m) - 1; m &= m - 1;
            int j = (w0 + lane) * 32 + b;
            if (pos < capacity) { l3d_collin_pair p; p.i = i; p.j = j; p.value = xcollinearity(a, segs[j], sigma_sqr); out[pos] = p; }
            ++pos;
        }
        o += tot;
    }
}

// ------------------------------------------------------------------------------------------
// hypotheses (view.cc:302-342) and affinity (line3D.cc:1600-1691), FP64
// ------------------------------------------------------------------------------------------
struct ViewD { double RtKinv[9]; double C[3]; float k_lower, k_upper; int seg_off; int n_seg; };
struct Hyp { double P1[3], P2[3], dir[3]; float d1, d2; int view; int pad; };

__device__ __forceinline__ void d_normalize(double* v)
{
    double n = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(v[0], v[0]), __dmul_rn(v[1], v[1])), __dmul_rn(v[2], v[2])));
    v[0] = __ddiv_rn(v[0], n); v[1] = __ddiv_rn(v[1], n); v[2] = __ddiv_rn(v[2], n);
}
__global__ void k_unproject(const float4* __restrict__ segs, const ViewD* __restrict__ views, const l3d_hypothesis_in* __restrict__ in, size_t n,
                            Hyp* __restrict__ hyp, l3d_segment3d* __restrict__ out)
{
    const size_t k = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const l3d_hypothesis_in h = in[k];
    const ViewD& v = views[h.view];
    const float4 s = segs[v.seg_off + h.seg];
    const double p[2][3] = {{double(s.x), double(s.y), 1.0}, {double(s.z), double(s.w), 1.0}};
    const double dep[2] = {double(h.depth1), double(h.depth2)};
    Hyp o; double P[2][3];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        double r[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < 3; ++c) acc = __dadd_rn(acc, __dmul_rn(v.RtKinv[a * 3 + c], p[e][c]));
            r[a] = acc;
        }
        d_normalize(r);
#pragma unroll
        for (int a = 0; a < 3; ++a) P[e][a] = __dadd_rn(v.C[a], __dmul_rn(r[a], dep[e]));
    }
    double d[3] = {__dsub_rn(P[1][0], P[0][0]), __dsub_rn(P[1][1], P[0][1]), __dsub_rn(P[1][2], P[0][2])};
    d_normalize(d);
    l3d_segment3d so;
#pragma unroll
    for (int a = 0; a < 3; ++a) { o.P1[a] = P[0][a]; o.P2[a] = P[1][a]; o.dir[a] = d[a]; so.P1[a] = P[0][a]; so.P2[a] = P[1][a]; so.dir[a] = d[a]; }
    o.d1 = h.depth1; o.d2 = h.depth2; o.view = h.view; o.pad = 0;
    hyp[k] = o; out[k] = so;
}

// line3D.cc:1684-1691: |P1 + (dir (X - P1)^T) dir - X|, the outer product evaluated first
__device__ __forceinline__ float d_point_line(const Hyp& s, const double* X)
{
    double v[3] = {__dsub_rn(X[0], s.P1[0]), __dsub_rn(X[1], s.P1[1]), __dsub_rn(X[2], s.P1[2])};
    double e[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < 3; ++j) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(s.dir[i], v[j]), s.dir[j]));
        e[i] = __dsub_rn(__dadd_rn(s.P1[i], acc), X[i]);
    }
    return float(sqrt(__dadd_rn(__dadd_rn(__dmul_rn(e[0], e[0]), __dmul_rn(e[1], e[1])), __dmul_rn(e[2], e[2]))));
}
// view.cc:353-377
__device__ __forceinline__ float d_sim_term(float d, float depth, float k_lower, float k_upper, float median, float neg2log)
{
    float lo = depth < median ? xm(k_lower, depth) : xm(k_lower, median);
    float up = depth < median ? xm(k_upper, depth) : xm(k_upper, median);
    if (d < lo) return 1.0f;
    float diff = xs(up, lo);
    float sig = xd(xm(-diff, diff), neg2log);                      // -(d2-d1)^2 / (2 logf(0.01f))
    float t = xs(d, lo);
    return xexpf(xd(xm(-t, t), xm(2.0f, sig)));
}
__global__ void k_affinity(const Hyp* __restrict__ hyp, const ViewD* __restrict__ views, const float* __restrict__ median,
                           const l3d_pair* __restrict__ pairs, size_t n, float sigma_a, float two_log001, float* __restrict__ out)
{
    const size_t k = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const Hyp a = hyp[pairs[k].a], b = hyp[pairs[k].b];
    const ViewD& va = views[a.view]; const ViewD& vb = views[b.view];
    const float ma = median[a.view], mb = median[b.view];
    float s1 = d_sim_term(d_point_line(b, a.P1), a.d1, va.k_lower, va.k_upper, ma, two_log001);
    float s2 = d_sim_term(d_point_line(b, a.P2), a.d2, va.k_lower, va.k_upper, ma, two_log001);
    float s3 = d_sim_term(d_point_line(a, b.P1), b.d1, vb.k_lower, vb.k_upper, mb, two_log001);
    float s4 = d_sim_term(d_point_line(a, b.P2), b.d2, vb.k_lower, vb.k_upper, mb, two_log001);
    float w_d = fminf(fminf(s1, s2), fminf(s3, s4));
    double dt = __dadd_rn(__dadd_rn(__dmul_rn(a.dir[0], b.dir[0]), __dmul_rn(a.dir[1], b.dir[1])), __dmul_rn(a.dir[2], b.dir[2]));
    float angle = float(acos(fmax(fmin(dt, 1.0), -1.0)) / 3.14159265358979323846 * 180.0);
    if (angle > 90.0f) angle = xs(180.0f, angle);
    float w_a = xexpf(xd(xm(-angle, angle), xm(xm(2.0f, sigma_a), sigma_a)));
    float sim = fminf(w_d, w_a);
    out[k] = (sim <= 0.01f) ? 0.0f : sim;
}

// Candidate pairs -> affinity edges on the device (line3D.cc:1014-1016, 1085-1087, 1163-1165): the similarity of the two
// hypotheses, the edge weight  mult * 0.5 * (score_a + score_b) * sim  in the host code's order of operations, the
// threshold of the candidate's kind, and an ordered compaction of the survivors -- only they travel back to the host.
constexpr int AE_T = 256;
__device__ __forceinline__ float affinity_of(const Hyp& a, const Hyp& b, const ViewD* __restrict__ views, const float* __restrict__ median, float sigma_a, float two_log001)
{
    const ViewD& va = views[a.view]; const ViewD& vb = views[b.view];
    const float ma = median[a.view], mb = median[b.view];
    float s1 = d_sim_term(d_point_line(b, a.P1), a.d1, va.k_lower, va.k_upper, ma, two_log001);
    float s2 = d_sim_term(d_point_line(b, a.P2), a.d2, va.k_lower, va.k_upper, ma, two_log001);
    float s3 = d_sim_term(d_point_line(a, b.P1), b.d1, vb.k_lower, vb.k_upper, mb, two_log001);
    float s4 = d_sim_term(d_point_line(a, b.P2), b.d2, vb.k_lower, vb.k_upper, mb, two_log001);
    float w_d = fminf(fminf(s1, s2), fminf(s3, s4));
    double dt = __dadd_rn(__dadd_rn(__dmul_rn(a.dir[0], b.dir[0]), __dmul_rn(a.dir[1], b.dir[1])), __dmul_rn(a.dir[2], b.dir[2]));
    float angle = float(acos(fmax(fmin(dt, 1.0), -1.0)) / 3.14159265358979323846 * 180.0);
    if (angle > 90.0f) angle = xs(180.0f, angle);
    float w_a = xexpf(xd(xm(-angle, angle), xm(xm(2.0f, sigma_a), sigma_a)));
    float sim = fminf(w_d, w_a);
    return (sim <= 0.01f) ? 0.0f : sim;
}
__global__ void __launch_bounds__(AE_T) k_affinity_edges(const Hyp* __restrict__ hyp, const ViewD* __restrict__ views, const float* __restrict__ median, const float* __restrict__ score,
                                                         const l3d_candidate* __restrict__ cands, size_t n, float sigma_a, float two_log001,
                                                         float* __restrict__ w_out, int* __restrict__ cta_kept)
{
    __shared__ int s_n[AE_T / 32];
    const size_t k = size_t(blockIdx.x) * AE_T + threadIdx.x;
    float w = 0.0f; bool keep = false;
    if (k < n) {
        const l3d_candidate c = cands[k];
        const float sim = affinity_of(hyp[c.a], hyp[c.b], views, median, sigma_a, two_log001);
        w = xm(xm(xm(c.mult, 0.5f), xa(score[c.a], score[c.b])), sim);
        keep = w > c.thr;
        w_out[k] = keep ? w : 0.0f;
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if ((threadIdx.x & 31) == 0) s_n[threadIdx.x >> 5] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int q = 0; q < AE_T / 32; ++q) t += s_n[q]; cta_kept[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(AE_T) k_affinity_compact(const float* __restrict__ w_in, size_t n, const int* __restrict__ cta_off, l3d_kept_edge* __restrict__ out)
{
    __shared__ int s_n[AE_T / 32];
    const size_t k = size_t(blockIdx.x) * AE_T + threadIdx.x;
    const float w = (k < n) ? w_in[k] : 0.0f;
    const bool keep = w != 0.0f;                       // (a kept weight exceeds a positive threshold)
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_n[warp] = __popc(m);
    __syncthreads();
    int before = 0;
    for (int q = 0; q < warp; ++q) before += s_n[q];
    if (keep) { l3d_kept_edge e; e.index = unsigned(k); e.w = w; out[cta_off[blockIdx.x] + before + __popc(m & ((1u << lane) - 1u))] = e; }
}

// The order in which the clustering takes the edges (clustering.cc:12: a stable sort by ascending weight): the kept
// candidates sorted by the 64-bit key (weight bits, position) -- weights are positive, so their bit patterns order like
// the values, and the position makes every key unique, so any sort gives the stable order.  Bitonic network over a
// power of two of keys (padding: all ones); strides below the tile size run in shared memory.
constexpr int BS_TILE = 2048;
__global__ void k_edge_keys(const l3d_kept_edge* __restrict__ kept, int n, int n_pad, unsigned long long* __restrict__ keys)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pad) return;
    keys[k] = (k < n) ? (((unsigned long long)__float_as_uint(kept[k].w) << 32) | (unsigned long long)unsigned(k)) : ~0ull;
}
__global__ void k_bitonic_global(unsigned long long* __restrict__ keys, int j, int k)
{
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned p = i ^ unsigned(j);
    if (p > i) {
        const unsigned long long a = keys[i], b = keys[p];
        const bool up = (i & unsigned(k)) == 0;
        if ((a > b) == up) { keys[i] = b; keys[p] = a; }
    }
}
__global__ void __launch_bounds__(BS_TILE / 2) k_bitonic_shared(unsigned long long* __restrict__ keys, int j_first, int k)
{
    __shared__ unsigned long long s[BS_TILE];
    const unsigned base = blockIdx.x * BS_TILE;
    s[threadIdx.x] = keys[base + threadIdx.x]; s[threadIdx.x + BS_TILE / 2] = keys[base + threadIdx.x + BS_TILE / 2];
    __syncthreads();
    for (int j = j_first; j > 0; j >>= 1) {
        const unsigned t = threadIdx.x;
        const unsigned i = ((t & ~unsigned(j - 1)) << 1) | (t & unsigned(j - 1));      // the lower index of the t-th pair at stride j
        const unsigned p = i | unsigned(j);
        const bool up = ((base + i) & unsigned(k)) == 0;
        const unsigned long long a = s[i], b = s[p];
        if ((a > b) == up) { s[i] = b; s[p] = a; }
        __syncthreads();
    }
    keys[base + threadIdx.x] = s[threadIdx.x]; keys[base + threadIdx.x + BS_TILE / 2] = s[threadIdx.x + BS_TILE / 2];
}
__global__ void k_edge_order(const unsigned long long* __restrict__ keys, int n, unsigned* __restrict__ order)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) order[k] = unsigned(keys[k] & 0xffffffffull);
}

// ------------------------------------------------------------------------------------------
// diffusion (cudawrapper.cu:717-829, 1131-1191)
// ------------------------------------------------------------------------------------------
// P: row-sorted entries (row, col, val); W: column-sorted.  rstart/cstart: first entry of a
// row of P / a column of W.  tpos[y] = position in P of the entry (col_y, row_y).
__global__ void k_rdd_normalise(const int* __restrict__ prow, float* __restrict__ pval, const int* __restrict__ rstart, int n, int nnz)
{
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= n) return;
    const int s = rstart[y];
    if (s < 0) return;
    float sum = 0.0f;
    int i = s;
    while (i < nnz && prow[i] == y) { sum = xa(sum, pval[i]); ++i; }
    if (sum < L3D_EPS_F) sum = L3D_EPS_F;
    for (int k = s; k < i; ++k) pval[k] = xd(pval[k], sum);
}
__global__ void k_rdd_step(const int* __restrict__ prow, const int* __restrict__ pcol, const float* __restrict__ pval,
                           const int* __restrict__ wcol, const float* __restrict__ wval,
                           const int* __restrict__ rstart, const int* __restrict__ cstart, const int* __restrict__ tpos,
                           float* __restrict__ pnext, int nnz)
{
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= nnz) return;
    const int r = pcol[y], c = prow[y];                     // transposed (cudawrapper.cu:778-780)
    float mul = 0.0f;
    int sp = rstart[r], sw = cstart[c];
    while (sp >= 0 && sw >= 0 && sp < nnz && sw < nnz) {    // lockstep walk (cudawrapper.cu:786-800)
        if (prow[sp] != r || wcol[sw] != c) break;
        mul = xa(mul, xm(pval[sp], wval[sw]));
        ++sp; ++sw;
    }
    mul = xm(mul, pval[y]);
    if (mul < L3D_EPS_F) mul = L3D_EPS_F;
    if (tpos[y] >= 0) pnext[tpos[y]] = mul;
}

// libm-compatible float functions, exposed for tests
__global__ void k_debug_math(int which, const float* __restrict__ in, float* __restrict__ out, int n)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    out[k] = which == 0 ? xexpf(in[k]) : xacosf(in[k]);
}

// FP32 throughput probes: register-only chains, operands in registers (not immediates), 16 independent chains per
// thread, 8 steps unrolled: 128 FFMA (64 FFMA2) per loop trip against 3 loop instructions.
__global__ void k_fma_probe(float* out, int iters, float b, float c)
{
    float a[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = threadIdx.x * 1e-3f + float(k);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int k = 0; k < 16; ++k) a[k] = fmaf(a[k], b, c);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fma2_probe(float* out, int iters, float b, float c)
{
    f32x2 a[8];
    const f32x2 B = pk2(b, b * 1.0000001f), C = pk2(c, c * 1.5f);
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = pk2(threadIdx.x * 1e-3f + float(k), threadIdx.x * 2e-3f + float(k));
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fma2(a[k], B, C);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { unsigned lo, hi; up2(a[k], lo, hi); s += __uint_as_float(lo) + __uint_as_float(hi); }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

// ==========================================================================================
// context
// ==========================================================================================
struct Delivered {            // records a source view handed to a later view
    int from_view = -1;
    int count = 0;            // records (a bound until the sending view has been collected; several GPUs: this rank's share)
    long long count_all = 0;  // several GPUs: a bound on all ranks' records together, the same number on every rank (it fixes the window layout)
    int* t_start = nullptr;   // n_seg(to) + 1, points into the owning buffer
    int* t_tgt = nullptr;
    float4* t_dep = nullptr;
    int owner = -1;           // index into ctx->deliver_bufs
    int max_group = 0;        // most records delivered to a single segment
    int slot = 0;             // position among the owner's deliveries
};
struct DeliverBuf { int* start = nullptr; int* tgt = nullptr; float4* dep = nullptr; int refs = 0; std::vector<int> counts, maxg; };   // counts/maxg: per slot, bounds until the owner is collected

struct PlannedView {           // one entry of l3d_plan_matching
    int view = -1;
    std::vector<l3d_neighbor> nb;
    uint64_t pair_tests = 0;     // of this process; with several GPUs on one scene: of the rank with most rows (the same number on every rank)
    int first_job = -1;         // valid while the view's batch is resident
    uint64_t n_candidates = 0;
};
struct JobHost { int src_view, tgt_view; float F[9]; };
struct ViewJob {                 // phase B of one view: on the stream, not yet collected
    bool active = false;
    int view = -1, parity = 0;
    std::vector<l3d_neighbor> nb;
    float sigma_p = 0.f, sigma_a = 0.f;
    uint64_t pair_tests = 0, cand_ub = 0;
    size_t M_ub = 0;
    bool scored = false, delivering = false;
    int n_deliver = 0, owner = -1;
    std::vector<std::pair<int, int> > consumed;      // (deliver buffer, slot) of every delivery the view took in
};
// The per-view work runs on groups of views (see VArgs): a slot holds one view's buffers, a group up to VG_MAX
// slots plus what the group's views share (one upload of tables and arguments, one memset, one result download).
// Two groups exist: the host prepares and launches the next one while the GPU works on the current one.
struct ViewSlot {
    ViewJob job;
    size_t plan_index = 0;
    DBuf<int> d_cnt, d_off, d_cta, d_F, d_boff;
    DBuf<int4> d_mids; DBuf<float4> d_mdep; DBuf<float> d_score, d_dsamples, d_sk, d_sd2; DBuf<unsigned> d_spl;
    DBuf<int2> d_items; DBuf<uint4> d_brange; DBuf<l3d_match> d_filtered;
    void release()
    {
        d_cnt.release(); d_off.release(); d_cta.release(); d_F.release(); d_boff.release(); d_mids.release(); d_mdep.release(); d_score.release();
        d_dsamples.release(); d_sk.release(); d_sd2.release(); d_spl.release(); d_items.release(); d_brange.release(); d_filtered.release();
    }
};
struct ViewGroup {
    int count = 0; size_t pos0 = 0;      // positions [pos0, pos0 + count) of the execution order
    ViewSlot slot[VG_MAX];
    HBuf<VArgs> h_args; DBuf<VArgs> d_args;
    HBuf<CamTable> h_cams; DBuf<CamTable> d_cams; size_t cam_off[VG_MAX + 1];
    DBuf<int> d_zero; DBuf<int> d_vres; HBuf<int> h_vres; int res_stride = 8;     // result ints per view: 8 scalars + one count per hand-over slot
    // several GPUs on one scene: depth samples of the group's views (one buffer: one all-reduce), the ranks' records,
    // the packed records of this rank, all ranks' packed records, all ranks' records view by view
    size_t hist_off = 0, hist_total = 0;
    HBuf<RankRec> h_allrec; MrLayout lay; DBuf<l3d_match> d_gather; int* d_fault = nullptr;
    cudaEvent_t ev_x[2] = {nullptr, nullptr};      // [0] exchange done on the main stream, [1] this group's records copied to the host
    cudaEvent_t kev[24] = {}; int kev_slot[24] = {}; int n_kev = 0;      // per-kernel events of the chain (l3d_debug_chain_times)
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // [0] results on the host, [2..5] stage boundaries
    void release()
    {
        for (int k = 0; k < VG_MAX; ++k) slot[k].release();
        h_args.release(); d_args.release(); h_cams.release(); d_cams.release(); d_zero.release(); d_vres.release(); h_vres.release();
        h_allrec.release(); d_gather.release();
        for (int k = 0; k < 2; ++k) if (ev_x[k]) { cudaEventDestroy(ev_x[k]); ev_x[k] = nullptr; }
        for (int k = 0; k < 24; ++k) if (kev[k]) { cudaEventDestroy(kev[k]); kev[k] = nullptr; }
        for (int k = 0; k < 6; ++k) if (ev[k]) { cudaEventDestroy(ev[k]); ev[k] = nullptr; }
    }
};
// Device memory for the hand-over buffers, recycled on the host: a buffer goes back when the last receiving view has
// been collected, so nothing on the GPU can still be reading it (stream-ordered allocation across streams cost 10 us
// per call here).
struct DevPool {
    struct Block { void* p; size_t cap; bool free; };
    std::vector<Block> blocks;
    cudaError_t get(size_t bytes, void** out)
    {
        size_t best = blocks.size();
        for (size_t k = 0; k < blocks.size(); ++k)
            if (blocks[k].free && blocks[k].cap >= bytes && blocks[k].cap <= 4 * std::max<size_t>(bytes, 4096) && (best == blocks.size() || blocks[k].cap < blocks[best].cap)) best = k;
        if (best == blocks.size()) {
            size_t cap = 4096; while (cap < bytes) cap <<= 1;
            void* p = nullptr;
            cudaError_t e = cudaMalloc(&p, cap);
            if (e != cudaSuccess) return e;
            Block b = {p, cap, true}; blocks.push_back(b);
        }
        blocks[best].free = false; *out = blocks[best].p;
        return cudaSuccess;
    }
    void put(void* p) { for (size_t k = 0; k < blocks.size(); ++k) if (blocks[k].p == p) { blocks[k].free = true; return; } }
    void put_all() { for (size_t k = 0; k < blocks.size(); ++k) blocks[k].free = true; }
    void release() { for (size_t k = 0; k < blocks.size(); ++k) cudaFree(blocks[k].p); blocks.clear(); }
};

// Everything one batch of pair tests (phase A) produces and needs: two of these exist, so that a helper thread can
// run the next batch's pair tests on a second stream while the per-view work (phase B) of the current batch, which
// leaves most of the GPU idle, goes on.
struct PhaseASet {
    cudaStream_t stream = nullptr;
    std::string err;
    uint64_t launches = 0;
    size_t begin = 0, end = 0;                 // plan entries covered
    DBuf<SrcHoist> d_sh; DBuf<SrcHoist1> d_sh1; DBuf<TgtHoist> d_th; DBuf<int> d_row_job;
    DBuf<unsigned> d_cand, d_l1; DBuf<int> d_tile_sums, d_total;
    DBuf<int> d_stage_col, d_stage_row; DBuf<float4> d_stage_dep; DBuf<unsigned char> d_stage_ok;
    DBuf<JobDev> d_jobs; DBuf<int> d_pa_cnt, d_pa_off, d_pa_raw, d_pa_idx, d_pa_tot;
    std::vector<JobDev> jobs_host; std::vector<uint64_t> job_cands, job_raw, job_raw_all /* summed over the ranks (several GPUs on one scene) */; std::vector<int> job_rowmax, job_raw_rowmax; DBuf<int> d_rm_begin, d_rm_len, d_rm_out;
    DBuf<unsigned long long> d_job_sum64; HBuf<unsigned long long> h_job_sum64;
    bool pa_ms_pending = false;
    bool background = false;      // this batch's pair tests run beside the per-view chain of the batch before (helper thread)
    cudaEvent_t pa_ev[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t pa_kev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // [0][1] around k_pretest, [1][4] around k_pretest2, [2][3] around k_exact_flat
    bool pa_kev_set[2] = {false, false};
    HBuf<int> h_res;
    void release()
    {
        d_sh.release(); d_sh1.release(); d_th.release(); d_row_job.release(); d_cand.release(); d_l1.release(); d_tile_sums.release(); d_total.release();
        d_stage_col.release(); d_stage_row.release(); d_stage_dep.release(); d_stage_ok.release(); d_jobs.release();
        d_pa_cnt.release(); d_pa_off.release(); d_pa_raw.release(); d_pa_idx.release(); d_pa_tot.release();
        d_rm_begin.release(); d_rm_len.release(); d_rm_out.release(); h_res.release(); d_job_sum64.release(); h_job_sum64.release();
        for (int k = 0; k < 3; ++k) if (pa_ev[k]) { cudaEventDestroy(pa_ev[k]); pa_ev[k] = nullptr; }
        for (int k = 0; k < 5; ++k) if (pa_kev[k]) { cudaEventDestroy(pa_kev[k]); pa_kev[k] = nullptr; }
        if (stream) { cudaStreamDestroy(stream); stream = nullptr; }
    }
};

struct l3d_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    // several GPUs on one scene (l3d_comm_init)
    NcclApi* nccl = nullptr; ncclComm_t comm = nullptr; int rank = 0, world = 1, result_root = -1;
    bool comm_broken = false;          // a call failed while several ranks were at work: no collective step is attempted any more (the others may be gone)
    bool copies_pending = false, want_records = true;      // want_records: the current calls copy the output records to the host
    bool chain_prof = false; double chain_ms[24] = {};      // device ms per kernel of the per-view chain, summed over the groups (with stage timing)
    double host_s[8] = {0, 0, 0, 0, 0, 0, 0, 0};      // host seconds in enqueue_group / finish_group (waiting) / finish_group (rest) / number of groups / waiting for the helper thread / pair tests in the foreground / totals over the ranks (l3d_debug_host_times)
    // how the pair tests of the next batch share the GPU with the per-view chain (read from the environment by l3d_plan_matching)
    int bg_blocks = 0;                              // L3D_BG_BLOCKS: blocks per SM the large pair-test kernels may take in the background (0: no cap)
    int p1_rows = 0;                                // L3D_P1_ROWS: source rows per block of k_pretest (0: P1_ROWS)
    bool no_overlap = false;                        // L3D_NO_OVERLAP: pair tests never run beside the chain
    int batch_div = 8;                              // L3D_BATCH_DIV: the schedule is cut into about this many batches
    int group_views = 8;                            // L3D_GROUP_VIEWS: views per group of the per-view chain (at most VG_MAX)
    // the record windows (one per group in flight): every rank's is mapped into every process, the ranks store their
    // output records straight into each other's
    unsigned char* win[2] = {nullptr, nullptr}; size_t win_cap = 0 /* bytes */; MrPeers peers[2]; unsigned win_seq[2] = {0u, 0u};
    DBuf<int> d_bar; DBuf<unsigned long long> d_tot; HBuf<unsigned long long> h_tot; DBuf<unsigned char> d_ipc; HBuf<unsigned char> h_ipc;
    uint64_t launches = 0;
    // views
    int V = 0;
    std::vector<l3d_view> views;          // host copies (segments pointer not kept)
    std::vector<int> seg_off;
    int total_segs = 0, max_seg = 0;
    DBuf<float4> d_segs;
    DBuf<ViewD> d_viewd;
    std::vector<std::vector<Delivered> > delivered;   // per destination view
    std::vector<DeliverBuf> deliver_bufs;
    // scratch
    DBuf<int> d_tile_sums, d_total;
    size_t score_smem = 48 * 1024; int n_sm = 148;
    // phase A (pair tests): the batch the per-view work reads (pa[cur]) and the one a helper thread prepares meanwhile
    PhaseASet pa[2]; int cur = 0;
    std::thread pa_worker; bool pa_worker_active = false; int pa_worker_rc = 0;
    std::vector<PlannedView> plan;                  // as announced (the order of the caller's calls)
    std::vector<size_t> order;                      // plan entries in the order they are executed: by dependency level
    struct Stashed { l3d_match_result res; bool ready; float ms[6], kms[3]; };
    std::vector<Stashed> results;                   // per plan entry, until the caller asks for it
    size_t call_next = 0; bool plan_started = false, run_copy_back = true;
    size_t batch_begin = 0, batch_end = 0;          // positions in `order` whose pair tests are resident
    uint64_t pair_budget = 24000000000ull;          // most pair tests of one batch (3 GB of candidate bits)
    uint64_t batch_pairs = 24000000000ull;          // pair tests per batch of the announced schedule (l3d_plan_matching)
    uint64_t batch_override = 0;                    // l3d_debug_set_batch_pairs
    int pretest_mode = 0;                           // l3d_debug_set_pretest
    int score_block = 0;                            // l3d_debug_set_score_block
    uint64_t cand_limit = 0x7fff0000ull;            // most candidates of one batch (32-bit candidate index); l3d_debug_set_candidate_limit
    float pa_kernel_ms[3] = {0.f, 0.f, 0.f};
    // views in flight: positions [fly_begin, fly_end) of `order` are on the GPU, as the groups [grp_begin, grp_end)
    ViewGroup grp[2]; size_t grp_begin = 0, grp_end = 0;
    size_t fly_begin = 0, fly_end = 0; std::vector<int> level;
    float run_sigma_p = 0.f, run_sigma_a = 0.f;
    DevPool pool;
    cudaStream_t copy_stream = nullptr, fin_stream = nullptr;
    bool stage_events = false;         // per-stage CUDA events inside a view (l3d_set_stage_timing): they separate kernels that otherwise overlap
    cudaEvent_t mark_ev[2] = {nullptr, nullptr};
    // retained results (l3d_retain_results): pinned chunks that are never moved, handed out front to back
    bool retain = false; std::vector<std::pair<l3d_match*, size_t> > arena; size_t arena_chunk = 0, arena_used = 0;
    HBuf<l3d_match> h_filtered; HBuf<int> h_ints;
    // collinearity
    DBuf<float4> d_csegs; DBuf<unsigned> d_cbits; DBuf<int> d_ccnt, d_coff; DBuf<l3d_collin_pair> d_cpairs; HBuf<l3d_collin_pair> h_cpairs;
    // hypotheses
    DBuf<Hyp> d_hyp; size_t n_hyp = 0; DBuf<l3d_hypothesis_in> d_hin; DBuf<l3d_segment3d> d_hout;
    DBuf<l3d_pair> d_pairs; DBuf<float> d_aff, d_medians;
    HBuf<l3d_candidate> h_cands; DBuf<l3d_candidate> d_cands; DBuf<float> d_hscore, d_ew; DBuf<int> d_ecnt, d_eoff; DBuf<l3d_kept_edge> d_kept; HBuf<l3d_kept_edge> h_kept; DBuf<unsigned long long> d_ekeys; DBuf<unsigned> d_eorder; HBuf<unsigned> h_eorder;
    DBuf<unsigned char> d_flush; int flush_value = 0;
    // diffusion: one device and one pinned buffer for all its arrays; the step kernel's duration and algorithmic bytes of the last call
    DBuf<int> d_rdd; HBuf<int> h_rdd; float rdd_step_ms = 0.f; double rdd_step_bytes = 0.0; int rdd_nnz = 0;
    // timing
    cudaEvent_t ev[2] = {nullptr, nullptr};          // around the FP32 peak probe
    float last_ms[6] = {0, 0, 0, 0, 0, 0};
};

namespace {

template <typename K, typename... Args>
inline cudaError_t pdl_launch(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

template <typename C>
int scan_exclusive(C* ctx, const int* in, int* out /* n + 1 */, int n)
{
    if (n <= SCAN_SMALL) {
        L3D_CUDA(pdl_launch(k_scan_small, dim3(1), dim3(SCAN_T), 0, ctx->stream, in, out, n));
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        return 0;
    }
    const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (tiles > SCAN_TILE) { ctx->err = "scan: input too long"; return 1; }
    L3D_CUDA(ctx->d_tile_sums.ensure(size_t(std::max(tiles, 1))));
    L3D_CUDA(ctx->d_total.ensure(1));
    L3D_CUDA(pdl_launch(k_scan_tiles, dim3(std::max(tiles, 1)), dim3(SCAN_T), 0, ctx->stream, in, out, ctx->d_tile_sums.p, n));
    L3D_CUDA(pdl_launch(k_scan_sums, dim3(1), dim3(SCAN_T), 0, ctx->stream, ctx->d_tile_sums.p, tiles, ctx->d_total.p));
    L3D_CUDA(pdl_launch(k_scan_add, dim3(std::max(tiles, 1)), dim3(SCAN_T), 0, ctx->stream, out, ctx->d_tile_sums.p, n, ctx->d_total.p));
    ctx->launches += 3;
    L3D_CUDA(cudaGetLastError());
    return 0;
}

// rows of a view this process holds: all of them, or every world-th segment starting at its rank
inline int local_rows(const l3d_ctx* ctx, int n_segments) { return ctx->world <= 1 ? n_segments : (n_segments > ctx->rank ? (n_segments - ctx->rank + ctx->world - 1) / ctx->world : 0); }

int fetch_int(l3d_ctx* ctx, const int* dptr, int* out)
{
    L3D_CUDA(ctx->h_ints.ensure(16));
    L3D_CUDA(cudaMemcpyAsync(ctx->h_ints.p, dptr, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    L3D_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = ctx->h_ints.p[0];
    return 0;
}

// Phase A for a set of (source view, neighbour) jobs: hoist, pre-test, scan, exact test.
// Results stay on the device: d_pa_off (candidate offsets per global row = staging ranges),
// d_pa_raw (raw matches per global row), d_stage_col / d_stage_dep.
int run_phase_a(l3d_ctx* cx, PhaseASet* ctx, const std::vector<JobHost>& jh)   // ctx: the set being filled (the error macro writes ctx->err)
{
    L3D_CUDA(cudaSetDevice(cx->device));
    cudaStream_t st = ctx->stream;
    const int n_jobs = int(jh.size());
    // Pair tests that run beside the per-view chain must leave it room: the chain is a sequence of short dependent
    // kernels whose blocks can only start when registers and shared memory are free, and kernels that fill every SM
    // make each of them wait for blocks to retire.  In the background the large kernels therefore ask for unused
    // dynamic shared memory, which caps them at 3-4 blocks per SM (about half of each SM stays free); alone they run
    // at full occupancy.
    size_t bg_p1 = 0, bg_p2 = 0, bg_ex = 0, bg_ef = 0;
    if (ctx->background && cx->bg_blocks > 0) {
        // what leaves room for bg_blocks blocks per SM: 228 KB of shared memory, 1 KB reserved per block
        auto cap = [&](const void* fn, size_t* dyn) -> cudaError_t {
            cudaFuncAttributes a; cudaError_t e = cudaFuncGetAttributes(&a, fn); if (e != cudaSuccess) return e;
            const long long per = (228ll * 1024) / (cx->bg_blocks + 1) + 1024;      // one block more would not fit
            long long d = per - 1024 - (long long)a.sharedSizeBytes;
            d = std::max(0ll, std::min(d, 227ll * 1024 - (long long)a.sharedSizeBytes));
            *dyn = size_t(d);
            return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(d));
        };
        L3D_CUDA(cap(reinterpret_cast<const void*>(&k_pretest<4, 6>), &bg_p1)); L3D_CUDA(cap(reinterpret_cast<const void*>(&k_pretest2), &bg_p2));
        L3D_CUDA(cap(reinterpret_cast<const void*>(&k_expand), &bg_ex)); L3D_CUDA(cap(reinterpret_cast<const void*>(&k_exact_flat), &bg_ef));
    }
    ctx->jobs_host.assign(size_t(n_jobs), JobDev());
    ctx->job_cands.assign(size_t(n_jobs), 0);
    ctx->job_rowmax.assign(size_t(n_jobs), 0);
    ctx->job_raw.assign(size_t(n_jobs), 0); ctx->job_raw_rowmax.assign(size_t(n_jobs), 0);
    ctx->pa_kev_set[0] = ctx->pa_kev_set[1] = false;
    if (n_jobs == 0) return 0;
    long long sh_tot = 0, th_tot = 0, words = 0, rows = 0, blocks = 0, l1_words = 0, l1_blocks = 0;
    const int p1_rows = cx->p1_rows > 0 ? cx->p1_rows : P1_ROWS;
    int max_n = 1;
    for (int j = 0; j < n_jobs; ++j) {
        const l3d_view& sv = cx->views[jh[j].src_view]; const l3d_view& tv = cx->views[jh[j].tgt_view];
        JobDev& d = ctx->jobs_host[j];
        memcpy(d.F, jh[j].F, sizeof d.F); memcpy(d.RtKinv_s, sv.RtKinv, sizeof d.RtKinv_s); memcpy(d.RtKinv_t, tv.RtKinv, sizeof d.RtKinv_t);
        memcpy(d.C_s, sv.C, sizeof d.C_s); memcpy(d.C_t, tv.C, sizeof d.C_t);
        d.src_off = cx->seg_off[jh[j].src_view]; d.n_src = local_rows(cx, sv.n_segments); d.src_first = cx->world > 1 ? cx->rank : 0; d.src_stride = cx->world > 1 ? cx->world : 1; d.tgt_off = cx->seg_off[jh[j].tgt_view]; d.n_tgt = tv.n_segments;
        d.sh_off = int(sh_tot); d.th_off = int(th_tot); d.cand_off = words; d.wpr = (d.n_tgt + 31) / 32;
        d.row_off = int(rows); d.blk_off = int(blocks); d.col_blocks = std::max((d.n_tgt + PT_COLS - 1) / PT_COLS, 1);
        sh_tot += d.n_src; th_tot += d.n_tgt; words += (long long)d.n_src * d.wpr; rows += d.n_src;
        blocks += (long long)((d.n_src + PT_ROWS - 1) / PT_ROWS) * d.col_blocks;
        d.l1_off = l1_words; d.l1_blk_off = int(l1_blocks);
        l1_words += (long long)((d.n_src + 31) / 32) * d.n_tgt;
        l1_blocks += (long long)((d.n_src + p1_rows - 1) / p1_rows) * ((d.n_tgt + P1_COLS - 1) / P1_COLS);
        max_n = std::max(max_n, std::max(d.n_src, d.n_tgt));
        if (sh_tot > 0x7fffffff || th_tot > 0x7fffffff || rows > 0x7ffffff0 || blocks > 0x7ffffff0 || l1_blocks > 0x7ffffff0) { ctx->err = "phase A: batch too large"; return 1; }
    }
    L3D_CUDA(ctx->d_jobs.ensure(size_t(n_jobs)));
    L3D_CUDA(cudaMemcpyAsync(ctx->d_jobs.p, ctx->jobs_host.data(), sizeof(JobDev) * size_t(n_jobs), cudaMemcpyHostToDevice, st));
    L3D_CUDA(ctx->d_sh.ensure(size_t(std::max<long long>(sh_tot, 1)))); L3D_CUDA(ctx->d_sh1.ensure(size_t(std::max<long long>(sh_tot, 1)))); L3D_CUDA(ctx->d_th.ensure(size_t(std::max<long long>(th_tot, 1))));
    L3D_CUDA(ctx->d_cand.ensure(size_t(std::max<long long>(words, 1)))); L3D_CUDA(ctx->d_row_job.ensure(size_t(rows) + 1));
    L3D_CUDA(ctx->d_l1.ensure(size_t(std::max<long long>(l1_words, 1))));
    L3D_CUDA(ctx->d_pa_cnt.ensure(size_t(rows) + 1)); L3D_CUDA(ctx->d_pa_off.ensure(size_t(rows) + 2)); L3D_CUDA(ctx->d_pa_raw.ensure(size_t(rows) + 1));
    L3D_CUDA(cudaMemsetAsync(ctx->d_pa_cnt.p, 0, sizeof(int) * (size_t(rows) + 1), st));
    L3D_CUDA(cudaEventRecord(ctx->pa_ev[0], st));
    if (rows > 0 && blocks > 0) {
        k_hoist<<<dim3((max_n + 127) / 128, n_jobs), 128, 0, st>>>(cx->d_segs.p, ctx->d_jobs.p, ctx->d_sh.p, ctx->d_sh1.p, ctx->d_th.p, ctx->d_row_job.p, cx->pretest_mode);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[0], st));
        k_pretest<4, 6><<<unsigned(l1_blocks), P1_BLOCK, bg_p1, st>>>(ctx->d_sh1.p, ctx->d_th.p, ctx->d_jobs.p, n_jobs, ctx->d_l1.p, p1_rows);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[1], st));
        k_pretest2<<<unsigned(blocks), PT_BLOCK, bg_p2, st>>>(ctx->d_sh.p, ctx->d_th.p, ctx->d_jobs.p, n_jobs, ctx->d_l1.p, ctx->d_cand.p, ctx->d_pa_cnt.p);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[4], st));
        ctx->pa_kev_set[0] = true;
        ctx->launches += 3;
        L3D_CUDA(cudaGetLastError());
    }
    if (scan_exclusive(ctx, ctx->d_pa_cnt.p, ctx->d_pa_off.p, int(rows))) return 1;
    L3D_CUDA(cudaEventRecord(ctx->pa_ev[1], st));
    // candidate totals per job (and overall) come back in one small copy
    L3D_CUDA(ctx->d_pa_idx.ensure(size_t(n_jobs) + 1)); L3D_CUDA(ctx->d_pa_tot.ensure(size_t(n_jobs) + 1)); L3D_CUDA(ctx->h_res.ensure(2 * size_t(n_jobs) + 64));
    {
        std::vector<int> idx(size_t(n_jobs) + 1);
        for (int j = 0; j < n_jobs; ++j) idx[j] = ctx->jobs_host[j].row_off;
        idx[n_jobs] = int(rows);
        L3D_CUDA(cudaMemcpyAsync(ctx->d_pa_idx.p, idx.data(), sizeof(int) * idx.size(), cudaMemcpyHostToDevice, st));
        k_gather_int<<<(n_jobs + 256) / 256, 256, 0, st>>>(ctx->d_pa_off.p, ctx->d_pa_idx.p, n_jobs + 1, ctx->d_pa_tot.p);
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(ctx->h_res.p, ctx->d_pa_tot.p, sizeof(int) * (size_t(n_jobs) + 1), cudaMemcpyDeviceToHost, st));
        // largest candidate count of a single row, per job
        std::vector<int> lens(static_cast<size_t>(n_jobs));
        for (int j = 0; j < n_jobs; ++j) lens[j] = ctx->jobs_host[j].n_src;
        L3D_CUDA(ctx->d_rm_begin.ensure(size_t(n_jobs))); L3D_CUDA(ctx->d_rm_len.ensure(size_t(n_jobs))); L3D_CUDA(ctx->d_rm_out.ensure(size_t(n_jobs)));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_rm_begin.p, idx.data(), sizeof(int) * size_t(n_jobs), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_rm_len.p, lens.data(), sizeof(int) * size_t(n_jobs), cudaMemcpyHostToDevice, st));
        k_range_max<<<n_jobs, 256, 0, st>>>(ctx->d_pa_cnt.p, ctx->d_rm_begin.p, ctx->d_rm_len.p, 0, ctx->d_rm_out.p);
        L3D_CUDA(ctx->d_job_sum64.ensure(size_t(n_jobs))); L3D_CUDA(ctx->h_job_sum64.ensure(size_t(n_jobs)));
        k_range_sum64<<<n_jobs, 256, 0, st>>>(ctx->d_pa_cnt.p, ctx->d_rm_begin.p, ctx->d_rm_len.p, ctx->d_job_sum64.p);
        ctx->launches += 2;
        L3D_CUDA(cudaMemcpyAsync(ctx->h_res.p + n_jobs + 1, ctx->d_rm_out.p, sizeof(int) * size_t(n_jobs), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->h_job_sum64.p, ctx->d_job_sum64.p, sizeof(unsigned long long) * size_t(n_jobs), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));        // idx / lens are stack vectors: must be consumed before they go away
    }
    {   // Candidate offsets are 32-bit: a batch whose candidates do not fit (a candidate rate far above the usual 2 %:
        // degenerate epipolar geometry, or the pre-test switched off) is reported so that the caller can cut it smaller.
        unsigned long long all = 0ull;
        for (int j = 0; j < n_jobs; ++j) all += ctx->h_job_sum64.p[j];
        if (all > cx->cand_limit) {
            char b[200]; snprintf(b, sizeof b, "pair tests: %llu candidates in one batch exceed the limit of %llu (32-bit candidate index)", all, (unsigned long long)cx->cand_limit);
            ctx->err = b;
            return 2;
        }
    }
    for (int j = 0; j < n_jobs; ++j) ctx->job_cands[j] = uint64_t(ctx->h_res.p[j + 1] - ctx->h_res.p[j]);
    for (int j = 0; j < n_jobs; ++j) ctx->job_rowmax[j] = ctx->h_res.p[n_jobs + 1 + j];
    const int n_cand = ctx->h_res.p[n_jobs];
    L3D_CUDA(ctx->d_stage_col.ensure(size_t(std::max(n_cand, 1)))); L3D_CUDA(ctx->d_stage_dep.ensure(size_t(std::max(n_cand, 1))));
    L3D_CUDA(ctx->d_stage_row.ensure(size_t(std::max(n_cand, 1)))); L3D_CUDA(ctx->d_stage_ok.ensure(size_t(std::max(n_cand, 1))));
    L3D_CUDA(cudaMemsetAsync(ctx->d_pa_raw.p, 0, sizeof(int) * (size_t(rows) + 1), st));
    if (rows > 0 && n_cand > 0) {
        k_expand<<<unsigned((rows * 32 + 255) / 256), 256, bg_ex, st>>>(ctx->d_jobs.p, ctx->d_row_job.p, int(rows), ctx->d_cand.p, ctx->d_pa_off.p, ctx->d_stage_col.p, ctx->d_stage_row.p);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[2], st));
        k_exact_flat<<<unsigned((n_cand + EF_T - 1) / EF_T), EF_T, bg_ef, st>>>(cx->d_segs.p, ctx->d_jobs.p, ctx->d_row_job.p, n_cand, ctx->d_stage_col.p, ctx->d_stage_row.p,
                                                                          ctx->d_stage_ok.p, ctx->d_stage_dep.p, ctx->d_pa_raw.p);
        L3D_CUDA(cudaEventRecord(ctx->pa_kev[3], st));
        ctx->pa_kev_set[1] = true;
        ctx->launches += 2;
        L3D_CUDA(cudaGetLastError());
    }
    L3D_CUDA(cudaEventRecord(ctx->pa_ev[2], st));
    ctx->pa_ms_pending = true;
    // raw matches per job and per row: tight bounds for the per-view work (the candidate counts are 2.5x larger).
    // Waiting for them here costs one idle gap per batch, not per view.
    if (rows > 0 && n_cand > 0) {
        L3D_CUDA(ctx->d_pa_idx.ensure(2 * size_t(n_jobs) + 2));
        k_range_sum_max<<<n_jobs, 256, 0, st>>>(ctx->d_pa_raw.p, ctx->d_rm_begin.p, ctx->d_rm_len.p, ctx->d_pa_idx.p, ctx->d_pa_idx.p + n_jobs);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(ctx->h_res.p, ctx->d_pa_idx.p, sizeof(int) * 2 * size_t(n_jobs), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        for (int j = 0; j < n_jobs; ++j) { ctx->job_raw[j] = uint64_t(ctx->h_res.p[j]); ctx->job_raw_rowmax[j] = ctx->h_res.p[n_jobs + j]; }
    }
    return 0;
}

bool same_neighbors(const std::vector<l3d_neighbor>& a, int n, const l3d_neighbor* b)
{
    if (int(a.size()) != n) return false;
    for (int c = 0; c < n; ++c)
        if (a[c].view != b[c].view || (a[c].to_be_matched != 0) != (b[c].to_be_matched != 0) || memcmp(a[c].F, b[c].F, sizeof a[c].F) != 0) return false;
    return true;
}

// Phase B of a group of views (all of one dependency level, see VArgs), split in two so that the next group can be
// put on the stream before the host waits for this one: enqueue_group does the host's bookkeeping for every view,
// uploads the group's tables and arguments once and launches every kernel once (sizes come from upper bounds, the
// real counts stay on the device); finish_group waits for the group's result records and fetches the views' output.
struct GroupView { int view; int n_nb; const l3d_neighbor* nb; int first_job; size_t plan_index; };

// Several GPUs on one scene: how many records the views of a group can put out at most, all ranks together (what the
// record window has to hold) -- the same bound enqueue_group computes, without touching anything.
size_t mr_group_need(l3d_ctx* ctx, const GroupView* gv, int count)
{
    PhaseASet& pa = ctx->pa[ctx->cur];
    size_t need = MR_HEADER_BYTES;
    for (int g = 0; g < count; ++g) {
        if (gv[g].view < 0 || gv[g].view >= ctx->V) continue;
        int n_slots = 0; size_t f_ub = 0;
        for (int c = 0; c < gv[g].n_nb; ++c) {
            const l3d_neighbor& n = gv[g].nb[c];
            if (n.to_be_matched) { const size_t j = size_t(gv[g].first_job + n_slots++); if (j < pa.job_raw_all.size()) f_ub += size_t(pa.job_raw_all[j]); }
            else {
                const std::vector<Delivered>& dl = ctx->delivered[gv[g].view];
                for (size_t k = 0; k < dl.size(); ++k) if (dl[k].from_view == n.view) { f_ub += size_t(dl[k].count_all); break; }
            }
        }
        need += (size_t(ctx->views[gv[g].view].n_segments) * 8 + 255) & ~size_t(255);
        need += 2 * size_t(ctx->world) * std::max<size_t>(f_ub, 1) * sizeof(l3d_match);
    }
    return need;
}

int enqueue_group(l3d_ctx* ctx, ViewGroup& gr, const GroupView* gv, int count, float sigma_p, float sigma_a)
{
    cudaStream_t st = ctx->stream;
    PhaseASet& pa = ctx->pa[ctx->cur];
    const bool mr = ctx->world > 1;
    gr.count = count;
    L3D_CUDA(gr.h_args.ensure(VG_MAX)); L3D_CUDA(gr.d_args.ensure(VG_MAX));
    if (mr) {
        L3D_CUDA(gr.h_allrec.ensure(size_t(VG_MAX) * MR_MAX_RANKS));
        // the records of the group two groups ago may still be on their way to the host out of this group's buffers
        if (gr.ev_x[1]) L3D_CUDA(cudaStreamWaitEvent(st, gr.ev_x[1], 0));
    }
    int res_stride = 8;
    for (int g = 0; g < count; ++g) { int nd = 0; for (int c = 0; c < gv[g].n_nb; ++c) nd += gv[g].nb[c].deliver ? 1 : 0; res_stride = std::max(res_stride, 8 + nd); }
    gr.res_stride = res_stride;
    L3D_CUDA(gr.d_vres.ensure(size_t(VG_MAX) * res_stride)); L3D_CUDA(gr.h_vres.ensure(size_t(VG_MAX) * res_stride));
    size_t cams_total = 0;
    for (int g = 0; g < count; ++g) {
        if (gv[g].view < 0 || gv[g].view >= ctx->V) { ctx->err = "l3d_match_view: view index out of range"; return 1; }
        for (int c = 0; c < gv[g].n_nb; ++c) {
            if (gv[g].nb[c].view < 0 || gv[g].nb[c].view >= ctx->V || gv[g].nb[c].view == gv[g].view) { ctx->err = "l3d_match_view: bad neighbour view index"; return 1; }
            if (c > 0 && gv[g].nb[c - 1].view >= gv[g].nb[c].view) { ctx->err = "l3d_match_view: neighbours must be in ascending view order"; return 1; }
        }
        gr.cam_off[g] = cams_total; cams_total += size_t(gv[g].n_nb);
    }
    gr.cam_off[count] = cams_total;
    L3D_CUDA(gr.h_cams.ensure(cams_total)); L3D_CUDA(gr.d_cams.ensure(cams_total));

    // ---- per view: tables, bounds, buffers, hand-over records, kernel arguments ------------------------------
    size_t zero_total = 0;
    int mx_src = 1, mx_cams = 1, mx_items = 1, mx_cta = 1, mn_sc_h = SC_T; long mx_warps = 1; size_t mx_M = 0, mx_smem = 0;
    for (int g = 0; g < count; ++g) mn_sc_h = std::min(mn_sc_h, gv[g].n_nb <= 192 ? 128 : (gv[g].n_nb <= 384 ? 64 : 32));
    if (ctx->score_block) mn_sc_h = std::min(mn_sc_h, ctx->score_block);
    bool any_scored = false, any_window = false, any_quadratic = false, any_deliver = false;
    struct Z { size_t items, iters, best, hist, cursor, end; } z[VG_MAX];
    size_t hist_total = 0, window_need = MR_HEADER_BYTES;
    const int slot_q = int(&gr - ctx->grp);
    memset(&gr.lay, 0, sizeof gr.lay);
    for (int g = 0; g < count; ++g) {
        const int view = gv[g].view, n_nb = gv[g].n_nb; const l3d_neighbor* nb = gv[g].nb; const int first_job = gv[g].first_job;
        ViewSlot& sl = gr.slot[g];
        ViewJob& job = sl.job;
        sl.plan_index = gv[g].plan_index;
        const l3d_view& sv = ctx->views[view];
        const int n_full = sv.n_segments;
        const int n_src = local_rows(ctx, n_full);         // the rows of the view this process works on
        CamTable* hc = gr.h_cams.p + gr.cam_off[g];
        job.view = view; job.nb.assign(nb, nb + n_nb); job.sigma_p = sigma_p; job.sigma_a = sigma_a;
        job.consumed.clear(); job.owner = -1;
        int n_slots = 0, n_deliver = 0, max_deliver_seg = 1;
        uint64_t pair_tests = 0, cand_ub = 0, raw_ub = 0, raw_ub_all = 0, delivered_ub = 0, delivered_ub_all = 0;      // raw_ub_all: over all ranks' rows
        size_t seg_ub = 0;                               // upper bound on the matches of one source segment
        std::vector<int> consumed;                       // indices into ctx->delivered[view]
        for (int c = 0; c < n_nb; ++c) {
            const l3d_neighbor& n = nb[c];
            const l3d_view& tv = ctx->views[n.view];
            CamTable& ct = hc[c];
            memcpy(ct.F, n.F, sizeof ct.F); memcpy(ct.RtKinv, tv.RtKinv, sizeof ct.RtKinv);
            memcpy(ct.P, tv.P, sizeof ct.P); memcpy(ct.C, tv.C, sizeof ct.C);
            ct.seg_off = ctx->seg_off[n.view]; ct.n_seg = tv.n_segments;
            ct.t_start = nullptr; ct.t_tgt = nullptr; ct.t_dep = nullptr; ct.slot = -1; ct.deliver = -1; ct.row_off = 0;
            if (n.to_be_matched) {
                const int j = first_job + n_slots;
                ct.mode = 0; ct.slot = n_slots++; ct.row_off = pa.jobs_host[j].row_off;
                cand_ub += pa.job_cands[j]; raw_ub += pa.job_raw[j]; raw_ub_all += pa.job_raw_all[j]; seg_ub += size_t(pa.job_raw_rowmax[j]);
                pair_tests += uint64_t(n_src) * uint64_t(tv.n_segments);
            } else {
                ct.mode = 2;
                std::vector<Delivered>& dl = ctx->delivered[view];
                for (size_t k = 0; k < dl.size(); ++k)
                    if (dl[k].from_view == n.view) {
                        ct.mode = 1; ct.t_start = dl[k].t_start; ct.t_tgt = dl[k].t_tgt; ct.t_dep = dl[k].t_dep;
                        delivered_ub += uint64_t(dl[k].count); delivered_ub_all += uint64_t(dl[k].count_all); seg_ub += size_t(dl[k].max_group);
                        consumed.push_back(int(k)); job.consumed.push_back(std::make_pair(dl[k].owner, dl[k].slot));
                        break;
                    }
            }
            if (n.deliver) { ct.deliver = n_deliver++; max_deliver_seg = std::max(max_deliver_seg, tv.n_segments); }
        }
        job.pair_tests = pair_tests; job.cand_ub = cand_ub;
        const size_t M_ub = size_t(raw_ub + delivered_ub);          // upper bound on the number of matches to score
        const size_t F_ub_all = size_t(raw_ub_all + (mr ? delivered_ub_all : delivered_ub));  // ... and on the output records of any rank (the same number on every rank)
        if (mr) {                                                    // the view's places in the group's window
            gr.lay.samp_off[g] = (long long)window_need; window_need += (size_t(n_full) * 8 + 255) & ~size_t(255);
            gr.lay.cap[g] = (long long)std::max<size_t>(F_ub_all, 1);
            gr.lay.rec_off[g] = (long long)window_need; window_need += size_t(ctx->world) * size_t(gr.lay.cap[g]) * sizeof(l3d_match);
            gr.lay.full_off[g] = (long long)window_need; window_need += size_t(ctx->world) * size_t(gr.lay.cap[g]) * sizeof(l3d_match);
        }
        const bool scored = n_slots > 0;                             // cudawrapper.cu:877-878
        const bool delivering = n_deliver > 0 && scored;
        const int stride = max_deliver_seg + 1;                      // one spare key per slot keeps t_start[n_seg] addressable
        const int n_keys = delivering ? n_deliver * stride : 0;
        const bool windowed = sv.spatial_k > 0.0f;
        const int SORT_MAX = 1 << 20;                                // the window kernel's index has 20 bits
        const bool quadratic = scored && (!windowed || seg_ub > size_t(SORT_MAX));
        // hypotheses per scoring block: the per-(camera, hypothesis) maxima must fit in shared memory whatever the number
        // of neighbours (k = -1: every co-visible view)
        const int sc_h = mn_sc_h;                        // (one launch covers the group: the same for all its views)
        const int max_items = int(M_ub / size_t(sc_h)) + n_src + 1;
        const int n_cta = int(std::max<size_t>((M_ub + RM_T - 1) / RM_T, 1));
        const size_t sm_bytes = size_t(n_nb) * size_t(sc_h) * sizeof(int);
        if (sm_bytes > 200 * 1024) { ctx->err = "l3d_match_view: more than 1600 neighbours in one view"; return 1; }
        // buffers of this view
        L3D_CUDA(sl.d_cnt.ensure(size_t(n_src) * n_nb + 1)); L3D_CUDA(sl.d_off.ensure(size_t(n_src) * n_nb + 2));
        L3D_CUDA(sl.d_mids.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_mdep.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_score.ensure(std::max<size_t>(M_ub, 1)));
        L3D_CUDA(sl.d_filtered.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_cta.ensure(size_t(n_cta) + 1)); L3D_CUDA(sl.d_F.ensure(4));
        if (!mr) L3D_CUDA(sl.d_dsamples.ensure(size_t(n_src) * 2 + 2));
        L3D_CUDA(sl.d_items.ensure(size_t(2 * max_items)));
        if (scored && windowed) {
            L3D_CUDA(sl.d_sk.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_sd2.ensure(std::max<size_t>(M_ub, 1))); L3D_CUDA(sl.d_spl.ensure(std::max<size_t>(M_ub, 1)));
            L3D_CUDA(sl.d_boff.ensure(size_t(n_src) * (BK_B + 1))); L3D_CUDA(sl.d_brange.ensure(size_t(std::max(n_src, 1))));
        }
        // what has to start from zero: one region per view in the group's zero buffer, then the hand-over group sizes of
        // all views (one block: with several GPUs on one scene they are summed over the ranks in one all-reduce)
        z[g].items = zero_total; z[g].iters = z[g].items + 4; z[g].best = z[g].iters + 4;
        z[g].cursor = z[g].best + 2 * size_t(n_src) + 2; z[g].end = (z[g].cursor + size_t(n_keys) + 2 + 3) & ~size_t(3);
        zero_total = z[g].end;
        z[g].hist = hist_total; hist_total += (size_t(n_keys) + 2 + 3) & ~size_t(3);
        // hand-over buffers
        DeliverBuf buf;
        if (delivering) {
            buf.refs = n_deliver;
            L3D_CUDA(ctx->pool.get(sizeof(int) * (size_t(n_keys) + 2), reinterpret_cast<void**>(&buf.start)));
            // (several GPUs on one scene: every rank holds the complete hand-over lists, all ranks' records)
            L3D_CUDA(ctx->pool.get(sizeof(int) * std::max<size_t>(mr ? F_ub_all : M_ub, 1), reinterpret_cast<void**>(&buf.tgt)));
            L3D_CUDA(ctx->pool.get(sizeof(float4) * std::max<size_t>(mr ? F_ub_all : M_ub, 1), reinterpret_cast<void**>(&buf.dep)));
        }
        job.M_ub = M_ub; job.scored = scored; job.delivering = delivering; job.n_deliver = n_deliver;
        // kernel arguments (pointers into the zero buffer are filled in below, when it has its size)
        VArgs& va = gr.h_args.p[g];
        memset(&va, 0, sizeof va);
        va.cams = gr.d_cams.p + gr.cam_off[g]; va.n_cams = n_nb; va.n_src = n_src;
        memcpy(va.src.RtKinv, sv.RtKinv, sizeof va.src.RtKinv); memcpy(va.src.C, sv.C, sizeof va.src.C);
        va.src.spatial_k = sv.spatial_k; va.src.cmax = std::max(std::fabs(sv.C[0]), std::max(std::fabs(sv.C[1]), std::fabs(sv.C[2])));
        va.src.seg_off = ctx->seg_off[view]; va.src.n_seg = n_src;
        va.src.row_first = mr ? ctx->rank : 0; va.src.row_stride = mr ? ctx->world : 1;
        va.mr = mr ? 1 : 0; va.n_seg_full = n_full;
        va.sp.sigma_p = sigma_p; va.sp.sigma_a = sigma_a;
        va.raw_cnt = pa.d_pa_raw.p; va.cand_off = pa.d_pa_off.p; va.stage_col = pa.d_stage_col.p; va.stage_dep = pa.d_stage_dep.p; va.stage_ok = pa.d_stage_ok.p;
        va.cnt = sl.d_cnt.p; va.off = sl.d_off.p; va.mids = sl.d_mids.p; va.mdep = sl.d_mdep.p; va.score = sl.d_score.p;
        va.sk = sl.d_sk.p; va.sd2 = sl.d_sd2.p; va.spl = sl.d_spl.p; va.boff = sl.d_boff.p; va.brange = sl.d_brange.p; va.items = sl.d_items.p; va.max_items = max_items;
        va.sc_h = sc_h;
        va.scored = scored ? 1 : 0; va.windowed = windowed ? 1 : 0; va.quadratic = quadratic ? 1 : 0; va.m_hi = SORT_MAX;
        va.cta_kept = sl.d_cta.p; va.n_cta = n_cta; va.Fout = sl.d_F.p; va.filtered = sl.d_filtered.p; va.dsamples = sl.d_dsamples.p;
        va.delivering = delivering ? 1 : 0; va.stride = stride; va.n_keys = n_keys; va.n_deliver = n_deliver;
        va.dl_start = buf.start; va.dl_tgt = buf.tgt; va.dl_dep = buf.dep;
        va.vres = gr.d_vres.p + size_t(g) * res_stride;
        // the hand-over records exist from now on (stream order); until this view is collected their sizes are bounds
        if (delivering) {
            job.owner = int(ctx->deliver_bufs.size());
            buf.counts.assign(size_t(n_deliver), 0); buf.maxg.assign(size_t(n_deliver), 0);
            for (int c = 0; c < n_nb; ++c) {
                const int d = hc[c].deliver;
                if (d < 0) continue;
                const uint64_t ub = std::min<uint64_t>(hc[c].slot >= 0 ? pa.job_raw_all[first_job + hc[c].slot] : uint64_t(mr ? F_ub_all : M_ub), 0x7fffffffull);
                Delivered dl; dl.from_view = view; dl.owner = job.owner; dl.slot = d;
                dl.count = int(ub); dl.max_group = int(std::min<uint64_t>(ub, uint64_t(n_src)));      // a source segment matches a target segment once
                dl.count_all = (long long)ub;
                if (mr) dl.max_group = int(std::min<uint64_t>(ub, uint64_t(n_full)));
                dl.t_start = buf.start + size_t(d) * stride; dl.t_tgt = buf.tgt; dl.t_dep = buf.dep;
                buf.counts[d] = dl.count; buf.maxg[d] = dl.max_group;
                ctx->delivered[nb[c].view].push_back(dl);
            }
            ctx->deliver_bufs.push_back(buf);
        }
        // the consumed deliveries leave the list; their memory is given back when this view has been collected
        {
            std::vector<Delivered>& dl = ctx->delivered[view];
            for (size_t k = 0; k < consumed.size(); ++k) dl[consumed[k]].from_view = -1;
            std::vector<Delivered> keep;
            for (size_t k = 0; k < dl.size(); ++k) if (dl[k].from_view >= 0) keep.push_back(dl[k]);
            dl.swap(keep);
        }
        job.active = true;
        mx_src = std::max(mx_src, n_src); mx_cams = std::max(mx_cams, n_nb); mx_warps = std::max(mx_warps, long(n_src) * n_nb);
        mx_M = std::max(mx_M, M_ub); mx_cta = std::max(mx_cta, n_cta);
        if (scored) { any_scored = true; mx_items = std::max(mx_items, max_items); mx_smem = std::max(mx_smem, sm_bytes); }
        any_window |= scored && windowed; any_quadratic |= quadratic; any_deliver |= delivering;
    }
    gr.hist_off = zero_total; gr.hist_total = hist_total;
    zero_total += hist_total;
    const size_t routed_off = zero_total;                     // several GPUs: records routed per (view, destination rank), and a fault word
    zero_total += size_t(VG_MAX) * MR_MAX_RANKS + 4;
    L3D_CUDA(gr.d_zero.ensure(zero_total + 4));
    if (mr && window_need > ctx->win_cap) { ctx->err = "l3d_match_view: internal error (window too small)"; return 1; }
    for (int g = 0; g < count; ++g) {
        VArgs& va = gr.h_args.p[g];
        va.nitems = gr.d_zero.p + z[g].items; va.iters = reinterpret_cast<unsigned long long*>(gr.d_zero.p + z[g].iters);
        va.best = reinterpret_cast<unsigned long long*>(gr.d_zero.p + z[g].best); va.hist = gr.d_zero.p + gr.hist_off + z[g].hist; va.cursor = gr.d_zero.p + z[g].cursor;
        if (mr) va.dsamples = reinterpret_cast<float*>(ctx->win[slot_q] + gr.lay.samp_off[g]);
    }
    if (mx_smem > ctx->score_smem) {
        L3D_CUDA(cudaFuncSetAttribute(k_score, cudaFuncAttributeMaxDynamicSharedMemorySize, int(mx_smem)));
        L3D_CUDA(cudaFuncSetAttribute(k_score_window, cudaFuncAttributeMaxDynamicSharedMemorySize, int(mx_smem)));
        ctx->score_smem = mx_smem;
    }

    // ---- one upload, one memset, one launch of every kernel for the whole group ------------------------------
    const unsigned G = unsigned(count);
    const VArgs* A = gr.d_args.p;
    L3D_CUDA(cudaMemcpyAsync(gr.d_cams.p, gr.h_cams.p, sizeof(CamTable) * cams_total, cudaMemcpyHostToDevice, st));
    L3D_CUDA(cudaMemcpyAsync(gr.d_args.p, gr.h_args.p, sizeof(VArgs) * size_t(count), cudaMemcpyHostToDevice, st));
    L3D_CUDA(cudaMemsetAsync(gr.d_zero.p, 0, sizeof(int) * zero_total, st));
    gr.n_kev = 0;
    // (per-kernel events, only when asked for: l3d_debug_chain_times)
#define L3D_KEV(slot) do { if (ctx->chain_prof && ctx->stage_events && gr.n_kev < 24) { if (!gr.kev[gr.n_kev]) cudaEventCreate(&gr.kev[gr.n_kev]); cudaEventRecord(gr.kev[gr.n_kev], st); gr.kev_slot[gr.n_kev] = (slot); ++gr.n_kev; } } while (0)
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[2], st));
    L3D_KEV(0);
    L3D_CUDA(pdl_launch(k_gather_counts, dim3((mx_src + 255) / 256, mx_cams, G), dim3(256), 0, st, A));
    L3D_KEV(1);
    L3D_CUDA(pdl_launch(k_scan_views, dim3(1, G), dim3(SCAN_T), 0, st, A, 0));
    L3D_KEV(2);
    L3D_CUDA(pdl_launch(k_fill, dim3(unsigned((mx_warps * 32 + 255) / 256), G), dim3(256), 0, st, A, ctx->d_segs.p));
    ctx->launches += 3;
    L3D_KEV(3);
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[3], st));
    if (any_window) {
        L3D_CUDA(pdl_launch(k_bucket_segments, dim3(mx_src, G), dim3(BK_T), 0, st, A));
        L3D_KEV(4);
        L3D_CUDA(pdl_launch(k_score_window, dim3(mx_items, G), dim3(mn_sc_h), mx_smem, st, A, ctx->d_segs.p));
        ctx->launches += 2;
    }
    if (any_quadratic) {
        L3D_CUDA(pdl_launch(k_score_items, dim3((mx_src + 255) / 256, G), dim3(256), 0, st, A));
        L3D_CUDA(pdl_launch(k_score, dim3(mx_items, G), dim3(mn_sc_h), mx_smem, st, A, ctx->d_segs.p));
        ctx->launches += 2;
    }
    L3D_KEV(5);
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[4], st));
    if (any_scored) { L3D_CUDA(pdl_launch(k_reduce_matches, dim3(mx_cta, G), dim3(RM_T), 0, st, A)); ctx->launches += 1; }
    L3D_KEV(6);
    L3D_CUDA(pdl_launch(k_filter_compact, dim3(mx_cta, G), dim3(RM_T), 0, st, A));
    ctx->launches += 1;
    L3D_KEV(7);
    if (mr) {
        // Every rank stores its records, depth samples and counts into all ranks' windows (regions fixed by shared
        // bounds), a one-word all-reduce tells that everybody's stores have landed, and the hand-over lists and the
        // medians are made from the own window.
        const unsigned pb = unsigned(std::min<size_t>(std::max<size_t>((mx_M + 255) / 256, 1), size_t(4 * ctx->n_sm)));
        const bool collect = ctx->want_records;
        const int root = !collect ? -1 : (ctx->result_root < 0 ? -2 : ctx->result_root);
        int* routed = gr.d_zero.p + routed_off; int* fault = routed + VG_MAX * MR_MAX_RANKS;
        const unsigned seq = ++ctx->win_seq[slot_q];
        const unsigned char* window = ctx->win[slot_q];
        L3D_CUDA(pdl_launch(k_mr_push, dim3(pb, G), dim3(256), 0, st, A, gr.lay, ctx->peers[slot_q], ctx->rank, ctx->world, root, routed));
        L3D_KEV(8);
        L3D_CUDA(pdl_launch(k_mr_signal, dim3(1), dim3(VG_MAX * MR_MAX_RANKS), 0, st, A, int(count), ctx->peers[slot_q], ctx->rank, ctx->world, (const int*)routed, seq));
        L3D_CUDA(pdl_launch(k_mr_wait, dim3(1), dim3(32), 0, st, window, ctx->world, seq, fault));
        ctx->launches += 3;
        L3D_KEV(9);
        gr.d_fault = fault;
        if (any_deliver) {
            const unsigned hb = unsigned(std::min<size_t>(std::max<size_t>(pb, 1), size_t(4 * ctx->n_sm)));
            L3D_CUDA(pdl_launch(k_mr_hist, dim3(hb, G), dim3(256), 0, st, A, window, gr.lay, ctx->world));
            L3D_KEV(10);
            L3D_CUDA(pdl_launch(k_scan_views, dim3(1, G), dim3(SCAN_T), 0, st, A, 1));
            L3D_KEV(11);
            L3D_CUDA(pdl_launch(k_mr_deliver_scatter, dim3(hb, G), dim3(256), 0, st, A, window, gr.lay, ctx->world));
            ctx->launches += 3;
            L3D_KEV(12);
        }
        L3D_CUDA(pdl_launch(k_finish_view, dim3(1, G), dim3(MS_T), 0, st, A));
        ctx->launches += 1;
        L3D_KEV(13);
        if (collect && (ctx->result_root < 0 || ctx->result_root == ctx->rank)) {
            // for the copy to the host: the ranks' regions of each view put end to end.  On this stream, so that the
            // window is only ever read in stream order (the other ranks write the group after next into it as soon as
            // this rank has joined the next group's all-reduce).
            size_t bound = 0;
            for (int g = 0; g < count; ++g) bound += size_t(gr.lay.cap[g]);
            L3D_CUDA(gr.d_gather.ensure(bound));
            k_mr_gather<<<dim3(std::min(2 * pb, unsigned(4 * ctx->n_sm)), G), 256, 0, st>>>(window, gr.lay, ctx->world, gr.d_gather.p);
            ctx->launches += 1;
            L3D_KEV(14);
        }
        L3D_CUDA(cudaEventRecord(gr.ev_x[0], st));      // the gathered records are complete
        L3D_CUDA(cudaGetLastError());
        if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[5], st));
        L3D_CUDA(cudaMemcpyAsync(gr.h_vres.p, gr.d_vres.p, sizeof(int) * size_t(count) * res_stride, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaMemcpyAsync(gr.h_allrec.p, window, sizeof(RankRec) * VG_MAX * size_t(ctx->world), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaMemcpyAsync(&gr.h_allrec.p[size_t(VG_MAX) * MR_MAX_RANKS - 1].pad, gr.d_fault, sizeof(int), cudaMemcpyDeviceToHost, st));      // (a spare word of the pinned table)
        L3D_CUDA(cudaEventRecord(gr.ev[0], st));
        return 0;
    }
    if (any_deliver) {
        L3D_CUDA(pdl_launch(k_scan_views, dim3(1, G), dim3(SCAN_T), 0, st, A, 1));
        L3D_CUDA(pdl_launch(k_deliver_scatter, dim3(unsigned((std::max<size_t>(mx_M, 1) + 255) / 256), G), dim3(256), 0, st, A));
        ctx->launches += 2;
        L3D_KEV(12);
    }
    // The last kernel (medians, result records) and the download of the records are needed by the host only: the next
    // group does not wait for them.  They go to a side stream (this group's buffers are not touched again before the
    // host has collected it), except when stages are being timed.
    cudaStream_t fs = st;
    if (!ctx->stage_events) {
        fs = ctx->fin_stream;
        L3D_CUDA(cudaEventRecord(gr.ev[1], st));
        L3D_CUDA(cudaStreamWaitEvent(fs, gr.ev[1], 0));
    }
    L3D_CUDA(pdl_launch(k_finish_view, dim3(1, G), dim3(MS_T), 0, fs, A));
    ctx->launches += 1;
    L3D_KEV(13);
    L3D_CUDA(cudaGetLastError());
    if (ctx->stage_events) L3D_CUDA(cudaEventRecord(gr.ev[5], st));
    L3D_CUDA(cudaMemcpyAsync(gr.h_vres.p, gr.d_vres.p, sizeof(int) * size_t(count) * res_stride, cudaMemcpyDeviceToHost, fs));
    L3D_CUDA(cudaEventRecord(gr.ev[0], fs));
    return 0;
}

// n records of pinned host memory that stay where they are until l3d_reset_matching
int arena_take(l3d_ctx* ctx, size_t n, l3d_match** out)
{
    while (ctx->arena_chunk < ctx->arena.size() && ctx->arena_used + n > ctx->arena[ctx->arena_chunk].second) { ++ctx->arena_chunk; ctx->arena_used = 0; }
    if (ctx->arena_chunk == ctx->arena.size()) {
        const size_t cap = std::max<size_t>(n, size_t(4) << 20);          // 128 MiB chunks
        l3d_match* p = nullptr;
        L3D_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&p), cap * sizeof(l3d_match), cudaHostAllocDefault));
        ctx->arena.push_back(std::make_pair(p, cap)); ctx->arena_used = 0;
    }
    *out = ctx->arena[ctx->arena_chunk].first + ctx->arena_used;
    ctx->arena_used += n;
    return 0;
}

// Several GPUs on one scene: the record windows.  Every rank allocates two (one per group in flight), exports them
// (CUDA IPC) and maps every other rank's; k_mr_push then stores a group's records into all of them.  Called by all
// ranks at the same point with the same size, with no group in flight.
int mr_ensure_window(l3d_ctx* ctx, size_t need)
{
    if (need <= ctx->win_cap) return 0;
    cudaStream_t st = ctx->stream;
    const int W = ctx->world;
    L3D_CUDA(ctx->d_bar.ensure(4)); L3D_CUDA(ctx->d_ipc.ensure(size_t(2) * sizeof(cudaIpcMemHandle_t) * W)); L3D_CUDA(ctx->h_ipc.ensure(size_t(2) * sizeof(cudaIpcMemHandle_t) * W));
    L3D_CUDA(cudaStreamSynchronize(st));
    for (int q = 0; q < 2; ++q)
        for (int r = 0; r < W; ++r) { if (r != ctx->rank && ctx->peers[q].base[r]) cudaIpcCloseMemHandle(ctx->peers[q].base[r]); ctx->peers[q].base[r] = nullptr; }
    // nobody frees a window that another rank still has mapped
    L3D_CUDA(cudaMemsetAsync(ctx->d_bar.p, 0, sizeof(int), st));
    L3D_NCCL(ctx->nccl->AllReduce(ctx->d_bar.p, ctx->d_bar.p, 1, ncclInt, ncclSum, ctx->comm, st));
    L3D_CUDA(cudaStreamSynchronize(st));
    const size_t cap = std::max<size_t>(need + need / 4, size_t(64) << 20);      // bytes; at least 64 MiB
    cudaIpcMemHandle_t mine[2];
    for (int q = 0; q < 2; ++q) {
        if (ctx->win[q]) cudaFree(ctx->win[q]);
        ctx->win[q] = nullptr;
        L3D_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->win[q]), cap));
        L3D_CUDA(cudaMemsetAsync(ctx->win[q], 0, MR_HEADER_BYTES, st));          // the arrival flags start below every sequence number
        L3D_CUDA(cudaIpcGetMemHandle(&mine[q], ctx->win[q]));
    }
    ctx->win_cap = 0;
    unsigned char* d_send = ctx->d_ipc.p; unsigned char* d_recv = ctx->d_ipc.p;      // in place: each rank's part at its own offset
    L3D_CUDA(cudaMemcpyAsync(d_send + size_t(ctx->rank) * sizeof mine, mine, sizeof mine, cudaMemcpyHostToDevice, st));
    L3D_NCCL(ctx->nccl->AllGather(d_send + size_t(ctx->rank) * sizeof mine, d_recv, sizeof mine, ncclChar, ctx->comm, st));
    L3D_CUDA(cudaMemcpyAsync(ctx->h_ipc.p, d_recv, sizeof mine * size_t(W), cudaMemcpyDeviceToHost, st));
    L3D_CUDA(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r)
        for (int q = 0; q < 2; ++q) {
            if (r == ctx->rank) { ctx->peers[q].base[r] = ctx->win[q]; continue; }
            cudaIpcMemHandle_t h; memcpy(&h, ctx->h_ipc.p + (size_t(r) * 2 + q) * sizeof h, sizeof h);
            void* ptr = nullptr;
            L3D_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
            ctx->peers[q].base[r] = static_cast<unsigned char*>(ptr);
        }
    ctx->win_cap = cap;
    return 0;
}

void mr_release_window(l3d_ctx* ctx)
{
    for (int q = 0; q < 2; ++q) {
        for (int r = 0; r < MR_MAX_RANKS; ++r) { if (r != ctx->rank && ctx->peers[q].base[r]) cudaIpcCloseMemHandle(ctx->peers[q].base[r]); ctx->peers[q].base[r] = nullptr; }
    }
    if (ctx->comm && ctx->d_bar.p && !ctx->comm_broken) {        // nobody frees a window that another rank still has mapped
        cudaMemsetAsync(ctx->d_bar.p, 0, sizeof(int), ctx->stream);
        ctx->nccl->AllReduce(ctx->d_bar.p, ctx->d_bar.p, 1, ncclInt, ncclSum, ctx->comm, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
    }
    for (int q = 0; q < 2; ++q) { if (ctx->win[q]) cudaFree(ctx->win[q]); ctx->win[q] = nullptr; }
    ctx->win_cap = 0;
}

// Raw matches per job summed over the ranks: the bounds every rank sizes the views' buffers with must cover what the
// other ranks may hand over.  One small all-reduce per batch of pair tests, on the main stream.
int mr_job_totals(l3d_ctx* ctx, PhaseASet& pa)
{
    const size_t n = pa.job_raw.size();
    pa.job_raw_all = pa.job_raw;
    if (ctx->world <= 1 || n == 0) return 0;
    L3D_CUDA(ctx->d_tot.ensure(n)); L3D_CUDA(ctx->h_tot.ensure(n));
    for (size_t j = 0; j < n; ++j) ctx->h_tot.p[j] = pa.job_raw[j];
    L3D_CUDA(cudaMemcpyAsync(ctx->d_tot.p, ctx->h_tot.p, sizeof(unsigned long long) * n, cudaMemcpyHostToDevice, ctx->stream));
    L3D_NCCL(ctx->nccl->AllReduce(ctx->d_tot.p, ctx->d_tot.p, n, ncclUint64, ncclSum, ctx->comm, ctx->stream));
    L3D_CUDA(cudaMemcpyAsync(ctx->h_tot.p, ctx->d_tot.p, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost, ctx->stream));
    L3D_CUDA(cudaStreamSynchronize(ctx->stream));
    for (size_t j = 0; j < n; ++j) pa.job_raw_all[j] = ctx->h_tot.p[j];
    return 0;
}

// Waits for a group, turns its result records into l3d_match_results (results[plan index], or *single for a view
// matched outside a schedule) and fetches the output records of all its views with one synchronise.
int finish_group(l3d_ctx* ctx, ViewGroup& gr, bool copy_back, l3d_match_result* single)
{
    const auto t_in = std::chrono::steady_clock::now();
    L3D_CUDA(cudaEventSynchronize(gr.ev[0]));
    const auto t_waited = std::chrono::steady_clock::now();
    ctx->host_s[1] += std::chrono::duration<double>(t_waited - t_in).count();
    struct Tail { l3d_ctx* c; std::chrono::steady_clock::time_point t; ~Tail() { c->host_s[2] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t).count(); } } tail = {ctx, t_waited};
    PhaseASet& pa = ctx->pa[ctx->cur];
    float ms_a[2] = {0.f, 0.f};
    if (pa.pa_ms_pending) {                            // phase A of a whole batch is booked on the first view collected after it
        cudaEventElapsedTime(&ms_a[0], pa.pa_ev[0], pa.pa_ev[1]); cudaEventElapsedTime(&ms_a[1], pa.pa_ev[1], pa.pa_ev[2]);
        for (int k = 0; k < 2; ++k) { float ms = 0.f; if (pa.pa_kev_set[k]) cudaEventElapsedTime(&ms, pa.pa_kev[2 * k], pa.pa_kev[2 * k + 1]); ctx->pa_kernel_ms[k] = ms; }
        { float ms = 0.f; if (pa.pa_kev_set[0]) cudaEventElapsedTime(&ms, pa.pa_kev[1], pa.pa_kev[4]); ctx->pa_kernel_ms[2] = ms; }
        pa.pa_ms_pending = false;
    } else ctx->pa_kernel_ms[0] = ctx->pa_kernel_ms[1] = ctx->pa_kernel_ms[2] = 0.f;
    float ms_b[3] = {0.f, 0.f, 0.f};
    if (ctx->stage_events) for (int k = 0; k < 3; ++k) cudaEventElapsedTime(&ms_b[k], gr.ev[k + 2], gr.ev[k + 3]);
    for (int k = 1; k < gr.n_kev; ++k) { float ms = 0.f; cudaEventElapsedTime(&ms, gr.kev[k - 1], gr.kev[k]); ctx->chain_ms[gr.kev_slot[k]] += double(ms); }
    gr.n_kev = 0;
    const bool mr = ctx->world > 1;
    int view_off[VG_MAX], view_cnt[VG_MAX];              // several GPUs: where each view's records are in this group's window
    if (mr) {
        long long total = 0; int max_cnt = 0;
        for (int v = 0; v < gr.count; ++v) {
            view_off[v] = int(total);
            for (int r = 0; r < ctx->world; ++r) {
                const int f = gr.h_allrec.p[size_t(r) * VG_MAX + v].F;
                if (f < 0 || (long long)f > gr.lay.cap[v]) { ctx->err = "l3d_match_view: internal error (record window overrun)"; return 1; }
                total += f; max_cnt = std::max(max_cnt, f);
            }
            view_cnt[v] = int(total - view_off[v]);
        }
        if (total > 0x7fffffffll) { ctx->err = "l3d_match_view: too many output records in one group of views"; return 1; }
        if (gr.h_allrec.p[size_t(VG_MAX) * MR_MAX_RANKS - 1].pad != 0) { ctx->err = "l3d_match_view: a rank did not deliver its records of a group of views within ten seconds"; return 1; }
        copy_back = copy_back && (ctx->result_root < 0 || ctx->result_root == ctx->rank);      // who gets the records on the host
        (void)max_cnt;
        if (copy_back && total > 0) L3D_CUDA(cudaStreamWaitEvent(ctx->copy_stream, gr.ev_x[0], 0));
    }
    bool copies = false;
    for (int g = 0; g < gr.count; ++g) {
        ViewSlot& sl = gr.slot[g];
        ViewJob& job = sl.job;
        job.active = false;
        const int* hr = gr.h_vres.p + size_t(g) * gr.res_stride;
        l3d_match_result* res = single ? single : &ctx->results[sl.plan_index].res;
        memset(res, 0, sizeof *res);
        res->pair_tests = job.pair_tests; res->n_candidates = job.cand_ub; res->scored = job.scored ? 1 : 0;
        int M = hr[0], F = (job.M_ub > 0) ? hr[1] : 0;
        float median = 1.0f;                              // line3D.cc:811
        if (mr) {                                          // counters: this rank's rows; records and median: the whole view
            long long m_all = 0;
            for (int r = 0; r < ctx->world; ++r) m_all += gr.h_allrec.p[size_t(r) * VG_MAX + g].M;
            const RankRec& mine = gr.h_allrec.p[size_t(ctx->rank) * VG_MAX + g];
            M = mine.M; F = view_cnt[g];
            if (job.scored && m_all > 0) memcpy(&median, &hr[2], sizeof(float));
            if (job.scored) { res->score_iters = mine.iters_all; res->score_iters_other_cam = mine.iters_all - mine.iters_same; }
        } else if (job.scored && M > 0) {
            memcpy(&median, &hr[2], sizeof(float));
            const unsigned long long all = (unsigned long long)(unsigned)hr[4] | ((unsigned long long)(unsigned)hr[5] << 32);
            const unsigned long long same = (unsigned long long)(unsigned)hr[6] | ((unsigned long long)(unsigned)hr[7] << 32);
            res->score_iters = all; res->score_iters_other_cam = all - same;
        }
        if (job.delivering) {                              // exact sizes of what this view handed over
            DeliverBuf& b = ctx->deliver_bufs[job.owner];
            for (int d = 0; d < job.n_deliver; ++d) { b.counts[d] = hr[8 + d]; b.maxg[d] = std::min(b.maxg[d], hr[8 + d]); }
            for (size_t c = 0; c < job.nb.size(); ++c) {
                if (!job.nb[c].deliver) continue;
                std::vector<Delivered>& dl = ctx->delivered[job.nb[c].view];
                for (size_t k = 0; k < dl.size(); ++k)
                    if (dl[k].owner == job.owner) { dl[k].count = b.counts[dl[k].slot]; dl[k].max_group = b.maxg[dl[k].slot]; }
            }
        }
        for (size_t k = 0; k < job.consumed.size(); ++k) {
            DeliverBuf& b = ctx->deliver_bufs[job.consumed[k].first];
            res->n_delivered += uint64_t(b.counts[job.consumed[k].second]);
            if (--b.refs == 0) {                           // the last receiver has been collected: nothing on the GPU reads it any more
                ctx->pool.put(b.start); ctx->pool.put(b.tgt); ctx->pool.put(b.dep);
                b.start = nullptr; b.tgt = nullptr; b.dep = nullptr;
            }
        }
        res->n_scored = uint64_t(M);
        res->n_raw = uint64_t(M) - res->n_delivered;
        res->median_depth = median;
        res->n_filtered = uint64_t(F);
        l3d_match* dst = nullptr;
        if (copy_back && F > 0) {                          // on its own stream: the next group is already running
            if (ctx->retain || !single) { if (arena_take(ctx, size_t(F), &dst)) return 1; }      // results of a schedule are collected ahead of their calls
            else { L3D_CUDA(ctx->h_filtered.ensure(size_t(F))); dst = ctx->h_filtered.p; }
            const l3d_match* from = mr ? gr.d_gather.p + view_off[g] : sl.d_filtered.p;
            L3D_CUDA(cudaMemcpyAsync(dst, from, sizeof(l3d_match) * size_t(F), cudaMemcpyDeviceToHost, ctx->copy_stream));
            copies = true;
        }
        res->filtered = dst;
        if (!single) {
            l3d_ctx::Stashed& stash = ctx->results[sl.plan_index];
            memset(stash.ms, 0, sizeof stash.ms); stash.kms[0] = stash.kms[1] = stash.kms[2] = 0.f;
            if (g == 0) {                                  // the group's device times are booked on its first view
                stash.ms[0] = ms_a[0]; stash.ms[1] = ms_a[1]; stash.ms[2] = ms_b[0]; stash.ms[3] = ms_b[1]; stash.ms[4] = ms_b[2];
                stash.ms[5] = ms_a[0] + ms_a[1] + ms_b[0] + ms_b[1] + ms_b[2];
                stash.kms[0] = ctx->pa_kernel_ms[0]; stash.kms[1] = ctx->pa_kernel_ms[1]; stash.kms[2] = ctx->pa_kernel_ms[2];
            }
            stash.ready = true;
        } else {
            ctx->last_ms[0] = ms_a[0]; ctx->last_ms[1] = ms_a[1]; ctx->last_ms[2] = ms_b[0]; ctx->last_ms[3] = ms_b[1]; ctx->last_ms[4] = ms_b[2];
            ctx->last_ms[5] = ms_a[0] + ms_a[1] + ms_b[0] + ms_b[1] + ms_b[2];
        }
    }
    if (mr) {
        // the next group can go onto the GPU before these copies have arrived: they are waited for when a result is
        // handed to the caller (copies_pending), and before this group's buffers are used again (ev_x[1])
        if (copies) { L3D_CUDA(cudaEventRecord(gr.ev_x[1], ctx->copy_stream)); ctx->copies_pending = true; }
    } else if (copies) L3D_CUDA(cudaStreamSynchronize(ctx->copy_stream));
    return 0;
}

int join_worker(l3d_ctx* ctx);

// Collects the views that were put on the GPU ahead of their calls and are not going to be asked for (and the helper thread).
void drain_pending(l3d_ctx* ctx)
{
    join_worker(ctx);
    while (ctx->grp_begin < ctx->grp_end) { ViewGroup& gr = ctx->grp[ctx->grp_begin % 2]; finish_group(ctx, gr, false, nullptr); ctx->fly_begin += size_t(gr.count); ++ctx->grp_begin; }
    ctx->fly_begin = ctx->fly_end = 0; ctx->grp_begin = ctx->grp_end = 0;
}

// The batch of the announced schedule that starts at position `from` of the execution order: its job list, and
// where it ends.
size_t plan_batch(l3d_ctx* ctx, size_t from, std::vector<JobHost>& jh)
{
    size_t e = from; uint64_t pairs = 0;
    while (e < ctx->order.size() && (e == from || pairs + ctx->plan[ctx->order[e]].pair_tests <= ctx->batch_pairs)) {
        PlannedView& pv = ctx->plan[ctx->order[e]];
        if (e > from && jh.size() + pv.nb.size() > 65535) break;        // one grid row per job in k_hoist
        pv.first_job = int(jh.size());
        for (size_t c = 0; c < pv.nb.size(); ++c)
            if (pv.nb[c].to_be_matched) { JobHost j; j.src_view = pv.view; j.tgt_view = pv.nb[c].view; memcpy(j.F, pv.nb[c].F, sizeof j.F); jh.push_back(j); }
        pairs += pv.pair_tests; ++e;
    }
    return e;
}

// Pair tests of the batch starting at `from`, on the helper thread and the second stream, into the set that is not
// in use.  The per-view work of a batch keeps only a fraction of the SMs busy (short dependent kernels); the pair
// tests of the next batch fill the rest.
void start_worker(l3d_ctx* ctx, size_t from)
{
    PhaseASet* set = &ctx->pa[ctx->cur ^ 1];
    std::vector<JobHost> jh;
    set->begin = from; set->end = plan_batch(ctx, from, jh);
    ctx->pa_worker_rc = 0;
    set->background = true;
    ctx->pa_worker = std::thread([ctx, set, jh]() {
        try { ctx->pa_worker_rc = run_phase_a(ctx, set, jh); }
        catch (const std::exception& e) { set->err = std::string("pair tests (helper thread): ") + e.what(); ctx->pa_worker_rc = 1; }
        catch (...) { set->err = "pair tests (helper thread): unknown exception"; ctx->pa_worker_rc = 1; }
    });
    ctx->pa_worker_active = true;
}

int join_worker(l3d_ctx* ctx)
{
    if (!ctx->pa_worker_active) return 0;
    ctx->pa_worker.join();
    ctx->pa_worker_active = false;
    return ctx->pa_worker_rc;
}

// More views of the schedule onto the GPU, in execution order, as far as slots and resident pair tests allow; starts
// the next batch when the current one has been collected completely.
int top_up(l3d_ctx* ctx)
{
    if (ctx->fly_end >= ctx->batch_end) {
        if (ctx->fly_begin < ctx->fly_end || ctx->fly_end >= ctx->order.size()) return 0;
        bool have = false;
        if (ctx->pa_worker_active) {                             // its pair tests were started while the previous batch was scored
            PhaseASet& nx = ctx->pa[ctx->cur ^ 1];
            const auto tj = std::chrono::steady_clock::now();
            const int rc = join_worker(ctx);
            ctx->host_s[4] += std::chrono::duration<double>(std::chrono::steady_clock::now() - tj).count();
            if (rc == 0) {
                if (nx.begin != ctx->fly_end) { ctx->err = "l3d_match_view: internal error (batch out of step)"; return 1; }
                ctx->cur ^= 1; have = true;
            } else if (rc != 2) { ctx->err = nx.err; return 1; }
            // rc == 2: more candidates than a batch can index -- the batch is redone below, cut smaller
        }
        while (!have) {
            std::vector<JobHost> jh;
            PhaseASet& set = ctx->pa[ctx->cur];
            set.begin = ctx->fly_end; set.end = plan_batch(ctx, ctx->fly_end, jh);
            set.background = false;
            const auto tf = std::chrono::steady_clock::now();
            const int rc = run_phase_a(ctx, &set, jh);
            ctx->host_s[5] += std::chrono::duration<double>(std::chrono::steady_clock::now() - tf).count();
            if (rc == 2 && set.end - set.begin > 1 && ctx->world <= 1) {             // halve the batch and try again (one GPU only: ranks must cut alike)
                uint64_t pairs = 0;
                for (size_t e = set.begin; e < set.end; ++e) pairs += ctx->plan[ctx->order[e]].pair_tests;
                ctx->batch_pairs = std::max<uint64_t>(pairs / 2, 1);
                continue;
            }
            if (rc) { ctx->err = set.err + (rc == 2 ? " (a single view: not supported)" : ""); return 1; }
            have = true;
        }
        PhaseASet& set = ctx->pa[ctx->cur];
        const auto tt = std::chrono::steady_clock::now();
        if (mr_job_totals(ctx, set)) return 1;
        ctx->host_s[6] += std::chrono::duration<double>(std::chrono::steady_clock::now() - tt).count();
        ctx->batch_begin = set.begin; ctx->batch_end = set.end;
        ctx->launches += set.launches; set.launches = 0;
        if (ctx->batch_end < ctx->order.size() && !ctx->stage_events && !ctx->no_overlap) start_worker(ctx, ctx->batch_end);      // the following batch, meanwhile (not when stages are timed: nothing may run beside anything)
    }
    // the next groups: consecutive views of one dependency level, two groups in flight
    while (ctx->fly_end < ctx->batch_end && ctx->grp_end - ctx->grp_begin < 2) {
        GroupView gv[VG_MAX]; int count = 0;
        const int lvl = ctx->level[ctx->order[ctx->fly_end]];
        while (count < ctx->group_views && ctx->fly_end + size_t(count) < ctx->batch_end && ctx->level[ctx->order[ctx->fly_end + size_t(count)]] == lvl) {
            PlannedView& pv = ctx->plan[ctx->order[ctx->fly_end + size_t(count)]];
            GroupView v = {pv.view, int(pv.nb.size()), pv.nb.data(), pv.first_job, ctx->order[ctx->fly_end + size_t(count)]};
            gv[count++] = v;
        }
        ViewGroup& gr = ctx->grp[ctx->grp_end % 2];
        gr.pos0 = ctx->fly_end;
        if (ctx->world > 1) {
            // the record windows must hold this group; growing them is a collective step with nothing in flight
            // (every rank computes the same bound at the same point of the same schedule)
            const size_t need = mr_group_need(ctx, gv, count);
            if (need > ctx->win_cap) {
                if (ctx->grp_begin < ctx->grp_end) return 0;          // collect what is in flight first; the caller's loop comes back
                if (mr_ensure_window(ctx, need)) return 1;
            }
        }
        const auto t0 = std::chrono::steady_clock::now();
        if (enqueue_group(ctx, gr, gv, count, ctx->run_sigma_p, ctx->run_sigma_a)) return 1;
        ctx->host_s[0] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); ctx->host_s[3] += 1.0;
        ctx->fly_end += size_t(count); ++ctx->grp_end;
    }
    return 0;
}

int match_view_impl(l3d_ctx* ctx, int view, int n_nb, const l3d_neighbor* nb, float sigma_p, float sigma_a, l3d_match_result* res, bool copy_back)
{
    if (!ctx) return 1;
    if (n_nb <= 0 || !nb || !res) { ctx->err = "l3d_match_view: no neighbours / null argument"; return 1; }
    if (view < 0 || view >= ctx->V) { ctx->err = "l3d_match_view: view index out of range"; return 1; }
    L3D_CUDA(cudaSetDevice(ctx->device));
    memset(res, 0, sizeof *res);
    const bool on_plan = ctx->call_next < ctx->plan.size() && ctx->plan[ctx->call_next].view == view && same_neighbors(ctx->plan[ctx->call_next].nb, n_nb, nb);
    if (ctx->plan_started && !(on_plan && ctx->run_sigma_p == sigma_p && ctx->run_sigma_a == sigma_a && ctx->run_copy_back == copy_back)) {
        ctx->err = "l3d_match_view: the call does not follow the schedule announced with l3d_plan_matching (same views, neighbours, sigmas and entry point, in order); call l3d_reset_matching to start over";
        return 1;
    }
    if (on_plan) {
        // ---- the next view of the announced schedule: the library works through the schedule in its own order
        // (by dependency level) and ahead of the caller; wait until this view's result has been collected
        if (!ctx->plan_started) { ctx->plan_started = true; ctx->run_sigma_p = sigma_p; ctx->run_sigma_a = sigma_a; ctx->run_copy_back = copy_back; }
        ctx->want_records = ctx->run_copy_back;
        const size_t p = ctx->call_next;
        while (!ctx->results[p].ready) {
            if (top_up(ctx)) { drain_pending(ctx); return 1; }
            if (ctx->grp_begin >= ctx->grp_end) { ctx->err = "l3d_match_view: internal error (schedule exhausted)"; return 1; }
            ViewGroup& gr = ctx->grp[ctx->grp_begin % 2];
            const int rc = finish_group(ctx, gr, ctx->run_copy_back, nullptr);
            ctx->fly_begin += size_t(gr.count); ++ctx->grp_begin;
            if (rc) { drain_pending(ctx); return 1; }
        }
        if (ctx->copies_pending) { L3D_CUDA(cudaStreamSynchronize(ctx->copy_stream)); ctx->copies_pending = false; }
        *res = ctx->results[p].res;
        memcpy(ctx->last_ms, ctx->results[p].ms, sizeof ctx->last_ms); memcpy(ctx->pa_kernel_ms, ctx->results[p].kms, sizeof ctx->pa_kernel_ms);
        ++ctx->call_next;
        if (top_up(ctx)) { drain_pending(ctx); return 1; }       // keep the GPU busy while the caller works
        return 0;
    }
    // ---- a call off the schedule (before it started) voids it: this view alone -------------------
    drain_pending(ctx);
    ctx->plan.clear(); ctx->order.clear(); ctx->results.clear(); ctx->call_next = 0; ctx->batch_begin = ctx->batch_end = 0; ctx->fly_begin = ctx->fly_end = 0;
    std::vector<JobHost> jh;
    for (int c = 0; c < n_nb; ++c)
        if (nb[c].to_be_matched) { JobHost j; j.src_view = view; j.tgt_view = nb[c].view; memcpy(j.F, nb[c].F, sizeof j.F); jh.push_back(j); }
    PhaseASet& set = ctx->pa[ctx->cur];
    set.background = false;
    ctx->want_records = copy_back;
    if (run_phase_a(ctx, &set, jh)) { ctx->err = set.err; return 1; }
    if (mr_job_totals(ctx, set)) return 1;
    ctx->launches += set.launches; set.launches = 0;
    GroupView one = {view, n_nb, nb, 0, 0};
    if (ctx->world > 1 && mr_ensure_window(ctx, mr_group_need(ctx, &one, 1))) return 1;
    if (enqueue_group(ctx, ctx->grp[0], &one, 1, sigma_p, sigma_a)) return 1;
    if (finish_group(ctx, ctx->grp[0], copy_back, res)) return 1;
    if (ctx->copies_pending) { L3D_CUDA(cudaStreamSynchronize(ctx->copy_stream)); ctx->copies_pending = false; }
    return 0;
}

// Nothing throws across the C ABI: the host-side containers can (out of memory).
template <typename F>
int guarded(l3d_ctx* ctx, F f)
{
    try { return f(); }
    catch (const std::exception& e) { if (ctx) ctx->err = std::string("exception: ") + e.what(); return 1; }
    catch (...) { if (ctx) ctx->err = "unknown exception"; return 1; }
}

}  // namespace

// ==========================================================================================
// C ABI
// ==========================================================================================
extern "C" {

int l3d_create(int device, l3d_ctx** out)
{
    if (!out) { g_create_error = "l3d_create: null output pointer"; return 1; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) { g_create_error = std::string("l3d_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path"; return 1; }
    if (device < 0 || device >= n) { g_create_error = "l3d_create: device index out of range"; return 1; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_error = std::string("l3d_create: cudaSetDevice: ") + cudaGetErrorString(e); return 1; }
    l3d_ctx* ctx = new l3d_ctx();
    ctx->device = device;
    { int sm = 0; if (cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sm > 0) ctx->n_sm = sm; }
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);        // the per-view chain is latency-bound: it goes first
    if ((e = cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, prio_hi)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); delete ctx; return 1; }
    {   // keep stream-ordered allocations cached across synchronisations (hand-over buffers are allocated per view)
        cudaMemPool_t pool; unsigned long long keep = ~0ull;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    for (int k = 0; k < 2; ++k) cudaEventCreate(&ctx->ev[k]);
    for (int q = 0; q < 2; ++q) for (int k = 0; k < 6; ++k) cudaEventCreate(&ctx->grp[q].ev[k]);
    for (int q = 0; q < 2; ++q) for (int k = 0; k < 2; ++k) cudaEventCreateWithFlags(&ctx->grp[q].ev_x[k], cudaEventDisableTiming);
    if ((e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); cudaStreamDestroy(ctx->stream); delete ctx; return 1; }
    if ((e = cudaStreamCreateWithPriority(&ctx->fin_stream, cudaStreamNonBlocking, prio_hi)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); l3d_destroy(ctx); return 1; }
    for (int q = 0; q < 2; ++q) {
        PhaseASet& set = ctx->pa[q];
        if ((e = cudaStreamCreateWithPriority(&set.stream, cudaStreamNonBlocking, prio_lo)) != cudaSuccess) { g_create_error = std::string("l3d_create: stream: ") + cudaGetErrorString(e); l3d_destroy(ctx); return 1; }
        for (int k = 0; k < 3; ++k) cudaEventCreate(&set.pa_ev[k]);
        for (int k = 0; k < 5; ++k) cudaEventCreate(&set.pa_kev[k]);
    }
    for (int k = 0; k < 2; ++k) cudaEventCreate(&ctx->mark_ev[k]);
    *out = ctx;
    return 0;
}

int l3d_comm_unique_id(void* id)
{
    std::string why;
    NcclApi* api = nccl_api(&why);
    if (!api || !id) { g_create_error = api ? "l3d_comm_unique_id: null argument" : "l3d_comm_unique_id: " + why; return 1; }
    ncclUniqueId u;
    static_assert(sizeof(ncclUniqueId) == L3D_COMM_ID_BYTES, "ncclUniqueId size");
    const ncclResult_t r = api->GetUniqueId(&u);
    if (r != ncclSuccess) { g_create_error = std::string("ncclGetUniqueId: ") + api->GetErrorString(r); return 1; }
    memcpy(id, &u, sizeof u);
    return 0;
}

int l3d_comm_init(l3d_ctx* ctx, int rank, int world, const void* id, int result_root)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (world < 1 || world > MR_MAX_RANKS || rank < 0 || rank >= world || (world > 1 && !id) || result_root >= world) { ctx->err = "l3d_comm_init: bad arguments (1 <= world <= 16)"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        l3d_reset_matching(ctx);
        if (l3d_comm_destroy(ctx)) return 1;
        if (world == 1) return 0;
        std::string why;
        ctx->nccl = nccl_api(&why);
        if (!ctx->nccl) { ctx->err = "l3d_comm_init: " + why; return 1; }
        ncclUniqueId u; memcpy(&u, id, sizeof u);
        L3D_NCCL(ctx->nccl->CommInitRank(&ctx->comm, world, u, rank));
        ctx->rank = rank; ctx->world = world; ctx->result_root = result_root;
        return 0;
    });
}

int l3d_comm_destroy(l3d_ctx* ctx)
{
    if (!ctx) return 1;
    if (ctx->comm) {
        cudaSetDevice(ctx->device);
        if (!ctx->comm_broken) { drain_pending(ctx); cudaStreamSynchronize(ctx->stream); }
        mr_release_window(ctx);
        if (ctx->comm_broken) ctx->nccl->CommAbort(ctx->comm); else ctx->nccl->CommDestroy(ctx->comm);
        ctx->comm = nullptr;
    }
    ctx->rank = 0; ctx->world = 1; ctx->result_root = -1; ctx->comm_broken = false;
    return 0;
}

int l3d_reset_matching(l3d_ctx* ctx)
{
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    drain_pending(ctx);
    for (size_t v = 0; v < ctx->delivered.size(); ++v) ctx->delivered[v].clear();
    ctx->deliver_bufs.clear();
    ctx->pool.put_all();                                 // hand-over buffers never taken, or not yet given back
    ctx->level.clear();
    ctx->plan.clear(); ctx->order.clear(); ctx->results.clear(); ctx->call_next = 0; ctx->plan_started = false; ctx->batch_begin = ctx->batch_end = 0; ctx->fly_begin = ctx->fly_end = 0;
    ctx->arena_chunk = 0; ctx->arena_used = 0;           // retained results are given up, their memory is reused
    cudaStreamSynchronize(ctx->stream);
    return 0;
}

void l3d_destroy(l3d_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    l3d_reset_matching(ctx);
    l3d_comm_destroy(ctx);
    ctx->d_segs.release(); ctx->d_viewd.release(); ctx->pa[0].release(); ctx->pa[1].release();
    ctx->d_tile_sums.release(); ctx->d_total.release();
    ctx->grp[0].release(); ctx->grp[1].release(); ctx->pool.release();
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->fin_stream) cudaStreamDestroy(ctx->fin_stream);
    for (size_t k = 0; k < ctx->arena.size(); ++k) cudaFreeHost(ctx->arena[k].first);
    for (int k = 0; k < 2; ++k) if (ctx->mark_ev[k]) cudaEventDestroy(ctx->mark_ev[k]);
    ctx->h_filtered.release(); ctx->h_ints.release();
    ctx->d_csegs.release(); ctx->d_cbits.release(); ctx->d_ccnt.release(); ctx->d_coff.release(); ctx->d_cpairs.release(); ctx->h_cpairs.release();
    ctx->h_cands.release(); ctx->d_cands.release(); ctx->d_hscore.release(); ctx->d_ew.release(); ctx->d_ecnt.release(); ctx->d_eoff.release(); ctx->d_kept.release(); ctx->h_kept.release(); ctx->d_ekeys.release(); ctx->d_eorder.release(); ctx->h_eorder.release();
    ctx->d_hyp.release(); ctx->d_hin.release(); ctx->d_hout.release(); ctx->d_pairs.release(); ctx->d_aff.release(); ctx->d_medians.release(); ctx->d_flush.release(); ctx->d_rdd.release(); ctx->h_rdd.release();
    for (int k = 0; k < 2; ++k) if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* l3d_last_error(const l3d_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }
uint64_t l3d_kernel_launches(const l3d_ctx* ctx) { return ctx ? ctx->launches : 0; }

int l3d_collinearity(l3d_ctx* ctx, const float* segments, int n, float collin_s, const l3d_collin_pair** pairs, size_t* count)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (!pairs || !count || (n > 0 && !segments)) { ctx->err = "l3d_collinearity: null argument"; return 1; }
        *pairs = nullptr; *count = 0;
        if (n <= 1) return 0;
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int wpr = (n + 31) / 32;
        L3D_CUDA(ctx->d_csegs.ensure(size_t(n))); L3D_CUDA(ctx->d_cbits.ensure(size_t(n) * wpr));
        L3D_CUDA(ctx->d_ccnt.ensure(size_t(n) + 1)); L3D_CUDA(ctx->d_coff.ensure(size_t(n) + 2));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_csegs.p, segments, sizeof(float4) * size_t(n), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemsetAsync(ctx->d_ccnt.p, 0, sizeof(int) * (size_t(n) + 1), st));
        const float s2 = collin_s * collin_s;                                       // cudawrapper.cu:850
        k_collin_count<<<dim3((wpr + 63) / 64, n), 64, 0, st>>>(ctx->d_csegs.p, n, s2, ctx->d_ccnt.p, ctx->d_cbits.p, wpr);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        if (scan_exclusive(ctx, ctx->d_ccnt.p, ctx->d_coff.p, n)) return 1;
        int total = 0;
        if (fetch_int(ctx, ctx->d_coff.p + n, &total)) return 1;
        if (total == 0) return 0;
        L3D_CUDA(ctx->d_cpairs.ensure(size_t(total))); L3D_CUDA(ctx->h_cpairs.ensure(size_t(total)));
        k_collin_fill<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_csegs.p, n, s2, ctx->d_coff.p, ctx->d_cbits.p, wpr, ctx->d_cpairs.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(ctx->h_cpairs.p, ctx->d_cpairs.p, sizeof(l3d_collin_pair) * size_t(total), cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        *pairs = ctx->h_cpairs.p; *count = size_t(total);
        return 0;
    });
}

int l3d_collinearity_prealloc(const float* d_segments, int n, float collin_s, l3d_collin_pair* d_pairs, int capacity,
                              int* d_count, void* cuda_stream, char* err, size_t err_len)
{
    // No cudaMalloc, cudaFree, cudaMallocAsync or any other allocating call may appear in this function.
    auto fail = [&](int code, const char* msg) { if (err && err_len) snprintf(err, err_len, "l3d_collinearity_prealloc: %s", msg); return code; };
    if (n < 0 || capacity < 0 || (n > 0 && !d_segments) || (capacity > 0 && !d_pairs) || !d_count) return fail(1, "null or negative argument");
    if (n > PREALLOC_MAX_SEGMENTS) {
        char b[160]; snprintf(b, sizeof b, "%d segments exceed the built-in limit of %d (use l3d_collinearity, which allocates)", n, PREALLOC_MAX_SEGMENTS);
        return fail(2, b);
    }
    // the kernels read the segments as float4 and write 12-byte records and an int: a misaligned pointer would raise a
    // sticky device fault that takes the caller's whole CUDA context with it, so it is refused here
    if ((reinterpret_cast<uintptr_t>(d_segments) & 15u) || (reinterpret_cast<uintptr_t>(d_pairs) & 3u) || (reinterpret_cast<uintptr_t>(d_count) & 3u))
        return fail(1, "d_segments must be 16-byte aligned, d_pairs and d_count 4-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // Set-up, once per device and with that device current: the array's name is only a host-side handle for the
    // symbol, the kernels need its address in device memory.  The per-device cache is a table of atomics: two threads
    // that make their first call at the same time both look the address up and store the same value.
    static std::atomic<unsigned int*> scratch_of_device[64];
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) return fail(3, e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    if (st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread) {      // the work must go to the device whose array was looked up
        int sdev = -1;
        e = cudaStreamGetDevice(st, &sdev);
        if (e != cudaSuccess) return fail(7, cudaGetErrorString(e));
        if (sdev != dev) return fail(7, "cuda_stream belongs to another device than the current one (cudaSetDevice to the stream's device first)");
    }
    unsigned int* scratch = scratch_of_device[dev].load(std::memory_order_acquire);
    if (!scratch) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_prealloc_scratch);
        if (e != cudaSuccess || !p) return fail(4, e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned null");
        scratch = static_cast<unsigned int*>(p);
        scratch_of_device[dev].store(scratch, std::memory_order_release);
    }
    unsigned* bits = scratch;
    int* cnt = reinterpret_cast<int*>(scratch + PREALLOC_CNT_OFF);
    int* off = reinterpret_cast<int*>(scratch + PREALLOC_OFF_OFF);
    if (n <= 1) {
        e = cudaMemsetAsync(d_count, 0, sizeof(int), st);
        return e == cudaSuccess ? 0 : fail(5, cudaGetErrorString(e));
    }
    const int wpr = (n + 31) / 32;
    const float s2 = collin_s * collin_s;                                       // cudawrapper.cu:850
    // the counts are accumulated with atomicAdd and must start from zero; the bitmask is fully rewritten
    e = cudaMemsetAsync(cnt, 0, sizeof(int) * (size_t(n) + 1), st);
    if (e != cudaSuccess) return fail(5, cudaGetErrorString(e));
    k_collin_count<<<dim3((wpr + 63) / 64, n), 64, 0, st>>>(reinterpret_cast<const float4*>(d_segments), n, s2, cnt, bits, wpr);
    k_scan_small<<<1, SCAN_T, 0, st>>>(cnt, off, n);
    k_collin_fill_capped<<<(n * 32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(d_segments), n, s2, off, bits, wpr, d_pairs, capacity, d_count);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(6, cudaGetErrorString(e));
    return 0;
}

int l3d_set_views(l3d_ctx* ctx, int n_views, const l3d_view* views)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n_views <= 0 || !views) { ctx->err = "l3d_set_views: no views"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        l3d_reset_matching(ctx);
        ctx->V = n_views;
        ctx->views.assign(views, views + n_views);
        ctx->seg_off.assign(size_t(n_views) + 1, 0);
        ctx->max_seg = 0;
        for (int v = 0; v < n_views; ++v) {
            if (views[v].n_segments < 0 || (views[v].n_segments > 0 && !views[v].segments && !(ctx->world > 1 && ctx->rank != 0))) { ctx->err = "l3d_set_views: bad segment array"; return 1; }
            if (views[v].n_segments > 65535) { ctx->err = "l3d_set_views: more than 65535 segments in a view"; return 1; }
            ctx->seg_off[v + 1] = ctx->seg_off[v] + views[v].n_segments;
            ctx->max_seg = std::max(ctx->max_seg, views[v].n_segments);
        }
        ctx->total_segs = ctx->seg_off[n_views];
        L3D_CUDA(ctx->d_segs.ensure(size_t(std::max(ctx->total_segs, 1))));
        L3D_CUDA(ctx->d_viewd.ensure(size_t(n_views)));
        std::vector<ViewD> vd(n_views);
        for (int v = 0; v < n_views; ++v) {
            // (several GPUs on one scene: a rank other than 0 may leave the segments out -- pointer NULL with n_segments
            // set -- and gets them from rank 0 in the broadcast below)
            if (views[v].n_segments && views[v].segments)
                L3D_CUDA(cudaMemcpyAsync(ctx->d_segs.p + ctx->seg_off[v], views[v].segments, sizeof(float4) * size_t(views[v].n_segments), cudaMemcpyHostToDevice, ctx->stream));
            memcpy(vd[v].RtKinv, views[v].RtKinv_d, sizeof vd[v].RtKinv); memcpy(vd[v].C, views[v].C_d, sizeof vd[v].C);
            vd[v].k_lower = views[v].k_lower; vd[v].k_upper = views[v].k_upper; vd[v].seg_off = ctx->seg_off[v]; vd[v].n_seg = views[v].n_segments;
            ctx->views[v].segments = nullptr;
        }
        L3D_CUDA(cudaMemcpyAsync(ctx->d_viewd.p, vd.data(), sizeof(ViewD) * size_t(n_views), cudaMemcpyHostToDevice, ctx->stream));
        // several GPUs on one scene: the segment table is replicated from rank 0 over NVLink (the camera tables are a few
        // hundred bytes per view and every rank has computed them from the same input)
        if (ctx->world > 1 && ctx->total_segs > 0)
            L3D_NCCL(ctx->nccl->Broadcast(ctx->d_segs.p, ctx->d_segs.p, size_t(ctx->total_segs) * sizeof(float4), ncclChar, 0, ctx->comm, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->delivered.assign(size_t(n_views), std::vector<Delivered>());
        ctx->n_hyp = 0;
        return 0;
    });
}

int l3d_deliver(l3d_ctx* ctx, int to_view, int from_view, size_t n, const l3d_match* records)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (to_view < 0 || to_view >= ctx->V || from_view < 0 || from_view >= ctx->V || to_view == from_view) { ctx->err = "l3d_deliver: view index out of range"; return 1; }
        if (n > 0 && !records) { ctx->err = "l3d_deliver: null argument"; return 1; }
        if (n > size_t(0x7fffffff)) { ctx->err = "l3d_deliver: too many records"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        drain_pending(ctx);
        const int n_to = ctx->views[to_view].n_segments, n_from = ctx->views[from_view].n_segments;
        // grouped by the receiving view's segment (counting sort), as k_deliver_scatter leaves them
        std::vector<int> start(size_t(n_to) + 2, 0), tgt(std::max<size_t>(n, 1)); std::vector<float4> dep(std::max<size_t>(n, 1));
        for (size_t k = 0; k < n; ++k) {
            if (records[k].src_seg < 0 || records[k].src_seg >= n_to || records[k].tgt_seg < 0 || records[k].tgt_seg >= n_from) { ctx->err = "l3d_deliver: segment index out of range"; return 1; }
            ++start[size_t(records[k].src_seg) + 1];
        }
        int max_group = 0;
        for (int q = 0; q < n_to; ++q) { max_group = std::max(max_group, start[size_t(q) + 1]); start[size_t(q) + 1] += start[q]; }
        start[size_t(n_to) + 1] = start[n_to];
        { std::vector<int> cur(start.begin(), start.begin() + n_to);
          for (size_t k = 0; k < n; ++k) { const int pos = cur[records[k].src_seg]++; tgt[pos] = records[k].tgt_seg; dep[pos] = make_float4(records[k].depths[0], records[k].depths[1], records[k].depths[2], records[k].depths[3]); } }
        DeliverBuf buf; buf.refs = 1; buf.counts.assign(1, int(n)); buf.maxg.assign(1, max_group);
        L3D_CUDA(ctx->pool.get(sizeof(int) * (size_t(n_to) + 2), reinterpret_cast<void**>(&buf.start)));
        L3D_CUDA(ctx->pool.get(sizeof(int) * std::max<size_t>(n, 1), reinterpret_cast<void**>(&buf.tgt)));
        L3D_CUDA(ctx->pool.get(sizeof(float4) * std::max<size_t>(n, 1), reinterpret_cast<void**>(&buf.dep)));
        L3D_CUDA(cudaMemcpyAsync(buf.start, start.data(), sizeof(int) * (size_t(n_to) + 2), cudaMemcpyHostToDevice, ctx->stream));
        if (n) {
            L3D_CUDA(cudaMemcpyAsync(buf.tgt, tgt.data(), sizeof(int) * n, cudaMemcpyHostToDevice, ctx->stream));
            L3D_CUDA(cudaMemcpyAsync(buf.dep, dep.data(), sizeof(float4) * n, cudaMemcpyHostToDevice, ctx->stream));
        }
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));          // the host vectors go away
        Delivered dl; dl.from_view = from_view; dl.count = int(n); dl.count_all = (long long)n; dl.max_group = max_group; dl.t_start = buf.start; dl.t_tgt = buf.tgt; dl.t_dep = buf.dep;
        dl.owner = int(ctx->deliver_bufs.size()); dl.slot = 0;
        ctx->deliver_bufs.push_back(buf);
        ctx->delivered[to_view].push_back(dl);
        return 0;
    });
}

int l3d_plan_matching(l3d_ctx* ctx, int n_views, const l3d_view_plan* plans)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        cudaSetDevice(ctx->device);
        drain_pending(ctx);
        ctx->plan.clear(); ctx->order.clear(); ctx->results.clear(); ctx->call_next = 0; ctx->plan_started = false; ctx->batch_begin = ctx->batch_end = 0; ctx->fly_begin = ctx->fly_end = 0;
        if (n_views <= 0) return 0;
        if (!plans) { ctx->err = "l3d_plan_matching: null argument"; return 1; }
        for (int k = 0; k < n_views; ++k) {
            const l3d_view_plan& p = plans[k];
            if (p.view < 0 || p.view >= ctx->V || p.n_neighbors <= 0 || !p.neighbors) { ctx->plan.clear(); ctx->err = "l3d_plan_matching: bad entry"; return 1; }
            PlannedView pv; pv.view = p.view; pv.nb.assign(p.neighbors, p.neighbors + p.n_neighbors);
            for (int c = 0; c < p.n_neighbors; ++c) {
                if (p.neighbors[c].view < 0 || p.neighbors[c].view >= ctx->V || p.neighbors[c].view == p.view) { ctx->plan.clear(); ctx->err = "l3d_plan_matching: bad neighbour"; return 1; }
                // (several GPUs on one scene: the rows of the rank with most of them -- the batches must be cut the same way on every rank)
                if (p.neighbors[c].to_be_matched) pv.pair_tests += uint64_t((ctx->views[p.view].n_segments + ctx->world - 1) / ctx->world) * uint64_t(ctx->views[p.neighbors[c].view].n_segments);
            }
            ctx->plan.push_back(pv);
        }
        // Execution order.  View b needs view a first only if a hands records over to b (a comes earlier and marks b
        // "deliver"), so the schedule is a dependency graph, usually of a few independent chains (views along a street;
        // C2: 200 views, longest chain 50).  The library works through it level by level -- the views of one level do not
        // depend on each other and run side by side -- and hands the results out in the order of the calls.
        {
            const size_t n = ctx->plan.size();
            std::vector<int> at(size_t(ctx->V), -1), level(n, 1);
            for (size_t k = 0; k < n; ++k) at[ctx->plan[k].view] = int(k);
            for (size_t k = 0; k < n; ++k)
                for (size_t c = 0; c < ctx->plan[k].nb.size(); ++c) {
                    const l3d_neighbor& nbr = ctx->plan[k].nb[c];
                    const int to = at[nbr.view];
                    if (nbr.deliver && to > int(k)) level[to] = std::max(level[to], level[k] + 1);
                }
            ctx->level = level;
            ctx->order.resize(n);
            for (size_t k = 0; k < n; ++k) ctx->order[k] = k;
            std::stable_sort(ctx->order.begin(), ctx->order.end(), [&level](size_t a, size_t b) { return level[a] < level[b]; });
            l3d_ctx::Stashed empty; memset(&empty, 0, sizeof empty);
            ctx->results.assign(n, empty);
        }
        // Batches: the pair tests of batch i + 1 run beside the per-view work of batch i, so the schedule is cut into
        // about eight of them (not smaller than 2.5e8 pair tests: a batch has fixed costs), none above the memory budget.
        uint64_t total = 0;
        for (size_t k = 0; k < ctx->plan.size(); ++k) total += ctx->plan[k].pair_tests;
        {   // tuning knobs, read once per schedule
            const char* e;
            ctx->bg_blocks = (e = getenv("L3D_BG_BLOCKS")) ? std::max(0, std::min(atoi(e), 16)) : 0;
            ctx->p1_rows = (e = getenv("L3D_P1_ROWS")) ? std::max(128, atoi(e) / 128 * 128) : 0;
            ctx->no_overlap = (e = getenv("L3D_NO_OVERLAP")) != nullptr && atoi(e) != 0;
            ctx->batch_div = (e = getenv("L3D_BATCH_DIV")) ? std::max(1, std::min(atoi(e), 256)) : 8;
            // (one scene on several GPUs: a rank's share of a view is small and every group costs an exchange, so larger groups)
            ctx->group_views = (e = getenv("L3D_GROUP_VIEWS")) ? std::max(1, std::min(atoi(e), VG_MAX)) : (ctx->world > 1 ? 16 : 8);
        }
        ctx->batch_pairs = ctx->batch_override ? ctx->batch_override : std::min<uint64_t>(ctx->pair_budget, std::max<uint64_t>(total / uint64_t(ctx->batch_div), 250000000ull));
        if (ctx->stage_events && !ctx->batch_override) ctx->batch_pairs = ctx->pair_budget;      // per-stage timing: nothing runs beside anything else
        return 0;
    });
}

int l3d_debug_set_batch_pairs(l3d_ctx* ctx, uint64_t pairs)
{
    if (!ctx) return 1;
    ctx->batch_override = pairs;
    return 0;
}

int l3d_debug_set_pretest(l3d_ctx* ctx, int mode)
{
    if (!ctx || mode < 0 || mode > 1) return 1;
    ctx->pretest_mode = mode;
    return 0;
}

int l3d_debug_set_score_block(l3d_ctx* ctx, int hypotheses)
{
    if (!ctx || !(hypotheses == 0 || hypotheses == 32 || hypotheses == 64 || hypotheses == 128)) return 1;
    ctx->score_block = hypotheses;
    return 0;
}

int l3d_debug_set_candidate_limit(l3d_ctx* ctx, uint64_t limit)
{
    if (!ctx) return 1;
    ctx->cand_limit = (limit == 0 || limit > 0x7fff0000ull) ? 0x7fff0000ull : limit;
    return 0;
}

int l3d_debug_chain_times(l3d_ctx* ctx, int on, double out[24])
{
    if (!ctx) return 1;
    if (out) for (int k = 0; k < 24; ++k) { out[k] = ctx->chain_ms[k]; ctx->chain_ms[k] = 0.0; }
    ctx->chain_prof = on != 0;
    return 0;
}

int l3d_debug_host_times(l3d_ctx* ctx, double out[4])
{
    if (!ctx || !out) return 1;
    for (int k = 0; k < 8; ++k) { out[k] = ctx->host_s[k]; ctx->host_s[k] = 0.0; }
    return 0;
}

int l3d_retain_results(l3d_ctx* ctx, int on)
{
    if (!ctx) return 1;
    ctx->retain = on != 0;
    return 0;
}

int l3d_match_view(l3d_ctx* ctx, int view, int n_neighbors, const l3d_neighbor* neighbors, float sigma_p, float sigma_a, l3d_match_result* result)
{
    const int rc = guarded(ctx, [&]() { return match_view_impl(ctx, view, n_neighbors, neighbors, sigma_p, sigma_a, result, true); });
    if (rc && ctx && ctx->world > 1) ctx->comm_broken = true;
    return rc;
}
int l3d_match_view_resident(l3d_ctx* ctx, int view, int n_neighbors, const l3d_neighbor* neighbors, float sigma_p, float sigma_a, l3d_match_result* result)
{
    const int rc = guarded(ctx, [&]() { return match_view_impl(ctx, view, n_neighbors, neighbors, sigma_p, sigma_a, result, false); });
    if (rc && ctx && ctx->world > 1) ctx->comm_broken = true;
    return rc;
}

int l3d_unproject(l3d_ctx* ctx, size_t n, const l3d_hypothesis_in* in, l3d_segment3d* out)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n == 0) { ctx->n_hyp = 0; return 0; }
        if (!in || !out) { ctx->err = "l3d_unproject: null argument"; return 1; }
        for (size_t k = 0; k < n; ++k)
            if (in[k].view < 0 || in[k].view >= ctx->V || in[k].seg < 0 || in[k].seg >= ctx->views[in[k].view].n_segments) { ctx->err = "l3d_unproject: view/segment out of range"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        L3D_CUDA(ctx->d_hin.ensure(n)); L3D_CUDA(ctx->d_hout.ensure(n)); L3D_CUDA(ctx->d_hyp.ensure(n));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_hin.p, in, sizeof(l3d_hypothesis_in) * n, cudaMemcpyHostToDevice, ctx->stream));
        k_unproject<<<unsigned((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_segs.p, ctx->d_viewd.p, ctx->d_hin.p, n, ctx->d_hyp.p, ctx->d_hout.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(out, ctx->d_hout.p, sizeof(l3d_segment3d) * n, cudaMemcpyDeviceToHost, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->n_hyp = n;
        return 0;
    });
}

int l3d_affinity(l3d_ctx* ctx, const float* median_depth, size_t n_pairs, const l3d_pair* pairs, float sigma_a, float* out)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n_pairs == 0) return 0;
        if (!median_depth || !pairs || !out) { ctx->err = "l3d_affinity: null argument"; return 1; }
        if (ctx->n_hyp == 0) { ctx->err = "l3d_affinity: no hypotheses resident (call l3d_unproject first)"; return 1; }
        for (size_t k = 0; k < n_pairs; ++k)
            if (pairs[k].a < 0 || size_t(pairs[k].a) >= ctx->n_hyp || pairs[k].b < 0 || size_t(pairs[k].b) >= ctx->n_hyp) { ctx->err = "l3d_affinity: pair index out of range"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        L3D_CUDA(ctx->d_pairs.ensure(n_pairs)); L3D_CUDA(ctx->d_aff.ensure(n_pairs)); L3D_CUDA(ctx->d_medians.ensure(size_t(ctx->V)));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_medians.p, median_depth, sizeof(float) * size_t(ctx->V), cudaMemcpyHostToDevice, ctx->stream));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_pairs.p, pairs, sizeof(l3d_pair) * n_pairs, cudaMemcpyHostToDevice, ctx->stream));
        const float two_log = 2.0f * logf(0.01f);                                   // view.cc:376
        k_affinity<<<unsigned((n_pairs + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_hyp.p, ctx->d_viewd.p, ctx->d_medians.p, ctx->d_pairs.p, n_pairs, sigma_a, two_log, ctx->d_aff.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(out, ctx->d_aff.p, sizeof(float) * n_pairs, cudaMemcpyDeviceToHost, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        return 0;
    });
}

int l3d_candidate_buffer(l3d_ctx* ctx, size_t n, l3d_candidate** buf)
{
    if (!ctx || !buf) return 1;
    return guarded(ctx, [&]() -> int {
        L3D_CUDA(cudaSetDevice(ctx->device));
        L3D_CUDA(ctx->h_cands.ensure(std::max<size_t>(n, 1)));
        *buf = ctx->h_cands.p;
        return 0;
    });
}

int l3d_affinity_edges(l3d_ctx* ctx, const float* median_depth, const float* hyp_score, size_t n, const l3d_candidate* cands, float sigma_a,
                       const l3d_kept_edge** kept, size_t* n_kept, const uint32_t** order)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (!kept || !n_kept) { ctx->err = "l3d_affinity_edges: null argument"; return 1; }
        *kept = nullptr; *n_kept = 0;
        if (order) *order = nullptr;
        if (n == 0) return 0;
        if (!median_depth || !hyp_score || !cands) { ctx->err = "l3d_affinity_edges: null argument"; return 1; }
        if (ctx->n_hyp == 0) { ctx->err = "l3d_affinity_edges: no hypotheses resident (call l3d_unproject first)"; return 1; }
        if (n > size_t(0x7fffffff) - AE_T) { ctx->err = "l3d_affinity_edges: too many candidates"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int blocks = int((n + AE_T - 1) / AE_T);
        L3D_CUDA(ctx->d_cands.ensure(n)); L3D_CUDA(ctx->d_ew.ensure(n)); L3D_CUDA(ctx->d_hscore.ensure(ctx->n_hyp)); L3D_CUDA(ctx->d_medians.ensure(size_t(ctx->V)));
        L3D_CUDA(ctx->d_ecnt.ensure(size_t(blocks) + 1)); L3D_CUDA(ctx->d_eoff.ensure(size_t(blocks) + 2));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_medians.p, median_depth, sizeof(float) * size_t(ctx->V), cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_hscore.p, hyp_score, sizeof(float) * ctx->n_hyp, cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(ctx->d_cands.p, cands, sizeof(l3d_candidate) * n, cudaMemcpyHostToDevice, st));
        const float two_log = 2.0f * logf(0.01f);                                   // view.cc:376
        k_affinity_edges<<<blocks, AE_T, 0, st>>>(ctx->d_hyp.p, ctx->d_viewd.p, ctx->d_medians.p, ctx->d_hscore.p, ctx->d_cands.p, n, sigma_a, two_log, ctx->d_ew.p, ctx->d_ecnt.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        if (scan_exclusive(ctx, ctx->d_ecnt.p, ctx->d_eoff.p, blocks)) return 1;
        int total = 0;
        if (fetch_int(ctx, ctx->d_eoff.p + blocks, &total)) return 1;
        if (total == 0) return 0;
        L3D_CUDA(ctx->d_kept.ensure(size_t(total))); L3D_CUDA(ctx->h_kept.ensure(size_t(total)));
        k_affinity_compact<<<blocks, AE_T, 0, st>>>(ctx->d_ew.p, n, ctx->d_eoff.p, ctx->d_kept.p);
        ctx->launches += 1;
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(ctx->h_kept.p, ctx->d_kept.p, sizeof(l3d_kept_edge) * size_t(total), cudaMemcpyDeviceToHost, st));
        if (order) {                                        // the clustering's edge order, while the records travel
            int n_pad = BS_TILE; while (n_pad < total) n_pad <<= 1;
            L3D_CUDA(ctx->d_ekeys.ensure(size_t(n_pad))); L3D_CUDA(ctx->d_eorder.ensure(size_t(total))); L3D_CUDA(ctx->h_eorder.ensure(size_t(total)));
            k_edge_keys<<<(n_pad + 255) / 256, 256, 0, st>>>(ctx->d_kept.p, total, n_pad, ctx->d_ekeys.p);
            ctx->launches += 1;
            for (int k = 2; k <= n_pad; k <<= 1) {
                int j = k >> 1;
                for (; j >= BS_TILE; j >>= 1) { k_bitonic_global<<<n_pad / 256, 256, 0, st>>>(ctx->d_ekeys.p, j, k); ctx->launches += 1; }
                k_bitonic_shared<<<n_pad / BS_TILE, BS_TILE / 2, 0, st>>>(ctx->d_ekeys.p, j, k);
                ctx->launches += 1;
            }
            k_edge_order<<<(total + 255) / 256, 256, 0, st>>>(ctx->d_ekeys.p, total, ctx->d_eorder.p);
            ctx->launches += 1;
            L3D_CUDA(cudaGetLastError());
            L3D_CUDA(cudaMemcpyAsync(ctx->h_eorder.p, ctx->d_eorder.p, sizeof(unsigned) * size_t(total), cudaMemcpyDeviceToHost, st));
        }
        L3D_CUDA(cudaStreamSynchronize(st));
        *kept = ctx->h_kept.p; *n_kept = size_t(total);
        if (order) *order = ctx->h_eorder.p;
        return 0;
    });
}

int l3d_diffuse(l3d_ctx* ctx, int n_nodes, size_t nnz, const l3d_edge* edges, l3d_edge* out)
{
    if (!ctx) return 1;
    return guarded(ctx, [&]() -> int {
        if (nnz == 0 || n_nodes <= 0) return 0;
        if (!edges || !out) { ctx->err = "l3d_diffuse: null argument"; return 1; }
        if (nnz > size_t(0x7ffffff0)) { ctx->err = "l3d_diffuse: more than 2^31 entries"; return 1; }
        L3D_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int N = int(nnz);
        for (int k = 0; k < N; ++k)
            if (edges[k].i < 0 || edges[k].i >= n_nodes || edges[k].j < 0 || edges[k].j >= n_nodes) { ctx->err = "l3d_diffuse: node id out of range"; return 1; }
        // W column-sorted (sparsematrix.cc:81-85), P row-sorted (sparsematrix.cc:158-161); the comparators order by the
        // integer ids only, stable, so equal keys (none for a valid input) keep their list order.  Two stable counting
        // passes per order (secondary key first) instead of comparison sorts: O(nnz + nodes).
        auto order_by = [&](const std::vector<int>& in, bool by_col_then_row, std::vector<int>& outv) {
            std::vector<int> tmp(N), cnt(size_t(n_nodes) + 1);
            for (int pass = 0; pass < 2; ++pass) {
                const bool use_i = by_col_then_row ? (pass == 0) : (pass == 1);          // secondary key first
                const std::vector<int>& src = pass == 0 ? in : tmp;
                std::vector<int>& dst = pass == 0 ? tmp : outv;
                std::fill(cnt.begin(), cnt.end(), 0);
                for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[src[k]]; ++cnt[size_t(use_i ? e.i : e.j) + 1]; }
                for (int d = 0; d < n_nodes; ++d) cnt[size_t(d) + 1] += cnt[d];
                for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[src[k]]; dst[cnt[use_i ? e.i : e.j]++] = src[k]; }
            }
        };
        std::vector<int> ident(N), wi(N), pi(N);
        for (int k = 0; k < N; ++k) ident[k] = k;
        order_by(ident, true, wi);                                                  // W: by (col, row)
        order_by(wi, false, pi);                                                    // P is built from W's stored order: by (row, col); pi holds input indices
        // pinned staging, one buffer: prow pcol wcol tpos (N ints each), rstart cstart (n_nodes each), pval wval (N floats each)
        const size_t n_int = size_t(4) * N + size_t(2) * n_nodes, n_all = n_int + size_t(2) * N;
        L3D_CUDA(ctx->h_rdd.ensure(n_all)); L3D_CUDA(ctx->d_rdd.ensure(n_all + size_t(N)));
        int* prow = ctx->h_rdd.p; int* pcol = prow + N; int* wcol = pcol + N; int* tpos = wcol + N; int* rstart = tpos + N; int* cstart = rstart + n_nodes;
        float* pval = reinterpret_cast<float*>(cstart + n_nodes); float* wval = pval + N;
        for (int d = 0; d < n_nodes; ++d) { rstart[d] = -1; cstart[d] = -1; }
        for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[wi[k]]; wcol[k] = e.j; wval[k] = e.w; if (cstart[e.j] < 0) cstart[e.j] = k; }
        for (int k = 0; k < N; ++k) { const l3d_edge& e = edges[pi[k]]; prow[k] = e.i; pcol[k] = e.j; pval[k] = e.w; if (rstart[e.i] < 0) rstart[e.i] = k; }
        // cudawrapper.cu:809-827: first entry (r, c) of row r = col_y with col == row_y (the transposed entry's place);
        // the rows are sorted by column, so it is found by bisection; and the algorithmic traffic of one iteration:
        // per entry the lock-step walk reads min(row length, column length) pairs of 16 bytes from each operand
        double walk = 0.0;
        {
            std::vector<int> rlen(size_t(n_nodes), 0), clen(size_t(n_nodes), 0);
            for (int k = 0; k < N; ++k) { ++rlen[prow[k]]; ++clen[wcol[k]]; }
            const size_t nt = std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), 16));
            std::vector<double> part(nt, 0.0);
            auto work = [&](size_t t) {
                const int y0 = int(size_t(N) * t / nt), y1 = int(size_t(N) * (t + 1) / nt);
                for (int y = y0; y < y1; ++y) {
                    const int r = pcol[y], c = prow[y];
                    int lo = rstart[r], hi = lo < 0 ? -1 : lo + rlen[r];
                    int found = -1;
                    while (lo >= 0 && lo < hi) { const int mid = (lo + hi) >> 1; if (pcol[mid] < c) lo = mid + 1; else hi = mid; }
                    if (lo >= 0 && lo < rstart[r] + rlen[r] && prow[lo] == r && pcol[lo] == c) found = lo;
                    tpos[y] = found;
                    part[t] += double(std::min(rlen[r], clen[c]));
                }
            };
            std::vector<std::thread> pool;
            for (size_t t = 1; t < nt; ++t) pool.emplace_back(work, t);
            work(0);
            for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
            for (size_t t = 0; t < nt; ++t) walk += part[t];
        }
        int* d_prow = ctx->d_rdd.p; int* d_pcol = d_prow + N; int* d_wcol = d_pcol + N; int* d_tpos = d_wcol + N; int* d_rstart = d_tpos + N; int* d_cstart = d_rstart + n_nodes;
        float* d_pa = reinterpret_cast<float*>(d_cstart + n_nodes); float* d_wval = d_pa + N; float* d_pb = d_wval + N;
        L3D_CUDA(cudaMemcpyAsync(ctx->d_rdd.p, ctx->h_rdd.p, sizeof(int) * n_all, cudaMemcpyHostToDevice, st));
        L3D_CUDA(cudaMemcpyAsync(d_pb, d_pa, sizeof(float) * N, cudaMemcpyDeviceToDevice, st));
        const int rb = (n_nodes + 255) / 256, eb = (N + 255) / 256;
        k_rdd_normalise<<<rb, 256, 0, st>>>(d_prow, d_pa, d_rstart, n_nodes, N);
        ctx->launches += 1;
        float* P = d_pa; float* Pn = d_pb;
        float step_ms = 0.f;
        for (int it = 0; it < 10; ++it) {                                           // L3D_RDD_MAX_ITER, cudawrapper.h:35
            if (it == 1) L3D_CUDA(cudaEventRecord(ctx->ev[0], st));                 // the step kernel's own duration: iteration 2 (warm)
            k_rdd_step<<<eb, 256, 0, st>>>(d_prow, d_pcol, P, d_wcol, d_wval, d_rstart, d_cstart, d_tpos, Pn, N);
            if (it == 1) L3D_CUDA(cudaEventRecord(ctx->ev[1], st));
            std::swap(P, Pn);
            ctx->launches += 1;
            if (it < 9) { k_rdd_normalise<<<rb, 256, 0, st>>>(d_prow, P, d_rstart, n_nodes, N); ctx->launches += 1; }
        }
        L3D_CUDA(cudaGetLastError());
        L3D_CUDA(cudaMemcpyAsync(pval, P, sizeof(float) * N, cudaMemcpyDeviceToHost, st));
        L3D_CUDA(cudaStreamSynchronize(st));
        cudaEventElapsedTime(&step_ms, ctx->ev[0], ctx->ev[1]);
        ctx->rdd_step_ms = step_ms; ctx->rdd_step_bytes = 32.0 * walk + 16.0 * double(N); ctx->rdd_nnz = N;
        for (int k = 0; k < N; ++k) { out[k].i = prow[k]; out[k].j = pcol[k]; out[k].w = pval[k]; }
        return 0;
    });
}

int l3d_diffuse_stats(l3d_ctx* ctx, double out[3])
{
    if (!ctx || !out) return 1;
    out[0] = ctx->rdd_step_ms; out[1] = ctx->rdd_step_bytes; out[2] = double(ctx->rdd_nnz);
    return 0;
}

int l3d_debug_math(l3d_ctx* ctx, int which, int n, const float* in, float* out)
{
    return guarded(ctx, [&]() -> int {
        if (!ctx) return 1;
        if (n <= 0) return 0;
        L3D_CUDA(cudaSetDevice(ctx->device));
        float *d_in, *d_out;
        L3D_CUDA(cudaMalloc(&d_in, sizeof(float) * n)); L3D_CUDA(cudaMalloc(&d_out, sizeof(float) * n));
        L3D_CUDA(cudaMemcpyAsync(d_in, in, sizeof(float) * n, cudaMemcpyHostToDevice, ctx->stream));
        k_debug_math<<<(n + 255) / 256, 256, 0, ctx->stream>>>(which, d_in, d_out, n);
        ctx->launches += 1;
        L3D_CUDA(cudaMemcpyAsync(out, d_out, sizeof(float) * n, cudaMemcpyDeviceToHost, ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        cudaFree(d_in); cudaFree(d_out);
        return 0;
    });
}

int l3d_last_timings(l3d_ctx* ctx, float ms[6])
{
    if (!ctx || !ms) return 1;
    for (int k = 0; k < 6; ++k) ms[k] = ctx->last_ms[k];
    return 0;
}

int l3d_set_stage_timing(l3d_ctx* ctx, int on)
{
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    drain_pending(ctx);
    ctx->stage_events = on != 0;
    return 0;
}

int l3d_mark(l3d_ctx* ctx, int which)
{
    if (!ctx || which < 0 || which > 1) return 1;
    L3D_CUDA(cudaSetDevice(ctx->device));
    if (which == 1) {            // the closing mark also comes after the side stream's last result records
        L3D_CUDA(cudaEventRecord(ctx->ev[1], ctx->fin_stream));
        L3D_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev[1], 0));
    }
    L3D_CUDA(cudaEventRecord(ctx->mark_ev[which], ctx->stream));
    return 0;
}

int l3d_marks_elapsed(l3d_ctx* ctx, float* ms)
{
    if (!ctx || !ms) return 1;
    L3D_CUDA(cudaSetDevice(ctx->device));
    L3D_CUDA(cudaEventSynchronize(ctx->mark_ev[1]));
    L3D_CUDA(cudaEventElapsedTime(ms, ctx->mark_ev[0], ctx->mark_ev[1]));
    return 0;
}

int l3d_last_kernel_timings(l3d_ctx* ctx, float ms[3])
{
    if (!ctx || !ms) return 1;
    ms[0] = ctx->pa_kernel_ms[0]; ms[1] = ctx->pa_kernel_ms[1]; ms[2] = ctx->pa_kernel_ms[2];
    return 0;
}

int l3d_flush_l2(l3d_ctx* ctx)
{
    if (!ctx) return 1;
    L3D_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = size_t(256) << 20;
    L3D_CUDA(ctx->d_flush.ensure(bytes));
    L3D_CUDA(cudaMemsetAsync(ctx->d_flush.p, ctx->flush_value++ & 0xff, bytes, ctx->stream));
    L3D_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// which = 0: scalar FFMA chains, 1: packed FFMA2 chains.  Best of five launches.
static int measure_fp32(l3d_ctx* ctx, int which, float* tflops)
{
    L3D_CUDA(cudaSetDevice(ctx->device));
    cudaDeviceProp prop; L3D_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 13;
    float* d = nullptr;
    L3D_CUDA(cudaMalloc(&d, sizeof(float) * size_t(blocks) * threads));
    const float b = 1.0000001f, c = 1e-7f;
    float best = 0.f;
    for (int r = 0; r < 6; ++r) {                       // the first launch warms up
        L3D_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
        if (which == 0) k_fma_probe<<<blocks, threads, 0, ctx->stream>>>(d, iters, b, c);
        else k_fma2_probe<<<blocks, threads, 0, ctx->stream>>>(d, iters, b, c);
        L3D_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
        L3D_CUDA(cudaStreamSynchronize(ctx->stream));
        float ms = 0.f; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        const double fl = double(blocks) * threads * double(iters) * 128.0 * 2.0;      // 128 FMAs per thread and loop trip in both variants
        if (r > 0) best = std::max(best, float(fl / (double(ms) * 1e-3) / 1e12));
    }
    ctx->launches += 6;
    cudaFree(d);
    *tflops = best;
    return 0;
}

int l3d_measure_fp32_peak(l3d_ctx* ctx, float* tflops)
{
    if (!ctx || !tflops) return 1;
    return measure_fp32(ctx, 0, tflops);
}

int l3d_measure_fp32x2_peak(l3d_ctx* ctx, float* tflops)
{
    if (!ctx || !tflops) return 1;
    return measure_fp32(ctx, 1, tflops);
}

}  // extern "C"
