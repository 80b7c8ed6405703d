// line3d_b200/csrc/l3d_common.cuh -- error handling, the NCCL loader, grow-only buffers, the per-view tables and the exclusive scans shared by every stage.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

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

}  // namespace
