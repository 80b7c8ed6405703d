// oracle/ref_shim/cuda_host_emu.cpp  --  TEST INFRASTRUCTURE, not product code.
// Host definitions of the CUDA runtime entry points the reference's DataArray<T>
// (dataArray.h:140-279) and line3D.cc:645 call, so that the reference's sources link and
// run on a machine without a GPU: "device" memory is ordinary heap memory.
#include "cuda_host_emu.h"
#include <cstdlib>
#include <cstring>

#undef threadIdx
#undef blockIdx
#undef blockDim
#undef gridDim
thread_local uint3 ref_emu_threadIdx;
thread_local uint3 ref_emu_blockIdx;
thread_local dim3 ref_emu_blockDim;
thread_local dim3 ref_emu_gridDim;

static int g_threads = 1;
extern "C" void ref_emu_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
extern "C" int ref_emu_get_threads() { return g_threads; }
int ref_emu_threads_internal() { return g_threads; }

extern "C" {
cudaError_t cudaMallocPitch(void** devPtr, size_t* pitch, size_t width, size_t height)
{
    // same 512-byte row alignment a real device uses, so strides differ from widths
    size_t p = (width + 511) / 512 * 512;
    void* m = std::calloc(p * (height ? height : 1), 1);
    if (!m) return cudaErrorMemoryAllocation;
    *devPtr = m; *pitch = p;
    return cudaSuccess;
}
cudaError_t cudaMemcpy2D(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t height, cudaMemcpyKind)
{
    for (size_t r = 0; r < height; ++r)
        std::memcpy(static_cast<char*>(dst) + r * dpitch, static_cast<const char*>(src) + r * spitch, width);
    return cudaSuccess;
}
cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
cudaError_t cudaDeviceSynchronize(void) { return cudaSuccess; }
const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA error"; }
cudaChannelFormatDesc cudaCreateChannelDesc(int x, int y, int z, int w, cudaChannelFormatKind f)
{
    cudaChannelFormatDesc d; d.x = x; d.y = y; d.z = z; d.w = w; d.f = f; return d;
}
cudaError_t cudaMemGetInfo(size_t* free_b, size_t* total_b) { *free_b = size_t(1) << 36; *total_b = size_t(1) << 36; return cudaSuccess; }
}
