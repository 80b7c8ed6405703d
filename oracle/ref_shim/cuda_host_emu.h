// oracle/ref_shim/cuda_host_emu.h  --  TEST INFRASTRUCTURE, not product code.
//
// Force-included (g++ -include) when oracle/Makefile compiles the reference's own
// cudawrapper.cu AS HOST C++, from where it lies under /root/reference, into
// oracle/_ref/.  The reference's CUDA file cannot be built by nvcc >= 12.0 (it uses the
// removed texture-reference API, cudawrapper.cu:5-10) and this container has no GPU, so
// the kernels are executed on the CPU instead: one C++ call per CUDA thread, in
// blockIdx/threadIdx order.  This header supplies what the device side of that file
// expects and g++ lacks:
//   * threadIdx / blockIdx / blockDim / gridDim            (thread-local variables)
//   * texture<T,2,mode>, cudaBindTexture2D, cudaUnbindTexture, tex2D
//     - tex2D() returns the texel whose centre is addressed.  The reference always
//       fetches at (index + 0.5) with linear filtering and unnormalised coordinates
//       (cudawrapper.cu:156,298,365,407,551-558), where the bilinear weights are
//       exactly 0/1, so this is the value the hardware returned.
//   * REF_LAUNCH(kernel, grid, block)(args...)  -- the Makefile's sed rewrites the six
//     `kernel <<< grid, block >>> (args)` launch sites (cudawrapper.cu:849,915,1013,
//     1151,1164,1178) into this form while piping the file to g++; nothing else in the
//     file is touched and no copy of it is written to disk.
// The CUDA runtime entry points DataArray<T> calls (dataArray.h:140-279) are defined in
// cuda_host_emu.cpp on top of malloc/memcpy.
#ifndef L3D_ORACLE_CUDA_HOST_EMU_H_
#define L3D_ORACLE_CUDA_HOST_EMU_H_

#include <cuda_runtime.h>
#include <cmath>
#include <cstddef>

extern thread_local uint3 ref_emu_threadIdx;
extern thread_local uint3 ref_emu_blockIdx;
extern thread_local dim3 ref_emu_blockDim;
extern thread_local dim3 ref_emu_gridDim;
#define threadIdx ref_emu_threadIdx
#define blockIdx ref_emu_blockIdx
#define blockDim ref_emu_blockDim
#define gridDim ref_emu_gridDim

template<class T, int Dim = 2, int Mode = cudaReadModeElementType>
struct texture {
    cudaTextureAddressMode addressMode[3];
    cudaTextureFilterMode filterMode;
    int normalized;
    const char* base; size_t pitch; size_t width, height;
    texture() : filterMode(cudaFilterModePoint), normalized(0), base(0), pitch(0), width(0), height(0) {}
};

template<class T, int Dim, int Mode>
inline cudaError_t cudaBindTexture2D(size_t* offset, texture<T, Dim, Mode>* tex, const void* devPtr,
                                     const cudaChannelFormatDesc*, size_t width, size_t height, size_t pitch)
{
    if (offset) *offset = 0;
    tex->base = static_cast<const char*>(devPtr); tex->pitch = pitch; tex->width = width; tex->height = height;
    return devPtr ? cudaSuccess : cudaErrorInvalidValue;
}
template<class T, int Dim, int Mode>
inline cudaError_t cudaUnbindTexture(texture<T, Dim, Mode>& tex) { tex.base = 0; return cudaSuccess; }

template<class T, int Dim, int Mode>
inline T tex2D(const texture<T, Dim, Mode>& tex, float x, float y)
{
    // texel-centre fetch with clamp addressing
    long ix = long(floorf(x)), iy = long(floorf(y));
    if (ix < 0) ix = 0; if (iy < 0) iy = 0;
    if (ix >= long(tex.width)) ix = long(tex.width) - 1;
    if (iy >= long(tex.height)) iy = long(tex.height) - 1;
    return *reinterpret_cast<const T*>(tex.base + size_t(iy) * tex.pitch + size_t(ix) * sizeof(T));
}

// number of host threads the emulated launches may use (1 = the reference's own,
// fully serial behaviour; >1 runs independent CUDA blocks concurrently)
extern "C" void ref_emu_set_threads(int n);
extern "C" int ref_emu_get_threads();
int ref_emu_threads_internal();

template<class K>
struct RefLaunch {
    K k; dim3 g, b;
    RefLaunch(K k_, dim3 g_, dim3 b_) : k(k_), g(g_), b(b_) {}
    template<class... A>
    void operator()(A... a) const
    {
        const long nblocks = long(g.x) * g.y * g.z;
        const int nt = ref_emu_threads_internal();
        #pragma omp parallel for schedule(dynamic, 4) num_threads(nt) if (nt > 1)
        for (long blk = 0; blk < nblocks; ++blk) {
            ref_emu_gridDim = g; ref_emu_blockDim = b;
            ref_emu_blockIdx.x = unsigned(blk % g.x);
            ref_emu_blockIdx.y = unsigned((blk / g.x) % g.y);
            ref_emu_blockIdx.z = unsigned(blk / (long(g.x) * g.y));
            for (unsigned tz = 0; tz < b.z; ++tz)
                for (unsigned ty = 0; ty < b.y; ++ty)
                    for (unsigned tx = 0; tx < b.x; ++tx) {
                        ref_emu_threadIdx.x = tx; ref_emu_threadIdx.y = ty; ref_emu_threadIdx.z = tz;
                        k(a...);
                    }
        }
    }
};
template<class K>
inline RefLaunch<K> ref_make_launch(K k, dim3 g, dim3 b) { return RefLaunch<K>(k, g, b); }
#define REF_LAUNCH(kernel, grid, block) ref_make_launch(kernel, grid, block)

#endif
