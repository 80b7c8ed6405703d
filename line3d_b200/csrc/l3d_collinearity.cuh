// line3d_b200/csrc/l3d_collinearity.cuh -- collinearity of the segments of one image, and the entry point that never allocates.
// Part of the single translation unit l3d_kernels.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

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

}  // namespace
