// rt_encode.cuh -- device helpers of the P3 text encoders (src/color.h:32-34: "r g b\n" with variable-width decimals),
// shared by the framebuffer passes (rt_frame.cu) and the allocation-free encoder (rt_static.cu).
#pragma once

#include <cuda_runtime.h>

namespace rt {

#define ENC_BLOCK 256
#define ENC_PPT 4                      // pixels per thread
#define ENC_TILE (ENC_BLOCK * ENC_PPT)

__device__ __forceinline__ int ndigits(unsigned v) { return v >= 100 ? 3 : (v >= 10 ? 2 : 1); }
__device__ __forceinline__ int line_len(const unsigned char* p) { return ndigits(p[0]) + ndigits(p[1]) + ndigits(p[2]) + 3; }

__device__ __forceinline__ unsigned block_exclusive_scan(unsigned v, unsigned* warp_sums, unsigned& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        unsigned w = lane < (ENC_BLOCK / 32) ? warp_sums[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        if (lane < (ENC_BLOCK / 32)) warp_sums[lane] = w;       // inclusive
    }
    __syncthreads();
    total = warp_sums[ENC_BLOCK / 32 - 1];
    const unsigned base = wid > 0 ? warp_sums[wid - 1] : 0;
    __syncthreads();
    return base + x - v;
}

__device__ __forceinline__ int put_number(char* dst, unsigned v) {
    int n = 0;
    if (v >= 100) { dst[n++] = char('0' + v / 100); v %= 100; dst[n++] = char('0' + v / 10); dst[n++] = char('0' + v % 10); }
    else if (v >= 10) { dst[n++] = char('0' + v / 10); dst[n++] = char('0' + v % 10); }
    else dst[n++] = char('0' + v);
    return n;
}

}  // namespace rt
