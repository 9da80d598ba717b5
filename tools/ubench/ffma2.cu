// Issue-rate probe for the packed fp32 FMA of sm_100a (FFMA2) against FFMA and FMNMX.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu && ./ffma2
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define UNROLL 8

template <int MODE>
__global__ void __launch_bounds__(1024, 1) probe(float* out, float seed) {
    float a[UNROLL], b[UNROLL];
    unsigned long long p[UNROLL];
    for (int k = 0; k < UNROLL; ++k) { a[k] = seed + k + threadIdx.x; b[k] = seed * k; p[k] = (unsigned long long)__float_as_uint(a[k]) << 32 | __float_as_uint(b[k]); }
    const float m = 1.0000001f, c = 1e-7f;
    const unsigned long long mm = (unsigned long long)__float_as_uint(m) << 32 | __float_as_uint(m), cc = (unsigned long long)__float_as_uint(c) << 32 | __float_as_uint(c);
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            if (MODE == 0) { a[k] = fmaf(a[k], m, c); b[k] = fmaf(b[k], m, c); }                         // 2 FFMA
            if (MODE == 1) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[k]) : "l"(mm), "l"(cc)); }   // 1 FFMA2 (same flops)
            if (MODE == 2) { a[k] = fminf(a[k], b[k] + 0.0f * i); b[k] = fmaxf(b[k], a[k]); }            // min/max
            if (MODE == 3) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[k]) : "l"(mm), "l"(cc)); a[k] = fminf(a[k], b[k]); b[k] = fmaxf(b[k], seed); }
        }
    }
    float s = 0;
    for (int k = 0; k < UNROLL; ++k) s += a[k] + b[k] + __uint_as_float((unsigned)(p[k] >> 32)) + __uint_as_float((unsigned)p[k]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE> void run(const char* name, double inst_per_iter) {
    float* out; cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<148, 1024>>>(out, 1.0f); cudaDeviceSynchronize();
    cudaEventRecord(e0); probe<MODE><<<148, 1024>>>(out, 1.0f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warp_inst = 148.0 * 32 * ITERS * UNROLL * inst_per_iter;      // 32 warps per SM
    printf("%-28s %8.3f ms  %7.1f G warp-inst/s  (%.2f per clk per SM at 1.965 GHz)\n", name, ms, warp_inst / ms / 1e6, warp_inst / ms / 1e6 / 148 / 1.965);
    cudaFree(out);
}

int main() {
    run<0>("2 x FFMA", 2);
    run<1>("1 x FFMA2 (same flops)", 1);
    run<2>("FADD + FMNMX + FMNMX", 3);
    run<3>("FFMA2 + 2 x FMNMX", 3);
    return 0;
}
