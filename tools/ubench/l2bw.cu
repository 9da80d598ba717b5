// L2 / HBM read bandwidth as the BVH walk sees it ("L2 bandwidth = to be measured on the box").
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2bw l2bw.cu && ./l2bw
// Two access shapes over a working set of S bytes, S from L2-resident to HBM-resident:
//   stream : every thread reads consecutive float4 (coalesced 512 B per warp instruction)
//   gather : every thread reads one random 64-byte record (4 x LDG.128, the node record of rt_trace.cuh) per step,
//            the next record index depends on the data just read (a walk: one record in flight per thread)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

__global__ void __launch_bounds__(1024, 1) stream_kernel(const float4* __restrict__ buf, size_t n_vec, int rounds, float* out) {
    float acc = 0.f;
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (int r = 0; r < rounds; ++r)
        for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
            const float4 v = __ldcg(buf + i);               // L2 only: L1 does not help a 48 MB walk
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 12345.678f) out[0] = acc;
}

__global__ void __launch_bounds__(1024, 1) gather_kernel(const float4* __restrict__ rec, unsigned n_rec, int steps, float* out) {
    unsigned idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u % n_rec;
    float acc = 0.f;
    for (int s = 0; s < steps; ++s) {
        const float4 a = __ldcg(rec + 4 * size_t(idx)), b = __ldcg(rec + 4 * size_t(idx) + 1), c = __ldcg(rec + 4 * size_t(idx) + 2), d = __ldcg(rec + 4 * size_t(idx) + 3);
        acc += a.x + b.y + c.z;
        idx = (__float_as_uint(d.w) + idx * 1664525u + 1013904223u) % n_rec;   // d.w holds a random 32-bit pattern
    }
    if (acc == 12345.678f) out[0] = acc;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("# %s, %d SMs, L2 %.0f MB\n", p.name, sms, p.l2CacheSize / 1048576.0);
    float* out; cudaMalloc(&out, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const size_t sizes_mb[] = {8, 16, 32, 48, 64, 96, 128, 256, 1024};
    for (size_t mb : sizes_mb) {
        const size_t bytes = mb << 20, n_vec = bytes / 16;
        float4* buf; cudaMalloc(&buf, bytes);
        std::vector<unsigned> h(bytes / 4);
        unsigned x = 12345u;
        for (auto& w : h) { x = x * 1664525u + 1013904223u; w = (x >> 9) | 0x3f800000u; }   // floats in [1,2): sums stay finite
        for (size_t r = 0; r < bytes / 64; ++r) { x = x * 1664525u + 1013904223u; h[r * 16 + 15] = x; }
        cudaMemcpy(buf, h.data(), bytes, cudaMemcpyHostToDevice);
        const int rounds = int(std::max<size_t>(2, (size_t(4) << 30) / bytes));
        float best_s = 1e30f, best_g = 1e30f;
        const int steps = 256;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0); stream_kernel<<<sms, 1024>>>(buf, n_vec, rounds, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep) best_s = ms < best_s ? ms : best_s;
            cudaEventRecord(e0); gather_kernel<<<sms * 2, 1024>>>(buf, unsigned(bytes / 64), steps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1); if (rep) best_g = ms < best_g ? ms : best_g;
        }
        const double gs = double(bytes) * rounds / best_s / 1e6, gg = double(sms) * 2 * 1024 * steps * 64 / best_g / 1e6;
        printf("working set %5zu MB  stream %8.0f GB/s  gather(64 B records, dependent) %8.0f GB/s = %7.0f M records/s\n", mb, gs, gg, gg / 64 * 1e3);
        cudaFree(buf);
    }
    if (cudaGetLastError() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    return 0;
}
