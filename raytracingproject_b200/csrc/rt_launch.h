// rt_launch.h -- host-visible launch interface of the per-precision kernel translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_trace.cuh"

namespace rt {

// megakernel shape: float 512 threads x 2 CTAs/SM (<= 64 regs), double 256 x 2 (<= 128 regs)
#define RT_MEGA_BLOCK(T) (sizeof(T) == 4 ? 512 : 256)
#define RT_MEGA_MINBLOCKS(T) 2
// warp-scheduled megakernel: fewer, fatter threads (more state is live across the sections)
// block shapes compiled (2 CTAs per SM each): 256 (128 regs), 320 (96), 384 (80), 512 (64)
#define RT_SCHED_SHAPES 4
inline int sched_block(int shape) { const int b[RT_SCHED_SHAPES] = { 256, 320, 384, 512 }; return b[shape < 0 || shape >= RT_SCHED_SHAPES ? 0 : shape]; }
#define RT_SCHED_DEFAULT_SHAPE(T) (sizeof(T) == 4 ? 2 : 0)

#define RT_MAX_SLICES 64

template <typename T> struct RenderParams {
    CameraFrame<T> cam;
    SceneView<T> scene;
    typename Quad<T>::type* accum;      // [slices][H][W] (sum.rgb, samples)
    unsigned int* job_counter;
    unsigned long long* ray_counter;
    uint64_t seed;
    PhiloxKeys keys;                    // the ten Philox round keys of `seed` (wavefront kernel: constant-bank operands)
    int sample_begin, sample_count, max_depth;
    int slices, tiles_x, tiles_y;
    int job_group;                      // wavefront: jobs fetched at a time = horizontally adjacent tiles of one layer
    int bulk_layers;                    // wavefront job order: tile-major over the first bulk_layers layers, then layer by layer
    int slice_begin[RT_MAX_SLICES + 1]; // wavefront: samples [slice_begin[k], slice_begin[k+1]) of a pixel form layer k (sizes may differ)
    int accumulate;
    int tile_rank, tile_world;          // wavefront: tile-interleaved sharding -- this launch renders the tile groups g with g % tile_world == tile_rank
    int th_s, th_l, th_n;               // section thresholds of the warp schedulers (render_sched_kernel, render_pool_kernel)
    int job_stride;                     // diagnostic (RT_JOB_STRIDE): a fetch advances the counter by job_group * job_stride -- this launch then renders every job_stride-th group only
    int jobs_shared;                    // wavefront: job_counter is shared with other GPUs (it may live in a peer's memory): system-scope atomics
};

struct HitRec;
struct MatRec;

// bytes of shared memory the staged scene needs
template <typename T> inline size_t staged_bytes(int n_nodes, int n_slots) {
    return size_t(n_nodes) * 64 + size_t(n_slots) * (3 * sizeof(typename Quad<T>::type) + sizeof(int));
}

// the same, rounded up so that what follows the staged scene in shared memory stays 16-byte aligned
template <typename T> __host__ __device__ inline size_t staged_bytes_aligned(int n_nodes, int n_slots) {
    return (size_t(n_nodes) * 64 + size_t(n_slots) * (3 * sizeof(typename Quad<T>::type) + sizeof(int)) + 15) & ~size_t(15);
}

// SM-resident wavefront (rt_wavefront.cuh): pool of P paths per warp, one CTA of BLOCK threads per SM
#ifndef RT_RING
#define RT_RING 2                      // jobs a warp keeps open (rt_wavefront.cuh)
#endif
#define RT_POOL_SHAPES 10
inline void pool_shape(int shape, int* paths, int* block) {
    const int P[RT_POOL_SHAPES] = { 128, 128, 128, 128, 128, 128, 96, 64, 128, 128 }, B[RT_POOL_SHAPES] = { 256, 512, 640, 768, 704, 736, 896, 1024, 896, 1024 };
    if (shape < 0 || shape >= RT_POOL_SHAPES) shape = 0;
    *paths = P[shape]; *block = B[shape];
}
#ifndef RT_DYNAMIC_ITEMS
#define RT_DYNAMIC_ITEMS 1             // (rt_wavefront.cuh)
#endif
template <typename T> inline size_t pool_warp_bytes(int paths) {      // + the warp's job descriptors (dynamic mode, float: 4 entries of 32 bytes)
    return size_t(paths) * (12 * sizeof(T) + 2 * 2 + 4) + ((RT_DYNAMIC_ITEMS != 0 && sizeof(T) == 4) ? 80 : 16 * RT_RING);      // three quads of T, two 16-bit words, four list bytes
}
#define RT_POOL_DEFAULT_SHAPE(T) (sizeof(T) == 4 ? 3 : 0)
#define RT_POOL_UNSTAGED_SHAPE(T) (sizeof(T) == 4 ? 9 : 0)
template <typename T> cudaError_t launch_render_pool(const RenderParams<T>& p, int shape, bool staged, int sm_count, cudaStream_t st);

template <typename T> cudaError_t launch_render_mega(const RenderParams<T>& p, bool staged, size_t smem_bytes, int sm_count, cudaStream_t st);
template <typename T> cudaError_t launch_render_sched(const RenderParams<T>& p, int shape, bool staged, size_t smem_bytes, int sm_count, cudaStream_t st);
template <typename T> cudaError_t launch_trace_closest(const SceneView<T>& sv, const double* rays, size_t n, double tmin, double tmax,
                                                       int brute, HitRec* out, cudaStream_t st);
template <typename T> cudaError_t launch_scatter(const MatRec* mats, const double* rays_in, const HitRec* hits, size_t n, uint64_t seed,
                                                 int* flag, double* att, double* rays_out, cudaStream_t st);
template <typename T> cudaError_t launch_gen_rays(const CameraFrame<T>& cam, const int* px_i, const int* px_j, const uint32_t* sample,
                                                  size_t n, uint64_t seed, double* rays_out, cudaStream_t st);
template <typename T> cudaError_t launch_bench_closest(const SceneView<T>& sv, const CameraFrame<T>& cam, size_t n, uint64_t seed, int brute,
                                                       bool staged, size_t smem_bytes, int sm_count, int* prim_out, float* t_out,
                                                       double* rays_out, size_t rays_out_n, cudaStream_t st);

}  // namespace rt
