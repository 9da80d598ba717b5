// rt_kernels_impl.inl -- launchers, included by rt_kernels_f32.cu / rt_kernels_f64.cu with RT_REAL set.
#include "rt_kernels.cuh"
#include "rt_wavefront.cuh"

namespace rt {

template <> cudaError_t launch_render_mega<RT_REAL>(const RenderParams<RT_REAL>& p, bool staged, size_t smem_bytes, int sm_count, cudaStream_t st) {
    const int block = RT_MEGA_BLOCK(RT_REAL);
    const int grid = sm_count * RT_MEGA_MINBLOCKS(RT_REAL);       // persistent: resident CTAs only
    if (staged) {
        cudaError_t e = cudaFuncSetAttribute(render_mega_kernel<RT_REAL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_bytes));
        if (e != cudaSuccess) return e;
        render_mega_kernel<RT_REAL, true><<<grid, block, smem_bytes, st>>>(p);
    } else {
        render_mega_kernel<RT_REAL, false><<<grid, block, 0, st>>>(p);
    }
    return cudaGetLastError();
}

template <bool STAGED, int BLOCK>
static cudaError_t launch_sched_shape(const RenderParams<RT_REAL>& p, size_t smem_bytes, int sm_count, cudaStream_t st) {
    if (STAGED) {
        cudaError_t e = cudaFuncSetAttribute(render_sched_kernel<RT_REAL, STAGED, BLOCK>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_bytes));
        if (e != cudaSuccess) return e;
    }
    render_sched_kernel<RT_REAL, STAGED, BLOCK><<<sm_count * 2, BLOCK, STAGED ? smem_bytes : 0, st>>>(p);   // persistent: resident CTAs only
    return cudaGetLastError();
}

template <> cudaError_t launch_render_sched<RT_REAL>(const RenderParams<RT_REAL>& p, int shape, bool staged, size_t smem_bytes, int sm_count, cudaStream_t st) {
    switch (sched_block(shape)) {
    case 256: return staged ? launch_sched_shape<true, 256>(p, smem_bytes, sm_count, st) : launch_sched_shape<false, 256>(p, smem_bytes, sm_count, st);
    case 320: return staged ? launch_sched_shape<true, 320>(p, smem_bytes, sm_count, st) : launch_sched_shape<false, 320>(p, smem_bytes, sm_count, st);
    case 384: return staged ? launch_sched_shape<true, 384>(p, smem_bytes, sm_count, st) : launch_sched_shape<false, 384>(p, smem_bytes, sm_count, st);
    default:  return staged ? launch_sched_shape<true, 512>(p, smem_bytes, sm_count, st) : launch_sched_shape<false, 512>(p, smem_bytes, sm_count, st);
    }
}

template <bool STAGED, int P, int BLOCK, bool WIDE = false>
static cudaError_t launch_pool_shape(const RenderParams<RT_REAL>& p, int sm_count, cudaStream_t st) {
    static_assert(Pool<RT_REAL, P>::bytes == int(P * (12 * sizeof(RT_REAL) + 2 * 2 + 4) + ((RT_DYNAMIC_ITEMS != 0 && sizeof(RT_REAL) == 4) ? 80 : 16 * RT_RING)), "pool_warp_bytes mirrors Pool::bytes");
    const size_t smem = (STAGED ? staged_bytes_aligned<RT_REAL>(p.scene.n_nodes, p.scene.n_slots) : (WIDE ? 0 : size_t(p.scene.n_top) * 64)) +
                        size_t(BLOCK / 32) * Pool<RT_REAL, P>::bytes;
    cudaError_t e = cudaFuncSetAttribute(render_pool_kernel<RT_REAL, STAGED, P, BLOCK, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    render_pool_kernel<RT_REAL, STAGED, P, BLOCK, WIDE><<<sm_count, BLOCK, smem, st>>>(p);      // persistent: one CTA per SM
    return cudaGetLastError();
}

template <> cudaError_t launch_render_pool<RT_REAL>(const RenderParams<RT_REAL>& p, int shape, bool staged, int sm_count, cudaStream_t st) {
    const bool wide = !staged && p.scene.nodes4 != nullptr;
#define RT_POOL_CASE(S, P, B) case S: return staged ? launch_pool_shape<true, P, B>(p, sm_count, st) : (wide ? launch_pool_shape<false, P, B, true>(p, sm_count, st) : launch_pool_shape<false, P, B>(p, sm_count, st))
    switch (shape) {
    RT_POOL_CASE(1, 128, 512);
    RT_POOL_CASE(2, 128, 640);
    RT_POOL_CASE(3, 128, 768);
    RT_POOL_CASE(4, 128, 704);
    RT_POOL_CASE(5, 128, 736);
    RT_POOL_CASE(6, 96, 896);
    RT_POOL_CASE(7, 64, 1024);
    RT_POOL_CASE(8, 128, 896);
    RT_POOL_CASE(9, 128, 1024);
    default: RT_POOL_CASE(0, 128, 256);
    }
#undef RT_POOL_CASE
}

template <> cudaError_t launch_trace_closest<RT_REAL>(const SceneView<RT_REAL>& sv, const double* rays, size_t n, double tmin, double tmax,
                                                      int brute, HitRec* out, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    trace_closest_kernel<RT_REAL><<<unsigned((n + 127) / 128), 128, 0, st>>>(sv, rays, n, tmin, tmax, brute, out);
    return cudaGetLastError();
}

template <> cudaError_t launch_scatter<RT_REAL>(const MatRec* mats, const double* rays_in, const HitRec* hits, size_t n, uint64_t seed,
                                                int* flag, double* att, double* rays_out, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    scatter_kernel<RT_REAL><<<unsigned((n + 127) / 128), 128, 0, st>>>(mats, rays_in, hits, n, seed, flag, att, rays_out);
    return cudaGetLastError();
}

template <> cudaError_t launch_gen_rays<RT_REAL>(const CameraFrame<RT_REAL>& cam, const int* px_i, const int* px_j, const uint32_t* sample,
                                                 size_t n, uint64_t seed, double* rays_out, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    gen_rays_kernel<RT_REAL><<<unsigned((n + 127) / 128), 128, 0, st>>>(cam, px_i, px_j, sample, n, seed, rays_out);
    return cudaGetLastError();
}

template <> cudaError_t launch_bench_closest<RT_REAL>(const SceneView<RT_REAL>& sv, const CameraFrame<RT_REAL>& cam, size_t n, uint64_t seed, int brute,
                                                      bool staged, size_t smem_bytes, int sm_count, int* prim_out, float* t_out,
                                                      double* rays_out, size_t rays_out_n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const int grid = sm_count * 4;
    if (staged) {
        cudaError_t e = cudaFuncSetAttribute(bench_closest_kernel<RT_REAL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_bytes));
        if (e != cudaSuccess) return e;
        bench_closest_kernel<RT_REAL, true><<<grid, 256, smem_bytes, st>>>(sv, cam, n, seed, brute, prim_out, t_out, rays_out, rays_out_n);
    } else {
        bench_closest_kernel<RT_REAL, false><<<grid, 256, 0, st>>>(sv, cam, n, seed, brute, prim_out, t_out, rays_out, rays_out_n);
    }
    return cudaGetLastError();
}

}  // namespace rt
