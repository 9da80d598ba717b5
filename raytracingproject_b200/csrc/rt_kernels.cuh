// rt_kernels.cuh -- kernels of the render path, templated on the arithmetic type.
// Instantiated twice: rt_kernels_f32.cu (FMA contraction on) and rt_kernels_f64.cu
// (--fmad=false, parity mode).
#pragma once

#include "rt_trace.cuh"
#include "rt_launch.h"

namespace rt {

// Cooperative copy of the scene into shared memory (the whole final scene -- 483 nodes and 484
// sphere slots -- is ~55 KB in float, ~78 KB in double).  Layout: nodes | sa | sb | sc | meta.
template <typename T> struct Staged {
    typedef typename Quad<T>::type Q;
    const float4* nodes; const Q* sa; const Q* sb; const Q* sc; const int* meta;
};

template <typename T>
__device__ __forceinline__ Staged<T> stage_scene(const SceneView<T>& sv, unsigned char* smem, bool staged) {
    typedef typename Quad<T>::type Q;
    Staged<T> st;
    if (!staged) {
        st.nodes = sv.nodes; st.sa = sv.sa; st.sb = sv.sb; st.sc = sv.sc; st.meta = sv.meta;
        return st;
    }
    // offsets come precomputed from the host (constant bank): one add per use when the compiler
    // rematerialises these pointers inside the hot loops instead of a chain of multiplies
    float4* n = reinterpret_cast<float4*>(smem);
    Q* a = reinterpret_cast<Q*>(smem + sv.off_sa);
    Q* b = reinterpret_cast<Q*>(smem + sv.off_sb);
    Q* c = reinterpret_cast<Q*>(smem + sv.off_sc);
    int* m = reinterpret_cast<int*>(smem + sv.off_meta);
    for (int i = threadIdx.x; i < 4 * sv.n_nodes; i += blockDim.x) n[i] = sv.nodes[i];
    for (int i = threadIdx.x; i < sv.n_slots; i += blockDim.x) { a[i] = sv.sa[i]; b[i] = sv.sb[i]; c[i] = sv.sc[i]; m[i] = sv.meta[i]; }
    __syncthreads();
    st.nodes = n; st.sa = a; st.sb = b; st.sc = c; st.meta = m;
    return st;
}

// ------------------------------------------------------------------------- megakernel --
// camera::render's pixel / sample loops (src/camera.h:37-47) and ray_color's recursion
// (src/camera_cpu.h:8-26) as ONE persistent kernel.  A warp owns an 8x4 pixel tile and one
// sample slice; each lane walks its pixel's samples in order (so the per-pixel sum is formed in
// the reference's order), and the bounce recursion is the iterative throughput loop -- it is
// tail-multiplicative.  The loop is flattened: one iteration = one path segment
// per lane, a lane whose path ended starts its next sample at once (no waiting for the longest
// path of the warp).
template <typename T, bool STAGED>
__global__ void __launch_bounds__(RT_MEGA_BLOCK(T), RT_MEGA_MINBLOCKS(T))
render_mega_kernel(const RenderParams<T> p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const Staged<T> st = stage_scene<T>(p.scene, smem, STAGED);
    const SceneView<T>& sv = p.scene;

    const int lane = threadIdx.x & 31;
    const int W = p.cam.image_width, H = p.cam.image_height;
    const int njobs = p.tiles_x * p.tiles_y * p.slices;
    unsigned long long my_rays = 0;

    for (;;) {
        int job = 0;
        if (lane == 0) job = int(atomicAdd(p.job_counter, 1u));
        job = __shfl_sync(0xffffffffu, job, 0);
        if (job >= njobs) break;
        const int slice = job % p.slices, tile = job / p.slices;
        const int i = (tile % p.tiles_x) * 8 + (lane & 7);
        const int j = (tile / p.tiles_x) * 4 + (lane >> 3);
        // samples [s, s_end) of this pixel
        int s = p.sample_begin + int((long long)p.sample_count * slice / p.slices);
        const int s_end = p.sample_begin + int((long long)p.sample_count * (slice + 1) / p.slices);
        const bool in_image = i < W && j < H;
        const uint32_t pixel = uint32_t(j) * uint32_t(W) + uint32_t(i);

        double acc_r = 0.0, acc_g = 0.0, acc_b = 0.0;
        Philox rng; rng.init(p.seed);
        Ray<T> ray; ray.o = mk<T>(0, 0, 0); ray.d = mk<T>(0, 0, 1); ray.tm = 0;
        Vec<T> thr = mk<T>(1, 1, 1);
        int depth = 0;                       // bounces left for the current path; 0 = need a new sample
        if (!in_image) s = s_end;

        while (true) {
            if (depth == 0) {
                if (s >= s_end) break;
                rng.set(pixel, uint32_t(s), 0u);
                ray = get_ray<T>(p.cam, i, j, rng);
                thr = mk<T>(1, 1, 1);
                depth = p.max_depth;
                ++s;
                if (depth <= 0) { depth = 0; continue; }    // src/camera_cpu.h:12-13: black, also for a negative max_depth
            }
            // one ray_color entry (depth > 0)
            ++my_rays;
            T t; int slot;
            trace_closest<T>(sv, st.nodes, st.sa, st.sb, st.meta, ray, T(0.001), T(INFINITY), t, slot);
            if (slot < 0) {                                           // miss: sky, path ends
                const Vec<T> c = thr * sky<T>(ray.d);
                acc_r += double(c.x); acc_g += double(c.y); acc_b += double(c.z);
                depth = 0;
                continue;
            }
            const Surface<T> sf = surface_at<T>(sv, st.sa, st.sb, st.sc, st.meta, ray, t, slot);
            rng.set(pixel, uint32_t(s - 1), uint32_t(p.max_depth - depth + 1));   // scatter at the b-th hit: stage b
            Vec<T> att, dnew;
            if (!scatter<T>(sf.type, sf.albedo, sf.param, ray.d, sf.normal, sf.front_face, rng, att, dnew)) {
                depth = 0;                                            // absorbed: contributes 0 (src/camera_cpu.h:20)
                continue;
            }
            thr = thr * att;
            ray.o = sf.p; ray.d = dnew;                               // time is inherited (src/material.h:22)
            --depth;                                                  // depth exhausted -> contributes 0 (:12-13)
        }

        if (in_image) {
            typedef typename Quad<T>::type Q;
            Q* dst = p.accum + (size_t(slice) * size_t(H) + size_t(j)) * size_t(W) + size_t(i);
            Q v;
            if (p.accumulate) { v = *dst; } else { v.x = 0; v.y = 0; v.z = 0; v.w = 0; }
            v.x += decltype(v.x)(acc_r); v.y += decltype(v.x)(acc_g); v.z += decltype(v.x)(acc_b);
            v.w += decltype(v.x)(s_end - (p.sample_begin + int((long long)p.sample_count * slice / p.slices)));
            *dst = v;
        }
    }
    // rays traced (every ray_color entry with depth > 0)
    for (int o = 16; o > 0; o >>= 1) my_rays += __shfl_xor_sync(0xffffffffu, my_rays, o);
    if (lane == 0 && my_rays) atomicAdd(p.ray_counter, my_rays);
}

// ------------------------------------------------------- warp-scheduled megakernel --
// Same work split as render_mega_kernel (warp = 8x4 pixel tile x one sample slice, lane = pixel,
// samples walked in order), but the bounce loop is cut into three SECTIONS and each warp runs, per
// iteration, the ONE section most of its lanes are waiting for:
//   N  visit one inner BVH node                       (lanes with cur >= 0)
//   L  test one leaf's spheres                        (lanes with a leaf in cur)
//   S  shade the finished segment / sky / start the next sample, then set up the next traversal
//      (lanes whose traversal ended: cur == RT_SENTINEL)
// A lane that is not in the chosen section idles for that iteration only.  In the lock-step kernel
// every lane waited for the slowest traversal of its warp and 7.4 of 32 lanes were active per
// issued instruction (profiles/r1_mega_v0_ncu_summary.txt); here N runs while most lanes are still
// traversing and S is postponed until `th_s` lanes wait for it (or nothing else is left), so the
// expensive section (Philox, normalisations, the double refinement) runs with a well-filled warp.
// The arithmetic per lane is unchanged, so parity mode stays bit-identical to the oracle.
template <typename T, bool STAGED, int BLOCK>
__global__ void __launch_bounds__(BLOCK, 2)
render_sched_kernel(const RenderParams<T> p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const Staged<T> st = stage_scene<T>(p.scene, smem, STAGED);
    const SceneView<T>& sv = p.scene;
    const unsigned FULL = 0xffffffffu;

    const int lane = threadIdx.x & 31;
    const int W = p.cam.image_width, H = p.cam.image_height;
    const int njobs = p.tiles_x * p.tiles_y * p.slices;
    const int th_s = p.th_s, th_l = p.th_l;
    unsigned long long my_rays = 0;
    StackEntry stack[RT_STACK];
    stack_init(stack);

    for (;;) {
        int job = 0;
        if (lane == 0) job = int(atomicAdd(p.job_counter, 1u));
        job = __shfl_sync(FULL, job, 0);
        if (job >= njobs) break;
        const int slice = job % p.slices, tile = job / p.slices;
        const int i = (tile % p.tiles_x) * 8 + (lane & 7);
        const int j = (tile / p.tiles_x) * 4 + (lane >> 3);
        const int s_begin = p.sample_begin + int((long long)p.sample_count * slice / p.slices);
        const int s_end = p.sample_begin + int((long long)p.sample_count * (slice + 1) / p.slices);
        int s = s_begin;
        const bool in_image = i < W && j < H;
        const uint32_t pixel = uint32_t(j) * uint32_t(W) + uint32_t(i);

        double acc_r = 0.0, acc_g = 0.0, acc_b = 0.0;
        Ray<T> ray; ray.o = mk<T>(0, 0, 0); ray.d = mk<T>(0, 0, 1); ray.tm = 0;
        Vec<T> thr = mk<T>(1, 1, 1);
        Trav<T> tv; tv.cur = RT_SENTINEL; tv.sp = 0; tv.tos = RT_SENTINEL; tv.tos_t = -INFINITY; tv.tbest = 0; tv.best_slot = -1; tv.qa = 1; slab_clear<T>(tv.slab);
        int depth = 0;                       // bounces left for the current path; 0 = start the next sample
        bool alive = in_image && s < s_end;

        for (;;) {
            const bool isN = alive && tv.cur >= 0;
            const bool isS = alive && tv.cur == RT_SENTINEL;
            const bool isL = alive && !isN && !isS;
            const int nN = __popc(__ballot_sync(FULL, isN));
            const int nL = __popc(__ballot_sync(FULL, isL));
            const int nS = __popc(__ballot_sync(FULL, isS));
            if ((nN | nL | nS) == 0) break;

            if (nS >= th_s || (nN | nL) == 0) {
                // ---------------------------------------------------------------- section S
                if (isS) {
                    bool need_ray = depth == 0;
                    if (!need_ray) {
                        // the segment traced last: src/camera_cpu.h:15-25
                        if (tv.best_slot < 0) {                                   // miss: sky, path ends
                            const Vec<T> c = thr * sky<T>(ray.d);
                            acc_r += double(c.x); acc_g += double(c.y); acc_b += double(c.z);
                            need_ray = true;
                        } else {
                            T t = tv.tbest;
                            const Surface<T> sf = surface_at<T>(sv, st.sa, st.sb, st.sc, st.meta, ray, t, tv.best_slot);
                            Philox rng; rng.init(p.seed);
                            rng.set(pixel, uint32_t(s - 1), uint32_t(p.max_depth - depth + 1));   // scatter at the b-th hit: stage b
                            Vec<T> att, dnew;
                            if (!scatter<T>(sf.type, sf.albedo, sf.param, ray.d, sf.normal, sf.front_face, rng, att, dnew)) {
                                need_ray = true;                                  // absorbed: contributes 0 (:20)
                            } else {
                                thr = thr * att;
                                ray.o = sf.p; ray.d = dnew;                       // time is inherited (src/material.h:22)
                                if (--depth == 0) need_ray = true;                // depth exhausted: contributes 0 (:12-13)
                            }
                        }
                    }
                    if (need_ray) {
                        depth = 0;
                        // camera::render's sample loop (src/camera.h:41-44); max_depth <= 0 -> every sample is black
                        if (p.max_depth <= 0) s = s_end;
                        if (s < s_end) {
                            Philox rng; rng.init(p.seed);
                            rng.set(pixel, uint32_t(s), 0u);
                            ray = get_ray<T>(p.cam, i, j, rng);
                            thr = mk<T>(1, 1, 1);
                            depth = p.max_depth;
                            ++s;
                        } else {
                            alive = false;
                        }
                    }
                    if (alive) {
                        ++my_rays;                                                // one ray_color entry with depth > 0
                        trav_init<T>(sv, ray, T(0.001), T(INFINITY), tv);
                    }
                }
            } else if (nL >= th_l || nL > nN) {
                // ---------------------------------------------------------------- section L
                if (isL) trav_leaf<T>(sv, st.sa, st.sb, st.meta, ray, T(0.001), tv, stack);
            } else {
                // ---------------------------------------------------------------- section N
                if (isN) trav_node<T>(st.nodes, ray, T(0.001), tv, stack);
            }
        }

        if (in_image) {
            typedef typename Quad<T>::type Q;
            Q* dst = p.accum + (size_t(slice) * size_t(H) + size_t(j)) * size_t(W) + size_t(i);
            Q v;
            if (p.accumulate) { v = *dst; } else { v.x = 0; v.y = 0; v.z = 0; v.w = 0; }
            v.x += decltype(v.x)(acc_r); v.y += decltype(v.x)(acc_g); v.z += decltype(v.x)(acc_b);
            v.w += decltype(v.x)(s_end - s_begin);
            *dst = v;
        }
    }
    for (int o = 16; o > 0; o >>= 1) my_rays += __shfl_xor_sync(FULL, my_rays, o);
    if (lane == 0 && my_rays) atomicAdd(p.ray_counter, my_rays);
}

// ------------------------------------------------------------------- parity-hook kernels --
template <typename T> __device__ __forceinline__ Ray<T> load_ray(const double* rays, size_t i) {
    const double* q = rays + 7 * i;
    Ray<T> r;
    r.o = mk<T>(T(q[0]), T(q[1]), T(q[2])); r.d = mk<T>(T(q[3]), T(q[4]), T(q[5])); r.tm = T(q[6]);
    return r;
}
template <typename T> __device__ __forceinline__ void store_ray(double* rays, size_t i, const Vec<T>& o, const Vec<T>& d, T tm) {
    double* q = rays + 7 * i;
    q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = d.x; q[4] = d.y; q[5] = d.z; q[6] = tm;
}

// rt_hit layout: p[3], normal[3], t (doubles), front_face, prim (int32) = 64 bytes
struct HitRec { double p[3], normal[3], t; int front_face, prim; };

template <typename T>
__global__ void trace_closest_kernel(const SceneView<T> sv, const double* __restrict__ rays, size_t n, double tmin, double tmax,
                                     int brute, HitRec* __restrict__ out) {
    const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const Ray<T> r = load_ray<T>(rays, idx);
    T t; int slot;
    if (brute) brute_closest<T>(sv, r, T(tmin), T(tmax), t, slot);
    else trace_closest<T>(sv, sv.nodes, sv.sa, sv.sb, sv.meta, r, T(tmin), T(tmax), t, slot);
    HitRec h;
    if (slot < 0) {
        h.p[0] = h.p[1] = h.p[2] = 0; h.normal[0] = h.normal[1] = h.normal[2] = 0; h.t = 0; h.front_face = 0; h.prim = -1;
    } else {
        const Surface<T> sf = surface_at<T>(sv, sv.sa, sv.sb, sv.sc, sv.meta, r, t, slot);
        h.p[0] = sf.p.x; h.p[1] = sf.p.y; h.p[2] = sf.p.z;
        h.normal[0] = sf.normal.x; h.normal[1] = sf.normal.y; h.normal[2] = sf.normal.z;
        h.t = t; h.front_face = sf.front_face; h.prim = sf.id;
    }
    out[idx] = h;
}

// materials: rt_material layout {int type, pad; double albedo[3], param} = 40 bytes
struct MatRec { int type, pad; double albedo[3], param; };

template <typename T>
__global__ void scatter_kernel(const MatRec* __restrict__ mats, const double* __restrict__ rays_in, const HitRec* __restrict__ hits,
                               size_t n, uint64_t seed, int* __restrict__ flag, double* __restrict__ att_out, double* __restrict__ rays_out) {
    const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const Ray<T> r = load_ray<T>(rays_in, idx);
    const MatRec m = mats[idx];
    const HitRec h = hits[idx];
    Philox rng; rng.init(seed); rng.set(uint32_t(idx), 0u, 1u);
    Vec<T> att = mk<T>(0, 0, 0), dnew = mk<T>(0, 0, 0);
    const bool ok = scatter<T>(m.type, mk<T>(T(m.albedo[0]), T(m.albedo[1]), T(m.albedo[2])), T(m.param), r.d,
                               mk<T>(T(h.normal[0]), T(h.normal[1]), T(h.normal[2])), h.front_face != 0, rng, att, dnew);
    flag[idx] = ok;
    att_out[3 * idx] = att.x; att_out[3 * idx + 1] = att.y; att_out[3 * idx + 2] = att.z;
    store_ray<T>(rays_out, idx, mk<T>(T(h.p[0]), T(h.p[1]), T(h.p[2])), dnew, r.tm);
}

template <typename T>
__global__ void gen_rays_kernel(const CameraFrame<T> cam, const int* __restrict__ px_i, const int* __restrict__ px_j,
                                const uint32_t* __restrict__ sample, size_t n, uint64_t seed, double* __restrict__ rays_out) {
    const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    Philox rng; rng.init(seed);
    rng.set(uint32_t(px_j[idx]) * uint32_t(cam.image_width) + uint32_t(px_i[idx]), sample[idx], 0u);
    const Ray<T> r = get_ray<T>(cam, px_i[idx], px_j[idx], rng);
    store_ray<T>(rays_out, idx, r.o, r.d, r.tm);
}

// BASELINE config 2 ray generator: even index = get_ray of a virtual 8192x8192 image,
// pixel = idx / 2, its own camera stream (pixel, sample 0, stage 0) -- the ray camera::get_ray would make;
// odd index = incoherent: origin uniform in [-12,12]x[0.05,3]x[-12,12], direction = uniform unit vector
// (ball rejection + normalise), time uniform, drawn in that order from the stream (idx / 2, sample 0, stage 1).
template <typename T>
__device__ __forceinline__ Ray<T> bench_ray(const CameraFrame<T>& cam, size_t idx, uint64_t seed) {
    Philox rng; rng.init(seed);
    const uint32_t pix = uint32_t(idx >> 1);
    if ((idx & 1) == 0) {
        rng.set(pix, 0u, 0u);
        return get_ray<T>(cam, int(pix % 8192u), int(pix / 8192u), rng);
    }
    rng.set(pix, 0u, 1u);
    Ray<T> r;
    const T ux = u01<T>(rng.next24()), uy = u01<T>(rng.next24()), uz = u01<T>(rng.next24());
    r.o = mk<T>(T(-12) + T(24) * ux, T(0.05) + T(2.95) * uy, T(-12) + T(24) * uz);
    r.d = unit_vector(random_in_unit_sphere<T, Philox>(rng));
    r.tm = u01<T>(rng.next24());
    return r;
}

template <typename T, bool STAGED>
__global__ void __launch_bounds__(256)
bench_closest_kernel(const SceneView<T> sv, const CameraFrame<T> cam, size_t n, uint64_t seed, int brute,
                     int* __restrict__ prim_out, float* __restrict__ t_out, double* __restrict__ rays_out, size_t rays_out_n) {
    extern __shared__ __align__(16) unsigned char smem[];
    const Staged<T> st = stage_scene<T>(sv, smem, STAGED);
    for (size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x; idx < n; idx += size_t(gridDim.x) * blockDim.x) {
        const Ray<T> r = bench_ray<T>(cam, idx, seed);
        T t; int slot;
        if (brute) brute_closest<T>(sv, r, T(0.001), T(INFINITY), t, slot);
        else trace_closest<T>(sv, st.nodes, st.sa, st.sb, st.meta, r, T(0.001), T(INFINITY), t, slot);
        int prim = -1;
        if (slot >= 0) prim = RT_META_ID(slot < sv.n_slots ? st.meta[slot] : sv.big_meta[slot - sv.n_slots]);
        prim_out[idx] = prim;
        t_out[idx] = slot >= 0 ? float(t) : 0.0f;
        if (idx < rays_out_n) store_ray<T>(rays_out, idx, r.o, r.d, r.tm);
    }
}

}  // namespace rt
