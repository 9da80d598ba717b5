// rt_scene.cu -- the world on the device without a per-sphere host loop (C5: 10 k .. 1 M spheres).
// The hittable_list of shared_ptr<sphere> (src/hittable_list.h:13, src/sphere.h:60-66, src/material.h:27-28,43-45,73-74)
// arrives as two AoS arrays; one kernel makes the conservative float boxes the LBVH is built over
// (src/sphere.h:12-13, :22-25: box at time 0 united with the box at time 1), one gathers the spheres into the
// SoA slots in leaf order once the tree exists.  Built with --fmad=false: same values as the host path.
#include "rt_scene.h"

namespace rt {

namespace {

__global__ void __launch_bounds__(256) scene_boxes_kernel(const rt_sphere* __restrict__ spheres, const int* __restrict__ ids, int n,
                                                          HostBox* __restrict__ boxes) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const rt_sphere s = spheres[ids[k]];
    const double r = fabs(s.radius);
    HostBox b;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        double lo = s.center1[a] - r, hi = s.center1[a] + r;
        if (s.is_moving) { lo = fmin(lo, s.center1[a] + s.center_vec[a] - r); hi = fmax(hi, s.center1[a] + s.center_vec[a] + r); }
        const double pad = fmax(fmax(fabs(lo), fabs(hi)), r) * 9.5367431640625e-07 + 1e-30;      // ldexp(.., -20): exact
        b.lo[a] = __double2float_rd(lo - pad);                                                   // rounded outwards
        b.hi[a] = __double2float_ru(hi + pad);
    }
    boxes[k] = b;
}

__global__ void __launch_bounds__(256) scene_fill_kernel(const rt_sphere* __restrict__ spheres, const rt_material* __restrict__ materials,
                                                         const int* __restrict__ ids, const int* __restrict__ order, int n,
                                                         float4* __restrict__ a32, float4* __restrict__ b32, float4* __restrict__ c32,
                                                         double4* __restrict__ a64, double4* __restrict__ b64, double4* __restrict__ c64,
                                                         int* __restrict__ meta) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    const int id = ids[order[slot]];
    const rt_sphere s = spheres[id];
    const rt_material mt = materials[s.material];
    const bool mov = s.is_moving != 0;
    const double4 a = make_double4(s.center1[0], s.center1[1], s.center1[2], s.radius);
    const double4 b = make_double4(mov ? s.center_vec[0] : 0.0, mov ? s.center_vec[1] : 0.0, mov ? s.center_vec[2] : 0.0, s.radius * s.radius);
    double param = mt.param;
    if (mt.type == RT_METAL) param = param < 1 ? param : 1;                                     // src/material.h:33
    const double4 c = make_double4(mt.albedo[0], mt.albedo[1], mt.albedo[2], param);
    a64[slot] = a; b64[slot] = b; c64[slot] = c;
    const float rf = float(a.w);
    a32[slot] = make_float4(float(a.x), float(a.y), float(a.z), rf);
    b32[slot] = make_float4(float(b.x), float(b.y), float(b.z), rf * rf);
    c32[slot] = make_float4(float(c.x), float(c.y), float(c.z), float(c.w));
    meta[slot] = (id << 3) | (mov ? 4 : 0) | mt.type;
}

}  // namespace

cudaError_t launch_scene_boxes(const rt_sphere* spheres, const int* ids, int n, HostBox* boxes, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    scene_boxes_kernel<<<(n + 255) / 256, 256, 0, st>>>(spheres, ids, n, boxes);
    return cudaGetLastError();
}

cudaError_t launch_scene_fill(const rt_sphere* spheres, const rt_material* materials, const int* ids, const int* order, int n,
                              float4* a32, float4* b32, float4* c32, double4* a64, double4* b64, double4* c64, int* meta, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    scene_fill_kernel<<<(n + 255) / 256, 256, 0, st>>>(spheres, materials, ids, order, n, a32, b32, c32, a64, b64, c64, meta);
    return cudaGetLastError();
}

}  // namespace rt
