// rt_demo.cu -- the compute stage of the reference's windowed demo as a CUDA kernel.
//
// assets/shaders/ray_tracing.comp:35-73 shades every pixel of an rgba8 storage image: a pinhole ray
// through the pixel (camera at the origin looking down +z, :39-59) against a sphere by the sign of the
// discriminant alone (`hit`, :65-73: no root, no depth), flat sphere colour or the background
// (0.5, 0, 0.25, 1).  The shader hard-codes one sphere (:49-52); the storage buffer of
// `Sphere{vec3 center; float radius; vec4 color}` it was meant to read (src/common_objects.h:17-22,
// commented out at :28-31) is taken here as a list: the statement `if (hit) pixel_color = sphere.color`
// runs once per sphere in list order, so the LAST sphere that is hit gives the colour.
//
// The window, swapchain and vkCmdDispatch around it (src/vulkan/*) need a graphics API: unavailable on
// this box.  The image goes to a headless device buffer instead.  float arithmetic, no FMA contraction
// (GLSL does not contract across statements; the oracle restates it with -ffp-contract=off).
#include "rt_demo.h"

namespace rt {

__device__ __forceinline__ float dot3_rn(float ax, float ay, float az, float bx, float by, float bz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}

__device__ __forceinline__ unsigned unorm8(float v) {                 // imageStore to rgba8: clamp, scale, round to nearest even
    v = fminf(fmaxf(v, 0.0f), 1.0f);
    return (unsigned)__float2int_rn(__fmul_rn(v, 255.0f));
}

__global__ void __launch_bounds__(256) demo_compute_kernel(const DemoSphere* __restrict__ spheres, int n, int width, int height,
                                                           uchar4* __restrict__ image) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= width || y >= height) return;
    // :38-39 -- note the vertical coefficient is divided by the WIDTH as well
    const float h = __fdiv_rn(__fsub_rn(__fmul_rn(float(x), 2.0f), float(width)), float(width));
    const float v = __fdiv_rn(__fsub_rn(__fmul_rn(float(y), 2.0f), float(height)), float(width));
    // ray.direction = forwards + h * right + v * up with the axes of :44-47
    const float dx = __fadd_rn(__fadd_rn(0.0f, __fmul_rn(h, 1.0f)), __fmul_rn(v, 0.0f));
    const float dy = __fadd_rn(__fadd_rn(0.0f, __fmul_rn(h, 0.0f)), __fmul_rn(v, 1.0f));
    const float dz = __fadd_rn(__fadd_rn(1.0f, __fmul_rn(h, 0.0f)), __fmul_rn(v, 0.0f));
    float r = 0.5f, g = 0.0f, b = 0.25f, a = 1.0f;
    for (int k = 0; k < n; ++k) {
        const DemoSphere s = spheres[k];
        const float ocx = __fsub_rn(0.0f, s.cx), ocy = __fsub_rn(0.0f, s.cy), ocz = __fsub_rn(0.0f, s.cz);
        const float qa = dot3_rn(dx, dy, dz, dx, dy, dz);
        const float qb = __fmul_rn(2.0f, dot3_rn(dx, dy, dz, ocx, ocy, ocz));
        const float qc = __fsub_rn(dot3_rn(ocx, ocy, ocz, ocx, ocy, ocz), __fmul_rn(s.radius, s.radius));
        const float disc = __fsub_rn(__fmul_rn(qb, qb), __fmul_rn(__fmul_rn(4.0f, qa), qc));
        if (disc > 0.0f) { r = s.r; g = s.g; b = s.b; a = s.a; }
    }
    image[size_t(y) * width + x] = make_uchar4(unorm8(r), unorm8(g), unorm8(b), unorm8(a));
}

cudaError_t launch_demo_compute(const DemoSphere* spheres, int n, int width, int height, unsigned char* rgba8, cudaStream_t st) {
    if (width <= 0 || height <= 0) return cudaSuccess;
    const dim3 grid((width + 31) / 32, (height + 7) / 8);
    demo_compute_kernel<<<grid, 256, 0, st>>>(spheres, n, width, height, reinterpret_cast<uchar4*>(rgba8));
    return cudaGetLastError();
}

}  // namespace rt
