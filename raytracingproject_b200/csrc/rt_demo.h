// rt_demo.h -- launch interface of rt_demo.cu (the demo's compute shader as a CUDA kernel).
#pragma once

#include <cuda_runtime.h>

namespace rt {

struct DemoSphere { float cx, cy, cz, radius, r, g, b, a; };      // src/common_objects.h:17-22, 32 bytes

cudaError_t launch_demo_compute(const DemoSphere* spheres, int n, int width, int height, unsigned char* rgba8, cudaStream_t st);

}  // namespace rt
