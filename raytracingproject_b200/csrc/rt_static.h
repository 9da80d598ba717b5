// rt_static.h -- launch interface of the allocation-free P3 encoder (rt_static.cu).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>

#include "rt_frame.h"

namespace rt {

// largest image the static scratch of rt_static.cu serves: 16 Mi pixels (a 4K frame is 8.3 M) = 16384 tiles of 1024
// pixels, one 64-bit word each + the ticket = 128 KB of device memory per GPU for the life of the process
#define RT_STATIC_MAX_PIXELS (1 << 24)

// rgb8 (device, 3 bytes per pixel) -> P3 text (device) + its length (device), asynchronous on `st`.  No allocation.
// Returns 0, -1 (bad argument; message in err) or -2 (CUDA error).
int encode_p3_static(int device, cudaStream_t st, const unsigned char* dev_rgb8, int width, int height, char* dev_ppm,
                     size_t ppm_capacity, unsigned long long* dev_ppm_len, char* err, size_t err_cap);

}  // namespace rt
