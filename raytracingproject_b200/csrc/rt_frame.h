// rt_frame.h -- launch interface of the framebuffer passes (rt_frame.cu).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>

namespace rt {

#define RT_MAX_PEERS 15

// other ranks' accumulation buffers mapped through CUDA IPC (read over NVLink in the resolve kernel)
struct PeerList {
    int n;
    int slices[RT_MAX_PEERS];
    const void* ptr[RT_MAX_PEERS];
};

struct EncodeHeader { int len; char text[48]; };     // "P3\n<W> <H>\n255\n" (src/camera.h:35)

template <typename Q>
cudaError_t launch_resolve(const Q* accum, int slices, const PeerList& peers, size_t n_pixels, double scale,
                           double* sum_out, unsigned char* rgb8, Q* reduced_out, cudaStream_t st);

int encode_blocks(size_t n_pixels);
cudaError_t launch_encode_p3(const unsigned char* rgb8, size_t n_pixels, const EncodeHeader& hdr, unsigned* block_sums,
                             unsigned long long* block_offsets, unsigned long long* total_len, char* text, cudaStream_t st);

}  // namespace rt
