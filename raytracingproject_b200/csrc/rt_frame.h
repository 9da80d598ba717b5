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

// optional extra views of the resolved frame, written by the same pass:
//   pfm    3 floats per pixel, the mean radiance (sum / spp, before gamma), rows bottom-to-top as a PFM file stores them
//   rgba8  the write_color bytes + alpha 255, 4 bytes per pixel: the headless stand-in for a presentable image
struct FrameViews { float* pfm; uchar4* rgba8; int width, height; int exact = 0; };     // exact: set by launch_resolve (A/B knob)

// tile-interleaved sharding: the resolve pass of a rank touches only the pixels of the tile groups it
// rendered -- (tile / group) % world == rank, tiles of tile_w x tile_h pixels, tiles_x per row of an image `width` wide.
// world <= 1: every pixel.
struct TileOwner { int world, rank, tiles_x, tile_w, tile_h, group, width; };

template <typename Q>
cudaError_t launch_resolve(const Q* accum, int slices, const PeerList& peers, size_t n_pixels, double scale,
                           double* sum_out, unsigned char* rgb8, Q* reduced_out, cudaStream_t st,
                           const FrameViews& views = FrameViews{nullptr, nullptr, 0, 0},
                           const TileOwner& owner = TileOwner{0, 0, 0, 0, 0, 0, 0});

// Stream-ordered hand-over of tile-sharded frames between GPUs: a peer that has written its tiles' bytes into the root's
// image (stores over NVLink inside its resolve kernel) bumps a counter in the root's memory (launch_frame_signal, after
// the resolve on the same stream); the root's stream waits until the counter reaches `target` (launch_frame_wait) before
// it encodes / copies the image.  *timed_out (device int, root) is set when the wait gives up after ~timeout_ms.
cudaError_t launch_frame_signal(unsigned* remote_counter, cudaStream_t st);
cudaError_t launch_frame_wait(const unsigned* counter, unsigned target, int timeout_ms, int* timed_out, cudaStream_t st);

// sums the sample layers of the accumulation buffer into layer 0 and zeroes the others (before a multi-GPU combine)
template <typename Q> cudaError_t launch_fold(Q* accum, int slices, size_t n_pixels, cudaStream_t st);

int encode_blocks(size_t n_pixels);
// One pass: accumulation layers (+ peers) -> rgb8 (optional) + P3 text.  tile_state: encode_blocks(n) + 1 64-bit words of
// scratch, *total_len = length of the text.  Same bytes as launch_resolve + launch_encode_p3.
template <typename Q>
cudaError_t launch_resolve_encode(const Q* accum, int slices, const PeerList& peers, size_t n_pixels, double scale, unsigned char* rgb8,
                                  char* text, const EncodeHeader& hdr, unsigned long long* tile_state, unsigned long long* total_len, cudaStream_t st);
// rgb8 must be readable up to 12 bytes past its last pixel (32-bit loads of 4-pixel groups)
// counted: block_sums already holds the character count of every 1024-pixel tile (launch_resolve_count_f32)
cudaError_t launch_encode_p3(const unsigned char* rgb8, size_t n_pixels, const EncodeHeader& hdr, unsigned* block_sums,
                             unsigned long long* block_offsets, unsigned long long* total_len, char* text, cudaStream_t st, bool counted = false);
// Float buffers with few layers and no peers: the bytes come from float sums + sqrt.approx, exact write_color only near a byte
// boundary (same bytes; rt_frame.cu).  launch_resolve takes that path by itself; this form also leaves the P3 character
// count of every tile in block_sums, so that resolve + launch_encode_p3(counted) is two passes, the second one over L2-resident bytes.
bool fast_f32_eligible(int slices, const PeerList& peers);
cudaError_t launch_resolve_count_f32(const float4* accum, int slices, size_t n_pixels, double scale, unsigned char* rgb8, unsigned* block_sums, cudaStream_t st);

}  // namespace rt
