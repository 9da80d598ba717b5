// rt_static.cu -- the P3 encoder for callers that must not allocate (code under CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc): rgb8 -> the text camera::render prints (src/camera.h:35, src/color.h:32-34),
// device buffers in, device buffers out, and the prefix-sum scratch in a file-scope __device__ array instead of
// cudaMalloc'ed memory.  No counterpart in the reference.
//
// The kernel is the encode half of resolve_encode_kernel (rt_frame.cu): one block per tile of ENC_TILE pixels, tiles
// numbered by a start ticket, text offsets by a single-pass prefix sum with decoupled look-back.  Its scratch -- one
// 64-bit state word per tile + the ticket -- is g_p3_scratch below: RT_STATIC_MAX_PIXELS / ENC_TILE + 1 words = 128 KB,
// held by every GPU that loads the library until the process exits.  Every call uses the SAME array: two calls that
// overlap on a GPU (different streams, or different host threads) race on it; calls on one stream are safe.  The state
// words and the ticket must start at zero, so every call clears them on its stream before the kernel.
#include "rt_static.h"
#include "rt_encode.cuh"

#include <cstdio>
#include <mutex>

namespace rt {

#define RT_STATIC_MAX_TILES (RT_STATIC_MAX_PIXELS / ENC_TILE)
static __device__ __align__(256) unsigned long long g_p3_scratch[RT_STATIC_MAX_TILES + 1];   // [tiles] look-back words, then the ticket

#define LB_AGG 1ull
#define LB_INC 2ull

// 4 pixels = 12 bytes from pixel p0 (p0 % 4 == 0).  Three 32-bit loads when the base is 4-byte aligned and the group is
// whole; byte loads otherwise -- unlike rt_frame.cu's load_pixels the caller's buffer is not padded, nothing past pixel
// n_pixels - 1 is touched.
__device__ __forceinline__ void load_pixels_exact(const unsigned char* __restrict__ rgb8, size_t p0, size_t n_pixels, bool words, unsigned char px[12]) {
    if (words && p0 + 4 <= n_pixels) {
        const unsigned* w = reinterpret_cast<const unsigned*>(rgb8 + 3 * p0);
        const unsigned a = w[0], b = w[1], c = w[2];
#pragma unroll
        for (int k = 0; k < 4; ++k) { px[k] = (unsigned char)(a >> (8 * k)); px[4 + k] = (unsigned char)(b >> (8 * k)); px[8 + k] = (unsigned char)(c >> (8 * k)); }
    } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) px[k] = p0 + k / 3 < n_pixels ? rgb8[3 * p0 + k] : (unsigned char)0;
    }
}

__global__ void __launch_bounds__(ENC_BLOCK) encode_p3_static_kernel(const unsigned char* __restrict__ rgb8, size_t n_pixels, bool words,
                                                                       char* __restrict__ text, EncodeHeader hdr,
                                                                       unsigned long long* __restrict__ tile_state, unsigned* __restrict__ ticket,
                                                                       unsigned long long* __restrict__ total_len, int n_tiles) {
    __shared__ unsigned warp_sums[ENC_BLOCK / 32];
    __shared__ __align__(16) char stage[ENC_TILE * 12 + 16];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_base;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const size_t p0 = size_t(tile) * ENC_TILE + size_t(threadIdx.x) * ENC_PPT;
    unsigned char px[12];
    unsigned len = 0;
    if (p0 < n_pixels) {
        load_pixels_exact(rgb8, p0, n_pixels, words, px);
#pragma unroll
        for (int k = 0; k < ENC_PPT; ++k) if (p0 + k < n_pixels) len += line_len(px + 3 * k);
    }
    unsigned total;
    const unsigned local = block_exclusive_scan(len, warp_sums, total);
    // where this tile's text starts: header + the totals of all tiles before it (a tile only waits for tiles that started earlier)
    if (threadIdx.x < 32) {
        const unsigned lane = threadIdx.x;
        volatile unsigned long long* st = tile_state;
        unsigned long long base = 0;
        if (tile == 0) {
            base = (unsigned long long)hdr.len;
            if (lane == 0) st[0] = (LB_INC << 62) | (base + total);
        } else {
            if (lane == 0) st[tile] = (LB_AGG << 62) | (unsigned long long)total;
            unsigned long long run = 0;
            int look = int(tile) - 1;                                  // the window [look - 31, look], lane 0 nearest
            for (;;) {
                const int idx = look - int(lane);
                unsigned long long v = LB_INC << 62;                   // before tile 0: an inclusive prefix of 0
                if (idx >= 0) { do { v = st[idx]; } while ((v >> 62) == 0ull); }
                const unsigned inc = __ballot_sync(0xffffffffu, (v >> 62) == LB_INC);
                const int first = inc ? __ffs(inc) - 1 : 32;           // aggregates nearer than the first inclusive prefix + that prefix
                unsigned long long mine = int(lane) <= first ? (v & ((1ull << 62) - 1)) : 0ull;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
                run += mine;
                if (inc) break;
                look -= 32;
            }
            base = run;
            if (lane == 0) { __threadfence(); st[tile] = (LB_INC << 62) | (base + total); }
        }
        if (lane == 0) {
            s_base = base;
            if (int(tile) == n_tiles - 1) *total_len = base + total;
        }
    }
    __syncthreads();
    char* dst = text + s_base;
    const unsigned mis = unsigned(reinterpret_cast<unsigned long long>(dst) & 15ull);
    unsigned off = mis + local;
    if (p0 < n_pixels) {
#pragma unroll
        for (int k = 0; k < ENC_PPT; ++k) {
            if (p0 + k >= n_pixels) break;
            off += put_number(stage + off, px[3 * k + 0]); stage[off++] = ' ';
            off += put_number(stage + off, px[3 * k + 1]); stage[off++] = ' ';
            off += put_number(stage + off, px[3 * k + 2]); stage[off++] = '\n';
        }
    }
    __syncthreads();
    // the stage has the destination's 16-byte misalignment: the body goes out as 128-bit stores, head and tail byte by byte
    const unsigned begin = mis, end = mis + total;
    const unsigned c0 = (begin + 15u) >> 4, c1 = end >> 4;
    char* dst0 = dst - mis;
    if (c0 < c1) {
        for (unsigned c = c0 + threadIdx.x; c < c1; c += ENC_BLOCK)
            reinterpret_cast<uint4*>(dst0)[c] = reinterpret_cast<const uint4*>(stage)[c];
        for (unsigned i = begin + threadIdx.x; i < c0 * 16u; i += ENC_BLOCK) dst0[i] = stage[i];
        for (unsigned i = c1 * 16u + threadIdx.x; i < end; i += ENC_BLOCK) dst0[i] = stage[i];
    } else {
        for (unsigned i = begin + threadIdx.x; i < end; i += ENC_BLOCK) dst0[i] = stage[i];
    }
    if (tile == 0 && threadIdx.x < hdr.len) text[threadIdx.x] = hdr.text[threadIdx.x];
}

// In host code `g_p3_scratch` is only a handle of the device array, not an address a kernel can use: its device address
// is asked for once per device, with that device current, and kept.
#define RT_STATIC_MAX_DEVICES 64
static std::mutex g_addr_mutex;
static unsigned long long* g_scratch_addr[RT_STATIC_MAX_DEVICES] = {};

static cudaError_t scratch_on_device(int device, unsigned long long** out) {
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (!g_scratch_addr[device]) {
        cudaError_t e = cudaSetDevice(device);
        if (e != cudaSuccess) return e;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_p3_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = static_cast<unsigned long long*>(p);
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

int encode_p3_static(int device, cudaStream_t st, const unsigned char* dev_rgb8, int width, int height, char* dev_ppm,
                     size_t ppm_capacity, unsigned long long* dev_ppm_len, char* err, size_t err_cap) {
    if (device < 0 || device >= RT_STATIC_MAX_DEVICES) { std::snprintf(err, err_cap, "rt_encode_p3_static: device %d out of range", device); return -1; }
    if (!dev_rgb8 || !dev_ppm || !dev_ppm_len || width < 1 || height < 1) { std::snprintf(err, err_cap, "rt_encode_p3_static: null buffer or empty image"); return -1; }
    const size_t npix = size_t(width) * size_t(height);
    if (npix > size_t(RT_STATIC_MAX_PIXELS)) {
        std::snprintf(err, err_cap, "rt_encode_p3_static: %d x %d = %zu pixels exceed the static scratch (RT_STATIC_MAX_PIXELS = %d); use rt_encode_p3",
                      width, height, npix, int(RT_STATIC_MAX_PIXELS));
        return -1;
    }
    EncodeHeader hdr;
    hdr.len = std::snprintf(hdr.text, sizeof hdr.text, "P3\n%d %d\n255\n", width, height);
    if (ppm_capacity < size_t(hdr.len) + 12 * npix) {
        std::snprintf(err, err_cap, "rt_encode_p3_static: ppm buffer of %zu bytes is below the worst case %zu (header + 12 per pixel)",
                      ppm_capacity, size_t(hdr.len) + 12 * npix);
        return -1;
    }
    unsigned long long* scratch = nullptr;
    cudaError_t e = scratch_on_device(device, &scratch);
    if (e == cudaSuccess) e = cudaSetDevice(device);
    const int nt = int((npix + ENC_TILE - 1) / ENC_TILE);
    // the last call left its state words and ticket behind: they must read zero when the first block starts
    if (e == cudaSuccess) e = cudaMemsetAsync(scratch, 0, (size_t(nt) + 1) * sizeof(unsigned long long), st);
    if (e == cudaSuccess) {
        const bool words = (reinterpret_cast<unsigned long long>(dev_rgb8) & 3ull) == 0ull;
        encode_p3_static_kernel<<<nt, ENC_BLOCK, 0, st>>>(dev_rgb8, npix, words, dev_ppm, hdr, scratch, reinterpret_cast<unsigned*>(scratch + nt),
                                                          dev_ppm_len, nt);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) { std::snprintf(err, err_cap, "rt_encode_p3_static: %s", cudaGetErrorString(e)); return -2; }
    return 0;
}

}  // namespace rt
