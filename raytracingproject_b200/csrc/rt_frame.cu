// rt_frame.cu -- framebuffer passes (the only HBM-bound part of the path):
//   resolve : sum of sample slices (+ peer partial buffers over NVLink) -> write_color bytes
//   encode  : bytes -> the P3 text camera::render prints, byte-identical
// Restates src/color.h:9-12,14-35 (1/spp, sqrt gamma, clamp [0,0.999], int(256*x), "r g b\n")
// and the header of src/camera.h:35.  Arithmetic is double as in the reference.
#include "rt_frame.h"
#include "rt_encode.cuh"

#include <cstdlib>

namespace rt {

__device__ __forceinline__ unsigned char write_color_byte(double sum, double scale) {
    double v = sum * scale;              // src/color.h:20-23
    v = sqrt(v);                         // :26-28 (linear_to_gamma)
    if (v < 0.000) v = 0.000;            // :31 interval(0.000, 0.999).clamp, src/interval.h:14-18
    if (v > 0.999) v = 0.999;
    return (unsigned char)(int)(256 * v);
}

// The same byte without the double-precision square root on the common path.  The passes are HBM-bound on paper, but at
// one sample layer the three IEEE double square roots per pixel (a ~30-instruction DFMA sequence each) cost more than
// the memory traffic (measured: 54 of 78 us at 4K).  Here the root is taken in float by the special-function unit
// (sqrt.approx: 2 ulp) and trusted only where it cannot decide wrongly: y = 256 * sqrt(float(v)) carries a relative error
// below 4e-7 (one rounding of v, the approximate root), i.e. less than 1e-4 absolute, so unless y lies within 5e-4 of an
// integer its integer part is the reference's; the few channels that do (0.1 %, plus NaN / negative input and exact zeros)
// take write_color_byte.  The clamp to 0.999 only matters through its byte, 255 = int(255.744).
__device__ __forceinline__ float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }   // ftz: a denormal argument is a byte 0 either way
__device__ __forceinline__ unsigned char write_color_byte_fast(double sum, double scale, bool exact) {
    const float y = 256.0f * sqrt_approx(float(sum * scale));
    if (exact || !(fabsf(y - rintf(y)) >= 5e-4f)) return write_color_byte(sum, scale);      // near a byte boundary, or NaN
    return (unsigned char)min(int(y), 255);
}
// RT_EXACT_WRITE_COLOR=1 in the environment: every channel through the double path (same-box A/B of the above)
static bool frame_exact() { const char* e = std::getenv("RT_EXACT_WRITE_COLOR"); static const bool v = e && *e && *e != '0'; return v; }

// One thread = 4 consecutive pixels: 4 x Quad loads per slice/peer (64 or 128 contiguous bytes),
// 3 x 32-bit stores of packed rgb.  Slices and peers are summed in a fixed order, so the image is
// reproducible.  peers[] are device pointers of other ranks' buffers opened through CUDA IPC: the
// loads cross NVLink inside this kernel (fused reduce + resolve).
__device__ __forceinline__ bool owns_pixel(const TileOwner& o, size_t pixel) {
    const int y = int(pixel / size_t(o.width)), x = int(pixel - size_t(y) * size_t(o.width));
    const int tile = (y / o.tile_h) * o.tiles_x + x / o.tile_w;
    return (tile / o.group) % o.world == o.rank;
}

template <typename Q>
__global__ void resolve_kernel(const Q* __restrict__ accum, int slices, PeerList peers, size_t n_pixels, double scale,
                               double* __restrict__ sum_out, unsigned char* __restrict__ rgb8, Q* __restrict__ reduced_out, FrameViews views,
                               TileOwner owner) {
    const size_t g = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (g >= n_pixels) return;
    // tile-sharded frame: only the pixels this rank rendered (the image the bytes go to may be another GPU's, over NVLink)
    bool mine[4] = { true, true, true, true };
    bool all_mine = true;
    if (owner.world > 1) {
        bool any = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) { mine[k] = g + k < n_pixels && owns_pixel(owner, g + k); any = any || mine[k]; all_mine = all_mine && mine[k]; }
        if (!any) return;
    }
    const int np = int(min(size_t(4), n_pixels - g));
    unsigned char b[12];
    double s[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k][0] = s[k][1] = s[k][2] = s[k][3] = 0.0;
    for (int sl = 0; sl < slices; ++sl) {
        const Q* src = accum + size_t(sl) * n_pixels + g;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (k < np) { const Q v = src[k]; s[k][0] += double(v.x); s[k][1] += double(v.y); s[k][2] += double(v.z); s[k][3] += double(v.w); }
    }
    for (int p = 0; p < peers.n; ++p) {
        const Q* base = reinterpret_cast<const Q*>(peers.ptr[p]);
        for (int sl = 0; sl < peers.slices[p]; ++sl) {
            const Q* src = base + size_t(sl) * n_pixels + g;
#pragma unroll
            for (int k = 0; k < 4; ++k) if (k < np) { const Q v = src[k]; s[k][0] += double(v.x); s[k][1] += double(v.y); s[k][2] += double(v.z); s[k][3] += double(v.w); }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (k < np && mine[k]) {
            const double sc = scale > 0.0 ? scale : 1.0 / s[k][3];       // spp_total = 0: samples actually accumulated
            b[3 * k + 0] = write_color_byte_fast(s[k][0], sc, views.exact != 0);
            b[3 * k + 1] = write_color_byte_fast(s[k][1], sc, views.exact != 0);
            b[3 * k + 2] = write_color_byte_fast(s[k][2], sc, views.exact != 0);
            if (sum_out) { sum_out[3 * (g + k) + 0] = s[k][0]; sum_out[3 * (g + k) + 1] = s[k][1]; sum_out[3 * (g + k) + 2] = s[k][2]; }
            if (views.pfm) {
                const size_t y = (g + k) / size_t(views.width), x = (g + k) - y * size_t(views.width);
                float* o = views.pfm + 3 * ((size_t(views.height) - 1 - y) * size_t(views.width) + x);
                o[0] = float(s[k][0] * sc); o[1] = float(s[k][1] * sc); o[2] = float(s[k][2] * sc);
            }
            if (reduced_out) { Q v; v.x = decltype(v.x)(s[k][0]); v.y = decltype(v.x)(s[k][1]); v.z = decltype(v.x)(s[k][2]); v.w = decltype(v.x)(s[k][3]); reduced_out[g + k] = v; }
        } else {
            b[3 * k] = b[3 * k + 1] = b[3 * k + 2] = 0;
        }
    }
    if (views.rgba8)
        for (int k = 0; k < np; ++k) if (mine[k]) views.rgba8[g + k] = make_uchar4(b[3 * k], b[3 * k + 1], b[3 * k + 2], 255);
    if (rgb8) {
        if (!all_mine) {                        // a 4-pixel group that straddles two owners (image width not a multiple of 4)
            for (int k = 0; k < np; ++k) if (mine[k]) { rgb8[3 * (g + k)] = b[3 * k]; rgb8[3 * (g + k) + 1] = b[3 * k + 1]; rgb8[3 * (g + k) + 2] = b[3 * k + 2]; }
        } else if (np == 4) {
            unsigned int* dst = reinterpret_cast<unsigned int*>(rgb8 + 3 * g);      // g % 4 == 0 -> 12-byte groups, 4-aligned
            dst[0] = b[0] | (b[1] << 8) | (b[2] << 16) | (unsigned(b[3]) << 24);
            dst[1] = b[4] | (b[5] << 8) | (b[6] << 16) | (unsigned(b[7]) << 24);
            dst[2] = b[8] | (b[9] << 8) | (b[10] << 16) | (unsigned(b[11]) << 24);
        } else {
            for (int k = 0; k < 3 * np; ++k) rgb8[3 * g + k] = b[k];
        }
    }
}

// ------------------------------------------------- float fast paths (few layers, no peers) --
// A float buffer with at most RT_FAST_LAYERS layers (every frame after rt_fold_slices, every short frame): the layers are
// summed in FLOAT and the byte comes from y = 256 * sqrt.approx(sum * scale): every step carries a relative error below
// 1e-7, y an absolute one below 1.5e-4, so unless y lies within RT_FAST_MARGIN of an integer its integer part is the
// reference's byte.  The channels that do (0.05 %), NaN and negative input take the reference arithmetic in double on the
// exact sum (write_color_byte, one copy of the code, driven by a bit mask).  No double adds, no conversions: the general
// kernels spend more issue slots on those than on memory (ncu: 42 % issue utilisation at 1.05 TB/s on a 4K frame).
#define RT_FAST_LAYERS 8
#define RT_FAST_MARGIN 2.5e-4f

// bytes of pixels p0 .. p0+3 packed little-endian into w[0..2] (= the 12 bytes of rgb8 + 3 * p0); np = how many exist
__device__ __forceinline__ void pixels4_bytes_f32(const float4* __restrict__ accum, int slices, size_t n_pixels, size_t p0, int np,
                                                  double scale, unsigned w[3]) {
    float s[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k][0] = s[k][1] = s[k][2] = s[k][3] = 0.0f;
    for (int sl = 0; sl < slices; ++sl) {
        const float4* src = accum + size_t(sl) * n_pixels + p0;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (k < np) { const float4 v = src[k]; s[k][0] += v.x; s[k][1] += v.y; s[k][2] += v.z; s[k][3] += v.w; }
    }
    unsigned need = 0u;
    w[0] = w[1] = w[2] = 0u;
    // y = 256 * sqrt(sum * scale) = sqrt(sum * (65536 * scale)), clamped to [0.25, 255.5]: below 0.5 the byte is 0 and from
    // 255 on it is 255 whatever the rounding (0.999 * 256 = 255.74, src/color.h:31), and neither clamp value is near an integer;
    // a NaN or negative sum becomes 0.25 -> byte 0, which is what write_color_byte makes of it
    const bool fixed = scale > 0.0;
    const float sc_fixed = float(scale * 65536.0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (k < np) {
            float sc = sc_fixed;
            if (!fixed) sc = 65536.0f / s[k][3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float y = fminf(fmaxf(sqrt_approx(s[k][c] * sc), 0.25f), 255.5f);
                const float r = (y + 12582912.0f) - 12582912.0f;               // nearest integer for 0 <= y < 2^22
                const int at = 3 * k + c;
                w[at >> 2] |= unsigned(int(y)) << (8 * (at & 3));
                if (!(fabsf(y - r) >= RT_FAST_MARGIN)) need |= 1u << at;
            }
        }
    }
    while (need) {                                                              // the reference's arithmetic on the exact sum
        const int at = __ffs(need) - 1;
        need &= need - 1u;
        const int k = at / 3, c = at - 3 * k;
        double sum = 0.0, cnt = 0.0;
        for (int sl = 0; sl < slices; ++sl) {
            const float4 v = accum[size_t(sl) * n_pixels + p0 + k];
            sum += double(c == 0 ? v.x : (c == 1 ? v.y : v.z)); cnt += double(v.w);
        }
        const unsigned b = write_color_byte(sum, scale > 0.0 ? scale : 1.0 / cnt);
        const unsigned sh = 8u * (at & 3), m = ~(0xffu << sh);
        if (at < 4) w[0] = (w[0] & m) | (b << sh); else if (at < 8) w[1] = (w[1] & m) | (b << sh); else w[2] = (w[2] & m) | (b << sh);
    }
}

__device__ __forceinline__ void store_bytes12(unsigned char* __restrict__ rgb8, size_t p0, int np, const unsigned w[3]) {
    if (np == 4) {
        unsigned* dst = reinterpret_cast<unsigned*>(rgb8 + 3 * p0);             // p0 % 4 == 0: 12-byte groups, 4-aligned
        dst[0] = w[0]; dst[1] = w[1]; dst[2] = w[2];
    } else {
        for (int i = 0; i < 3 * np; ++i) rgb8[3 * p0 + i] = (unsigned char)(w[i >> 2] >> (8 * (i & 3)));
    }
}

// bytes (+ optionally the number of P3 characters of every block's 1024 pixels = one tile of the encoder, so that the encoder
// needs no counting pass): 16 B per layer read, 3 B written per pixel, one thread = 4 pixels
__global__ void __launch_bounds__(ENC_BLOCK) resolve_bytes_f32_kernel(const float4* __restrict__ accum, int slices, size_t n_pixels, double scale,
                                                                      unsigned char* __restrict__ rgb8, unsigned* __restrict__ block_sums) {
    __shared__ unsigned warp_sums[ENC_BLOCK / 32];
    const size_t p0 = (size_t(blockIdx.x) * ENC_BLOCK + threadIdx.x) * 4;
    const int np = p0 < n_pixels ? int(min(size_t(4), n_pixels - p0)) : 0;
    unsigned len = 0;
    if (np > 0) {
        unsigned w[3];
        pixels4_bytes_f32(accum, slices, n_pixels, p0, np, scale, w);
        store_bytes12(rgb8, p0, np, w);
        if (block_sums) {
#pragma unroll
            for (int i = 0; i < 12; ++i) if (i < 3 * np) { const unsigned v = (w[i >> 2] >> (8 * (i & 3))) & 0xffu; len += 1u + (v >= 10u) + (v >= 100u); }
            len += 3u * unsigned(np);                                           // two blanks and the newline of every line
        }
    }
    if (block_sums) {                                                           // block-uniform
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) len += __shfl_xor_sync(0xffffffffu, len, o);
        if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = len;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned t = 0;
#pragma unroll
            for (int q = 0; q < ENC_BLOCK / 32; ++q) t += warp_sums[q];
            block_sums[blockIdx.x] = t;
        }
    }
}

// digits of a byte for the text: entry v = its decimal characters in the low bytes (first character lowest) | count << 24
__device__ __forceinline__ unsigned digits_entry(unsigned v) {
    const unsigned h = v / 100u, t = (v / 10u) % 10u, o = v % 10u;
    if (v >= 100u) return ('0' + h) | (('0' + t) << 8) | (('0' + o) << 16) | (3u << 24);
    if (v >= 10u) return ('0' + t) | (('0' + o) << 8) | (2u << 24);
    return ('0' + o) | (1u << 24);
}
// one number and the separator that follows it: at most four byte stores, no division, no branch on the digit count
__device__ __forceinline__ unsigned put_entry(char* dst, unsigned e, unsigned sep) {
    const unsigned n = e >> 24;
    const unsigned word = (e & 0xffffffu) | (sep << (8u * n));
    dst[0] = char(word);
    dst[1] = char(word >> 8);
    if (n >= 2u) dst[2] = char(word >> 16);
    if (n >= 3u) dst[3] = char(word >> 24);
    return n + 1u;
}

template <typename Q>
cudaError_t launch_resolve(const Q* accum, int slices, const PeerList& peers, size_t n_pixels, double scale,
                           double* sum_out, unsigned char* rgb8, Q* reduced_out, cudaStream_t st, const FrameViews& views, const TileOwner& owner) {
    if (n_pixels == 0) return cudaSuccess;
    const size_t threads = (n_pixels + 3) / 4;
    FrameViews v = views;
    v.exact = frame_exact();
    if (sizeof(Q) == sizeof(float4) && !v.exact && rgb8 && !sum_out && !reduced_out && !v.pfm && !v.rgba8 && owner.world <= 1 && peers.n == 0 &&
        slices <= RT_FAST_LAYERS) {
        resolve_bytes_f32_kernel<<<unsigned((threads + ENC_BLOCK - 1) / ENC_BLOCK), ENC_BLOCK, 0, st>>>(reinterpret_cast<const float4*>(accum), slices, n_pixels, scale, rgb8, nullptr);
        return cudaGetLastError();
    }
    resolve_kernel<Q><<<unsigned((threads + 255) / 256), 256, 0, st>>>(accum, slices, peers, n_pixels, scale, sum_out, rgb8, reduced_out, v, owner);
    return cudaGetLastError();
}
template cudaError_t launch_resolve<float4>(const float4*, int, const PeerList&, size_t, double, double*, unsigned char*, float4*, cudaStream_t, const FrameViews&, const TileOwner&);
template cudaError_t launch_resolve<double4>(const double4*, int, const PeerList&, size_t, double, double*, unsigned char*, double4*, cudaStream_t, const FrameViews&, const TileOwner&);

// ---------------------------------------------------------- tile-sharded frame hand-over --
// (see rt_frame.h)  The signal runs AFTER the peer's resolve kernel on the same stream, so that kernel's stores to the
// root's image have been performed; the fence + system-scope atomic publish them to the root GPU.
__global__ void frame_signal_kernel(unsigned* remote_counter) {
    __threadfence_system();
    atomicAdd_system(remote_counter, 1u);
}
__global__ void frame_wait_kernel(const unsigned* counter, unsigned target, long long timeout_ns, int* timed_out) {
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        const unsigned v = *reinterpret_cast<const volatile unsigned*>(counter);
        if (int(v - target) >= 0) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if ((long long)(t - t0) > timeout_ns) { *timed_out = 1; break; }
        __nanosleep(200);
    }
    __threadfence_system();
}
cudaError_t launch_frame_signal(unsigned* remote_counter, cudaStream_t st) {
    frame_signal_kernel<<<1, 1, 0, st>>>(remote_counter);
    return cudaGetLastError();
}
cudaError_t launch_frame_wait(const unsigned* counter, unsigned target, int timeout_ms, int* timed_out, cudaStream_t st) {
    frame_wait_kernel<<<1, 1, 0, st>>>(counter, target, (long long)timeout_ms * 1000000ll, timed_out);
    return cudaGetLastError();
}

// -------------------------------------------------------------------------- slice fold --
// The wavefront kernel renders a pixel's samples in `slices` groups, each into its own [H][W] layer
// (different warps, no atomics).  Before a multi-GPU combine the layers are summed into layer 0 and
// the others are zeroed, so that ONE layer per rank crosses NVLink (33 MB instead of 133 MB at 1080p
// with 4 slices); the buffer keeps its meaning (resolve sums all layers, accumulate passes add to any).
// One thread = one pixel of every layer: 16-byte (float4) or 32-byte (double4) coalesced accesses,
// layers summed in double in layer order, the order resolve_kernel uses.
template <typename Q>
__global__ void fold_kernel(Q* __restrict__ accum, int slices, size_t n_pixels) {
    const size_t g = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (g >= n_pixels) return;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    Q zero; zero.x = 0; zero.y = 0; zero.z = 0; zero.w = 0;
    for (int sl = 0; sl < slices; ++sl) {
        Q* src = accum + size_t(sl) * n_pixels + g;
        const Q v = *src;
        s0 += double(v.x); s1 += double(v.y); s2 += double(v.z); s3 += double(v.w);
        if (sl > 0) *src = zero;
    }
    Q v; v.x = decltype(v.x)(s0); v.y = decltype(v.x)(s1); v.z = decltype(v.x)(s2); v.w = decltype(v.x)(s3);
    accum[g] = v;
}

template <typename Q>
cudaError_t launch_fold(Q* accum, int slices, size_t n_pixels, cudaStream_t st) {
    if (n_pixels == 0 || slices <= 1) return cudaSuccess;
    fold_kernel<Q><<<unsigned((n_pixels + 255) / 256), 256, 0, st>>>(accum, slices, n_pixels);
    return cudaGetLastError();
}
template cudaError_t launch_fold<float4>(float4*, int, size_t, cudaStream_t);
template cudaError_t launch_fold<double4>(double4*, int, size_t, cudaStream_t);

// ------------------------------------------------------------------------- P3 encoder --
// "r g b\n" with variable-width decimals (src/color.h:32-34) needs a prefix sum of line lengths:
// pass 1 per-block totals, pass 2 scan of the totals (one block), pass 3 block-local scan,
// text staged in shared memory and copied out with consecutive threads writing consecutive bytes.
// the ENC_PPT = 4 pixels of a thread are 12 consecutive bytes at a 4-byte aligned address: three 32-bit loads
// (a warp reads 384 contiguous bytes) instead of twelve byte loads.  The caller's rgb8 buffer is padded by
// >= 12 bytes, so the last thread may read past pixel n_pixels - 1; it never uses those bytes.
__device__ __forceinline__ void load_pixels(const unsigned char* __restrict__ rgb8, size_t p0, unsigned char px[12]) {
    const unsigned* w = reinterpret_cast<const unsigned*>(rgb8 + 3 * p0);
    const unsigned a = w[0], b = w[1], c = w[2];
#pragma unroll
    for (int k = 0; k < 4; ++k) { px[k] = (unsigned char)(a >> (8 * k)); px[4 + k] = (unsigned char)(b >> (8 * k)); px[8 + k] = (unsigned char)(c >> (8 * k)); }
}

__global__ void __launch_bounds__(ENC_BLOCK) encode_count_kernel(const unsigned char* __restrict__ rgb8, size_t n_pixels, unsigned* __restrict__ block_sums) {
    __shared__ unsigned warp_sums[ENC_BLOCK / 32];
    const size_t p0 = size_t(blockIdx.x) * ENC_TILE + size_t(threadIdx.x) * ENC_PPT;
    unsigned len = 0;
    if (p0 < n_pixels) {
        unsigned char px[12];
        load_pixels(rgb8, p0, px);
#pragma unroll
        for (int k = 0; k < ENC_PPT; ++k) if (p0 + k < n_pixels) len += line_len(px + 3 * k);
    }
    unsigned total;
    block_exclusive_scan(len, warp_sums, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// exclusive scan of block_sums (64-bit offsets out), single block: every thread takes one contiguous chunk, all its loads in
// flight at once (the first version walked the array 256 entries at a time, one dependent load + three barriers per step:
// 27 us at 4K, as long as the pass that writes 95 MB of text), then one scan over the 256 chunk totals
#define ENC_SCAN_BLOCK 1024             // 8100 tiles of a 4K frame: 8 entries per thread (256 threads took 18 us, all latency)
__global__ void __launch_bounds__(ENC_SCAN_BLOCK) encode_scan_kernel(const unsigned* __restrict__ block_sums, int n_blocks,
                                                                  unsigned long long* __restrict__ block_offsets, unsigned long long header_len,
                                                                  unsigned long long* __restrict__ total_len) {
    __shared__ unsigned long long warp_tot[ENC_SCAN_BLOCK / 32];
    const int per = (n_blocks + ENC_SCAN_BLOCK - 1) / ENC_SCAN_BLOCK;
    const int lo = min(n_blocks, int(threadIdx.x) * per), hi = min(n_blocks, lo + per);
    unsigned long long mine = 0;
    for (int i = lo; i < hi; ++i) mine += block_sums[i];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_tot[wid] = x;
    __syncthreads();
    unsigned long long base = header_len;
    for (int w = 0; w < wid; ++w) base += warp_tot[w];
    unsigned long long run = base + x - mine;                       // exclusive prefix of this thread's chunk
    for (int i = lo; i < hi; ++i) { block_offsets[i] = run; run += block_sums[i]; }
    if (threadIdx.x == ENC_SCAN_BLOCK - 1) *total_len = base + x;
}

// The block's text is staged in shared memory at the SAME 16-byte misalignment its destination has, so that
// every aligned 16-byte chunk of the destination is an aligned 16-byte chunk of the stage: the body goes out
// as 128-bit stores (512 contiguous bytes per warp instruction), only the ragged head and tail byte by byte.
__global__ void __launch_bounds__(ENC_BLOCK) encode_write_kernel(const unsigned char* __restrict__ rgb8, size_t n_pixels,
                                                                   const unsigned long long* __restrict__ block_offsets,
                                                                   char* __restrict__ text, EncodeHeader hdr) {
    __shared__ unsigned warp_sums[ENC_BLOCK / 32];
    __shared__ __align__(16) char stage[ENC_TILE * 12 + 16];
    __shared__ unsigned s_lut[256];
    s_lut[threadIdx.x] = digits_entry(threadIdx.x);                             // ENC_BLOCK == 256
    const size_t p0 = size_t(blockIdx.x) * ENC_TILE + size_t(threadIdx.x) * ENC_PPT;
    const int np = p0 < n_pixels ? int(min(size_t(ENC_PPT), n_pixels - p0)) : 0;
    unsigned char px[12];
    if (np > 0) load_pixels(rgb8, p0, px);
    __syncthreads();
    unsigned e[12];                                                             // characters | count << 24 of every number
    unsigned len = 3u * unsigned(np);                                           // two blanks and the newline of every line
#pragma unroll
    for (int i = 0; i < 12; ++i) { e[i] = i < 3 * np ? s_lut[px[i]] : 0u; len += e[i] >> 24; }
    unsigned total;
    char* dst = text + block_offsets[blockIdx.x];
    const unsigned mis = unsigned(reinterpret_cast<unsigned long long>(dst) & 15ull);
    unsigned off = mis + block_exclusive_scan(len, warp_sums, total);
#pragma unroll
    for (int i = 0; i < 12; ++i)
        if (i < 3 * np) off += put_entry(stage + off, e[i], (i % 3) == 2 ? unsigned('\n') : unsigned(' '));
    __syncthreads();
    const unsigned begin = mis, end = mis + total;
    const unsigned c0 = (begin + 15u) >> 4, c1 = end >> 4;              // full 16-byte chunks [c0, c1)
    char* dst0 = dst - mis;                                             // 16-byte aligned
    if (c0 < c1) {
        for (unsigned c = c0 + threadIdx.x; c < c1; c += ENC_BLOCK)
            reinterpret_cast<uint4*>(dst0)[c] = reinterpret_cast<const uint4*>(stage)[c];
        for (unsigned i = begin + threadIdx.x; i < c0 * 16u; i += ENC_BLOCK) dst0[i] = stage[i];
        for (unsigned i = c1 * 16u + threadIdx.x; i < end; i += ENC_BLOCK) dst0[i] = stage[i];
    } else {
        for (unsigned i = begin + threadIdx.x; i < end; i += ENC_BLOCK) dst0[i] = stage[i];
    }
    if (blockIdx.x == 0 && threadIdx.x < hdr.len) text[threadIdx.x] = hdr.text[threadIdx.x];   // "P3\nW H\n255\n"
}

int encode_blocks(size_t n_pixels) { return int((n_pixels + ENC_TILE - 1) / ENC_TILE); }

// ----------------------------------------------------------- fused resolve + P3 encode --
// ONE pass from the accumulation buffer to the image AND its P3 text: a block resolves a tile of ENC_TILE pixels
// (write_color, as resolve_kernel), keeps the bytes in registers, counts the characters of its lines, finds where its
// text starts with a single-pass prefix sum over the tiles (decoupled look-back: a tile publishes its own total at once
// and adds up its predecessors' totals until it meets one whose inclusive prefix is known), and writes the lines.
// Traffic = the accumulation layers read once + 3 B of rgb8 + <= 12 B of text written per pixel; the three-kernel
// encoder (count / scan / write) read the bytes twice more and paid three launches on a 30 us job.
// Tiles are numbered by a ticket in the order the blocks START, so a tile only ever waits for tiles that are already running.
#define LB_AGG 1ull
#define LB_INC 2ull
template <typename Q>
__global__ void __launch_bounds__(ENC_BLOCK) resolve_encode_kernel(const Q* __restrict__ accum, int slices, PeerList peers, size_t n_pixels, double scale,
                                                                   unsigned char* __restrict__ rgb8, char* __restrict__ text, EncodeHeader hdr,
                                                                   unsigned* __restrict__ ticket, unsigned long long* __restrict__ tile_state,
                                                                   unsigned long long* __restrict__ total_len, int n_tiles, bool exact) {
    __shared__ unsigned warp_sums[ENC_BLOCK / 32];
    __shared__ __align__(16) char stage[ENC_TILE * 12 + 16];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_part[ENC_BLOCK / 32];
    __shared__ int s_inc[ENC_BLOCK / 32];
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const size_t p0 = size_t(tile) * ENC_TILE + size_t(threadIdx.x) * ENC_PPT;
    unsigned char px[12];
    unsigned len = 0;
    if (p0 < n_pixels) {
        const int np = int(min(size_t(4), n_pixels - p0));
        double s[4][4];
#pragma unroll
        for (int k = 0; k < 4; ++k) s[k][0] = s[k][1] = s[k][2] = s[k][3] = 0.0;
        for (int sl = 0; sl < slices; ++sl) {
            const Q* src = accum + size_t(sl) * n_pixels + p0;
#pragma unroll
            for (int k = 0; k < 4; ++k) if (k < np) { const Q v = src[k]; s[k][0] += double(v.x); s[k][1] += double(v.y); s[k][2] += double(v.z); s[k][3] += double(v.w); }
        }
        for (int p = 0; p < peers.n; ++p) {
            const Q* base = reinterpret_cast<const Q*>(peers.ptr[p]);
            for (int sl = 0; sl < peers.slices[p]; ++sl) {
                const Q* src = base + size_t(sl) * n_pixels + p0;
#pragma unroll
                for (int k = 0; k < 4; ++k) if (k < np) { const Q v = src[k]; s[k][0] += double(v.x); s[k][1] += double(v.y); s[k][2] += double(v.z); s[k][3] += double(v.w); }
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (k < np) {
                const double sc = scale > 0.0 ? scale : 1.0 / s[k][3];
                px[3 * k + 0] = write_color_byte_fast(s[k][0], sc, exact);
                px[3 * k + 1] = write_color_byte_fast(s[k][1], sc, exact);
                px[3 * k + 2] = write_color_byte_fast(s[k][2], sc, exact);
                len += line_len(px + 3 * k);
            } else {
                px[3 * k] = px[3 * k + 1] = px[3 * k + 2] = 0;
            }
        }
    }
    if (rgb8 && p0 < n_pixels) {
        const int np = int(min(size_t(4), n_pixels - p0));
        if (np == 4) {
            unsigned int* dst = reinterpret_cast<unsigned int*>(rgb8 + 3 * p0);
            dst[0] = px[0] | (px[1] << 8) | (px[2] << 16) | (unsigned(px[3]) << 24);
            dst[1] = px[4] | (px[5] << 8) | (px[6] << 16) | (unsigned(px[7]) << 24);
            dst[2] = px[8] | (px[9] << 8) | (px[10] << 16) | (unsigned(px[11]) << 24);
        } else {
            for (int k = 0; k < 3 * np; ++k) rgb8[3 * p0 + k] = px[k];
        }
    }
    unsigned total;
    const unsigned local = block_exclusive_scan(len, warp_sums, total);
    // ---- where does this tile's text start: header + the totals of all tiles before it.  The WHOLE block looks back, 256
    // tiles per step (thread k reads the state of tile - 1 - k): all tiles of a wave start together, so none of the ~600
    // tiles in front of a tile has an inclusive prefix yet when it gets here, and a 32-wide window walked that distance in
    // ~18 dependent L2 round trips with seven warps waiting at the barrier (45 % of the kernel's stall samples at 4K)
    unsigned long long base = (unsigned long long)hdr.len;
    {
        volatile unsigned long long* st = tile_state;
        const unsigned lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
        if (tile == 0) {
            if (threadIdx.x == 0) st[0] = (LB_INC << 62) | (base + total);
        } else {
            if (threadIdx.x == 0) st[tile] = (LB_AGG << 62) | (unsigned long long)total;
            unsigned long long run = 0;
            int look = int(tile) - 1;
            for (;;) {
                const int idx = look - int(threadIdx.x);
                unsigned long long v = LB_INC << 62;                   // before tile 0: an inclusive prefix of 0
                if (idx >= 0) { do { v = st[idx]; } while ((v >> 62) == 0ull); }
                const unsigned inc = __ballot_sync(0xffffffffu, (v >> 62) == LB_INC);
                const int first = inc ? __ffs(inc) - 1 : 32;           // aggregates nearer than the first inclusive prefix + that prefix
                unsigned long long mine = int(lane) <= first ? (v & ((1ull << 62) - 1)) : 0ull;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
                if (lane == 0) { s_part[wid] = mine; s_inc[wid] = inc != 0u; }
                __syncthreads();
                bool done = false;
#pragma unroll
                for (int w = 0; w < ENC_BLOCK / 32; ++w) if (!done) { run += s_part[w]; done = s_inc[w] != 0; }
                __syncthreads();                                       // s_part / s_inc are rewritten by the next step
                if (done) break;
                look -= ENC_BLOCK;
            }
            base = run;
            if (threadIdx.x == 0) { __threadfence(); st[tile] = (LB_INC << 62) | (base + total); }
        }
        if (threadIdx.x == 0 && int(tile) == n_tiles - 1) *total_len = base + total;
    }
    char* dst = text + base;
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

template <typename Q>
cudaError_t launch_resolve_encode(const Q* accum, int slices, const PeerList& peers, size_t n_pixels, double scale, unsigned char* rgb8,
                                  char* text, const EncodeHeader& hdr, unsigned long long* tile_state, unsigned long long* total_len, cudaStream_t st) {
    const int nb = encode_blocks(n_pixels);
    if (nb <= 0) return cudaSuccess;
    // tile_state: [nb] look-back words, then the ticket (one 64-bit word)
    cudaError_t e = cudaMemsetAsync(tile_state, 0, (size_t(nb) + 1) * sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    resolve_encode_kernel<Q><<<nb, ENC_BLOCK, 0, st>>>(accum, slices, peers, n_pixels, scale, rgb8, text, hdr,
                                                       reinterpret_cast<unsigned*>(tile_state + nb), tile_state, total_len, nb, frame_exact());
    return cudaGetLastError();
}
template cudaError_t launch_resolve_encode<float4>(const float4*, int, const PeerList&, size_t, double, unsigned char*, char*, const EncodeHeader&, unsigned long long*, unsigned long long*, cudaStream_t);
template cudaError_t launch_resolve_encode<double4>(const double4*, int, const PeerList&, size_t, double, unsigned char*, char*, const EncodeHeader&, unsigned long long*, unsigned long long*, cudaStream_t);


bool fast_f32_eligible(int slices, const PeerList& peers) { return slices <= RT_FAST_LAYERS && peers.n == 0 && !frame_exact(); }

cudaError_t launch_resolve_count_f32(const float4* accum, int slices, size_t n_pixels, double scale, unsigned char* rgb8, unsigned* block_sums, cudaStream_t st) {
    const int nb = encode_blocks(n_pixels);
    if (nb > 0) resolve_bytes_f32_kernel<<<nb, ENC_BLOCK, 0, st>>>(accum, slices, n_pixels, scale, rgb8, block_sums);
    return cudaGetLastError();
}

cudaError_t launch_encode_p3(const unsigned char* rgb8, size_t n_pixels, const EncodeHeader& hdr, unsigned* block_sums,
                             unsigned long long* block_offsets, unsigned long long* total_len, char* text, cudaStream_t st, bool counted) {
    const int nb = encode_blocks(n_pixels);
    if (nb > 0 && !counted) encode_count_kernel<<<nb, ENC_BLOCK, 0, st>>>(rgb8, n_pixels, block_sums);
    encode_scan_kernel<<<1, ENC_SCAN_BLOCK, 0, st>>>(block_sums, nb, block_offsets, (unsigned long long)hdr.len, total_len);
    if (nb > 0) encode_write_kernel<<<nb, ENC_BLOCK, 0, st>>>(rgb8, n_pixels, block_offsets, text, hdr);
    return cudaGetLastError();
}

}  // namespace rt
