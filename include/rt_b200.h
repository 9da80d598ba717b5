/*
 * rt_b200.h -- C ABI of librt_b200.so: the B200-native (sm_100a CUDA) replacement for the
 * CPU path-tracing hot path of ShuvalovIvan/RayTracingProject.
 *
 * The reference has no process/device boundary: `cam.render(world)` (src/main.cpp:70 ->
 * src/camera.h:32-50) runs everything on one host thread.  This header is the boundary the
 * replacement inserts at that call; north_star names its three core entry
 * points rt_upload_scene / rt_render / rt_readback.  Each declaration cites the reference
 * interface it replaces (paths relative to the reference tree).
 *
 * Conventions: plain C, POD structs with fixed layout (doubles, as the reference computes in
 * double), every function returns 0 on success and a negative RT_E* code on failure with a
 * message in rt_last_error().  Nothing here throws, nothing falls back to the CPU: without a
 * CUDA device rt_create fails.  A context is bound to ONE device and is externally
 * synchronised (one caller at a time), like the reference, which is not re-entrant
 * (function-local static RNG, src/rtweekend.h:26-27).
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2u

enum {
    RT_OK = 0,
    RT_E_INVALID = -1,   /* bad argument / ABI mismatch */
    RT_E_CUDA = -2,      /* CUDA runtime error (message has the cudaError string) */
    RT_E_STATE = -3,     /* call order (e.g. render before upload) */
    RT_E_NOMEM = -4
};

/* material tags: src/material.h:15 (lambertian), :31 (metal), :48 (dielectric) */
enum { RT_LAMBERTIAN = 0, RT_METAL = 1, RT_DIELECTRIC = 2 };

/* arithmetic type of the device path.  F64 = "parity mode": IEEE double, same operation order
 * as the reference, no FMA contraction.  F32 = fast mode (big spheres intersected in double). */
enum { RT_F32 = 0, RT_F64 = 1 };

/* bounce-loop variants (north_star item 3).  RT_MEGAKERNEL: one persistent kernel whose warps
 * schedule node / leaf / shade sections among their lanes; RT_MEGA_LOCKSTEP: the first, lock-step
 * form of it (kept for the ncu A/B in profiles/); RT_WAVEFRONT: queue-based stages. */
enum { RT_MEGAKERNEL = 0, RT_WAVEFRONT = 1, RT_MEGA_LOCKSTEP = 2 };

/* closest-hit back ends for rt_trace_closest / rt_bench_closest */
enum { RT_TRACE_BVH = 0, RT_TRACE_BRUTE = 1 };

/* acceleration-structure builders */
enum { RT_BVH_SAH_HOST = 0, RT_BVH_LBVH_GPU = 1 };

/* src/sphere.h:60-66 (private fields of `sphere`; static :9-14, moving :19-28) */
typedef struct {
    double  center1[3];
    double  center_vec[3];   /* center2 - center1; ignored unless is_moving */
    double  radius;
    int32_t is_moving;
    int32_t material;        /* index into the material array */
} rt_sphere;

/* src/material.h:27-28 (lambertian albedo), :43-45 (metal albedo, fuzz<=1), :73-74 (ir) */
typedef struct {
    int32_t type;
    int32_t pad;
    double  albedo[3];
    double  param;           /* metal: fuzz, dielectric: index of refraction */
} rt_material;

/* public fields of `camera`, src/camera.h:15-26 */
typedef struct {
    double  aspect_ratio;
    int32_t image_width;
    int32_t samples_per_pixel;
    int32_t max_depth;
    int32_t pad;
    double  vfov;
    double  lookfrom[3], lookat[3], vup[3];
    double  defocus_angle, focus_dist;
} rt_camera;

/* src/ray.h:23-26 */
typedef struct { double orig[3], dir[3], tm; } rt_ray;

/* src/hittable.h:7-14 plus the list index of the sphere that was hit (-1 = miss) */
typedef struct {
    double  p[3], normal[3], t;
    int32_t front_face;
    int32_t prim;
} rt_hit;

typedef struct {
    uint32_t abi;            /* RT_ABI_VERSION */
    int32_t  builder;        /* RT_BVH_SAH_HOST | RT_BVH_LBVH_GPU */
    int32_t  max_leaf;       /* spheres per leaf, 1..8 (0 = default: 4 for the host SAH builder, 1 for the GPU LBVH) */
    int32_t  wide;           /* 1: walk the tree in global memory through 4-wide node records (one 128-byte fetch = two
                                levels).  0 (default) / -1: binary records -- measured 2 % faster on C5 */
    double   big_radius;     /* spheres with radius >= this are kept out of the tree and tested first,
                                in double (the r = 1000 ground sphere; SURVEY.md 7 hard parts 1, 5).
                                0 = default 100; negative = never */
} rt_bvh_opts;

typedef struct {
    uint32_t abi;            /* RT_ABI_VERSION */
    int32_t  precision;      /* RT_F32 | RT_F64 */
    int32_t  variant;        /* RT_MEGAKERNEL | RT_WAVEFRONT */
    int32_t  sample_begin;   /* first sample index of every pixel rendered by this call ... */
    int32_t  sample_count;   /* ... and how many (0 = all of camera.samples_per_pixel): the shard
                                a rank renders under sample-range sharding */
    int32_t  accumulate;     /* 0: overwrite the accumulation buffer, 1: add to it (progressive) */
    uint64_t seed;           /* Philox key; stream = (pixel, sample, stage), see DESIGN.md */
    int32_t  slices;         /* sample slices per pixel rendered by different warps (0 = auto) */
    int32_t  sched;          /* tuning: th_s | th_l << 8 | shape+1 << 16 | th_n << 20 of the warp scheduler, bulk jobs per warp << 24 (0 = defaults) */
    int32_t  tile_rank;      /* tile-interleaved sharding: with tile_world > 1 this call renders ALL its samples, but only of the */
    int32_t  tile_world;     /* tile pairs whose index is tile_rank modulo tile_world (RT_WAVEFRONT only; 0 or 1 = the whole image) */
    int32_t  top_cap;        /* tuning: cap on the tree nodes staged in shared memory for a scene in global memory (0 = all that fits, -1 = none) */
    int32_t  shared_jobs;    /* N >= 1: this call renders whatever (tile, sample range) jobs of the WHOLE frame it gets from the job counter it */
} rt_render_opts;            /*    shares with N - 1 other GPUs (rt_jobs_export / rt_jobs_open / rt_jobs_share); 0: the call's own counter */

typedef struct {
    uint64_t rays;           /* path segments traced by the last render (ray_color entries, depth > 0) */
    uint64_t samples;        /* W*H*sample_count of the last render */
    double   render_ms;      /* device time of the last render (CUDA events on the ctx stream) */
    double   resolve_ms;     /* device time of the last resolve/encode */
    double   build_ms;       /* BVH build (host wall or device time) of the last upload */
    double   upload_ms;
    int32_t  n_spheres, n_big, n_nodes, bvh_depth;
    int32_t  image_width, image_height;
    int32_t  slices, kernel_launches;   /* launches of this library's kernels in the last render+resolve */
    int32_t  scene_in_smem;  /* last render: 1 = whole scene staged in shared memory, 2 = top levels of the tree only, 0 = global */
    int32_t  big_in_tree;    /* spheres with radius >= big_radius beyond the 16 largest: they are in the tree like any sphere */
    int32_t  top_nodes_in_smem;   /* last render, scene_in_smem == 2: how many nodes of the tree's first levels were staged */
    int32_t  pad;
} rt_stats;

typedef struct rt_ctx rt_ctx;

/* ---- lifetime -------------------------------------------------------------------------- */
int         rt_create(rt_ctx **ctx, int device);      /* fails (RT_E_CUDA) when no CUDA device */
int         rt_destroy(rt_ctx *ctx);
const char *rt_last_error(const rt_ctx *ctx);          /* ctx may be NULL: last creation error */
uint32_t    rt_abi_version(void);

/* ---- the render path: replaces camera::render(world), src/camera.h:32-50 ------------------ */

/* world -> device.  Replaces the hittable_list of shared_ptr<sphere> (src/hittable_list.h:13,
 * src/main.cpp:12-53): spheres/materials are copied into tagged SoA buffers and a BVH is built
 * (bvh.h is a stub in the reference; src/bvh.h:16-24 fixes only the traversal contract). */
int rt_upload_scene(rt_ctx *ctx, const rt_sphere *spheres, size_t n_spheres,
                    const rt_material *materials, size_t n_materials, const rt_bvh_opts *opts);

/* Next row 8f-4, dynamic scenes: the world of the last rt_upload_scene with its spheres moved -- same count, same order,
 * other centres / motion vectors (src/sphere.h:19-28) / radii / materials.  The tree keeps its topology and is REFITTED on the
 * device (every box recomputed bottom-up); no rebuild.  Hits stay exact; the tree's quality ages with the distance moved.
 * RT_E_STATE when the count changed or a sphere crossed big_radius (the split into tree / outsized spheres is fixed at upload). */
int rt_update_spheres(rt_ctx *ctx, const rt_sphere *spheres, size_t n_spheres,
                      const rt_material *materials, size_t n_materials);

/* camera::initialize + the pixel/sample/bounce loops (src/camera.h:37-47, src/camera_cpu.h:8-26).
 * Asynchronous: enqueues on the context's stream. */
int rt_render(rt_ctx *ctx, const rt_camera *cam, const rt_render_opts *opts);

/* write_color for every pixel + the P3 stream camera::render prints (src/color.h:14-35,
 * src/camera.h:35,45).  Synchronises.  Every output pointer is optional:
 *   sum_rgb  H*W*3 doubles, per-pixel radiance sums (pixel_color before write_color)
 *   rgb8     H*W*3 bytes, exactly the integers write_color prints
 *   ppm      P3 text, byte-identical to the reference's stdout; *ppm_len in = capacity, out = length
 * spp_total = divisor used by write_color (0 = samples rendered into the buffer so far). */
int rt_readback(rt_ctx *ctx, int spp_total, double *sum_rgb, uint8_t *rgb8, char *ppm, size_t *ppm_len);

/* The same hand-over in two halves, for a caller that renders frame after frame (the loop around camera::render in a caller
 * of its own; the reference prints one frame, src/main.cpp:70): rt_readback_begin resolves the frame and the P3 text on the
 * device, waits for that, starts the device-to-host copies of rgb8 (H*W*3 bytes) and / or the text (ppm_capacity >= its
 * length, at most H*W*12 + 64) on a copy stream of the context and returns; rt_readback_end waits for the copies and gives the
 * text's length.  Between the two the caller may rt_upload_scene / rt_render the next frame -- the copies overlap its render --
 * but must leave the host buffers alone; pinned host memory makes the copies asynchronous.  One split read-back at a time
 * (RT_E_STATE otherwise); on a tile-sharded frame (rt_frame_export) the peers must not deliver the next frame before
 * rt_readback_end has returned. */
int rt_readback_begin(rt_ctx *ctx, int spp_total, uint8_t *rgb8, char *ppm, size_t ppm_capacity);
int rt_readback_end(rt_ctx *ctx, size_t *ppm_len);

/* ---- next row 8f-2: output formats next to P3, progressive preview ------------------------------ *
 * Progressive accumulation itself is rt_render with accumulate = 1 over further sample ranges
 * (sample_begin / sample_count): the streams are keyed by (pixel, sample), so any split of [0, spp)
 * into passes sums the same samples.  These two read the buffer between passes. */
enum { RT_IMG_P3 = 0,      /* the text camera::render prints (src/camera.h:35, src/color.h:32-34) */
       RT_IMG_P6 = 1,      /* "P6\nW H\n255\n" + the same write_color bytes, 3 per pixel */
       RT_IMG_PFM = 2 };   /* "PF\nW H\n-1.0\n" + float32 mean radiance (pixel_color / spp, src/color.h:20-23,
                              before gamma and clamp), little endian, rows bottom to top */

/* write_color for every pixel in one of the formats above.  buf == NULL: *len = bytes needed.
 * Otherwise *len in = capacity, out = bytes written.  Synchronises. */
int rt_readback_image(rt_ctx *ctx, int spp_total, int format, void *buf, size_t *len);

/* The frame as write_color bytes + alpha 255 (4 bytes per pixel, row-major) in a DEVICE buffer the
 * context owns: the compute-side substitute for the image start_interactive_loop would present
 * (src/graphical_environment.h:31; window and swapchain: unavailable on this box).  Asynchronous
 * on the context's stream unless host_rgba8 (optional, W*H*4 bytes) is given.  *dev_rgba8 stays
 * valid until the next rt_preview with another image size, or rt_destroy. */
int rt_preview(rt_ctx *ctx, int spp_total, void **dev_rgba8, uint8_t *host_rgba8);

int rt_image_size(const rt_camera *cam, int *width, int *height);   /* src/camera.h:53-54 */
int rt_synchronize(rt_ctx *ctx);
int rt_get_stats(rt_ctx *ctx, rt_stats *out);
/* profiling aid: counters of the wavefront kernel's scheduler when the library was built with
 * -DRT_POOL_STATS (zeros otherwise); out[0] = rays of the last render */
int rt_debug_counters(rt_ctx *ctx, uint64_t out[32]);
/* profiling aid: per-warp timeline records of a -DRT_WARP_TIMES build (8 words per warp; zeros otherwise) */
int rt_debug_warp_times(rt_ctx *ctx, uint64_t *out, size_t n_words);

/* host logic, no device needed: the sample layers rt_render cuts a wavefront frame into when rt_render_opts.slices = 0
 * (bulk layers, then layers that halve).  sizes[] has room for 64 entries; returns how many were written
 * (they sum to sample_count), *n_bulk = the leading bulk layers.  jobs_per_warp 0 = default. */
int rt_plan_layers(int sample_count, long long tiles, long long resident_warps, int jobs_per_warp, int *sizes, int *n_bulk);

/* ---- multi-GPU plumbing (one process per GPU) -------------------------------- */

/* Device pointer / geometry of the accumulation buffer so the host framework can reduce it
 * across ranks (NCCL) or map it into a peer (CUDA IPC).  elem_bytes 4 (RT_F32, float4 per
 * pixel) or 8 (RT_F64, double4 per pixel); n_elems = slices*W*H*4
 * (layer-major: the first W*H*4 elements are layer 0). */
int rt_accum_buffer(rt_ctx *ctx, void **dev_ptr, size_t *n_elems, int *elem_bytes);
/* Sums the sample layers (rt_render_opts.slices) of the accumulation buffer into layer 0 and zeroes
 * the others, on the context's stream: after it the first W*H*4 elements are this rank's whole partial
 * sum, so one layer per rank crosses NVLink in the combine.  Every rank of a combine calls it (the
 * fused resolve then reads layer 0 of each peer only).  The next rt_render ends the folded state. */
int rt_fold_slices(rt_ctx *ctx);
int rt_stream(rt_ctx *ctx, void **cuda_stream);
/* The fused peer reduce + resolve: rank 0 reads every peer's partial buffer over NVLink inside the resolve
 * kernel.  rt_ipc_export writes RT_IPC_HANDLE_BYTES: the CUDA IPC handle of the accumulation buffer + its
 * image size and precision; rt_ipc_open_peers (n_peers blobs back to back) refuses peers whose frame differs
 * from this context's.  Only layer 0 of a peer is read, so every rank must rt_fold_slices before rank 0
 * resolves (rt_readback fails with RT_E_STATE otherwise).  A context whose frame changes shape drops its
 * mappings; a rank whose buffer was reallocated (larger frame or more layers) must export again. */
#define RT_IPC_HANDLE_BYTES 80
int rt_ipc_export(rt_ctx *ctx, void *handle);
int rt_ipc_open_peers(rt_ctx *ctx, const void *handles, int n_peers);   /* peers other than self */
int rt_ipc_close_peers(rt_ctx *ctx);

/* Tile-interleaved sharding (rt_render_opts.tile_rank / tile_world): every rank renders ALL samples of its share of
 * the tiles, so the combine has no arithmetic -- each rank's resolve kernel writes the write_color bytes of its tiles
 * straight into the ROOT's image over NVLink (CUDA IPC mapping) and bumps a counter there; the root's rt_readback
 * resolves its own tiles, waits on its stream for the counter (every peer delivered), then encodes / copies.
 *   root:  rt_render(tile_rank = r0, tile_world = N) once, rt_frame_export -> blob to every peer; per frame
 *          rt_render(...), rt_readback(...)
 *   peer:  rt_frame_open(blob); per frame rt_render(tile_rank = r, tile_world = N), rt_resolve_tiles(spp_total)
 * A peer must not start frame k + 1 before the root's rt_readback of frame k has returned (one barrier per frame).
 * The per-pixel sums (sum_rgb, PFM) of the root cover its own tiles only. */
/* Dynamic sharding (a third split next to sample ranges and tiles): the GPUs of a node take the jobs of ONE frame
 * -- (tile, sample layer) pairs, the unit the wavefront kernel's warps already fetch -- from one counter that lives in the root
 * GPU's memory; every warp of every GPU advances it with a system-scope atomic over NVLink.  A GPU that is a few per cent slower
 * (clocks, neighbours) simply takes fewer jobs, so all of them finish together: no rank skew in front of the combine.  Every rank
 * renders into its own accumulation buffer (zeroed first: it holds only the jobs this rank took); the buffers are then combined
 * like sample shards (rt_fold_slices + reduce, or the fused IPC resolve).
 *   root:  rt_jobs_export(blob) once -> blob to every peer;   peer: rt_jobs_open(blob) once
 *   every rank, every frame: rt_render(ALL samples, shared_jobs = number of GPUs), rt_fold_slices, combine
 * All ranks must render the same frames in the same order (the counters are used round-robin by frame number, the root clears
 * them four frames ahead) and stay within three frames of each other -- any per-frame combine guarantees that.
 * RT_WAVEFRONT + RT_F32 only (the order of a pixel's samples is free there); scene and camera must be the same on all ranks.
 * rt_jobs_share: the same for several contexts of ONE process (their devices must be able to reach the root's memory). */
int rt_jobs_export(rt_ctx *ctx, void *handle);             /* RT_IPC_HANDLE_BYTES */
int rt_jobs_open(rt_ctx *ctx, const void *handle);
int rt_jobs_share(rt_ctx *ctx, rt_ctx *root);
int rt_jobs_close(rt_ctx *ctx);

int rt_frame_export(rt_ctx *ctx, void *handle);            /* RT_IPC_HANDLE_BYTES; re-export after the image size changes */
int rt_frame_open(rt_ctx *ctx, const void *handle);
int rt_frame_close(rt_ctx *ctx);
int rt_resolve_tiles(rt_ctx *ctx, int spp_total);          /* asynchronous on the context's stream */

/* ---- parity / microbenchmark hooks (same code the render path runs) ------------------------- */

/* closest hit in (tmin, tmax): hittable_list::hit, src/hittable_list.h:25-39 (ground truth,
 * RT_TRACE_BRUTE) or the BVH traversal (src/bvh.h:16-24 contract, RT_TRACE_BVH). */
int rt_trace_closest(rt_ctx *ctx, const rt_ray *rays, size_t n, double tmin, double tmax,
                     int precision, int backend, rt_hit *out);

/* material::scatter (src/material.h:19-71) for n independent (material, ray, hit) triples;
 * element i draws from the stream (pixel = i, sample = 0, stage = 1) of `seed`. */
int rt_scatter(rt_ctx *ctx, const rt_material *materials, const rt_ray *rays_in, const rt_hit *hits,
               size_t n, uint64_t seed, int precision,
               int32_t *scattered_flag, double *attenuation_rgb, rt_ray *scattered);

/* camera::get_ray (src/camera.h:87-113) for n (i, j, sample) triples, stream stage 0. */
int rt_gen_rays(rt_ctx *ctx, const rt_camera *cam, const int32_t *px_i, const int32_t *px_j,
                const uint32_t *sample, size_t n, uint64_t seed, int precision, rt_ray *out);

/* BASELINE config 2: n synthetic rays generated in-kernel (even index: get_ray of a virtual
 * 8192x8192 image; odd: incoherent origin/direction) traced against the uploaded
 * scene.  prim_out/t_out (host, optional) receive per-ray results; rays_out (host, optional)
 * receives the rays in double so an oracle can re-trace them.  *ms = device time of the trace. */
int rt_bench_closest(rt_ctx *ctx, const rt_camera *cam, size_t n, uint64_t seed, int precision,
                     int backend, int repeats, double *ms, int32_t *prim_out, float *t_out,
                     rt_ray *rays_out, size_t rays_out_n);

/* write_color (src/color.h:14-35) for n per-pixel sums (3 doubles each) through the resolve kernel,
 * and the P3 encoder (src/camera.h:35 header + src/color.h:32-34 lines) for a w*h rgb8 image. */
int rt_write_color(rt_ctx *ctx, const double *sums_rgb, size_t n, int samples_per_pixel, uint8_t *rgb8);
int rt_encode_p3(rt_ctx *ctx, const uint8_t *rgb8, int width, int height, char *ppm, size_t *ppm_len);

/* The P3 encoder for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid allocation).  NO COUNTERPART IN THE REFERENCE.  Same text as
 * rt_encode_p3, but every buffer is the caller's and lives on the context's device: dev_rgb8 (width*height*3 bytes, read
 * exactly, no padding needed), dev_ppm (ppm_capacity >= 32 + 12*width*height bytes: the worst case is required up front,
 * the length is not known before the pass), *dev_ppm_len (device, 8 bytes: the text's length).  Asynchronous on the
 * context's stream (rt_stream / rt_synchronize).  The prefix-sum scratch is a static __device__ array inside the library
 * (128 KB on every GPU that loads it, until exit) sized for RT_STATIC_MAX_PIXELS; larger images are refused with
 * RT_E_INVALID and a message.  The call allocates nothing on host or device (no cudaMalloc / cudaFree, no library that
 * does).  All calls share the one array per GPU: calls on the same stream are ordered and safe, calls that overlap on a
 * GPU (two contexts, two streams) race on it -- serialise them. */
#define RT_STATIC_MAX_PIXELS (1 << 24)
int rt_encode_p3_static(rt_ctx *ctx, const uint8_t *dev_rgb8, int width, int height, char *dev_ppm, size_t ppm_capacity,
                        unsigned long long *dev_ppm_len);

/* ---- next row 8f-3: the demo's compute shader as a CUDA kernel ---------------------------------- */

/* `Sphere` of the windowed demo, src/common_objects.h:17-22 (glm::vec3 center; float radius; glm::vec4 color) */
typedef struct { float center[3]; float radius; float color[4]; } rt_demo_sphere;

/* assets/shaders/ray_tracing.comp:35-73 for every pixel of a width x height rgba8 image (host buffer,
 * 4 bytes per pixel, row-major from y = 0): pinhole ray against the spheres by the sign of the
 * discriminant, flat colour, background (0.5, 0, 0.25, 1).  spheres == NULL: the sphere the shader
 * hard-codes (:49-52).  Headless: replaces imageStore into the swapchain-bound storage image and the
 * vkCmdDispatch of src/vulkan/command_buffers.cpp:143-158 (graphics API: unavailable on this box). */
int rt_demo_compute(rt_ctx *ctx, const rt_demo_sphere *spheres, size_t n, int width, int height, uint8_t *rgba8);

/* Host SAH builder on its own (no device needed): nodes_out receives 16 floats per inner node
 * (layout in DESIGN.md), order_out the slot -> sphere index permutation.  Capacity: n-1 nodes, n slots. */
int rt_build_bvh_host(const rt_sphere *spheres, size_t n, int max_leaf, float *nodes_out, int32_t *order_out,
                      int32_t *root_ref, int32_t *depth, int32_t *n_inner);

/* src/main.cpp:11-68 scene + camera generators (default-seeded mt19937, g++ draw order),
 * generalised for BASELINE configs 4 and 5: grid half-width g (11), thresholds p_diffuse (0.8),
 * p_metal (0.95).  Returns the sphere count (call with cap = 0 to size the arrays). */
size_t rt_scene_rtiow(int g, double p_diffuse, double p_metal, uint32_t mt_seed,
                      rt_sphere *spheres, rt_material *materials, size_t cap);
int    rt_camera_rtiow(rt_camera *cam, int image_width, int spp, int max_depth, double scale);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
