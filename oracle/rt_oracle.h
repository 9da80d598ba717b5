/*
 * rt_oracle.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C, IEEE-double restatement of the CPU path-tracing hot path of
 * ShuvalovIvan/RayTracingProject (reference tree: /root/reference, citations
 * below are relative to it).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this.  The product
 * library (raytracingproject_b200/csrc) never links or calls it.
 *
 * Parity pinning: PINNED.  tests/test_oracle_golden.py checks this file
 * against (i) the reference's committed golden image image.ppm (md5 of the
 * normalised text, 30 spp, mt19937), (ii) the PixelMatch known answer of
 * tests/tests.cpp:35-45, (iii) the reference itself compiled in place
 * (oracle/_ref/libref_harness.so, see ref_harness.cpp) in both RNG modes.
 *
 * Build: gcc -O2 -ffp-contract=off (FMA contraction changes the image, see
 * SURVEY.md section 0 item 9).
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- POD mirrors of the reference's classes ------------------------------ */

typedef struct { double x, y, z; } o_vec3;              /* src/vec3.h:8-70   */

typedef struct {                                        /* src/ray.h:4-27    */
    o_vec3 orig, dir;
    double tm;
} o_ray;

enum { O_LAMBERTIAN = 0, O_METAL = 1, O_DIELECTRIC = 2 };

typedef struct {                                        /* src/material.h    */
    int32_t type;        /* O_LAMBERTIAN / O_METAL / O_DIELECTRIC            */
    int32_t pad;
    o_vec3  albedo;      /* lambertian :28, metal :44 (unused for glass)     */
    double  param;       /* metal fuzz (:45, already clamped) / dielectric ir (:74) */
} o_material;

typedef struct {                                        /* src/sphere.h:60-66 */
    o_vec3  center1;
    o_vec3  center_vec;  /* center2 - center1 (zero when !is_moving)         */
    double  radius;
    int32_t is_moving;
    int32_t mat;         /* index into the material array                    */
} o_sphere;

typedef struct {                                        /* src/hittable.h:7-22 */
    o_vec3  p, normal;
    double  t;
    int32_t front_face;
    int32_t prim;        /* index of the sphere in list order, -1 = miss     */
} o_hit;

typedef struct {                                        /* src/camera.h:15-26,118-125 */
    /* public configuration */
    double aspect_ratio;
    int32_t image_width, samples_per_pixel, max_depth, pad0;
    double vfov;
    o_vec3 lookfrom, lookat, vup;
    double defocus_angle, focus_dist;
    /* derived by o_camera_initialize */
    int32_t image_height, pad1;
    o_vec3 center, pixel00_loc, pixel_delta_u, pixel_delta_v;
    o_vec3 u, v, w, defocus_disk_u, defocus_disk_v;
} o_camera;

/* ---- random numbers ------------------------------------------------------ */

enum { O_RNG_MT19937 = 0, O_RNG_PHILOX = 1 };

typedef struct {
    int32_t  mode;
    /* mt19937 (libstdc++ std::mt19937, default seed 5489) */
    uint32_t mt[624];
    int32_t  mti;
    /* keyed Philox4x32-10: counter = (pixel, sample, stage, k>>2), word k&3 */
    uint32_t key[2];
    uint32_t pixel, sample, stage, k;
    uint32_t block[4];
    uint32_t block_idx;      /* which k>>2 is cached in block[], ~0u = none  */
    uint64_t draws;          /* total draws served (statistics)              */
} o_rng;

void     o_rng_init_mt(o_rng *r, uint32_t seed);         /* seed 5489 = default-constructed std::mt19937 */
void     o_rng_init_philox(o_rng *r, uint64_t seed);
void     o_rng_set_stream(o_rng *r, uint32_t pixel, uint32_t sample, uint32_t stage);
uint32_t o_mt_next_u32(o_rng *r);
void     o_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
double   o_random_double(o_rng *r);                      /* src/rtweekend.h:25-29 */

/* ---- hot path ------------------------------------------------------------ */

void   o_camera_initialize(o_camera *c);                 /* src/camera.h:52-85  */
o_ray  o_get_ray(const o_camera *c, int i, int j, o_rng *r); /* src/camera.h:87-113 */

int    o_sphere_hit(const o_sphere *s, const o_ray *r, double tmin, double tmax, o_hit *rec); /* src/sphere.h:30-57 */
int    o_list_hit(const o_sphere *s, size_t n, const o_ray *r, double tmin, double tmax, o_hit *rec); /* src/hittable_list.h:25-39 */
int    o_scatter(const o_material *m, const o_ray *r_in, const o_hit *rec,
                 o_vec3 *attenuation, o_ray *scattered, o_rng *rng);   /* src/material.h:19-71 */

/* Oracle-side BVH (the reference's bvh.h is a non-compiling stub; only its
 * traversal contract src/bvh.h:16-24 and the slab test src/aabb.h:35-53 are
 * restated).  Used to make big oracle renders affordable; ground truth for
 * closest hit stays o_list_hit. */
typedef struct o_bvh o_bvh;
o_bvh *o_bvh_build(const o_sphere *s, size_t n);
void   o_bvh_free(o_bvh *b);
int    o_bvh_hit(const o_bvh *b, const o_sphere *s, const o_ray *r, double tmin, double tmax, o_hit *rec);
int    o_aabb_hit(const double lo[3], const double hi[3], const o_ray *r, double tmin, double tmax); /* src/aabb.h:35-53 */

typedef struct {
    const o_sphere   *spheres;
    const o_material *materials;
    size_t            n_spheres;
    const o_bvh      *bvh;      /* NULL -> linear list like src/main.cpp:12,70 */
} o_world;

/* src/camera_cpu.h:8-26 (recursive, exactly as written).  rays_out (nullable)
 * counts every entry with depth > 0. */
o_vec3 o_ray_color(const o_ray *r, int depth, const o_world *w, o_rng *rng,
                   int max_depth, uint64_t *rays_out);

/* src/color.h:14-35: sum over samples -> three bytes. */
void   o_write_color(o_vec3 pixel_sum, int spp, uint8_t rgb[3]);

/* src/camera.h:32-50.  Renders rows [row0,row1) into sum[H*W*3] (double,
 * un-normalised per-pixel sums) and rgb8[H*W*3].  mode MT: one sequential
 * stream in pixel order (single thread only).  mode PHILOX: keyed per
 * (pixel,sample,stage); nthreads rows-interleaved pthreads.
 * sample range [s0, s0+ns) of each pixel is rendered (ns = spp normally). */
typedef struct {
    int32_t  rng_mode;
    int32_t  nthreads;
    uint64_t seed;
    int32_t  sample0, nsamples;
    int32_t  row0, row1;       /* row1 <= 0 -> image_height */
    uint64_t mt_skip;          /* MT mode: random_double() draws to discard first.  The reference
                                  builds its scene and renders from ONE global mt19937
                                  (src/main.cpp:19-34 then :70), so the golden image needs the
                                  render to start where scene generation stopped. */
} o_render_opts;

int    o_render(const o_world *w, o_camera *cam, const o_render_opts *opt,
                double *sum, uint8_t *rgb8, uint64_t *rays_out);

/* P3 text exactly as src/camera.h:35 + src/color.h:32-34 emit it. Returns
 * number of bytes written (excluding NUL); buf may be NULL to query size. */
size_t o_ppm_p3(const uint8_t *rgb8, int w, int h, char *buf, size_t cap);

/* src/main.cpp:11-53 scene, generalised:  grid half-width g (11 in the
 * reference), thresholds p_diffuse (0.8) and p_metal (0.95), mt19937 seed
 * (5489 = default).  Draw order is the g++ one.  Returns the
 * number of spheres written (one material per sphere, same index). */
size_t o_scene_rtiow(int g, double p_diffuse, double p_metal, uint32_t mt_seed,
                     o_sphere *s, o_material *m, size_t cap, uint64_t *draws_out /* nullable */);
void   o_camera_rtiow(o_camera *c, int image_width, int spp, int max_depth); /* src/main.cpp:57-68 */

/* assets/shaders/ray_tracing.comp restated (see rt_oracle.c) */
typedef struct { float center[3]; float radius; float color[4]; } o_demo_sphere;   /* src/common_objects.h:17-22 */
int    o_demo_hit(const float origin[3], const float direction[3], const float center[3], float radius);
void   o_demo_compute(const o_demo_sphere *spheres /* NULL: the shader's own */, size_t n, int width, int height, uint8_t *rgba8);

#ifdef __cplusplus
}
#endif
#endif
