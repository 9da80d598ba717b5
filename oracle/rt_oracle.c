/*
 * rt_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 * See rt_oracle.h for the scope statement.  Every function cites the
 * reference lines (relative to /root/reference) whose arithmetic AND
 * operation order it restates.  All arithmetic is IEEE double; compile with
 * -ffp-contract=off.
 */
#include "rt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ vec3 -- */
/* src/vec3.h:81-119 -- note operator/(v,t) is (1/t)*v, not v.e/t (:101-103) */

static inline o_vec3 v3(double x, double y, double z) { o_vec3 r = {x, y, z}; return r; }
static inline o_vec3 vadd(o_vec3 a, o_vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline o_vec3 vsub(o_vec3 a, o_vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline o_vec3 vmul(o_vec3 a, o_vec3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline o_vec3 vscale(double t, o_vec3 a) { return v3(t * a.x, t * a.y, t * a.z); }
static inline o_vec3 vdiv(o_vec3 a, double t) { return vscale(1 / t, a); }
static inline o_vec3 vneg(o_vec3 a) { return v3(-a.x, -a.y, -a.z); }
static inline double vdot(o_vec3 a, o_vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline double vlen2(o_vec3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
static inline double vlen(o_vec3 a) { return sqrt(vlen2(a)); }
static inline o_vec3 vunit(o_vec3 a) { return vdiv(a, vlen(a)); }
static inline o_vec3 vcross(o_vec3 u, o_vec3 v) {
    return v3(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
static inline o_vec3 ray_at(const o_ray *r, double t) { return vadd(r->orig, vscale(t, r->dir)); } /* src/ray.h:19-21 */

/* ------------------------------------------------------------------- rng -- */

/* std::mt19937 (MT19937, 32 bit) as libstdc++ implements it; default seed 5489. */
void o_rng_init_mt(o_rng *r, uint32_t seed) {
    memset(r, 0, sizeof *r);
    r->mode = O_RNG_MT19937;
    r->mt[0] = seed;
    for (int i = 1; i < 624; ++i)
        r->mt[i] = 1812433253u * (r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) + (uint32_t)i;
    r->mti = 624;
    r->block_idx = ~0u;
}

uint32_t o_mt_next_u32(o_rng *r) {
    uint32_t *mt = r->mt;
    if (r->mti >= 624) {
        for (int k = 0; k < 624; ++k) {
            uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
            mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        r->mti = 0;
    }
    uint32_t y = mt[r->mti++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
}

/* Philox4x32-10 (Salmon et al., SC'11), the counter-based generator that
 * replaces src/rtweekend.h:25-29 in the keyed-stream contract. */
void o_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int round = 0; round < 10; ++round) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void o_rng_init_philox(o_rng *r, uint64_t seed) {
    memset(r, 0, sizeof *r);
    r->mode = O_RNG_PHILOX;
    r->key[0] = (uint32_t)seed;
    r->key[1] = (uint32_t)(seed >> 32);
    r->block_idx = ~0u;
}

void o_rng_set_stream(o_rng *r, uint32_t pixel, uint32_t sample, uint32_t stage) {
    if (r->mode != O_RNG_PHILOX) return;   /* mt19937: one global sequential stream */
    r->pixel = pixel; r->sample = sample; r->stage = stage;
    r->k = 0;
    r->block_idx = ~0u;
}

/* src/rtweekend.h:25-29.  MT mode: uniform_real_distribution<double>(0,1) on
 * mt19937 == libstdc++ generate_canonical<double,53>: two 32-bit draws,
 * (g1 + g2*2^32) / 2^64, clamped below 1.  PHILOX mode: u = (word>>8)*2^-24,
 * exactly representable in float and double. */
double o_random_double(o_rng *r) {
    r->draws++;
    if (r->mode == O_RNG_MT19937) {
        double sum = 0.0, tmp = 1.0;
        sum += (double)o_mt_next_u32(r) * tmp; tmp *= 4294967296.0;
        sum += (double)o_mt_next_u32(r) * tmp; tmp *= 4294967296.0;
        double ret = sum / tmp;
        if (ret >= 1.0) ret = nextafter(1.0, 0.0);
        return ret;
    }
    uint32_t blk = r->k >> 2;
    if (blk != r->block_idx) {
        uint32_t ctr[4] = { r->pixel, r->sample, r->stage, blk };
        o_philox4x32_10(ctr, r->key, r->block);
        r->block_idx = blk;
    }
    uint32_t word = r->block[r->k & 3u];
    r->k++;
    return (double)(word >> 8) * (1.0 / 16777216.0);
}

/* src/rtweekend.h:31-34 */
static inline double random_range(o_rng *r, double min, double max) {
    return min + (max - min) * o_random_double(r);
}

/* src/vec3.h:63-69.  g++ evaluates constructor arguments right to left, so the
 * FIRST draw lands in z, the second in y, the third in x. */
static o_vec3 vec3_random(o_rng *r) {
    double z = o_random_double(r);
    double y = o_random_double(r);
    double x = o_random_double(r);
    return v3(x, y, z);
}
static o_vec3 vec3_random_range(o_rng *r, double min, double max) {
    double z = random_range(r, min, max);
    double y = random_range(r, min, max);
    double x = random_range(r, min, max);
    return v3(x, y, z);
}

/* src/vec3.h:121-127 (first draw -> y, second -> x) */
static o_vec3 random_in_unit_disk(o_rng *r) {
    for (;;) {
        double y = random_range(r, -1, 1);
        double x = random_range(r, -1, 1);
        o_vec3 p = v3(x, y, 0);
        if (vlen2(p) < 1) return p;
    }
}

/* src/vec3.h:129-135 */
static o_vec3 random_in_unit_sphere(o_rng *r) {
    for (;;) {
        o_vec3 p = vec3_random_range(r, -1, 1);
        if (vlen2(p) < 1) return p;
    }
}

/* src/vec3.h:137-139 */
static o_vec3 random_unit_vector(o_rng *r) { return vunit(random_in_unit_sphere(r)); }

/* src/vec3.h:149-151 */
static o_vec3 reflect(o_vec3 v, o_vec3 n) { return vsub(v, vscale(2 * vdot(v, n), n)); }

/* src/vec3.h:153-158 */
static o_vec3 refract(o_vec3 uv, o_vec3 n, double etai_over_etat) {
    double cos_theta = fmin(vdot(vneg(uv), n), 1.0);
    o_vec3 r_out_perp = vscale(etai_over_etat, vadd(uv, vscale(cos_theta, n)));
    o_vec3 r_out_parallel = vscale(-sqrt(fabs(1.0 - vlen2(r_out_perp))), n);
    return vadd(r_out_perp, r_out_parallel);
}

/* ---------------------------------------------------------------- camera -- */

/* src/camera.h:52-85 */
void o_camera_initialize(o_camera *c) {
    c->image_height = (int)(c->image_width / c->aspect_ratio);
    c->image_height = (c->image_height < 1) ? 1 : c->image_height;

    c->center = c->lookfrom;

    double theta = c->vfov * 3.1415926535897932385 / 180.0;          /* src/rtweekend.h:17,21-23 */
    double h = tan(theta / 2);
    double viewport_height = 2 * h * c->focus_dist;
    double viewport_width = viewport_height * ((double)c->image_width / c->image_height);

    c->w = vunit(vsub(c->lookfrom, c->lookat));
    c->u = vunit(vcross(c->vup, c->w));
    c->v = vcross(c->w, c->u);

    o_vec3 viewport_u = vscale(viewport_width, c->u);
    o_vec3 viewport_v = vscale(viewport_height, vneg(c->v));

    c->pixel_delta_u = vdiv(viewport_u, c->image_width);
    c->pixel_delta_v = vdiv(viewport_v, c->image_height);

    o_vec3 viewport_upper_left =
        vsub(vsub(vsub(c->center, vscale(c->focus_dist, c->w)), vdiv(viewport_u, 2)), vdiv(viewport_v, 2));
    c->pixel00_loc = vadd(viewport_upper_left, vscale(0.5, vadd(c->pixel_delta_u, c->pixel_delta_v)));

    double defocus_radius = c->focus_dist * tan((c->defocus_angle / 2) * 3.1415926535897932385 / 180.0);
    c->defocus_disk_u = vscale(defocus_radius, c->u);
    c->defocus_disk_v = vscale(defocus_radius, c->v);
}

/* src/camera.h:87-113.  Draw order: px, py (:104-105), disk loop (:111),
 * time (:97). */
o_ray o_get_ray(const o_camera *c, int i, int j, o_rng *rng) {
    o_vec3 pixel_center = vadd(vadd(c->pixel00_loc, vscale((double)i, c->pixel_delta_u)),
                               vscale((double)j, c->pixel_delta_v));
    double px = -0.5 + o_random_double(rng);
    double py = -0.5 + o_random_double(rng);
    o_vec3 pixel_sample = vadd(pixel_center, vadd(vscale(px, c->pixel_delta_u), vscale(py, c->pixel_delta_v)));

    o_vec3 ray_origin;
    if (c->defocus_angle <= 0) {
        ray_origin = c->center;
    } else {
        o_vec3 p = random_in_unit_disk(rng);
        ray_origin = vadd(vadd(c->center, vscale(p.x, c->defocus_disk_u)), vscale(p.y, c->defocus_disk_v));
    }
    o_ray r;
    r.dir = vsub(pixel_sample, ray_origin);
    r.orig = ray_origin;
    r.tm = o_random_double(rng);
    return r;
}

/* ---------------------------------------------------------------- sphere -- */

/* src/sphere.h:30-57 + src/hittable.h:15-21; open interval per src/interval.h:33-35 */
int o_sphere_hit(const o_sphere *s, const o_ray *r, double tmin, double tmax, o_hit *rec) {
    o_vec3 center = s->is_moving ? vadd(s->center1, vscale(r->tm, s->center_vec)) : s->center1; /* :31,:68-72 */
    o_vec3 oc = vsub(r->orig, center);
    double a = vlen2(r->dir);
    double half_b = vdot(oc, r->dir);
    double c = vlen2(oc) - s->radius * s->radius;

    double discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return 0;

    double sqrtd = sqrt(discriminant);
    double root = (-half_b - sqrtd) / a;
    if (!(tmin < root && root < tmax)) {
        root = (-half_b + sqrtd) / a;
        if (!(tmin < root && root < tmax)) return 0;
    }

    rec->t = root;
    rec->p = ray_at(r, rec->t);
    o_vec3 outward_normal = vdiv(vsub(rec->p, center), s->radius);
    rec->front_face = vdot(r->dir, outward_normal) < 0;
    rec->normal = rec->front_face ? outward_normal : vneg(outward_normal);
    return 1;
}

/* src/hittable_list.h:25-39: first in list wins exact ties (strict <). */
int o_list_hit(const o_sphere *s, size_t n, const o_ray *r, double tmin, double tmax, o_hit *rec) {
    o_hit temp;
    int hit_anything = 0;
    double closest_so_far = tmax;
    for (size_t i = 0; i < n; ++i) {
        if (o_sphere_hit(&s[i], r, tmin, closest_so_far, &temp)) {
            hit_anything = 1;
            closest_so_far = temp.t;
            temp.prim = (int32_t)i;
            *rec = temp;
        }
    }
    if (!hit_anything) rec->prim = -1;
    return hit_anything;
}

/* ------------------------------------------------------------- aabb / bvh -- */

/* src/aabb.h:35-53 exactly as written, including 1/0 = inf and NaN compares. */
int o_aabb_hit(const double lo[3], const double hi[3], const o_ray *r, double tmin, double tmax) {
    const double d[3] = { r->dir.x, r->dir.y, r->dir.z };
    const double o[3] = { r->orig.x, r->orig.y, r->orig.z };
    for (int a = 0; a < 3; a++) {
        double invD = 1 / d[a];
        double orig = o[a];
        double t0 = (lo[a] - orig) * invD;
        double t1 = (hi[a] - orig) * invD;
        if (invD < 0) { double tt = t0; t0 = t1; t1 = tt; }
        if (t0 > tmin) tmin = t0;
        if (t1 < tmax) tmax = t1;
        if (tmax <= tmin) return 0;
    }
    return 1;
}

typedef struct {
    double lo[3], hi[3];
    int32_t left, right;     /* inner: child node indices; leaf: left = -1 - sphere index */
} o_bvh_node;

struct o_bvh {
    o_bvh_node *nodes;
    size_t n_nodes, cap;
};

/* src/sphere.h:9-14,19-28 + src/aabb.h:12-27: bounding box of a sphere. */
static void sphere_box(const o_sphere *s, double lo[3], double hi[3]) {
    const double c1[3] = { s->center1.x, s->center1.y, s->center1.z };
    const double cv[3] = { s->center_vec.x, s->center_vec.y, s->center_vec.z };
    for (int a = 0; a < 3; ++a) {
        double l = fmin(c1[a] - s->radius, c1[a] + s->radius);
        double h = fmax(c1[a] - s->radius, c1[a] + s->radius);
        if (s->is_moving) {
            double c2 = c1[a] + cv[a];
            l = fmin(l, fmin(c2 - s->radius, c2 + s->radius));
            h = fmax(h, fmax(c2 - s->radius, c2 + s->radius));
        }
        lo[a] = l; hi[a] = h;
    }
}

typedef struct { const o_sphere *s; int axis; } sort_ctx;
static sort_ctx g_sort_ctx;            /* build is single-threaded */
static int cmp_centroid(const void *pa, const void *pb) {
    int ia = *(const int *)pa, ib = *(const int *)pb;
    double la[3], ha[3], lb[3], hb[3];
    sphere_box(&g_sort_ctx.s[ia], la, ha);
    sphere_box(&g_sort_ctx.s[ib], lb, hb);
    double ca = la[g_sort_ctx.axis] + ha[g_sort_ctx.axis];
    double cb = lb[g_sort_ctx.axis] + hb[g_sort_ctx.axis];
    if (ca < cb) return -1;
    if (ca > cb) return 1;
    return (ia > ib) - (ia < ib);
}

static int32_t bvh_build_rec(o_bvh *b, const o_sphere *s, int *idx, size_t start, size_t end) {
    int32_t me = (int32_t)b->n_nodes++;
    o_bvh_node *nd = &b->nodes[me];
    for (int a = 0; a < 3; ++a) { nd->lo[a] = INFINITY; nd->hi[a] = -INFINITY; }
    double clo[3] = { INFINITY, INFINITY, INFINITY }, chi[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (size_t i = start; i < end; ++i) {
        double lo[3], hi[3];
        sphere_box(&s[idx[i]], lo, hi);
        for (int a = 0; a < 3; ++a) {
            nd->lo[a] = fmin(nd->lo[a], lo[a]); nd->hi[a] = fmax(nd->hi[a], hi[a]);
            double c = lo[a] + hi[a];
            clo[a] = fmin(clo[a], c); chi[a] = fmax(chi[a], c);
        }
    }
    if (end - start == 1) {
        nd->left = -1 - idx[start];
        nd->right = 0;
        return me;
    }
    int axis = 0;
    if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
    if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
    g_sort_ctx.s = s; g_sort_ctx.axis = axis;
    qsort(idx + start, end - start, sizeof(int), cmp_centroid);
    size_t mid = start + (end - start) / 2;
    int32_t l = bvh_build_rec(b, s, idx, start, mid);
    int32_t r = bvh_build_rec(b, s, idx, mid, end);
    b->nodes[me].left = l;       /* nodes[] never reallocates: capacity 2n */
    b->nodes[me].right = r;
    return me;
}

o_bvh *o_bvh_build(const o_sphere *s, size_t n) {
    if (n == 0) return NULL;
    o_bvh *b = (o_bvh *)calloc(1, sizeof *b);
    b->cap = 2 * n;
    b->nodes = (o_bvh_node *)calloc(b->cap, sizeof(o_bvh_node));
    int *idx = (int *)malloc(n * sizeof(int));
    for (size_t i = 0; i < n; ++i) idx[i] = (int)i;
    bvh_build_rec(b, s, idx, 0, n);
    free(idx);
    return b;
}

void o_bvh_free(o_bvh *b) {
    if (!b) return;
    free(b->nodes);
    free(b);
}

/* src/bvh.h:16-24: box test, then left, then right with t_max narrowed to the
 * left hit.  Exact-t ties are resolved to the lowest list index so that the
 * answer equals o_list_hit (the reference's ground truth) on ties as well. */
static int bvh_hit_rec(const o_bvh *b, int32_t ni, const o_sphere *s, const o_ray *r,
                       double tmin, double tmax, o_hit *rec) {
    const o_bvh_node *nd = &b->nodes[ni];
    if (!o_aabb_hit(nd->lo, nd->hi, r, tmin, tmax)) return 0;
    if (nd->left < 0) {
        int si = -1 - nd->left;
        o_hit temp;
        /* nextafter: keep an equal-t candidate alive so the tie rule can see it */
        if (!o_sphere_hit(&s[si], r, tmin, nextafter(tmax, INFINITY), &temp)) return 0;
        if (temp.t > tmax) return 0;
        if (temp.t == tmax && !(rec->prim >= 0 && rec->t == tmax && si < rec->prim)) return 0;
        temp.prim = si;
        *rec = temp;
        return 1;
    }
    int hit_left = bvh_hit_rec(b, nd->left, s, r, tmin, tmax, rec);
    int hit_right = bvh_hit_rec(b, nd->right, s, r, tmin, hit_left ? rec->t : tmax, rec);
    return hit_left || hit_right;
}

int o_bvh_hit(const o_bvh *b, const o_sphere *s, const o_ray *r, double tmin, double tmax, o_hit *rec) {
    rec->prim = -1;
    rec->t = tmax;
    if (!b) return 0;
    return bvh_hit_rec(b, 0, s, r, tmin, tmax, rec);
}

/* -------------------------------------------------------------- materials -- */

/* src/material.h:76-81 */
static double reflectance(double cosine, double ref_idx) {
    double r0 = (1 - ref_idx) / (1 + ref_idx);
    r0 = r0 * r0;
    return r0 + (1 - r0) * pow((1 - cosine), 5);
}

/* src/material.h:19-25 (lambertian), :35-41 (metal), :52-71 (dielectric) */
int o_scatter(const o_material *m, const o_ray *r_in, const o_hit *rec,
              o_vec3 *attenuation, o_ray *scattered, o_rng *rng) {
    switch (m->type) {
    case O_LAMBERTIAN: {
        o_vec3 scatter_direction = vadd(rec->normal, random_unit_vector(rng)); /* no near_zero guard (:21) */
        scattered->orig = rec->p; scattered->dir = scatter_direction; scattered->tm = r_in->tm;
        *attenuation = m->albedo;
        return 1;
    }
    case O_METAL: {
        o_vec3 reflected = reflect(vunit(r_in->dir), rec->normal);
        o_vec3 dir = vadd(reflected, vscale(m->param, random_in_unit_sphere(rng))); /* drawn even if fuzz==0 */
        scattered->orig = rec->p; scattered->dir = dir; scattered->tm = r_in->tm;
        *attenuation = m->albedo;
        return vdot(scattered->dir, rec->normal) > 0;
    }
    default: {
        *attenuation = v3(1.0, 1.0, 1.0);
        double ir = m->param;
        double refraction_ratio = rec->front_face ? (1.0 / ir) : ir;

        o_vec3 unit_direction = vunit(r_in->dir);
        double cos_theta = fmin(vdot(vneg(unit_direction), rec->normal), 1.0);
        double sin_theta = sqrt(1.0 - cos_theta * cos_theta);

        int cannot_refract = refraction_ratio * sin_theta > 1.0;
        o_vec3 direction;
        /* `||` short-circuit: the uniform is drawn only when refraction is possible (:64) */
        if (cannot_refract || reflectance(cos_theta, refraction_ratio) > o_random_double(rng))
            direction = reflect(unit_direction, rec->normal);
        else
            direction = refract(unit_direction, rec->normal, refraction_ratio);

        scattered->orig = rec->p; scattered->dir = direction; scattered->tm = r_in->tm;
        return 1;
    }
    }
}

/* ------------------------------------------------------------- integrator -- */

static int world_hit(const o_world *w, const o_ray *r, double tmin, double tmax, o_hit *rec) {
    if (w->bvh) return o_bvh_hit(w->bvh, w->spheres, r, tmin, tmax, rec);
    return o_list_hit(w->spheres, w->n_spheres, r, tmin, tmax, rec);
}

/* src/camera_cpu.h:8-26, recursive exactly as written (right-nested products).
 * Keyed-stream contract: the scatter at the b-th hit of a path draws from
 * stage b = max_depth - depth + 1. */
o_vec3 o_ray_color(const o_ray *r, int depth, const o_world *w, o_rng *rng,
                   int max_depth, uint64_t *rays_out) {
    o_hit rec;

    if (depth <= 0) return v3(0, 0, 0);
    if (rays_out) ++*rays_out;
    o_rng_set_stream(rng, rng->pixel, rng->sample, (uint32_t)(max_depth - depth + 1));

    if (world_hit(w, r, 0.001, INFINITY, &rec)) {
        o_ray scattered;
        o_vec3 attenuation;
        const o_material *m = &w->materials[w->spheres[rec.prim].mat];
        if (o_scatter(m, r, &rec, &attenuation, &scattered, rng))
            return vmul(attenuation, o_ray_color(&scattered, depth - 1, w, rng, max_depth, rays_out));
        return v3(0, 0, 0);
    }

    o_vec3 unit_direction = vunit(r->dir);
    double a = 0.5 * (unit_direction.y + 1.0);
    return vadd(vscale(1.0 - a, v3(1.0, 1.0, 1.0)), vscale(a, v3(0.5, 0.7, 1.0)));
}

/* src/color.h:9-12,14-35 + src/interval.h:14-18 */
void o_write_color(o_vec3 pixel_sum, int spp, uint8_t rgb[3]) {
    double c[3] = { pixel_sum.x, pixel_sum.y, pixel_sum.z };
    double scale = 1.0 / spp;
    for (int k = 0; k < 3; ++k) {
        double v = c[k] * scale;
        v = sqrt(v);
        if (v < 0.000) v = 0.000;
        if (v > 0.999) v = 0.999;
        rgb[k] = (uint8_t)(int)(256 * v);
    }
}

typedef struct {
    const o_world *w;
    const o_camera *cam;
    const o_render_opts *opt;
    double *sum;
    uint8_t *rgb8;
    int tid, nthreads, row0, row1;
    uint64_t rays;
} render_job;

/* src/camera.h:37-47: rows top to bottom, pixels left to right, samples in
 * order, double accumulation in sample order. */
static void *render_rows(void *arg) {
    render_job *jb = (render_job *)arg;
    const o_camera *cam = jb->cam;
    const o_render_opts *opt = jb->opt;
    o_rng rng;
    if (opt->rng_mode == O_RNG_MT19937) {
        o_rng_init_mt(&rng, (uint32_t)opt->seed);
        for (uint64_t k = 0; k < opt->mt_skip; ++k) (void)o_random_double(&rng);
    } else {
        o_rng_init_philox(&rng, opt->seed);
    }
    const int W = cam->image_width;
    for (int j = jb->row0 + jb->tid; j < jb->row1; j += jb->nthreads) {
        for (int i = 0; i < W; ++i) {
            o_vec3 pixel_color = v3(0, 0, 0);
            uint32_t pixel = (uint32_t)j * (uint32_t)W + (uint32_t)i;
            for (int s = opt->sample0; s < opt->sample0 + opt->nsamples; ++s) {
                o_rng_set_stream(&rng, pixel, (uint32_t)s, 0);
                o_ray r = o_get_ray(cam, i, j, &rng);
                o_vec3 c = o_ray_color(&r, cam->max_depth, jb->w, &rng, cam->max_depth, &jb->rays);
                pixel_color = vadd(pixel_color, c);
            }
            size_t o = ((size_t)j * W + i) * 3;
            if (jb->sum) { jb->sum[o] = pixel_color.x; jb->sum[o + 1] = pixel_color.y; jb->sum[o + 2] = pixel_color.z; }
            if (jb->rgb8) o_write_color(pixel_color, opt->nsamples, jb->rgb8 + o);
        }
    }
    return NULL;
}

int o_render(const o_world *w, o_camera *cam, const o_render_opts *opt,
             double *sum, uint8_t *rgb8, uint64_t *rays_out) {
    o_camera_initialize(cam);
    int nthreads = opt->nthreads < 1 ? 1 : opt->nthreads;
    if (opt->rng_mode == O_RNG_MT19937) nthreads = 1;   /* one sequential stream */
    if (nthreads > 256) nthreads = 256;
    int row0 = opt->row0 < 0 ? 0 : opt->row0;
    int row1 = (opt->row1 <= 0 || opt->row1 > cam->image_height) ? cam->image_height : opt->row1;
    render_job jobs[256];
    pthread_t th[256];
    for (int t = 0; t < nthreads; ++t) {
        render_job jb = { w, cam, opt, sum, rgb8, t, nthreads, row0, row1, 0 };
        jobs[t] = jb;
    }
    if (nthreads == 1) {
        render_rows(&jobs[0]);
    } else {
        for (int t = 0; t < nthreads; ++t) pthread_create(&th[t], NULL, render_rows, &jobs[t]);
        for (int t = 0; t < nthreads; ++t) pthread_join(th[t], NULL);
    }
    uint64_t rays = 0;
    for (int t = 0; t < nthreads; ++t) rays += jobs[t].rays;
    if (rays_out) *rays_out = rays;
    return 0;
}

/* src/camera.h:35 header + src/color.h:32-34 body ("r g b\n"). */
size_t o_ppm_p3(const uint8_t *rgb8, int w, int h, char *buf, size_t cap) {
    char line[64];
    size_t n = 0;
    int len = snprintf(line, sizeof line, "P3\n%d %d\n255\n", w, h);
    if (buf && n + (size_t)len < cap) memcpy(buf + n, line, (size_t)len);
    n += (size_t)len;
    for (size_t p = 0; p < (size_t)w * (size_t)h; ++p) {
        len = snprintf(line, sizeof line, "%d %d %d\n", rgb8[3 * p], rgb8[3 * p + 1], rgb8[3 * p + 2]);
        if (buf && n + (size_t)len < cap) memcpy(buf + n, line, (size_t)len);
        n += (size_t)len;
    }
    if (buf && n < cap) buf[n] = 0;
    return n;
}

/* ------------------------------------------------------------------ scene -- */

/* src/main.cpp:11-53 with the g++ draw order:
 *   choose_mat; centre: FIRST draw -> z term, second -> x term (:20);
 *   diffuse: RIGHT operand of color::random()*color::random() first, each
 *   vec3::random() filling z,y,x (:27); then center2.y (:29);
 *   metal: color::random(.5,1) (z,y,x) then fuzz (:33-34); glass: none. */
size_t o_scene_rtiow(int g, double p_diffuse, double p_metal, uint32_t mt_seed,
                     o_sphere *s, o_material *m, size_t cap, uint64_t *draws_out) {
    o_rng rng;
    o_rng_init_mt(&rng, mt_seed);
    size_t n = 0;
#define PUSH(C1, CV, R, MOV, TYPE, ALB, PAR)                                        \
    do {                                                                            \
        if (n < cap) {                                                              \
            s[n].center1 = (C1); s[n].center_vec = (CV); s[n].radius = (R);         \
            s[n].is_moving = (MOV); s[n].mat = (int32_t)n;                          \
            m[n].type = (TYPE); m[n].pad = 0; m[n].albedo = (ALB); m[n].param = (PAR); \
        }                                                                           \
        ++n;                                                                        \
    } while (0)

    PUSH(v3(0, -1000, 0), v3(0, 0, 0), 1000.0, 0, O_LAMBERTIAN, v3(0.5, 0.5, 0.5), 0.0);

    for (int a = -g; a < g; a++) {
        for (int b = -g; b < g; b++) {
            double choose_mat = o_random_double(&rng);
            double zz = b + 0.9 * o_random_double(&rng);
            double xx = a + 0.9 * o_random_double(&rng);
            o_vec3 center = v3(xx, 0.2, zz);

            if (vlen(vsub(center, v3(4, 0.2, 0))) > 0.9) {
                if (choose_mat < p_diffuse) {
                    o_vec3 right = vec3_random(&rng);
                    o_vec3 left = vec3_random(&rng);
                    o_vec3 albedo = vmul(left, right);
                    o_vec3 center2 = vadd(center, v3(0, random_range(&rng, 0, .5), 0));
                    PUSH(center, vsub(center2, center), 0.2, 1, O_LAMBERTIAN, albedo, 0.0);
                } else if (choose_mat < p_metal) {
                    o_vec3 albedo = vec3_random_range(&rng, 0.5, 1);
                    double fuzz = random_range(&rng, 0, 0.5);
                    PUSH(center, v3(0, 0, 0), 0.2, 0, O_METAL, albedo, fuzz < 1 ? fuzz : 1); /* src/material.h:33 */
                } else {
                    PUSH(center, v3(0, 0, 0), 0.2, 0, O_DIELECTRIC, v3(0, 0, 0), 1.5);
                }
            }
        }
    }

    PUSH(v3(0, 1, 0), v3(0, 0, 0), 1.0, 0, O_DIELECTRIC, v3(0, 0, 0), 1.5);
    PUSH(v3(-4, 1, 0), v3(0, 0, 0), 1.0, 0, O_LAMBERTIAN, v3(0.4, 0.2, 0.1), 0.0);
    PUSH(v3(4, 1, 0), v3(0, 0, 0), 1.0, 0, O_METAL, v3(0.7, 0.6, 0.5), 0.0);
#undef PUSH
    if (draws_out) *draws_out = rng.draws;
    return n;
}

/* src/main.cpp:57-68 */
void o_camera_rtiow(o_camera *c, int image_width, int spp, int max_depth) {
    memset(c, 0, sizeof *c);
    c->aspect_ratio = 16.0 / 9.0;
    c->image_width = image_width;
    c->samples_per_pixel = spp;
    c->max_depth = max_depth;
    c->vfov = 20;
    c->lookfrom = v3(13, 2, 3);
    c->lookat = v3(0, 0, 0);
    c->vup = v3(0, 1, 0);
    c->defocus_angle = 0.6;
    c->focus_dist = 10.0;
}

/* --------------------------------------------- batch entry points (tests) -- */
/* Same conventions as the ref_* functions of ref_harness.cpp. */

void ob_closest_hit(const o_world *w, const o_ray *rays, size_t n, double tmin, double tmax, o_hit *out) {
    for (size_t i = 0; i < n; ++i) {
        o_hit rec;
        memset(&rec, 0, sizeof rec);
        if (!world_hit(w, &rays[i], tmin, tmax, &rec)) { memset(&rec, 0, sizeof rec); rec.prim = -1; }
        out[i] = rec;
    }
}

/* element i draws from stream (pixel = i, sample = 0, stage = 1) */
void ob_scatter(const o_material *m, const o_ray *rin, const o_hit *rec, size_t n, uint64_t seed,
                int32_t *scattered_flag, o_vec3 *attenuation, o_ray *scattered) {
    o_rng rng;
    o_rng_init_philox(&rng, seed);
    for (size_t i = 0; i < n; ++i) {
        o_rng_set_stream(&rng, (uint32_t)i, 0, 1);
        attenuation[i] = v3(0, 0, 0);
        memset(&scattered[i], 0, sizeof scattered[i]);
        scattered_flag[i] = o_scatter(&m[i], &rin[i], &rec[i], &attenuation[i], &scattered[i], &rng);
    }
}

void ob_get_rays(o_camera *c, const int32_t *px_i, const int32_t *px_j, const uint32_t *sample,
                 size_t n, uint64_t seed, o_ray *out) {
    o_rng rng;
    o_camera_initialize(c);
    o_rng_init_philox(&rng, seed);
    for (size_t k = 0; k < n; ++k) {
        o_rng_set_stream(&rng, (uint32_t)px_j[k] * (uint32_t)c->image_width + (uint32_t)px_i[k], sample[k], 0);
        out[k] = o_get_ray(c, px_i[k], px_j[k], &rng);
    }
}

/* tests/tests.cpp:11-45 */
void ob_pixel_match(double out[3]) {
    o_rng rng;
    o_rng_init_mt(&rng, 5489u);
    o_camera cam;
    o_camera_rtiow(&cam, 400, 30, 50);
    o_camera_initialize(&cam);
    o_sphere ground = { {0, -1000, 0}, {0, 0, 0}, 1000.0, 0, 0 };
    o_material gm = { O_LAMBERTIAN, 0, {0.5, 0.5, 0.5}, 0.0 };
    o_world w = { &ground, &gm, 1, NULL };
    o_ray r = o_get_ray(&cam, cam.image_width / 2, cam.image_height / 2, &rng);
    o_vec3 c = o_ray_color(&r, cam.max_depth, &w, &rng, cam.max_depth, NULL);
    out[0] = c.x; out[1] = c.y; out[2] = c.z;
}

void ob_write_color(const double *sums, size_t n, int spp, uint8_t *rgb) {
    for (size_t i = 0; i < n; ++i) o_write_color(v3(sums[3 * i], sums[3 * i + 1], sums[3 * i + 2]), spp, rgb + 3 * i);
}

/* ---- the windowed demo's compute shader -------------------------------------------
 * assets/shaders/ray_tracing.comp:35-73, float arithmetic, one rounding per operation.  The shader
 * hard-codes one sphere (:49-52); here the `if (hit) pixel_color = sphere.color` statement (:61-63) runs
 * once per sphere of a list laid out as src/common_objects.h:17-22, in list order (the last sphere hit
 * gives the colour -- the shader has no depth test).  Pinned against the shader itself compiled as
 * C++ (oracle/_ref/libref_glsl.so) in tests/test_demo_shader.py. */
static float demo_dot(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

int o_demo_hit(const float origin[3], const float direction[3], const float center[3], float radius) {   /* :65-73 */
    float oc[3] = { origin[0] - center[0], origin[1] - center[1], origin[2] - center[2] };
    float a = demo_dot(direction, direction);
    float b = 2.0f * demo_dot(direction, oc);
    float c = demo_dot(oc, oc) - radius * radius;
    float discriminant = b * b - 4.0f * a * c;
    return discriminant > 0;
}

static uint8_t demo_unorm8(float v) {
    v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    return (uint8_t)nearbyintf(v * 255.0f);
}

void o_demo_compute(const o_demo_sphere *spheres, size_t n, int width, int height, uint8_t *rgba8) {
    static const o_demo_sphere hard_coded = { {3.0f, 0.0f, 0.0f}, 1.0f, {1.0f, 0.3f, 0.7f, 1.0f} };        /* :49-52 */
    const float position[3] = {0.0f, 0.0f, 0.0f}, forwards[3] = {0.0f, 0.0f, 1.0f}, right[3] = {1.0f, 0.0f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    if (!spheres) { spheres = &hard_coded; n = 1; }
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            float hc = ((float)x * 2 - width) / width;                 /* :38 */
            float vc = ((float)y * 2 - height) / width;                /* :39 (divided by the width too) */
            float color[4] = {0.5f, 0.0f, 0.25f, 1.0f};                /* :41 */
            float dir[3];
            for (int k = 0; k < 3; ++k) dir[k] = forwards[k] + hc * right[k] + vc * up[k];       /* :57 */
            for (size_t s = 0; s < n; ++s)
                if (o_demo_hit(position, dir, spheres[s].center, spheres[s].radius))
                    for (int k = 0; k < 4; ++k) color[k] = spheres[s].color[k];
            uint8_t *t = rgba8 + ((size_t)y * width + x) * 4;
            for (int k = 0; k < 4; ++k) t[k] = demo_unorm8(color[k]);
        }
}
