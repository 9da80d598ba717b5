// ref_harness.cpp -- ORACLE SIDE (test infrastructure, NOT product code).
//
// Compiles the REFERENCE ITSELF (ShuvalovIvan/RayTracingProject) from the headers
// where they lie under /root/reference/src -- nothing is copied or patched on
// disk -- and exposes it through a small C interface so the C restatement
// (rt_oracle.c) and the CUDA path can be checked against the real thing.
//
// How the two edits the parity harness needs are made without touching the
// reference sources:
//   1. random_double() (src/rtweekend.h:25-29) is redirected by renaming, for the
//      duration of the reference includes, the two std types it instantiates:
//        std::mt19937                  -> std::refhook_engine
//        std::uniform_real_distribution-> std::refhook_dist
//      refhook_dist::operator() serves either a REAL std::mt19937 +
//      std::uniform_real_distribution<double> pair (mode MT: reproduces the golden
//      image bit for bit) or the keyed Philox4x32-10 stream (mode PHILOX).
//   2. src/bvh.h:17 names an undeclared member `box` (the member is `bbox`) and its
//      constructor is a stub (:12-14): `#define box bbox` around that include makes
//      it compile, `#define private public` lets this harness fill left/right/bbox
//      with a median-split build.  Parity is defined on closest hit, not the tree.
//
// Built by oracle/Makefile into oracle/_ref/libref_harness.so (git-ignored).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <limits>
#include <memory>
#include <random>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

// ----------------------------------------------------------------- rng hook --
namespace refhook {

inline void philox4x32_10(const uint32_t c_in[4], const uint32_t k_in[2], uint32_t out[4]) {
    uint32_t c[4] = { c_in[0], c_in[1], c_in[2], c_in[3] };
    uint32_t k[2] = { k_in[0], k_in[1] };
    for (int r = 0; r < 10; ++r) {
        const uint64_t a = 0xD2511F53ull * c[0];
        const uint64_t b = 0xCD9E8D57ull * c[2];
        const uint32_t t0 = uint32_t(b >> 32) ^ c[1] ^ k[0];
        const uint32_t t2 = uint32_t(a >> 32) ^ c[3] ^ k[1];
        c[0] = t0; c[1] = uint32_t(b); c[2] = t2; c[3] = uint32_t(a);
        k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
    }
    for (int i = 0; i < 4; ++i) out[i] = c[i];
}

struct stream {
    bool philox = false;
    // MT mode: the real thing
    std::mt19937 mt;
    std::uniform_real_distribution<double> dist{0.0, 1.0};
    // PHILOX mode
    uint32_t key[2] = {0, 0};
    uint32_t pixel = 0, sample = 0, stage = 0, k = 0;
    uint64_t draws = 0;

    void set(uint32_t px, uint32_t s, uint32_t st) { pixel = px; sample = s; stage = st; k = 0; }
    double next() {
        ++draws;
        if (!philox) return dist(mt);
        const uint32_t ctr[4] = { pixel, sample, stage, k >> 2 };
        uint32_t w[4];
        philox4x32_10(ctr, key, w);
        const uint32_t word = w[k & 3u];
        ++k;
        return double(word >> 8) * (1.0 / 16777216.0);
    }
};

thread_local stream tls;

struct engine {};
template <class T> struct dist {
    dist(T, T) {}
    T operator()(engine&) { return tls.next(); }
};

}  // namespace refhook

namespace std {
using refhook_engine = ::refhook::engine;
template <class T> using refhook_dist = ::refhook::dist<T>;
}

// ------------------------------------------------- the reference, in place --
#define mt19937 refhook_engine
#define uniform_real_distribution refhook_dist
#define private public
#include "rtweekend.h"
#include "camera.h"
#include "color.h"
#include "hittable_list.h"
#include "material.h"
#include "sphere.h"
#define box bbox
#include "bvh.h"
#undef box

// The reference's scene lives inside main() (src/main.cpp:11-71).  Include it
// (it pulls in camera_cpu.h, which has no include guard, exactly once) with main renamed and cam.render(world) intercepted so the harness can either
// let it run (golden image) or capture the world and leave.
namespace refhook {
struct scene_captured {};
thread_local hittable_list* capture_into = nullptr;
inline const hittable& intercept(const hittable_list& w) {
    if (capture_into) { *capture_into = w; throw scene_captured{}; }
    return w;
}
}
// `int main() { ... }` has no return statement; renamed to an ordinary function that
// would be undefined behaviour at -O2, so it is turned into the body of a lambda:
//   int ref_main_decl(); static auto ref_main_body = []() { ... };
#define main ref_main_decl(); static auto ref_main_body = []
#define render(w) render(::refhook::intercept(w))
#include REF_MAIN_CPP
;
#undef render
#undef main
static void ref_main() { ref_main_body(); }
#undef private
#undef mt19937
#undef uniform_real_distribution

#include "rt_oracle.h"   // POD layouts only (o_sphere, o_material, o_ray, o_hit, o_camera)

namespace {

vec3 V(const o_vec3& a) { return vec3(a.x, a.y, a.z); }
o_vec3 O(const vec3& a) { return o_vec3{a.x(), a.y(), a.z()}; }

struct RefWorld {
    hittable_list list;
    std::vector<shared_ptr<material>> mats;             // one per sphere
    std::unordered_map<const material*, int> prim_of;   // rec.mat -> list index
    shared_ptr<bvh_node> bvh;
};

shared_ptr<material> make_material(const o_material& m) {
    switch (m.type) {
        case O_LAMBERTIAN: return make_shared<lambertian>(V(m.albedo));
        case O_METAL:      return make_shared<metal>(V(m.albedo), m.param);
        default:           return make_shared<dielectric>(m.param);
    }
}

// Median split on the longest centroid axis over bounding_box() -- fills the
// reference's own bvh_node (src/bvh.h:28-31) so its own hit() does the traversal.
shared_ptr<hittable> build_bvh(std::vector<shared_ptr<hittable>>& objs, size_t start, size_t end) {
    if (end - start == 1) return objs[start];
    aabb cb;
    aabb bb;
    for (size_t i = start; i < end; ++i) {
        aabb b = objs[i]->bounding_box();
        bb = aabb(bb, b);
        point3 c(b.x.min + b.x.max, b.y.min + b.y.max, b.z.min + b.z.max);
        cb = aabb(cb, aabb(c, c));
    }
    int axis = 0;
    if (cb.y.size() > cb.axis(axis).size()) axis = 1;
    if (cb.z.size() > cb.axis(axis).size()) axis = 2;
    std::stable_sort(objs.begin() + start, objs.begin() + end,
                     [axis](const shared_ptr<hittable>& a, const shared_ptr<hittable>& b) {
                         const interval& ia = a->bounding_box().axis(axis);
                         const interval& ib = b->bounding_box().axis(axis);
                         return ia.min + ia.max < ib.min + ib.max;
                     });
    size_t mid = start + (end - start) / 2;
    auto node = make_shared<bvh_node>(objs, size_t(0), size_t(0));   // stub ctor (src/bvh.h:12-14)
    node->left = build_bvh(objs, start, mid);
    node->right = build_bvh(objs, mid, end);
    node->bbox = bb;
    return node;
}

struct KeyedCamera : CPUImpl::Camera {
    // The recursion at src/camera_cpu.h:19 is a virtual call, so this override is
    // re-entered at every bounce: set the stage key, count the ray, delegate.
    mutable uint64_t rays = 0;
    color ray_color(const ray& r, int depth, const hittable& world) const override {
        if (depth > 0) {
            ++rays;
            refhook::tls.set(refhook::tls.pixel, refhook::tls.sample, uint32_t(max_depth - depth + 1));
        }
        return CPUImpl::Camera::ray_color(r, depth, world);
    }
};

void load_camera(KeyedCamera& cam, const o_camera& c) {
    cam.aspect_ratio = c.aspect_ratio;
    cam.image_width = c.image_width;
    cam.samples_per_pixel = c.samples_per_pixel;
    cam.max_depth = c.max_depth;
    cam.vfov = c.vfov;
    cam.lookfrom = V(c.lookfrom);
    cam.lookat = V(c.lookat);
    cam.vup = V(c.vup);
    cam.defocus_angle = c.defocus_angle;
    cam.focus_dist = c.focus_dist;
}

void init_stream(int mode, uint64_t seed) {
    auto& s = refhook::tls;
    s = refhook::stream();
    s.philox = (mode == O_RNG_PHILOX);
    if (s.philox) { s.key[0] = uint32_t(seed); s.key[1] = uint32_t(seed >> 32); }
    else s.mt.seed(uint32_t(seed));
}

}  // namespace

extern "C" {

// Runs the reference's own main() (src/main.cpp) up to cam.render and exports the
// world it built.  mt_seed 5489 == default-constructed std::mt19937.
size_t ref_scene_main(uint32_t mt_seed, o_sphere* s, o_material* m, size_t cap, uint64_t* draws_out) {
    init_stream(O_RNG_MT19937, mt_seed);
    hittable_list captured;
    refhook::capture_into = &captured;
    try { ref_main(); } catch (const refhook::scene_captured&) {}
    refhook::capture_into = nullptr;
    if (draws_out) *draws_out = refhook::tls.draws;
    size_t n = 0;
    for (const auto& obj : captured.objects) {
        const sphere* sp = dynamic_cast<const sphere*>(obj.get());
        if (!sp) continue;
        if (n < cap) {
            s[n].center1 = O(sp->center1);
            s[n].center_vec = sp->is_moving ? O(sp->center_vec) : o_vec3{0, 0, 0};
            s[n].radius = sp->radius;
            s[n].is_moving = sp->is_moving;
            s[n].mat = int32_t(n);
            m[n].pad = 0;
            if (auto* l = dynamic_cast<const lambertian*>(sp->mat.get())) {
                m[n].type = O_LAMBERTIAN; m[n].albedo = O(l->albedo); m[n].param = 0;
            } else if (auto* mt = dynamic_cast<const metal*>(sp->mat.get())) {
                m[n].type = O_METAL; m[n].albedo = O(mt->albedo); m[n].param = mt->fuzz;
            } else if (auto* d = dynamic_cast<const dielectric*>(sp->mat.get())) {
                m[n].type = O_DIELECTRIC; m[n].albedo = o_vec3{0, 0, 0}; m[n].param = d->ir;
            }
        }
        ++n;
    }
    return n;
}

// The reference's main() unmodified, stdout captured: the golden-image path.
size_t ref_main_ppm(char* buf, size_t cap) {
    init_stream(O_RNG_MT19937, 5489u);
    std::ostringstream out, log;
    auto* old_out = std::cout.rdbuf(out.rdbuf());
    auto* old_log = std::clog.rdbuf(log.rdbuf());
    ref_main();
    std::cout.rdbuf(old_out);
    std::clog.rdbuf(old_log);
    const std::string s = out.str();
    if (buf && cap) { size_t n = std::min(cap - 1, s.size()); memcpy(buf, s.data(), n); buf[n] = 0; }
    return s.size();
}

void* ref_world_create(const o_sphere* s, const o_material* m, size_t n, int with_bvh) {
    auto* w = new RefWorld;
    for (size_t i = 0; i < n; ++i) {
        auto mat = make_material(m[s[i].mat]);
        w->mats.push_back(mat);
        w->prim_of[mat.get()] = int(i);
        if (s[i].is_moving)
            w->list.add(make_shared<sphere>(V(s[i].center1), V(s[i].center1) + V(s[i].center_vec), s[i].radius, mat));
        else
            w->list.add(make_shared<sphere>(V(s[i].center1), s[i].radius, mat));
    }
    if (with_bvh && n > 0) {
        auto objs = w->list.objects;
        auto root = build_bvh(objs, 0, objs.size());
        if (n == 1) {
            w->bvh = make_shared<bvh_node>(objs, size_t(0), size_t(0));
            w->bvh->left = root; w->bvh->right = root; w->bvh->bbox = root->bounding_box();
        } else {
            w->bvh = std::dynamic_pointer_cast<bvh_node>(root);
        }
    }
    return w;
}

void ref_world_destroy(void* h) { delete static_cast<RefWorld*>(h); }

// hittable_list::hit (src/hittable_list.h:25-39) or bvh_node::hit (src/bvh.h:16-24)
// over n rays, interval (tmin, tmax).
void ref_closest_hit(void* h, const o_ray* rays, size_t n, double tmin, double tmax, int use_bvh, o_hit* out) {
    auto* w = static_cast<RefWorld*>(h);
    const hittable& world = (use_bvh && w->bvh) ? static_cast<const hittable&>(*w->bvh)
                                                 : static_cast<const hittable&>(w->list);
    for (size_t i = 0; i < n; ++i) {
        ray r(V(rays[i].orig), V(rays[i].dir), rays[i].tm);
        hit_record rec;
        o_hit& o = out[i];
        if (world.hit(r, interval(tmin, tmax), rec)) {
            o.p = O(rec.p); o.normal = O(rec.normal); o.t = rec.t; o.front_face = rec.front_face;
            o.prim = w->prim_of.at(rec.mat.get());
        } else {
            std::memset(&o, 0, sizeof o);
            o.prim = -1;
        }
    }
}

// material::scatter (src/material.h) for n (ray, hit) pairs; randoms come from the
// keyed stream (seed; pixel = i, sample = 0, stage = 1).
void ref_scatter(const o_material* m, const o_ray* rin, const o_hit* rec, size_t n, uint64_t seed,
                 int32_t* scattered_flag, o_vec3* attenuation, o_ray* scattered) {
    init_stream(O_RNG_PHILOX, seed);
    for (size_t i = 0; i < n; ++i) {
        auto mat = make_material(m[i]);
        hit_record hr;
        hr.p = V(rec[i].p); hr.normal = V(rec[i].normal); hr.t = rec[i].t; hr.front_face = rec[i].front_face != 0;
        hr.mat = mat;
        ray r(V(rin[i].orig), V(rin[i].dir), rin[i].tm);
        refhook::tls.set(uint32_t(i), 0, 1);
        color att(0, 0, 0);
        ray sc(point3(0, 0, 0), vec3(0, 0, 0), 0.0);
        bool ok = mat->scatter(r, hr, att, sc);
        scattered_flag[i] = ok;
        attenuation[i] = O(att);
        scattered[i].orig = O(sc.origin()); scattered[i].dir = O(sc.direction()); scattered[i].tm = sc.time();
    }
}

// camera::initialize + get_ray (src/camera.h:52-113) for pixel/sample lists, keyed
// stream stage 0.  Fills the derived fields of *c as well.
void ref_get_rays(o_camera* c, const int32_t* px_i, const int32_t* px_j, const uint32_t* sample,
                  size_t n, uint64_t seed, o_ray* out) {
    KeyedCamera cam;
    load_camera(cam, *c);
    cam.initialize();
    c->image_height = cam.image_height;
    c->center = O(cam.center); c->pixel00_loc = O(cam.pixel00_loc);
    c->pixel_delta_u = O(cam.pixel_delta_u); c->pixel_delta_v = O(cam.pixel_delta_v);
    c->u = O(cam.u); c->v = O(cam.v); c->w = O(cam.w);
    c->defocus_disk_u = O(cam.defocus_disk_u); c->defocus_disk_v = O(cam.defocus_disk_v);
    init_stream(O_RNG_PHILOX, seed);
    for (size_t k = 0; k < n; ++k) {
        refhook::tls.set(uint32_t(px_j[k]) * uint32_t(cam.image_width) + uint32_t(px_i[k]), sample[k], 0);
        ray r = cam.get_ray(px_i[k], px_j[k]);
        out[k].orig = O(r.origin()); out[k].dir = O(r.direction()); out[k].tm = r.time();
    }
}

// camera::render's triple loop (src/camera.h:37-47) re-driven through the public
// initialize/get_ray/ray_color so the (pixel, sample) key can be set and rows
// spread over threads.  Same signature as o_render.
int ref_render(void* h, o_camera* c, const o_render_opts* opt, int use_bvh,
               double* sum, uint8_t* rgb8, uint64_t* rays_out) {
    auto* w = static_cast<RefWorld*>(h);
    const hittable& world = (use_bvh && w->bvh) ? static_cast<const hittable&>(*w->bvh)
                                                 : static_cast<const hittable&>(w->list);
    int nthreads = std::max(1, opt->nthreads);
    if (opt->rng_mode == O_RNG_MT19937) nthreads = 1;
    KeyedCamera probe;
    load_camera(probe, *c);
    probe.initialize();
    c->image_height = probe.image_height;
    const int W = probe.image_width, H = probe.image_height;
    const int row0 = std::max(0, opt->row0);
    const int row1 = (opt->row1 <= 0 || opt->row1 > H) ? H : opt->row1;
    std::vector<uint64_t> rays(nthreads, 0);
    auto work = [&](int tid) {
        KeyedCamera cam;
        load_camera(cam, *c);
        cam.initialize();
        init_stream(opt->rng_mode, opt->seed);
        if (opt->rng_mode == O_RNG_MT19937)
            for (uint64_t k = 0; k < opt->mt_skip; ++k) (void)random_double();
        for (int j = row0 + tid; j < row1; j += nthreads) {
            for (int i = 0; i < W; ++i) {
                color pixel_color(0, 0, 0);
                for (int s = opt->sample0; s < opt->sample0 + opt->nsamples; ++s) {
                    refhook::tls.set(uint32_t(j) * uint32_t(W) + uint32_t(i), uint32_t(s), 0);
                    ray r = cam.get_ray(i, j);
                    pixel_color += cam.ray_color(r, cam.max_depth, world);
                }
                size_t o = (size_t(j) * W + i) * 3;
                if (sum) { sum[o] = pixel_color.x(); sum[o + 1] = pixel_color.y(); sum[o + 2] = pixel_color.z(); }
                if (rgb8) {
                    std::ostringstream txt;
                    write_color(txt, pixel_color, opt->nsamples);        // src/color.h:14-35
                    int r8, g8, b8;
                    std::istringstream(txt.str()) >> r8 >> g8 >> b8;
                    rgb8[o] = uint8_t(r8); rgb8[o + 1] = uint8_t(g8); rgb8[o + 2] = uint8_t(b8);
                }
            }
        }
        rays[tid] = cam.rays;
    };
    if (nthreads == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
        for (auto& t : th) t.join();
    }
    uint64_t total = 0;
    for (auto r : rays) total += r;
    if (rays_out) *rays_out = total;
    return 0;
}

// The PixelMatch known answer (tests/tests.cpp:11-45): main.cpp camera, ground
// sphere only, fresh mt19937, centre pixel.
void ref_pixel_match(double out[3]) {
    init_stream(O_RNG_MT19937, 5489u);
    KeyedCamera cam;
    cam.aspect_ratio = 16.0 / 9.0; cam.image_width = 400; cam.samples_per_pixel = 30; cam.max_depth = 50;
    cam.vfov = 20; cam.lookfrom = point3(13, 2, 3); cam.lookat = point3(0, 0, 0); cam.vup = vec3(0, 1, 0);
    cam.defocus_angle = 0.6; cam.focus_dist = 10.0;
    hittable_list world;
    world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));
    cam.initialize();
    auto sz = cam.image_size();
    ray r = cam.get_ray(sz.first / 2, sz.second / 2);
    color c = cam.ray_color(r, cam.max_depth, world);
    out[0] = c.x(); out[1] = c.y(); out[2] = c.z();
}

}  // extern "C"
