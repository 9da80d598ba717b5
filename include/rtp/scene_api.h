// scene_api.h -- host-side mirror of the reference's scene-description API (C++17, header only).
//
// ShuvalovIvan/RayTracingProject describes a scene with vec3 / ray / interval / aabb / hittable /
// hittable_list / sphere / material objects (src/vec3.h, src/ray.h, src/interval.h, src/aabb.h,
// src/hittable.h, src/hittable_list.h, src/sphere.h, src/material.h) and renders it on the CPU through
// their virtual hit() / scatter().  Here the SAME names, constructors and public members only DESCRIBE
// the scene; tracing happens on the B200 behind the C ABI (include/rt_b200.h).  What the reference
// keeps private (sphere geometry src/sphere.h:60-66, material parameters src/material.h:27-28,43-45,
// 73-74) is exported through one virtual per hierarchy -- hittable::describe / material::describe --
// into the POD arrays rt_upload_scene takes.  There is deliberately NO hit() and NO scatter() here:
// the product has no CPU render path (the CPU oracle lives in oracle/, test infrastructure only).
//
// The small same-named headers next to this file (vec3.h, sphere.h, ...) just include it, so a
// caller written against the reference's headers compiles unchanged with -Iinclude/rtp.
#ifndef RTP_SCENE_API_H
#define RTP_SCENE_API_H

#include <cmath>
#include <cstdint>
#include <iostream>
#include <limits>
#include <memory>
#include <random>
#include <stdexcept>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../rt_b200.h"

using std::make_shared;
using std::shared_ptr;
using std::sqrt;

// ---- src/rtweekend.h:16-34 ------------------------------------------------------------------
const double infinity = std::numeric_limits<double>::infinity();
const double pi = 3.1415926535897932385;

inline double degrees_to_radians(double degrees) { return degrees * pi / 180.0; }

// Host-side scene generation keeps the reference's generator: ONE default-seeded std::mt19937 read
// through uniform_real_distribution<double>(0,1), so `main.cpp` builds the same 485 spheres.  (The
// device never sees this stream: rendering draws from the keyed Philox stream, DESIGN.md.)
namespace rtp_detail {
inline std::mt19937& host_engine() { static std::mt19937 engine; return engine; }
}  // namespace rtp_detail

inline double random_double() {
    static std::uniform_real_distribution<double> unit(0.0, 1.0);
    return unit(rtp_detail::host_engine());
}
inline double random_double(double lo, double hi) { return lo + (hi - lo) * random_double(); }

// ---- src/vec3.h -------------------------------------------------------------------------------
class vec3 {
  public:
    double e[3];

    vec3() : e{0, 0, 0} {}
    vec3(double x, double y, double z) : e{x, y, z} {}

    double x() const { return e[0]; }
    double y() const { return e[1]; }
    double z() const { return e[2]; }

    vec3 operator-() const { return vec3(-e[0], -e[1], -e[2]); }
    double operator[](int i) const { return e[i]; }
    double& operator[](int i) { return e[i]; }

    vec3& operator+=(const vec3& o) { for (int k = 0; k < 3; ++k) e[k] += o.e[k]; return *this; }
    vec3& operator*=(double s) { for (int k = 0; k < 3; ++k) e[k] *= s; return *this; }
    vec3& operator/=(double s) { return *this *= 1 / s; }

    double length_squared() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    double length() const { return std::sqrt(length_squared()); }
    bool near_zero() const { const double s = 1e-8; return std::fabs(e[0]) < s && std::fabs(e[1]) < s && std::fabs(e[2]) < s; }
    // tolerance helper of the reference's unit test (src/vec3.h:56-61): every component within 1e-3
    bool similar_to(const vec3& o) const {
        for (int k = 0; k < 3; ++k) if (!(std::fabs(e[k] - o.e[k]) < 1e-3)) return false;
        return true;
    }

    // The reference writes vec3(random_double(), random_double(), random_double()); g++ evaluates the
    // arguments right to left, so the FIRST draw lands in z.  Sequenced explicitly here
    // so that any compiler reproduces the golden scene.
    static vec3 random() { const double z = random_double(), y = random_double(), x = random_double(); return vec3(x, y, z); }
    static vec3 random(double lo, double hi) {
        const double z = random_double(lo, hi), y = random_double(lo, hi), x = random_double(lo, hi);
        return vec3(x, y, z);
    }
};

using point3 = vec3;
using color = vec3;

inline std::ostream& operator<<(std::ostream& out, const vec3& v) { return out << v.e[0] << ' ' << v.e[1] << ' ' << v.e[2]; }
inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]); }
inline vec3 operator*(double s, const vec3& v) { return vec3(s * v.e[0], s * v.e[1], s * v.e[2]); }
inline vec3 operator*(const vec3& v, double s) { return s * v; }
inline vec3 operator/(vec3 v, double s) { return (1 / s) * v; }
inline double dot(const vec3& a, const vec3& b) { return a.e[0] * b.e[0] + a.e[1] * b.e[1] + a.e[2] * b.e[2]; }
inline vec3 cross(const vec3& a, const vec3& b) {
    return vec3(a.e[1] * b.e[2] - a.e[2] * b.e[1], a.e[2] * b.e[0] - a.e[0] * b.e[2], a.e[0] * b.e[1] - a.e[1] * b.e[0]);
}
inline vec3 unit_vector(vec3 v) { return v / v.length(); }
inline vec3 reflect(const vec3& v, const vec3& n) { return v - 2 * dot(v, n) * n; }
inline vec3 refract(const vec3& uv, const vec3& n, double etai_over_etat) {
    const double cos_theta = std::fmin(dot(-uv, n), 1.0);
    const vec3 perp = etai_over_etat * (uv + cos_theta * n);
    return perp + (-std::sqrt(std::fabs(1.0 - perp.length_squared()))) * n;
}
// host-side samplers (scene authoring only; the render path samples on the device)
inline vec3 random_in_unit_disk() {
    for (;;) { const double y = random_double(-1, 1), x = random_double(-1, 1); const vec3 p(x, y, 0); if (p.length_squared() < 1) return p; }
}
inline vec3 random_in_unit_sphere() {
    for (;;) { const vec3 p = vec3::random(-1, 1); if (p.length_squared() < 1) return p; }
}
inline vec3 random_unit_vector() { return unit_vector(random_in_unit_sphere()); }
inline vec3 random_on_hemisphere(const vec3& normal) { const vec3 u = random_unit_vector(); return dot(u, normal) > 0.0 ? u : -u; }

// ---- src/ray.h ----------------------------------------------------------------------------------
class ray {
  public:
    ray() {}
    ray(const point3& origin, const vec3& direction, double time = 0.0) : orig(origin), dir(direction), tm(time) {}
    point3 origin() const { return orig; }
    vec3 direction() const { return dir; }
    double time() const { return tm; }
    point3 at(double t) const { return orig + t * dir; }

  private:
    point3 orig;
    vec3 dir;
    double tm = 0;
};

// ---- src/interval.h -------------------------------------------------------------------------------
class interval {
  public:
    double min, max;
    interval() : min(+infinity), max(-infinity) {}
    interval(double lo, double hi) : min(lo), max(hi) {}
    interval(const interval& a, const interval& b) : min(std::fmin(a.min, b.min)), max(std::fmax(a.max, b.max)) {}
    double size() const { return max - min; }
    interval expand(double delta) const { return interval(min - delta / 2, max + delta / 2); }
    bool contains(double x) const { return min <= x && x <= max; }
    bool surrounds(double x) const { return min < x && x < max; }
    double clamp(double x) const { return x < min ? min : (x > max ? max : x); }
    static const interval empty, universe;
};

// ---- src/color.h:9-35 ------------------------------------------------------------------------------
// linear_to_gamma and write_color for CALLER code that formats a colour of its own (a legend, a test value).  The frame
// itself never passes through them: B200Impl::Camera::render takes its bytes from the device (rt_readback: resolve + P3
// encode kernels, byte-identical to this text); these two exist so that code written against the reference's color.h compiles.
inline double linear_to_gamma(double linear_component) { return std::sqrt(linear_component); }
inline void write_color(std::ostream& out, color pixel_color, int samples_per_pixel) {
    const double scale = 1.0 / samples_per_pixel;                       // :20-23
    const double r = linear_to_gamma(pixel_color.x() * scale), g = linear_to_gamma(pixel_color.y() * scale), b = linear_to_gamma(pixel_color.z() * scale);
    static const interval intensity(0.000, 0.999);                      // :31
    out << static_cast<int>(256 * intensity.clamp(r)) << ' ' << static_cast<int>(256 * intensity.clamp(g)) << ' '
        << static_cast<int>(256 * intensity.clamp(b)) << '\n';          // :32-34
}
inline const interval interval::empty(+infinity, -infinity);
inline const interval interval::universe(-infinity, +infinity);

// ---- src/aabb.h (construction only; the slab test runs on the device, rt_trace.cuh) ----------------
class aabb {
  public:
    interval x, y, z;
    aabb() {}
    aabb(const interval& ix, const interval& iy, const interval& iz) : x(ix), y(iy), z(iz) {}
    aabb(const point3& a, const point3& b)
        : x(std::fmin(a[0], b[0]), std::fmax(a[0], b[0])), y(std::fmin(a[1], b[1]), std::fmax(a[1], b[1])),
          z(std::fmin(a[2], b[2]), std::fmax(a[2], b[2])) {}
    aabb(const aabb& a, const aabb& b) : x(a.x, b.x), y(a.y, b.y), z(a.z, b.z) {}
    const interval& axis(int n) const { return n == 1 ? y : (n == 2 ? z : x); }
};

// ---- scene export: what rt_upload_scene takes (include/rt_b200.h) -----------------------------------
// Filled by hittable::describe / material::describe.  A material shared by several spheres is
// exported once (keyed by address), like the shared_ptr it replaces.
struct rt_scene_builder {
    std::vector<rt_sphere> spheres;
    std::vector<rt_material> materials;
    std::unordered_map<const void*, int32_t> material_ids;
};

// ---- src/material.h -----------------------------------------------------------------------------------
class hit_record;

class material {
  public:
    virtual ~material() = default;
    // index of this material in b.materials (appended on first use)
    int32_t describe(rt_scene_builder& b) const {
        auto it = b.material_ids.find(this);
        if (it != b.material_ids.end()) return it->second;
        const int32_t id = int32_t(b.materials.size());
        b.materials.push_back(pod());
        b.material_ids.emplace(this, id);
        return id;
    }

  protected:
    virtual rt_material pod() const = 0;
    static rt_material make_pod(int32_t type, const color& albedo, double param) {
        rt_material m;
        m.type = type; m.pad = 0;
        m.albedo[0] = albedo.x(); m.albedo[1] = albedo.y(); m.albedo[2] = albedo.z();
        m.param = param;
        return m;
    }
};

class lambertian : public material {          // src/material.h:15-29
  public:
    lambertian(const color& a) : albedo(a) {}
  protected:
    rt_material pod() const override { return make_pod(RT_LAMBERTIAN, albedo, 0.0); }
  private:
    color albedo;
};

class metal : public material {               // src/material.h:31-46 (fuzz clamped to <= 1 in the ctor)
  public:
    metal(const color& a, double f) : albedo(a), fuzz(f < 1 ? f : 1) {}
  protected:
    rt_material pod() const override { return make_pod(RT_METAL, albedo, fuzz); }
  private:
    color albedo;
    double fuzz;
};

class dielectric : public material {          // src/material.h:48-82
  public:
    dielectric(double index_of_refraction) : ir(index_of_refraction) {}
  protected:
    rt_material pod() const override { return make_pod(RT_DIELECTRIC, color(0, 0, 0), ir); }
  private:
    double ir;
};

// ---- src/hittable.h --------------------------------------------------------------------------------------
// Kept for source compatibility (rt_trace_closest fills the same fields, rt_hit in rt_b200.h).
class hit_record {
  public:
    point3 p;
    vec3 normal;
    shared_ptr<material> mat;
    double t = 0;
    bool front_face = false;
    void set_face_normal(const ray& r, const vec3& outward_normal) {
        front_face = dot(r.direction(), outward_normal) < 0;
        normal = front_face ? outward_normal : -outward_normal;
    }
};

class hittable {
  public:
    virtual ~hittable() = default;
    virtual aabb bounding_box() const = 0;
    // appends this object's primitives (and their materials) to the export arrays
    virtual void describe(rt_scene_builder& b) const = 0;
};

// ---- src/sphere.h --------------------------------------------------------------------------------------
class sphere : public hittable {
  public:
    // stationary (src/sphere.h:9-14)
    sphere(point3 center, double r, shared_ptr<material> m) : center1(center), radius(r), mat(std::move(m)), is_moving(false) {
        const vec3 rvec(radius, radius, radius);
        bbox = aabb(center1 - rvec, center1 + rvec);
    }
    // moving linearly from c1 at time 0 to c2 at time 1 (src/sphere.h:19-28)
    sphere(point3 c1, point3 c2, double r, shared_ptr<material> m) : center1(c1), radius(r), mat(std::move(m)), is_moving(true) {
        const vec3 rvec(radius, radius, radius);
        bbox = aabb(aabb(c1 - rvec, c1 + rvec), aabb(c2 - rvec, c2 + rvec));
        center_vec = c2 - c1;
    }
    aabb bounding_box() const override { return bbox; }
    void describe(rt_scene_builder& b) const override {
        if (!mat) throw std::runtime_error("sphere without a material cannot be rendered");
        rt_sphere s;
        for (int k = 0; k < 3; ++k) { s.center1[k] = center1[k]; s.center_vec[k] = is_moving ? center_vec[k] : 0.0; }
        s.radius = radius;
        s.is_moving = is_moving ? 1 : 0;
        s.material = mat->describe(b);
        b.spheres.push_back(s);
    }

  private:
    point3 center1;
    double radius;
    shared_ptr<material> mat;
    bool is_moving;
    vec3 center_vec;
    aabb bbox;
};

// ---- src/hittable_list.h -----------------------------------------------------------------------------------
class hittable_list : public hittable {
  public:
    std::vector<shared_ptr<hittable>> objects;

    hittable_list() {}
    hittable_list(shared_ptr<hittable> object) { add(std::move(object)); }
    void clear() { objects.clear(); bbox = aabb(); }
    void add(shared_ptr<hittable> object) {
        bbox = aabb(bbox, object->bounding_box());
        objects.push_back(std::move(object));
    }
    aabb bounding_box() const override { return bbox; }
    // list order is kept: it is the tie-break of the closest-hit search (src/hittable_list.h:31)
    void describe(rt_scene_builder& b) const override { for (const auto& o : objects) o->describe(b); }

  private:
    aabb bbox;
};

// ---- src/bvh.h ------------------------------------------------------------------------------------------------
// The book's idiom `world = hittable_list(make_shared<bvh_node>(world));` keeps working: the node only remembers the objects
// (in list order, the tie-break of the closest-hit search) and their box; the tree that is walked is the one rt_upload_scene
// builds on the device side (the reference's constructor, src/bvh.h:12-14, is an empty stub).
class bvh_node : public hittable {
  public:
    bvh_node(const hittable_list& list) : bvh_node(list.objects, 0, list.objects.size()) {}
    bvh_node(const std::vector<shared_ptr<hittable>>& src_objects, size_t start, size_t end) {
        if (start > end || end > src_objects.size()) throw std::out_of_range("bvh_node: object range outside the list");
        for (size_t i = start; i < end; ++i) {
            bbox = aabb(bbox, src_objects[i]->bounding_box());
            members.push_back(src_objects[i]);
        }
    }
    aabb bounding_box() const override { return bbox; }
    void describe(rt_scene_builder& b) const override { for (const auto& o : members) o->describe(b); }

  private:
    std::vector<shared_ptr<hittable>> members;
    aabb bbox;
};

#endif  // RTP_SCENE_API_H
