// rt_device.cuh -- device-side math of the path-tracing hot path, templated on the arithmetic
// type T (float = fast mode, double = parity mode).
//
// Parity mode contract: every function below performs the SAME IEEE operations in the SAME
// order as the reference lines it cites (paths relative to the reference tree), and the
// translation unit that instantiates T = double is compiled with --fmad=false, so results are
// bit-identical to the g++ -ffp-contract=off build of the reference wherever only + - * / sqrt
// are involved.  The one deliberate exception is pow(x,5) (src/material.h:80), evaluated as
// x*x*x*x*x; it only feeds a comparison against a 24-bit uniform.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rt {

// ------------------------------------------------------------------------------- vec3 --
// src/vec3.h:8-119.  Note operator/(v,t) is (1/t)*v (src/vec3.h:101-103), not v.e/t.
template <typename T> struct Vec { T x, y, z; };

template <typename T> __host__ __device__ __forceinline__ Vec<T> mk(T x, T y, T z) { Vec<T> v; v.x = x; v.y = y; v.z = z; return v; }
template <typename T> __host__ __device__ __forceinline__ Vec<T> operator+(Vec<T> a, Vec<T> b) { return mk<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> __host__ __device__ __forceinline__ Vec<T> operator-(Vec<T> a, Vec<T> b) { return mk<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> __host__ __device__ __forceinline__ Vec<T> operator-(Vec<T> a) { return mk<T>(-a.x, -a.y, -a.z); }
template <typename T> __host__ __device__ __forceinline__ Vec<T> operator*(Vec<T> a, Vec<T> b) { return mk<T>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename T> __host__ __device__ __forceinline__ Vec<T> operator*(T t, Vec<T> a) { return mk<T>(t * a.x, t * a.y, t * a.z); }
template <typename T> __host__ __device__ __forceinline__ T dot(Vec<T> a, Vec<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __host__ __device__ __forceinline__ T len2(Vec<T> a) { return a.x * a.x + a.y * a.y + a.z * a.z; }

__device__ __forceinline__ float  rt_sqrt(float x)  { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }   // fast mode: one MUFU (sqrtf adds a range-scaling sequence even with -prec-sqrt=false)
__device__ __forceinline__ double rt_sqrt(double x) { return sqrt(x); }
__device__ __forceinline__ float  rt_fmin(float a, float b)   { return fminf(a, b); }
__device__ __forceinline__ double rt_fmin(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ float  rt_fabs(float a)  { return fabsf(a); }
__device__ __forceinline__ double rt_fabs(double a) { return fabs(a); }

// fast mode: reciprocal and reciprocal square root straight from the special-function unit (1-2 ulp, one instruction; `1.0f / x`
// and `1.0f / sqrtf(x)` compile to range-scaled sequences of 7 and 11 even with -prec-div=false).  A denormal argument gives
// +-inf: the callers clamp (slab_setup) or cannot meet one (squared lengths of directions)
__device__ __forceinline__ float  rt_rcp(float x)  { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ double rt_rcp(double x) { return 1.0 / x; }
__device__ __forceinline__ float  rt_rsqrt(float x)  { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ double rt_rsqrt(double x) { return 1.0 / sqrt(x); }     // parity mode: the reference's 1 / sqrt (src/vec3.h:117-119, :101-103)

template <typename T> __device__ __forceinline__ Vec<T> unit_vector(Vec<T> v) { return rt_rsqrt(len2(v)) * v; }   // src/vec3.h:117-119
template <typename T> __device__ __forceinline__ Vec<T> reflect(Vec<T> v, Vec<T> n) { return v - (T(2) * dot(v, n)) * n; }   // src/vec3.h:149-151

// src/vec3.h:153-158
template <typename T> __device__ __forceinline__ Vec<T> refract(Vec<T> uv, Vec<T> n, T etai_over_etat) {
    T cos_theta = rt_fmin(dot(-uv, n), T(1));
    Vec<T> r_out_perp = etai_over_etat * (uv + cos_theta * n);
    Vec<T> r_out_parallel = (-rt_sqrt(rt_fabs(T(1) - len2(r_out_perp)))) * n;
    return r_out_perp + r_out_parallel;
}

// ------------------------------------------------------------------------------ philox --
// Replaces the global mt19937 of src/rtweekend.h:25-29 (keyed-stream contract):
// Philox4x32-10, key = seed, counter = (pixel, sample, stage, k >> 2), draw k uses word k & 3,
// u = (word >> 8) * 2^-24 -- exactly representable in float and double.
// TABLE = true: the ten round keys (key + r * Weyl constant, the same for every block of a launch) come precomputed
// from the kernel's parameter block -- a constant-bank operand of the XOR, no instruction -- instead of two additions
// per round: 4 instead of 6 issue slots per round in the render kernels, whose hottest lines these are.
#ifndef RT_PHILOX_UNROLL
#define RT_PHILOX_UNROLL 10            // rounds per pass of the loop in generate() (10 = fully unrolled; the render kernels hold four copies of it)
#endif
#define RT_PRAGMA_(x) _Pragma(#x)
#define RT_PRAGMA(x) RT_PRAGMA_(x)
struct PhiloxKeys { uint32_t k[20]; };     // k[2r], k[2r + 1] = round keys of round r
inline void philox_round_keys(uint64_t seed, PhiloxKeys& out) {
    uint32_t k0 = uint32_t(seed), k1 = uint32_t(seed >> 32);
    for (int r = 0; r < 10; ++r) { out.k[2 * r] = k0; out.k[2 * r + 1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
}

template <bool TABLE> struct PhiloxT {
    uint32_t key0, key1;
    const uint32_t* rk;
    uint32_t pixel, sample, stage, k;
    uint32_t w0, w1, w2, w3;

    __device__ __forceinline__ void init(uint64_t seed) { key0 = uint32_t(seed); key1 = uint32_t(seed >> 32); rk = nullptr; k = 0; pixel = sample = stage = 0; }
    __device__ __forceinline__ void init(const PhiloxKeys& keys) { key0 = key1 = 0; rk = keys.k; k = 0; pixel = sample = stage = 0; }
    __device__ __forceinline__ void set(uint32_t px, uint32_t s, uint32_t st) { pixel = px; sample = s; stage = st; k = 0; }

    __device__ __forceinline__ void generate(uint32_t block) {
        uint32_t c0 = pixel, c1 = sample, c2 = stage, c3 = block;
        uint32_t k0 = key0, k1 = key1;
        RT_PRAGMA(unroll RT_PHILOX_UNROLL)
        for (int r = 0; r < 10; ++r) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            if (TABLE) { c0 = hi1 ^ c1 ^ rk[2 * r]; c1 = lo1; c2 = hi0 ^ c3 ^ rk[2 * r + 1]; c3 = lo0; }
            else { c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
        }
        w0 = c0; w1 = c1; w2 = c2; w3 = c3;
    }

    // 24-bit integer m of the next draw; u = m * 2^-24
    __device__ __forceinline__ uint32_t next24() {
        const uint32_t lane = k & 3u;
        if (lane == 0u) generate(k >> 2);
        ++k;
        const uint32_t w = lane == 0u ? w0 : (lane == 1u ? w1 : (lane == 2u ? w2 : w3));
        return w >> 8;
    }
};
typedef PhiloxT<false> Philox;

template <typename T> __device__ __forceinline__ T u01(uint32_t m) { return T(m) * T(1.0 / 16777216.0); }   // random_double()

// random_double(-1,1) = -1 + 2*u (src/rtweekend.h:31-34): with u = m*2^-24 this is the exact
// value a * 2^-23, a = m - 2^23.  The rejection tests |p|^2 < 1 of src/vec3.h:124,132 are exact
// in double (sums < 2^48 in units of 2^-46); evaluating them on the integers makes the float
// path take the same accept/reject decision as the reference.
__device__ __forceinline__ int32_t sym24(uint32_t m) { return int32_t(m) - (1 << 23); }

// src/vec3.h:129-135 with the g++ draw order z, y, x.  Generic form: continues the
// stream wherever it stands (used by the C2 ray generator).
template <typename T> __device__ __forceinline__ bool ball_accept(uint32_t mz, uint32_t my, uint32_t mx, Vec<T>& out) {
    const int32_t az = sym24(mz), ay = sym24(my), ax = sym24(mx);
    const int64_t l2 = int64_t(ax) * ax + int64_t(ay) * ay + int64_t(az) * az;
    if (l2 >= (int64_t(1) << 46)) return false;
    out = mk<T>(T(ax) * T(1.0 / 8388608.0), T(ay) * T(1.0 / 8388608.0), T(az) * T(1.0 / 8388608.0));
    return true;
}
template <typename T, typename R> __device__ __forceinline__ Vec<T> random_in_unit_sphere(R& rng) {
    for (;;) {
        const uint32_t mz = rng.next24(), my = rng.next24(), mx = rng.next24();
        Vec<T> v;
        if (ball_accept<T>(mz, my, mx, v)) return v;
    }
}

// The same sampler for a stream that stands at draw 0 of its stage (every scatter does): identical
// draws, but the Philox rounds are emitted ONCE, in a loop over 4-word blocks, instead of once per
// draw site -- the triplets (z, y, x) of iteration i are draws 3i..3i+2, so block b completes one
// triplet (b % 3 == 0, 1) or two (b % 3 == 2).  Smaller code (the kernels are I-cache bound) and
// the lanes of a warp stay converged on the expensive part.
// BLOCK0_READY: the caller has generated block 0 already (scatter does, for opaque and glass hits in one go).
template <typename T, typename R, bool BLOCK0_READY = false> __device__ __forceinline__ Vec<T> random_in_unit_sphere_from0(R& rng) {
    uint32_t c0 = 0, c1 = 0;
    int nc = 0;
    Vec<T> v = mk<T>(T(0), T(0), T(0));
    for (uint32_t b = 0;; ++b) {
        if (!BLOCK0_READY || b > 0u) rng.generate(b);
        const uint32_t w0 = rng.w0 >> 8, w1 = rng.w1 >> 8, w2 = rng.w2 >> 8, w3 = rng.w3 >> 8;
        if (nc == 0) {
            if (ball_accept<T>(w0, w1, w2, v)) break;
            c0 = w3; nc = 1;
        } else if (nc == 1) {
            if (ball_accept<T>(c0, w0, w1, v)) break;
            c0 = w2; c1 = w3; nc = 2;
        } else {
            if (ball_accept<T>(c0, c1, w0, v)) break;
            if (ball_accept<T>(w1, w2, w3, v)) break;
            nc = 0;
        }
    }
    return v;
}

// ------------------------------------------------------------------------------ camera --
// Derived camera frame (src/camera.h:118-125), computed on the host in double by the exact
// formulas of camera::initialize (src/camera.h:52-85) and narrowed to T.
template <typename T> struct CameraFrame {
    Vec<T> center, pixel00_loc, pixel_delta_u, pixel_delta_v, defocus_disk_u, defocus_disk_v;
    int image_width, image_height;
    int defocus;     // defocus_angle > 0
};

template <typename T> struct Ray { Vec<T> o, d; T tm; };

// src/camera.h:87-113.  Draw order: px, py (:104-105); disk iterations, first draw -> y,
// second -> x (src/vec3.h:123 under g++); time (:97).  Stream stage 0.
template <typename T, typename R> __device__ __forceinline__ Ray<T> get_ray(const CameraFrame<T>& c, int i, int j, R& rng) {
    const Vec<T> pixel_center = (c.pixel00_loc + T(i) * c.pixel_delta_u) + T(j) * c.pixel_delta_v;
    rng.generate(0u);                                   // draws 0..3: px, py, first disk pair (or time)
    const T px = T(-0.5) + u01<T>(rng.w0 >> 8);
    const T py = T(-0.5) + u01<T>(rng.w1 >> 8);
    const Vec<T> pixel_sample = pixel_center + (px * c.pixel_delta_u + py * c.pixel_delta_v);
    Ray<T> r;
    uint32_t mt;                                        // the draw that follows the accepted disk pair
    if (!c.defocus) {
        r.o = c.center;
        mt = rng.w2 >> 8;
    } else {
        int32_t ay = sym24(rng.w2 >> 8), ax = sym24(rng.w3 >> 8);
        bool ok = int64_t(ax) * ax + int64_t(ay) * ay < (int64_t(1) << 46);
        for (uint32_t b = 1u;; ++b) {                   // pairs never straddle a block: draws 4b, 4b+1 | 4b+2, 4b+3
            rng.generate(b);
            if (ok) { mt = rng.w0 >> 8; break; }
            ay = sym24(rng.w0 >> 8); ax = sym24(rng.w1 >> 8);
            if (int64_t(ax) * ax + int64_t(ay) * ay < (int64_t(1) << 46)) { mt = rng.w2 >> 8; break; }
            ay = sym24(rng.w2 >> 8); ax = sym24(rng.w3 >> 8);
            ok = int64_t(ax) * ax + int64_t(ay) * ay < (int64_t(1) << 46);
        }
        const T dx = T(ax) * T(1.0 / 8388608.0), dy = T(ay) * T(1.0 / 8388608.0);
        r.o = (c.center + dx * c.defocus_disk_u) + dy * c.defocus_disk_v;
    }
    r.d = pixel_sample - r.o;
    r.tm = u01<T>(mt);
    return r;
}

// ------------------------------------------------------------------------------ sphere --
// One sphere slot of the tagged SoA scene (replaces class sphere, src/sphere.h:60-66, and the
// material it points to): a = (center1.xyz, radius), b = (center_vec.xyz, radius^2),
// c = (albedo.rgb, param); meta = list_index << 3 | is_moving << 2 | material tag.
template <typename T> struct Quad;
template <> struct Quad<float>  { typedef float4 type; };
template <> struct Quad<double> { typedef double4 type; };

#define RT_META_TYPE(m) ((m) & 3)
#define RT_META_ID(m) ((m) >> 3)

// src/sphere.h:68-72 -- center1 + time*center_vec.  Static spheres carry center_vec = 0, for
// which the expression returns center1 exactly, so the is_moving branch (:31) is not needed.
template <typename T> __device__ __forceinline__ Vec<T> sphere_center(const typename Quad<T>::type& a, const typename Quad<T>::type& b, T tm) {
    return mk<T>(a.x, a.y, a.z) + tm * mk<T>(b.x, b.y, b.z);
}

// src/sphere.h:30-48 in the reference's own form (parity mode; also the double path the fast
// mode uses for big spheres).  `a` = direction.length_squared() is hoisted per ray.  Returns the
// accepted root or a negative value.  Tie rule of src/hittable_list.h:31 (strict <, first in
// list wins) is restated order-independently: a root equal to tbest is accepted only from a
// sphere with a smaller list index than the current best.  The walk carries SLOTS, not list indices: `tie()` looks the two
// indices up and is called only when a root equals tbest exactly -- almost never -- so a sphere test pays neither the
// load of its index nor a register for the best one.
template <typename T, typename Tie> __device__ __forceinline__ bool root_ok(T root, T tmin, T tbest, const Tie& tie) {
    return tmin < root && (root < tbest || (root == tbest && tie()));
}

// list index of `slot` < that of `best_slot` (tree slots: meta; slots >= n_slots: the big spheres kept out of the tree;
// best_slot < 0: there is no best yet, t_best is the caller's t_max and an equal root is not inside (tmin, tmax))
struct TieBySlot {
    const int* meta; const int* big_meta; int n_slots, slot, best_slot;
    __device__ __forceinline__ int id_of(int s) const { return RT_META_ID(s < n_slots ? meta[s] : big_meta[s - n_slots]); }
    __device__ __forceinline__ bool operator()() const { return best_slot >= 0 && id_of(slot) < id_of(best_slot); }
};

template <typename Tie>
__device__ __forceinline__ bool sphere_hit_ref(const Vec<double>& o, const Vec<double>& d, double a, const Vec<double>& center,
                                               double rr, double tmin, double tbest, const Tie& tie, double& t_out) {
    const Vec<double> oc = o - center;
    const double half_b = dot(oc, d);
    const double c = len2(oc) - rr;
    const double discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return false;
    const double sqrtd = sqrt(discriminant);
    double root = (-half_b - sqrtd) / a;
    if (!root_ok(root, tmin, tbest, tie)) {
        root = (-half_b + sqrtd) / a;
        if (!root_ok(root, tmin, tbest, tie)) return false;
    }
    t_out = root;
    return true;
}

// Fast-mode double helpers: reciprocal and square root good to ~1e-15 relative from a float seed and
// two Newton steps (8 and 10 instructions instead of the ~70 / ~85 of the IEEE-rounded routines).
// Only the fast mode uses them; parity mode keeps `/` and sqrt().
__device__ __forceinline__ double rcp_newton(double x) {
    double r = double(rt_rcp(float(x)));
    r = fma(r, fma(-x, r, 1.0), r);
    return fma(r, fma(-x, r, 1.0), r);
}
__device__ __forceinline__ double sqrt_newton(double x) {                 // x >= 0
    const double y0 = double(rt_rsqrt(fmaxf(float(x), 1e-36f)));
    const double y = y0 * fma(-0.5 * x * y0, y0, 1.5);                    // 1/sqrt(x) to ~1e-14
    const double s = x * y;
    return fma(0.5 * y, fma(-s, s, x), s);
}

// Fast-mode test of a BIG sphere (kept out of the tree, e.g. the r = 1000 ground): the quadratic of
// sphere_hit_ref in double -- |oc|^2 - r^2 must not cancel in float and the self-intersection root
// has to stay below tmin = 0.001 -- with inv_a = 1/|d|^2 hoisted per ray.
template <typename Tie>
__device__ __forceinline__ bool sphere_hit_big(const Vec<double>& o, const Vec<double>& d, double a, double inv_a, const Vec<double>& center,
                                               double rr, double tmin, double tbest, const Tie& tie, double& t_out) {
    const Vec<double> oc = o - center;
    const double half_b = dot(oc, d);
    const double c = len2(oc) - rr;
    const double discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return false;
    const double sqrtd = sqrt_newton(discriminant);
    double root = (-half_b - sqrtd) * inv_a;
    if (!root_ok(root, tmin, tbest, tie)) {
        root = (-half_b + sqrtd) * inv_a;
        if (!root_ok(root, tmin, tbest, tie)) return false;
    }
    t_out = root;
    return true;
}

// Fast-mode sphere test (float): the same quadratic, solved around the closest-approach point
// so that r^2 - |l|^2 does not cancel catastrophically (l = oc - (oc.d/a) d).  inv_a = 1/a.
template <typename Tie>
__device__ __forceinline__ bool sphere_hit_f32(const Vec<float>& o, const Vec<float>& d, float inv_a, const Vec<float>& center,
                                               float rr, float tmin, float tbest, const Tie& tie, float& t_out) {
    const Vec<float> oc = o - center;
    const float tca = -dot(oc, d) * inv_a;
    const Vec<float> l = oc + tca * d;
    const float disc = rr - len2(l);
    if (disc < 0.0f) return false;
    const float s = rt_sqrt(disc * inv_a);
    float root = tca - s;
    if (!root_ok(root, tmin, tbest, tie)) {
        root = tca + s;
        if (!root_ok(root, tmin, tbest, tie)) return false;
    }
    t_out = root;
    return true;
}

// ---------------------------------------------------------------------------- materials --
// material::scatter for the three materials (src/material.h:19-25, :35-41, :52-71).  `n` is the
// face-corrected normal (src/hittable.h:15-21).  Returns false when the ray is absorbed.
template <typename T> __device__ __forceinline__ T reflectance(T cosine, T ref_idx) {   // src/material.h:76-81
    T r0 = (T(1) - ref_idx) / (T(1) + ref_idx);
    r0 = r0 * r0;
    const T x = T(1) - cosine;
    const T x2 = x * x;
    return r0 + (T(1) - r0) * ((x2 * x2) * x);
}

template <typename T, typename R> __device__ __forceinline__ bool scatter(int type, Vec<T> albedo, T param, const Vec<T>& d_in, const Vec<T>& n,
                                                                          bool front_face, R& rng, Vec<T>& attenuation, Vec<T>& d_out) {
    // block 0 of the stage's stream for every lane at once: the first ball attempt of an opaque hit, the one uniform of a glass
    // hit (drawn below only if the ray can refract; generating it commits to nothing, the stream is a pure function of its
    // counter).  In a batch that mixes the two, glass lanes used to run their own ten rounds at 2-3 lanes.
    rng.generate(0u);
    if (type < 2) {
        // both opaque materials draw one point in the unit ball first (3k draws, metal even when
        // fuzz == 0, :38); the sampler is instantiated once for the two of them
        const Vec<T> ball = random_in_unit_sphere_from0<T, R, true>(rng);
        attenuation = albedo;
        if (type == 0) {                               // lambertian, no near_zero guard (:21)
            d_out = n + unit_vector(ball);
            return true;
        }
        const Vec<T> reflected = reflect(unit_vector(d_in), n);   // metal (:37-40)
        d_out = reflected + param * ball;
        return dot(d_out, n) > T(0);
    } else {                                           // dielectric
        attenuation = mk<T>(T(1), T(1), T(1));
        const T refraction_ratio = front_face ? (T(1) / param) : param;
        const Vec<T> unit_direction = unit_vector(d_in);
        const T cos_theta = rt_fmin(dot(-unit_direction, n), T(1));
        const T sin_theta = rt_sqrt(T(1) - cos_theta * cos_theta);
        const bool cannot_refract = refraction_ratio * sin_theta > T(1);
        // `||` short-circuit (:64): the uniform is drawn only when refraction is possible
        bool do_reflect = cannot_refract;
        if (!do_reflect) do_reflect = reflectance(cos_theta, refraction_ratio) > u01<T>(rng.w0 >> 8);
        if (do_reflect)
            d_out = reflect(unit_direction, n);
        else
            d_out = refract(unit_direction, n, refraction_ratio);
        return true;
    }
}

// src/camera_cpu.h:23-25
template <typename T> __device__ __forceinline__ Vec<T> sky(const Vec<T>& d) {
    const Vec<T> unit_direction = unit_vector(d);
    const T a = T(0.5) * (unit_direction.y + T(1));
    return (T(1) - a) * mk<T>(T(1), T(1), T(1)) + a * mk<T>(T(0.5), T(0.7), T(1));
}

}  // namespace rt
