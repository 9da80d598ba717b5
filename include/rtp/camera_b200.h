// camera_b200.h -- B200Impl::Camera, the drop-in for CPUImpl::Camera (src/camera_cpu.h:4-28).
//
//     CPUImpl::Camera cam;  ...fields...;  cam.render(world);          // reference, src/main.cpp:55-70
//     B200Impl::Camera cam; ...fields...;  cam.render(world);          // this repository
//
// render() walks `world` into POD arrays (hittable::describe), hands them to librt_b200.so through
// the C ABI (rt_upload_scene / rt_render / rt_readback, include/rt_b200.h) and writes the P3 stream the
// reference prints (src/camera.h:35,45) to std::cout.  camera::render is non-virtual in the reference
// (src/camera.h:32), so hiding it here is what a statically typed caller -- all that exist -- binds to.
// Errors surface as std::runtime_error (the reference has no error path at all); there is no CPU
// fallback: without a CUDA device rt_create fails and render() throws.
#ifndef RTP_CAMERA_B200_H
#define RTP_CAMERA_B200_H

#include <cstring>
#include <string>

#include "camera.h"

namespace B200Impl {

class Camera : public camera {
  public:
    // knobs the reference does not have
    uint64_t seed = 0;                    // Philox key (the reference's global mt19937 has no seed knob)
    int precision = RT_F32;               // RT_F64 = parity mode (bit-identical to the harness build of the reference)
    int variant = RT_WAVEFRONT;           // RT_MEGAKERNEL | RT_WAVEFRONT | RT_MEGA_LOCKSTEP
    int device = 0;
    int bvh_builder = RT_BVH_SAH_HOST;    // RT_BVH_LBVH_GPU for large scenes
    int output_format = RT_IMG_P3;        // RT_IMG_P6 / RT_IMG_PFM: binary image on std::cout instead of the P3 text

    Camera() {}
    Camera(const Camera&) = delete;
    Camera& operator=(const Camera&) = delete;
    ~Camera() override { if (ctx_) rt_destroy(ctx_); }

    // camera::render(world), src/camera.h:32-50
    void render(const hittable& world) {
        initialize();
        upload(world);
        const rt_camera c = pod();
        rt_render_opts o = {};
        o.abi = RT_ABI_VERSION; o.precision = precision; o.variant = variant; o.seed = seed;
        check(rt_render(ctx_, &c, &o), "rt_render");
        write_image(std::cout, samples_per_pixel);
        std::clog << "\rDone.                 \n";
    }

    // Progressive use (the compute-side stand-in for start_interactive_loop,
    // src/graphical_environment.h:31): render samples [first, first + count) into the accumulation buffer
    // (first == 0 starts a new frame), then look at the frame so far -- preview() resolves it into a
    // device rgba8 buffer (and optionally a host copy), write_image() streams it in output_format.
    void render_pass(const hittable& world, int first, int count) {
        initialize();
        upload(world);
        const rt_camera c = pod();
        rt_render_opts o = {};
        o.abi = RT_ABI_VERSION; o.precision = precision; o.variant = variant; o.seed = seed;
        o.sample_begin = first; o.sample_count = count; o.accumulate = first > 0;
        check(rt_render(ctx_, &c, &o), "rt_render");
    }
    void* preview(uint8_t* host_rgba8 = nullptr) const {
        void* dev = nullptr;
        check(rt_preview(ctx_, 0, &dev, host_rgba8), "rt_preview");
        return dev;
    }
    void write_image(std::ostream& out, int spp_total = 0) const {
        size_t len = 0;
        if (output_format == RT_IMG_P3) len = size_t(image_width) * size_t(image_height) * 12 + 64;
        else check(rt_readback_image(ctx_, spp_total, output_format, nullptr, &len), "rt_readback_image");
        std::string buf(len, '\0');
        check(rt_readback_image(ctx_, spp_total, output_format, &buf[0], &len), "rt_readback_image");
        out.write(buf.data(), std::streamsize(len));
    }

    // camera::get_ray(i, j), src/camera.h:87-100, on the device.  The reference draws from its global
    // sequential generator; here successive calls walk the samples 0, 1, 2, ... of the keyed stream.
    ray get_ray(int i, int j) const {
        ensure_ctx();
        const rt_camera c = pod();
        const int32_t pi = i, pj = j;
        const uint32_t s = next_sample_++;
        rt_ray r;
        check(rt_gen_rays(ctx_, &c, &pi, &pj, &s, 1, seed, precision, &r), "rt_gen_rays");
        return ray(point3(r.orig[0], r.orig[1], r.orig[2]), vec3(r.dir[0], r.dir[1], r.dir[2]), r.tm);
    }

    // CPUImpl::Camera::ray_color, src/camera_cpu.h:8-26, for ONE ray: the bounce loop is driven from the
    // host through the parity hooks (closest hit, then scatter, both on the device).  Meant for
    // test-style calls (tests/tests.cpp:38-42); render() runs the whole loop inside one kernel.
    color ray_color(const ray& r_in, int depth, const hittable& world) const override {
        upload(world);
        rt_ray r = to_pod(r_in);
        color throughput(1, 1, 1);
        for (int bounce = 1; depth > 0; --depth, ++bounce) {
            rt_hit h;
            check(rt_trace_closest(ctx_, &r, 1, 0.001, infinity, precision, RT_TRACE_BVH, &h), "rt_trace_closest");
            if (h.prim < 0) {                                                 // sky, src/camera_cpu.h:23-25
                const vec3 u = unit_vector(vec3(r.dir[0], r.dir[1], r.dir[2]));
                const double a = 0.5 * (u.y() + 1.0);
                return throughput * ((1.0 - a) * color(1.0, 1.0, 1.0) + a * color(0.5, 0.7, 1.0));
            }
            const rt_material m = scene_.materials[size_t(scene_.spheres[size_t(h.prim)].material)];
            int32_t ok = 0; double att[3]; rt_ray out;
            check(rt_scatter(ctx_, &m, &r, &h, 1, seed + uint64_t(bounce), precision, &ok, att, &out), "rt_scatter");
            if (!ok) return color(0, 0, 0);
            throughput = throughput * color(att[0], att[1], att[2]);
            r = out;
        }
        return color(0, 0, 0);                                                // depth exhausted, :12-13
    }

    rt_stats stats() const { rt_stats s = {}; if (ctx_) rt_get_stats(ctx_, &s); return s; }

  private:
    mutable rt_ctx* ctx_ = nullptr;
    mutable rt_scene_builder scene_;
    mutable bool have_scene_ = false;
    mutable uint32_t next_sample_ = 0;

    template <typename V> static bool same_pod(const V& a, const V& b) {
        return a.size() == b.size() && (a.empty() || std::memcmp(a.data(), b.data(), a.size() * sizeof(a[0])) == 0);
    }
    static rt_ray to_pod(const ray& r) {
        rt_ray p;
        for (int k = 0; k < 3; ++k) { p.orig[k] = r.origin()[k]; p.dir[k] = r.direction()[k]; }
        p.tm = r.time();
        return p;
    }
    void check(int rc, const char* what) const {
        if (rc != RT_OK) throw std::runtime_error(std::string(what) + ": " + rt_last_error(ctx_));
    }
    void ensure_ctx() const {
        if (!ctx_ && rt_create(&ctx_, device) != RT_OK) throw std::runtime_error(std::string("rt_create: ") + rt_last_error(nullptr));
    }
    void upload(const hittable& world) const {
        ensure_ctx();
        rt_scene_builder b;
        world.describe(b);
        // skip the upload only when the world DESCRIBES the same as the one on the device: a world edited in place (or
        // another list at the same address) must be rendered as it is now, as the reference does (src/camera.h:32)
        if (have_scene_ && same_pod(b.spheres, scene_.spheres) && same_pod(b.materials, scene_.materials)) return;
        rt_bvh_opts o = {};
        o.abi = RT_ABI_VERSION; o.builder = bvh_builder;
        check(rt_upload_scene(ctx_, b.spheres.data(), b.spheres.size(), b.materials.data(), b.materials.size(), &o), "rt_upload_scene");
        scene_ = std::move(b);
        have_scene_ = true;
    }
};

}  // namespace B200Impl

#endif  // RTP_CAMERA_B200_H
