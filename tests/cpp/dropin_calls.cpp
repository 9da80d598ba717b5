// dropin_calls.cpp -- the call sequence of the reference's unit test (tests/tests.cpp:35-45: initialize(),
// image_size(), get_ray(i, j), ray_color(r, depth, world) on a camera fixture) and of a caller that EDITS its
// world between two render() calls, written against include/rtp with B200Impl::Camera.  Values are printed with
// 17 digits on stdout as "name v0 v1 v2" lines; tests/test_cpp_dropin.py compares them with the CPU oracle
// driven through the same keyed streams.  (The reference's own tests.cpp is compiled against the same
// headers, unmodified but for the camera type, by oracle/Makefile where the reference tree exists.)
#include <cstdio>
#include <cstdlib>
#include <sstream>

#include "bvh.h"
#include "camera.h"
#include "camera_b200.h"
#include "hittable_list.h"
#include "material.h"
#include "sphere.h"

#include <gtest/gtest.h>

namespace {

void print3(const char* name, const vec3& v) { std::printf("%s %.17g %.17g %.17g\n", name, v.x(), v.y(), v.z()); }

class DropInFixture : public ::testing::Test {
  protected:
    void SetUp() override {
        cam.aspect_ratio = 16.0 / 9.0; cam.image_width = 400; cam.samples_per_pixel = 30; cam.max_depth = 50;
        cam.vfov = 20; cam.lookfrom = point3(13, 2, 3); cam.lookat = point3(0, 0, 0); cam.vup = vec3(0, 1, 0);
        cam.defocus_angle = 0.6; cam.focus_dist = 10.0;
        cam.precision = RT_F64;
    }
    hittable_list world;
    B200Impl::Camera cam;
};

// centre-pixel radiance over the ground sphere alone
TEST_F(DropInFixture, CentrePixelCalls) {
    world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));
    cam.initialize();
    const auto size = cam.image_size();
    EXPECT_EQ(size.first, 400);
    EXPECT_EQ(size.second, 225);
    const ray r = cam.get_ray(size.first / 2, size.second / 2);
    print3("ray_origin", r.origin());
    print3("ray_direction", r.direction());
    std::printf("ray_time %.17g\n", r.time());
    const color c = cam.ray_color(r, cam.max_depth, world);
    print3("ray_color", c);
    EXPECT_TRUE(c.x() >= 0 && c.x() <= 1 && c.z() <= 1) << c;
    const color black = cam.ray_color(r, 0, world);                  // depth exhausted: src/camera_cpu.h:12-13
    EXPECT_TRUE(black.similar_to(color(0, 0, 0))) << black;
}

// src/color.h as callers see it: write_color's text for a colour of their own (the frame's bytes come from the device)
TEST(ColorHeader, WriteColorFormatsLikeTheReference) {
    std::ostringstream o;
    write_color(o, color(0.5 * 8, 2.0 * 8, 0.0), 8);                  // sqrt(0.5) * 256 = 181.02; above 0.999 -> 255; 0
    write_color(o, color(0.0625, 0.25, 1.0), 1);                      // 64, 128, 255 (exact roots on byte boundaries)
    EXPECT_EQ(o.str(), std::string("181 255 0\n64 128 255\n"));
    EXPECT_EQ(linear_to_gamma(0.25), 0.5);
}

// a world edited in place between two render() calls is rendered as it is NOW (same list object, same count)
TEST_F(DropInFixture, EditedWorldIsRenderedAgain) {
    cam.image_width = 64; cam.samples_per_pixel = 2; cam.max_depth = 8;
    world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));
    world.add(make_shared<sphere>(point3(0, 1, 0), 1.0, make_shared<lambertian>(color(0.8, 0.1, 0.1))));
    std::ostringstream first, second, third;
    std::streambuf* old = std::cout.rdbuf(first.rdbuf());
    cam.render(world);
    world.clear();                                                   // same object, same number of spheres, other contents
    world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));
    world.add(make_shared<sphere>(point3(2, 1, 0), 1.0, make_shared<metal>(color(0.1, 0.8, 0.1), 0.0)));
    std::cout.rdbuf(second.rdbuf());
    cam.render(world);
    std::cout.rdbuf(third.rdbuf());
    cam.render(world);                                               // unchanged this time: same image again
    std::cout.rdbuf(old);
    EXPECT_NE(first.str(), second.str());
    EXPECT_EQ(second.str() == third.str(), true);
    std::printf("first_ppm_bytes %zu\n", first.str().size());
    std::fwrite(first.str().data(), 1, first.str().size(), stdout);
    std::printf("second_ppm_bytes %zu\n", second.str().size());
    std::fwrite(second.str().data(), 1, second.str().size(), stdout);
}

// the book's `world = hittable_list(make_shared<bvh_node>(world))` (src/bvh.h:10): same spheres in the same order -> same image
TEST_F(DropInFixture, WorldWrappedInABvhNodeRendersTheSameImage) {
    cam.image_width = 64; cam.samples_per_pixel = 2; cam.max_depth = 8;
    world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));
    world.add(make_shared<sphere>(point3(0, 1, 0), 1.0, make_shared<dielectric>(1.5)));
    world.add(make_shared<sphere>(point3(-4, 1, 0), point3(-4, 1.5, 0), 1.0, make_shared<lambertian>(color(0.4, 0.2, 0.1))));
    world.add(make_shared<sphere>(point3(4, 1, 0), 1.0, make_shared<metal>(color(0.7, 0.6, 0.5), 0.0)));
    std::ostringstream plain, wrapped, part;
    std::streambuf* old = std::cout.rdbuf(plain.rdbuf());
    cam.render(world);
    const auto node = make_shared<bvh_node>(world);
    EXPECT_EQ(node->bounding_box().x.min, world.bounding_box().x.min);
    EXPECT_EQ(node->bounding_box().y.max, world.bounding_box().y.max);
    hittable_list tree(node);
    std::cout.rdbuf(wrapped.rdbuf());
    cam.render(tree);
    hittable_list front(make_shared<bvh_node>(world.objects, 0, 2));      // the first two objects only
    std::cout.rdbuf(part.rdbuf());
    cam.render(front);
    std::cout.rdbuf(old);
    EXPECT_GT(plain.str().size(), size_t(64 * 36 * 6));
    EXPECT_EQ(plain.str() == wrapped.str(), true);
    EXPECT_NE(plain.str(), part.str());
    bool refused = false;
    try { bvh_node beyond(world.objects, 3, 9); } catch (const std::out_of_range&) { refused = true; }
    EXPECT_TRUE(refused);
}

}  // namespace
