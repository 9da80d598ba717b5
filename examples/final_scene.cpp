// final_scene.cpp -- a caller written against the reference's API (the program of src/main.cpp:11-71,
// restated): the only B200-specific lines are the camera type and the optional knobs read from argv.
//   usage: final_scene [image_width spp max_depth [f32|f64 [p3|p6|pfm [passes]]]]  > image.ppm
//          (passes > 1: the samples are rendered in that many progressive passes with a preview after each)
//          final_scene --dump-scene > scene.bin      (the exported POD arrays, no GPU needed; used by tests)
#include <cstdlib>
#include <cstring>

#include "rtweekend.h"

#include "camera.h"
#include "camera_b200.h"
#include "color.h"
#include "hittable_list.h"
#include "material.h"
#include "sphere.h"

int main(int argc, char** argv) {
    hittable_list world;
    world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));

    for (int a = -11; a < 11; ++a) {
        for (int b = -11; b < 11; ++b) {
            const double which = random_double();
            // two draws inside one constructor call: g++ evaluates them right to left
            const double cz = b + 0.9 * random_double();
            const double cx = a + 0.9 * random_double();
            const point3 center(cx, 0.2, cz);
            if ((center - point3(4, 0.2, 0)).length() <= 0.9) continue;
            if (which < 0.8) {
                const color rhs = color::random(), lhs = color::random();     // right operand of `*` first
                const point3 center2 = center + vec3(0, random_double(0, .5), 0);
                world.add(make_shared<sphere>(center, center2, 0.2, make_shared<lambertian>(lhs * rhs)));
            } else if (which < 0.95) {
                const color albedo = color::random(0.5, 1);
                const double fuzz = random_double(0, 0.5);
                world.add(make_shared<sphere>(center, 0.2, make_shared<metal>(albedo, fuzz)));
            } else {
                world.add(make_shared<sphere>(center, 0.2, make_shared<dielectric>(1.5)));
            }
        }
    }
    world.add(make_shared<sphere>(point3(0, 1, 0), 1.0, make_shared<dielectric>(1.5)));
    world.add(make_shared<sphere>(point3(-4, 1, 0), 1.0, make_shared<lambertian>(color(0.4, 0.2, 0.1))));
    world.add(make_shared<sphere>(point3(4, 1, 0), 1.0, make_shared<metal>(color(0.7, 0.6, 0.5), 0.0)));

    if (argc > 1 && !std::strcmp(argv[1], "--dump-scene")) {
        rt_scene_builder b;
        world.describe(b);
        const uint64_t n[2] = { b.spheres.size(), b.materials.size() };
        std::cout.write(reinterpret_cast<const char*>(n), sizeof n);
        std::cout.write(reinterpret_cast<const char*>(b.spheres.data()), std::streamsize(b.spheres.size() * sizeof(rt_sphere)));
        std::cout.write(reinterpret_cast<const char*>(b.materials.data()), std::streamsize(b.materials.size() * sizeof(rt_material)));
        return 0;
    }

    B200Impl::Camera cam;
    cam.aspect_ratio = 16.0 / 9.0;
    cam.image_width = argc > 1 ? std::atoi(argv[1]) : 400;
    cam.samples_per_pixel = argc > 2 ? std::atoi(argv[2]) : 30;
    cam.max_depth = argc > 3 ? std::atoi(argv[3]) : 50;
    cam.vfov = 20;
    cam.lookfrom = point3(13, 2, 3);
    cam.lookat = point3(0, 0, 0);
    cam.vup = vec3(0, 1, 0);
    cam.defocus_angle = 0.6;
    cam.focus_dist = 10.0;
    if (argc > 4 && !std::strcmp(argv[4], "f64")) cam.precision = RT_F64;

    if (argc > 5 && !std::strcmp(argv[5], "p6")) cam.output_format = RT_IMG_P6;
    if (argc > 5 && !std::strcmp(argv[5], "pfm")) cam.output_format = RT_IMG_PFM;
    const int passes = argc > 6 ? std::atoi(argv[6]) : 1;
    if (passes <= 1) {
        cam.render(world);
        return 0;
    }
    // progressive: equal sample ranges into one buffer, the frame so far resolved on the device after each
    for (int k = 0, first = 0; k < passes; ++k) {
        const int count = (cam.samples_per_pixel * (k + 1)) / passes - first;
        if (count <= 0) continue;
        cam.render_pass(world, first, count);
        first += count;
        cam.preview();
        std::clog << "\rSamples so far: " << first << ' ' << std::flush;
    }
    cam.write_image(std::cout, cam.samples_per_pixel);
    std::clog << "\rDone.                 \n";
    return 0;
}
