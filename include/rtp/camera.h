// camera.h -- the configuration half of the reference's `camera` (src/camera.h:10-126): the ten public
// fields, image_size() and initialize().  What the reference's render loop does with them
// (src/camera.h:32-50: pixel / sample loops, get_ray, ray_color, write_color) is the hot path this
// repository moves onto the B200; see camera_b200.h for the class callers instantiate.
#ifndef RTP_CAMERA_H
#define RTP_CAMERA_H

#include <utility>

#include "scene_api.h"

class camera {
  protected:
    virtual ~camera() = default;                 // like the reference: only subclasses are instantiable (src/camera.h:12)

  public:
    double aspect_ratio = 1.0;                   // src/camera.h:15-26, same names and defaults
    int    image_width = 100;
    int    samples_per_pixel = 10;
    int    max_depth = 10;
    double vfov = 90;
    point3 lookfrom = point3(0, 0, -1);
    point3 lookat = point3(0, 0, 0);
    vec3   vup = vec3(0, 1, 0);
    double defocus_angle = 0;
    double focus_dist = 10;

    auto image_size() const { return std::make_pair(image_width, image_height); }

    // src/camera.h:52-54.  The derived frame (basis, pixel deltas, defocus disk; :56-85) is recomputed
    // by the library from the public fields with the reference's double formulas (rt_api.cu camera_frame).
    void initialize() {
        rt_camera c = pod();
        int w = 0, h = 0;
        if (rt_image_size(&c, &w, &h) != RT_OK) throw std::runtime_error("camera: bad image_width / aspect_ratio");
        image_height = h;
    }

    virtual color ray_color(const ray& r, int depth, const hittable& world) const = 0;   // src/camera.h:115

  protected:
    int image_height = 0;

    rt_camera pod() const {
        rt_camera c;
        c.aspect_ratio = aspect_ratio; c.image_width = image_width; c.samples_per_pixel = samples_per_pixel;
        c.max_depth = max_depth; c.pad = 0; c.vfov = vfov;
        for (int k = 0; k < 3; ++k) { c.lookfrom[k] = lookfrom[k]; c.lookat[k] = lookat[k]; c.vup[k] = vup[k]; }
        c.defocus_angle = defocus_angle; c.focus_dist = focus_dist;
        return c;
    }
};

#endif  // RTP_CAMERA_H
