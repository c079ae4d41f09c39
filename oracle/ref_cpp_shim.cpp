// ORACLE — test infrastructure only. C entry points over the REFERENCE'S OWN C++ renderer
// (/root/reference/C++: Book-1 spheres, compiled from where it lies into oracle/_ref/ by
// oracle/Makefile; no reference source is copied into this repo). Used to pin the oracle's
// camera, sphere intersection, normal and mirror-reflection arithmetic against code the
// reference's author wrote: Viewport::Render_no_rand (C++/src/viewport.cpp:59-105),
// Camera (C++/headers/viewport.h:24-46), Sphere::collisionNormal (C++/src/sphere.cpp:12-36),
// Material::onHit (C++/headers/materials.h:84-124).
#include <cstring>
#include <sstream>
#include "viewport.h"
#include "sphere.h"
#include "materials.h"

namespace {
int g_mode = 0;
float g_mint = 0.001f, g_maxt = 1000.0f;

Hit closest(const Ray& r, const Scene& scene, bool* mirror) {
    Hit min_hit = NO_HIT;
    for (const auto& s : scene.spheres) {   // brute-force closest hit, as C++/src/tests.cpp:228-239 does
        Hit hit = s.collisionNormal(r, g_mint, g_maxt);
        if (hit.isHit() && (!min_hit.isHit() || hit.t < min_hit.t)) {
            min_hit = hit;
            if (mirror) *mirror = s.material.metallicness == 1.0;
        }
    }
    return min_hit;
}
RGB_float sky(const Ray& r) {
    auto t = 0.5 * (r.direction.unit_vector().y + 1.0);
    return (1.0 - t) * RGB_float(1.0, 1.0, 1.0) + t * RGB_float(0.5, 0.7, 1.0);
}
// mode 0: first-hit normal shading   1: first-hit (t/maxt, 0, hit?)   2: the glass-test closure
// (tests.cpp:217-252 / glass_tests.rs:8-55): non-mirror hit -> yellow, miss -> blue, mirror -> follow
// `next`   3: follow mirrors, sky gradient on a miss (ray_color_d restricted to mirrors)
RGB_float cb(const Ray& r, const Scene& scene, uint depth, uint max_depth) {
    if (g_mode >= 2 && depth >= max_depth) return RGB_float(0, 0, 0);
    bool mirror = false;
    Hit h = closest(r, scene, &mirror);
    if (g_mode == 0) {
        if (!h.isHit()) return sky(r);
        return RGB_float(0.5 * (h.normal.x + 1.0), 0.5 * (h.normal.y + 1.0), 0.5 * (h.normal.z + 1.0));
    }
    if (g_mode == 1) return h.isHit() ? RGB_float(h.t / g_maxt, 0.0, 1.0) : RGB_float(0, 0, 0);
    if (!h.isHit()) return g_mode == 2 ? RGB_float(0.0, 0.0, 1.0) : sky(r);
    if (!mirror) return RGB_float(1.0, 1.0, 0.0);
    RGB_float c = cb(h.next, scene, depth + 1, max_depth);
    return RGB_float(c.R * h.col_mod.x, c.G * h.col_mod.y, c.B * h.col_mod.z);
}
}  // namespace

extern "C" int ref_cpp_render_no_rand(const float* spheres /* n x (ox,oy,oz,r, cr,cg,cb, metallicness,opacity,ir) */,
                                      int n, unsigned width, float aspect, float vfov, const float origin[3],
                                      const float vup[3], const float dir[3], int mode, float mint, float maxt,
                                      unsigned max_depth, float* out, unsigned* out_h) {
    try {
        Scene scene;
        for (int i = 0; i < n; ++i) {
            const float* s = spheres + 10 * i;
            scene.spheres.push_back(Sphere(vec3(s[0], s[1], s[2]), s[3], materials::Material(s[7], s[8], s[9]),
                                           vec3(s[4], s[5], s[6])));
        }
        Camera cam(width, aspect, vfov, vec3(origin[0], origin[1], origin[2]), vec3(vup[0], vup[1], vup[2]),
                   vec3(dir[0], dir[1], dir[2]), 0.0f);
        Viewport vp(cam, 1, max_depth);
        vp.gamma = 1.0f;
        g_mode = mode; g_mint = mint; g_maxt = maxt;
        std::streambuf* old = std::cout.rdbuf();
        std::ostringstream sink;
        std::cout.rdbuf(sink.rdbuf());            // Render_no_rand prints "Done"
        Img img = vp.Render_no_rand(cb, scene);
        std::cout.rdbuf(old);
        *out_h = (unsigned)img.size();
        for (size_t j = 0; j < img.size(); ++j)
            for (size_t i = 0; i < img[j].size(); ++i) {
                float* d = out + (j * img[j].size() + i) * 3;
                d[0] = (float)img[j][i].R; d[1] = (float)img[j][i].G; d[2] = (float)img[j][i].B;
            }
        return 0;
    } catch (...) {
        return -1;
    }
}
