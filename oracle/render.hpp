// ORACLE — test infrastructure only. Restates the integrators
// (Rust/src/viewport/ray_color.rs:12-92, main.rs:17-46, viewport.rs:524-554, glass_tests.rs:8-55)
// and the render loops + camera (Rust/src/viewport.rs:207-213, 270-305, 308-401, 430-516).
#pragma once
#include <atomic>
#include <cmath>
#include <thread>
#include <vector>
#include "../include/rtw.h"
#include "scene.hpp"

namespace oracle {

struct IntegratorCfg { uint32_t kind; float mint, maxt; float light_weight = 1.0f; };

inline IntegratorCfg integrator_cfg(const rtw_render_opts& o) {
    IntegratorCfg c{o.integrator, 0.001f, 100000.0f};
    switch (o.integrator) {
        case RTW_INT_GRADIENT:   c.mint = 0.001f; c.maxt = 100000.0f; break;  // ray_color.rs:17-18
        case RTW_INT_D:          c.mint = 0.001f; c.maxt = 1000.0f; break;    // main.rs:22-23
        case RTW_INT_BG_COLOR:   c.mint = 0.001f; c.maxt = 10000.0f; break;   // ray_color.rs:48-49
        case RTW_INT_NORMAL:     c.mint = 0.0f;   c.maxt = 1000.0f; break;    // viewport.rs:525-526
        case RTW_INT_ALBEDO:     c.mint = 0.001f; c.maxt = 100000.0f; break;
        case RTW_INT_GLASS_TEST: c.mint = 0.001f; c.maxt = 1000.0f; break;    // glass_tests.rs:13-14
        case RTW_INT_LIGHT_CAST:
        case RTW_INT_LIGHT_BIASED: c.mint = 0.001f; c.maxt = 1000.0f; break;  // Scene::new(.., 0.001, 1000.0), Rust2 tests.rs:639-644
    }
    c.light_weight = o.light_weight;
    if (o.maxt > 0.0f) { c.mint = o.mint; c.maxt = o.maxt; }
    return c;
}

// the sky of ray_color_gradient / ray_color_d / the normal closure (ray_color.rs:38-40)
inline Vec3 sky_gradient(const Ray& r) {
    Vec3 unit_direction = r.direction.unit();
    float t = 0.5f * (unit_direction.y + 1.0f);
    return Vec3((1.0f - t) + t * 0.5f, (1.0f - t) + t * 0.7f, 1.0f);
}

// ---- Rust2's light-biased estimators (Rust2/src/viewport/ray_color.rs:55-164) over the Rust/src scene model.
// Material::material_pdf (Rust2/src/objects/material.rs:17): the uber-material's diffuse branch answers like
// Lambertian (material.rs:44-63); Mirror (:93-101) and MirrorGlass (:199-235) are 0 for every direction but their own
// reflection / refraction, so metallicness == 1 and opacity > 0 give 0 towards any light.
inline float material_pdf(const Hit& h, const Ray& in, const Ray& r) {
    if (h.mat.opacity > 0.0f || h.mat.metallicness == 1.0f) return 0.0f;
    float cos = r.direction.unit().dot(h.normal.unit());
    float v = in.direction.dot(h.normal) >= 0.0f ? -cos : cos;
    v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    return v * 0.31830988618379067154f;
}
// the light loop (ray_color.rs:67-92, 130-152). `biased`: the threshold on pdf and the weight of :141-147.
inline Vec3 light_gather(const Hit& h, const Ray& in, const Scene& scene, const IntegratorCfg& ig, TraceCtx& cx, bool biased,
                         float& count) {
    Vec3 color(0.f, 0.f, 0.f);
    const uint32_t bounce = cx.bounce;
    for (const Scene::Light& l : scene.lights) {
        // middle of l.get_aabb() (Interval::mid_point, Rust2/src/objects/aabb.rs:52-54)
        Box3 b = l.quad ? PrimTraits<Quad>::leaf_box(scene.quads[l.index]) : PrimTraits<Sphere>::leaf_box(scene.spheres[l.index]);
        Vec3 middle_l((b.x.min + b.x.max) * 0.5f, (b.y.min + b.y.max) * 0.5f, (b.z.min + b.z.max) * 0.5f);
        Vec3 to_light = (middle_l - h.point).unit();
        Ray r(h.point, to_light);                                   // Ray::new: time 0
        cx.bounce = 0x8000u | (bounce & 0x7FFFu);                   // volume streams of the shadow ray (shared with the kernels)
        OptHit hr = scene.collision_normal(r, ig.mint, ig.maxt, cx);
        cx.bounce = bounce;
        OptHit hl = l.quad ? scene.quads[l.index].collision_normal(r, ig.mint, ig.maxt)
                           : scene.spheres[l.index].collision_normal(r, ig.mint, ig.maxt);
        if (!hr || !hl) continue;
        // `hr.0 == hl` (Hit: PartialEq over r, p, n, t - Rust2/src/objects/hit.rs:3-9)
        if (!(hr->t == hl->t && hr->point.x == hl->point.x && hr->point.y == hl->point.y && hr->point.z == hl->point.z &&
              hr->normal.x == hl->normal.x && hr->normal.y == hl->normal.y && hr->normal.z == hl->normal.z)) continue;
        Vec3 emmited = hl->mat.emmited;
        float pdf = material_pdf(h, in, r);
        float distance2 = hl->t * hl->t * r.direction.dot(r.direction);
        if (biased) {
            if (pdf <= 1.0f / (255.0f * maxf(maxf(emmited.x, emmited.y), emmited.z))) continue;
            count += ig.light_weight;
            color += emmited * pdf / distance2 * ig.light_weight;
        } else {
            count += 1.0f;
            color += emmited * pdf / distance2;
        }
    }
    return color;
}
// light_biased_ray_cast (ray_color.rs:55-109): first hit only.
inline Vec3 light_biased_ray_cast(const Ray& r, const Scene& scene, const IntegratorCfg& ig, TraceCtx& cx) {
    OptHit hit = scene.collision_normal(r, ig.mint, ig.maxt, cx);
    if (!hit) return scene.background_color;
    float count = 0.0f;
    Vec3 color = light_gather(*hit, r, scene, ig, cx, false, count);
    return (count != 0.0f ? color * hit->col_mod / count : Vec3(0.f, 0.f, 0.f)) + hit->mat.emmited;
}
// light_biased_ray_color (ray_color.rs:111-164). `o.reflect(&h)` is Material::on_hit of Rust/src with the zero-direction
// repair every Rust/src integrator applies (ray_color.rs:62-64).
inline Vec3 light_biased_ray_color(const Ray& r, const Scene& scene, uint32_t depth, const IntegratorCfg& ig, TraceCtx& cx) {
    if (depth <= 0) return scene.background_color;
    OptHit hit = scene.collision_normal(r, ig.mint, ig.maxt, cx);
    if (!hit) return scene.background_color;
    float count = 1.0f;
    bool front = !(r.direction.dot(hit->normal) > 0.0f);
    OnHit oh = on_hit(hit->mat, *hit, r, *cx.rng);
    Ray next = oh.next;
    if (next.direction.close_to_zero()) next.direction = front ? hit->normal : hit->normal * -1.0f;
    const uint32_t bounce = cx.bounce;
    cx.bounce = bounce + 1;
    Vec3 color = light_biased_ray_color(next, scene, depth - 1, ig, cx);
    cx.bounce = bounce;
    color += light_gather(*hit, r, scene, ig, cx, true, count);
    cx.bounce = bounce + 1;
    return color * hit->col_mod / count + hit->mat.emmited;
}

// Recursive, exactly as the reference. `depth` counts down (ray_color.rs:14-16).
inline Vec3 ray_color(const Ray& r, const Scene& scene, uint32_t depth, const IntegratorCfg& ig, TraceCtx& cx) {
    if (ig.kind == RTW_INT_NORMAL || ig.kind == RTW_INT_ALBEDO) {
        // first-hit shading (viewport.rs:524-554); the closure there brute-forces scene.spheres,
        // here every primitive class is searched so quads/instances can be checked too.
        OptHit hit = scene.collision_normal(r, ig.mint, ig.maxt, cx);
        if (hit) {
            if (ig.kind == RTW_INT_NORMAL) {
                Vec3 n = hit->normal;
                return Vec3(0.5f * (n.x + 1.0f), 0.5f * (n.y + 1.0f), 0.5f * (n.z + 1.0f));
            }
            return hit->col_mod;
        }
        return ig.kind == RTW_INT_NORMAL ? sky_gradient(r) : scene.background_color;
    }
    if (ig.kind == RTW_INT_LIGHT_CAST) return light_biased_ray_cast(r, scene, ig, cx);
    if (ig.kind == RTW_INT_LIGHT_BIASED) return light_biased_ray_color(r, scene, depth, ig, cx);
    if (depth < 1) return Vec3(0.f, 0.f, 0.f);
    OptHit hit = scene.collision_normal(r, ig.mint, ig.maxt, cx);
    if (hit) {
        if (ig.kind == RTW_INT_GLASS_TEST && hit->mat.metallicness != 1.0f) return Vec3(1.f, 1.f, 0.f);
        Vec3 cm = hit->col_mod;
        bool front = !(r.direction.dot(hit->normal) > 0.0f);
        OnHit oh = on_hit(hit->mat, *hit, r, *cx.rng);
        Ray next = oh.next;
        if (next.direction.close_to_zero()) next.direction = front ? hit->normal : hit->normal * -1.0f;
        cx.bounce += 1;
        Vec3 color = ray_color(next, scene, depth - 1, ig, cx) * cm;
        if (ig.kind != RTW_INT_BG_COLOR) return color;   // gradient / d / glass_test
        // ray_color.rs:67-88
        Vec3 scattered_color;
        if (hit->mat.metallicness != 1.0f) {
            float scattering_pdf = lambertian_scatter_pdf(oh.cos_theta);
            float pdf = scattering_pdf;
            scattered_color = color * scattering_pdf / pdf;    // 0/0 -> NaN when cos_theta <= 0
        }
        return color * hit->mat.metallicness + scattered_color * (1.0f - hit->mat.metallicness) + oh.emitted;
    }
    switch (ig.kind) {
        case RTW_INT_BG_COLOR: return scene.background_color;       // ray_color.rs:91
        case RTW_INT_GLASS_TEST: return Vec3(0.f, 0.f, 1.f);        // glass_tests.rs:54
        default: return sky_gradient(r);
    }
}

// viewport.rs:207-213
inline Vec3 gamma_correct(const Vec3& col, float g) {
    return Vec3(std::pow(col.x, g), std::pow(col.y, g), std::pow(col.z, g));
}

// viewport.rs:308-401 (Viewport::new)
inline void camera_init(rtw_camera* cam, uint32_t width, float aspect_ratio, float vfov, Vec3 c_origin,
                        Vec3 c_dir, Vec3 c_vup, float lens_radius) {
    Vec3 w = -c_dir;
    Vec3 u = c_vup.cross(w).unit();
    Vec3 v = w.cross(u);
    uint64_t height = (uint64_t)((float)width / aspect_ratio);
    float h = std::tan(vfov * 3.14159265358979323846f / 360.0f);
    float viewport_height = 2.0f * h;
    float viewport_width = aspect_ratio * viewport_height;
    Vec3 viewport_u = u * viewport_width;
    Vec3 viewport_v = -v * viewport_height;
    Vec3 pixel_delta_u = viewport_u / (float)width;
    Vec3 pixel_delta_v = viewport_v / (float)height;
    Vec3 viewport_upper_left = -w - viewport_u / 2.0f - viewport_v / 2.0f;
    Vec3 pixel00_loc = viewport_upper_left + (pixel_delta_u + pixel_delta_v) * 0.5f;
    cam->width = width; cam->height = (uint32_t)height;
    auto put = [](float* d, const Vec3& s) { d[0] = s.x; d[1] = s.y; d[2] = s.z; };
    put(cam->origin, c_origin); put(cam->pixel00, pixel00_loc);
    put(cam->delta_u, pixel_delta_u); put(cam->delta_v, pixel_delta_v);
    put(cam->u, u); put(cam->v, v);
    cam->lens_radius = lens_radius;
}

struct RenderResult { uint64_t paths = 0, rays = 0, nan_paths = 0; Counters ctr; };

inline Vec3 v3(const float* p) { return Vec3(p[0], p[1], p[2]); }

// One camera sample of pixel (i,j): restates the loop bodies of render_row (viewport.rs:287-300),
// Viewport::render (:455-470) and render_no_rand (:502-508).
inline Vec3 sample_pixel(const Scene& scene, const rtw_camera& cam, const rtw_render_opts& o,
                         const IntegratorCfg& ig, uint32_t i, uint32_t j, uint32_t s, uint32_t s_root,
                         Counters* ctr, bool count, bool* tie = nullptr) {
    PathRng rng;
    rng.k0 = (uint32_t)o.seed; rng.k1 = (uint32_t)(o.seed >> 32);
    rng.pixel = j * cam.width + i; rng.sample = s; rng.block = 0;
    TraceCtx cx; cx.rng = &rng; cx.bounce = 0; cx.ctr = count ? ctr : nullptr;
    Vec3 origin = v3(cam.origin), p00 = v3(cam.pixel00), du = v3(cam.delta_u), dv = v3(cam.delta_v);
    Ray r;
    if (o.mode == RTW_MODE_NO_RAND) {
        r = Ray(origin, p00 + du * (float)i + dv * (float)j);
    } else {
        Rand4 c = rng.next4();
        Vec3 random_point(0.f, 0.f, 0.f);
        if (cam.lens_radius != 0.0f) random_point = random_in_unit_disk(rng);
        Vec3 ro = origin + (v3(cam.u) * random_point.x + v3(cam.v) * random_point.y) * cam.lens_radius;
        if (o.mode == RTW_MODE_RANDOM) {
            float time = (float)o.frame / o.fps;
            r = Ray(ro, p00 + du * ((float)i + c.f[0]) + dv * ((float)j + c.f[1]), time + o.shutter_speed * c.f[2]);
        } else {  // STRATIFIED: x outer, y inner (viewport.rs:455-456); Ray::new -> time 0
            uint32_t x = s / s_root, y = s % s_root;
            r = Ray(ro, p00 + du * ((float)i + (((float)x + c.f[0]) / (float)s_root)) +
                            dv * ((float)j + (((float)y + c.f[1]) / (float)s_root)));
        }
    }
    Vec3 c = ray_color(r, scene, o.depth, ig, cx);
    if (tie && cx.tie) *tie = true;
    return c;
}

// Effective sample count of a mode (Viewport::render rounds up to a square, viewport.rs:443).
inline uint32_t effective_samples(const rtw_render_opts& o, uint32_t* s_root_out) {
    uint32_t s_root = 1, n = o.samples;
    if (o.mode == RTW_MODE_NO_RAND) n = 1;
    else if (o.mode == RTW_MODE_STRATIFIED) {
        s_root = (uint32_t)std::ceil(std::sqrt((float)o.samples));
        n = s_root * s_root;
    }
    if (s_root_out) *s_root_out = s_root;
    return n;
}

// Row-parallel over `threads` std::threads, one task per image row like async_render
// (viewport.rs:236-244).
inline RenderResult render(const Scene& scene, const rtw_camera& cam, const rtw_render_opts& o, float* out,
                           unsigned threads, uint8_t* tie_mask = nullptr) {
    IntegratorCfg ig = integrator_cfg(o);
    uint32_t s_root = 1;
    uint32_t total = effective_samples(o, &s_root);
    uint32_t s0 = o.sample_count ? o.sample_begin : 0, sn = o.sample_count ? o.sample_count : total;
    uint32_t r0 = o.row_count ? o.row_begin : 0, rn = o.row_count ? o.row_count : cam.height;
    float inv_g = 1.0f / o.gamma;
    if (threads == 0) threads = std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> next_row{0};
    std::vector<RenderResult> partial(threads);
    auto worker = [&](unsigned tid) {
        RenderResult& res = partial[tid];
        for (;;) {
            uint32_t jr = next_row.fetch_add(1);
            if (jr >= rn) break;
            uint32_t j = r0 + jr;
            if (o.row_stride != 0u && o.row_band != 0u) { uint32_t b = jr / o.row_band; j = r0 + b * o.row_stride + (jr - b * o.row_band); }   // interleaved bands (rtw_render_opts)
            for (uint32_t i = 0; i < cam.width; ++i) {
                Vec3 color(0.f, 0.f, 0.f);
                bool tie = false;
                for (uint32_t s = s0; s < s0 + sn; ++s) {
                    Vec3 c = sample_pixel(scene, cam, o, ig, i, j, s, s_root, &res.ctr, true, &tie);
                    if (!(std::isfinite(c.x) && std::isfinite(c.y) && std::isfinite(c.z))) ++res.nan_paths;
                    color += c;
                    ++res.paths;
                }
                Vec3 px = color;
                if (o.output == RTW_OUT_RESOLVED) {
                    if (o.mode == RTW_MODE_NO_RAND) px = gamma_correct(color, inv_g);           // viewport.rs:508
                    else px = gamma_correct(color / (float)total, inv_g);                        // :301, :472
                }
                if (tie_mask) tie_mask[(size_t)jr * cam.width + i] = tie ? 1 : 0;
                float* d = out + ((size_t)jr * cam.width + i) * 3;
                d[0] = px.x; d[1] = px.y; d[2] = px.z;
            }
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t) pool.emplace_back(worker, t);
    worker(0);
    for (auto& t : pool) t.join();
    RenderResult tot;
    for (auto& p : partial) {
        tot.paths += p.paths; tot.nan_paths += p.nan_paths;
        tot.ctr.rays += p.ctr.rays; tot.ctr.slab_tests += p.ctr.slab_tests; tot.ctr.sphere_tests += p.ctr.sphere_tests;
        tot.ctr.quad_tests += p.ctr.quad_tests; tot.ctr.instance_tests += p.ctr.instance_tests;
    }
    tot.rays = tot.ctr.rays;
    return tot;
}

}  // namespace oracle
