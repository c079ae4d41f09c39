// ORACLE — test infrastructure only (see vec3.hpp header). C entry points over the CPU restatement
// so pytest / bench.py can drive it through ctypes with the very same rtw_scene_desc / rtw_camera /
// rtw_render_opts structs the product ABI takes (include/rtw.h).
//
// Parity pin status: Vec3::rotated is pinned by the reference's rotation_tests KATs
// (Rust/src/vec3.rs:364-404) and the sphere/metal path by the reference C++ tree's own
// Render_no_rand run (oracle/_ref, built from /root/reference/C++ by oracle/Makefile). Everything
// else has no asserting test in the reference => "parity unpinned" beyond those.
#include <chrono>
#include <cstring>
#include <memory>
#include "render.hpp"
#include "postprocessing.hpp"

using namespace oracle;

struct oracle_scene { Scene scene; };

namespace {
Material to_mat(const rtw_material& m) {
    Material o; o.metallicness = m.metallicness; o.opacity = m.opacity; o.ir = m.ir; o.emmited = v3(m.emitted);
    return o;
}
}  // namespace

extern "C" {

typedef struct oracle_stats {
    uint64_t paths, rays, nan_paths;
    uint64_t slab_tests, sphere_tests, quad_tests, instance_tests;
    double seconds;
    uint32_t threads;
} oracle_stats;

int oracle_scene_create(const rtw_scene_desc* d, uint32_t build_seed, oracle_scene** out) {
    if (!d || !out || d->abi_version != RTW_ABI_VERSION) return RTW_E_INVALID;
    auto os = std::make_unique<oracle_scene>();
    Scene& sc = os->scene;
    std::vector<std::shared_ptr<const PerlinNoise>> noises;
    for (uint32_t i = 0; i < d->n_noises; ++i) {
        auto n = std::make_shared<PerlinNoise>();
        for (int k = 0; k < 256; ++k) {
            n->ranvec[k] = v3(d->noises[i].ranvec[k]);
            n->perm_x[k] = d->noises[i].perm_x[k]; n->perm_y[k] = d->noises[i].perm_y[k]; n->perm_z[k] = d->noises[i].perm_z[k];
        }
        noises.push_back(n);
    }
    std::vector<ImageTexture> texs;
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const rtw_texture& t = d->textures[i];
        if (t.width == 0 || t.height == 0 || !t.rgb) return RTW_E_INVALID;
        auto img = std::make_shared<std::vector<Vec3>>((size_t)t.width * t.height);
        for (size_t k = 0; k < img->size(); ++k) (*img)[k] = v3(t.rgb + 3 * k);
        ImageTexture it; it.img = img; it.row = t.width; it.col = t.height; it.noise_scale = t.noise_scale;
        if (t.noise_id >= 0) { if ((uint32_t)t.noise_id >= d->n_noises) return RTW_E_INVALID; it.noise = noises[t.noise_id]; }
        texs.push_back(std::move(it));
    }
    auto mk_sphere = [&](const rtw_sphere& s, Sphere& o) -> bool {
        if (s.mat_id >= d->n_materials || s.tex_id >= d->n_textures) return false;
        o.origin = v3(s.origin); o.radius = s.radius; o.col_mod = v3(s.col_mod); o.velocity = v3(s.velocity);
        o.mat = to_mat(d->materials[s.mat_id]); o.texture = texs[s.tex_id];
        return true;
    };
    auto mk_quad = [&](const rtw_quad& q, Quad& o) -> bool {
        if (q.mat_id >= d->n_materials || q.tex_id >= d->n_textures) return false;
        if (q.shape > RTW_SHAPE_TRIANGLE) return false;
        o = Quad::make(v3(q.origin), v3(q.u), v3(q.v), to_mat(d->materials[q.mat_id]), v3(q.velocity), texs[q.tex_id]);
        o.triangle = q.shape == RTW_SHAPE_TRIANGLE;
        return true;
    };
    if (d->n_top_spheres > d->n_spheres || d->n_top_quads > d->n_quads) return RTW_E_INVALID;
    sc.spheres.resize(d->n_top_spheres);
    for (uint32_t i = 0; i < d->n_top_spheres; ++i) if (!mk_sphere(d->spheres[i], sc.spheres[i])) return RTW_E_INVALID;
    sc.quads.resize(d->n_top_quads);
    for (uint32_t i = 0; i < d->n_top_quads; ++i) if (!mk_quad(d->quads[i], sc.quads[i])) return RTW_E_INVALID;
    sc.instances.resize(d->n_instances);
    for (uint32_t k = 0; k < d->n_instances; ++k) {
        const rtw_instance& in = d->instances[k];
        Instance& o = sc.instances[k];
        if ((uint64_t)in.first_sphere + in.n_spheres > d->n_spheres || (uint64_t)in.first_quad + in.n_quads > d->n_quads) return RTW_E_INVALID;
        o.spheres.resize(in.n_spheres);
        for (uint32_t i = 0; i < in.n_spheres; ++i) if (!mk_sphere(d->spheres[in.first_sphere + i], o.spheres[i])) return RTW_E_INVALID;
        o.quads.resize(in.n_quads);
        for (uint32_t i = 0; i < in.n_quads; ++i) if (!mk_quad(d->quads[in.first_quad + i], o.quads[i])) return RTW_E_INVALID;
        o.translation = v3(in.translation); o.rotation = v3(in.rotation);
        o.const_density = in.density_kind == RTW_DENSITY_CONST; o.density = in.density;
    }
    sc.background_color = v3(d->background);
    if (d->n_lights && !d->lights) return RTW_E_INVALID;
    for (uint32_t i = 0; i < d->n_lights; ++i) {
        const rtw_light& l = d->lights[i];
        const bool quad = l.kind == RTW_LIGHT_QUAD;
        if (l.kind > RTW_LIGHT_QUAD || l.index >= (quad ? d->n_top_quads : d->n_top_spheres)) return RTW_E_INVALID;
        sc.lights.push_back(Scene::Light{quad, l.index});
    }
    sc.build(build_seed);
    *out = os.release();
    return RTW_OK;
}

void oracle_scene_destroy(oracle_scene* s) { delete s; }

// `tie_mask` (optional, rows*width bytes): 1 where some closest-hit decision of the pixel was taken
// between exactly equal t — the reference's result there depends on its random build order.
int oracle_render_ex(oracle_scene* s, const rtw_camera* cam, const rtw_render_opts* opts, float* out,
                     uint32_t threads, oracle_stats* st, uint8_t* tie_mask) {
    if (!s || !cam || !opts || !out) return RTW_E_INVALID;
    auto t0 = std::chrono::steady_clock::now();
    RenderResult r = render(s->scene, *cam, *opts, out, threads, tie_mask);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (st) {
        st->paths = r.paths; st->rays = r.rays; st->nan_paths = r.nan_paths;
        st->slab_tests = r.ctr.slab_tests; st->sphere_tests = r.ctr.sphere_tests;
        st->quad_tests = r.ctr.quad_tests; st->instance_tests = r.ctr.instance_tests;
        st->seconds = sec;
        st->threads = threads ? threads : std::max(1u, std::thread::hardware_concurrency());
    }
    return RTW_OK;
}

int oracle_render(oracle_scene* s, const rtw_camera* cam, const rtw_render_opts* opts, float* out,
                  uint32_t threads, oracle_stats* st) {
    return oracle_render_ex(s, cam, opts, out, threads, st, nullptr);
}

int oracle_camera_init(rtw_camera* cam, uint32_t width, float aspect_ratio, float vfov_deg, const float origin[3],
                       const float direction[3], const float vup[3], float lens_radius) {
    if (!cam) return RTW_E_INVALID;
    camera_init(cam, width, aspect_ratio, vfov_deg, v3(origin), v3(direction), v3(vup), lens_radius);
    return RTW_OK;
}

// ---- known-answer helpers for unit tests
void oracle_vec3_rotated(const float v[3], const float rot[3], float out[3]) {
    Vec3 r = v3(v).rotated(v3(rot)); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    Philox4 p = philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    std::memcpy(out, p.w, sizeof p.w);
}
float oracle_perlin_noise(const rtw_noise* n, const float p[3]) {
    PerlinNoise pn;
    for (int k = 0; k < 256; ++k) { pn.ranvec[k] = v3(n->ranvec[k]); pn.perm_x[k] = n->perm_x[k]; pn.perm_y[k] = n->perm_y[k]; pn.perm_z[k] = n->perm_z[k]; }
    return pn.noise(v3(p));
}
float oracle_reflectance(float cosine, float ref_idx) { return reflectance(cosine, ref_idx); }
void oracle_refract(const float uv[3], const float n[3], float ratio, float out[3]) {
    Vec3 r = refract(v3(uv), v3(n), ratio); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// Interval::intersect (objects.rs:125-140): returns 0 for None, else 1 with out = {min,max}
int oracle_interval_intersect(float imin, float imax, float a, float b, float out[2]) {
    auto r = Interval{imin, imax}.intersect(a, b);
    if (!r) return 0;
    out[0] = r->min; out[1] = r->max; return 1;
}

typedef struct oracle_hit { int32_t hit; float t; float normal[3]; float point[3]; float col_mod[3]; float metallicness, opacity, ir; float emitted[3]; } oracle_hit;
// One Scene::collision_normal call (viewport.rs:136-150), for single-ray known-answer tests.
int oracle_collide(oracle_scene* s, const float origin[3], const float dir[3], float time, float mint, float maxt,
                   uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, oracle_hit* out) {
    if (!s || !out) return RTW_E_INVALID;
    PathRng rng; rng.k0 = (uint32_t)seed; rng.k1 = (uint32_t)(seed >> 32); rng.pixel = pixel; rng.sample = sample;
    TraceCtx cx; cx.rng = &rng; cx.bounce = bounce;
    OptHit h = s->scene.collision_normal(Ray(v3(origin), v3(dir), time), mint, maxt, cx);
    std::memset(out, 0, sizeof *out);
    if (!h) return RTW_OK;
    out->hit = 1; out->t = h->t;
    auto put = [](float* d, const Vec3& v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; };
    put(out->normal, h->normal); put(out->point, h->point); put(out->col_mod, h->col_mod);
    out->metallicness = h->mat.metallicness; out->opacity = h->mat.opacity; out->ir = h->mat.ir; put(out->emitted, h->mat.emmited);
    return RTW_OK;
}

// write_img_f32's per-channel conversion (Rust/src/write_img.rs:12-14):
//   (x * 255.0).clamp(0.0, 255.0).round() as u8
// f32::clamp keeps NaN, f32::round rounds half away from zero, `as u8` saturates and maps NaN to 0.
void oracle_quantize(const float* rgb, uint8_t* out, uint64_t n_values) {
    for (uint64_t k = 0; k < n_values; ++k) {
        float x = rgb[k] * 255.0f;
        if (x < 0.0f) x = 0.0f;          // clamp: NaN fails both comparisons and stays NaN
        if (x > 255.0f) x = 255.0f;
        x = std::round(x);               // half away from zero, like f32::round
        out[k] = (x != x) ? (uint8_t)0 : (uint8_t)x;
    }
}

// Rust2/src/postprocessing.rs:70-135 bilateral_filter (oracle/postprocessing.hpp). params_out (may be
// null) receives {spatial, avg_gradient, inv_range, inv_spatial}.
void oracle_bilateral_filter(const uint8_t* rgb8, uint32_t width, uint32_t height, uint32_t size, int kind,
                             int exp_mode, uint8_t* out_rgb8, float* params_out) {
    oracle_post::Params p;
    oracle_post::bilateral_filter(rgb8, width, height, size, kind, exp_mode, out_rgb8, &p);
    if (params_out) { params_out[0] = p.spatial; params_out[1] = p.avg_gradient; params_out[2] = p.inv_range; params_out[3] = p.inv_spatial; }
}

}  // extern "C"
