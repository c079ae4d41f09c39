// Test harness for include/rtw.hpp (the C++ host mirror of the reference's scene / render API):
// the same scenes that raytracing-in-a-weekend_b200/scenes.py builds with the Python mirror, written
// line for line against rtw.hpp as a caller of the reference would write them. tests/test_cpp_host.py
// compares the flattened rtw_scene_desc / rtw_camera / rtw_render_opts with the Python mirror's
// (CPU) and the rendered images (GPU). Links only librtw_b200.so.
#include "../../include/rtw.hpp"
#include "../../include/rtw_json.hpp"

using namespace rtw;

static ImageTexture::Ptr solid(float r, float g, float b) { return ImageTexture::from_color(Vec3(r, g, b)); }
static const float PI = 3.14159274101257324f;   // f32(pi), as std::f32::consts::PI

struct JobCpp {
    Scene scene;
    Viewport viewport;
    Integrator ray_color;
    uint32_t mode;
};

// scenes.config_c4(smoke=True): SURVEY §8d C4, Book-2 Cornell box with two constant-density boxes
static JobCpp cornell_smoke(size_t width, size_t height, size_t samples) {
    const Vec3 Z;
    auto white = solid(0.73f, 0.73f, 0.73f), red = solid(0.65f, 0.05f, 0.05f), green = solid(0.12f, 0.45f, 0.15f);
    std::vector<Quad> quads = {
        Quad::new_(Vec3(555.f, 0.f, 0.f), Vec3(0.f, 0.f, 555.f), Vec3(0.f, 555.f, 0.f), SCATTER_M, Z, green),
        Quad::new_(Vec3(0.f, 0.f, 0.f), Vec3(0.f, 555.f, 0.f), Vec3(0.f, 0.f, 555.f), SCATTER_M, Z, red),
        Quad::new_(Vec3(343.f, 554.f, 332.f), Vec3(-130.f, 0.f, 0.f), Vec3(0.f, 0.f, -105.f),
                   Material::new_emmiting(0.f, 0.f, 1.f, Vec3(15.f, 15.f, 15.f)), Z, solid(1.f, 1.f, 1.f)),
        Quad::new_(Vec3(0.f, 0.f, 0.f), Vec3(0.f, 0.f, 555.f), Vec3(555.f, 0.f, 0.f), SCATTER_M, Z, white),
        Quad::new_(Vec3(555.f, 555.f, 555.f), Vec3(-555.f, 0.f, 0.f), Vec3(0.f, 0.f, -555.f), SCATTER_M, Z, white),
        Quad::new_(Vec3(0.f, 0.f, 555.f), Vec3(0.f, 555.f, 0.f), Vec3(555.f, 0.f, 0.f), SCATTER_M, Z, white)};
    Instance b1 = Instance::new_box(Vec3(0.f, 0.f, 0.f), Vec3(165.f, 330.f, 165.f), solid(0.f, 0.f, 0.f), SCATTER_M);
    b1.rotate(Vec3(0.f, 15.0f * PI / 180.0f, 0.f));
    b1.translate(Vec3(265.f, 0.f, 295.f));
    Instance b2 = Instance::new_box(Vec3(0.f, 0.f, 0.f), Vec3(165.f, 165.f, 165.f), solid(1.f, 1.f, 1.f), SCATTER_M);
    b2.rotate(Vec3(0.f, -18.0f * PI / 180.0f, 0.f));
    b2.translate(Vec3(130.f, 0.f, 65.f));
    for (Instance* b : {&b1, &b2}) { b->dist_fn = DistFn::const_density; b->density = 0.01f; }
    JobCpp j{Scene::new_({}, quads, {b1, b2}),
             Viewport::new_from_res(width, height, samples, 50, 2.0f, 40.0f, Vec3(278.f, 278.f, -800.f), Vec3(0.f, 0.f, 1.f),
                                    Vec3(0.f, 1.f, 0.f), std::nullopt, 0.0f),
             Integrator::BgColor, RTW_MODE_RANDOM};
    return j;
}

// tests/test_cpp_host.py::mixed_scene: every other constructor — Sphere::new's c*c albedo, a Perlin
// texture, a moving sphere, a glass pair, a generated checker image, a top-level triangle, a
// one-sphere volume instance, a mixed instance; stratified render loop, lens, shutter.
static JobCpp mixed(size_t width, size_t height, size_t samples) {
    std::vector<float> checker((size_t)16 * 8 * 3);
    for (uint32_t y = 0; y < 8; ++y)
        for (uint32_t x = 0; x < 16; ++x) {
            const bool odd = ((x / 2) + (y / 2)) & 1u;
            float* t = &checker[((size_t)y * 16 + x) * 3];
            t[0] = odd ? 0.9f : 0.2f; t[1] = odd ? 0.9f : 0.3f; t[2] = odd ? 0.9f : 0.1f;
        }
    auto noise = PerlinNoise::new_(11);
    std::vector<Sphere> spheres = {
        Sphere::new_(Vec3(-1.f, 0.f, -1.f), 0.5f, Vec3(0.8f, 0.8f, 0.8f), METALLIC_M),
        Sphere::new_with_texture(Vec3(0.f, -100.5f, -1.f), 100.f, Vec3(1.f, 1.f, 1.f), SCATTER_M,
                                 ImageTexture::from_color_noise(Vec3(1.f, 1.f, 1.f), 0.01f, noise)),
        Sphere::new_moving(Vec3(1.f, 0.f, -1.f), 0.4f, Vec3(0.9f, 0.6f, 0.2f), FUZZY3_M, Vec3(0.5f, 0.25f, 0.f)),
        Sphere::new_with_texture(Vec3(0.f, 0.f, -1.2f), 0.5f, std::nullopt, GLASS_M, solid(1.f, 1.f, 1.f)),
        Sphere::new_with_texture(Vec3(0.f, 0.f, -1.2f), 0.4f, std::nullopt, GLASSR_M, solid(1.f, 1.f, 1.f)),
        Sphere::new_moving_with_texture(Vec3(0.f, 1.1f, -1.5f), 0.45f, std::nullopt, SCATTER_M, Vec3(0.f, 0.f, 0.f),
                                        ImageTexture::new_(checker, 16, 8)),
        Sphere::new_(Vec3(2.f, 0.3f, -2.f), 0.6f)};
    std::vector<Quad> quads = {
        Triangle::new_(Vec3(-2.5f, -0.4f, -2.5f), Vec3(1.5f, 0.1f, 0.2f), Vec3(0.2f, 1.6f, 0.1f), SCATTER_M, ImageTexture::new_(checker, 16, 8)),
        Quad::new_(Vec3(-0.5f, 2.4f, -2.f), Vec3(1.f, 0.f, 0.f), Vec3(0.f, 0.f, 1.f),
                   Material::new_emmiting(0.f, 0.f, 1.f, Vec3(4.f, 4.f, 4.f)), Vec3(), solid(1.f, 1.f, 1.f))};
    Instance fog = Instance::new_sphere({Sphere::new_with_texture(Vec3(0.f, 0.f, 0.f), 0.7f, std::nullopt, SCATTER_M, solid(0.2f, 0.4f, 0.9f))});
    fog.translate(Vec3(-2.f, 0.4f, -1.6f));
    fog.dist_fn = DistFn::const_density;
    fog.density = 2.0f;
    Instance both = Instance::new_({Sphere::new_(Vec3(0.f, 0.f, 0.f), 0.25f, Vec3(0.9f, 0.1f, 0.1f), SCATTER_M)},
                                   {Quad::new_(Vec3(-0.5f, -0.3f, 0.f), Vec3(1.f, 0.f, 0.f), Vec3(0.f, 0.f, -1.f), METALLIC_M, Vec3(), solid(0.7f, 0.7f, 0.7f))});
    both.rotate(Vec3(0.1f, PI / 4.0f, -0.2f));
    both.translate(Vec3(1.2f, 0.9f, -1.4f));
    Scene sc = Scene::new_(spheres, quads, {fog, both});
    sc.background_color = Vec3(0.3f, 0.4f, 0.6f);
    Viewport vp = Viewport::new_(width, (float)width / (float)height, samples, 12, 2.0f, 70.0f, Vec3(0.2f, 0.5f, 1.5f),
                                 Vec3(-0.05f, -0.1f, -1.f), Vec3(0.f, 1.f, 0.f), "Rendering", 0.02f);
    vp.shutter_speed = 0.5f; vp.fps = 24.f; vp.frame = 3;
    return JobCpp{sc, vp, Integrator::BgColor, RTW_MODE_STRATIFIED};
}

static JobCpp make(int which, uint32_t w, uint32_t h, uint32_t spp) { return which == 0 ? cornell_smoke(w, h, spp) : mixed(w, h, spp); }

static std::unique_ptr<FlatScene> g_flat;   // keeps the last flattened scene alive for the caller
static std::string g_text;

extern "C" {
// Flatten scene `which` with the C++ mirror; the pointer stays valid until the next call.
const rtw_scene_desc* hm_flatten(int which, uint32_t w, uint32_t h, uint32_t spp, rtw_camera* cam, rtw_render_opts* opts) {
    try {
        JobCpp j = make(which, w, h, spp);
        g_flat = j.scene.flatten();
        if (cam) *cam = j.viewport.camera();
        if (opts) *opts = j.viewport.opts(j.ray_color, j.mode);
        return &g_flat->desc;
    } catch (const std::exception&) { return nullptr; }
}
// Render scene `which` on the device through rtw.hpp: out = H*W*3 floats, out8 = the quantised image.
// Returns 0, or the rtw error code (e.g. RTW_E_CUDA without a device: no CPU fallback).
int hm_render(int which, uint32_t w, uint32_t h, uint32_t spp, uint32_t flags, float* out, uint8_t* out8) {
    try {
        JobCpp j = make(which, w, h, spp);
        j.viewport.flags = flags;
        GpuScene gs(j.scene);
        Img img = j.mode == RTW_MODE_STRATIFIED ? j.viewport.render(j.ray_color, gs) : async_render(j.viewport, j.ray_color, gs);
        std::memcpy(out, img.rgb.data(), img.rgb.size() * sizeof(float));
        if (out8) { auto q = quantize(img); std::memcpy(out8, q.data(), q.size()); }
        return 0;
    } catch (const Error& e) { return e.code; }
}
// The JSON wire format (include/rtw_json.hpp): text of the sphere list of scene 1 ...
const char* hm_json_dumps(void) {
    JobCpp j = mixed(8, 8, 1);
    g_text = json::dumps(Scene::new_sphere(j.scene.spheres));
    return g_text.c_str();
}
// ... and a scene that came in as text, flattened. NULL + *parse_error = 1 on a ParseError.
const rtw_scene_desc* hm_json_flatten(const char* text, int* parse_error) {
    *parse_error = 0;
    try {
        g_flat = json::loads(text).flatten();
        return &g_flat->desc;
    } catch (const json::ParseError&) { *parse_error = 1; return nullptr; }
      catch (const std::exception&) { return nullptr; }
}
}
