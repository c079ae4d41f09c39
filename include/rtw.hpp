// rtw.hpp — C++17 host-side mirror of the reference's scene-building and render API,
// header-only, on top of the C ABI of rtw.h (librtw_b200.so). The reference is compiled code (Rust)
// and this image has no Rust toolchain, so this is the host side a compiled caller uses today;
// rust_shim/ is the same thing written against the reference's own structs (source only).
//
// Same type names, constructor names, argument order and quirks as Rust/src of
// Terence-23/RayTracing-in-a-weekend, so scenes written against the reference port line by line
// (`new` is a C++ keyword: the reference's `T::new(..)` is `T::new_(..)` here; `Option<X>` arguments
// are `std::optional<X>`):
//
//     Vec3                      Rust/src/vec3.rs:10-20,182-184
//     Material + presets        Rust/src/objects/materials.rs:15-20, 56-87, 157-212
//     PerlinNoise, ImageTexture Rust/src/texture.rs:21-27, 63-69, 110-149, 196-257
//     Sphere                    Rust/src/objects/sphere.rs:13-20, 151-247
//     Quad                      Rust/src/objects/quad.rs:8-20, 85-106
//     Triangle                  Rust2/src/objects/triangle.rs:11-48
//     Instance                  Rust/src/objects/instance.rs:21-39, 83-242
//     Scene                     Rust/src/viewport.rs:79-135
//     Viewport, render fns      Rust/src/viewport.rs:48-77, 215-269, 308-516
//     write_img_f32's quantiser Rust/src/write_img.rs:6-19
//
// Nothing here computes pixels: Scene::flatten() produces the plain arrays of rtw_scene_desc,
// rtw_scene_create builds the BVH and uploads, rtw_render* run the CUDA kernels. There is no CPU
// fallback: every call throws rtw::Error when the library reports a failure (no device, bad scene).
#ifndef RTW_HPP
#define RTW_HPP

#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "rtw.h"

namespace rtw {

// The reference unwraps / panics (viewport.rs:242); across the C ABI failures are codes, here exceptions.
struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};
inline void check(int rc) {
    if (rc != RTW_OK) throw Error(rc, std::string("rtw error ") + std::to_string(rc) + ": " + rtw_last_error());
}

// ---------------------------------------------------------------------------------------- vec3.rs
struct Vec3 {
    float x = 0.f, y = 0.f, z = 0.f;
    constexpr Vec3() = default;
    constexpr Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    static constexpr Vec3 new_(float x, float y, float z) { return Vec3(x, y, z); }   // vec3.rs:182-184
    friend constexpr Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
    friend constexpr Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
    friend constexpr Vec3 operator-(Vec3 a) { return {-a.x, -a.y, -a.z}; }
    friend constexpr Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
    friend constexpr bool operator==(Vec3 a, Vec3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
    void store(float out[3]) const { out[0] = x; out[1] = y; out[2] = z; }
};

// Scene-authoring randomness: the reference draws scenes and Perlin tables from rand 0.8.5's unseeded
// thread_rng (texture.rs:111-125); the harness needs the same scene everywhere, so it is a
// Philox4x32-10 stream with an explicit key: counter = (n, 0, 0, 0), key = (key, 0), f32 = 24 bits
// (rand's Standard). Bit-identical to raytracing-in-a-weekend_b200/philox.py. Never on the render path.
class PhiloxStream {
public:
    explicit PhiloxStream(uint64_t key) : k0_((uint32_t)key), k1_((uint32_t)(key >> 32)) {}
    uint32_t u32() {
        if ((n_ & 3u) == 0u || !have_) { block(n_ >> 2); have_ = true; }
        return buf_[n_++ & 3u];
    }
    float random() { return (float)(u32() >> 8) * (1.0f / 16777216.0f); }          // random::<f32>()
    float uniform(float lo, float hi) { return random() * (hi - lo) + lo; }
private:
    void block(uint64_t index) {
        uint32_t c0 = (uint32_t)index, c1 = 0, c2 = 0, c3 = 0, k0 = k0_, k1 = k1_;
        for (int r = 0; r < 10; ++r) {
            const uint64_t p0 = 0xD2511F53ull * c0, p1 = 0xCD9E8D57ull * c2;
            const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        buf_[0] = c0; buf_[1] = c1; buf_[2] = c2; buf_[3] = c3;
    }
    uint32_t k0_, k1_, buf_[4] = {0, 0, 0, 0};
    uint64_t n_ = 0;
    bool have_ = false;
};

// ----------------------------------------------------------------------------------- materials.rs
struct Material {   // materials.rs:15-20; `emmited` keeps the reference's spelling
    float metallicness = 0.f, opacity = 0.f, ir = 1.f;
    Vec3 emmited;
    static constexpr Material new_(float m, float opacity, float ir) { return Material{m, opacity, ir, Vec3()}; }   // :57-67
    static constexpr Material new_m(float m) { return Material{m, 0.0f, 1.0f, Vec3()}; }                             // :68-79
    static constexpr Material new_emmiting(float m, float opacity, float ir, Vec3 e) { return Material{m, opacity, ir, e}; }   // :80-87
    auto key() const { return std::make_tuple(metallicness, opacity, ir, emmited.x, emmited.y, emmited.z); }
};
inline constexpr Material METALLIC_M{1.0f, 0.0f, 1.0f, Vec3()};        // materials.rs:157-166
inline constexpr Material SCATTER_M{0.0f, 0.0f, 1.0f, Vec3()};         // :168-177
inline constexpr Material FUZZY3_M{0.7f, 0.0f, 1.0f, Vec3()};          // :179-188
inline constexpr Material GLASS_M{1.0f, 1.0f, 1.50f, Vec3()};          // :190-199
inline constexpr Material GLASSR_M{1.0f, 1.0f, 1.0f / 1.50f, Vec3()};  // :201-210  ir = 1.0 / GLASS_M.ir
inline constexpr Material EMPTY_M = SCATTER_M;                         // :212

// ------------------------------------------------------------------------------------- texture.rs
// texture.rs:63-69. `new_` (texture.rs:110-135): ranvec = unit(U[-1,1]^3) with no ball rejection; the
// three permutations come from create_permute, whose shuffle loop runs over the empty range 255..0,
// so they are the identity (texture.rs:137-149).
struct PerlinNoise {
    std::array<std::array<float, 3>, 256> ranvec{};
    std::array<uint32_t, 256> perm_x{}, perm_y{}, perm_z{};
    // The reference seeds from thread_rng; here the table stream is an explicit Philox key. Without one, keys are
    // dealt 7, 8, 9, ... in construction order, as the Python mirror does (scene.py PerlinNoise._auto_key).
    static std::shared_ptr<PerlinNoise> new_() { static uint64_t auto_key = 7; return new_(auto_key++); }
    static std::shared_ptr<PerlinNoise> new_(uint64_t key) {
        auto n = std::make_shared<PerlinNoise>();
        PhiloxStream rs(key);
        for (int i = 0; i < 256; ++i) {
            float v[3];
            for (float& c : v) c = rs.random() * 2.0f + -1.0f;              // Vec3::random(-1,1): x*(max-min)+min
            const float len = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
            n->ranvec[i] = {v[0] / len, v[1] / len, v[2] / len};           // .unit()
            n->perm_x[i] = n->perm_y[i] = n->perm_z[i] = (uint32_t)i;
        }
        return n;
    }
};

// texture.rs:21-27. `row` is the image width and `col` its height (texture.rs:196-204);
// texel (x, y) = img[y*row + x] (texture.rs:265). Shared, as every Sphere::clone in the reference
// deep-copies it and the device wants one copy.
struct ImageTexture {
    std::vector<float> img;   // row*col*3
    uint32_t row = 1, col = 1;
    std::shared_ptr<PerlinNoise> noise;
    float noise_scale = 1.0f;
    using Ptr = std::shared_ptr<const ImageTexture>;

    static Ptr make(std::vector<float> img, uint32_t w, uint32_t h, std::shared_ptr<PerlinNoise> noise, float scale) {
        if (img.size() != (size_t)w * h * 3) throw Error(RTW_E_INVALID, "ImageTexture: img must hold width*height texels");
        auto t = std::make_shared<ImageTexture>();
        t->img = std::move(img); t->row = w; t->col = h; t->noise = std::move(noise); t->noise_scale = scale;
        return t;
    }
    static Ptr new_(std::vector<float> img, uint32_t width, uint32_t height) { return make(std::move(img), width, height, nullptr, 1.0f); }   // texture.rs:197-205
    static Ptr from_color(Vec3 c) { return make({c.x, c.y, c.z}, 1, 1, nullptr, 1.0f); }                                                       // :206-214
    // from_path (texture.rs:215-226) decodes with the `image` crate; no decoder is linked here: pass the
    // decoded 8-bit RGB rows, converted as `into_rgb32f` does (u8 / 255, no sRGB curve).
    static Ptr from_rgb8(const uint8_t* rgb, uint32_t width, uint32_t height) {
        std::vector<float> f((size_t)width * height * 3);
        for (size_t i = 0; i < f.size(); ++i) f[i] = (float)rgb[i] / 255.0f;
        return make(std::move(f), width, height, nullptr, 1.0f);
    }
    // texture.rs:228-245: the reference builds a fresh PerlinNoise::new() inside; pass one to share or to fix its key
    static Ptr new_with_noise(std::vector<float> img, uint32_t w, uint32_t h, float scale, std::shared_ptr<PerlinNoise> noise = nullptr) {
        return make(std::move(img), w, h, noise ? std::move(noise) : PerlinNoise::new_(), scale);
    }
    static Ptr from_color_noise(Vec3 c, float scale, std::shared_ptr<PerlinNoise> noise = nullptr) {
        return make({c.x, c.y, c.z}, 1, 1, noise ? std::move(noise) : PerlinNoise::new_(), scale);
    }
};

// -------------------------------------------------------------------------------------- sphere.rs
struct Sphere {   // sphere.rs:13-20
    Vec3 origin;
    float radius = 0.f;
    Vec3 col_mod{1.f, 1.f, 1.f};
    Material mat = EMPTY_M;
    Vec3 velocity;
    ImageTexture::Ptr texture;

    // sphere.rs:151-172. NOTE: Some(c) sets BOTH col_mod = c and a 1x1 texture of colour c,
    // so the effective albedo is c*c.
    static Sphere new_(Vec3 origin, float r, std::optional<Vec3> col_mod = {}, std::optional<Material> mat = {}) {
        const Vec3 c = col_mod.value_or(Vec3(1.f, 1.f, 1.f));
        return Sphere{origin, r, c, mat.value_or(EMPTY_M), Vec3(), ImageTexture::from_color(c)};
    }
    static Sphere new_moving(Vec3 origin, float r, std::optional<Vec3> col_mod, std::optional<Material> mat, Vec3 velocity) {   // :173-196
        const Vec3 c = col_mod.value_or(Vec3(1.f, 1.f, 1.f));
        return Sphere{origin, r, c, mat.value_or(EMPTY_M), velocity, ImageTexture::from_color(c)};
    }
    static Sphere new_with_texture(Vec3 origin, float r, std::optional<Vec3> col_mod, std::optional<Material> mat, ImageTexture::Ptr tex) {   // :197-221
        return Sphere{origin, r, col_mod.value_or(Vec3(1.f, 1.f, 1.f)), mat.value_or(EMPTY_M), Vec3(), std::move(tex)};
    }
    static Sphere new_moving_with_texture(Vec3 origin, float r, std::optional<Vec3> col_mod, std::optional<Material> mat, Vec3 velocity,
                                          ImageTexture::Ptr tex) {   // :222-246
        return Sphere{origin, r, col_mod.value_or(Vec3(1.f, 1.f, 1.f)), mat.value_or(EMPTY_M), velocity, std::move(tex)};
    }
};

// ---------------------------------------------------------------------------------------- quad.rs
// quad.rs:8-20; normal / d / w (quad.rs:93-105) are derived by the backend from origin / u / v.
struct Quad {
    Vec3 origin, u, v;
    Material mat;
    Vec3 velocity;                     // stored, never read by the reference (quad.rs:13)
    ImageTexture::Ptr texture;
    uint32_t shape = RTW_SHAPE_PARALLELOGRAM;
    static Quad new_(Vec3 origin, Vec3 u, Vec3 v, Material mat, Vec3 velocity, ImageTexture::Ptr texture) {   // quad.rs:85-106
        return Quad{origin, u, v, mat, velocity, std::move(texture), RTW_SHAPE_PARALLELOGRAM};
    }
};
// Rust2/src/objects/triangle.rs:11-48 in this object model: the half of the parallelogram next to
// `origin`; goes wherever a Quad goes.
struct Triangle {
    static Quad new_(Vec3 origin, Vec3 u, Vec3 v, Material mat, ImageTexture::Ptr texture) {                  // triangle.rs:27-48
        return Quad{origin, u, v, mat, Vec3(), std::move(texture), RTW_SHAPE_TRIANGLE};
    }
};

// ------------------------------------------------------------------------------------ instance.rs
// instance.rs:21-26: the reference stores a `fn(f32) -> f32`; only these two exist and a function
// pointer cannot cross the ABI, so the field is the choice between them.
enum class DistFn : uint32_t { surface = RTW_DENSITY_SURFACE, const_density = RTW_DENSITY_CONST };

struct Instance {   // instance.rs:28-39
    std::vector<Sphere> spheres;
    std::vector<Quad> quads;
    Vec3 translation, rotation;        // rotation: the Euler angles fed to Vec3::rotated (vec3.rs:161-181)
    DistFn dist_fn = DistFn::surface;  // pub field, instance.rs:37
    float density = 0.f;               // pub field, instance.rs:38

    // instance.rs:83-165: six quads, in the reference's order and winding
    static Instance new_box(Vec3 a, Vec3 b, ImageTexture::Ptr tex, Material mat) {
        const Vec3 mn(std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z));
        const Vec3 mx(std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z));
        const Vec3 dx(mx.x - mn.x, 0.f, 0.f), dy(0.f, mx.y - mn.y, 0.f), dz(0.f, 0.f, mx.z - mn.z), z;
        return new_quads({Quad::new_(Vec3(mn.x, mn.y, mx.z), dx, dy, mat, z, tex),
                          Quad::new_(Vec3(mx.x, mn.y, mx.z), -dz, dy, mat, z, tex),
                          Quad::new_(Vec3(mx.x, mn.y, mn.z), -dx, dy, mat, z, tex),
                          Quad::new_(Vec3(mn.x, mn.y, mn.z), dz, dy, mat, z, tex),
                          Quad::new_(Vec3(mn.x, mx.y, mx.z), dx, -dz, mat, z, tex),
                          Quad::new_(Vec3(mn.x, mn.y, mn.z), dx, dz, mat, z, tex)});
    }
    static Instance new_sphere(std::vector<Sphere> s) { Instance i; i.spheres = std::move(s); return i; }   // :166-186
    static Instance new_quads(std::vector<Quad> q) { Instance i; i.quads = std::move(q); return i; }        // :187-207
    static Instance new_(std::vector<Sphere> s, std::vector<Quad> q) {                                      // :208-228
        Instance i; i.spheres = std::move(s); i.quads = std::move(q); return i;
    }
    void rotate(Vec3 rot) { rotation = rotation + rot; }        // :230-232
    Vec3 getr() const { return rotation; }
    void translate(Vec3 vec) { translation = translation + vec; }   // :237-239
    Vec3 gett() const { return translation; }
};

// ------------------------------------------------------------------------------- viewport.rs Scene
// The plain arrays of rtw_scene_desc plus what keeps their pointers alive.
struct FlatScene {
    rtw_scene_desc desc{};
    std::vector<rtw_sphere> spheres;
    std::vector<rtw_quad> quads;
    std::vector<rtw_instance> instances;
    std::vector<rtw_material> materials;
    std::vector<rtw_texture> textures;
    std::vector<rtw_noise> noises;
    std::vector<rtw_light> lights;
    std::vector<ImageTexture::Ptr> keep;
};

class GpuScene;

struct Scene {   // viewport.rs:79-88. BVH construction happens in the backend at rtw_scene_create.
    std::vector<Sphere> spheres;
    std::vector<Quad> quads;
    std::vector<Instance> instances;
    Vec3 background_color;             // pub field, viewport.rs:87
    // the `lights` argument of Rust2's light-biased integrators (Rust2/src/viewport/ray_color.rs:59,115): top-level
    // primitives of this scene, named by position: {RTW_LIGHT_SPHERE, i} = spheres[i], {RTW_LIGHT_QUAD, i} = quads[i]
    std::vector<rtw_light> lights;

    static Scene new_(std::vector<Sphere> s, std::vector<Quad> q, std::vector<Instance> i) {   // viewport.rs:120-135
        Scene sc; sc.spheres = std::move(s); sc.quads = std::move(q); sc.instances = std::move(i); return sc;
    }
    static Scene new_sphere(std::vector<Sphere> s) { return new_(std::move(s), {}, {}); }       // :90-104
    static Scene new_quad(std::vector<Quad> q) { return new_({}, std::move(q), {}); }           // :105-119

    // Same numbering rules as the Python mirror's Scene.flatten(): materials deduplicated by value,
    // textures by identity (1x1 un-noised colours by value), noises by identity, instance primitives
    // appended after the top-level ones in instance order.
    std::unique_ptr<FlatScene> flatten() const {
        auto fs = std::make_unique<FlatScene>();
        std::map<decltype(std::declval<Material>().key()), uint32_t> mat_ids;
        std::map<const ImageTexture*, uint32_t> tex_ids;
        std::map<std::tuple<float, float, float>, uint32_t> colour_ids;
        std::map<const PerlinNoise*, int32_t> noise_ids;
        auto mat_id = [&](const Material& m) {
            auto it = mat_ids.find(m.key());
            if (it != mat_ids.end()) return it->second;
            rtw_material o{m.metallicness, m.opacity, m.ir, {m.emmited.x, m.emmited.y, m.emmited.z}};
            fs->materials.push_back(o);
            return mat_ids[m.key()] = (uint32_t)fs->materials.size() - 1u;
        };
        auto noise_id = [&](const std::shared_ptr<PerlinNoise>& n) -> int32_t {
            if (!n) return -1;
            auto it = noise_ids.find(n.get());
            if (it != noise_ids.end()) return it->second;
            rtw_noise o;
            std::memcpy(o.ranvec, n->ranvec.data(), sizeof o.ranvec);
            std::memcpy(o.perm_x, n->perm_x.data(), sizeof o.perm_x);
            std::memcpy(o.perm_y, n->perm_y.data(), sizeof o.perm_y);
            std::memcpy(o.perm_z, n->perm_z.data(), sizeof o.perm_z);
            fs->noises.push_back(o);
            return noise_ids[n.get()] = (int32_t)fs->noises.size() - 1;
        };
        auto tex_id = [&](const ImageTexture::Ptr& t) {
            if (!t) throw Error(RTW_E_INVALID, "primitive without a texture");
            auto it = tex_ids.find(t.get());
            if (it != tex_ids.end()) return it->second;
            const bool colour = t->row == 1 && t->col == 1 && !t->noise;
            if (colour) {
                auto c = colour_ids.find({t->img[0], t->img[1], t->img[2]});
                if (c != colour_ids.end()) return tex_ids[t.get()] = c->second;
            }
            rtw_texture o{t->row, t->col, t->img.data(), noise_id(t->noise), t->noise_scale};
            fs->textures.push_back(o);
            fs->keep.push_back(t);
            const uint32_t id = (uint32_t)fs->textures.size() - 1u;
            if (colour) colour_ids[{t->img[0], t->img[1], t->img[2]}] = id;
            return tex_ids[t.get()] = id;
        };
        auto put_sphere = [&](const Sphere& s) {
            rtw_sphere o{};
            s.origin.store(o.origin); o.radius = s.radius; s.col_mod.store(o.col_mod); s.velocity.store(o.velocity);
            o.mat_id = mat_id(s.mat); o.tex_id = tex_id(s.texture);
            fs->spheres.push_back(o);
        };
        auto put_quad = [&](const Quad& q) {
            rtw_quad o{};
            q.origin.store(o.origin); q.u.store(o.u); q.v.store(o.v); q.velocity.store(o.velocity);
            o.mat_id = mat_id(q.mat); o.tex_id = tex_id(q.texture); o.shape = q.shape;
            fs->quads.push_back(o);
        };
        // numbering follows the Python mirror: all spheres (top level, then per instance), then all quads
        for (const Sphere& s : spheres) put_sphere(s);
        for (const Instance& in : instances) for (const Sphere& s : in.spheres) put_sphere(s);
        for (const Quad& q : quads) put_quad(q);
        for (const Instance& in : instances) for (const Quad& q : in.quads) put_quad(q);
        uint32_t first_sphere = (uint32_t)spheres.size(), first_quad = (uint32_t)quads.size();
        for (const Instance& in : instances) {
            rtw_instance o{};
            o.first_sphere = first_sphere; o.n_spheres = (uint32_t)in.spheres.size();
            o.first_quad = first_quad; o.n_quads = (uint32_t)in.quads.size();
            in.translation.store(o.translation); in.rotation.store(o.rotation);
            o.density_kind = (uint32_t)in.dist_fn; o.density = in.density;
            first_sphere += o.n_spheres; first_quad += o.n_quads;
            fs->instances.push_back(o);
        }
        rtw_scene_desc& d = fs->desc;
        d.abi_version = RTW_ABI_VERSION;
        d.n_spheres = (uint32_t)fs->spheres.size(); d.n_top_spheres = (uint32_t)spheres.size();
        d.n_quads = (uint32_t)fs->quads.size(); d.n_top_quads = (uint32_t)quads.size();
        d.n_instances = (uint32_t)fs->instances.size();
        d.n_materials = (uint32_t)fs->materials.size(); d.n_textures = (uint32_t)fs->textures.size();
        d.n_noises = (uint32_t)fs->noises.size();
        d.spheres = fs->spheres.data(); d.quads = fs->quads.data(); d.instances = fs->instances.data();
        d.materials = fs->materials.data(); d.textures = fs->textures.data(); d.noises = fs->noises.data();
        background_color.store(d.background);
        for (const rtw_light& l : lights) {
            const size_t lim = l.kind == RTW_LIGHT_SPHERE ? spheres.size() : quads.size();
            if (l.kind > RTW_LIGHT_QUAD || l.index >= lim) throw Error(RTW_E_INVALID, "light is not a top-level primitive of the scene");
        }
        fs->lights = lights;
        d.n_lights = (uint32_t)fs->lights.size(); d.lights = fs->lights.empty() ? nullptr : fs->lights.data();
        return fs;
    }
};

// A Scene flattened, BVH built and uploaded (rtw_scene_create); rtw_scene_destroy on destruction.
// Keep one per Scene across renders: render_multi clones the scene per frame in the reference
// (viewport.rs:263), here the device copy is reused.
class GpuScene {
public:
    explicit GpuScene(const Scene& scene, int device = -1) {
        auto fs = scene.flatten();
        check(rtw_scene_create(&fs->desc, device, &h_));
    }
    ~GpuScene() { if (h_) rtw_scene_destroy(h_); }
    GpuScene(const GpuScene&) = delete;
    GpuScene& operator=(const GpuScene&) = delete;
    GpuScene(GpuScene&& o) noexcept : h_(o.h_) { o.h_ = nullptr; }
    rtw_scene* get() const { return h_; }
    rtw_stats stats() const { rtw_stats s{}; check(rtw_get_stats(h_, &s)); return s; }
    rtw_scene_info info() const { rtw_scene_info i{}; check(rtw_scene_get_info(h_, &i)); return i; }
    float rebuild_lbvh() { float ms = 0.f; check(rtw_scene_rebuild_lbvh(h_, &ms)); return ms; }
    void use_sah() { check(rtw_scene_use_sah(h_)); }
private:
    rtw_scene* h_ = nullptr;
};

// ---------------------------------------------------------------------------- viewport.rs Viewport
// type Img = Vec<Vec<Rgb<f32>>> (viewport.rs:46): rows top to bottom, [j][i]; stored flat here.
struct Img {
    uint32_t width = 0, height = 0;
    std::vector<float> rgb;            // height * width * 3
    const float* at(uint32_t j, uint32_t i) const { return &rgb[((size_t)j * width + i) * 3]; }
};

// The closures the reference passes as `ray_color`.
enum class Integrator : uint32_t {
    Gradient = RTW_INT_GRADIENT,       // ray_color_gradient  viewport/ray_color.rs:12-41
    D = RTW_INT_D,                     // ray_color_d         main.rs:17-46
    BgColor = RTW_INT_BG_COLOR,        // ray_color_bg_color  viewport/ray_color.rs:43-92
    Normal = RTW_INT_NORMAL,           // viewport.rs:524-554
    Albedo = RTW_INT_ALBEDO,
    GlassTest = RTW_INT_GLASS_TEST,    // viewport/glass_tests.rs:8-55
    LightCast = RTW_INT_LIGHT_CAST,    // Rust2/src/viewport/ray_color.rs:55-109 (lights = Scene::lights)
    LightBiased = RTW_INT_LIGHT_BIASED // Rust2/src/viewport/ray_color.rs:111-164 (weight = Viewport::light_weight)
};

struct Viewport {   // viewport.rs:48-77; pub fields keep the reference's names
    size_t samples = 1;
    float aspect_ratio = 1.f;
    size_t width = 0, height = 0;
    size_t depth = 10;
    float gamma = 2.f;
    float shutter_speed = 0.f, fps = 30.f;      // viewport.rs:394-398 defaults
    size_t frame = 0, number_of_frames = 1, start_frame = 0;
    std::string msg;
    // constructor arguments kept for the camera derivation (private in the reference)
    float vfov = 90.f, lens_radius = 0.f;
    Vec3 origin, direction{0.f, 0.f, -1.f}, vup{0.f, 1.f, 0.f};
    uint64_t seed = 1;                          // the reference is unseeded; SURVEY §8d "Render RNG"
    uint32_t flags = 0;                         // RTW_FLAG_*
    float light_weight = 1.0f;                  // `biased_weight` of light_biased_ray_color (Rust2 ray_color.rs:116)

    // Viewport::new (viewport.rs:308-401)
    static Viewport new_(size_t width, float aspect_ratio, size_t samples, size_t depth, float gamma,
                         std::optional<float> vfov = {}, std::optional<Vec3> origin = {}, std::optional<Vec3> direction = {},
                         std::optional<Vec3> vup = {}, std::optional<std::string> msg = {}, std::optional<float> lens_radius = {}) {
        Viewport v;
        v.samples = samples; v.aspect_ratio = aspect_ratio; v.width = width;
        v.height = (size_t)((float)width / aspect_ratio);                   // viewport.rs:346
        v.depth = depth; v.gamma = gamma;
        v.vfov = vfov.value_or(90.f); v.origin = origin.value_or(Vec3());
        v.direction = direction.value_or(Vec3(0.f, 0.f, -1.f)); v.vup = vup.value_or(Vec3(0.f, 1.f, 0.f));
        v.msg = msg.value_or(""); v.lens_radius = lens_radius.value_or(0.f);
        return v;
    }
    // Viewport::new_from_res (viewport.rs:402-428): aspect = width as f32 / height as f32
    static Viewport new_from_res(size_t width, size_t height, size_t samples, size_t depth, float gamma,
                                 std::optional<float> vfov = {}, std::optional<Vec3> origin = {}, std::optional<Vec3> direction = {},
                                 std::optional<Vec3> vup = {}, std::optional<std::string> msg = {}, std::optional<float> lens_radius = {}) {
        return new_(width, (float)width / (float)height, samples, depth, gamma, vfov, origin, direction, vup, std::move(msg), lens_radius);
    }

    rtw_camera camera() const {
        rtw_camera cam{};
        float o[3], d[3], u[3];
        origin.store(o); direction.store(d); vup.store(u);
        check(rtw_camera_init(&cam, (uint32_t)width, aspect_ratio, vfov, o, d, u, lens_radius));
        return cam;
    }
    rtw_render_opts opts(Integrator ray_color, uint32_t mode) const {
        rtw_render_opts o{};
        rtw_render_opts_default(&o);
        o.samples = (uint32_t)samples; o.depth = (uint32_t)depth; o.gamma = gamma;
        o.integrator = (uint32_t)ray_color; o.mode = mode;
        o.shutter_speed = shutter_speed; o.fps = fps; o.frame = (uint32_t)frame; o.seed = seed;
        o.mint = o.maxt = 0.f;
        o.sample_begin = o.sample_count = o.row_begin = o.row_count = 0;
        o.output = RTW_OUT_RESOLVED; o.flags = flags; o.light_weight = light_weight;
        return o;
    }
    Img run(Integrator ray_color, const GpuScene& scene, uint32_t mode) const {
        const rtw_camera cam = camera();
        const rtw_render_opts o = opts(ray_color, mode);
        Img img; img.width = cam.width; img.height = cam.height;
        img.rgb.resize((size_t)cam.width * cam.height * 3);
        check(rtw_render(scene.get(), &cam, &o, img.rgb.data()));
        return img;
    }
    // Viewport::render (viewport.rs:430-478): ceil(sqrt(samples))^2 stratified samples, time = 0
    Img render(Integrator ray_color, const GpuScene& scene) const { return run(ray_color, scene, RTW_MODE_STRATIFIED); }
    // Viewport::render_no_rand (viewport.rs:479-516)
    Img render_no_rand(Integrator ray_color, const GpuScene& scene) const { return run(ray_color, scene, RTW_MODE_NO_RAND); }
    // render + write_img_f32's quantiser fused on the device (write_img.rs:6-19): H*W*3 bytes
    std::vector<uint8_t> render_rgb8(Integrator ray_color, const GpuScene& scene, uint32_t mode = RTW_MODE_RANDOM) const {
        const rtw_camera cam = camera();
        const rtw_render_opts o = opts(ray_color, mode);
        std::vector<uint8_t> out((size_t)cam.width * cam.height * 3);
        check(rtw_render_rgb8(scene.get(), &cam, &o, out.data()));
        return out;
    }
};

// async_render (viewport.rs:215-248). Blocking: the tokio row tasks become the GPU grid.
inline Img async_render(const Viewport& viewport, Integrator ray_color, const GpuScene& scene) {
    return viewport.run(ray_color, scene, RTW_MODE_RANDOM);
}
// render_multi (viewport.rs:249-269): one Img per frame in start_frame..start_frame+number_of_frames.
inline std::vector<Img> render_multi(const Viewport& viewport, Integrator ray_color, const GpuScene& scene) {
    const rtw_camera cam = viewport.camera();
    const rtw_render_opts o = viewport.opts(ray_color, RTW_MODE_RANDOM);
    const size_t per = (size_t)cam.width * cam.height * 3, n = viewport.number_of_frames;
    std::vector<float> all(per * n);
    check(rtw_render_frames(scene.get(), &cam, &o, (uint32_t)viewport.start_frame, (uint32_t)n, all.data()));
    std::vector<Img> out(n);
    for (size_t k = 0; k < n; ++k) {
        out[k].width = cam.width; out[k].height = cam.height;
        out[k].rgb.assign(all.begin() + k * per, all.begin() + (k + 1) * per);
    }
    return out;
}
// write_img_f32 up to `img.save` (write_img.rs:6-19): (x*255).clamp(0,255).round() as u8, NaN -> 0, on the device.
inline std::vector<uint8_t> quantize(const Img& img, int device = -1) {
    std::vector<uint8_t> out(img.rgb.size());
    check(rtw_quantize(img.rgb.data(), out.data(), img.rgb.size(), device));
    return out;
}

}  // namespace rtw
#endif  // RTW_HPP
