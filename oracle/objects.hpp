// ORACLE — test infrastructure only. Restates Rust/src/objects.rs (Hit :16-48, Interval :78-141),
// objects/materials.rs, objects/sphere.rs:99-147 and objects/quad.rs:37-106.
#pragma once
#include <cmath>
#include <limits>
#include <optional>
#include "rng.hpp"
#include "texture.hpp"
#include "vec3.hpp"

namespace oracle {

// ---------------------------------------------------------------- materials.rs:15-20
struct Material {
    float metallicness = 0.f, opacity = 0.f, ir = 1.f;
    Vec3 emmited;
};

// objects.rs:16-22
struct Hit {
    float t = -1.f;
    Vec3 normal, point, col_mod{1.f, 1.f, 1.f};
    Material mat;
};
using OptHit = std::optional<Hit>;

// `min_hit == None || min_hit > i` (aabb.rs:136,148; viewport.rs:145): Option<Hit> ordering via
// Hit::partial_cmp on t (objects.rs:44-48); strict, so the first of two equal t wins.
// `tie` (optional) is raised when two candidates have exactly equal t: the reference's winner then
// depends on its random BVH build order, so such pixels are unpinned.
inline void keep_closer(OptHit& min_hit, const OptHit& i, bool* tie = nullptr) {
    if (!i) return;
    if (tie && min_hit && min_hit->t == i->t) *tie = true;
    if (!min_hit || min_hit->t > i->t) min_hit = i;
}

// materials.rs:5-13
inline float lambertian_scatter_pdf(float cos_theta) {
    return cos_theta > 0.0f ? cos_theta * 0.31830988618379067154f /* FRAC_1_PI */ : 0.0f;
}

// materials.rs:89-97
inline Vec3 refract(const Vec3& uv, const Vec3& n, float etai_over_etat) {
    float cos_theta = (-uv).dot(n);
    if (cos_theta > 1.0f) cos_theta = 1.0f;
    Vec3 r_out_perp = (uv + n * cos_theta) * etai_over_etat;
    Vec3 r_out_parallel = n * -std::sqrt(std::fabs(1.0f - r_out_perp.length2()));
    return r_out_perp + r_out_parallel;
}
// materials.rs:98-103. powi(5) = x * ((x*x)*(x*x)) (compiler-rt __powisf2 order).
inline float reflectance(float cosine, float ref_idx) {
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    float x = 1.0f - cosine;
    float x2 = x * x;
    float x4 = x2 * x2;
    return r0 + (1.0f - r0) * (x * x4);
}

// materials.rs:213-228
inline Ray diffuse(const Hit& hit, const Ray& r, PathRng& rng) {
    Vec3 target = hit.normal + random_unit_vec(rng);
    if (target.close_to_zero()) return Ray(hit.point, hit.normal, r.time);
    return Ray(hit.point, target, r.time);
}

struct OnHit { Ray next; Vec3 emitted; float cos_theta; };

// materials.rs:105-154
inline OnHit on_hit(const Material& m, const Hit& h, const Ray& r, PathRng& rng) {
    if (m.opacity > 0.0f) {
        Vec3 n;
        bool front_face;
        if (r.direction.dot(h.normal) > 0.0f) { n = -h.normal; front_face = false; }
        else { n = h.normal; front_face = true; }
        float refraction_ratio = front_face ? 1.0f / m.ir : m.ir;
        Vec3 unit_direction = r.direction.unit();
        float cos_theta = (-unit_direction).dot(n);
        if (cos_theta > 1.0f) cos_theta = 1.0f;
        float sin_theta = std::sqrt(1.0f - cos_theta * cos_theta);
        bool cannot_refract = refraction_ratio * sin_theta > 1.0f;
        float refl = reflectance(cos_theta, refraction_ratio);
        // `cannot_refract || reflectance > random()` short-circuits in Rust: the coin is not drawn
        // on total internal reflection. The stream discipline (rng.hpp) draws one block per
        // dielectric bounce regardless, which changes no distribution.
        float coin = rng.next4().f[0];
        Vec3 direction = (cannot_refract || refl > coin) ? unit_direction.reflect(n)
                                                        : refract(unit_direction, n, refraction_ratio);
        return OnHit{Ray(h.point, direction, r.time), m.emmited, 0.0f};
    }
    Vec3 sc = diffuse(h, r, rng).direction;
    Ray reflect_(h.point, r.direction.unit().reflect(h.normal), r.time);
    reflect_.direction = reflect_.direction * m.metallicness + sc * (1.0f - m.metallicness);
    return OnHit{reflect_, m.emmited, m.metallicness != 1.0f ? sc.dot(h.normal) : 0.0f};
}

// ---------------------------------------------------------------- objects.rs:78-141
inline float minf(float a, float b) { return a <= b ? a : b; }
inline float maxf(float a, float b) { return a >= b ? a : b; }

struct Interval {
    float min = 0.f, max = 0.f;
    static Interval make(float x1, float x2) { return Interval{minf(x1, x2), maxf(x1, x2)}; }  // Interval::new
    Interval operator+(const Interval& r) const { return Interval{minf(min, r.min), maxf(max, r.max)}; }
    Interval operator+(float r) const { return Interval{min + r, max + r}; }
    // objects.rs:125-140: intersection of Y = aX + b with [min,max]
    std::optional<Interval> intersect(float a, float b) const {
        if (a == 0.0f) {
            if (min < b && b < max)
                return Interval{-std::numeric_limits<float>::infinity(), std::numeric_limits<float>::infinity()};
            return std::nullopt;
        }
        float inv_a = 1.0f / a;
        float x1 = (min - b) * inv_a;
        float x2 = (max - b) * inv_a;
        return Interval::make(x1, x2);
    }
};

struct Box3 { Interval x, y, z; };

// The slab test shared by AABB / QuadAABB / IAABB::collision_normal (aabb.rs:109-126,
// qaabb.rs:202-225, iaabb.rs:156-179): false = "return None".
inline bool slab_test(const Box3& b, const Ray& r, float mint, float maxt, uint64_t* counter = nullptr) {
    if (counter) ++*counter;
    auto xh = b.x.intersect(r.direction.x, r.origin.x);
    if (!xh) return false;
    auto yh = b.y.intersect(r.direction.y, r.origin.y);
    if (!yh) return false;
    auto zh = b.z.intersect(r.direction.z, r.origin.z);
    if (!zh) return false;
    float mn = maxf(maxf(xh->min, yh->min), maxf(zh->min, mint));
    float mx = minf(minf(xh->max, yh->max), minf(zh->max, maxt));
    return !(mn > mx);
}

// ---------------------------------------------------------------- sphere.rs:13-20
struct Sphere {
    Vec3 origin;
    float radius = 0.f;
    Vec3 col_mod{1.f, 1.f, 1.f};
    Material mat;
    Vec3 velocity;
    ImageTexture texture;

    // sphere.rs:99-147
    OptHit collision_normal(const Ray& r, float mint, float maxt) const {
        Vec3 origin_ = origin + velocity * r.time;
        Vec3 oc = r.origin - origin_;
        float a = r.direction.dot(r.direction);
        float b = oc.dot(r.direction);
        float c = oc.dot(oc) - radius * radius;
        float d = b * b - a * c;
        if (d < 0.0f) return std::nullopt;
        float x = (-b - std::sqrt(d)) / a;
        if (x < mint) x = (-b + std::sqrt(d)) / a;
        if (x < mint || x > maxt) return std::nullopt;
        Vec3 normal = (r.at(x) - origin_).unit();
        const float PI = 3.14159265358979323846f, FRAC_1_PI = 0.31830988618379067154f;
        float u = (std::atan2(-normal.z, normal.x) + PI) * FRAC_1_PI * 0.5f;
        float v = 1.0f - (FRAC_1_PI * std::acos(-normal.y));
        size_t tex_x = to_usize(std::floor(u * (float)(texture.row - 1)));
        size_t tex_y = to_usize(std::floor(v * (float)(texture.col - 1)));
        Hit h;
        h.t = x;
        h.normal = normal;
        h.point = r.at(x);
        h.mat = mat;
        h.col_mod = texture.color_at(tex_x, tex_y, r.at(x)) * col_mod;
        return h;
    }
    // Rust `f32 as usize` saturates: negative/NaN -> 0.
    static size_t to_usize(float f) { return (f > 0.0f) ? (size_t)f : 0; }
    // aabb.rs:27-38: leaf box = origin ± radius at time 0 — velocity ignored
    Box3 leaf_box() const {
        return Box3{Interval::make(origin.x - radius, origin.x + radius),
                    Interval::make(origin.y - radius, origin.y + radius),
                    Interval::make(origin.z - radius, origin.z + radius)};
    }
};

// ---------------------------------------------------------------- quad.rs:8-20
struct Quad {
    Vec3 origin, u, v;
    Material mat;
    Vec3 velocity;  // stored, never read (quad.rs:13)
    ImageTexture texture;
    Vec3 normal;
    float d = 0.f;
    Vec3 w;
    // Rust2/src/objects/triangle.rs in this object model: the same record (triangle.rs:36-47 =
    // quad.rs:93-105) with the triangle's inside test and leaf box.
    bool triangle = false;

    // quad.rs:85-106
    static Quad make(Vec3 origin, Vec3 u, Vec3 v, Material mat, Vec3 velocity, ImageTexture tex) {
        Quad q;
        Vec3 n = u.cross(v);
        q.origin = origin; q.u = u; q.v = v; q.mat = mat; q.velocity = velocity; q.texture = std::move(tex);
        q.normal = n.unit();
        q.d = q.normal.dot(origin);
        q.w = n / n.dot(n);
        return q;
    }
    // quad.rs:37-81
    OptHit collision_normal(const Ray& r, float mint, float maxt) const {
        float denominator = normal.dot(r.direction);
        if (std::fabs(denominator) <= 1e-8f) return std::nullopt;
        float t = (d - normal.dot(r.origin)) / denominator;
        if (t < mint || t > maxt) return std::nullopt;
        Vec3 point = r.at(t);
        Vec3 planar = point - origin;
        float alfa = w.dot(planar.cross(v));
        float beta = w.dot(u.cross(planar));
        if (triangle) {   // triangle.rs:107-113
            if (alfa < 0.0f || beta < 0.0f || alfa + beta > 1.0f) return std::nullopt;
        } else if (alfa < 0.0f || alfa > 1.0f || beta < 0.0f || beta > 1.0f) return std::nullopt;
        Hit h;
        h.t = t;
        h.normal = normal;
        h.point = point;
        size_t tx = alfa != 1.0f ? Sphere::to_usize(std::floor(alfa * (float)texture.row)) : texture.row - 1;
        size_t ty = beta != 1.0f ? Sphere::to_usize(std::floor(beta * (float)texture.col)) : texture.col - 1;
        h.col_mod = texture.color_at(tx, ty, point);
        h.mat = mat;
        return h;
    }
    // qaabb.rs:19-61 (From<Quad>) + pad() :107-126
    Box3 leaf_box() const {
        Vec3 op = origin + u + v, vc = origin + v, uc = origin + u;
        Box3 b;
        if (triangle) {   // triangle.rs:50-89: three corners, each axis at least MIN_AABB_WIDTH wide
            b.x = Interval{minf(vc.x, minf(uc.x, origin.x)), maxf(vc.x, maxf(uc.x, origin.x))};
            b.y = Interval{minf(vc.y, minf(uc.y, origin.y)), maxf(vc.y, maxf(uc.y, origin.y))};
            b.z = Interval{minf(vc.z, minf(uc.z, origin.z)), maxf(vc.z, maxf(uc.z, origin.z))};
            const float MIN_AABB_WIDTH = 0.005f;
            for (Interval* iv : {&b.x, &b.y, &b.z}) {
                if (iv->max - iv->min < MIN_AABB_WIDTH) {
                    float c = 0.5f * (iv->max + iv->min);
                    iv->max = c + MIN_AABB_WIDTH * 0.5f;
                    iv->min = c - MIN_AABB_WIDTH * 0.5f;
                }
            }
            return b;
        }
        b.x = Interval{minf(minf(op.x, vc.x), minf(uc.x, origin.x)), maxf(maxf(op.x, vc.x), maxf(uc.x, origin.x))};
        b.y = Interval{minf(minf(op.y, vc.y), minf(uc.y, origin.y)), maxf(maxf(op.y, vc.y), maxf(uc.y, origin.y))};
        b.z = Interval{minf(minf(op.z, vc.z), minf(uc.z, origin.z)), maxf(maxf(op.z, vc.z), maxf(uc.z, origin.z))};
        const float MIN_SIZE_DIR = 0.00001f;
        for (Interval* iv : {&b.x, &b.y, &b.z}) {
            if (iv->max - iv->min < MIN_SIZE_DIR) {
                float center = (iv->max - iv->min) / 2.0f + iv->min;
                iv->min = center - 0.5f * MIN_SIZE_DIR;
                iv->max = center + 0.5f * MIN_SIZE_DIR;
            }
        }
        return b;
    }
};

}  // namespace oracle
