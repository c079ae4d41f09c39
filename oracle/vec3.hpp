// ORACLE — test infrastructure only. CPU restatement of Rust/src/vec3.rs (reference
// Terence-23/RayTracing-in-a-weekend). Never linked into the product library; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// f32 throughout, same expression order as the reference; build with -ffp-contract=off so a*b+c
// is not fused (rustc does not contract).
#pragma once
#include <cmath>
#include <cstdint>

namespace oracle {

// vec3.rs:10-15
struct Vec3 {
    float x = 0.f, y = 0.f, z = 0.f;
    Vec3() = default;
    Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}

    // vec3.rs:51-153 operator impls
    Vec3 operator-() const { return {-x, -y, -z}; }
    Vec3 operator*(float r) const { return {x * r, y * r, z * r}; }
    Vec3 operator*(const Vec3& r) const { return {x * r.x, y * r.y, z * r.z}; }
    Vec3 operator/(float r) const { return {x / r, y / r, z / r}; }
    Vec3 operator+(const Vec3& r) const { return {x + r.x, y + r.y, z + r.z}; }
    Vec3 operator-(const Vec3& r) const { return {x - r.x, y - r.y, z - r.z}; }
    Vec3& operator+=(const Vec3& r) { *this = *this + r; return *this; }
    Vec3& operator-=(const Vec3& r) { *this = *this - r; return *this; }
    Vec3& operator*=(float r) { *this = *this * r; return *this; }
    Vec3& operator/=(float r) { *this = *this / r; return *this; }

    // vec3.rs:161-181 — the reference's Euler map. Not a rotation matrix for x/z angles
    // (SURVEY N13); restated term by term, left-to-right products.
    Vec3 rotated(const Vec3& rot) const {
        float asin_ = std::sin(rot.x), acos_ = std::cos(rot.x);
        float bsin = std::sin(rot.y), bcos = std::cos(rot.y);
        float csin = std::sin(rot.z), ccos = std::cos(rot.z);
        Vec3 o;
        o.x = x * bcos * ccos + y * (asin_ * bsin * ccos - asin_ * ccos) +
              z * (acos_ * bsin * ccos + asin_ * csin);
        o.y = x * bcos * csin + y * (asin_ * bsin * csin + acos_ * ccos) +
              z * (acos_ * bsin * csin - asin_ * ccos);
        o.z = x * -bsin + y * asin_ * bcos + z * acos_ * bcos;
        return o;
    }
    void rotate(const Vec3& rot) { *this = rotated(rot); }

    float length2() const { return x * x + y * y + z * z; }              // vec3.rs:188-190
    float length() const { return std::sqrt(x * x + y * y + z * z); }    // vec3.rs:191-193
    float dot(const Vec3& o) const { return x * o.x + y * o.y + z * o.z; }  // vec3.rs:194-196
    Vec3 cross(const Vec3& o) const {                                    // vec3.rs:197-203
        return {y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x};
    }
    Vec3 unit() const { return *this / length(); }                       // vec3.rs:204-206
    Vec3 reflect(const Vec3& n) const { return *this - n * 2.0f * dot(n); }  // vec3.rs:256-258
    bool close_to_zero() const {                                         // vec3.rs:259-261
        return std::fabs(x) < 1e-7f && std::fabs(y) < 1e-7f && std::fabs(z) < 1e-7f;
    }
    // vec3.rs:17-21: PartialEq is "difference is close to zero"
    bool approx_eq(const Vec3& o) const { return (*this - o).close_to_zero(); }
};

// vec3.rs:269-302
struct Ray {
    Vec3 origin, direction;
    float time = 0.f;
    Ray() = default;
    Ray(Vec3 o, Vec3 d) : origin(o), direction(d), time(0.f) {}           // Ray::new
    Ray(Vec3 o, Vec3 d, float t) : origin(o), direction(d), time(t) {}    // Ray::new_with_time
    Vec3 at(float t) const { return origin + direction * t; }
    Ray rotated(const Vec3& rot) const { return Ray(origin.rotated(rot), direction.rotated(rot), time); }
};

}  // namespace oracle
