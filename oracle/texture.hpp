// ORACLE — test infrastructure only. Restates Rust/src/texture.rs (PerlinNoise :63-194,
// ImageTexture :196-267).
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <vector>
#include "vec3.hpp"

namespace oracle {

// texture.rs:63-69. Tables arrive as data through the C ABI (rtw_noise); the reference's
// create_permute loops over the empty range 255..0 and so returns the identity (texture.rs:137-149,
// SURVEY N12) — that quirk lives in whoever fills the tables, not here.
struct PerlinNoise {
    Vec3 ranvec[256];
    uint32_t perm_x[256], perm_y[256], perm_z[256];

    // texture.rs:86-108
    static float perlin_interp(const Vec3 c[2][2][2], float u, float v, float w) {
        float uu = u * u * (3.0f - 2.0f * u);
        float vv = v * v * (3.0f - 2.0f * v);
        float ww = w * w * (3.0f - 2.0f * w);
        float accum = 0.0f;
        for (int ii = 0; ii < 2; ++ii)
            for (int jj = 0; jj < 2; ++jj)
                for (int kk = 0; kk < 2; ++kk) {
                    float i = (float)ii, j = (float)jj, k = (float)kk;
                    Vec3 weight_v(u - i, v - j, w - k);
                    accum += (i * uu + (1.0f - i) * (1.0f - uu)) * (j * vv + (1.0f - j) * (1.0f - vv)) *
                             (k * ww + (1.0f - k) * (1.0f - ww)) * c[ii][jj][kk].dot(weight_v);
                }
        return accum;
    }
    // texture.rs:154-179
    float noise(const Vec3& p) const {
        float u = p.x - std::floor(p.x);
        float v = p.y - std::floor(p.y);
        float w = p.z - std::floor(p.z);
        long i = (long)std::floor(p.x), j = (long)std::floor(p.y), k = (long)std::floor(p.z);
        Vec3 c[2][2][2];
        for (long di = 0; di < 2; ++di)
            for (long dj = 0; dj < 2; ++dj)
                for (long dk = 0; dk < 2; ++dk)
                    c[di][dj][dk] = ranvec[(perm_x[(i + di) & 255] ^ perm_y[(j + dj) & 255] ^
                                            perm_z[(k + dk) & 255]) & 255u];
        return perlin_interp(c, u, v, w);
    }
};

// texture.rs:21-27, 196-267. `row` = width, `col` = height.
struct ImageTexture {
    std::shared_ptr<const std::vector<Vec3>> img;  // shared, the reference deep-clones (sphere.rs:12)
    size_t row = 1, col = 1;
    std::shared_ptr<const PerlinNoise> noise;      // Option<PerlinNoise>
    float noise_scale = 1.0f;

    // texture.rs:259-267
    Vec3 color_at(size_t x, size_t y, const Vec3& p) const {
        float noise_mult = noise ? noise->noise(p / noise_scale) : 1.0f;
        size_t idx = y * row + x;
        if (idx >= img->size()) idx = img->size() - 1;   // the reference would panic; never hit in tests
        return (*img)[idx] * noise_mult;
    }
};

}  // namespace oracle
