// oracle/postprocessing.hpp — TEST INFRASTRUCTURE (CPU checker), not a product path.
// CPU restatement of the reference's 8-bit bilateral filter, Rust2/src/postprocessing.rs:11-135
// (SURVEY 8f-4 tail). Plain f32 loops in the reference's own order; compiled with
// -ffp-contract=off like the rest of the oracle (rustc does not contract a*b+c).
//
// Reference behaviour kept on purpose:
//   * Proximity::get_pixels (postprocessing.rs:30-61) walks `left..right` x `up..down`, both
//     half-open, so the window is [x-s, x+s) x [y-s, y+s) clipped to the image: the column x+s and
//     the row y+s are never visited, and on the last column / last row the pixel itself is not in
//     its own window. x is the OUTER loop, y the inner one (the f32 sums depend on that order).
//   * ProximityType::Edges keeps a pixel when |dx| + |dy| < size (a diamond inside that square).
//   * sigma_range is the mean luminance gradient of the interior pixels, accumulated in one f32
//     in row-major order (:75-87); a flat image gives 0, inv_range = inf, every weight NaN, and the
//     `as u8` casts (:127-131) turn NaN into 0: a constant image comes out black.
//   * `as u8` saturates and truncates towards zero.
// Parity pin: the reference's only test of this function writes a PNG and asserts nothing
// (postprocessing.rs:396-449) => "parity unpinned" by the reference; pinned here by an independent
// pure-Python restatement on small images (tests/test_postprocessing.py).
//
// exp: the reference calls f32::exp (the platform's expf). exp_mode 0 does the same (glibc expf,
// < 1 ulp but not always correctly rounded, and FMA-dependent between CPUs); exp_mode 1 rounds the
// double-precision exp to f32, which is what the CUDA path computes, so that GPU and oracle can be
// compared byte for byte. The two modes are compared with each other in the CPU tests.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace oracle_post {

inline uint8_t as_u8(float v) {                       // Rust `f32 as u8`
    if (!(v == v)) return 0;
    if (v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint8_t)v;
}

inline float intensity(const uint8_t* p) {            // postprocessing.rs:64-68
    return 0.2989f * (float)p[0] / 255.0f + 0.5870f * (float)p[1] / 255.0f + 0.1140f * (float)p[2] / 255.0f;
}

inline float exp_f32(float x, int exp_mode) {
    return exp_mode == 0 ? std::exp(x) : (float)std::exp((double)x);
}

struct Params { float spatial, avg_gradient, inv_range, inv_spatial; };

inline Params bilateral_params(const uint8_t* img, uint32_t w, uint32_t h) {   // :71-90
    Params p;
    p.spatial = std::ceil(0.02f * std::sqrt((float)(w * w + h * h)));
    float gradient_sum = 0.0f;
    for (uint32_t y = 1; y + 1 < h; ++y)
        for (uint32_t x = 1; x + 1 < w; ++x) {
            const float Iu = intensity(img + 3 * ((size_t)(y - 1) * w + x));
            const float Id = intensity(img + 3 * ((size_t)(y + 1) * w + x));
            const float Ir = intensity(img + 3 * ((size_t)y * w + x + 1));
            const float Il = intensity(img + 3 * ((size_t)y * w + x - 1));
            gradient_sum += std::sqrt((Iu - Id) * (Iu - Id) + (Il - Ir) * (Il - Ir));
        }
    p.avg_gradient = gradient_sum / (float)((w - 2u) * (h - 2u));
    p.inv_range = 0.5f / (p.avg_gradient * p.avg_gradient);
    p.inv_spatial = 0.5f / (p.spatial * p.spatial);
    return p;
}

// kind 0 = ProximityType::Square, 1 = ProximityType::Edges. img/out: h rows of w RGB8 pixels.
inline void bilateral_filter(const uint8_t* img, uint32_t w, uint32_t h, uint32_t size, int kind, int exp_mode,
                             uint8_t* out, Params* params_out) {
    const Params P = bilateral_params(img, w, h);
    if (params_out) *params_out = P;
    for (uint32_t y = 0; y < h; ++y)
        for (uint32_t x = 0; x < w; ++x) {
            const uint32_t left = x - (x < size ? x : size), up = y - (y < size ? y : size);
            const uint32_t right = x + (w - x - 1 < size ? w - x - 1 : size);
            const uint32_t down = y + (h - y - 1 < size ? h - y - 1 : size);
            float w_sum[3] = {0.0f, 0.0f, 0.0f}, col_sum[3] = {0.0f, 0.0f, 0.0f};
            const uint8_t* p = img + 3 * ((size_t)y * w + x);
            for (uint32_t i = left; i < right; ++i)
                for (uint32_t j = up; j < down; ++j) {
                    const uint32_t dx = x < i ? i - x : x - i, dy = y < j ? j - y : y - j;
                    if (kind == 1 && !(dx + dy < size)) continue;
                    const uint8_t* pi = img + 3 * ((size_t)j * w + i);
                    const float d2 = (float)((int64_t)dx * dx + (int64_t)dy * dy);
                    for (int c = 0; c < 3; ++c) {
                        const float t = (float)((int)pi[c] - (int)p[c]) / 255.0f;
                        const float wgt = exp_f32(-P.inv_spatial * d2 - P.inv_range * (t * t), exp_mode);
                        col_sum[c] += (float)pi[c] * wgt / 255.0f;
                        w_sum[c] += wgt;
                    }
                }
            uint8_t* o = out + 3 * ((size_t)y * w + x);
            for (int c = 0; c < 3; ++c) o[c] = as_u8(col_sum[c] * 255.0f / w_sum[c]);
        }
}

}  // namespace oracle_post
