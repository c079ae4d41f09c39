// ORACLE — test infrastructure only.
//
// The reference draws every random number from rand 0.8.5's unseeded thread_rng (ChaCha12):
// vec3.rs:223-254, materials.rs:127, instance.rs:25, viewport.rs:278,295-297,447,465-467.
// "Same seed" parity is therefore impossible; oracle and CUDA path instead share
// one counter-based stream, Philox4x32-10 (Salmon et al., SC'11), with this discipline:
//
//   key     = (seed lo32, seed hi32)
//   counter = (pixel index j*W+i, global sample index, block, stream)
//   stream 0 ("main"): blocks 0,1,2,... consumed in program order along the path:
//       camera:   one block  -> (xi_i, xi_j, xi_time, -)           [RANDOM / STRATIFIED only]
//                 lens disk (only when lens_radius != 0): one block per loop turn holding two
//                 candidate points (w0,w1) then (w2,w3)            [vec3.rs:240-254]
//       bounce:   dielectric branch: one block, w0 = the Schlick coin  [materials.rs:127]
//                 opaque branch: random_unit_vec, one block per rejection turn (w0,w1,w2)
//   stream 1+k ("volume of instance k"): block = (bounce << 16) | j
//       j = 0:    w0 = the free-flight xi of const_density         [instance.rs:24-26]
//       j >= 1:   rejection turns of the scatter normal's random_unit_vec [instance.rs:297]
//     Keyed by (bounce, instance) rather than drawn in visit order so the result does not depend
//     on BVH traversal order (the reference visits in random-build order).
//
//   u32 -> f32: (w >> 8) * 2^-24, identical to rand 0.8.5's Standard f32 (24-bit, [0,1)).
#pragma once
#include <cstdint>
#include "vec3.hpp"

namespace oracle {

struct Philox4 { uint32_t w[4]; };

inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return Philox4{{c0, c1, c2, c3}};
}

inline float u32_to_unit_f32(uint32_t w) { return (float)(w >> 8) * (1.0f / 16777216.0f); }

struct Rand4 { float f[4]; };

// Per-path generator. Copyable; recursion passes it by reference so draws are sequential.
struct PathRng {
    uint32_t k0 = 0, k1 = 0;
    uint32_t pixel = 0, sample = 0;
    uint32_t block = 0;   // next main-stream block

    Rand4 block_at(uint32_t blk, uint32_t stream) const {
        Philox4 p = philox4x32_10(pixel, sample, blk, stream, k0, k1);
        return Rand4{{u32_to_unit_f32(p.w[0]), u32_to_unit_f32(p.w[1]), u32_to_unit_f32(p.w[2]),
                      u32_to_unit_f32(p.w[3])}};
    }
    Rand4 next4() { return block_at(block++, 0u); }
};

// Vec3::random(-1,1) then ball rejection then unit (vec3.rs:216-238), over one stream.
// `draw(turn)` returns the block for rejection turn `turn`.
template <class Draw>
inline Vec3 random_unit_vec_from(Draw&& draw) {
    for (uint32_t turn = 0;; ++turn) {
        Rand4 r = draw(turn);
        Vec3 p(r.f[0] * (1.0f - -1.0f) + -1.0f, r.f[1] * (1.0f - -1.0f) + -1.0f,
               r.f[2] * (1.0f - -1.0f) + -1.0f);
        if (p.x * p.x + p.y * p.y + p.z * p.z <= 1.0f) return p.unit();
    }
}
inline Vec3 random_unit_vec(PathRng& rng) {
    return random_unit_vec_from([&](uint32_t) { return rng.next4(); });
}
// vec3.rs:240-254
inline Vec3 random_in_unit_disk(PathRng& rng) {
    for (;;) {
        Rand4 r = rng.next4();
        Vec3 p(r.f[0] * 2.0f - 1.0f, r.f[1] * 2.0f - 1.0f, 0.f);
        if (p.x * p.x + p.y * p.y <= 1.0f) return p;
        Vec3 q(r.f[2] * 2.0f - 1.0f, r.f[3] * 2.0f - 1.0f, 0.f);
        if (q.x * q.x + q.y * q.y <= 1.0f) return q;
    }
}

}  // namespace oracle
