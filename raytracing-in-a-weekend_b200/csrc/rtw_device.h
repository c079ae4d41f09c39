// rtw_device.h — plain-data layout of the device-resident scene, f32 vector algebra and the
// counter-based RNG shared by every kernel of librtw_b200.so.
//
// Everything here is `__host__ __device__` so that the host-side scene builder (scene_build.cpp)
// and the kernels (rtw_kernels.cu) agree on one definition. The arithmetic restates
// Rust/src/vec3.rs of Terence-23/RayTracing-in-a-weekend operation by operation (left-to-right
// products and sums); rtw_kernels.cu is compiled twice, once with -fmad=false (RTW_FLAG_STRICT_FP:
// no a*b+c contraction, like rustc) and once with nvcc's default contraction (the fast build).
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>
#include <cuda_runtime.h>   // float4, uint4, cudaTextureObject_t (host and device)

#if defined(__CUDACC__)
#define RTW_HD __host__ __device__ __forceinline__
#define RTW_HD_NOINL inline __host__ __device__ __noinline__
#else
#define RTW_HD inline
#define RTW_HD_NOINL inline
#endif

namespace rtw {

// The fast device build (nvcc's default contraction, no -DRTW_STRICT_FP) may reorder and
// approximate within ~1 ulp; the strict build and the host keep the reference's operation order.
#if defined(__CUDA_ARCH__) && !defined(RTW_STRICT_FP)
#define RTW_FAST_NODES 1
// 1/x, sqrt(x) and 1/sqrt(x) from the SFU approximations plus one Newton step (<= 1 ulp, no slow-path branch).
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return fmaf(fmaf(-x, r, 1.0f), r, r);
}
__device__ __forceinline__ float fast_rsqrt(float x) {
    float rs;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(x));
    return rs * fmaf(-0.5f * x, rs * rs, 1.5f);
}
__device__ __forceinline__ float fast_sqrt_pos(float x) {   // x > 0
    float rs;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(x));
    const float s = x * rs;
    return fmaf(fmaf(-s, s, x), 0.5f * rs, s);
}
#endif

// ------------------------------------------------------------------ Vec3 (vec3.rs:10-262)
struct V3 { float x, y, z; };
RTW_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
RTW_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
RTW_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RTW_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
RTW_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
RTW_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_FASTVEC)
RTW_HD V3 operator/(V3 a, float s) { const float r = fast_rcp(s); return v3(a.x * r, a.y * r, a.z * r); }
#else
RTW_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
#endif
RTW_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }            // vec3.rs:194-196
RTW_HD V3 cross(V3 a, V3 b) {                                                           // vec3.rs:197-203
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
RTW_HD float length(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }          // vec3.rs:191-193
RTW_HD float length2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }                // vec3.rs:188-190
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_FASTVEC)
RTW_HD V3 unit(V3 a) { const float r = fast_rsqrt(a.x * a.x + a.y * a.y + a.z * a.z); return v3(a.x * r, a.y * r, a.z * r); }
#else
RTW_HD V3 unit(V3 a) { return a / length(a); }                                           // vec3.rs:204-206
#endif
RTW_HD V3 reflect(V3 v, V3 n) { return v - n * 2.0f * dot(v, n); }                       // vec3.rs:256-258
RTW_HD bool close_to_zero(V3 a) {                                                        // vec3.rs:259-261
    return fabsf(a.x) < 1e-7f && fabsf(a.y) < 1e-7f && fabsf(a.z) < 1e-7f;
}
RTW_HD bool finite3(V3 a) { return isfinite(a.x) && isfinite(a.y) && isfinite(a.z); }

// bit casts
RTW_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
RTW_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

// ------------------------------------------------------------------ child / primitive references
// A reference names either an internal BVH node or one primitive: top 2 bits = kind.
enum : uint32_t { KIND_NODE = 0u, KIND_SPHERE = 1u, KIND_QUAD = 2u, KIND_INSTANCE = 3u };
static const uint32_t REF_NONE = 0xFFFFFFFFu;
RTW_HD uint32_t ref_kind(uint32_t r) { return r >> 30; }
RTW_HD uint32_t ref_index(uint32_t r) { return r & 0x1FFFFFFFu; }
RTW_HD uint32_t make_ref(uint32_t kind, uint32_t idx) { return (kind << 30) | idx; }
// The reference of a quad record whose inside test is the triangle's (alfa + beta <= 1,
// Rust2/src/objects/triangle.rs:107-113) carries this bit, in search results too: shape is part of
// the primitive's name, so the 64-byte record needs no seventeenth word. Indices stay below 2^29.
static const uint32_t REF_TRIANGLE = 0x20000000u;
static const uint32_t REF_INDEX_LIMIT = 0x10000000u;   // spheres, quads, instances, nodes: fewer than this
// Inside the BVH (node children, traversal stack) the reference of an instance that holds exactly
// one static sphere carries this bit, so that the walk can treat it as a primitive test rather than
// a level change. Search results always carry the plain reference.
static const uint32_t REF_ONE_SPHERE_INSTANCE = 0x20000000u;
// ... and the reference of a constant-density volume instance carries this one: such children are
// visited last among their siblings (see trav_step_node).
static const uint32_t REF_LATE_INSTANCE = 0x10000000u;
static const uint32_t REF_INSTANCE_FLAGS = 0x30000000u;
RTW_HD uint32_t instance_index(uint32_t r) { return r & 0x0FFFFFFFu; }

// ------------------------------------------------------------------ scene records (all 16-byte aligned)
// BVH node, 4-wide: the boxes of up to four children, one component per float4 (lane c of each
// vector belongs to child c), and their references: 128 B = one L1/L2 line, eight 128-bit loads.
// An unused slot has reference REF_NONE and the box [+inf,+inf]^3, which fails the slab test for
// every ray. A child box is the child's own reference box (a primitive's leaf box, an instance's
// world box, or the union for an inner node).
struct Node { float4 lox, loy, loz, hix, hiy, hiz; uint4 ref; uint4 pad; };
// The same node with the boxes as IEEE half floats, rounded outwards (lo down, hi up): 64 B, four
// 128-bit loads. a = (lox[4], loy[4]) b = (loz[4], hix[4]) c = (hiy[4], hiz[4]), two halves per
// word, child 0 in the low half of the first word. Used by the fast kernels: what bounds the walk
// is the number of 16-byte L1 accesses per node (every lane reads a different node), so halving
// the node halves the dominant cost. The boxes only grow, so no hit is lost; the strict kernels
// keep the exact f32 boxes (the reference's own leaf boxes).
struct Node16 { uint4 a, b, c; uint4 ref; };

// One traversable level (the scene top level, or the inside of one instance): root reference +
// root box (the reference slab-tests every BVH root: aabb.rs:109-126) + the level's moving
// spheres, which live outside the tree (their t=0 box is tested with the un-shrunk
// interval).
struct Level {
    float lo[3]; uint32_t root;        // REF_NONE when the level is empty (aabb.rs:106)
    float hi[3]; uint32_t mov_first;   // range in Scene::moving
    uint32_t mov_count, pad0, pad1, pad2;
};

// Instance (instance.rs:28-39): translation, the reference's Euler map for -rotation (world ->
// local) and +rotation (local -> world) as coefficient rows (see rot_apply), its level.
struct Instance {
    float t[3]; float density;
    float inv_rows[12];
    float fwd_rows[12];
    uint32_t flags;                    // bit0: const-density volume, bit1: rotation == 0
    uint32_t index;                    // position in Scene.instances (selects the volume RNG stream)
    uint32_t pad0, pad1;
    Level level;
};
enum : uint32_t { INST_VOLUME = 1u, INST_ROT_IDENTITY = 2u,
                  INST_ONE_SPHERE = 4u };   // the level is exactly one static sphere (Instance::new_sphere of one sphere)

struct Quad {   // quad.rs:8-20 with normal/d/w derived as quad.rs:93-105. 64 B.
    float4 q0;  // normal.xyz, d
    float4 q1;  // origin.xyz, u.x
    float4 q2;  // u.y u.z v.x v.y
    float4 q3;  // v.z, w.xyz
};

struct Material {  // materials.rs:15-20
    float metallicness, opacity, ir, pad;
    float emitted[3]; float pad1;
};

struct Texture {   // texture.rs:21-27 (`row` = width, `col` = height)
    uint32_t width, height;
    int32_t noise_id; float noise_scale;
    float solid[3]; uint32_t n_texels;
    cudaTextureObject_t obj;           // float4 texels, point filter, unnormalised coords; 0 for 1x1
    const float4* texels;              // linear copy (host builds; device keeps it for 1-row fallbacks)
};

struct Noise {     // texture.rs:63-69; perms stored & 255 (index is masked anyway, texture.rs:171-174)
    float4 ranvec[256];
    uint8_t perm[3][256];
};

struct Scene {
    const Node* nodes;
    const Node16* nodes16;        // same topology, same indices
    const float4* spheres;        // origin.xyz, radius
    const float4* sphere_vel;     // velocity.xyz, 0   (read only for moving spheres)
    const float4* sphere_col;     // col_mod.xyz, bits(flags: bit0 moving)
    const uint2* sphere_ids;      // mat_id, tex_id
    const float4* sphere_box;     // 2 per sphere: t=0 leaf box lo, hi (aabb.rs:27-38), moving only
    const Quad* quads;
    const uint2* quad_ids;
    const Instance* instances;
    const float4* one_sphere;     // 4 per instance, valid for INST_ONE_SPHERE ones: (T, density) (centre, radius) (box lo, bits flags) (box hi, bits index)
    const uint32_t* moving;       // sphere indices
    const Material* materials;
    const Texture* textures;
    const Noise* noises;
    Level top;
    float background[3];
    uint32_t n_noises;
    // Baked instances (rtw_host.h): world-space copies of the primitives of rigidly placed surface instances sit
    // behind the scene's own spheres / quads (index >= n_*_orig); *_src[index - n_*_orig] = (instance reference,
    // reference of the primitive inside it), which is what a search result names (trav_result).
    uint32_t n_spheres_orig, n_quads_orig;
    const uint2* baked_sphere_src;
    const uint2* baked_quad_src;
    // light list of the light-biased integrators (rtw_scene_desc.lights): 2 float4 per light,
    // (middle of the light's box, bits(reference of the primitive)) (emitted, 0)
    const float4* lights;
    uint32_t n_lights;
};

// ------------------------------------------------------------------ Vec3::rotated (vec3.rs:161-181)
// The reference's Euler map is linear; per instance the host stores, for each output row, the
// factors in the reference's own evaluation order:
//   x' = (x*r0)*r1 + y*r2 + z*r3            r = (cb, cc, sa*sb*cc - sa*cc, ca*sb*cc + sa*sc)
//   y' = (x*r4)*r5 + y*r6 + z*r7            r = (cb, sc, sa*sb*sc + ca*cc, ca*sb*sc - sa*cc)
//   z' = x*r8 + (y*r9)*r10 + (z*r11)*r10    r = (-sb, sa, cb, ca)
RTW_HD V3 rot_apply(const float* r, V3 p) {
    V3 o;
    o.x = p.x * r[0] * r[1] + p.y * r[2] + p.z * r[3];
    o.y = p.x * r[4] * r[5] + p.y * r[6] + p.z * r[7];
    o.z = p.x * r[8] + p.y * r[9] * r[10] + p.z * r[11] * r[10];
    return o;
}

// ------------------------------------------------------------------ Philox4x32-10
// Stream discipline (shared with the test oracle so renders can be compared sample for sample;
// the reference itself is unseeded):
//   key = seed; counter = (pixel index, global sample index, block, stream)
//   stream 0: camera block, lens-disk blocks, then per bounce one block (dielectric coin) or one
//             block per random_unit_vec rejection turn — consumed in path order;
//   stream 1+k: volume of instance k, block = (bounce << 16) | j, j = 0 free-flight xi,
//             j >= 1 rejection turns of the scatter normal.
struct U4 { uint32_t x, y, z, w; };
RTW_HD void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#if defined(__CUDA_ARCH__)
    hi = __umulhi(a, b); lo = a * b;
#else
    uint64_t p = (uint64_t)a * b; hi = (uint32_t)(p >> 32); lo = (uint32_t)p;
#endif
}
RTW_HD U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        mulhilo(0xD2511F53u, c0, hi0, lo0);
        mulhilo(0xCD9E8D57u, c2, hi1, lo1);
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    U4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3; return o;
}
// rand 0.8.5 Standard f32: 24 random bits in [0,1)
RTW_HD float u32_to_unit(uint32_t w) { return (float)(w >> 8) * (1.0f / 16777216.0f); }

struct PathRng {
    uint32_t k0, k1, pixel, sample, block;
    RTW_HD float4 at(uint32_t blk, uint32_t stream) const {
        U4 p = philox4x32_10(pixel, sample, blk, stream, k0, k1);
        return make_float4(u32_to_unit(p.x), u32_to_unit(p.y), u32_to_unit(p.z), u32_to_unit(p.w));
    }
    RTW_HD float4 next4() { return at(block++, 0u); }
};

// vec3.rs:216-238: Vec3::random(-1,1) -> reject outside the unit ball -> unit()
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_DIRECT_SPHERE)
// Fast build: the same distribution without the rejection loop. A point of U[-1,1]^3 kept inside the
// unit ball and normalised is uniform on the sphere; so is (sqrt(1-z^2) cos phi, sqrt(1-z^2) sin phi, z)
// with z = 1-2u, phi = 2 pi v. The loop makes a warp wait for its unluckiest lane (about five Philox
// blocks for 16 diffuse lanes against 1.9 on average per lane); one block always suffices here. The
// strict build keeps the reference's loop and with it the oracle's Philox stream, sample for sample.
__device__ __forceinline__ V3 sphere_point(float u, float v) {
    const float z = fmaf(-2.0f, u, 1.0f);
    const float r = sqrtf(fmaxf(0.0f, fmaf(-z, z, 1.0f)));
    float sn, cs;
    __sincosf(6.283185307179586f * (v - 0.5f), &sn, &cs);
    return v3(r * cs, r * sn, z);
}
RTW_HD V3 random_unit_vec(PathRng& rng) {
    const float4 r = rng.next4();
    return sphere_point(r.x, r.y);
}
#else
RTW_HD V3 random_unit_vec(PathRng& rng) {
    for (;;) {
        float4 r = rng.next4();
        V3 p = v3(r.x * (1.0f - -1.0f) + -1.0f, r.y * (1.0f - -1.0f) + -1.0f, r.z * (1.0f - -1.0f) + -1.0f);
        if (p.x * p.x + p.y * p.y + p.z * p.z <= 1.0f) return unit(p);
    }
}
#endif
// vec3.rs:240-254: two candidate points per block
RTW_HD V3 random_in_unit_disk(PathRng& rng) {
    for (;;) {
        float4 r = rng.next4();
        V3 p = v3(r.x * 2.0f - 1.0f, r.y * 2.0f - 1.0f, 0.f);
        if (p.x * p.x + p.y * p.y <= 1.0f) return p;
        V3 q = v3(r.z * 2.0f - 1.0f, r.w * 2.0f - 1.0f, 0.f);
        if (q.x * q.x + q.y * q.y <= 1.0f) return q;
    }
}

// ------------------------------------------------------------------ read-only loads
#if defined(__CUDA_ARCH__)
#define RTW_LDG(p) __ldg(p)
#else
#define RTW_LDG(p) (*(p))
#endif

}  // namespace rtw
