// rtw_trace.h — "extend": closest-hit search of one ray against the two-level BVH.
//
// Replaces Scene::collision_normal (Rust/src/viewport.rs:136-150) and the three recursive BVH
// walks below it (objects/aabb/aabb.rs:105-155, qaabb.rs:197-259, iaabb.rs:150-212). The reference
// visits both children of every node whose box the ray crosses, evaluates a full Hit (normal, UV,
// texel, Perlin) for every candidate and keeps the smallest t. Here the walk is iterative,
// front-to-back with the search interval shrunk to the best t so far, and only (t, primitive) is
// kept; the Hit of the winner is evaluated afterwards (rtw_shade.h). The closest hit is the same
// because
//   * every primitive is reached through its own reference leaf box (the child box stored in its
//     parent, or the level's root box), slab-tested with the reference's rule (objects.rs:125-140),
//     and inner boxes are supersets of it, exactly as in the reference's trees;
//   * moving spheres stay outside the tree: their t=0 leaf box is tested against the un-shrunk
//     interval;
//   * equal t is resolved by the lowest reference (class S < Q < I, then index), where the
//     reference's own order depends on its random build axis.
//
// The walk is ONE state machine over both levels (Trav): entering an instance pushes the world
// frame of the ray on the traversal stack and carries on in the same loop with the local ray, so
// that all 32 lanes of a warp execute the same node / leaf code whatever level each of them is in
// (the first version recursed into a second loop per instance and ran at 7 of 32 lanes active,
// profiles/r1a_*). The kernels drive it warp-synchronously (trav_step_node for every lane that
// sits on a node, then trav_step_other); trace() below is the scalar driver the host-side
// emulator (tests/emu) uses.
#pragma once
#include "rtw_device.h"
#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

#ifndef RTW_NODE_SORT_EXCHANGES
#define RTW_NODE_SORT_EXCHANGES 3   // C5 +1.5 %, C2 +1.9 %, C4 +1.1 % over the full network; node visits +0.6 % (profiles/r2j_ab_*)
#endif

namespace rtw {

struct Ray { V3 o, d; float time; };

// objects.rs:78-141 minf/maxf
RTW_HD float ref_minf(float a, float b) { return a <= b ? a : b; }
RTW_HD float ref_maxf(float a, float b) { return a >= b ? a : b; }

// Interval::intersect (objects.rs:125-140), literal: used when a direction component is zero.
RTW_HD bool axis_interval(float mn, float mx, float a, float b, float& lo, float& hi) {
    if (a == 0.0f) {
        if (mn < b && b < mx) { lo = -INFINITY; hi = INFINITY; return true; }
        return false;
    }
    float inv_a = 1.0f / a;
    float x1 = (mn - b) * inv_a, x2 = (mx - b) * inv_a;
    lo = ref_minf(x1, x2); hi = ref_maxf(x1, x2);
    return true;
}
RTW_HD_NOINL bool slab_slow(V3 lo, V3 hi, V3 o, V3 d, float mint, float maxt, float& tnear) {
    float x0, x1, y0, y1, z0, z1;
    if (!axis_interval(lo.x, hi.x, d.x, o.x, x0, x1)) return false;
    if (!axis_interval(lo.y, hi.y, d.y, o.y, y0, y1)) return false;
    if (!axis_interval(lo.z, hi.z, d.z, o.z, z0, z1)) return false;
    float mn = ref_maxf(ref_maxf(x0, y0), ref_maxf(z0, mint));
    float mx = ref_minf(ref_minf(x1, y1), ref_minf(z1, maxt));
    tnear = mn;
    return !(mn > mx);
}

// Sphere::collision_normal up to the accepted root (sphere.rs:100-123). `c` = centre at ray time.
RTW_HD bool sphere_t(V3 c, float radius, V3 o, V3 d, float a, float mint, float maxt, float& t) {
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_FASTPRIM)
    // same roots, branch-free: both candidates from one reciprocal, selected afterwards, so that the
    // lanes of a warp run one straight instruction sequence instead of four nested branches
    const V3 oc = o - c;
    const float b = dot(oc, d);
    const float cc = dot(oc, oc) - radius * radius;
    const float disc = b * b - a * cc;
    const float sq = fast_sqrt_pos(fmaxf(disc, 1e-30f));
    const float ia = fast_rcp(a);
    const float x1 = (-b - sq) * ia, x2 = (sq - b) * ia;
    const float x = x1 < mint ? x2 : x1;
    t = x;
    return disc >= 0.0f && x >= mint && x <= maxt;
#else
    V3 oc = o - c;
    float b = dot(oc, d);
    float cc = dot(oc, oc) - radius * radius;
    float disc = b * b - a * cc;
    if (disc < 0.0f) return false;
    float sq = sqrtf(disc);
    float x = (-b - sq) / a;
    if (x < mint) x = (-b + sq) / a;
    if (x < mint || x > maxt) return false;
    t = x;
    return true;
#endif
}

// Quad::collision_normal up to the inside test (quad.rs:38-60). `lim` bounds alfa + beta: 1 for a
// triangle record (Rust2/src/objects/triangle.rs:107-113: `alfa + beta > 1.0` rejects; its upper bounds
// on alfa and beta follow from the sum), 2 for a parallelogram, where it can never reject.
RTW_HD bool quad_t(const Quad* qp, V3 o, V3 d, float mint, float maxt, float lim, float& t) {
    float4 q0 = RTW_LDG(&qp->q0), q1 = RTW_LDG(&qp->q1), q2 = RTW_LDG(&qp->q2), q3 = RTW_LDG(&qp->q3);
    V3 n = v3(q0.x, q0.y, q0.z);
    float den = dot(n, d);
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_FASTPRIM)
    {   // the same tests, evaluated without early exits (one straight sequence per warp)
        const float tt = (q0.w - dot(n, o)) * fast_rcp(den);
        const V3 p = o + d * tt;
        const V3 planar = p - v3(q1.x, q1.y, q1.z);
        const V3 u = v3(q1.w, q2.x, q2.y), v = v3(q2.z, q2.w, q3.x), w = v3(q3.y, q3.z, q3.w);
        const float al = dot(w, cross(planar, v));
        const float be = dot(w, cross(u, planar));
        t = tt;
        return fabsf(den) > 1e-8f && tt >= mint && tt <= maxt && al >= 0.0f && al <= 1.0f && be >= 0.0f && be <= 1.0f && al + be <= lim;
    }
#endif
    if (fabsf(den) <= 1e-8f) return false;
    float tt = (q0.w - dot(n, o)) / den;
    if (tt < mint || tt > maxt) return false;
    V3 p = o + d * tt;
    V3 planar = p - v3(q1.x, q1.y, q1.z);
    V3 u = v3(q1.w, q2.x, q2.y), v = v3(q2.z, q2.w, q3.x), w = v3(q3.y, q3.z, q3.w);
    float al = dot(w, cross(planar, v));
    float be = dot(w, cross(u, planar));
    if (al < 0.0f || al > 1.0f || be < 0.0f || be > 1.0f) return false;
    if (al + be > lim) return false;
    t = tt;
    return true;
}

// Result of a search. At top level an instance hit is (KIND_INSTANCE, k) + the primitive inside.
struct Best {
    float t;
    uint32_t ref;     // REF_NONE = miss
    uint32_t inner;
};
RTW_HD bool closer(float t, uint32_t ref, float bt, uint32_t bref) { return t < bt || (t == bt && ref < bref); }

// Algorithmic work of a traversal, counted only by the instrumented kernel build
// (-DRTW_COUNT_WORK, RTW_FLAG_COUNT_WORK): the per-ray figures bench.py's roofline uses.
struct Work { uint32_t nodes, spheres, quads, instances, hits; };
#if defined(RTW_COUNT_WORK)
#define RTW_COUNT(cx, field) (++(cx).work->field)
#else
#define RTW_COUNT(cx, field) ((void)0)
#endif

// What the search needs besides the ray: the scene, the path's RNG identity (the volume free
// flight of instance.rs:24-26 draws from a per-instance Philox stream) and the integrator's interval.
struct TraceCtx {
    const Scene* sc;
    uint32_t k0, k1, pixel, sample;   // Philox key and counter words 0/1 of this path
    uint32_t bounce;
    float mint, maxt;                 // the integrator's interval (un-shrunk)
    Work* work;                       // null unless RTW_COUNT_WORK
};
RTW_HD float4 ctx_rand(const TraceCtx& cx, uint32_t blk, uint32_t stream) {
    U4 p = philox4x32_10(cx.pixel, cx.sample, blk, stream, cx.k0, cx.k1);
    return make_float4(u32_to_unit(p.x), u32_to_unit(p.y), u32_to_unit(p.z), u32_to_unit(p.w));
}

// random_unit_vec (vec3.rs:216-238) drawn from a side stream: block = base | turn, turn = 1, 2, ...
RTW_HD V3 random_unit_vec_stream(const TraceCtx& cx, uint32_t base_block, uint32_t stream) {
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_DIRECT_SPHERE)
    {
        const float4 r = ctx_rand(cx, base_block | 1u, stream);
        return sphere_point(r.x, r.y);
    }
#endif
    for (uint32_t turn = 0;; ++turn) {
        float4 r = ctx_rand(cx, base_block | (turn + 1u), stream);
        V3 p = v3(r.x * (1.0f - -1.0f) + -1.0f, r.y * (1.0f - -1.0f) + -1.0f, r.z * (1.0f - -1.0f) + -1.0f);
        if (p.x * p.x + p.y * p.y + p.z * p.z <= 1.0f) return unit(p);
    }
}

// The traversal stack: per-thread column of 32-bit words. On the device it lives in shared memory,
// word w of thread t at base[w * stride] with base = &smem[t], stride = block size (conflict-free);
// the host emulator passes a plain array (stride 1). Depth needed: see stack_words().
#ifndef RTW_STACK_ADDR
#define RTW_STACK_ADDR 1   // device builds: Trav::sp is the shared-memory ADDRESS of the next free word, so a push is one store + one add
                           // instead of address arithmetic (shift-add) + store + increment: 3.5 % fewer instructions in wf_extend,
                           // +2.0 % on C5, +1.7 % on C2, +1.2 % on C4 (profiles/r2m_ab_stack_address.txt); 0 = word index
#endif
struct Stack {
    uint32_t* base;
    uint32_t stride;
#if defined(__CUDA_ARCH__) && RTW_STACK_ADDR
    __device__ __forceinline__ uint32_t origin() const { return (uint32_t)__cvta_generic_to_shared(base); }
    __device__ __forceinline__ void push(uint32_t& sp, uint32_t v) const {
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(sp), "r"(v));
        sp += stride * 4u;
    }
    __device__ __forceinline__ uint32_t pop(uint32_t& sp) const {
        sp -= stride * 4u;
        uint32_t v;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(sp));
        return v;
    }
#else
    RTW_HD uint32_t origin() const { return 0u; }
    RTW_HD void push(uint32_t& sp, uint32_t v) const { base[sp * stride] = v; ++sp; }
    RTW_HD uint32_t pop(uint32_t& sp) const { --sp; return base[sp * stride]; }
#endif
};
// bottom sentinel + top pushes + world frame (10) + sentinel + probe frame (2 + sentinel) + inner pushes
#define RTW_STACK 32   // deepest node chain the builder allows per level (scene_build.cpp)
RTW_HD uint32_t stack_words(uint32_t top_depth, uint32_t inner_depth) { return top_depth + inner_depth + 16u; }
// ... and when no instance is walked as a level (only primitives, baked copies and one-sphere records, which are
// resolved in place): the bottom sentinel and the top level's pushes
RTW_HD uint32_t stack_words_flat(uint32_t top_depth) { return top_depth + 4u; }

static const uint32_t CUR_DONE = 0xFFFFFFFFu;   // nothing left to do
static const uint32_t CUR_EXIT = 0xFFFFFFFEu;   // sentinel: the level being walked is exhausted
static const uint32_t NODE_LIMIT = 0x40000000u; // refs below this are internal nodes (KIND_NODE == 0)

enum : uint32_t { TRAV_TOP = 0u, TRAV_INNER = 1u, TRAV_PROBE = 2u };

// One ray's search state.
struct Trav {
    V3 o, d;              // the ray in the frame being walked (world, or instance-local)
    V3 inv;               // 1/d per axis (objects.rs:136)
    float a;              // d.d (sphere.rs:103)
    float t;              // best t of the level being walked == upper end of its search interval
    uint32_t ref;         // ... and its primitive (REF_NONE: none yet)
    float top_t;          // while inside an instance: the top level's best
    uint32_t top_ref;
    uint32_t inner;       // the top level's best's primitive inside its instance
    uint32_t cur;         // next reference to process, CUR_EXIT or CUR_DONE
    uint32_t inst;        // instance being walked (reference), REF_NONE at top level
    uint32_t sp;
    uint32_t mode;        // TRAV_*  | 4 when some 1/d is not finite (literal slab rule)
};
RTW_HD bool trav_slow(const Trav& T) { return (T.mode & 4u) != 0u; }
RTW_HD void trav_set_ray(Trav& T, V3 o, V3 d) {
    T.o = o; T.d = d;
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_FASTINV)
    T.inv = v3(fast_rcp(d.x), fast_rcp(d.y), fast_rcp(d.z));    // a zero component gives NaN here, inf below: both take the slow slab
#else
    T.inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
#endif
    T.a = dot(d, d);
    bool slow = !(isfinite(T.inv.x) && isfinite(T.inv.y) && isfinite(T.inv.z));
    T.mode = (T.mode & 3u) | (slow ? 4u : 0u);
}

// The slab test of AABB/QuadAABB/IAABB::collision_normal (aabb.rs:109-126): miss iff
// max(tx0,ty0,tz0,mint) > min(tx1,ty1,tz1,maxt). With finite 1/d no NaN can arise, so
// fminf/fmaxf equal the reference's minf/maxf.
RTW_HD bool slab_ray(V3 lo, V3 hi, V3 o, V3 d, V3 inv, bool slow, float mint, float maxt, float& tnear) {
    if (slow) return slab_slow(lo, hi, o, d, mint, maxt, tnear);
    float x1 = (lo.x - o.x) * inv.x, x2 = (hi.x - o.x) * inv.x;
    float y1 = (lo.y - o.y) * inv.y, y2 = (hi.y - o.y) * inv.y;
    float z1 = (lo.z - o.z) * inv.z, z2 = (hi.z - o.z) * inv.z;
    float mn = fmaxf(fmaxf(fminf(x1, x2), fminf(y1, y2)), fmaxf(fminf(z1, z2), mint));
    float mx = fminf(fminf(fmaxf(x1, x2), fmaxf(y1, y2)), fminf(fmaxf(z1, z2), maxt));
    tnear = mn;
    return !(mn > mx);
}
RTW_HD bool slab(V3 lo, V3 hi, const Trav& T, float mint, float maxt, float& tnear) {
    return slab_ray(lo, hi, T.o, T.d, T.inv, trav_slow(T), mint, maxt, tnear);
}

// Moving spheres of a level: t=0 leaf box against the un-shrunk interval (aabb.rs:27-38,109-126),
// then the sphere at its position at ray time (sphere.rs:100).
RTW_HD void test_moving(const TraceCtx& cx, const Level& lv, float time, Trav& T) {
    const Scene& sc = *cx.sc;
    for (uint32_t m = 0; m < lv.mov_count; ++m) {
        RTW_COUNT(cx, spheres);
        uint32_t idx = RTW_LDG(&sc.moving[lv.mov_first + m]);
        float4 lo = RTW_LDG(&sc.sphere_box[2 * idx]), hi = RTW_LDG(&sc.sphere_box[2 * idx + 1]);
        float tn;
        if (!slab(v3(lo.x, lo.y, lo.z), v3(hi.x, hi.y, hi.z), T, cx.mint, cx.maxt, tn)) continue;
        float4 s = RTW_LDG(&sc.spheres[idx]);
        float4 vel = RTW_LDG(&sc.sphere_vel[idx]);
        V3 c = v3(s.x, s.y, s.z) + v3(vel.x, vel.y, vel.z) * time;
        float t;
        uint32_t ref = make_ref(KIND_SPHERE, idx);
        if (sphere_t(c, s.w, T.o, T.d, T.a, cx.mint, T.t, t) && closer(t, ref, T.t, T.ref)) { T.t = t; T.ref = ref; }
    }
}

// Start walking a level with the ray currently in T: its moving spheres, then its root box
// (the reference slab-tests every BVH root, aabb.rs:109-126; an empty BVH is a miss, aabb.rs:106).
// The level's sentinel must already be on the stack.
RTW_HD void trav_begin_level(const TraceCtx& cx, const Level& lv, float time, Trav& T, const Stack& S) {
    if (lv.mov_count) {
        test_moving(cx, lv, time, T);
        if ((T.mode & 3u) == TRAV_PROBE && T.ref != REF_NONE) { T.cur = S.pop(T.sp); return; }
    }
    float tn;
    if (lv.root == REF_NONE ||
        !slab(v3(lv.lo[0], lv.lo[1], lv.lo[2]), v3(lv.hi[0], lv.hi[1], lv.hi[2]), T, cx.mint, T.t, tn)) {
        T.cur = S.pop(T.sp);
        return;
    }
    T.cur = lv.root;
}

// Scene::collision_normal (viewport.rs:136-150): start the search of `r`.
RTW_HD void trav_begin(const TraceCtx& cx, const Ray& r, Trav& T, const Stack& S) {
    T.mode = TRAV_TOP;
    trav_set_ray(T, r.o, r.d);
    T.t = cx.maxt; T.ref = REF_NONE; T.inner = REF_NONE;
    T.top_t = 0.0f; T.top_ref = REF_NONE;
    T.inst = REF_NONE; T.sp = S.origin();
    S.push(T.sp, CUR_DONE);               // bottom of the stack: popping it ends the search
    trav_begin_level(cx, cx.sc->top, r.time, T, S);
}

// One static sphere / quad leaf against the level's best so far.
RTW_HD void trav_test_prim(const TraceCtx& cx, uint32_t ref, Trav& T) {
    const Scene& sc = *cx.sc;
    const uint32_t idx = ref_index(ref);
    float t;
    bool hit;
    if (ref_kind(ref) == KIND_SPHERE) {
        RTW_COUNT(cx, spheres);
        float4 s = RTW_LDG(&sc.spheres[idx]);
        hit = sphere_t(v3(s.x, s.y, s.z), s.w, T.o, T.d, T.a, cx.mint, T.t, t);
    } else {
        RTW_COUNT(cx, quads);
        hit = quad_t(&sc.quads[idx], T.o, T.d, cx.mint, T.t, (ref & REF_TRIANGLE) ? 1.0f : 2.0f, t);
    }
    if (hit && closer(t, ref, T.t, T.ref)) { T.t = t; T.ref = ref; }
}
RTW_HD bool ref_is_prim(uint32_t r) { return (r - NODE_LIMIT) < 0x80000000u; }   // KIND_SPHERE or KIND_QUAD
RTW_HD bool ref_is_one_sphere_instance(uint32_t r) { return r >= 0xE0000000u && r < CUR_EXIT; }
// what trav_step_prim handles: sphere, quad, one-sphere instance
RTW_HD bool ref_is_leaf_test(uint32_t r) { return ref_is_prim(r) || ref_is_one_sphere_instance(r); }

// Slab tests of the four child boxes of node `idx`: tn[c] = entry parameter if child c is crossed
// within [mint, maxt], else +inf; rr = the child references.
//   strict build / host: exact f32 boxes, the reference's own operation order, (bound - o) * (1/d);
//   fast device build: half-float boxes (Node16, rounded outwards) and one FFMA per plane,
//   bound * (1/d) - o * (1/d).
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_SIGN_SELECT)
// Fast device build: the ray's direction signs pick each axis' entry and exit planes while they are still packed
// halves (12 selects per node), so that a child costs max3 / min3 + the interval clamp instead of six two-way
// min / max. Same values as min(x1, x2) / max(x1, x2): the FFMA is monotone in the plane coordinate.
__device__ __forceinline__ void slab4_signed(const uint4 a, const uint4 b, const uint4 c4, const Trav& T, float mint,
                                             float maxt, float tn[4]) {
    const bool sx = T.inv.x < 0.0f, sy = T.inv.y < 0.0f, sz = T.inv.z < 0.0f;
    const uint32_t nx01 = sx ? b.z : a.x, nx23 = sx ? b.w : a.y, fx01 = sx ? a.x : b.z, fx23 = sx ? a.y : b.w;
    const uint32_t ny01 = sy ? c4.x : a.z, ny23 = sy ? c4.y : a.w, fy01 = sy ? a.z : c4.x, fy23 = sy ? a.w : c4.y;
    const uint32_t nz01 = sz ? c4.z : b.x, nz23 = sz ? c4.w : b.y, fz01 = sz ? b.x : c4.z, fz23 = sz ? b.y : c4.w;
    const float ox = -T.o.x * T.inv.x, oy = -T.o.y * T.inv.y, oz = -T.o.z * T.inv.z;
    float NX[4], NY[4], NZ[4], FX[4], FY[4], FZ[4];
    float2 f;
    f = __half22float2(*reinterpret_cast<const __half2*>(&nx01)); NX[0] = f.x; NX[1] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&nx23)); NX[2] = f.x; NX[3] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&ny01)); NY[0] = f.x; NY[1] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&ny23)); NY[2] = f.x; NY[3] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&nz01)); NZ[0] = f.x; NZ[1] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&nz23)); NZ[2] = f.x; NZ[3] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&fx01)); FX[0] = f.x; FX[1] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&fx23)); FX[2] = f.x; FX[3] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&fy01)); FY[0] = f.x; FY[1] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&fy23)); FY[2] = f.x; FY[3] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&fz01)); FZ[0] = f.x; FZ[1] = f.y;
    f = __half22float2(*reinterpret_cast<const __half2*>(&fz23)); FZ[2] = f.x; FZ[3] = f.y;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float x1 = fmaf(NX[c], T.inv.x, ox), y1 = fmaf(NY[c], T.inv.y, oy), z1 = fmaf(NZ[c], T.inv.z, oz);
        const float x2 = fmaf(FX[c], T.inv.x, ox), y2 = fmaf(FY[c], T.inv.y, oy), z2 = fmaf(FZ[c], T.inv.z, oz);
        const float mn = fmaxf(fmaxf(fmaxf(x1, y1), z1), mint);
        const float mx = fminf(fminf(fminf(x2, y2), z2), maxt);
        tn[c] = (mn > mx) ? INFINITY : mn;
    }
}
#endif

RTW_HD void slab4(const Scene& sc, uint32_t idx, const Trav& T, float mint, float maxt, float tn[4], uint4& rr) {
#if defined(RTW_FAST_NODES)
    const Node16* np = &sc.nodes16[idx];
    const uint4 a = RTW_LDG(&np->a), b = RTW_LDG(&np->b), c4 = RTW_LDG(&np->c);
    rr = RTW_LDG(&np->ref);
#if !defined(RTW_NO_SIGN_SELECT)
    if (!trav_slow(T)) { slab4_signed(a, b, c4, T, mint, maxt, tn); return; }
#endif
    float LX[4], LY[4], LZ[4], HX[4], HY[4], HZ[4];
    {
        float2 f;
        f = __half22float2(*reinterpret_cast<const __half2*>(&a.x)); LX[0] = f.x; LX[1] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&a.y)); LX[2] = f.x; LX[3] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&a.z)); LY[0] = f.x; LY[1] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&a.w)); LY[2] = f.x; LY[3] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&b.x)); LZ[0] = f.x; LZ[1] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&b.y)); LZ[2] = f.x; LZ[3] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&b.z)); HX[0] = f.x; HX[1] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&b.w)); HX[2] = f.x; HX[3] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&c4.x)); HY[0] = f.x; HY[1] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&c4.y)); HY[2] = f.x; HY[3] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&c4.z)); HZ[0] = f.x; HZ[1] = f.y;
        f = __half22float2(*reinterpret_cast<const __half2*>(&c4.w)); HZ[2] = f.x; HZ[3] = f.y;
    }
#else
    const Node* np = &sc.nodes[idx];
    const float4 lox = RTW_LDG(&np->lox), loy = RTW_LDG(&np->loy), loz = RTW_LDG(&np->loz);
    const float4 hix = RTW_LDG(&np->hix), hiy = RTW_LDG(&np->hiy), hiz = RTW_LDG(&np->hiz);
    rr = RTW_LDG(&np->ref);
    const float LX[4] = {lox.x, lox.y, lox.z, lox.w}, LY[4] = {loy.x, loy.y, loy.z, loy.w}, LZ[4] = {loz.x, loz.y, loz.z, loz.w};
    const float HX[4] = {hix.x, hix.y, hix.z, hix.w}, HY[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, HZ[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
#endif
    if (trav_slow(T)) {
        for (int c = 0; c < 4; ++c) {
            float t;
            bool h = slab_slow(v3(LX[c], LY[c], LZ[c]), v3(HX[c], HY[c], HZ[c]), T.o, T.d, mint, maxt, t);
            tn[c] = h ? t : INFINITY;
        }
        return;
    }
#if defined(RTW_FAST_NODES)
    const float nx = -T.o.x * T.inv.x, ny = -T.o.y * T.inv.y, nz = -T.o.z * T.inv.z;
#endif
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int c = 0; c < 4; ++c) {
#if defined(RTW_FAST_NODES)
        float x1 = fmaf(LX[c], T.inv.x, nx), x2 = fmaf(HX[c], T.inv.x, nx);
        float y1 = fmaf(LY[c], T.inv.y, ny), y2 = fmaf(HY[c], T.inv.y, ny);
        float z1 = fmaf(LZ[c], T.inv.z, nz), z2 = fmaf(HZ[c], T.inv.z, nz);
#else
        float x1 = (LX[c] - T.o.x) * T.inv.x, x2 = (HX[c] - T.o.x) * T.inv.x;
        float y1 = (LY[c] - T.o.y) * T.inv.y, y2 = (HY[c] - T.o.y) * T.inv.y;
        float z1 = (LZ[c] - T.o.z) * T.inv.z, z2 = (HZ[c] - T.o.z) * T.inv.z;
#endif
        float mn = fmaxf(fmaxf(fminf(x1, x2), fminf(y1, y2)), fmaxf(fminf(z1, z2), mint));
        float mx = fminf(fminf(fmaxf(x1, x2), fmaxf(y1, y2)), fminf(fmaxf(z1, z2), maxt));
        tn[c] = (mn > mx) ? INFINITY : mn;
    }
}

RTW_HD void cswap(float& ta, uint32_t& ra, float& tb, uint32_t& rb) {
    const bool sw = tb < ta;
    const float t = sw ? tb : ta; const uint32_t r = sw ? rb : ra;
    tb = sw ? ta : tb; rb = sw ? ra : rb;
    ta = t; ra = r;
}

// T.cur is an internal node: four child boxes. The crossed children are visited nearest first, the
// others stacked farthest first. INLINE_PRIMS: sphere / quad children whose box is crossed are
// tested on the spot (they are leaves: nothing to descend into); otherwise they are visited like
// any other child (trav_step_prim), which lets a warp batch them.
template <bool INLINE_PRIMS>
RTW_HD void trav_step_node(const TraceCtx& cx, Trav& T, const Stack& S) {
    RTW_COUNT(cx, nodes);
    uint4 rr;
    float tn[4];
    slab4(*cx.sc, T.cur, T, cx.mint, T.t, tn, rr);
    uint32_t rf[4] = {rr.x, rr.y, rr.z, rr.w};
    if (INLINE_PRIMS) {
        bool any = false;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int c = 0; c < 4; ++c)
            if (tn[c] < INFINITY && ref_is_prim(rf[c])) { trav_test_prim(cx, rf[c], T); tn[c] = INFINITY; any = true; }
        if (any && (T.mode & 3u) == TRAV_PROBE && T.ref != REF_NONE) {   // instance.rs:284-294 only asks "anything?"
            while (S.pop(T.sp) != CUR_EXIT) {}
            T.cur = CUR_EXIT;
            return;
        }
    }
    // -DRTW_LATE_VOLUMES: constant-density volumes (REF_LATE_INSTANCE) go last whatever their distance
    // (their hit costs a free-flight draw, which the shrunk interval of a nearer surface hit usually
    // spares). It paid while C5's global fog was walked by every ray; with enclosing volumes dropped
    // from the top level (rtw_host.h) the four extra tests per node step cost more than they save
    // (+4 % without, profiles/r1h_ab_fastvec_nolate.txt).
#if defined(RTW_LATE_VOLUMES)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int c = 0; c < 4; ++c)
        if ((rf[c] & 0xD0000000u) == 0xD0000000u && tn[c] < INFINITY) tn[c] = 3.0e38f;
#endif
    // RTW_NODE_SORT_EXCHANGES (fast device build; A/B): 5 = the full network (nearest first, the rest stacked farthest first);
    // 3 = the nearest child only is brought to the front (the others are stacked in whatever partial order that leaves), 4 = also
    // the farthest to the back. Any order is correct - crossed children all get visited; the order only decides how soon
    // the search interval shrinks - and an exchange is five ALU-pipe instructions (FSETP + 2 SEL + 2 FSEL) in a kernel whose
    // busiest pipe is the ALU.
    cswap(tn[0], rf[0], tn[1], rf[1]); cswap(tn[2], rf[2], tn[3], rf[3]);
    cswap(tn[0], rf[0], tn[2], rf[2]);
#if !defined(RTW_FAST_NODES) || RTW_NODE_SORT_EXCHANGES >= 4
    cswap(tn[1], rf[1], tn[3], rf[3]);
#endif
#if !defined(RTW_FAST_NODES) || RTW_NODE_SORT_EXCHANGES >= 5
    cswap(tn[1], rf[1], tn[2], rf[2]);
#endif
    if (tn[3] < INFINITY) S.push(T.sp, rf[3]);
    if (tn[2] < INFINITY) S.push(T.sp, rf[2]);
    if (tn[1] < INFINITY) S.push(T.sp, rf[1]);
    T.cur = tn[0] < INFINITY ? rf[0] : S.pop(T.sp);
}

// const_density (instance.rs:24-26): ln(xi) / -density, xi from the instance's volume stream.
RTW_HD float volume_distance(const Instance& in, const TraceCtx& cx) {
    float xi = ctx_rand(cx, (cx.bounce << 16) | 0u, 1u + in.index).x;
    return logf(xi) / -in.density;
}

// Instance::collision_normal for an instance that holds exactly one static sphere (C5's two
// volumes; Instance::new_sphere of one sphere): the same steps as the general walk — frame change,
// root (= leaf) box, sphere, constant-density probe (instance.rs:250-302) — without touching the
// traversal state.
RTW_HD void trav_one_sphere_instance(const TraceCtx& cx, uint32_t iref, Trav& T) {
    const uint32_t k = instance_index(iref);
    const float4* rec = cx.sc->one_sphere + 4u * k;       // one 64-B record: no dependent loads
    const float4 r0 = RTW_LDG(rec), s = RTW_LDG(rec + 1), r2 = RTW_LDG(rec + 2), r3 = RTW_LDG(rec + 3);
    const uint32_t flags = f2u(r2.w);
    V3 o = T.o - v3(r0.x, r0.y, r0.z);
    V3 d = T.d, inv = T.inv;
    float a = T.a;
    bool slow = trav_slow(T);
    if (!(flags & INST_ROT_IDENTITY)) {
        const Instance& in = cx.sc->instances[k];
        o = rot_apply(in.inv_rows, o); d = rot_apply(in.inv_rows, d);
        inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
        a = dot(d, d);
        slow = !(isfinite(inv.x) && isfinite(inv.y) && isfinite(inv.z));
    }
    const V3 lo = v3(r2.x, r2.y, r2.z), hi = v3(r3.x, r3.y, r3.z);
    const V3 c = v3(s.x, s.y, s.z);
    float tn, t;
#if defined(RTW_FAST_NODES) && !defined(RTW_NO_VOLSKIP)
    // A ray that starts inside a constant-density sphere meets the exit face first; the probe then
    // starts on or beyond that face heading outwards and finds nothing, so the reference returns
    // None (instance.rs:280-300, SURVEY N14: only rays entering from outside can scatter). The
    // fast build takes that exit before the square root, the free-flight draw and the logarithm;
    // the strict build walks the literal sequence.
    if ((flags & INST_VOLUME) && r0.w > 0.0f) {
        const V3 oc = o - c;
        if (dot(oc, oc) - s.w * s.w < 0.0f) return;
    }
#endif
    // (the fast build leaves out the two box tests: the box holds nothing but the sphere, whose own test
    // has the same interval, and the parent node has just tested this box)
#if !defined(RTW_FAST_NODES) || defined(RTW_NO_VOLSKIP)
    if (!slab_ray(lo, hi, o, d, inv, slow, cx.mint, T.t, tn)) return;
#endif
    RTW_COUNT(cx, spheres);
    if (!sphere_t(c, s.w, o, d, a, cx.mint, T.t, t)) return;
    if (flags & INST_VOLUME) {
        // const_density (instance.rs:24-26), as volume_distance()
        float xi = ctx_rand(cx, (cx.bounce << 16) | 0u, 1u + f2u(r3.w)).x;
        float distance = logf(xi) / -r0.w;
        if (distance >= 0.0f) {
            V3 o2 = (o + d * t) + d * distance;
            float t2;
#if !defined(RTW_FAST_NODES) || defined(RTW_NO_VOLSKIP)
            if (!slab_ray(lo, hi, o2, d, inv, slow, cx.mint, cx.maxt, tn)) return;
#endif
            RTW_COUNT(cx, spheres);
            if (!sphere_t(c, s.w, o2, d, a, cx.mint, cx.maxt, t2)) return;
        }
    }
    if (closer(t, iref, T.t, T.ref)) { T.t = t; T.ref = iref; T.inner = cx.sc->instances[k].level.root; }
}

// Instance::collision_normal, first half (instance.rs:250-271): ray -> local frame, walk the inside.
RTW_HD void trav_enter_instance(const TraceCtx& cx, float time, Trav& T, const Stack& S) {
    RTW_COUNT(cx, instances);
    const Instance& in = cx.sc->instances[instance_index(T.cur)];
    // world frame of the ray: restored by trav_level_finished
    S.push(T.sp, f2u(T.o.x)); S.push(T.sp, f2u(T.o.y)); S.push(T.sp, f2u(T.o.z));
    S.push(T.sp, f2u(T.d.x)); S.push(T.sp, f2u(T.d.y)); S.push(T.sp, f2u(T.d.z));
    S.push(T.sp, f2u(T.inv.x)); S.push(T.sp, f2u(T.inv.y)); S.push(T.sp, f2u(T.inv.z));
    S.push(T.sp, f2u(T.a));
    S.push(T.sp, CUR_EXIT);
    V3 o = T.o - v3(in.t[0], in.t[1], in.t[2]);
    T.inst = T.cur & ~REF_INSTANCE_FLAGS;
    T.top_t = T.t; T.top_ref = T.ref;     // T.t stays: the inner search starts from the same bound
    T.ref = REF_NONE;
    if (in.flags & INST_ROT_IDENTITY) {
        T.o = o;                           // same direction: inv, a and the slow flag stand
        T.mode = (T.mode & 4u) | TRAV_INNER;
    } else {
        T.mode = TRAV_INNER;
        trav_set_ray(T, rot_apply(in.inv_rows, o), rot_apply(in.inv_rows, T.d));
    }
    trav_begin_level(cx, in.level, time, T, S);
}

// The sentinel of the level being walked came off the stack.
//   top level:  the search is over.
//   instance:   instance.rs:272-302 — no inner hit => miss; a constant-density volume re-traverses
//               from a point `distance` behind the entry face and only counts if that probe hits
//               anything (the entry-face t is kept); then back to the world frame.
RTW_HD void trav_level_finished(const TraceCtx& cx, float time, Trav& T, const Stack& S) {
    const uint32_t mode = T.mode & 3u;
    if (mode == TRAV_TOP) { T.cur = CUR_DONE; return; }
    const Instance& in = cx.sc->instances[instance_index(T.inst)];
    bool hit;
    if (mode == TRAV_INNER) {
        hit = T.ref != REF_NONE;
        if (hit && (in.flags & INST_VOLUME)) {
            float distance = volume_distance(in, cx);
            if (distance >= 0.0f) {
                S.push(T.sp, f2u(T.t)); S.push(T.sp, T.ref); S.push(T.sp, CUR_EXIT);
                T.o = (T.o + T.d * T.t) + T.d * distance;       // same direction: inv, a unchanged
                T.t = cx.maxt; T.ref = REF_NONE;
                T.mode = (T.mode & 4u) | TRAV_PROBE;
                trav_begin_level(cx, in.level, time, T, S);
                return;
            }
        }
    } else {
        hit = T.ref != REF_NONE;                                  // the probe found something
        T.ref = S.pop(T.sp); T.t = u2f(S.pop(T.sp));              // the inner closest hit again
    }
    // leave the instance
    T.a = u2f(S.pop(T.sp));
    T.inv.z = u2f(S.pop(T.sp)); T.inv.y = u2f(S.pop(T.sp)); T.inv.x = u2f(S.pop(T.sp));
    T.d.z = u2f(S.pop(T.sp)); T.d.y = u2f(S.pop(T.sp)); T.d.x = u2f(S.pop(T.sp));
    T.o.z = u2f(S.pop(T.sp)); T.o.y = u2f(S.pop(T.sp)); T.o.x = u2f(S.pop(T.sp));
    bool slow = !(isfinite(T.inv.x) && isfinite(T.inv.y) && isfinite(T.inv.z));
    T.mode = TRAV_TOP | (slow ? 4u : 0u);
    if (hit && closer(T.t, T.inst, T.top_t, T.top_ref)) { T.inner = T.ref; T.ref = T.inst; }
    else { T.t = T.top_t; T.ref = T.top_ref; }
    T.inst = REF_NONE;
    T.cur = S.pop(T.sp);
}

// T.cur is a leaf test: a sphere / quad (a child left on the stack by the non-inlining node step,
// or the root of a one-primitive level) or an instance that holds exactly one sphere.
RTW_HD void trav_step_prim(const TraceCtx& cx, Trav& T, const Stack& S) {
    if (ref_is_one_sphere_instance(T.cur)) {
        RTW_COUNT(cx, instances);
        const uint32_t iref = T.cur & ~REF_INSTANCE_FLAGS;
        trav_one_sphere_instance(cx, iref, T);
        T.cur = S.pop(T.sp);
        return;
    }
    trav_test_prim(cx, T.cur, T);
    if ((T.mode & 3u) == TRAV_PROBE && T.ref != REF_NONE) {     // instance.rs:284-294 only asks "anything?"
        while (S.pop(T.sp) != CUR_EXIT) {}
        T.cur = CUR_EXIT;
    } else {
        T.cur = S.pop(T.sp);
    }
}

// The same for a leaf test that was taken off the stack earlier (`ref`), T.cur being whatever came
// off the stack after it (wf_extend parks leaf tests so that its lanes keep taking node steps).
RTW_HD void trav_test_parked(const TraceCtx& cx, uint32_t ref, Trav& T, const Stack& S) {
    if (ref_is_one_sphere_instance(ref)) {
        RTW_COUNT(cx, instances);
        trav_one_sphere_instance(cx, ref & ~REF_INSTANCE_FLAGS, T);
        return;
    }
    trav_test_prim(cx, ref, T);
    if ((T.mode & 3u) == TRAV_PROBE && T.ref != REF_NONE && T.cur != CUR_EXIT) {   // instance.rs:284-294 only asks "anything?"
        while (S.pop(T.sp) != CUR_EXIT) {}
        T.cur = CUR_EXIT;
    }
}

// T.cur is an instance leaf (a level change) or the sentinel. Such events are chained until the
// lane is back on a node or a leaf test (or done).
// One-sphere instances (C5's volumes: frame change, box, sphere, free-flight probe — some 200 instructions) run
// in the event steps rather than the leaf steps, which keeps the leaf steps to the two short tests (+0.7 % on C5
// with the event weight at 8, profiles/r1l_ab_one_sphere_event.txt). -DRTW_ONE_SPHERE_EVENT=0 puts them back.
#ifndef RTW_ONE_SPHERE_EVENT
#define RTW_ONE_SPHERE_EVENT 1
#endif
#if RTW_ONE_SPHERE_EVENT
RTW_HD bool ref_is_event(uint32_t r) { return r >= 0xC0000000u && r != CUR_DONE; }
#else
RTW_HD bool ref_is_event(uint32_t r) { return (r >= 0xC0000000u && r < 0xE0000000u) || r == CUR_EXIT; }
#endif
RTW_HD void trav_step_event(const TraceCtx& cx, float time, Trav& T, const Stack& S) {
    do {
        if (T.cur == CUR_EXIT) trav_level_finished(cx, time, T, S);
#if RTW_ONE_SPHERE_EVENT
        else if (ref_is_one_sphere_instance(T.cur)) {
            RTW_COUNT(cx, instances);
            trav_one_sphere_instance(cx, T.cur & ~REF_INSTANCE_FLAGS, T);
            T.cur = S.pop(T.sp);
        }
#endif
        else trav_enter_instance(cx, time, T, S);
    } while (ref_is_event(T.cur));
}

// Anything but an internal node, in one call (drivers that do not schedule step kinds separately).
RTW_HD void trav_step_other(const TraceCtx& cx, float time, Trav& T, const Stack& S) {
    do {
        if (ref_is_leaf_test(T.cur)) trav_step_prim(cx, T, S);
        else trav_step_event(cx, time, T, S);
    } while (T.cur >= NODE_LIMIT && T.cur != CUR_DONE);
}

// The result of a finished search. A hit on the world-space copy of a baked instance primitive (rtw_host.h) is
// reported as what it stands for - (instance, primitive inside it) - so the Hit is evaluated through the
// instance's frame exactly as for a walked instance (instance.rs:257-307).
RTW_HD Best trav_result(const Scene& sc, const Trav& T) {
    Best b; b.t = T.t; b.ref = T.ref; b.inner = T.inner;
    const uint32_t kind = ref_kind(b.ref);
    if (kind == KIND_SPHERE && ref_index(b.ref) >= sc.n_spheres_orig) {
        const uint2 src = RTW_LDG(&sc.baked_sphere_src[ref_index(b.ref) - sc.n_spheres_orig]);
        b.ref = src.x; b.inner = src.y;
    } else if (kind == KIND_QUAD && ref_index(b.ref) >= sc.n_quads_orig) {
        const uint2 src = RTW_LDG(&sc.baked_quad_src[ref_index(b.ref) - sc.n_quads_orig]);
        b.ref = src.x; b.inner = src.y;
    }
    return b;
}

// Scalar driver (host emulator, and any caller that does not schedule the steps itself).
RTW_HD Best trace(const TraceCtx& cx, const Ray& r, const Stack& S) {
    Trav T;
    trav_begin(cx, r, T, S);
    while (T.cur != CUR_DONE) {
        if (T.cur < NODE_LIMIT) trav_step_node<true>(cx, T, S);
        else trav_step_other(cx, r.time, T, S);
    }
    return trav_result(*cx.sc, T);
}

// Instance::collision_normal's frame change for the Hit evaluation (instance.rs:257-258).
RTW_HD Ray to_local(const Instance& in, const Ray& r) {
    Ray l;
    V3 o = r.o - v3(in.t[0], in.t[1], in.t[2]);
    if (in.flags & INST_ROT_IDENTITY) { l.o = o; l.d = r.d; }
    else { l.o = rot_apply(in.inv_rows, o); l.d = rot_apply(in.inv_rows, r.d); }
    l.time = r.time;
    return l;
}

}  // namespace rtw
