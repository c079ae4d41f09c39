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
#pragma once
#include "rtw_device.h"

namespace rtw {

struct Ray { V3 o, d; float time; };

// per-ray constants of one frame (world or instance-local)
struct RayK {
    V3 inv;       // 1/d per axis (objects.rs:136)
    float a;      // d.d (sphere.rs:103)
    bool slow;    // some |1/d| is not finite (d == 0 or denormal): use the literal slab rule
};
RTW_HD RayK ray_constants(const Ray& r) {
    RayK k;
    k.inv = v3(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z);
    k.a = dot(r.d, r.d);
    k.slow = !(isfinite(k.inv.x) && isfinite(k.inv.y) && isfinite(k.inv.z));
    return k;
}

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
RTW_HD_NOINL bool slab_slow(V3 lo, V3 hi, const Ray& r, float mint, float maxt, float& tnear) {
    float x0, x1, y0, y1, z0, z1;
    if (!axis_interval(lo.x, hi.x, r.d.x, r.o.x, x0, x1)) return false;
    if (!axis_interval(lo.y, hi.y, r.d.y, r.o.y, y0, y1)) return false;
    if (!axis_interval(lo.z, hi.z, r.d.z, r.o.z, z0, z1)) return false;
    float mn = ref_maxf(ref_maxf(x0, y0), ref_maxf(z0, mint));
    float mx = ref_minf(ref_minf(x1, y1), ref_minf(z1, maxt));
    tnear = mn;
    return !(mn > mx);
}
// The slab test of AABB/QuadAABB/IAABB::collision_normal (aabb.rs:109-126): miss iff
// max(tx0,ty0,tz0,mint) > min(tx1,ty1,tz1,maxt). With finite 1/d no NaN can arise, so
// fminf/fmaxf equal the reference's minf/maxf.
RTW_HD bool slab(V3 lo, V3 hi, const Ray& r, const RayK& k, float mint, float maxt, float& tnear) {
    if (k.slow) return slab_slow(lo, hi, r, mint, maxt, tnear);
    float x1 = (lo.x - r.o.x) * k.inv.x, x2 = (hi.x - r.o.x) * k.inv.x;
    float y1 = (lo.y - r.o.y) * k.inv.y, y2 = (hi.y - r.o.y) * k.inv.y;
    float z1 = (lo.z - r.o.z) * k.inv.z, z2 = (hi.z - r.o.z) * k.inv.z;
    float mn = fmaxf(fmaxf(fminf(x1, x2), fminf(y1, y2)), fmaxf(fminf(z1, z2), mint));
    float mx = fminf(fminf(fmaxf(x1, x2), fmaxf(y1, y2)), fminf(fmaxf(z1, z2), maxt));
    tnear = mn;
    return !(mn > mx);
}

// Sphere::collision_normal up to the accepted root (sphere.rs:100-123). `c` = centre at ray time.
RTW_HD bool sphere_t(V3 c, float radius, const Ray& r, float a, float mint, float maxt, float& t) {
    V3 oc = r.o - c;
    float b = dot(oc, r.d);
    float cc = dot(oc, oc) - radius * radius;
    float d = b * b - a * cc;
    if (d < 0.0f) return false;
    float sq = sqrtf(d);
    float x = (-b - sq) / a;
    if (x < mint) x = (-b + sq) / a;
    if (x < mint || x > maxt) return false;
    t = x;
    return true;
}

// Quad::collision_normal up to the inside test (quad.rs:38-60).
RTW_HD bool quad_t(const Quad* qp, const Ray& r, float mint, float maxt, float& t, float& alfa, float& beta) {
    float4 q0 = RTW_LDG(&qp->q0), q1 = RTW_LDG(&qp->q1), q2 = RTW_LDG(&qp->q2), q3 = RTW_LDG(&qp->q3);
    V3 n = v3(q0.x, q0.y, q0.z);
    float den = dot(n, r.d);
    if (fabsf(den) <= 1e-8f) return false;
    float tt = (q0.w - dot(n, r.o)) / den;
    if (tt < mint || tt > maxt) return false;
    V3 p = r.o + r.d * tt;
    V3 planar = p - v3(q1.x, q1.y, q1.z);
    V3 u = v3(q1.w, q2.x, q2.y), v = v3(q2.z, q2.w, q3.x), w = v3(q3.y, q3.z, q3.w);
    float al = dot(w, cross(planar, v));
    float be = dot(w, cross(u, planar));
    if (al < 0.0f || al > 1.0f || be < 0.0f || be > 1.0f) return false;
    t = tt; alfa = al; beta = be;
    return true;
}

// best candidate so far. `t` doubles as the shrunk upper bound of the search interval.
struct Best {
    float t;
    uint32_t ref;     // REF_NONE = nothing yet; at top level an instance hit is (KIND_INSTANCE, k)
    uint32_t inner;   // for instance hits: the primitive inside the instance
};
RTW_HD bool closer(float t, uint32_t ref, const Best& b) { return t < b.t || (t == b.t && ref < b.ref); }

// Algorithmic work of a traversal, counted only by the instrumented kernel build
// (-DRTW_COUNT_WORK, RTW_FLAG_COUNT_WORK): the per-ray figures bench.py's roofline uses.
struct Work { uint32_t nodes, spheres, quads, instances, hits; };
#if defined(RTW_COUNT_WORK)
#define RTW_COUNT(cx, field) (++(cx).work->field)
#else
#define RTW_COUNT(cx, field) ((void)0)
#endif

struct TraceCtx {
    const Scene* sc;
    const PathRng* rng;
    uint32_t bounce;
    float mint, maxt;   // the integrator's interval (un-shrunk)
    Work* work;         // null unless RTW_COUNT_WORK
};

#define RTW_STACK 32

// Leaf test of one sphere/quad against the current best (static spheres only; moving ones are
// handled by test_moving).
RTW_HD void test_prim(const TraceCtx& cx, uint32_t ref, const Ray& r, const RayK& k, Best& best) {
    const Scene& sc = *cx.sc;
    const float mint = cx.mint;
    uint32_t idx = ref_index(ref);
    float t;
    if (ref_kind(ref) == KIND_SPHERE) {
        RTW_COUNT(cx, spheres);
        float4 s = RTW_LDG(&sc.spheres[idx]);
        if (sphere_t(v3(s.x, s.y, s.z), s.w, r, k.a, mint, best.t, t) && closer(t, ref, best)) {
            best.t = t; best.ref = ref; best.inner = REF_NONE;
        }
    } else {
        RTW_COUNT(cx, quads);
        float al, be;
        if (quad_t(&sc.quads[idx], r, mint, best.t, t, al, be) && closer(t, ref, best)) {
            best.t = t; best.ref = ref; best.inner = REF_NONE;
        }
    }
}

// Moving spheres of a level: t=0 leaf box against the un-shrunk interval (aabb.rs:27-38,109-126),
// then the sphere at its position at ray time (sphere.rs:100).
RTW_HD void test_moving(const TraceCtx& cx, const Level& lv, const Ray& r, const RayK& k, Best& best) {
    const Scene& sc = *cx.sc;
    const float mint = cx.mint, maxt = cx.maxt;
    for (uint32_t m = 0; m < lv.mov_count; ++m) {
        RTW_COUNT(cx, spheres);
        uint32_t idx = RTW_LDG(&sc.moving[lv.mov_first + m]);
        float4 lo = RTW_LDG(&sc.sphere_box[2 * idx]), hi = RTW_LDG(&sc.sphere_box[2 * idx + 1]);
        float tn;
        if (!slab(v3(lo.x, lo.y, lo.z), v3(hi.x, hi.y, hi.z), r, k, mint, maxt, tn)) continue;
        float4 s = RTW_LDG(&sc.spheres[idx]);
        float4 vel = RTW_LDG(&sc.sphere_vel[idx]);
        V3 c = v3(s.x, s.y, s.z) + v3(vel.x, vel.y, vel.z) * r.time;
        float t;
        uint32_t ref = make_ref(KIND_SPHERE, idx);
        if (sphere_t(c, s.w, r, k.a, mint, best.t, t) && closer(t, ref, best)) {
            best.t = t; best.ref = ref; best.inner = REF_NONE;
        }
    }
}

// Walk one instance-local level (spheres and quads only). ANY = stop at the first hit (the volume
// re-traversal of instance.rs:284-294 only asks whether anything is hit).
template <bool ANY>
RTW_HD bool walk_inner(const TraceCtx& cx, const Level& lv, const Ray& r, const RayK& k, Best& best) {
    const Scene& sc = *cx.sc;
    const float mint = cx.mint;
    if (lv.mov_count) {
        test_moving(cx, lv, r, k, best);
        if (ANY && best.ref != REF_NONE) return true;
    }
    if (lv.root == REF_NONE) return best.ref != REF_NONE;
    float tn;
    if (!slab(v3(lv.lo[0], lv.lo[1], lv.lo[2]), v3(lv.hi[0], lv.hi[1], lv.hi[2]), r, k, mint, best.t, tn))
        return best.ref != REF_NONE;
    uint32_t stack[RTW_STACK];
    int sp = 0;
    uint32_t cur = lv.root;
    for (;;) {
        if (ref_kind(cur) == KIND_NODE) {
            RTW_COUNT(cx, nodes);
            const Node* np = &sc.nodes[cur];
            float4 n0 = RTW_LDG(&np->n0), n1 = RTW_LDG(&np->n1), n2 = RTW_LDG(&np->n2);
            uint4 n3 = RTW_LDG(&np->n3);
            float t0, t1;
            bool h0 = slab(v3(n0.x, n0.y, n0.z), v3(n0.w, n1.x, n1.y), r, k, mint, best.t, t0);
            bool h1 = slab(v3(n1.z, n1.w, n2.x), v3(n2.y, n2.z, n2.w), r, k, mint, best.t, t1);
            if (h0 && h1) {
                bool swap = t1 < t0;
                stack[sp++] = swap ? n3.x : n3.y;
                cur = swap ? n3.y : n3.x;
                continue;
            }
            if (h0) { cur = n3.x; continue; }
            if (h1) { cur = n3.y; continue; }
        } else {
            test_prim(cx, cur, r, k, best);
            if (ANY && best.ref != REF_NONE) return true;
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return best.ref != REF_NONE;
}

// Instance::collision_normal up to "is there a hit and at what t" (instance.rs:250-302).
RTW_HD Ray to_local(const Instance& in, const Ray& r) {
    Ray l;
    V3 o = r.o - v3(in.t[0], in.t[1], in.t[2]);
    if (in.flags & INST_ROT_IDENTITY) { l.o = o; l.d = r.d; }
    else { l.o = rot_apply(in.inv_rows, o); l.d = rot_apply(in.inv_rows, r.d); }
    l.time = r.time;
    return l;
}
// const_density (instance.rs:24-26): ln(xi) / -density, xi from the instance's volume stream.
RTW_HD float volume_distance(const Instance& in, const TraceCtx& cx) {
    float xi = cx.rng->at((cx.bounce << 16) | 0u, 1u + in.index).x;
    return logf(xi) / -in.density;
}

RTW_HD void test_instance(const TraceCtx& cx, uint32_t ref, const Ray& r, Best& best) {
    const Scene& sc = *cx.sc;
    RTW_COUNT(cx, instances);
    const Instance& in = sc.instances[ref_index(ref)];
    Ray l = to_local(in, r);
    RayK k = ray_constants(l);
    Best ib; ib.t = best.t; ib.ref = REF_NONE; ib.inner = REF_NONE;
    if (!walk_inner<false>(cx, in.level, l, k, ib)) return;
    if (in.flags & INST_VOLUME) {
        float distance = volume_distance(in, cx);
        if (distance >= 0.0f) {
            Ray l2 = l;
            l2.o = (l.o + l.d * ib.t) + l.d * distance;
            Best any; any.t = cx.maxt; any.ref = REF_NONE; any.inner = REF_NONE;
            if (!walk_inner<true>(cx, in.level, l2, k, any)) return;
        }
    }
    if (closer(ib.t, ref, best)) { best.t = ib.t; best.ref = ref; best.inner = ib.ref; }
}

// Scene::collision_normal (viewport.rs:136-150): closest of everything.
RTW_HD Best trace(const TraceCtx& cx, const Ray& r) {
    const Scene& sc = *cx.sc;
    const Level& lv = sc.top;
    Best best; best.t = cx.maxt; best.ref = REF_NONE; best.inner = REF_NONE;
    RayK k = ray_constants(r);
    if (lv.mov_count) test_moving(cx, lv, r, k, best);
    if (lv.root == REF_NONE) return best;
    float tn;
    if (!slab(v3(lv.lo[0], lv.lo[1], lv.lo[2]), v3(lv.hi[0], lv.hi[1], lv.hi[2]), r, k, cx.mint, best.t, tn))
        return best;
    uint32_t stack[RTW_STACK];
    int sp = 0;
    uint32_t cur = lv.root;
    for (;;) {
        uint32_t kind = ref_kind(cur);
        if (kind == KIND_NODE) {
            RTW_COUNT(cx, nodes);
            const Node* np = &sc.nodes[cur];
            float4 n0 = RTW_LDG(&np->n0), n1 = RTW_LDG(&np->n1), n2 = RTW_LDG(&np->n2);
            uint4 n3 = RTW_LDG(&np->n3);
            float t0, t1;
            bool h0 = slab(v3(n0.x, n0.y, n0.z), v3(n0.w, n1.x, n1.y), r, k, cx.mint, best.t, t0);
            bool h1 = slab(v3(n1.z, n1.w, n2.x), v3(n2.y, n2.z, n2.w), r, k, cx.mint, best.t, t1);
            if (h0 && h1) {
                bool swap = t1 < t0;
                stack[sp++] = swap ? n3.x : n3.y;
                cur = swap ? n3.y : n3.x;
                continue;
            }
            if (h0) { cur = n3.x; continue; }
            if (h1) { cur = n3.y; continue; }
        } else if (kind == KIND_INSTANCE) {
            test_instance(cx, cur, r, best);
        } else {
            test_prim(cx, cur, r, k, best);
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return best;
}

}  // namespace rtw
