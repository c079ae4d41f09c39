// ORACLE — test infrastructure only. Restates the three reference BVHs
// (Rust/src/objects/aabb/{aabb,qaabb,iaabb}.rs — one template here, they differ only in the leaf
// box and the sort key), Instance (objects/instance.rs:21-26,83-165,230-310) and Scene
// (viewport.rs:79-151).
#pragma once
#include <algorithm>
#include <memory>
#include <vector>
#include "objects.hpp"

namespace oracle {

struct Counters {
    uint64_t rays = 0;        // Scene::collision_normal calls
    uint64_t slab_tests = 0;  // box tests
    uint64_t sphere_tests = 0, quad_tests = 0, instance_tests = 0;
};

// Context threaded through collision calls: RNG + bounce index for the volume sub-streams,
// instrumentation counters, traversal mode.
struct TraceCtx {
    PathRng* rng = nullptr;
    uint32_t bounce = 0;
    Counters* ctr = nullptr;
    bool tie = false;         // set when a closest-hit decision was made between equal t (unpinned by the reference)
};

// Build-time randomness (`random::<Axis>()`, aabb.rs:42, qaabb.rs:134, iaabb.rs:87): a private
// Philox stream; the choice changes traversal cost and tie order only.
struct BuildRng {
    uint32_t n = 0, seed = 0;
    int axis() { Philox4 p = philox4x32_10(n++, 0x4158u, 0u, 0u, seed, 0xB17Du); return (int)(p.w[0] % 3u); }
};

template <class Prim> struct PrimTraits;

// aabb.rs:11-17 / qaabb.rs:11-18 / iaabb.rs:10-16
template <class Prim>
struct Bvh {
    Box3 box;                         // x, y, z intervals
    std::vector<const Prim*> prims;   // the reference clones the list at every level (aabb.rs:68)
    std::vector<Bvh> aabbs;           // 0 or 2 children

    static Bvh empty() { return Bvh{}; }   // all-zero intervals (aabb.rs:75-83)

    // aabb.rs:20-71, qaabb.rs:127-163, iaabb.rs:80-116
    static Bvh build(std::vector<const Prim*> ps, BuildRng& br) {
        if (ps.empty()) return empty();
        if (ps.size() == 1) {
            Bvh b;
            b.box = PrimTraits<Prim>::leaf_box(*ps[0]);
            b.prims = ps;
            return b;
        }
        int axis = br.axis();
        std::sort(ps.begin(), ps.end(), [axis](const Prim* a, const Prim* b) {
            return PrimTraits<Prim>::sort_key(*a, axis) < PrimTraits<Prim>::sort_key(*b, axis);
        });
        size_t len = ps.size() / 2;
        std::vector<const Prim*> p1(ps.begin(), ps.begin() + len), p2(ps.begin() + len, ps.end());
        Bvh a1 = build(std::move(p1), br), a2 = build(std::move(p2), br);
        Bvh b;
        b.box = Box3{a1.box.x + a2.box.x, a1.box.y + a2.box.y, a1.box.z + a2.box.z};
        b.prims = std::move(ps);
        b.aabbs.push_back(std::move(a1));
        b.aabbs.push_back(std::move(a2));
        return b;
    }

    // aabb.rs:105-155 / qaabb.rs:197-259 / iaabb.rs:150-212
    OptHit collision_normal(const Ray& r, float mint, float maxt, TraceCtx& cx) const {
        if (prims.empty() && aabbs.empty()) return std::nullopt;
        if (!slab_test(box, r, mint, maxt, cx.ctr ? &cx.ctr->slab_tests : nullptr)) return std::nullopt;
        OptHit min_hit;
        if (!aabbs.empty()) {
            for (const Bvh& c : aabbs) keep_closer(min_hit, c.collision_normal(r, mint, maxt, cx), &cx.tie);
        } else {
            for (const Prim* p : prims) keep_closer(min_hit, PrimTraits<Prim>::collide(*p, r, mint, maxt, cx), &cx.tie);
        }
        return min_hit;
    }
};

template <> struct PrimTraits<Sphere> {
    static Box3 leaf_box(const Sphere& s) { return s.leaf_box(); }
    static float sort_key(const Sphere& s, int axis) {   // aabb.rs:44-52: centre + radius
        return (axis == 0 ? s.origin.x : axis == 1 ? s.origin.y : s.origin.z) + s.radius;
    }
    static OptHit collide(const Sphere& s, const Ray& r, float mint, float maxt, TraceCtx& cx) {
        if (cx.ctr) ++cx.ctr->sphere_tests;
        return s.collision_normal(r, mint, maxt);
    }
};
template <> struct PrimTraits<Quad> {
    static Box3 leaf_box(const Quad& q) { return q.leaf_box(); }
    static float sort_key(const Quad& q, int axis) {     // qaabb.rs:134-146: box min
        Box3 b = q.leaf_box();
        return axis == 0 ? b.x.min : axis == 1 ? b.y.min : b.z.min;
    }
    static OptHit collide(const Quad& q, const Ray& r, float mint, float maxt, TraceCtx& cx) {
        if (cx.ctr) ++cx.ctr->quad_tests;
        return q.collision_normal(r, mint, maxt);
    }
};

// ---------------------------------------------------------------- instance.rs:28-39
struct Instance {
    std::vector<Quad> quads;
    std::vector<Sphere> spheres;
    Bvh<Quad> qaabb;
    Bvh<Sphere> saabb;
    Vec3 translation, rotation;
    bool const_density = false;   // dist_fn = &const_density (instance.rs:24-26) vs &surface (:21-23)
    float density = 0.f;
    uint32_t index = 0;           // position in Scene.instances: selects the volume RNG sub-stream

    void build(BuildRng& br) {
        std::vector<const Sphere*> sp; for (auto& s : spheres) sp.push_back(&s);
        std::vector<const Quad*> qp;   for (auto& q : quads) qp.push_back(&q);
        saabb = Bvh<Sphere>::build(sp, br);
        qaabb = Bvh<Quad>::build(qp, br);
    }

    OptHit inner(const Ray& r, float mint, float maxt, TraceCtx& cx) const {   // instance.rs:262-271
        OptHit min_hit;
        keep_closer(min_hit, saabb.collision_normal(r, mint, maxt, cx), &cx.tie);
        keep_closer(min_hit, qaabb.collision_normal(r, mint, maxt, cx), &cx.tie);
        return min_hit;
    }

    // instance.rs:250-310
    OptHit collision_normal(const Ray& r_world, float mint, float maxt, TraceCtx& cx) const {
        Ray r = Ray(r_world.origin - translation, r_world.direction, r_world.time).rotated(-rotation);
        OptHit min_hit = inner(r, mint, maxt, cx);
        if (!min_hit) return std::nullopt;
        Hit hit = *min_hit;
        // (self.dist_fn)(self.density): surface -> -1.0; const_density -> ln(xi) / -d
        float distance = -1.0f;
        if (const_density) {
            float xi = cx.rng->block_at((cx.bounce << 16) | 0u, 1u + index).f[0];
            distance = std::log(xi) / -density;
        }
        if (distance >= 0.0f) {
            r.origin = hit.point + r.direction * distance;
            OptHit again = inner(r, mint, maxt, cx);
            if (!again) return std::nullopt;
            hit.point = r.origin;
            const PathRng* rng = cx.rng;
            uint32_t base = cx.bounce << 16, stream = 1u + index;
            hit.normal = random_unit_vec_from([&](uint32_t turn) { return rng->block_at(base | (turn + 1u), stream); });
        }
        hit.point.rotate(rotation);
        hit.point += translation;
        hit.normal.rotate(rotation);
        return hit;
    }

    // iaabb.rs:23-76: leaf box = AABB of the rotated corners of (qaabb ∪ saabb) + translation.
    // An empty inner BVH contributes the interval [0,0] (aabb.rs:75-83), so the local origin is
    // always inside the box.
    Box3 leaf_box() const {
        Interval x = qaabb.box.x + saabb.box.x, y = qaabb.box.y + saabb.box.y, z = qaabb.box.z + saabb.box.z;
        Vec3 pts[8] = {{x.min, y.min, z.min}, {x.min, y.min, z.max}, {x.min, y.max, z.min}, {x.min, y.max, z.max},
                       {x.max, y.min, z.min}, {x.max, y.min, z.max}, {x.max, y.max, z.min}, {x.max, y.max, z.max}};
        for (auto& p : pts) p = p.rotated(rotation);
        float mxx = pts[0].x, mxy = pts[0].y, mxz = pts[0].z, mnx = pts[0].x, mny = pts[0].y, mnz = pts[0].z;
        for (int i = 1; i < 8; ++i) {
            mxx = maxf(mxx, pts[i].x); mxy = maxf(mxy, pts[i].y); mxz = maxf(mxz, pts[i].z);
            mnx = minf(mnx, pts[i].x); mny = minf(mny, pts[i].y); mnz = minf(mnz, pts[i].z);
        }
        return Box3{Interval{mnx, mxx} + translation.x, Interval{mny, mxy} + translation.y,
                    Interval{mnz, mxz} + translation.z};
    }
};

template <> struct PrimTraits<Instance> {
    static Box3 leaf_box(const Instance& i) { return i.leaf_box(); }
    static float sort_key(const Instance& i, int axis) {   // iaabb.rs:87-99: box min
        Box3 b = i.leaf_box();
        return axis == 0 ? b.x.min : axis == 1 ? b.y.min : b.z.min;
    }
    static OptHit collide(const Instance& i, const Ray& r, float mint, float maxt, TraceCtx& cx) {
        if (cx.ctr) ++cx.ctr->instance_tests;
        return i.collision_normal(r, mint, maxt, cx);
    }
};

// ---------------------------------------------------------------- viewport.rs:79-151
struct Scene {
    std::vector<Sphere> spheres;
    std::vector<Quad> quads;
    std::vector<Instance> instances;
    Bvh<Sphere> aabb;
    Bvh<Quad> qaabb;
    Bvh<Instance> iaabb;
    Vec3 background_color;
    // the `lights` of Rust2's light-biased integrators (Rust2/src/viewport/ray_color.rs:59,115): top-level primitives
    struct Light { bool quad; uint32_t index; };
    std::vector<Light> lights;

    void build(uint32_t build_seed) {   // Scene::new, viewport.rs:120-135
        BuildRng br; br.seed = build_seed;
        for (uint32_t k = 0; k < instances.size(); ++k) { instances[k].index = k; instances[k].build(br); }
        std::vector<const Sphere*> sp; for (auto& s : spheres) sp.push_back(&s);
        std::vector<const Quad*> qp;   for (auto& q : quads) qp.push_back(&q);
        std::vector<const Instance*> ip; for (auto& i : instances) ip.push_back(&i);
        aabb = Bvh<Sphere>::build(sp, br);
        qaabb = Bvh<Quad>::build(qp, br);
        iaabb = Bvh<Instance>::build(ip, br);
    }
    // viewport.rs:136-150: order S, Q, I; strict > keeps the earlier class on equal t
    OptHit collision_normal(const Ray& r, float mint, float maxt, TraceCtx& cx) const {
        if (cx.ctr) ++cx.ctr->rays;
        OptHit min_hit;
        keep_closer(min_hit, aabb.collision_normal(r, mint, maxt, cx), &cx.tie);
        keep_closer(min_hit, qaabb.collision_normal(r, mint, maxt, cx), &cx.tie);
        keep_closer(min_hit, iaabb.collision_normal(r, mint, maxt, cx), &cx.tie);
        return min_hit;
    }
};

}  // namespace oracle
