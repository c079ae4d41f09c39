// scene_build.cpp — host side of rtw_scene_create: validate the rtw_scene_desc, derive the
// per-primitive quantities the reference derives in its constructors, compute every primitive's
// reference leaf box, build the two-level SAH BVH and lay everything out as the SoA arrays of
// rtw_device.h. Host-only C++ (no CUDA calls); compiled with -ffp-contract=off so the derived
// f32 values equal the reference's (rustc never fuses a*b+c).
//
//   Quad::new            Rust/src/objects/quad.rs:85-106     normal, d, w
//   AABB leaf box        objects/aabb/aabb.rs:27-38          origin +- radius at time 0
//   QuadAABB::from, pad  objects/aabb/qaabb.rs:19-61,107-126 corners, padded to 1e-5
//   IAABB::from          objects/aabb/iaabb.rs:23-76         rotated corners of the inner boxes + T
//   Vec3::rotated        vec3.rs:161-181                     the Euler map (not a rotation for x/z)
//   Viewport::new        viewport.rs:308-401                 camera basis
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>
#include <cuda_fp16.h>
#include "rtw_host.h"
#include "rtw_shade.h"

namespace rtw {

namespace {

inline float minf(float a, float b) { return a <= b ? a : b; }   // objects.rs:78-91
inline float maxf(float a, float b) { return a >= b ? a : b; }

struct Box { float lo[3], hi[3]; };
inline Box box_empty() {
    Box b;
    for (int a = 0; a < 3; ++a) { b.lo[a] = std::numeric_limits<float>::infinity(); b.hi[a] = -b.lo[a]; }
    return b;
}
inline void box_grow(Box& b, const Box& o) {
    for (int a = 0; a < 3; ++a) { b.lo[a] = minf(b.lo[a], o.lo[a]); b.hi[a] = maxf(b.hi[a], o.hi[a]); }
}
inline float box_area(const Box& b) {
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    if (!(dx >= 0.f) || !(dy >= 0.f) || !(dz >= 0.f)) return 0.f;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}

// two floats -> two IEEE halves in one word (first in the low half), rounded towards -inf / +inf
inline uint32_t pack_half2_rd(float a, float b) {
    return (uint32_t)__half_as_ushort(__float2half_rd(a)) | ((uint32_t)__half_as_ushort(__float2half_rd(b)) << 16);
}
inline uint32_t pack_half2_ru(float a, float b) {
    return (uint32_t)__half_as_ushort(__float2half_ru(a)) | ((uint32_t)__half_as_ushort(__float2half_ru(b)) << 16);
}

struct BuildPrim { uint32_t ref; Box box; float c[3]; };

// vec3.rs:161-181 as coefficient rows (see rot_apply in rtw_device.h)
void euler_rows(const float rot[3], float r[12]) {
    float sa = std::sin(rot[0]), ca = std::cos(rot[0]);
    float sb = std::sin(rot[1]), cb = std::cos(rot[1]);
    float sc = std::sin(rot[2]), cc = std::cos(rot[2]);
    r[0] = cb; r[1] = cc; r[2] = sa * sb * cc - sa * cc; r[3] = ca * sb * cc + sa * sc;
    r[4] = cb; r[5] = sc; r[6] = sa * sb * sc + ca * cc; r[7] = ca * sb * sc - sa * cc;
    r[8] = -sb; r[9] = sa; r[10] = cb; r[11] = ca;
}

bool finite3(const float* p) { return std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]); }

// Two-step build per level: a binary top-down SAH tree (full sweep on each axis, one primitive per
// leaf), then a collapse into 4-wide nodes: a node adopts its grandchildren, largest box first,
// until it has four children. Leaves are never merged: every primitive stays reachable through its
// own reference box, which is what the traversal slab-tests before the primitive (N16).
struct Builder {
    struct BNode { Box b[2]; uint32_t c[2]; };        // c: REF of a leaf, or BIN | index into bin
    static const uint32_t BIN = 0x80000000u;          // marks a binary inner node in BNode::c (temporary)
    std::vector<Node>& nodes;
    std::vector<BNode> bin;
    uint32_t max_pushes = 0;                           // deepest stack use of the level built last
    explicit Builder(std::vector<Node>& n) : nodes(n) {}

    // Top-down SAH over the primitives at positions [begin, end) of the three index lists `ord[axis]`, each of which
    // holds the same primitives sorted by centroid along its axis (sorted once per level, partitioned stably on the
    // way down: O(n log n) for the whole level instead of three sorts per node). Full sweep on every axis; falls
    // back to a median split when the remaining depth budget (RTW_STACK binary levels) gets tight. Returns a leaf
    // REF, or the index of the binary node.
    std::vector<uint32_t> ord[3], part_tmp;
    std::vector<uint8_t> side;
    std::vector<float> right_area;
    uint32_t build(const std::vector<BuildPrim>& p, size_t begin, size_t end, uint32_t depth, Box& out_box, bool& is_bin) {
        const size_t n = end - begin;
        if (n == 1) { const BuildPrim& q = p[ord[0][begin]]; out_box = q.box; is_bin = false; return q.ref; }
        Box cb = box_empty();
        for (size_t i = begin; i < end; ++i) {
            const BuildPrim& q = p[ord[0][i]];
            for (int a = 0; a < 3; ++a) { cb.lo[a] = minf(cb.lo[a], q.c[a]); cb.hi[a] = maxf(cb.hi[a], q.c[a]); }
        }
        uint32_t levels_needed = 1; while (((size_t)1 << levels_needed) < n) ++levels_needed;
        const bool force_median = depth + levels_needed + 2 >= (uint32_t)RTW_STACK;
        int best_axis = -1; size_t best_k = 0; float best_cost = std::numeric_limits<float>::infinity();
        if (!force_median) {
            for (int axis = 0; axis < 3; ++axis) {
                if (!(cb.hi[axis] > cb.lo[axis])) continue;
                const uint32_t* o = ord[axis].data() + begin;
                Box rb = box_empty();
                for (size_t i = n; i-- > 1;) { box_grow(rb, p[o[i]].box); right_area[i] = box_area(rb); }
                Box lb = box_empty();
                for (size_t k = 1; k < n; ++k) {
                    box_grow(lb, p[o[k - 1]].box);
                    const float cost = box_area(lb) * (float)k + right_area[k] * (float)(n - k);
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_k = k; }
                }
            }
        }
        if (best_axis < 0) {   // median split along the widest centroid axis (or index order if all equal)
            int axis = 0; float w = -1.f;
            for (int a = 0; a < 3; ++a) { float e = cb.hi[a] - cb.lo[a]; if (e > w) { w = e; axis = a; } }
            best_axis = axis; best_k = n / 2;
        }
        // the first best_k primitives of the chosen axis' order go left; the other two lists are partitioned stably
        for (size_t i = 0; i < n; ++i) side[ord[best_axis][begin + i]] = i < best_k ? 0 : 1;
        for (int axis = 0; axis < 3; ++axis) {
            if (axis == best_axis) continue;
            uint32_t* o = ord[axis].data() + begin;
            size_t l = 0, r = 0;
            for (size_t i = 0; i < n; ++i) { if (side[o[i]] == 0) o[l++] = o[i]; else part_tmp[r++] = o[i]; }
            for (size_t i = 0; i < r; ++i) o[l + i] = part_tmp[i];
        }
        uint32_t me = (uint32_t)bin.size();
        bin.emplace_back();
        binflag.resize(bin.size() * 2, 0);
        Box b0, b1; bool bin0, bin1;
        uint32_t c0 = build(p, begin, begin + best_k, depth + 1, b0, bin0);
        uint32_t c1 = build(p, begin + best_k, end, depth + 1, b1, bin1);
        bin[me].b[0] = b0; bin[me].b[1] = b1;
        bin[me].c[0] = c0; bin[me].c[1] = c1;
        binflag.resize(bin.size() * 2, 0);
        binflag[2 * me] = bin0; binflag[2 * me + 1] = bin1;
        out_box = b0; box_grow(out_box, b1);
        is_bin = true;
        return me;
    }
    std::vector<uint8_t> binflag;                      // binflag[2*i+k]: child k of bin[i] is a binary node

    // Emit the 4-wide node for binary node `b`; returns its reference and the stack words a walk
    // below it can need (siblings left on the stack at every level).
    uint32_t emit4(uint32_t b, uint32_t& pushes) {
        struct Child { Box box; uint32_t c; bool is_bin; };
        Child ch[4]; int n = 2;
        for (int k = 0; k < 2; ++k) { ch[k].box = bin[b].b[k]; ch[k].c = bin[b].c[k]; ch[k].is_bin = binflag[2 * b + k] != 0; }
        while (n < 4) {
            int pick = -1; float area = -1.f;
            for (int k = 0; k < n; ++k) if (ch[k].is_bin) { float a = box_area(ch[k].box); if (a > area) { area = a; pick = k; } }
            if (pick < 0) break;
            uint32_t g = ch[pick].c;
            Child a, c;
            a.box = bin[g].b[0]; a.c = bin[g].c[0]; a.is_bin = binflag[2 * g] != 0;
            c.box = bin[g].b[1]; c.c = bin[g].c[1]; c.is_bin = binflag[2 * g + 1] != 0;
            ch[pick] = a; ch[n++] = c;
        }
        uint32_t me = (uint32_t)nodes.size();
        nodes.emplace_back();
        const float inf = std::numeric_limits<float>::infinity();
        float lo[3][4], hi[3][4]; uint32_t ref[4];
        uint32_t below = 0;
        for (int k = 0; k < 4; ++k) {
            if (k < n) {
                for (int a = 0; a < 3; ++a) { lo[a][k] = ch[k].box.lo[a]; hi[a][k] = ch[k].box.hi[a]; }
                if (ch[k].is_bin) { uint32_t pb = 0; ref[k] = emit4(ch[k].c, pb); below = std::max(below, pb); }
                else ref[k] = ch[k].c;
            } else {
                for (int a = 0; a < 3; ++a) { lo[a][k] = inf; hi[a][k] = inf; }
                ref[k] = REF_NONE;
            }
        }
        Node& nd = nodes[me];
        nd.lox = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]); nd.loy = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
        nd.loz = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]); nd.hix = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
        nd.hiy = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]); nd.hiz = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
        nd.ref = make_uint4(ref[0], ref[1], ref[2], ref[3]);
        nd.pad = make_uint4(0u, 0u, 0u, 0u);
        pushes = (uint32_t)(n - 1) + below;
        return make_ref(KIND_NODE, me);
    }

    // Experimental collapse (RTW_BVH_COLLAPSE=dp; default: the greedy one above): choose, for every binary node,
    // which descendants become the up-to-four children of its wide node so that the summed surface area of all
    // wide nodes — the expected number of node visits of a random ray — is minimal. Leaves cost the same under
    // every cut (each keeps its own box), so only node areas enter. cut[b][k] = least cost of handing the subtree
    // of binary node b to its parent in at most k child slots; choice 0 = as one wide node, i = i slots to the left.
    std::vector<float> dp_cut;            // [4 * b + (k - 1)]
    std::vector<uint8_t> dp_choice;       // [4 * b + (k - 1)]; [4 * b + 0] of a wide node's root = its own split
    std::vector<uint8_t> dp_root;
    float dp_cutk(uint32_t c, bool is_bin, int k) const { return is_bin ? dp_cut[4 * c + (k - 1)] : 0.0f; }
    void dp_solve() {
        const size_t nb = bin.size();
        dp_cut.assign(4 * nb, 0.f); dp_choice.assign(4 * nb, 0); dp_root.assign(nb, 1);
        for (size_t b = nb; b-- > 0;) {       // children were created after their parent: solve them first
            const uint32_t L = bin[b].c[0], R = bin[b].c[1];
            const bool lb = binflag[2 * b] != 0, rb = binflag[2 * b + 1] != 0;
            Box bb = bin[b].b[0]; box_grow(bb, bin[b].b[1]);
            float best = std::numeric_limits<float>::infinity(); int bi = 1;
            for (int i = 1; i <= 3; ++i) { float c = dp_cutk(L, lb, i) + dp_cutk(R, rb, 4 - i); if (c < best) { best = c; bi = i; } }
            const float wide = box_area(bb) + best;
            dp_root[b] = (uint8_t)bi;
            dp_cut[4 * b] = wide; dp_choice[4 * b] = 0;
            for (int k = 2; k <= 4; ++k) {
                float bestk = wide; int ck = 0;
                for (int i = 1; i < k; ++i) { float c = dp_cutk(L, lb, i) + dp_cutk(R, rb, k - i); if (c < bestk) { bestk = c; ck = i; } }
                dp_cut[4 * b + (k - 1)] = bestk; dp_choice[4 * b + (k - 1)] = (uint8_t)ck;
            }
        }
    }
    // RTW_BVH_STATS=1 (tuning aid): summed surface area of the wide nodes, relative to the root's - the expected number of
    // node visits of a random ray crossing the level - for the best collapse into nodes of at most `width` children.
    double dp_visits(int width) const {
        const size_t nb = bin.size();
        if (nb == 0) return 0.0;
        std::vector<double> cut((size_t)width * nb, 0.0);          // cut[width * b + (k - 1)]: subtree of b handed up in <= k slots
        auto cutk = [&](uint32_t c, bool is_bin, int k) { return is_bin ? cut[(size_t)width * c + (k - 1)] : 0.0; };
        for (size_t b = nb; b-- > 0;) {
            const uint32_t L = bin[b].c[0], R = bin[b].c[1];
            const bool lb = binflag[2 * b] != 0, rb = binflag[2 * b + 1] != 0;
            Box bb = bin[b].b[0]; box_grow(bb, bin[b].b[1]);
            double best = std::numeric_limits<double>::infinity();
            for (int i = 1; i < width; ++i) best = std::min(best, cutk(L, lb, i) + cutk(R, rb, width - i));
            const double wide = (double)box_area(bb) + best;
            cut[(size_t)width * b] = wide;
            for (int k = 2; k <= width; ++k) {
                double bestk = wide;
                for (int i = 1; i < k; ++i) bestk = std::min(bestk, cutk(L, lb, i) + cutk(R, rb, k - i));
                cut[(size_t)width * b + (k - 1)] = bestk;
            }
        }
        Box rb = bin[0].b[0]; box_grow(rb, bin[0].b[1]);
        return cut[0] / std::max(1e-30, (double)box_area(rb));
    }
    struct DpChild { Box box; uint32_t c; bool is_bin; };
    void dp_expand(const Box& box, uint32_t c, bool is_bin, int k, DpChild* out, int& n) const {
        const int ch = (is_bin && k > 1) ? dp_choice[4 * c + (k - 1)] : 0;
        if (!is_bin || ch == 0) { out[n].box = box; out[n].c = c; out[n].is_bin = is_bin; ++n; return; }
        dp_expand(bin[c].b[0], bin[c].c[0], binflag[2 * c] != 0, ch, out, n);
        dp_expand(bin[c].b[1], bin[c].c[1], binflag[2 * c + 1] != 0, k - ch, out, n);
    }
    uint32_t emit4_dp(uint32_t b, uint32_t& pushes) {
        DpChild ch[4]; int n = 0;
        const int i = dp_root[b];
        dp_expand(bin[b].b[0], bin[b].c[0], binflag[2 * b] != 0, i, ch, n);
        dp_expand(bin[b].b[1], bin[b].c[1], binflag[2 * b + 1] != 0, 4 - i, ch, n);
        uint32_t me = (uint32_t)nodes.size();
        nodes.emplace_back();
        const float inf = std::numeric_limits<float>::infinity();
        float lo[3][4], hi[3][4]; uint32_t ref[4];
        uint32_t below = 0;
        for (int k = 0; k < 4; ++k) {
            if (k < n) {
                for (int a = 0; a < 3; ++a) { lo[a][k] = ch[k].box.lo[a]; hi[a][k] = ch[k].box.hi[a]; }
                if (ch[k].is_bin) { uint32_t pb = 0; ref[k] = emit4_dp(ch[k].c, pb); below = std::max(below, pb); }
                else ref[k] = ch[k].c;
            } else {
                for (int a = 0; a < 3; ++a) { lo[a][k] = inf; hi[a][k] = inf; }
                ref[k] = REF_NONE;
            }
        }
        Node& nd = nodes[me];
        nd.lox = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]); nd.loy = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
        nd.loz = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]); nd.hix = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
        nd.hiy = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]); nd.hiz = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
        nd.ref = make_uint4(ref[0], ref[1], ref[2], ref[3]);
        nd.pad = make_uint4(0u, 0u, 0u, 0u);
        pushes = (uint32_t)(n - 1) + below;
        return make_ref(KIND_NODE, me);
    }

    // A level made of an already built level plus a few more leaves: nodes on top of its root, each holding the
    // level below and up to three of the extra leaves. Used for the top level WITH the enclosing volumes, which
    // differs from the one without them by those one or two leaves: one SAH build serves both.
    Level wrap(const Level& inner, uint32_t inner_pushes, const std::vector<BuildPrim>& extra) {
        Level lv = inner;
        uint32_t pushes = inner_pushes;
        const float inf = std::numeric_limits<float>::infinity();
        for (size_t e = 0; e < extra.size();) {
            Box cb[4]; uint32_t ref[4]; int n = 0;
            if (lv.root != REF_NONE) {
                for (int a = 0; a < 3; ++a) { cb[0].lo[a] = lv.lo[a]; cb[0].hi[a] = lv.hi[a]; }
                ref[n++] = lv.root;
            }
            while (n < 4 && e < extra.size()) { cb[n] = extra[e].box; ref[n++] = extra[e].ref; ++e; }
            if (n == 1) { lv.root = ref[0]; for (int a = 0; a < 3; ++a) { lv.lo[a] = cb[0].lo[a]; lv.hi[a] = cb[0].hi[a]; } continue; }
            const uint32_t me = (uint32_t)nodes.size();
            nodes.emplace_back();
            float lo[3][4], hi[3][4]; uint32_t rf[4];
            Box ub = box_empty();
            for (int k = 0; k < 4; ++k) {
                if (k < n) { for (int a = 0; a < 3; ++a) { lo[a][k] = cb[k].lo[a]; hi[a][k] = cb[k].hi[a]; } rf[k] = ref[k]; box_grow(ub, cb[k]); }
                else { for (int a = 0; a < 3; ++a) { lo[a][k] = inf; hi[a][k] = inf; } rf[k] = REF_NONE; }
            }
            Node& nd = nodes[me];
            nd.lox = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]); nd.loy = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
            nd.loz = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]); nd.hix = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
            nd.hiy = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]); nd.hiz = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
            nd.ref = make_uint4(rf[0], rf[1], rf[2], rf[3]);
            nd.pad = make_uint4(0u, 0u, 0u, 0u);
            lv.root = make_ref(KIND_NODE, me);
            for (int a = 0; a < 3; ++a) { lv.lo[a] = ub.lo[a]; lv.hi[a] = ub.hi[a]; }
            pushes += (uint32_t)(n - 1);
        }
        max_pushes = pushes;
        return lv;
    }

    Level level(std::vector<BuildPrim>& prims, uint32_t mov_first, uint32_t mov_count) {
        Level lv; std::memset(&lv, 0, sizeof lv);
        lv.mov_first = mov_first; lv.mov_count = mov_count;
        max_pushes = 0;
        if (prims.empty()) { lv.root = REF_NONE; return lv; }
        bin.clear(); binflag.clear();
        const size_t n = prims.size();
        for (int axis = 0; axis < 3; ++axis) {
            ord[axis].resize(n);
            for (size_t i = 0; i < n; ++i) ord[axis][i] = (uint32_t)i;
            std::stable_sort(ord[axis].begin(), ord[axis].end(),
                             [&prims, axis](uint32_t a, uint32_t b) { return prims[a].c[axis] < prims[b].c[axis]; });
        }
        side.assign(n, 0); part_tmp.resize(n); right_area.resize(n);
        Box rb; bool is_bin;
        uint32_t r = build(prims, 0, n, 0, rb, is_bin);
        // the DP-chosen collapse is the default (C5 +0.5 %, C4 +5.5 %, C2 +2.7 % on the B200,
        // profiles/r2a_ab_bake_dp_fastvote.txt); RTW_BVH_COLLAPSE=greedy brings the largest-child-first one back
        static const bool use_dp = [] { const char* e = std::getenv("RTW_BVH_COLLAPSE"); return !(e && std::string(e) == "greedy"); }();
        static const bool stats = [] { const char* e = std::getenv("RTW_BVH_STATS"); return e && e[0] == '1'; }();
        if (is_bin && stats)
            std::fprintf(stderr, "bvh level: %zu leaves; expected node visits of a random ray (sum of node areas / root area): "
                                 "binary %.3f, 4-wide %.3f, 8-wide %.3f, 16-wide %.3f\n", n, dp_visits(2), dp_visits(4), dp_visits(8), dp_visits(16));
        if (is_bin && use_dp) { dp_solve(); lv.root = emit4_dp(r, max_pushes); }
        else lv.root = is_bin ? emit4(r, max_pushes) : r;
        for (int a = 0; a < 3; ++a) { lv.lo[a] = rb.lo[a]; lv.hi[a] = rb.hi[a]; }
        return lv;
    }
};

// Quad::new (quad.rs:93-105): the 64-byte record from origin, u, v
Quad make_quad(V3 Q, V3 u, V3 v) {
    V3 n = cross(u, v);                      // quad.rs:93
    V3 normal = unit(n);                     // :99
    float dd = dot(normal, Q);               // :100
    V3 w = n / dot(n, n);                    // :101
    Quad o;
    o.q0 = make_float4(normal.x, normal.y, normal.z, dd);
    o.q1 = make_float4(Q.x, Q.y, Q.z, u.x);
    o.q2 = make_float4(u.y, u.z, v.x, v.y);
    o.q3 = make_float4(v.z, w.x, w.y, w.z);
    return o;
}
// qaabb.rs:19-61 + pad() :107-126 for a parallelogram; Rust2 triangle.rs:50-89 for a triangle
Box quad_box(V3 Q, V3 u, V3 v, bool tri) {
    V3 op = Q + u + v, vc = Q + v, uc = Q + u;
    const float P4[4][3] = {{op.x, op.y, op.z}, {vc.x, vc.y, vc.z}, {uc.x, uc.y, uc.z}, {Q.x, Q.y, Q.z}};
    Box b;
    for (int a = 0; a < 3; ++a) {
        float mn, mx;
        if (tri) {
            mn = minf(P4[1][a], minf(P4[2][a], P4[3][a]));
            mx = maxf(P4[1][a], maxf(P4[2][a], P4[3][a]));
            const float MIN_AABB_WIDTH = 0.005f;
            if (mx - mn < MIN_AABB_WIDTH) {
                float c = 0.5f * (mx + mn);
                mx = c + MIN_AABB_WIDTH * 0.5f; mn = c - MIN_AABB_WIDTH * 0.5f;
            }
        } else {
            mn = minf(minf(P4[0][a], P4[1][a]), minf(P4[2][a], P4[3][a]));
            mx = maxf(maxf(P4[0][a], P4[1][a]), maxf(P4[2][a], P4[3][a]));
            const float MIN_SIZE_DIR = 0.00001f;
            if (mx - mn < MIN_SIZE_DIR) {
                float center = (mx - mn) / 2.0f + mn;
                mn = center - 0.5f * MIN_SIZE_DIR; mx = center + 0.5f * MIN_SIZE_DIR;
            }
        }
        b.lo[a] = mn; b.hi[a] = mx;
    }
    return b;
}

// The instance's world -> local map (rot_apply with inv_rows) as a matrix, and whether it is orthogonal: then the
// frame change is a rigid motion and the instance's primitives can be copied into world space (rtw_host.h).
bool rigid_frame(const float* r, double M[3][3]) {
    M[0][0] = (double)(r[0] * r[1]); M[0][1] = r[2]; M[0][2] = r[3];
    M[1][0] = (double)(r[4] * r[5]); M[1][1] = r[6]; M[1][2] = r[7];
    M[2][0] = r[8]; M[2][1] = (double)(r[9] * r[10]); M[2][2] = (double)(r[11] * r[10]);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double d = 0.0;
            for (int k = 0; k < 3; ++k) d += M[i][k] * M[j][k];
            if (!(std::fabs(d - (i == j ? 1.0 : 0.0)) < 2e-6)) return false;
        }
    return true;
}

BuildPrim make_prim(uint32_t ref, const Box& b) {
    BuildPrim p; p.ref = ref; p.box = b;
    for (int a = 0; a < 3; ++a) p.c[a] = 0.5f * b.lo[a] + 0.5f * b.hi[a];
    return p;
}

}  // namespace

Scene HostScene::host_view() {
    Scene s; std::memset(&s, 0, sizeof s);
    s.nodes = nodes.data(); s.nodes16 = nodes16.data();
    s.spheres = spheres.data(); s.sphere_vel = sphere_vel.data(); s.sphere_col = sphere_col.data();
    s.sphere_ids = sphere_ids.data(); s.sphere_box = sphere_box.data();
    s.quads = quads.data(); s.quad_ids = quad_ids.data();
    s.instances = instances.data(); s.one_sphere = one_sphere.data(); s.moving = moving.data();
    s.materials = materials.data();
    for (size_t i = 0; i < textures.size(); ++i) { textures[i].texels = texels[i].data(); textures[i].obj = 0; }
    s.textures = textures.data(); s.noises = noises.data();
    s.top = top;
    std::memcpy(s.background, background, sizeof background);
    s.n_noises = (uint32_t)noises.size();
    s.n_spheres_orig = n_spheres_orig; s.n_quads_orig = n_quads_orig;
    s.baked_sphere_src = baked_sphere_src.data(); s.baked_quad_src = baked_quad_src.data();
    s.lights = lights.data(); s.n_lights = n_lights;
    return s;
}

int build_host_scene(const rtw_scene_desc* d, HostScene& hs, std::string& err) {
    if (!d) { err = "null scene description"; return RTW_E_INVALID; }
    // RTW_BUILD_TIMING=1: phase times of this function on stderr (tuning aid)
    static const bool timing = [] { const char* e = std::getenv("RTW_BUILD_TIMING"); return e && e[0] == '1'; }();
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "build_host_scene: %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    if (d->abi_version != RTW_ABI_VERSION) { err = "abi_version mismatch"; return RTW_E_INVALID; }
    if (d->n_top_spheres > d->n_spheres || d->n_top_quads > d->n_quads) { err = "n_top_* exceeds n_*"; return RTW_E_INVALID; }
    if ((d->n_spheres && !d->spheres) || (d->n_quads && !d->quads) || (d->n_instances && !d->instances) ||
        (d->n_materials && !d->materials) || (d->n_textures && !d->textures) || (d->n_noises && !d->noises)) {
        err = "null array with non-zero count"; return RTW_E_INVALID;
    }
    if (d->n_spheres >= REF_INDEX_LIMIT || d->n_quads >= REF_INDEX_LIMIT || d->n_instances >= REF_INDEX_LIMIT - 2u) {
        err = "too many primitives"; return RTW_E_UNSUPPORTED;
    }
    if ((d->n_spheres || d->n_quads) && (d->n_materials == 0 || d->n_textures == 0)) {
        err = "primitives need at least one material and one texture"; return RTW_E_INVALID;
    }

    // ---- materials, noises, textures
    hs.materials.resize(d->n_materials);
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const rtw_material& m = d->materials[i];
        Material& o = hs.materials[i]; std::memset(&o, 0, sizeof o);
        o.metallicness = m.metallicness; o.opacity = m.opacity; o.ir = m.ir;
        o.emitted[0] = m.emitted[0]; o.emitted[1] = m.emitted[1]; o.emitted[2] = m.emitted[2];
    }
    hs.noises.resize(d->n_noises);
    for (uint32_t i = 0; i < d->n_noises; ++i) {
        for (int k = 0; k < 256; ++k) {
            hs.noises[i].ranvec[k] = make_float4(d->noises[i].ranvec[k][0], d->noises[i].ranvec[k][1], d->noises[i].ranvec[k][2], 0.f);
            hs.noises[i].perm[0][k] = (uint8_t)(d->noises[i].perm_x[k] & 255u);
            hs.noises[i].perm[1][k] = (uint8_t)(d->noises[i].perm_y[k] & 255u);
            hs.noises[i].perm[2][k] = (uint8_t)(d->noises[i].perm_z[k] & 255u);
        }
    }
    hs.textures.resize(d->n_textures);
    hs.texels.resize(d->n_textures);
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const rtw_texture& t = d->textures[i];
        if (t.width == 0 || t.height == 0 || !t.rgb) { err = "texture with zero size or null texels"; return RTW_E_INVALID; }
        if ((uint64_t)t.width * t.height >= (1ull << 31)) { err = "texture too large"; return RTW_E_UNSUPPORTED; }
        if (t.noise_id >= 0 && (uint32_t)t.noise_id >= d->n_noises) { err = "texture noise_id out of range"; return RTW_E_INVALID; }
        Texture& o = hs.textures[i]; std::memset(&o, 0, sizeof o);
        o.width = t.width; o.height = t.height; o.noise_id = t.noise_id < 0 ? -1 : t.noise_id; o.noise_scale = t.noise_scale;
        o.n_texels = t.width * t.height;
        o.solid[0] = t.rgb[0]; o.solid[1] = t.rgb[1]; o.solid[2] = t.rgb[2];
    }
    // RGB f32 -> the float4 texels the CUDA arrays hold: 1.3 ms for C5's earth map, done beside the BVH builds
    struct Joiner { std::thread t; ~Joiner() { if (t.joinable()) t.join(); } } texel_job;
    auto convert_texels = [&hs, d]() {
        for (uint32_t i = 0; i < d->n_textures; ++i) {
            const rtw_texture& t = d->textures[i];
            std::vector<float4>& tv = hs.texels[i];
            const uint32_t n = hs.textures[i].n_texels;
            tv.resize(n);
            for (uint32_t k = 0; k < n; ++k) tv[k] = make_float4(t.rgb[3 * k], t.rgb[3 * k + 1], t.rgb[3 * k + 2], 0.f);
        }
    };
    {
        uint64_t texels = 0;
        for (uint32_t i = 0; i < d->n_textures; ++i) texels += hs.textures[i].n_texels;
        if (texels >= 65536u) texel_job.t = std::thread(convert_texels); else convert_texels();
    }

    lap("materials, noises, textures");
    // ---- spheres
    const uint32_t ns = d->n_spheres, nq = d->n_quads;
    hs.spheres.resize(ns); hs.sphere_vel.resize(ns); hs.sphere_col.resize(ns); hs.sphere_ids.resize(ns);
    hs.sphere_box.resize(2 * (size_t)ns);
    std::vector<Box> sbox(ns), qbox(nq);
    std::vector<uint32_t> qtri(nq);   // REF_TRIANGLE for triangle records, 0 otherwise
    std::vector<uint8_t> smoving(ns);
    for (uint32_t i = 0; i < ns; ++i) {
        const rtw_sphere& s = d->spheres[i];
        if (s.mat_id >= d->n_materials || s.tex_id >= d->n_textures) { err = "sphere material/texture id out of range"; return RTW_E_INVALID; }
        if (!finite3(s.origin) || !std::isfinite(s.radius) || !finite3(s.velocity)) { err = "non-finite sphere"; return RTW_E_INVALID; }
        bool mv = s.velocity[0] != 0.f || s.velocity[1] != 0.f || s.velocity[2] != 0.f;
        smoving[i] = mv;
        hs.spheres[i] = make_float4(s.origin[0], s.origin[1], s.origin[2], s.radius);
        hs.sphere_vel[i] = make_float4(s.velocity[0], s.velocity[1], s.velocity[2], 0.f);
        hs.sphere_col[i] = make_float4(s.col_mod[0], s.col_mod[1], s.col_mod[2], u2f(mv ? 1u : 0u));
        hs.sphere_ids[i] = make_uint2(s.mat_id, s.tex_id);
        for (int a = 0; a < 3; ++a) {   // aabb.rs:27-38: Interval::new(o - r, o + r)
            float x1 = s.origin[a] - s.radius, x2 = s.origin[a] + s.radius;
            sbox[i].lo[a] = minf(x1, x2); sbox[i].hi[a] = maxf(x1, x2);
        }
        hs.sphere_box[2 * i] = make_float4(sbox[i].lo[0], sbox[i].lo[1], sbox[i].lo[2], 0.f);
        hs.sphere_box[2 * i + 1] = make_float4(sbox[i].hi[0], sbox[i].hi[1], sbox[i].hi[2], 0.f);
    }
    // ---- quads
    hs.quads.resize(nq); hs.quad_ids.resize(nq);
    for (uint32_t i = 0; i < nq; ++i) {
        const rtw_quad& q = d->quads[i];
        if (q.mat_id >= d->n_materials || q.tex_id >= d->n_textures) { err = "quad material/texture id out of range"; return RTW_E_INVALID; }
        if (!finite3(q.origin) || !finite3(q.u) || !finite3(q.v)) { err = "non-finite quad"; return RTW_E_INVALID; }
        V3 Q = v3(q.origin[0], q.origin[1], q.origin[2]), u = v3(q.u[0], q.u[1], q.u[2]), v = v3(q.v[0], q.v[1], q.v[2]);
        hs.quads[i] = make_quad(Q, u, v);
        hs.quad_ids[i] = make_uint2(q.mat_id, q.tex_id);
        if (q.shape > RTW_SHAPE_TRIANGLE) { err = "quad shape is neither parallelogram nor triangle"; return RTW_E_INVALID; }
        const bool tri = q.shape == RTW_SHAPE_TRIANGLE;
        qtri[i] = tri ? REF_TRIANGLE : 0u;
        qbox[i] = quad_box(Q, u, v, tri);
    }
    hs.n_spheres_orig = ns; hs.n_quads_orig = nq;
    // ---- light list (Rust2/src/viewport/ray_color.rs:68-75,131-137: the ray goes to the middle of l.get_aabb())
    if (d->n_lights && !d->lights) { err = "null light list with non-zero count"; return RTW_E_INVALID; }
    for (uint32_t i = 0; i < d->n_lights; ++i) {
        const rtw_light& l = d->lights[i];
        const bool quad = l.kind == RTW_LIGHT_QUAD;
        if (l.kind > RTW_LIGHT_QUAD || l.index >= (quad ? d->n_top_quads : d->n_top_spheres)) {
            err = "light is not a top-level sphere / quad of the scene"; return RTW_E_INVALID;
        }
        const Box& b = quad ? qbox[l.index] : sbox[l.index];
        const uint32_t ref = quad ? make_ref(KIND_QUAD, l.index | qtri[l.index]) : make_ref(KIND_SPHERE, l.index);
        const Material& m = hs.materials[quad ? hs.quad_ids[l.index].x : hs.sphere_ids[l.index].x];
        hs.lights.push_back(make_float4((b.lo[0] + b.hi[0]) * 0.5f, (b.lo[1] + b.hi[1]) * 0.5f, (b.lo[2] + b.hi[2]) * 0.5f, u2f(ref)));
        hs.lights.push_back(make_float4(m.emitted[0], m.emitted[1], m.emitted[2], 0.f));
    }
    hs.n_lights = d->n_lights;
    if (hs.lights.empty()) { hs.lights.push_back(make_float4(0.f, 0.f, 0.f, u2f(REF_NONE))); hs.lights.push_back(make_float4(0.f, 0.f, 0.f, 0.f)); }

    lap("spheres, quads");
    // ---- instances
    hs.instances.resize(d->n_instances);
    std::vector<BuildPrim> top;
    std::vector<uint32_t> top_moving;
    auto add_sphere = [&](uint32_t i, std::vector<BuildPrim>& prims, std::vector<uint32_t>& mov) {
        if (smoving[i]) mov.push_back(i); else prims.push_back(make_prim(make_ref(KIND_SPHERE, i), sbox[i]));
    };
    for (uint32_t i = 0; i < d->n_top_spheres; ++i) add_sphere(i, top, top_moving);
    for (uint32_t i = 0; i < d->n_top_quads; ++i) top.push_back(make_prim(make_ref(KIND_QUAD, i | qtri[i]), qbox[i]));
    uint32_t n_flat = 0, blas_nodes_before = 0;
    std::vector<std::pair<uint32_t, Box>> inst_leaves;
    std::vector<std::vector<uint32_t>> inst_moving(d->n_instances);
    std::vector<std::vector<BuildPrim>> inst_prims(d->n_instances);
    for (uint32_t k = 0; k < d->n_instances; ++k) {
        const rtw_instance& in = d->instances[k];
        if ((uint64_t)in.first_sphere + in.n_spheres > ns || (uint64_t)in.first_quad + in.n_quads > nq) {
            err = "instance primitive range out of bounds"; return RTW_E_INVALID;
        }
        if (in.n_spheres && in.first_sphere < d->n_top_spheres) { err = "instance sphere range overlaps top-level spheres"; return RTW_E_INVALID; }
        if (in.n_quads && in.first_quad < d->n_top_quads) { err = "instance quad range overlaps top-level quads"; return RTW_E_INVALID; }
        if (!finite3(in.translation) || !finite3(in.rotation) || std::isnan(in.density)) { err = "non-finite instance"; return RTW_E_INVALID; }
        if (in.density_kind != RTW_DENSITY_SURFACE && in.density_kind != RTW_DENSITY_CONST) { err = "unknown density_kind"; return RTW_E_INVALID; }
        Instance& o = hs.instances[k]; std::memset(&o, 0, sizeof o);
        o.index = k; o.density = in.density;
        for (int a = 0; a < 3; ++a) o.t[a] = in.translation[a];
        bool rot0 = in.rotation[0] == 0.f && in.rotation[1] == 0.f && in.rotation[2] == 0.f;
        bool t0 = in.translation[0] == 0.f && in.translation[1] == 0.f && in.translation[2] == 0.f;
        float neg[3] = {-in.rotation[0], -in.rotation[1], -in.rotation[2]};
        euler_rows(neg, o.inv_rows);              // instance.rs:258  .rotated(-self.rotation)
        euler_rows(in.rotation, o.fwd_rows);      // instance.rs:303-306
        if (rot0) o.flags |= INST_ROT_IDENTITY;
        if (in.density_kind == RTW_DENSITY_CONST) o.flags |= INST_VOLUME;
        if (in.n_spheres == 1 && in.n_quads == 0 && !smoving[in.first_sphere]) o.flags |= INST_ONE_SPHERE;
        o.level.root = REF_NONE;
        if (rot0 && t0 && in.density_kind == RTW_DENSITY_SURFACE) {
            // An untransformed surface instance is exactly its primitives: rotated(0) and "- 0" are
            // identities in f32, so they join the top level (C5's 400 ground boxes).
            for (uint32_t i = 0; i < in.n_spheres; ++i) add_sphere(in.first_sphere + i, top, top_moving);
            for (uint32_t i = 0; i < in.n_quads; ++i) top.push_back(make_prim(make_ref(KIND_QUAD, (in.first_quad + i) | qtri[in.first_quad + i]), qbox[in.first_quad + i]));
            ++n_flat;
            continue;
        }
        for (uint32_t i = 0; i < in.n_spheres; ++i) add_sphere(in.first_sphere + i, inst_prims[k], inst_moving[k]);
        for (uint32_t i = 0; i < in.n_quads; ++i) inst_prims[k].push_back(make_prim(make_ref(KIND_QUAD, (in.first_quad + i) | qtri[in.first_quad + i]), qbox[in.first_quad + i]));
        // iaabb.rs:23-76: (qaabb box U saabb box); an empty inner BVH contributes [0,0] (aabb.rs:75-83)
        Box sb, qb;
        if (in.n_spheres) { sb = box_empty(); for (uint32_t i = 0; i < in.n_spheres; ++i) box_grow(sb, sbox[in.first_sphere + i]); }
        else { for (int a = 0; a < 3; ++a) sb.lo[a] = sb.hi[a] = 0.f; }
        if (in.n_quads) { qb = box_empty(); for (uint32_t i = 0; i < in.n_quads; ++i) box_grow(qb, qbox[in.first_quad + i]); }
        else { for (int a = 0; a < 3; ++a) qb.lo[a] = qb.hi[a] = 0.f; }
        Box ub = qb; box_grow(ub, sb);
        Box lb = box_empty();
        for (int c = 0; c < 8; ++c) {
            V3 p = v3((c & 4) ? ub.hi[0] : ub.lo[0], (c & 2) ? ub.hi[1] : ub.lo[1], (c & 1) ? ub.hi[2] : ub.lo[2]);
            V3 r = rot_apply(o.fwd_rows, p);
            lb.lo[0] = minf(lb.lo[0], r.x); lb.lo[1] = minf(lb.lo[1], r.y); lb.lo[2] = minf(lb.lo[2], r.z);
            lb.hi[0] = maxf(lb.hi[0], r.x); lb.hi[1] = maxf(lb.hi[1], r.y); lb.hi[2] = maxf(lb.hi[2], r.z);
        }
        for (int a = 0; a < 3; ++a) { lb.lo[a] = lb.lo[a] + in.translation[a]; lb.hi[a] = lb.hi[a] + in.translation[a]; }
        if (!finite3(lb.lo) || !finite3(lb.hi)) { err = "instance bounds are not finite"; return RTW_E_INVALID; }
        inst_leaves.push_back({k, lb});
    }
    lap("instances");
    // ---- baked instances (rtw_host.h): world-space copies of the primitives of rigidly placed surface instances
    std::vector<BuildPrim> baked_prims;                  // leaves that replace the baked instances' leaves
    std::vector<char> inst_baked(d->n_instances, 0);
    for (auto& il : inst_leaves) {
        const uint32_t k = il.first;
        const rtw_instance& in = d->instances[k];
        const Instance& o = hs.instances[k];
        if ((o.flags & INST_VOLUME) || !inst_moving[k].empty() || inst_prims[k].empty()) continue;
        double M[3][3];
        if (!rigid_frame(o.inv_rows, M)) continue;
        // local -> world: p_w = M^T p_l + T (M orthogonal: the inverse of the world -> local map p_l = M (p_w - T))
        auto to_world = [&](V3 p, bool point) {
            double q[3];
            for (int a = 0; a < 3; ++a) q[a] = M[0][a] * p.x + M[1][a] * p.y + M[2][a] * p.z + (point ? (double)o.t[a] : 0.0);
            return v3((float)q[0], (float)q[1], (float)q[2]);
        };
        bool ok = true;
        std::vector<BuildPrim> mine;
        std::vector<float4> new_spheres; std::vector<Quad> new_quads;
        std::vector<uint2> src_s, src_q;
        const uint32_t iref = make_ref(KIND_INSTANCE, k);
        for (uint32_t i = 0; i < in.n_spheres && ok; ++i) {
            const uint32_t si = in.first_sphere + i;
            const float4 sp = hs.spheres[si];
            const V3 c = to_world(v3(sp.x, sp.y, sp.z), true);
            if (!std::isfinite(c.x) || !std::isfinite(c.y) || !std::isfinite(c.z)) { ok = false; break; }
            Box b;
            const float cc[3] = {c.x, c.y, c.z};
            for (int a = 0; a < 3; ++a) {
                // the copy's centre is rounded to f32: a few ulps of slack around centre +- radius
                const float pad = 4.0f * 1.1920929e-7f * (std::fabs(cc[a]) + std::fabs(sp.w));
                b.lo[a] = cc[a] - std::fabs(sp.w) - pad; b.hi[a] = cc[a] + std::fabs(sp.w) + pad;
            }
            const uint32_t idx = (uint32_t)(hs.spheres.size() + new_spheres.size());
            new_spheres.push_back(make_float4(c.x, c.y, c.z, sp.w));
            src_s.push_back(make_uint2(iref, make_ref(KIND_SPHERE, si)));
            mine.push_back(make_prim(make_ref(KIND_SPHERE, idx), b));
        }
        for (uint32_t i = 0; i < in.n_quads && ok; ++i) {
            const uint32_t qi = in.first_quad + i;
            const Quad& ql = hs.quads[qi];
            const V3 Q = to_world(v3(ql.q1.x, ql.q1.y, ql.q1.z), true);
            const V3 u = to_world(v3(ql.q1.w, ql.q2.x, ql.q2.y), false), v = to_world(v3(ql.q2.z, ql.q2.w, ql.q3.x), false);
            const Quad qw = make_quad(Q, u, v);
            const float* f = &qw.q0.x;
            for (int a = 0; a < 16; ++a) if (!std::isfinite(f[a])) ok = false;
            if (!ok) break;
            Box b = quad_box(Q, u, v, qtri[qi] != 0u);
            for (int a = 0; a < 3; ++a) {
                const float pad = 4.0f * 1.1920929e-7f * std::max(std::fabs(b.lo[a]), std::fabs(b.hi[a]));
                b.lo[a] -= pad; b.hi[a] += pad;
            }
            const uint32_t idx = (uint32_t)(hs.quads.size() + new_quads.size());
            new_quads.push_back(qw);
            src_q.push_back(make_uint2(iref, make_ref(KIND_QUAD, qi | qtri[qi])));
            mine.push_back(make_prim(make_ref(KIND_QUAD, idx | qtri[qi]), b));
        }
        if (!ok || hs.spheres.size() + new_spheres.size() >= REF_INDEX_LIMIT || hs.quads.size() + new_quads.size() >= REF_INDEX_LIMIT) continue;
        hs.spheres.insert(hs.spheres.end(), new_spheres.begin(), new_spheres.end());
        hs.quads.insert(hs.quads.end(), new_quads.begin(), new_quads.end());
        hs.baked_sphere_src.insert(hs.baked_sphere_src.end(), src_s.begin(), src_s.end());
        hs.baked_quad_src.insert(hs.baked_quad_src.end(), src_q.begin(), src_q.end());
        baked_prims.insert(baked_prims.end(), mine.begin(), mine.end());
        inst_baked[k] = 1;
        hs.has_baked = true;
    }
    if (hs.baked_sphere_src.empty()) hs.baked_sphere_src.push_back(make_uint2(REF_NONE, REF_NONE));
    if (hs.baked_quad_src.empty()) hs.baked_quad_src.push_back(make_uint2(REF_NONE, REF_NONE));

    auto record_level = [&](const std::vector<BuildPrim>& prims, int32_t owner) {
        LbvhLevelDesc lv; lv.first = (uint32_t)hs.lbvh_leaves.size(); lv.count = (uint32_t)prims.size(); lv.owner = owner; lv.pad = 0u;
        for (const BuildPrim& p : prims) {
            LbvhLeaf l; l.ref = p.ref; l.pad = 0u;
            for (int a = 0; a < 3; ++a) { l.lo[a] = p.box.lo[a]; l.hi[a] = p.box.hi[a]; }
            hs.lbvh_leaves.push_back(l);
        }
        hs.lbvh_levels.push_back(lv);
    };
    // ---- levels: top first (its nodes lead the array and stay hot), then each instance
    for (auto& il : inst_leaves)
        top.push_back(make_prim(make_ref(KIND_INSTANCE, il.first) |
                                ((hs.instances[il.first].flags & INST_ONE_SPHERE) ? REF_ONE_SPHERE_INSTANCE : 0u) |
                                ((hs.instances[il.first].flags & INST_VOLUME) ? REF_LATE_INSTANCE : 0u), il.second));
    hs.info = rtw_scene_info{};
    hs.info.n_tlas_leaves = (uint32_t)top.size();
    hs.moving = top_moving;
    lap("baked copies");
    uint32_t top_depth = 0, inner_depth = 0;
    uint32_t plain_top_depth = 0, baked_top_depth = 0, inner_depth_unbaked = 0;
    // ---- enclosing volumes (see rtw_host.h): one-sphere constant-density instances, not rotated,
    // whose sphere contains every other leaf box of the top level with a 0.1 % margin
    std::vector<char> top_drop(top.size(), 0);
    {
        std::vector<char>& drop = top_drop;
        auto far_corner = [](const Box& b, const float c[3]) {
            float d2 = 0.f;
            for (int a = 0; a < 3; ++a) { float e = std::max(std::fabs(b.lo[a] - c[a]), std::fabs(b.hi[a] - c[a])); d2 += e * e; }
            return std::sqrt(d2);
        };
        for (size_t i = 0; i < top.size(); ++i) {
            if (ref_kind(top[i].ref) != KIND_INSTANCE) continue;
            const Instance& in = hs.instances[instance_index(top[i].ref)];
            const uint32_t need = INST_ONE_SPHERE | INST_VOLUME | INST_ROT_IDENTITY;
            if ((in.flags & need) != need || !(in.density > 0.f)) continue;
            const rtw_instance& di = d->instances[in.index];
            const float4 sp = hs.spheres[di.first_sphere];
            HostScene::Enclosing e;
            for (int a = 0; a < 3; ++a) e.c[a] = (&sp.x)[a] + in.t[a];
            e.r = std::fabs(sp.w) * 0.999f;
            e.mov_far = 0.f; e.mov_speed = 0.f;
            bool ok = true;
            for (size_t j = 0; j < top.size() && ok; ++j)
                if (j != i && !(far_corner(top[j].box, e.c) < e.r)) ok = false;
            auto moving_ok = [&](const std::vector<uint32_t>& mv) {
                for (uint32_t m : mv) {
                    e.mov_far = std::max(e.mov_far, far_corner(sbox[m], e.c));
                    const float4 v = hs.sphere_vel[m];
                    e.mov_speed = std::max(e.mov_speed, std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z));
                }
            };
            moving_ok(top_moving);
            // moving spheres inside transformed instances: their boxes are not in world space; such a
            // scene keeps the volume (nothing in the reference's scenes does this)
            for (auto& il : inst_leaves) if (!inst_moving[il.first].empty()) ok = false;
            if (!ok || !(e.mov_far < e.r)) continue;
            drop[i] = 1;
            hs.enclosing.push_back(e);
        }
    }
    // The top level without the enclosing volumes is built by SAH; the one with them (what the strict kernels and
    // renders from outside walk) is that tree under one more node holding the dropped leaves (Builder::wrap).
    // The three groups of levels - the reference's top levels, the baked ones, the instances' own - are independent:
    // each is built by its own thread into its own node array, and the arrays are then laid end to end (top levels
    // first: their nodes lead the array and stay hot) with the node references shifted (rtw_scene_create is part of
    // the end-to-end frame: 4.6 ms of builds become the 2.3 ms of the longest).
    std::vector<BuildPrim> dropped;
    for (size_t i = 0; i < top.size(); ++i) if (top_drop[i]) dropped.push_back(top[i]);
    const uint32_t n_mov = (uint32_t)top_moving.size();
    const bool has_enclosing = !hs.enclosing.empty();
    for (auto& il : inst_leaves)
        if (!(hs.instances[il.first].flags & INST_ONE_SPHERE)) {
            ++hs.n_walked_instances;
            if (!inst_baked[il.first]) ++hs.n_walked_instances_baked;
        }
    std::vector<BuildPrim> rest, all, inside;
    if (has_enclosing) for (size_t i = 0; i < top.size(); ++i) if (!top_drop[i]) rest.push_back(top[i]);
    if (hs.has_baked) {
        // the same two top levels with the baked instances' leaves replaced by their primitives' world-space copies
        for (size_t i = 0; i < top.size(); ++i) {
            const bool is_inst = ref_kind(top[i].ref) == KIND_INSTANCE;
            if (is_inst && inst_baked[instance_index(top[i].ref)]) continue;
            all.push_back(top[i]);
            if (!top_drop[i]) inside.push_back(top[i]);
        }
        all.insert(all.end(), baked_prims.begin(), baked_prims.end());
        inside.insert(inside.end(), baked_prims.begin(), baked_prims.end());
    }
    std::vector<uint32_t> inst_mov_first(d->n_instances, 0u);
    for (auto& il : inst_leaves) {
        inst_mov_first[il.first] = (uint32_t)hs.moving.size();
        hs.moving.insert(hs.moving.end(), inst_moving[il.first].begin(), inst_moving[il.first].end());
    }
    std::vector<Node> nodes_plain, nodes_baked, nodes_inst;
    auto build_plain = [&]() {
        Builder b(nodes_plain);
        if (has_enclosing) {
            hs.top_inside = b.level(rest, 0u, n_mov);
            plain_top_depth = b.max_pushes;
            hs.top = b.wrap(hs.top_inside, b.max_pushes, dropped);
            plain_top_depth = std::max(plain_top_depth, b.max_pushes);
        } else {
            hs.top = b.level(top, 0u, n_mov);
            plain_top_depth = b.max_pushes;
        }
    };
    auto build_baked = [&]() {
        if (!hs.has_baked) return;
        Builder b(nodes_baked);
        if (has_enclosing) {
            hs.top_baked_inside = b.level(inside, 0u, n_mov);
            baked_top_depth = b.max_pushes;
            hs.top_baked = b.wrap(hs.top_baked_inside, b.max_pushes, dropped);
            baked_top_depth = std::max(baked_top_depth, b.max_pushes);
        } else {
            hs.top_baked = b.level(all, 0u, n_mov);
            baked_top_depth = b.max_pushes;
        }
    };
    auto build_instances = [&]() {
        Builder b(nodes_inst);
        for (auto& il : inst_leaves) {
            const uint32_t k = il.first;
            hs.instances[k].level = b.level(inst_prims[k], inst_mov_first[k], (uint32_t)inst_moving[k].size());
            inner_depth = std::max(inner_depth, b.max_pushes);
            if (!inst_baked[k]) inner_depth_unbaked = std::max(inner_depth_unbaked, b.max_pushes);
        }
    };
    {
        const size_t work = top.size() + all.size() + hs.spheres.size() + hs.quads.size();
        if (work >= 512u) {                       // small scenes: a thread costs more than the build
            std::thread t1(build_baked), t2(build_instances);
            build_plain();
            t1.join(); t2.join();
        } else { build_plain(); build_baked(); build_instances(); }
    }
    // lay the three node arrays end to end; references below NODE_LIMIT are node indices
    auto shift_level = [](Level& lv, uint32_t off) { if (lv.root < NODE_LIMIT) lv.root += off; };
    auto append_nodes = [&](std::vector<Node>& src) {
        const uint32_t off = (uint32_t)hs.nodes.size();
        for (Node& n : src) {
            uint32_t* r = &n.ref.x;
            for (int k = 0; k < 4; ++k) if (r[k] < NODE_LIMIT) r[k] += off;
        }
        hs.nodes.insert(hs.nodes.end(), src.begin(), src.end());
        return off;
    };
    {
        const uint32_t off = append_nodes(nodes_plain);
        shift_level(hs.top, off); if (has_enclosing) shift_level(hs.top_inside, off);
    }
    hs.info.n_tlas_nodes = (uint32_t)hs.nodes.size();
    if (hs.has_baked) {
        const uint32_t off = append_nodes(nodes_baked);
        shift_level(hs.top_baked, off); if (has_enclosing) shift_level(hs.top_baked_inside, off);
    }
    blas_nodes_before = (uint32_t)hs.nodes.size();
    {
        const uint32_t off = append_nodes(nodes_inst);
        for (auto& il : inst_leaves) shift_level(hs.instances[il.first].level, off);
    }
    if (has_enclosing) record_level(rest, -2);
    record_level(top, -1);
    if (hs.has_baked) { if (has_enclosing) record_level(inside, -4); record_level(all, -3); }
    for (auto& il : inst_leaves) record_level(inst_prims[il.first], (int32_t)il.first);
    top_depth = std::max(plain_top_depth, baked_top_depth);
    lap("instance levels");
    // the half-float copy of the nodes (fast kernels): boxes rounded outwards
    hs.nodes16.resize(hs.nodes.size());
    for (size_t i = 0; i < hs.nodes.size(); ++i) {
        const Node& n = hs.nodes[i];
        auto dn = [](float4 v) { return make_uint2(pack_half2_rd(v.x, v.y), pack_half2_rd(v.z, v.w)); };
        auto up = [](float4 v) { return make_uint2(pack_half2_ru(v.x, v.y), pack_half2_ru(v.z, v.w)); };
        uint2 lx = dn(n.lox), ly = dn(n.loy), lz = dn(n.loz), hx = up(n.hix), hy = up(n.hiy), hz = up(n.hiz);
        Node16& o = hs.nodes16[i];
        o.a = make_uint4(lx.x, lx.y, ly.x, ly.y); o.b = make_uint4(lz.x, lz.y, hx.x, hx.y); o.c = make_uint4(hy.x, hy.y, hz.x, hz.y);
        o.ref = n.ref;
    }
    hs.one_sphere.assign(4 * (size_t)std::max<uint32_t>(d->n_instances, 1u), make_float4(0.f, 0.f, 0.f, 0.f));
    for (uint32_t k = 0; k < d->n_instances; ++k) {
        const Instance& in = hs.instances[k];
        if (!(in.flags & INST_ONE_SPHERE)) continue;
        const float4 sp = hs.spheres[ref_index(in.level.root)];
        hs.one_sphere[4 * k + 0] = make_float4(in.t[0], in.t[1], in.t[2], in.density);
        hs.one_sphere[4 * k + 1] = sp;
        hs.one_sphere[4 * k + 2] = make_float4(in.level.lo[0], in.level.lo[1], in.level.lo[2], u2f(in.flags));
        hs.one_sphere[4 * k + 3] = make_float4(in.level.hi[0], in.level.hi[1], in.level.hi[2], u2f(in.index));
    }
    hs.info.n_blas_nodes = (uint32_t)hs.nodes.size() - blas_nodes_before;
    hs.info.n_flattened_instances = n_flat;
    for (char b : inst_baked) hs.info.n_baked_instances += b ? 1u : 0u;
    hs.inst_baked.assign(inst_baked.begin(), inst_baked.end());
    hs.info.bvh_max_depth = top_depth + inner_depth;
    hs.top_depth = top_depth; hs.inner_depth = inner_depth;
    hs.any_levels = hs.n_walked_instances > 0;
    hs.fast_top_depth = hs.has_baked ? baked_top_depth : plain_top_depth;
    hs.fast_inner_depth = hs.has_baked ? inner_depth_unbaked : inner_depth;
    hs.fast_levels = (hs.has_baked ? hs.n_walked_instances_baked : hs.n_walked_instances) > 0;
    if (stack_words(top_depth, inner_depth) > 160u) { err = "BVH deeper than the traversal stack"; return RTW_E_UNSUPPORTED; }
    if (hs.nodes.empty()) { hs.nodes.emplace_back(); std::memset(&hs.nodes[0], 0, sizeof(Node)); hs.nodes16.emplace_back(); std::memset(&hs.nodes16[0], 0, sizeof(Node16)); }
    if (hs.moving.empty()) hs.moving.push_back(0u);
    for (int a = 0; a < 3; ++a) hs.background[a] = d->background[a];
    if (texel_job.t.joinable()) texel_job.t.join();
    lap("half-float nodes, records");
    return RTW_OK;
}

// Viewport::new (viewport.rs:308-401). `upper_left_corner` is a direction: the origin is not added.
void camera_init(rtw_camera* cam, uint32_t width, float aspect_ratio, float vfov, const float origin[3],
                 const float direction[3], const float vup[3], float lens_radius) {
    V3 c_dir = v3(direction[0], direction[1], direction[2]), c_vup = v3(vup[0], vup[1], vup[2]);
    V3 w = -c_dir;
    V3 u = unit(cross(c_vup, w));
    V3 v = cross(w, u);
    uint64_t height = (uint64_t)((float)width / aspect_ratio);
    float h = std::tan(vfov * 3.14159265358979323846f / 360.0f);
    float viewport_height = 2.0f * h;
    float viewport_width = aspect_ratio * viewport_height;
    V3 viewport_u = u * viewport_width;
    V3 viewport_v = -v * viewport_height;
    V3 pixel_delta_u = viewport_u / (float)width;
    V3 pixel_delta_v = viewport_v / (float)height;
    V3 viewport_upper_left = -w - viewport_u / 2.0f - viewport_v / 2.0f;
    V3 pixel00_loc = viewport_upper_left + (pixel_delta_u + pixel_delta_v) * 0.5f;
    std::memset(cam, 0, sizeof *cam);
    cam->width = width; cam->height = (uint32_t)height;
    auto put = [](float* d, V3 s) { d[0] = s.x; d[1] = s.y; d[2] = s.z; };
    put(cam->origin, v3(origin[0], origin[1], origin[2])); put(cam->pixel00, pixel00_loc);
    put(cam->delta_u, pixel_delta_u); put(cam->delta_v, pixel_delta_v);
    put(cam->u, u); put(cam->v, v);
    cam->lens_radius = lens_radius;
}

int make_render_params(const rtw_camera* cam, const rtw_render_opts* o, RenderParams& P, uint32_t& total,
                       std::string& err) {
    if (!cam || !o) { err = "null camera / options"; return RTW_E_INVALID; }
    if (cam->width == 0 || cam->height == 0) { err = "empty image"; return RTW_E_INVALID; }
    if ((uint64_t)cam->width * cam->height >= (1ull << 32)) { err = "image too large"; return RTW_E_UNSUPPORTED; }
    std::memset(&P, 0, sizeof P);
    P.cam = *cam;
    P.integrator = o->integrator; P.mode = o->mode; P.depth = o->depth;
    switch (o->integrator) {   // the constants each reference closure hard-codes
        case RTW_INT_GRADIENT:   P.mint = 0.001f; P.maxt = 100000.0f; break;   // ray_color.rs:17-18
        case RTW_INT_D:          P.mint = 0.001f; P.maxt = 1000.0f; break;     // main.rs:22-23
        case RTW_INT_BG_COLOR:   P.mint = 0.001f; P.maxt = 10000.0f; break;    // ray_color.rs:48-49
        case RTW_INT_NORMAL:     P.mint = 0.0f;   P.maxt = 1000.0f; break;     // viewport.rs:525-526
        case RTW_INT_ALBEDO:     P.mint = 0.001f; P.maxt = 100000.0f; break;
        case RTW_INT_GLASS_TEST: P.mint = 0.001f; P.maxt = 1000.0f; break;     // glass_tests.rs:13-14
        case RTW_INT_LIGHT_CAST:
        case RTW_INT_LIGHT_BIASED: P.mint = 0.001f; P.maxt = 1000.0f; break;   // Scene::new(.., 0.001, 1000.0), Rust2 tests.rs:639-644
        default: err = "unknown integrator"; return RTW_E_INVALID;
    }
    if (o->maxt > 0.0f) { P.mint = o->mint; P.maxt = o->maxt; }
    P.s_root = 1; total = o->samples;
    if (o->mode == RTW_MODE_NO_RAND) total = 1;
    else if (o->mode == RTW_MODE_STRATIFIED) {    // viewport.rs:443
        P.s_root = (uint32_t)std::ceil(std::sqrt((float)o->samples));
        total = P.s_root * P.s_root;
    } else if (o->mode != RTW_MODE_RANDOM) { err = "unknown mode"; return RTW_E_INVALID; }
    if (total == 0) { err = "samples must be > 0"; return RTW_E_INVALID; }
    if (!(o->gamma > 0.0f) && o->output == RTW_OUT_RESOLVED) { err = "gamma must be > 0"; return RTW_E_INVALID; }
    P.time0 = (float)o->frame / o->fps;           // viewport.rs:279
    P.shutter_speed = o->shutter_speed;
    P.k0 = (uint32_t)o->seed; P.k1 = (uint32_t)(o->seed >> 32);
    P.light_weight = o->light_weight;
    if (o->integrator == RTW_INT_LIGHT_BIASED && !(o->light_weight >= 0.0f)) { err = "light_weight must be >= 0"; return RTW_E_INVALID; }
    P.sample_begin = o->sample_count ? o->sample_begin : 0u;
    P.sample_count = o->sample_count ? o->sample_count : total;
    P.row_begin = o->row_count ? o->row_begin : 0u;
    P.row_count = o->row_count ? o->row_count : cam->height;
    if ((uint64_t)P.sample_begin + P.sample_count > total) { err = "sample range exceeds the sample count"; return RTW_E_INVALID; }
    P.row_band = 0u; P.row_stride = 0u;
    if (o->row_stride != 0u) {        // interleaved bands: the last output row must still be an image row
        if (o->row_count == 0u || o->row_band == 0u || o->row_stride < o->row_band) { err = "row_stride needs row_count > 0 and 0 < row_band <= row_stride"; return RTW_E_INVALID; }
        P.row_band = o->row_band; P.row_stride = o->row_stride;
        const uint32_t last = P.row_count - 1u, b = last / P.row_band;
        if ((uint64_t)P.row_begin + (uint64_t)b * P.row_stride + (last - b * P.row_band) >= cam->height) { err = "row bands exceed the image"; return RTW_E_INVALID; }
    } else
    if ((uint64_t)P.row_begin + P.row_count > cam->height) { err = "row range exceeds the image"; return RTW_E_INVALID; }
    if (o->output != RTW_OUT_RESOLVED && o->output != RTW_OUT_SUM) { err = "unknown output kind"; return RTW_E_INVALID; }
    return RTW_OK;
}

}  // namespace rtw
