// rtw_lbvh.h — the optional on-GPU rebuild of the BVH (BASELINE north_star: "host SAH build, plus an
// optional on-GPU LBVH rebuild"). It replaces the reference's builders AABB::new /
// QuadAABB::new / IAABB::new (Rust/src/objects/aabb/aabb.rs:20-71, qaabb.rs:127-163,
// iaabb.rs:80-116: random-axis median split, sort + deep clone at every level) for callers that
// rebuild per frame.
//
// Per level (the scene's top level, the top level without enclosing volumes, every instance):
//   1. Morton code (30 bits) of every leaf's box centre inside the level's centroid bounds
//   2. radix sort of (code, leaf)
//   3. Karras 2012 binary radix tree over the sorted codes (ties broken by position)
//   4. bottom-up box fit (second arrival at a node computes it) with treelet restructuring on the way (lbvh_treelet_node;
//      LBVH_FIT_PASSES passes)
//   5. collapse to the 4-wide nodes the traversal walks (rtw_device.h Node / Node16), top down and breadth first: the
//      children of a wide node are the descendants of its binary node that a dynamic program, solved during the fit
//      (lbvh_dp_node), chose for the least summed wide-node area; the internal ones become the next wide nodes
// Leaves keep the primitive's own reference leaf box as the child box stored in their parent, exactly
// as in the host SAH build, so the closest hit of a ray is the same whatever tree it walks
// (rtw_trace.h: every primitive is reached through its own box; ties go to the lowest reference).
//
// The per-element steps are __host__ __device__: rtw_lbvh.cu runs them as kernels, the host-side
// test emulator (tests/emu) runs them in plain loops.
#pragma once
#include "rtw_device.h"
#if defined(__CUDACC__) || defined(RTW_LBVH_HOST_FP16)
#include <cuda_fp16.h>
#endif

namespace rtw {

struct LbvhLeaf {          // one traversable leaf of a level: reference + its reference leaf box
    float lo[3]; uint32_t ref;
    float hi[3]; uint32_t pad;
};
struct LbvhLevelDesc {     // a level's range in the concatenated leaf array
    uint32_t first, count;
    int32_t owner;         // -1: scene top level, -2: ... without enclosing volumes, -3 / -4: the same two with baked instances, k >= 0: instance k
    uint32_t pad;
};

RTW_HD uint32_t lbvh_expand10(uint32_t v) {      // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
// centre of the leaf box, normalised to the level's centroid bounds [clo, clo + 1/cinv]
RTW_HD uint32_t lbvh_morton(const LbvhLeaf& l, const float clo[3], const float cinv[3]) {
    uint32_t q[3];
    for (int a = 0; a < 3; ++a) {
        float c = 0.5f * l.lo[a] + 0.5f * l.hi[a];
        float u = (c - clo[a]) * cinv[a];
        u = !(u > 0.0f) ? 0.0f : (u > 1.0f ? 1.0f : u);      // NaN -> 0
        float s = u * 1024.0f;
        q[a] = s >= 1023.0f ? 1023u : (uint32_t)s;
    }
    return (lbvh_expand10(q[0]) << 2) | (lbvh_expand10(q[1]) << 1) | lbvh_expand10(q[2]);
}

RTW_HD int lbvh_clz(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
// common-prefix length of sorted keys i and j (Karras 2012, with the index as tie-break); -1 outside
RTW_HD int lbvh_delta(const uint32_t* codes, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = codes[i], b = codes[j];
    return a != b ? lbvh_clz(a ^ b) : 32 + lbvh_clz((uint32_t)i ^ (uint32_t)j);
}

// A child of a binary node: >= 0 is an internal node, < 0 is the sorted leaf ~child.
// Internal node i of n - 1 (n >= 2 leaves): its children and their parent links.
RTW_HD void lbvh_karras(const uint32_t* codes, int n, int i, int* left, int* right, int* parent_internal,
                        int* parent_leaf) {
    const int d = lbvh_delta(codes, n, i, i + 1) - lbvh_delta(codes, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(codes, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(codes, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(codes, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (lbvh_delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    const int lc = lo == gamma ? ~gamma : gamma;
    const int rc = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    left[i] = lc; right[i] = rc;
    if (lc < 0) parent_leaf[~lc] = i; else parent_internal[lc] = i;
    if (rc < 0) parent_leaf[~rc] = i; else parent_internal[rc] = i;
    if (i == 0) parent_internal[0] = -1;
}

struct LbvhBox { float lo[3], hi[3]; };
RTW_HD LbvhBox lbvh_union(const LbvhBox& a, const LbvhBox& b) {
    LbvhBox r;
    for (int k = 0; k < 3; ++k) { r.lo[k] = a.lo[k] <= b.lo[k] ? a.lo[k] : b.lo[k]; r.hi[k] = a.hi[k] >= b.hi[k] ? a.hi[k] : b.hi[k]; }
    return r;
}
RTW_HD LbvhBox lbvh_leaf_box(const LbvhLeaf& l) {
    LbvhBox r;
    for (int k = 0; k < 3; ++k) { r.lo[k] = l.lo[k]; r.hi[k] = l.hi[k]; }
    return r;
}

// What the fit / emit steps read: the level's sorted leaves and its binary tree.
struct LbvhTree {
    const LbvhLeaf* leaves;     // the level's leaves (unsorted), indexed through `order`
    const uint32_t* order;      // sorted position -> leaf index
    const int* left; const int* right;
    const LbvhBox* ibox;        // box of every internal node
    const uint32_t* dp_choice;  // the collapse chosen for every internal node (lbvh_dp_node)
    uint32_t node_base;         // index of the level's first wide node in Scene::nodes / nodes16
};
RTW_HD LbvhBox lbvh_child_box(const LbvhTree& t, int c) {
    return c < 0 ? lbvh_leaf_box(t.leaves[t.order[~c]]) : t.ibox[c];
}

// Loads of what another thread of the same kernel has written (the fit step): past the non-coherent L1 on the device.
#if defined(__CUDA_ARCH__)
#define RTW_LBVH_LD(p) __ldcg(p)
#else
#define RTW_LBVH_LD(p) (*(p))
#endif

RTW_HD float lbvh_half_area(const LbvhBox& b) {
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// The binary tree while it is being fitted: links and boxes are written by the thread that owns the subtree.
struct LbvhBuild {
    const LbvhLeaf* leaves; const uint32_t* order;
    int* left; int* right; int* parent_i; int* parent_l;
    LbvhBox* ibox;
    float* sah_cost;            // summed box area of the internal nodes of every node's subtree (lbvh_treelet_node)
    float* dp_cut;              // [4 * node + (k - 1)]: least summed wide-node area of the subtree handed up in <= k slots
    uint32_t* dp_choice;        // per node: bits 0-1 the wide node's own split (slots to the left, 1..3), bits 2k-2 .. 2k-1
                                // (k = 2..4) the split when the subtree is handed up in k slots (0 = as one wide node)
};
RTW_HD LbvhBox lbvh_build_child_box(const LbvhBuild& b, int c) {
    if (c < 0) return lbvh_leaf_box(b.leaves[b.order[~c]]);
    LbvhBox r;
    const float* f = reinterpret_cast<const float*>(&b.ibox[c]);
    r.lo[0] = RTW_LBVH_LD(f + 0); r.lo[1] = RTW_LBVH_LD(f + 1); r.lo[2] = RTW_LBVH_LD(f + 2);
    r.hi[0] = RTW_LBVH_LD(f + 3); r.hi[1] = RTW_LBVH_LD(f + 4); r.hi[2] = RTW_LBVH_LD(f + 5);
    return r;
}

// The collapse into 4-wide nodes as a dynamic program over the binary tree, solved on the way up (the host SAH build
// does the same, scene_build.cpp dp_solve): choose, for every binary node, which descendants become the up-to-four
// children of its wide node so that the summed surface area of all wide nodes — the expected number of node visits
// of a random ray — is least. cut[k] = least cost of handing the node's subtree to its parent in at most k child slots
// (k = 1: as one wide node). Leaves cost nothing under any cut (each keeps its own box). Against opening the child with
// the largest box first: node visits per ray C4 1.54 -> 1.15 of the SAH tree's, C2 1.123 -> 1.106, C5 1.151 -> 1.166
// (tools/lbvh_quality.py).
RTW_HD float lbvh_dp_cut(const LbvhBuild& b, int c, int k) { return c >= 0 ? RTW_LBVH_LD(&b.dp_cut[4 * c + (k - 1)]) : 0.0f; }
RTW_HD void lbvh_dp_node(const LbvhBuild& b, int N, float area) {
    const int L = RTW_LBVH_LD(&b.left[N]), R = RTW_LBVH_LD(&b.right[N]);
    float cl[3], cr[3];
    for (int k = 1; k <= 3; ++k) { cl[k - 1] = lbvh_dp_cut(b, L, k); cr[k - 1] = lbvh_dp_cut(b, R, k); }
    float best = INFINITY; uint32_t bi = 1u;
    for (int i = 1; i <= 3; ++i) { const float c = cl[i - 1] + cr[3 - i]; if (c < best) { best = c; bi = (uint32_t)i; } }
    const float wide = area + best;
    uint32_t packed = bi;
    b.dp_cut[4 * N] = wide;
    for (int k = 2; k <= 4; ++k) {
        float bestk = wide; uint32_t ck = 0u;
        for (int i = 1; i < k; ++i) { const float c = cl[i - 1] + cr[k - i - 1]; if (c < bestk) { bestk = c; ck = (uint32_t)i; } }
        b.dp_cut[4 * N + (k - 1)] = bestk;
        packed |= ck << (2 * k - 2);
    }
    b.dp_choice[N] = packed;
}

// Leaves of the treelets lbvh_treelet_node restructures (0: one rotation per node instead, lbvh_fit_node)
#ifndef LBVH_TREELET
#define LBVH_TREELET 6
#endif
// Fit of internal node N, by the second thread to arrive there: every box below N is final and no other thread will
// touch N's subtree again, so this thread may also restructure it. One tree rotation first (Kensler 2008): a child A of
// N changes places with a grandchild G on the other side (child of B) when that shrinks B's box — the Morton order
// pairs leaves by the bits of their centres, not by their extent, and the bottom-up pass undoes the worst of it
// (node visits per ray against the host SAH tree, tools/lbvh_quality.py: C5 1.31 -> 1.17, C2 1.24 -> 1.13). The largest
// gain of the four candidates wins, ties and NaN boxes leave the tree alone. Then N's box. (Grandchild <-> grandchild
// exchanges as two more candidates changed nothing measurable: C5 1.151 -> 1.158, C2 1.123 -> 1.118.)
RTW_HD void lbvh_treelet_node(const LbvhBuild& b, int N);
RTW_HD void lbvh_fit_node(const LbvhBuild& b, int N) {
    if (LBVH_TREELET != 0) { lbvh_treelet_node(b, N); return; }    // (below; the rotation that follows is the A/B alternative)
    int ch[2] = {RTW_LBVH_LD(&b.left[N]), RTW_LBVH_LD(&b.right[N])};
    LbvhBox cb[2] = {lbvh_build_child_box(b, ch[0]), lbvh_build_child_box(b, ch[1])};
    int gc[2][2] = {{0, 0}, {0, 0}}; LbvhBox gb[2][2];
    float best_gain = 0.0f; int bs = -1, bg = -1;
    for (int s = 0; s < 2; ++s) {                    // A = ch[s] against the children of B = ch[1 - s]
        const int B = ch[1 - s];
        if (B < 0) continue;
        gc[s][0] = RTW_LBVH_LD(&b.left[B]); gc[s][1] = RTW_LBVH_LD(&b.right[B]);
        gb[s][0] = lbvh_build_child_box(b, gc[s][0]); gb[s][1] = lbvh_build_child_box(b, gc[s][1]);
        const float area_b = lbvh_half_area(cb[1 - s]);
        for (int g = 0; g < 2; ++g) {                // grandchild g goes up, the other one stays and is joined by A
            const float gain = area_b - lbvh_half_area(lbvh_union(cb[s], gb[s][1 - g]));
            if (gain > best_gain) { best_gain = gain; bs = s; bg = g; }
        }
    }
    if (bs >= 0) {
        const int A = ch[bs], B = ch[1 - bs], G = gc[bs][bg];
        if (bg == 0) b.left[B] = A; else b.right[B] = A;
        if (bs == 0) b.left[N] = G; else b.right[N] = G;
        if (A < 0) b.parent_l[~A] = B; else b.parent_i[A] = B;
        if (G < 0) b.parent_l[~G] = N; else b.parent_i[G] = N;
        const LbvhBox nb = lbvh_union(cb[bs], gb[bs][1 - bg]);
        b.ibox[B] = nb;
        cb[bs] = gb[bs][bg]; cb[1 - bs] = nb;
    }
    b.ibox[N] = lbvh_union(cb[0], cb[1]);
    if (bs >= 0) lbvh_dp_node(b, ch[1 - bs], lbvh_half_area(cb[1 - bs]));   // B has new children: its table first
    lbvh_dp_node(b, N, lbvh_half_area(b.ibox[N]));
}

// Treelet restructuring (Karras & Aila 2013) in place of the single rotation, by the thread that owns N's subtree: the
// treelet is N plus the internal nodes below it opened largest box first until it has LBVH_TREELET leaves (subtrees or
// primitives); every way of joining those leaves into a binary tree is priced by the summed area of its internal nodes
// (a dynamic program over the 2^n leaf subsets: cost[S] = area[S] + min over the splits of S), and the cheapest replaces
// the treelet's internal nodes when it beats the present one. Subtree costs below the treelet's leaves are carried in
// `sah_cost`. The node ids of the treelet are reused; N stays its root.
// Node visits per ray against the host SAH tree (tools/lbvh_quality.py; C5 / C2 / C4): one rotation per node, four passes
// 1.127 / 1.106 / 1.149; treelets of 6 leaves, three passes 1.042 / 1.100 / 1.027; of 7 leaves, four passes 1.033 / 1.134 /
// 1.027 (C2, a 1000-unit ground sphere among 0.2-unit ones, is where the area heuristic and the camera's rays part ways);
// of 5 leaves, four passes 1.088 / 1.069 / 1.160. LBVH_TREELET 0 keeps the rotation.
#if !LBVH_TREELET
RTW_HD void lbvh_treelet_node(const LbvhBuild&, int) {}
#else
RTW_HD int lbvh_ctz(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
RTW_HD void lbvh_treelet_node(const LbvhBuild& b, int N) {
    constexpr int NL = LBVH_TREELET, NS = 1 << NL;
    int leaf[NL]; LbvhBox lbox[NL]; float lcost[NL]; int nl = 2;
    int inner[NL]; int ni = 0;
    float inner_area = 0.0f;
    leaf[0] = RTW_LBVH_LD(&b.left[N]); leaf[1] = RTW_LBVH_LD(&b.right[N]);
    lbox[0] = lbvh_build_child_box(b, leaf[0]); lbox[1] = lbvh_build_child_box(b, leaf[1]);
    while (nl < NL) {
        int pick = -1; float pa = -1.0f;
        for (int k = 0; k < nl; ++k) if (leaf[k] >= 0) { const float a = lbvh_half_area(lbox[k]); if (a > pa) { pa = a; pick = k; } }
        if (pick < 0) break;
        const int c = leaf[pick];
        inner[ni++] = c; inner_area += pa;
        leaf[pick] = RTW_LBVH_LD(&b.left[c]); leaf[nl] = RTW_LBVH_LD(&b.right[c]);
        lbox[pick] = lbvh_build_child_box(b, leaf[pick]); lbox[nl] = lbvh_build_child_box(b, leaf[nl]);
        ++nl;
    }
    float leaf_sum = 0.0f;
    for (int k = 0; k < nl; ++k) { lcost[k] = leaf[k] >= 0 ? RTW_LBVH_LD(&b.sah_cost[leaf[k]]) : 0.0f; leaf_sum += lcost[k]; }
    const uint32_t full = (1u << nl) - 1u;
    // boxes of all leaf subsets, each from the subset without its lowest leaf (one union per subset)
    LbvhBox sbox[NS];
    float cost[NS]; uint8_t part[NS];
    for (uint32_t S = 1u; S <= full; ++S) {
        const uint32_t low = S & (0u - S), rest = S ^ low;
        const int k = lbvh_ctz(low);
        sbox[S] = rest ? lbvh_union(sbox[rest], lbox[k]) : lbox[k];
    }
    const float narea = lbvh_half_area(sbox[full]);
    const float present = narea + inner_area + leaf_sum;
    bool restructure = false;
    if (nl >= 3) {
        for (uint32_t S = 1u; S <= full; ++S) {
            const uint32_t low = S & (0u - S), rest = S ^ low;
            if (rest == 0u) { cost[S] = lcost[lbvh_ctz(low)]; part[S] = 0; continue; }
            float best = INFINITY; uint32_t bp = low;
            for (uint32_t q = (rest - 1u) & rest;; q = (q - 1u) & rest) {          // P = low | q, q a proper subset of rest
                const uint32_t P = low | q;
                const float c = cost[P] + cost[S ^ P];
                if (c < best) { best = c; bp = P; }
                if (q == 0u) break;
            }
            cost[S] = lbvh_half_area(sbox[S]) + best; part[S] = (uint8_t)bp;
        }
        restructure = cost[full] < present;
    }
    if (!restructure) {
        b.ibox[N] = sbox[full];
        b.sah_cost[N] = present;
        lbvh_dp_node(b, N, narea);
        return;
    }
    // the new topology, parents before children; then boxes / costs / collapse tables children first
    uint32_t ord[NL]; int id_of[NS]; int no = 0, next = 0;
    uint32_t st[NL]; int sp = 0;
    st[sp++] = full; id_of[full] = N;
    while (sp > 0) {
        const uint32_t S = st[--sp];
        ord[no++] = S;
        const uint32_t P = part[S], Q = S ^ P;
        if (P & (P - 1u)) { id_of[P] = inner[next++]; st[sp++] = P; }
        if (Q & (Q - 1u)) { id_of[Q] = inner[next++]; st[sp++] = Q; }
    }
    for (int k = no - 1; k >= 0; --k) {
        const uint32_t S = ord[k], P = part[S], Q = S ^ P;
        const int me = id_of[S];
        const int cl = (P & (P - 1u)) ? id_of[P] : leaf[lbvh_ctz(P)];
        const int cr = (Q & (Q - 1u)) ? id_of[Q] : leaf[lbvh_ctz(Q)];
        b.left[me] = cl; b.right[me] = cr;
        if (cl < 0) b.parent_l[~cl] = me; else b.parent_i[cl] = me;
        if (cr < 0) b.parent_l[~cr] = me; else b.parent_i[cr] = me;
        b.ibox[me] = sbox[S];
        b.sah_cost[me] = cost[S];
        lbvh_dp_node(b, me, lbvh_half_area(sbox[S]));
    }
}
#endif

#ifndef LBVH_FIT_PASSES
#define LBVH_FIT_PASSES 3      // bottom-up passes of lbvh_fit_node over the tree, each one more round of restructuring
#endif

// The children of the wide node that stands for binary node i, as the dynamic program chose them (lbvh_dp_node): the
// node's own split gives its left subtree s slots and its right one 4 - s; a subtree handed k slots either stays closed
// (one child: a wide node of its own, or a leaf) or splits them between its two children in turn. Returns their number
// (2..4), left to right; a candidate >= 0 is an internal node that becomes a wide node, < 0 is the sorted leaf ~c.
RTW_HD int lbvh_pick(const LbvhTree& t, int i, int cand[4]) {
    int n = 0;
    int sc[6], sk[6]; int sp = 0;                       // (subtree, slots) still to place, leftmost on top
    const int s0 = (int)(t.dp_choice[i] & 3u);
    sc[sp] = t.right[i]; sk[sp++] = 4 - s0;
    sc[sp] = t.left[i]; sk[sp++] = s0;
    while (sp > 0) {
        const int c = sc[--sp], k = sk[sp];
        const int ch = (c >= 0 && k > 1) ? (int)((t.dp_choice[c] >> (2 * k - 2)) & 3u) : 0;
        if (c < 0 || ch == 0) { cand[n++] = c; continue; }
        sc[sp] = t.right[c]; sk[sp++] = k - ch;
        sc[sp] = t.left[c]; sk[sp++] = ch;
    }
    return n;
}

#if defined(__CUDACC__) || defined(RTW_LBVH_HOST_FP16)
RTW_HD uint32_t lbvh_half2_rd(float a, float b) {   // two floats -> two halves rounded towards -inf
    return (uint32_t)__half_as_ushort(__float2half_rd(a)) | ((uint32_t)__half_as_ushort(__float2half_rd(b)) << 16);
}
RTW_HD uint32_t lbvh_half2_ru(float a, float b) {   // ... towards +inf
    return (uint32_t)__half_as_ushort(__float2half_ru(a)) | ((uint32_t)__half_as_ushort(__float2half_ru(b)) << 16);
}

// Writes the wide node `me` of the level (both node formats) over the candidates lbvh_pick chose; child_pos[k] is the
// position among the level's wide nodes that internal candidate k was given.
RTW_HD void lbvh_emit(const LbvhTree& t, const int cand[4], int n, const uint32_t child_pos[4], uint32_t me_pos, Node* nodes,
                      Node16* nodes16) {
    const float inf = INFINITY;
    float lo[3][4], hi[3][4]; uint32_t ref[4];
    for (int k = 0; k < 4; ++k) {
        if (k < n) {
            const int c = cand[k];
            const LbvhBox b = lbvh_child_box(t, c);
            for (int a = 0; a < 3; ++a) { lo[a][k] = b.lo[a]; hi[a][k] = b.hi[a]; }
            ref[k] = c < 0 ? t.leaves[t.order[~c]].ref : make_ref(KIND_NODE, t.node_base + child_pos[k]);
        } else {
            for (int a = 0; a < 3; ++a) { lo[a][k] = inf; hi[a][k] = inf; }
            ref[k] = REF_NONE;
        }
    }
    const uint32_t me = t.node_base + me_pos;
    Node& nd = nodes[me];
    nd.lox = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]); nd.loy = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
    nd.loz = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]); nd.hix = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
    nd.hiy = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]); nd.hiz = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
    nd.ref = make_uint4(ref[0], ref[1], ref[2], ref[3]);
    nd.pad = make_uint4(0u, 0u, 0u, 0u);
    Node16& h = nodes16[me];
    h.a = make_uint4(lbvh_half2_rd(lo[0][0], lo[0][1]), lbvh_half2_rd(lo[0][2], lo[0][3]),
                     lbvh_half2_rd(lo[1][0], lo[1][1]), lbvh_half2_rd(lo[1][2], lo[1][3]));
    h.b = make_uint4(lbvh_half2_rd(lo[2][0], lo[2][1]), lbvh_half2_rd(lo[2][2], lo[2][3]),
                     lbvh_half2_ru(hi[0][0], hi[0][1]), lbvh_half2_ru(hi[0][2], hi[0][3]));
    h.c = make_uint4(lbvh_half2_ru(hi[1][0], hi[1][1]), lbvh_half2_ru(hi[1][2], hi[1][3]),
                     lbvh_half2_ru(hi[2][0], hi[2][1]), lbvh_half2_ru(hi[2][2], hi[2][3]));
    h.ref = nd.ref;
}
#endif

// Stack words a walk below a level built this way can need: three siblings per wide node on the way down.
RTW_HD uint32_t lbvh_pushes(uint32_t wide_depth) { return 3u * wide_depth; }

}  // namespace rtw
