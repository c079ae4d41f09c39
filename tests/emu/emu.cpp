// TEST INFRASTRUCTURE ONLY — never part of librtw_b200.so, never loaded by the package.
// Compiles the very headers the CUDA kernels are made of (rtw_trace.h, rtw_shade.h) plus the host
// scene builder for the CPU, so `-m "not gpu"` tests can check the flattening, the SAH BVH and the
// traversal/shading logic against the oracle without a GPU. One thread, pixel by pixel.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#define RTW_LBVH_HOST_FP16 1
#include <algorithm>
#include "../../raytracing-in-a-weekend_b200/csrc/rtw_host.h"
#include "../../raytracing-in-a-weekend_b200/csrc/rtw_shade.h"

using namespace rtw;

// The LBVH rebuild of rtw_lbvh.cu, step for step, in plain loops over the same __host__ __device__
// element functions (rtw_lbvh.h): every level with two or more leaves is rebuilt behind the SAH nodes
// and becomes the level that is walked.
static void lbvh_rebuild_host(HostScene& hs) {
    uint32_t top_depth = hs.top_depth, inner_depth = hs.inner_depth;
    for (const LbvhLevelDesc& lv : hs.lbvh_levels) {
        const int n = (int)lv.count;
        if (n < 2) continue;
        const LbvhLeaf* leaves = hs.lbvh_leaves.data() + lv.first;
        float clo[3], chi[3], cinv[3];
        for (int a = 0; a < 3; ++a) { clo[a] = INFINITY; chi[a] = -INFINITY; }
        for (int i = 0; i < n; ++i)
            for (int a = 0; a < 3; ++a) {
                float c = 0.5f * leaves[i].lo[a] + 0.5f * leaves[i].hi[a];
                if (c == c) { clo[a] = std::min(clo[a], c); chi[a] = std::max(chi[a], c); }
            }
        for (int a = 0; a < 3; ++a) cinv[a] = chi[a] > clo[a] ? 1.0f / (chi[a] - clo[a]) : 0.0f;
        std::vector<uint32_t> codes(n), order(n), sorted(n);
        for (int i = 0; i < n; ++i) { codes[i] = lbvh_morton(leaves[i], clo, cinv); order[i] = (uint32_t)i; }
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return codes[a] < codes[b]; });
        for (int i = 0; i < n; ++i) sorted[i] = codes[order[i]];
        std::vector<int> left(n - 1), right(n - 1), parent_i(n - 1, -1), parent_l(n, -1);
        for (int i = 0; i < n - 1; ++i) lbvh_karras(sorted.data(), n, i, left.data(), right.data(), parent_i.data(), parent_l.data());
        std::vector<LbvhBox> ibox(n - 1);
        std::vector<uint32_t> flag(n - 1, 0u);
        LbvhTree t; t.leaves = leaves; t.order = order.data(); t.left = left.data(); t.right = right.data();
        t.ibox = ibox.data();
        LbvhBuild bt; bt.leaves = leaves; bt.order = order.data(); bt.left = left.data(); bt.right = right.data();
        bt.parent_i = parent_i.data(); bt.parent_l = parent_l.data(); bt.ibox = ibox.data();
        std::vector<float> dp_cut(4 * (size_t)(n - 1), 0.f); std::vector<uint32_t> dp_choice(n - 1, 0u);
        std::vector<float> sah_cost(n - 1, 0.f);
        bt.dp_cut = dp_cut.data(); bt.dp_choice = dp_choice.data(); bt.sah_cost = sah_cost.data();
        t.dp_choice = dp_choice.data();
        for (int pass = 0; pass < LBVH_FIT_PASSES; ++pass) {
            std::fill(flag.begin(), flag.end(), 0u);
            for (int p = 0; p < n; ++p)
                for (int cur = parent_l[p]; cur >= 0; cur = parent_i[cur]) {
                    if (flag[cur]++ == 0u) break;
                    lbvh_fit_node(bt, cur);
                }
        }
        // breadth-first collapse: queue position = position among the level's wide nodes (the binary root is wide node 0)
        t.node_base = (uint32_t)hs.nodes.size();
        hs.nodes.resize(hs.nodes.size() + (size_t)(n - 1)); hs.nodes16.resize(hs.nodes.size());
        std::vector<int> queue(1, 0);
        uint32_t wide_depth = 0;
        for (size_t head = 0; head < queue.size();) {
            const size_t level_end = queue.size();
            ++wide_depth;
            for (; head < level_end; ++head) {
                int cand[4]; uint32_t pos[4] = {0u, 0u, 0u, 0u};
                const int nc = lbvh_pick(t, queue[head], cand);
                for (int k = 0; k < nc; ++k) if (cand[k] >= 0) { pos[k] = (uint32_t)queue.size(); queue.push_back(cand[k]); }
                lbvh_emit(t, cand, nc, pos, (uint32_t)head, hs.nodes.data(), hs.nodes16.data());
            }
        }
        hs.nodes.resize(t.node_base + queue.size()); hs.nodes16.resize(hs.nodes.size());
        Level& L = lv.owner == -1 ? hs.top : lv.owner == -2 ? hs.top_inside : lv.owner == -3 ? hs.top_baked
                   : lv.owner == -4 ? hs.top_baked_inside : hs.instances[lv.owner].level;
        L.root = make_ref(KIND_NODE, t.node_base);
        for (int a = 0; a < 3; ++a) { L.lo[a] = ibox[0].lo[a]; L.hi[a] = ibox[0].hi[a]; }
        const uint32_t pushes = lbvh_pushes(wide_depth);
        if (lv.owner < 0) top_depth = std::max(top_depth, pushes); else inner_depth = std::max(inner_depth, pushes);
    }
    hs.top_depth = top_depth; hs.inner_depth = inner_depth;
}

extern "C" int emu_render_ex(const rtw_scene_desc* d, const rtw_camera* cam, const rtw_render_opts* o, float* out,
                             uint64_t* stats, rtw_scene_info* info, int use_lbvh);
extern "C" int emu_render(const rtw_scene_desc* d, const rtw_camera* cam, const rtw_render_opts* o, float* out,
                          uint64_t* stats /* paths, rays, nan, nodes, spheres, quads, instances, hits */, rtw_scene_info* info) {
    return emu_render_ex(d, cam, o, out, stats, info, 0);
}
extern "C" int emu_render_ex(const rtw_scene_desc* d, const rtw_camera* cam, const rtw_render_opts* o, float* out,
                             uint64_t* stats /* paths, rays, nan, nodes, spheres, quads, instances, hits */, rtw_scene_info* info, int use_lbvh) {
    HostScene hs; std::string err;
    int rc = build_host_scene(d, hs, err);
    if (rc) { std::fprintf(stderr, "emu: %s\n", err.c_str()); return rc; }
    if (use_lbvh & 1) lbvh_rebuild_host(hs);
    if (info) *info = hs.info;
    Scene sc = hs.host_view();
    if ((use_lbvh & 2) && hs.has_baked) sc.top = hs.top_baked;     // the top level with baked instances (rtw_host.h), as the fast kernels walk it
    // bit 2 (tools/lbvh_quality.py only): the top level without the enclosing volumes, which is what the kernels walk
    // when the camera sits inside all of them (rtw_api.cu); the volumes themselves are then missing from the image
    if ((use_lbvh & 4) && !hs.enclosing.empty()) sc.top = ((use_lbvh & 2) && hs.has_baked) ? hs.top_baked_inside : hs.top_inside;
    RenderParams P; uint32_t total = 0;
    rc = make_render_params(cam, o, P, total, err);
    if (rc) { std::fprintf(stderr, "emu: %s\n", err.c_str()); return rc; }
    uint64_t paths = 0, nan = 0; uint32_t rays32 = 0; uint64_t rays = 0;
    uint64_t work_total[5] = {0, 0, 0, 0, 0};   // nodes, spheres, quads, instances, hits (the strict walk's; tools/bvh_proxy.py)
    std::vector<uint32_t> stack_mem(stack_words(hs.top_depth, hs.inner_depth));
    Stack S; S.base = stack_mem.data(); S.stride = 1u;
    for (uint32_t jr = 0; jr < P.row_count; ++jr)
        for (uint32_t i = 0; i < cam->width; ++i) {
            V3 acc = v3(0, 0, 0);
            for (uint32_t s = P.sample_begin; s < P.sample_begin + P.sample_count; ++s) {
                Path p;
                generate(P, i, image_row(P, jr), s, p);
                rays32 = 0;
                Work w{0u, 0u, 0u, 0u, 0u};
                if (use_lbvh & 8) { while (!path_step<false>(sc, sc.noises, P, p, rays32, S, &w)) {} }   // the megakernel's Hit flavour
                else { while (!path_step<true>(sc, sc.noises, P, p, rays32, S, &w)) {} }
                rays += rays32;
                work_total[0] += w.nodes; work_total[1] += w.spheres; work_total[2] += w.quads; work_total[3] += w.instances; work_total[4] += w.hits;
                V3 c = path_result(p);
                if (!finite3(c)) ++nan;
                acc = acc + c; ++paths;
            }
            float* px = out + ((size_t)jr * cam->width + i) * 3;
            if (o->output == RTW_OUT_RESOLVED) {
                float ig = 1.0f / o->gamma;
                px[0] = std::pow(acc.x / (float)total, ig); px[1] = std::pow(acc.y / (float)total, ig); px[2] = std::pow(acc.z / (float)total, ig);
            } else { px[0] = acc.x; px[1] = acc.y; px[2] = acc.z; }
        }
    if (stats) { stats[0] = paths; stats[1] = rays; stats[2] = nan; for (int k = 0; k < 5; ++k) stats[3 + k] = work_total[k]; }
    return 0;
}
