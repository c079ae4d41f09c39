// TEST INFRASTRUCTURE ONLY — never part of librtw_b200.so, never loaded by the package.
// Compiles the very headers the CUDA kernels are made of (rtw_trace.h, rtw_shade.h) plus the host
// scene builder for the CPU, so `-m "not gpu"` tests can check the flattening, the SAH BVH and the
// traversal/shading logic against the oracle without a GPU. One thread, pixel by pixel.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "../../raytracing-in-a-weekend_b200/csrc/rtw_host.h"
#include "../../raytracing-in-a-weekend_b200/csrc/rtw_shade.h"

using namespace rtw;

extern "C" int emu_render(const rtw_scene_desc* d, const rtw_camera* cam, const rtw_render_opts* o, float* out,
                          uint64_t* stats /* paths, rays, nan */, rtw_scene_info* info) {
    HostScene hs; std::string err;
    int rc = build_host_scene(d, hs, err);
    if (rc) { std::fprintf(stderr, "emu: %s\n", err.c_str()); return rc; }
    if (info) *info = hs.info;
    Scene sc = hs.host_view();
    RenderParams P; uint32_t total = 0;
    rc = make_render_params(cam, o, P, total, err);
    if (rc) { std::fprintf(stderr, "emu: %s\n", err.c_str()); return rc; }
    uint64_t paths = 0, nan = 0; uint32_t rays32 = 0; uint64_t rays = 0;
    std::vector<uint32_t> stack_mem(stack_words(hs.top_depth, hs.inner_depth));
    Stack S; S.base = stack_mem.data(); S.stride = 1u;
    for (uint32_t jr = 0; jr < P.row_count; ++jr)
        for (uint32_t i = 0; i < cam->width; ++i) {
            V3 acc = v3(0, 0, 0);
            for (uint32_t s = P.sample_begin; s < P.sample_begin + P.sample_count; ++s) {
                Path p;
                generate(P, i, P.row_begin + jr, s, p);
                rays32 = 0;
                while (!path_step(sc, sc.noises, P, p, rays32, S)) {}
                rays += rays32;
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
    if (stats) { stats[0] = paths; stats[1] = rays; stats[2] = nan; }
    return 0;
}
