// rtw_multi.cu — several CUDA devices behind one call of the C ABI (include/rtw.h, "several devices behind one
// call"; SURVEY 8b device list + split, 8e). The reference's caller is one Rust process that calls async_render /
// render_multi (Rust/src/viewport.rs:215-269, main.rs:76-80,412-414); this file gives that one process every GPU
// of the box without torchrun: the scene replicated per device, one host thread and one stream per share (the
// wavefront host loop polls, so a share needs a thread of its own), and
//   samples : per-share f32 sum buffers, combined on devices[0] by ONE kernel that reads the other shares' buffers
//             through NVLink peer memory, adds them in share order (deterministic) and resolves in the same pass
//             (staged peer copies where peer access is unavailable);
//   tiles   : row bands dealt round-robin, resolved where they were rendered, copied straight into out_rgb;
//   frames  : render_multi's frame k -> share k mod n.
// Built only on the public entry points of rtw_api.cu (rtw_scene_create, rtw_render_device, rtw_get_stats).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/rtw.h"

namespace rtw { void set_last_error(const std::string& msg); }

namespace {

struct Share {
    int device = 0;
    rtw_scene* scene = nullptr;
    cudaStream_t stream = nullptr;
    float* d_buf = nullptr; size_t d_floats = 0;
    float* d_stage = nullptr; size_t stage_floats = 0;       // devices[0] only: landing zone when a peer cannot be mapped
    bool peer_ok = true;                                      // devices[0] can read this share's buffer directly
    rtw_stats stats{};
    int rc = RTW_OK;
    std::string err;
};

// sum_k src[k][i], k in share order, then the resolve of render_row's last line (viewport.rs:301, :207-213)
__global__ void combine_resolve_kernel(float* __restrict__ out, const float* const* __restrict__ srcs, int n,
                                       unsigned long long n_values, float total, float inv_gamma) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (; i < n_values; i += stride) {
        float acc = srcs[0][i];
        for (int k = 1; k < n; ++k) acc += srcs[k][i];          // peers: loads over NVLink
        out[i] = inv_gamma != 0.0f ? powf(acc / total, inv_gamma) : acc;
    }
}

int cuda_rc(cudaError_t e) { return e == cudaErrorMemoryAllocation ? RTW_E_NOMEM : RTW_E_CUDA; }

}  // namespace

struct rtw_scene_multi {
    std::vector<Share> shares;
    const float** d_srcs = nullptr;        // on devices[0]
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float combine_ms = 0.f;
    bool have_stats = false;
};

namespace {

int fail(int code, const std::string& msg) { rtw::set_last_error(msg); return code; }

// Make `floats` floats of scratch available on the share's device.
int ensure_buf(Share& sh, size_t floats) {
    if (sh.d_floats >= floats) return RTW_OK;
    cudaError_t e = cudaSetDevice(sh.device);
    if (e == cudaSuccess && sh.d_buf) { cudaStreamSynchronize(sh.stream); e = cudaFree(sh.d_buf); sh.d_buf = nullptr; sh.d_floats = 0; }
    if (e == cudaSuccess) e = cudaMalloc(&sh.d_buf, floats * sizeof(float));
    if (e != cudaSuccess) { sh.err = std::string("share buffer: ") + cudaGetErrorString(e); cudaGetLastError(); return cuda_rc(e); }
    sh.d_floats = floats;
    return RTW_OK;
}

// Runs fn(k) for every share on its own thread (share 0 on the caller's); returns the first failure.
template <class F>
int for_each_share(rtw_scene_multi* ms, F fn) {
    const int n = (int)ms->shares.size();
    std::vector<std::thread> th;
    auto body = [&](int k) {
        Share& sh = ms->shares[k];
        sh.rc = RTW_OK; sh.err.clear();
        cudaError_t e = cudaSetDevice(sh.device);
        if (e != cudaSuccess) { sh.rc = RTW_E_CUDA; sh.err = cudaGetErrorString(e); return; }
        sh.rc = fn(k);
        if (sh.rc != RTW_OK && sh.err.empty()) sh.err = rtw_last_error();     // this thread's message
    };
    for (int k = 1; k < n; ++k) th.emplace_back(body, k);
    body(0);
    for (auto& t : th) t.join();
    for (int k = 0; k < n; ++k)
        if (ms->shares[k].rc != RTW_OK) return fail(ms->shares[k].rc, "share " + std::to_string(k) + " (device " +
                                                    std::to_string(ms->shares[k].device) + "): " + ms->shares[k].err);
    return RTW_OK;
}

struct DevGuard {
    int prev = -1;
    DevGuard() { if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; } }
    ~DevGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

uint32_t effective_samples(const rtw_render_opts* o) {      // viewport.rs:443: render() rounds up to a square
    if (o->mode == RTW_MODE_NO_RAND) return 1u;
    if (o->mode == RTW_MODE_STRATIFIED) { uint32_t r = (uint32_t)std::ceil(std::sqrt((float)o->samples)); return r * r; }
    return o->samples;
}

}  // namespace

extern "C" {

int rtw_scene_multi_device_count(const rtw_scene_multi* ms) { return ms ? (int)ms->shares.size() : 0; }

void rtw_scene_multi_destroy(rtw_scene_multi* ms) {
    if (!ms) return;
    DevGuard guard;
    for (auto& sh : ms->shares) {
        if (cudaSetDevice(sh.device) != cudaSuccess) { cudaGetLastError(); continue; }
        if (sh.stream) cudaStreamSynchronize(sh.stream);
        if (sh.scene) rtw_scene_destroy(sh.scene);
        if (sh.d_buf) cudaFree(sh.d_buf);
        if (sh.d_stage) cudaFree(sh.d_stage);
        if (sh.stream) cudaStreamDestroy(sh.stream);
    }
    if (!ms->shares.empty() && cudaSetDevice(ms->shares[0].device) == cudaSuccess) {
        if (ms->d_srcs) cudaFree(ms->d_srcs);
        if (ms->ev0) cudaEventDestroy(ms->ev0);
        if (ms->ev1) cudaEventDestroy(ms->ev1);
    }
    cudaGetLastError();
    delete ms;
}

int rtw_scene_multi_create(const rtw_scene_desc* desc, const int* devices, int n_devices, rtw_scene_multi** out) {
    if (!out) return fail(RTW_E_INVALID, "null out pointer");
    *out = nullptr;
    if (!desc || !devices || n_devices < 1 || n_devices > 64) return fail(RTW_E_INVALID, "bad device list");
    const int ndev = rtw_device_count();
    if (ndev == 0) return fail(RTW_E_CUDA, "no usable CUDA device; this library has no CPU path");
    for (int k = 0; k < n_devices; ++k)
        if (devices[k] < 0 || devices[k] >= ndev) return fail(RTW_E_INVALID, "device ordinal out of range");
    DevGuard guard;
    rtw_scene_multi* ms = new (std::nothrow) rtw_scene_multi();
    if (!ms) return fail(RTW_E_NOMEM, "out of host memory");
    ms->shares.resize(n_devices);
    for (int k = 0; k < n_devices; ++k) ms->shares[k].device = devices[k];
    // every share flattens, builds and uploads on its own thread (the host SAH build is per share: each
    // rtw_scene owns its arrays; n builds side by side cost the wall time of one)
    int rc = for_each_share(ms, [&](int k) {
        Share& sh = ms->shares[k];
        int r = rtw_scene_create(desc, sh.device, &sh.scene);
        if (r) return r;
        cudaError_t e = cudaStreamCreateWithFlags(&sh.stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { sh.err = cudaGetErrorString(e); return cuda_rc(e); }
        return (int)RTW_OK;
    });
    if (rc) { const std::string msg = rtw_last_error(); rtw_scene_multi_destroy(ms); return fail(rc, msg); }
    // devices[0] maps the other devices' memory (NVLink / NVSwitch on the B200 box)
    const int d0 = devices[0];
    cudaSetDevice(d0);
    for (int k = 1; k < n_devices; ++k) {
        Share& sh = ms->shares[k];
        if (sh.device == d0) { sh.peer_ok = true; continue; }
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, d0, sh.device) != cudaSuccess) { cudaGetLastError(); can = 0; }
        if (can) {
            cudaError_t e = cudaDeviceEnablePeerAccess(sh.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) can = 0;
            cudaGetLastError();
        }
        sh.peer_ok = can != 0;
    }
    cudaError_t e = cudaMalloc(&ms->d_srcs, 64 * sizeof(float*));
    if (e == cudaSuccess) e = cudaEventCreate(&ms->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ms->ev1);
    if (e != cudaSuccess) { const std::string msg = cudaGetErrorString(e); cudaGetLastError(); rtw_scene_multi_destroy(ms); return fail(cuda_rc(e), msg); }
    *out = ms;
    return RTW_OK;
}

int rtw_render_multi(rtw_scene_multi* ms, const rtw_camera* cam, const rtw_render_opts* opts, uint32_t split,
                     uint32_t band_rows, float* out_rgb) {
    if (!ms || !cam || !opts || !out_rgb) return fail(RTW_E_INVALID, "null argument");
    if (split != RTW_SPLIT_SAMPLES && split != RTW_SPLIT_TILES) return fail(RTW_E_INVALID, "rtw_render_multi: split must be samples or tiles");
    if (opts->output != RTW_OUT_RESOLVED && opts->output != RTW_OUT_SUM) return fail(RTW_E_INVALID, "unknown output kind");
    DevGuard guard;
    const int n = (int)ms->shares.size();
    const uint32_t W = cam->width;
    const uint32_t row0 = opts->row_count ? opts->row_begin : 0u, rows = opts->row_count ? opts->row_count : cam->height;
    const size_t n_values = (size_t)rows * W * 3u;
    ms->combine_ms = 0.f; ms->have_stats = false;
    for (auto& sh : ms->shares) std::memset(&sh.stats, 0, sizeof sh.stats);
    int rc;
    if (split == RTW_SPLIT_SAMPLES) {
        const uint32_t total = effective_samples(opts);
        const uint32_t s0 = opts->sample_count ? opts->sample_begin : 0u, sn = opts->sample_count ? opts->sample_count : total;
        if (sn == 0u || (uint64_t)s0 + sn > total) return fail(RTW_E_INVALID, "sample range exceeds the sample count");
        rc = for_each_share(ms, [&](int k) {
            Share& sh = ms->shares[k];
            int r = ensure_buf(sh, n_values);
            if (r) return r;
            const uint32_t base = sn / (uint32_t)n, rem = sn % (uint32_t)n;           // remainders to the first shares
            const uint32_t begin = s0 + (uint32_t)k * base + std::min<uint32_t>((uint32_t)k, rem);
            const uint32_t count = base + ((uint32_t)k < rem ? 1u : 0u);
            cudaError_t e;
            if (count == 0u) {                                                        // more shares than samples
                e = cudaMemsetAsync(sh.d_buf, 0, n_values * sizeof(float), sh.stream);
            } else {
                rtw_render_opts o = *opts;
                o.sample_begin = begin; o.sample_count = count; o.output = RTW_OUT_SUM;
                if ((r = rtw_render_device(sh.scene, cam, &o, sh.d_buf, sh.stream))) return r;
                e = cudaSuccess;
            }
            if (e == cudaSuccess) e = cudaStreamSynchronize(sh.stream);
            if (e != cudaSuccess) { sh.err = cudaGetErrorString(e); return cuda_rc(e); }
            if (count) rtw_get_stats(sh.scene, &sh.stats);
            return (int)RTW_OK;
        });
        if (rc) return rc;
        // ---- combine on devices[0]
        Share& s0h = ms->shares[0];
        cudaError_t e = cudaSetDevice(s0h.device);
        std::vector<const float*> srcs(n);
        size_t need_stage = 0;
        for (int k = 0; k < n; ++k) if (!ms->shares[k].peer_ok) need_stage += n_values;
        if (e == cudaSuccess && s0h.stage_floats < need_stage) {
            if (s0h.d_stage) cudaFree(s0h.d_stage);
            s0h.d_stage = nullptr; s0h.stage_floats = 0;
            e = cudaMalloc(&s0h.d_stage, need_stage * sizeof(float));
            if (e == cudaSuccess) s0h.stage_floats = need_stage;
        }
        size_t off = 0;
        for (int k = 0; k < n && e == cudaSuccess; ++k) {
            Share& sh = ms->shares[k];
            if (sh.peer_ok) { srcs[k] = sh.d_buf; continue; }
            e = cudaMemcpyPeerAsync(s0h.d_stage + off, s0h.device, sh.d_buf, sh.device, n_values * sizeof(float), s0h.stream);
            srcs[k] = s0h.d_stage + off; off += n_values;
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(ms->d_srcs, srcs.data(), n * sizeof(float*), cudaMemcpyHostToDevice, s0h.stream);
        if (e == cudaSuccess) e = cudaEventRecord(ms->ev0, s0h.stream);
        if (e == cudaSuccess) {
            const float inv_gamma = opts->output == RTW_OUT_RESOLVED ? 1.0f / opts->gamma : 0.0f;
            const int block = 256;
            const int grid = (int)std::min<unsigned long long>(148ull * 8ull, (n_values + block - 1) / block);
            combine_resolve_kernel<<<std::max(grid, 1), block, 0, s0h.stream>>>(s0h.d_buf, ms->d_srcs, n, n_values, (float)total, inv_gamma);
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaEventRecord(ms->ev1, s0h.stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(out_rgb, s0h.d_buf, n_values * sizeof(float), cudaMemcpyDeviceToHost, s0h.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s0h.stream);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms->combine_ms, ms->ev0, ms->ev1);
        if (e != cudaSuccess) { const std::string msg = std::string("combine: ") + cudaGetErrorString(e); cudaGetLastError(); return fail(cuda_rc(e), msg); }
    } else {
        const uint32_t band = band_rows ? band_rows : 8u;
        const uint32_t n_bands = (rows + band - 1u) / band;
        rc = for_each_share(ms, [&](int k) {
            Share& sh = ms->shares[k];
            // this share's bands k, k + n, k + 2n, ... as ONE render of interleaved rows (rtw_render_opts::row_stride):
            // a render per band would pay the pipeline's ramp and drain once per band
            uint32_t mine = 0, my_rows = 0;
            for (uint32_t b = (uint32_t)k; b < n_bands; b += (uint32_t)n) { ++mine; my_rows += std::min(band, rows - b * band); }
            if (mine == 0u) return (int)RTW_OK;
            int r = ensure_buf(sh, (size_t)my_rows * W * 3u);
            if (r) return r;
            rtw_render_opts o = *opts;
            o.row_begin = row0 + (uint32_t)k * band; o.row_count = my_rows;
            o.row_band = band; o.row_stride = (uint32_t)n * band;
            if ((r = rtw_render_device(sh.scene, cam, &o, sh.d_buf, sh.stream))) return r;
            // bands back to their places in the host image: one strided copy for the full bands, one for a short last band
            const size_t band_bytes = (size_t)band * W * 3u * sizeof(float);
            const uint32_t last_rows = my_rows - (mine - 1u) * band;          // rows of this share's last band (<= band)
            const uint32_t full = last_rows == band ? mine : mine - 1u;
            float* dst = out_rgb + (size_t)k * band * W * 3u;
            cudaError_t e = cudaSuccess;
            if (full) e = cudaMemcpy2DAsync(dst, (size_t)n * band_bytes, sh.d_buf, band_bytes, band_bytes, full, cudaMemcpyDeviceToHost, sh.stream);
            if (e == cudaSuccess && full < mine)
                e = cudaMemcpyAsync(dst + (size_t)full * n * band * W * 3u, sh.d_buf + (size_t)full * band * W * 3u,
                                    (size_t)last_rows * W * 3u * sizeof(float), cudaMemcpyDeviceToHost, sh.stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(sh.stream);
            if (e != cudaSuccess) { sh.err = cudaGetErrorString(e); return cuda_rc(e); }
            rtw_get_stats(sh.scene, &sh.stats);
            return (int)RTW_OK;
        });
        if (rc) return rc;
    }
    ms->have_stats = true;
    return RTW_OK;
}

int rtw_render_frames_multi(rtw_scene_multi* ms, const rtw_camera* cam, const rtw_render_opts* opts, uint32_t first_frame,
                            uint32_t n_frames, float* out_rgb) {
    if (!ms || !cam || !opts || !out_rgb) return fail(RTW_E_INVALID, "null argument");
    DevGuard guard;
    const int n = (int)ms->shares.size();
    const uint32_t rows = opts->row_count ? opts->row_count : cam->height;
    const size_t n_values = (size_t)rows * cam->width * 3u;
    ms->combine_ms = 0.f; ms->have_stats = false;
    for (auto& sh : ms->shares) std::memset(&sh.stats, 0, sizeof sh.stats);
    int rc = for_each_share(ms, [&](int k) {                 // render_multi (viewport.rs:256-267): only `frame` changes
        Share& sh = ms->shares[k];
        if ((uint32_t)k >= n_frames) return (int)RTW_OK;
        int r = ensure_buf(sh, n_values);
        if (r) return r;
        for (uint32_t f = (uint32_t)k; f < n_frames; f += (uint32_t)n) {
            float* d = sh.d_buf;
            rtw_render_opts o = *opts;
            o.frame = first_frame + f;
            if ((r = rtw_render_device(sh.scene, cam, &o, d, sh.stream))) return r;
            cudaError_t e = cudaMemcpyAsync(out_rgb + (size_t)f * n_values, d, n_values * sizeof(float), cudaMemcpyDeviceToHost, sh.stream);
            if (e != cudaSuccess) { sh.err = cudaGetErrorString(e); return cuda_rc(e); }
            rtw_stats st; std::memset(&st, 0, sizeof st);
            rtw_get_stats(sh.scene, &st);
            sh.stats.paths += st.paths; sh.stats.rays += st.rays; sh.stats.nan_paths += st.nan_paths;
            sh.stats.gpu_launches += st.gpu_launches; sh.stats.kernel_ms += st.kernel_ms; sh.stats.total_ms += st.total_ms;
            sh.stats.iterations += st.iterations;
        }
        cudaError_t e = cudaStreamSynchronize(sh.stream);
        if (e != cudaSuccess) { sh.err = cudaGetErrorString(e); return cuda_rc(e); }
        return (int)RTW_OK;
    });
    if (rc) return rc;
    ms->have_stats = true;
    return RTW_OK;
}

int rtw_multi_get_stats(rtw_scene_multi* ms, rtw_stats* total, rtw_stats* per_device, float* combine_ms) {
    if (!ms || !total) return fail(RTW_E_INVALID, "null argument");
    std::memset(total, 0, sizeof *total);
    if (combine_ms) *combine_ms = ms->have_stats ? ms->combine_ms : 0.f;
    if (!ms->have_stats) return RTW_OK;
    int n_timed = 0;
    for (size_t k = 0; k < ms->shares.size(); ++k) {
        const rtw_stats& s = ms->shares[k].stats;
        if (per_device) per_device[k] = s;
        total->paths += s.paths; total->rays += s.rays; total->nan_paths += s.nan_paths; total->gpu_launches += s.gpu_launches;
        total->nodes += s.nodes; total->sphere_tests += s.sphere_tests; total->quad_tests += s.quad_tests;
        total->instance_tests += s.instance_tests; total->hits += s.hits; total->iterations += s.iterations;
        total->kernel_ms = std::max(total->kernel_ms, s.kernel_ms); total->total_ms = std::max(total->total_ms, s.total_ms);
        if (s.extend_ms > 0.f) { total->extend_ms += s.extend_ms; total->shade_ms += s.shade_ms; ++n_timed; }
    }
    if (n_timed) { total->extend_ms /= (float)n_timed; total->shade_ms /= (float)n_timed; }
    if (ms->combine_ms > 0.f) total->gpu_launches += 1;
    return RTW_OK;
}

}  // extern "C"
