// rtw_api.cu — the extern "C" boundary declared in include/rtw.h: scene upload, render calls,
// resolve/quantise kernels, stats. The only product entry points; nothing here (or anywhere in
// this library) can compute a pixel without a CUDA device.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>
#include "rtw_host.h"
#include "rtw_kernels.h"
#include "rtw_lbvh_driver.h"
#include "rtw_post.h"

using namespace rtw;

namespace {

thread_local std::string g_err;
// Entry points that have to make a particular device current put the caller's device back when they return.
struct DeviceGuard {
    int prev = -1;
    DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; } }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
int fail(int code, const std::string& msg) { g_err = msg; return code; }
int cuda_fail(cudaError_t e, const char* what) {
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? RTW_E_NOMEM : RTW_E_CUDA;
}
#define CK(call)                                              \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

// Device and pinned memory of destroyed scenes is kept for the next one (per device, up to RTW_MEM_CACHE_MB,
// default 4096). A caller that builds a scene per frame (render_multi clones the scene per frame,
// viewport.rs:263) would otherwise pay cudaMalloc / cudaFree of the 0.5 GB path pool every frame, and those
// calls serialise on a driver lock with anything else that touches the device: with an nvidia-smi poll running
// beside the render, one rtw_scene_destroy in four stalled for 0.1-0.7 s (profiles/r1l_e2e_trace.txt).
// rtw_trim_memory() gives the cached blocks back.
struct MemCache {
    struct Block { void* p; size_t bytes; int device; bool pinned; };
    struct Tex { cudaArray_t arr; cudaTextureObject_t obj; uint32_t w, h; int device; size_t bytes; };   // float4 texels, point filter
    std::mutex m;
    std::vector<Block> idle;
    std::vector<Tex> idle_tex;
    std::unordered_map<void*, Block> live;
    size_t held = 0;
};
MemCache& mem_cache() { static MemCache* c = new MemCache; return *c; }      // outlives the CUDA context teardown at exit
size_t cache_limit() {
    const char* e = std::getenv("RTW_MEM_CACHE_MB");
    return (size_t)(e ? std::strtoull(e, nullptr, 10) : 4096ull) << 20;
}
size_t round_bytes(size_t b) {
    if (b < 512u) b = 512u;
    return b <= (1u << 20) ? (b + 511u) & ~(size_t)511u : (b + (2u << 20) - 1u) & ~(size_t)((2u << 20) - 1u);
}
void raw_free(const MemCache::Block& b) {
    int cur = 0; cudaGetDevice(&cur);
    if (cur != b.device) cudaSetDevice(b.device);
    if (b.pinned) cudaFreeHost(b.p); else cudaFree(b.p);
    if (cur != b.device) cudaSetDevice(cur);
}
void raw_free(const MemCache::Tex& t) {
    int cur = 0; cudaGetDevice(&cur);
    if (cur != t.device) cudaSetDevice(t.device);
    cudaDestroyTextureObject(t.obj); cudaFreeArray(t.arr);
    if (cur != t.device) cudaSetDevice(cur);
}
void trim_locked(MemCache& c, int device /* -1: all */) {
    for (size_t i = 0; i < c.idle_tex.size();) {
        if (device < 0 || c.idle_tex[i].device == device) { raw_free(c.idle_tex[i]); c.held -= c.idle_tex[i].bytes; c.idle_tex[i] = c.idle_tex.back(); c.idle_tex.pop_back(); }
        else ++i;
    }
    for (size_t i = 0; i < c.idle.size();) {
        if (device < 0 || c.idle[i].device == device) { raw_free(c.idle[i]); c.held -= c.idle[i].bytes; c.idle[i] = c.idle.back(); c.idle.pop_back(); }
        else ++i;
    }
}
cudaError_t cached_alloc(void** out, size_t bytes, bool pinned = false) {
    MemCache& c = mem_cache();
    int dev = 0; cudaGetDevice(&dev);
    const size_t want = round_bytes(bytes);
    std::lock_guard<std::mutex> lk(c.m);
    size_t best = c.idle.size();
    for (size_t i = 0; i < c.idle.size(); ++i) {
        const MemCache::Block& b = c.idle[i];
        if (b.device != dev || b.pinned != pinned || b.bytes < want || b.bytes > 2u * want) continue;
        if (best == c.idle.size() || b.bytes < c.idle[best].bytes) best = i;
    }
    if (best != c.idle.size()) {
        const MemCache::Block b = c.idle[best];
        c.idle[best] = c.idle.back(); c.idle.pop_back();
        c.held -= b.bytes;
        c.live[b.p] = b;
        *out = b.p;
        return cudaSuccess;
    }
    void* p = nullptr;
    cudaError_t e = pinned ? cudaMallocHost(&p, want) : cudaMalloc(&p, want);
    if (e == cudaErrorMemoryAllocation) {        // give the cached blocks back and try once more
        cudaGetLastError();
        trim_locked(c, pinned ? -1 : dev);
        e = pinned ? cudaMallocHost(&p, want) : cudaMalloc(&p, want);
    }
    if (e != cudaSuccess) return e;
    c.live[p] = MemCache::Block{p, want, dev, pinned};
    *out = p;
    return cudaSuccess;
}
// The caller has made sure that nothing on the device still uses `p`.
void cached_free(void* p) {
    if (!p) return;
    MemCache& c = mem_cache();
    std::lock_guard<std::mutex> lk(c.m);
    auto it = c.live.find(p);
    if (it == c.live.end()) { cudaFree(p); return; }
    const MemCache::Block b = it->second;
    c.live.erase(it);
    const size_t limit = cache_limit();
    if (b.bytes > limit) { raw_free(b); return; }
    c.idle.push_back(b); c.held += b.bytes;
    while (c.held > limit && !c.idle.empty()) {  // oldest first
        raw_free(c.idle.front()); c.held -= c.idle.front().bytes; c.idle.erase(c.idle.begin());
    }
}
// A width x height float4 array with its point-sampling texture object, from the cache if one of that shape is idle.
cudaError_t cached_texture(uint32_t w, uint32_t h, MemCache::Tex* out) {
    MemCache& c = mem_cache();
    int dev = 0; cudaGetDevice(&dev);
    {
        std::lock_guard<std::mutex> lk(c.m);
        for (size_t i = 0; i < c.idle_tex.size(); ++i)
            if (c.idle_tex[i].device == dev && c.idle_tex[i].w == w && c.idle_tex[i].h == h) {
                *out = c.idle_tex[i];
                c.held -= out->bytes;
                c.idle_tex[i] = c.idle_tex.back(); c.idle_tex.pop_back();
                return cudaSuccess;
            }
    }
    MemCache::Tex t{}; t.w = w; t.h = h; t.device = dev; t.bytes = (size_t)w * h * sizeof(float4);
    cudaChannelFormatDesc cd = cudaCreateChannelDesc<float4>();
    cudaError_t e = cudaMallocArray(&t.arr, &cd, w, h);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        { std::lock_guard<std::mutex> lk(c.m); trim_locked(c, dev); }
        e = cudaMallocArray(&t.arr, &cd, w, h);
    }
    if (e != cudaSuccess) return e;
    cudaResourceDesc rd; std::memset(&rd, 0, sizeof rd);
    rd.resType = cudaResourceTypeArray; rd.res.array.array = t.arr;
    cudaTextureDesc td; std::memset(&td, 0, sizeof td);
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    if ((e = cudaCreateTextureObject(&t.obj, &rd, &td, nullptr)) != cudaSuccess) { cudaFreeArray(t.arr); return e; }
    *out = t;
    return cudaSuccess;
}
void cached_texture_release(const MemCache::Tex& t) {
    MemCache& c = mem_cache();
    std::lock_guard<std::mutex> lk(c.m);
    if (c.held + t.bytes > cache_limit() || c.idle_tex.size() >= 256u) { raw_free(t); return; }
    c.idle_tex.push_back(t); c.held += t.bytes;
}
template <class T> cudaError_t cached_alloc_t(T** out, size_t bytes, bool pinned = false) {
    void* p = nullptr;
    cudaError_t e = cached_alloc(&p, bytes, pinned);
    *out = static_cast<T*>(p);
    return e;
}

// viewport.rs:207-213 + :301 — mean over the samples, then powf(1/gamma) per channel, no clamp.
__global__ void resolve_kernel(float* __restrict__ rgb, unsigned long long n_values, float total, float inv_gamma) {
    unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (; k < n_values; k += stride) rgb[k] = powf(rgb[k] / total, inv_gamma);
}

// write_img_f32 (Rust/src/write_img.rs:6-19): (x*255).clamp(0,255).round() as u8; NaN -> 0.
__global__ void quantize_kernel(const float* __restrict__ rgb, uint8_t* __restrict__ out, unsigned long long n_values) {
    unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (; k < n_values; k += stride) {
        float x = rgb[k] * 255.0f;
        uint8_t q = 0;
        if (x == x) {
            x = x < 0.0f ? 0.0f : (x > 255.0f ? 255.0f : x);
            q = (uint8_t)roundf(x);
        }
        out[k] = q;
    }
}

// Seeds the wavefront counters on the device. (The first version copied them from slot 0 of the pinned ring the
// host polls; a second render enqueued right behind the first could have that slot overwritten by the first
// render's last, still pending, D2H copy before its own H2D read it: a frame that found the sample supply
// "exhausted" and came back black.)
__global__ void wf_init_kernel(WfCounters* wc, unsigned long long total_paths) {
    wc->head = 0u; wc->live_mark = 0u; wc->next_path = 0ull; wc->total_paths = total_paths; wc->iterations = 0ull;
}

// ---- micro-benchmarks (rtw_microbench): denominators for the FP32 and L2 rooflines
__global__ void bench_fp32_kernel(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f,
          x6 = x0 + 6.f, x7 = x0 + 7.f;
    for (int i = 0; i < iters; ++i) {   // 8 independent FFMA chains per thread
        x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
        x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 123.456f) out[0] = s;
}
__global__ void bench_read_kernel(const uint4* __restrict__ buf, unsigned long long n_vec, int passes, unsigned* out) {
    unsigned acc = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (int p = 0; p < passes; ++p) {
        // rotate the block -> slice mapping every pass so that no SM re-reads what its L1 holds
        const unsigned blk = (blockIdx.x + (unsigned)p * 61u) % gridDim.x;
        for (unsigned long long k = (unsigned long long)blk * blockDim.x + threadIdx.x; k < n_vec; k += stride) {
            uint4 v = __ldg(&buf[k]);
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

int grid_for(unsigned long long n, int block) {
    unsigned long long g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > 148ull * 16) g = 148ull * 16;
    return (int)g;
}

}  // namespace

namespace rtw { void set_last_error(const std::string& msg) { g_err = msg; } }   // rtw_multi.cu reports through the same string

struct rtw_scene {
    int device = 0;
    int sm_count = 0;
    // Traversal stack depth, per kernel family: [0] what the fast / count kernels walk (baked top levels: usually no
    // instance levels at all), [1] the reference's structure (strict kernels, RTW_NO_BAKE, rebuilt LBVH levels)
    uint32_t stack_words[2] = {0, 0};
    size_t smem_bytes[2] = {0, 0};
    int occ_mega[3][2][2] = {};           // [variant][stack kind][plain / light-biased instance]
    int cur_stack = 1;                    // stack kind of the current render
    Scene d{};                            // device pointers
    Scene d_cur{};                        // what the current render walks: d, or d with top = top_inside
    Level tops[4] = {};                   // [0] the reference's structure, [1] without enclosing volumes, [2] / [3] the same with baked instances (rtw_host.h)
    bool has_baked = false;
    uint32_t n_walked_instances_baked = 0;
    // optional on-device LBVH rebuild (rtw_lbvh.h): the leaves of every level, the node arrays with
    // room for the rebuilt levels behind the SAH nodes, and the SAH levels to go back to
    std::vector<LbvhLeaf> lb_leaves_h;
    std::vector<LbvhLevelDesc> lb_levels;
    std::vector<Level> sah_inst_levels;       // per instance
    Level sah_tops[4] = {};
    uint32_t sah_top_depth = 0, sah_inner_depth = 0, n_sah_nodes = 0;
    uint32_t sah_fast_top_depth = 0, sah_fast_inner_depth = 0;   // the same for what the fast kernels walk (HostScene)
    bool fast_levels = false, any_levels = false;
    std::vector<uint8_t> inst_baked;
    uint32_t sah_stack_words[2] = {0, 0};
    LbvhLeaf* lb_leaves_d = nullptr;
    Node* lb_nodes = nullptr; Node16* lb_nodes16 = nullptr;
    LbvhLevelOut* lb_out_d = nullptr;
    // the levels of a rebuild are independent: each gets its own scratch and one of up to kLbLanes streams, forked from and
    // joined to the scene's stream with events, so that their latency-bound fit passes overlap (C5: five levels)
    static const int kLbLanes = 8;
    std::vector<LbvhScratch> lb_scratch;
    cudaStream_t lb_stream[kLbLanes] = {};
    cudaEvent_t lb_fork = nullptr, lb_join[kLbLanes] = {};
    std::vector<HostScene::Enclosing> enclosing;
    std::vector<void*> allocs;
    std::vector<MemCache::Tex> texs;
    rtw_scene_info info{};
    cudaStream_t stream = nullptr;        // used by rtw_render / rtw_render_frames
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    Counters* d_ctr = nullptr;
    uint32_t* d_work = nullptr;
    Counters* h_ctr = nullptr;            // pinned
    float* d_img = nullptr; size_t d_img_floats = 0;
    bool have_stats = false;
    uint64_t launches = 0;
    // wavefront pipeline state (allocated on first use)
    WfPool pool{};
    uint32_t n_walked_instances = 0;      // instance leaves of the top level that are walked as a level (not one-sphere records)
    uint32_t cur_walked_instances = 0;    // ... of the top level the current render walks
    uint32_t pool_allocated = 0;          // slots the pool arrays hold; pool.capacity = slots the current render uses
    WfCounters* d_wc = nullptr;
    WfCounters* h_wc = nullptr;           // pinned, 4 slots
    cudaEvent_t wf_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    int occ_wf[3][2] = {};
    bool last_wavefront = false;
    uint64_t wf_paths = 0, wf_iterations = 0;
    static const int kTimed = 32;             // sampled (extend, shade) launches timed with events per render
    cudaEvent_t wf_t[kTimed][3] = {};
    int wf_timed = 0;                         // samples recorded this render
    int wf_harvested = 0;                     // ... of which already folded into the sums below
    double wf_extend_sum = 0.0, wf_shade_sum = 0.0;

    ~rtw_scene() {
        DeviceGuard guard;
        cudaSetDevice(device);
        cudaDeviceSynchronize();          // the blocks below go back to the cache: nothing may still be using them
        for (const auto& t : texs) cached_texture_release(t);
        for (auto p : allocs) cached_free(p);
        cached_free(d_img);
        cached_free(lb_leaves_d);
        cached_free(lb_nodes);
        cached_free(lb_nodes16);
        cached_free(lb_out_d);
        cached_free(d_wc);
        cached_free(h_wc);
        for (auto e : wf_ev) if (e) cudaEventDestroy(e);
        for (auto& t : wf_t) for (auto e : t) if (e) cudaEventDestroy(e);
        cached_free(d_ctr);
        cached_free(d_work);
        cached_free(h_ctr);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (ev2) cudaEventDestroy(ev2);
        for (auto st : lb_stream) if (st) cudaStreamDestroy(st);
        for (auto e : lb_join) if (e) cudaEventDestroy(e);
        if (lb_fork) cudaEventDestroy(lb_fork);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

template <class T>
int upload(rtw_scene* s, const std::vector<T>& v, const T** out) {
    size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    void* p = nullptr;
    CK(cached_alloc(&p, bytes));
    s->allocs.push_back(p);
    s->info.device_bytes += bytes;
    if (!v.empty()) { CK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice)); s->info.h2d_bytes += v.size() * sizeof(T); }
    *out = static_cast<const T*>(p);
    return RTW_OK;
}

int upload_scene(rtw_scene* s, HostScene& hs) {
    int rc;
    // image textures -> cudaArray + texture object (point filter, unnormalised coords)
    for (size_t i = 0; i < hs.textures.size(); ++i) {
        Texture& t = hs.textures[i];
        t.obj = 0; t.texels = nullptr;
        if (t.n_texels == 1u) continue;
        cudaDeviceProp* unused = nullptr; (void)unused;
        int max_w = 0, max_h = 0;
        CK(cudaDeviceGetAttribute(&max_w, cudaDevAttrMaxTexture2DWidth, s->device));
        CK(cudaDeviceGetAttribute(&max_h, cudaDevAttrMaxTexture2DHeight, s->device));
        if ((int)t.width > max_w || (int)t.height > max_h) return fail(RTW_E_UNSUPPORTED, "texture exceeds the 2D texture limits");
        MemCache::Tex ct{};
        CK(cached_texture(t.width, t.height, &ct));
        s->texs.push_back(ct);
        s->info.device_bytes += (uint64_t)t.n_texels * sizeof(float4);
        s->info.h2d_bytes += (uint64_t)t.n_texels * sizeof(float4);
        CK(cudaMemcpy2DToArray(ct.arr, 0, 0, hs.texels[i].data(), t.width * sizeof(float4), t.width * sizeof(float4),
                               t.height, cudaMemcpyHostToDevice));
        t.obj = ct.obj;
    }
    Scene& d = s->d;
    if ((rc = upload(s, hs.nodes, &d.nodes))) return rc;
    if ((rc = upload(s, hs.nodes16, &d.nodes16))) return rc;
    if ((rc = upload(s, hs.spheres, &d.spheres))) return rc;
    if ((rc = upload(s, hs.sphere_vel, &d.sphere_vel))) return rc;
    if ((rc = upload(s, hs.sphere_col, &d.sphere_col))) return rc;
    if ((rc = upload(s, hs.sphere_ids, &d.sphere_ids))) return rc;
    if ((rc = upload(s, hs.sphere_box, &d.sphere_box))) return rc;
    if ((rc = upload(s, hs.quads, &d.quads))) return rc;
    if ((rc = upload(s, hs.quad_ids, &d.quad_ids))) return rc;
    if ((rc = upload(s, hs.instances, &d.instances))) return rc;
    if ((rc = upload(s, hs.one_sphere, &d.one_sphere))) return rc;
    if ((rc = upload(s, hs.moving, &d.moving))) return rc;
    if ((rc = upload(s, hs.materials, &d.materials))) return rc;
    if ((rc = upload(s, hs.textures, &d.textures))) return rc;
    if ((rc = upload(s, hs.noises, &d.noises))) return rc;
    if ((rc = upload(s, hs.baked_sphere_src, &d.baked_sphere_src))) return rc;
    if ((rc = upload(s, hs.baked_quad_src, &d.baked_quad_src))) return rc;
    d.n_spheres_orig = hs.n_spheres_orig; d.n_quads_orig = hs.n_quads_orig;
    if ((rc = upload(s, hs.lights, &d.lights))) return rc;
    d.n_lights = hs.n_lights;
    d.top = hs.top;
    s->tops[0] = hs.top; s->tops[1] = hs.top_inside; s->tops[2] = hs.top_baked; s->tops[3] = hs.top_baked_inside;
    s->has_baked = hs.has_baked; s->n_walked_instances_baked = hs.n_walked_instances_baked;
    s->enclosing = hs.enclosing;
    s->lb_leaves_h = hs.lbvh_leaves; s->lb_levels = hs.lbvh_levels;
    s->n_walked_instances = hs.n_walked_instances;
    for (int k = 0; k < 4; ++k) s->sah_tops[k] = s->tops[k];
    s->sah_inst_levels.resize(hs.instances.size());
    for (size_t k = 0; k < hs.instances.size(); ++k) s->sah_inst_levels[k] = hs.instances[k].level;
    s->sah_top_depth = hs.top_depth; s->sah_inner_depth = hs.inner_depth; s->n_sah_nodes = (uint32_t)hs.nodes.size();
    s->sah_fast_top_depth = hs.fast_top_depth; s->sah_fast_inner_depth = hs.fast_inner_depth;
    s->fast_levels = hs.fast_levels; s->any_levels = hs.any_levels; s->inst_baked = hs.inst_baked;
    std::memcpy(d.background, hs.background, sizeof d.background);
    d.n_noises = (uint32_t)hs.noises.size();
    return RTW_OK;
}

uint32_t env_u32(const char* name, uint32_t dflt) {
    const char* e = std::getenv(name);
    return e ? (uint32_t)std::strtoul(e, nullptr, 10) : dflt;
}

// Work decomposition (see MegaLaunch): enough items to balance, chunks short enough that the tail
// is small, long enough that the three REDs per lane per item are noise.
// Tuning knob of the extend phase (see rtw_kernels.cu); RTW_REFILL_BELOW in the environment overrides.
uint32_t g_refill_below = 0xFFFFFFFFu;
uint32_t refill_below_default() {
    if (g_refill_below == 0xFFFFFFFFu) {
        const char* e = std::getenv("RTW_REFILL_BELOW");
        g_refill_below = e ? (uint32_t)std::atoi(e) : 1u;
    }
    return g_refill_below;
}

MegaLaunch plan_mega(const rtw_scene* s, const RenderParams& P, int variant) {
    MegaLaunch L;
    L.block = RTW_MEGA_BLOCK;
    int occ = s->occ_mega[variant][s->cur_stack][light_integrator(P.integrator) ? 1 : 0];
    if (occ < 1) occ = 1;
    L.grid = s->sm_count * occ;
    L.stack_words = s->stack_words[s->cur_stack];
    L.smem_bytes = s->smem_bytes[s->cur_stack];
    L.refill_below = refill_below_default();
    L.tiles_x = (P.cam.width + 7u) / 8u;
    L.tiles_y = (P.row_count + 3u) / 4u;
    const uint64_t tiles = (uint64_t)L.tiles_x * L.tiles_y;
    const uint64_t resident_warps = (uint64_t)L.grid * (RTW_MEGA_BLOCK / 32);
    uint64_t want_chunks = (16 * resident_warps + tiles - 1) / tiles;       // >= 16 items per resident warp
    if (want_chunks < 1) want_chunks = 1;
    if (want_chunks > P.sample_count) want_chunks = P.sample_count;
    uint32_t chunk = (uint32_t)((P.sample_count + want_chunks - 1) / want_chunks);
    // at most 128 samples of a tile per item: C3 at 1 000 spp 5 240 -> 5 606 Mpaths/s against the 256 of before (its tiles
    // differ in cost and four items per tile left the frame's tail unbalanced; 64: 5 554, 32: 5 422, 16: 5 171; the
    // other configs do not move: profiles/r2u_ab_mega_chunk.txt)
    static const uint32_t chunk_max = std::max(1u, env_u32("RTW_MEGA_CHUNK_MAX", 128u));
    if (chunk > chunk_max) chunk = chunk_max;
    if (chunk < 1u) chunk = 1u;
    L.chunk = chunk;
    L.n_chunks = (P.sample_count + chunk - 1u) / chunk;
    L.n_items = (uint32_t)std::min<uint64_t>(tiles * L.n_chunks, 0xFFFFFFF0ull);
    const uint64_t max_useful_blocks = ((uint64_t)L.n_items + (RTW_MEGA_BLOCK / 32) - 1) / (RTW_MEGA_BLOCK / 32);
    if ((uint64_t)L.grid > max_useful_blocks) L.grid = (int)std::max<uint64_t>(1, max_useful_blocks);
    return L;
}


// Which pipeline a render takes when the caller names none (profiles/r2r_crossover.txt: both pipelines forced on every
// config from 0.1 M to 200 M paths). The wavefront pipeline buys coherence — every warp runs one kind of step with most
// of its lanes — and pays for it with a path record that crosses HBM twice per ray (56 B), two launches per bounce and a
// ramp and a drain of about 2 x depth iterations over a mostly dead pool: 4-20 ms per frame that the megakernel, one
// launch with the paths in registers, does not have. So:
//  * a render below 2^24 paths takes the megakernel (C5 at 10 M paths: 13.8 ms against 21.5; C2 at 0.8 M: 0.58 against
//    4.5 ms; the wavefront pipeline leads from 23 M paths on C4 and 41 M on C5, and on C2 only ties at 200 M);
//  * a scene whose top level is a handful of primitives with nothing below (C1: five spheres, C3: four) takes it at any
//    size: one node fetch and one or two leaf tests per ray keep the megakernel's lanes in step by themselves (C1 at
//    23 M paths 4 681 against 2 198 Mpaths/s, C3 at 92 M 4 741 against 2 647, C3 at its full 360 M 4 807 against 3 556);
//  * everything else takes the wavefront pipeline (C5 at 164 M paths 1 576 against 860 Mpaths/s, at 6.4 G 1 990).
// RTW_WF_MIN_PATHS / RTW_MEGA_FLAT_LEAVES move the two thresholds (tuning only).
bool wavefront_pays(const rtw_scene* s, uint64_t n_paths) {
    static const uint64_t min_paths = env_u32("RTW_WF_MIN_PATHS", 1u << 24);
    static const uint32_t flat_leaves = env_u32("RTW_MEGA_FLAT_LEAVES", 8u);
    if (n_paths < min_paths) return false;
    const bool flat = s->info.n_tlas_leaves <= flat_leaves && s->info.n_blas_nodes == 0u && s->info.n_baked_instances == 0u;
    return !flat;
}

// The occupancies depend on the dynamic shared memory (the stack columns). Largest size first: the attribute
// cudaFuncAttributeMaxDynamicSharedMemorySize the queries set must end up covering every launch.
int refresh_mega_occupancy(rtw_scene* s) {
    for (int k = 0; k < 2; ++k) {          // stack_words[0] <= stack_words[1]
        CK(mega_occupancy_fast(s->occ_mega[0][k], s->smem_bytes[k]));
        CK(mega_occupancy_strict(s->occ_mega[1][k], s->smem_bytes[k]));
        CK(mega_occupancy_count(s->occ_mega[2][k], s->smem_bytes[k]));
    }
    return RTW_OK;
}
int refresh_wf_occupancy(rtw_scene* s) {
    for (int k = 0; k < 2; ++k) {
        CK(wf_extend_occupancy_fast(&s->occ_wf[0][k], s->smem_bytes[k]));
        CK(wf_extend_occupancy_strict(&s->occ_wf[1][k], s->smem_bytes[k]));
        CK(wf_extend_occupancy_count(&s->occ_wf[2][k], s->smem_bytes[k]));
    }
    return RTW_OK;
}

template <class T>
int pool_alloc(rtw_scene* s, T** p, size_t n) {
    void* q = nullptr;
    CK(cached_alloc(&q, n * sizeof(T)));
    s->allocs.push_back(q);
    s->info.device_bytes += n * sizeof(T);
    *p = static_cast<T*>(q);
    return RTW_OK;
}

// Pool of path slots for one render of `n_paths` paths. The bigger the pool the fewer, longer launches (each pays a
// tail and the launch gaps: cost ~ rays / pool) but the longer the frame drains at the end (cost ~ pool), so the best
// size grows with the square root of the frame's RAYS: 128 * sqrt(rays) fits what was measured - C5 at 1 250 spp
// (2.2 G rays) 6 Mi slots best, 9 Mi -0.5 %, 12 Mi -4 %; C5 at 10 000 spp 12 Mi > 8 Mi > 6 Mi; C4 at 625 spp (1.5 G rays,
// 6.9 rays per path) 3.5-5 Mi +5.4 % over the 1.8 Mi of the old paths / 128 rule; C2 6 Mi +2.3 % over 3.2 Mi
// (profiles/r2f_sweep.txt, r2h_ab_*). Rays per path are the scene's own from its previous render (3 before there is one).
// Between 1 Mi and 16 Mi slots, never more than there are paths. RTW_WF_POOL overrides. Arrays only ever grow.
int ensure_pool(rtw_scene* s, unsigned long long n_paths) {
    unsigned long long want = env_u32("RTW_WF_POOL", 0u);
    if (want == 0ull) {
        double rpp = 3.0;
        if (s->have_stats && s->ev2 && cudaEventQuery(s->ev2) == cudaSuccess && s->last_wavefront && s->wf_paths > 0 && s->h_ctr->rays > 0)
            rpp = std::min(64.0, std::max(1.0, (double)s->h_ctr->rays / (double)s->wf_paths));
        cudaGetLastError();
        want = (unsigned long long)(128.0 * std::sqrt((double)n_paths * rpp));
        want = std::min<unsigned long long>(16ull << 20, std::max<unsigned long long>(1ull << 20, want));
        want = std::min(want, std::max<unsigned long long>(16384ull, n_paths));      // never more slots than paths
    }
    const uint32_t cap = (uint32_t)((std::max<unsigned long long>(want, 4096ull) + 255ull) & ~255ull);   // whole shade blocks
    int rc;
    if (s->pool_allocated < cap) {
        WfQueue& q = s->pool.q;     // (the smaller arrays stay on the scene's allocation list until it is destroyed)
        if ((rc = pool_alloc(s, &q.o, cap)) || (rc = pool_alloc(s, &q.d, cap)) || (rc = pool_alloc(s, &q.id, cap)) ||
            (rc = pool_alloc(s, &q.thr, cap))) return rc;
        if ((rc = pool_alloc(s, &s->pool.hit, cap))) return rc;
        if ((rc = pool_alloc(s, &s->pool.supply, (size_t)cap / 16u))) return rc;      // 2 words per warp of 32 slots
        if ((rc = pool_alloc(s, &s->pool.warp_live, (size_t)cap / 32u))) return rc;
        s->pool_allocated = cap;
    }
    if (!s->d_wc) {
        CK(cached_alloc_t(&s->d_wc, sizeof(WfCounters)));
        CK(cached_alloc_t(&s->h_wc, 4 * sizeof(WfCounters), true));
        for (auto& e : s->wf_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto& t : s->wf_t) for (auto& e : t) CK(cudaEventCreate(&e));
        int rc2 = refresh_wf_occupancy(s);
        if (rc2) return rc2;
    }
    s->pool.capacity = cap;
    return RTW_OK;
}

// The wavefront render: fill the pool, then extend + shade until every slot has run dry. The host
// stays two batches of iterations ahead of the device and reads the queue length of finished
// batches from pinned memory, so the call returns only when the frame is complete.
int enqueue_wavefront(rtw_scene* s, const RenderParams& P, int variant, float* d_out, cudaStream_t st) {
    int rc = ensure_pool(s, (unsigned long long)P.cam.width * P.row_count * P.sample_count);
    if (rc) return rc;
    WfLaunch L{};
    L.tiles_x = (P.cam.width + 7u) / 8u;
    L.tiles_y = (P.row_count + 3u) / 4u;
    L.stack_words = s->stack_words[s->cur_stack];
    L.smem_bytes = s->smem_bytes[s->cur_stack];
    L.fetch_min = std::min(32u, std::max(1u, env_u32("RTW_WF_FETCH_MIN", 16u)));
    L.chunk = std::max(32u, env_u32("RTW_WF_CHUNK", 64u));
    L.supply_chunk = std::max(32u, env_u32("RTW_WF_SUPPLY_CHUNK", 256u)) & ~31u;
    L.sample_major = env_u32("RTW_WF_SAMPLE_MAJOR", 0u) ? 1u : 0u;
    L.index32 = ((unsigned long long)L.tiles_x * L.tiles_y * P.sample_count < (1ull << 32)) ? 1u : 0u;
    CK(cudaMemsetAsync(s->pool.supply, 0, (size_t)s->pool.capacity / 16u * sizeof(unsigned long long), st));
    L.extend_grid = s->sm_count * std::max(1, std::min(s->occ_wf[variant][s->cur_stack], (int)env_u32("RTW_WF_MAX_BLOCKS", 64u)));
    L.shade_grid = (int)(s->pool.capacity / RTW_WF_SHADE_BLOCK);
    WfCounters init{};
    init.total_paths = 32ull * L.tiles_x * L.tiles_y * P.sample_count;
    L.total_paths = init.total_paths;
    {
        const bool events_rare = (unsigned long long)s->cur_walked_instances * 16ull <= (unsigned long long)s->info.n_tlas_leaves;
        L.prim_w = std::max(1u, env_u32("RTW_WF_PRIM_W", events_rare ? 8u : 4u));
        L.event_w = std::max(1u, env_u32("RTW_WF_EVENT_W", events_rare ? 8u : 4u));   // (16 before one-sphere instances joined the event steps)
        const uint32_t w = std::max(L.prim_w, L.event_w);
        L.node_sure = (32u * w + (4u + w) - 1u) / (4u + w);
    }
    wf_init_kernel<<<1, 1, 0, st>>>(s->d_wc, init.total_paths);
    CK(cudaGetLastError());
    auto extend = [&]() {
        return variant == 1   ? launch_wf_extend_strict(s->d_cur, P, s->pool, s->d_wc, s->d_ctr, L, st)
               : variant == 2 ? launch_wf_extend_count(s->d_cur, P, s->pool, s->d_wc, s->d_ctr, L, st)
                              : launch_wf_extend_fast(s->d_cur, P, s->pool, s->d_wc, s->d_ctr, L, st);
    };
    auto shade = [&](uint32_t mark, int init_pass) {
        return variant == 1   ? launch_wf_shade_strict(s->d_cur, P, s->pool, s->d_wc, s->d_ctr, d_out, mark, init_pass, L, st)
               : variant == 2 ? launch_wf_shade_count(s->d_cur, P, s->pool, s->d_wc, s->d_ctr, d_out, mark, init_pass, L, st)
                              : launch_wf_shade_fast(s->d_cur, P, s->pool, s->d_wc, s->d_ctr, d_out, mark, init_pass, L, st);
    };
    s->wf_timed = 0; s->wf_harvested = 0; s->wf_extend_sum = 0.0; s->wf_shade_sum = 0.0;
    CK(shade(0u, 1));                       // first fill
    s->launches += 1;
    // the host runs at most three batches ahead and a ring slot is reused after kTimed * 8 rounds: 3 * batch must stay below that
    const uint32_t batch = std::min(64u, std::max(1u, env_u32("RTW_WF_BATCH", 8u)));
    uint32_t it = 0;
    for (uint32_t k = 0;; ++k) {
        for (uint32_t b = 0; b < batch; ++b, ++it) {
            // every 8th round, from the 4th on, is bracketed with events. A ring slot about to be reused
            // was recorded kTimed*8 rounds ago (the host runs at most three batches ahead), so its
            // events are complete: its times go into the running sums => the mean covers the whole render.
            const bool timed = (it & 7u) == 3u;
            cudaEvent_t* te = s->wf_t[s->wf_timed % rtw_scene::kTimed];
            if (timed && s->wf_timed >= rtw_scene::kTimed) {
                float a = 0.f, b = 0.f;
                CK(cudaEventElapsedTime(&a, te[0], te[1]));
                CK(cudaEventElapsedTime(&b, te[1], te[2]));
                s->wf_extend_sum += a; s->wf_shade_sum += b; ++s->wf_harvested;
            }
            if (timed) CK(cudaEventRecord(te[0], st));
            CK(extend());
            if (timed) CK(cudaEventRecord(te[1], st));
            CK(shade(it + 1u, 0));
            if (timed) { CK(cudaEventRecord(te[2], st)); ++s->wf_timed; }
        }
        s->launches += 2ull * batch;
        CK(cudaMemcpyAsync(&s->h_wc[k & 3u], s->d_wc, sizeof(WfCounters), cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(s->wf_ev[k & 3u], st));
        if (k >= 2u) {
            const uint32_t j = (k - 2u) & 3u;
            CK(cudaEventSynchronize(s->wf_ev[j]));
            // the sample supply has run out: from here on slots only die (WfLaunch::drain)
            if (!L.drain && s->h_wc[j].next_path >= init.total_paths && !env_u32("RTW_WF_NO_DRAIN", 0u)) {
                L.drain = 1u;
                // fewer reservations of a mostly dead pool, but still >= 8 chunks per resident warp: the first drain
                // iterations are mostly live and need the balance (2048-slot chunks cost 6 % at 200 spp, profiles/r2b_ab_*)
                const uint32_t warps = (uint32_t)L.extend_grid * (RTW_WF_BLOCK / 32u);
                const uint32_t c = (s->pool.capacity / std::max(1u, warps * 8u)) & ~63u;
                L.chunk = std::max(L.chunk, std::min(1024u, c));
            }
            // the last shade of batch k-2 carried mark (k-1)*batch: no live slot after it => the frame is done
            if (s->h_wc[j].live_mark != (k - 1u) * batch) break;
        }
    }
    s->wf_iterations = it;
    s->wf_paths = (uint64_t)P.cam.width * P.row_count * P.sample_count;
    if (!first_hit_integrator(P.integrator) && P.depth < 1u) s->wf_paths = s->wf_paths;   // black frame, still counted
    return RTW_OK;
}

// Enqueue one render on `st` into device buffer d_out (row_count*W*3 floats).
int enqueue_render(rtw_scene* s, const rtw_camera* cam, const rtw_render_opts* o, float* d_out, cudaStream_t st) {
    RenderParams P; uint32_t total = 0; std::string err;
    int rc = make_render_params(cam, o, P, total, err);
    if (rc) return fail(rc, err);
    const int variant = (o->flags & RTW_FLAG_STRICT_FP) ? 1 : (o->flags & RTW_FLAG_COUNT_WORK) ? 2 : 0;
    // Enclosing volumes (rtw_host.h): when the camera (lens included) and every moving sphere over the
    // shutter interval stay inside each of them, no ray of this render starts outside => they cannot
    // return a hit (instance.rs:280-300) and the fast kernels walk the top level built without them.
    s->d_cur = s->d;
    s->d_cur.top = s->tops[0];
    // the fast kernels walk the top level with baked instances (rtw_host.h) unless RTW_NO_BAKE=1
    const int baked = (variant != 1 && s->has_baked && !env_u32("RTW_NO_BAKE", 0u)) ? 2 : 0;
    s->d_cur.top = s->tops[baked];
    s->cur_walked_instances = baked ? s->n_walked_instances_baked : s->n_walked_instances;
    s->cur_stack = (variant != 1 && (baked || !s->has_baked)) ? 0 : 1;
    if (variant != 1 && !s->enclosing.empty() && !env_u32("RTW_KEEP_ENCLOSING", 0u)) {
        const float lens = std::fabs(cam->lens_radius) *
                           (std::sqrt(cam->u[0] * cam->u[0] + cam->u[1] * cam->u[1] + cam->u[2] * cam->u[2]) +
                            std::sqrt(cam->v[0] * cam->v[0] + cam->v[1] * cam->v[1] + cam->v[2] * cam->v[2]));
        const float tmax = std::fabs(P.time0) + std::fabs(P.shutter_speed);
        bool inside = true;
        for (const auto& e : s->enclosing) {
            float d2 = 0.f;
            for (int a = 0; a < 3; ++a) { const float x = cam->origin[a] - e.c[a]; d2 += x * x; }
            if (!(std::sqrt(d2) + lens < e.r) || !(e.mov_far + e.mov_speed * tmax < e.r)) inside = false;
        }
        if (inside) s->d_cur.top = s->tops[baked + 1];
    }
    MegaLaunch L = plan_mega(s, P, variant);
    if ((uint64_t)L.tiles_x * L.tiles_y * L.n_chunks > 0xFFFFFFF0ull) return fail(RTW_E_UNSUPPORTED, "too many work items; shard the render");
    const size_t n_values = (size_t)P.row_count * cam->width * 3u;
    CK(cudaMemsetAsync(d_out, 0, n_values * sizeof(float), st));
    CK(cudaMemsetAsync(s->d_ctr, 0, sizeof(Counters), st));
    CK(cudaMemsetAsync(s->d_work, 0, sizeof(uint32_t), st));
    CK(cudaEventRecord(s->ev0, st));
    if ((o->flags & RTW_FLAG_WAVEFRONT) && (o->flags & RTW_FLAG_MEGAKERNEL)) return fail(RTW_E_INVALID, "both pipeline flags set");
    const uint64_t n_paths = (uint64_t)cam->width * P.row_count * P.sample_count;
    if ((o->flags & RTW_FLAG_WAVEFRONT) && P.depth > WF_MAX_DEPTH) return fail(RTW_E_UNSUPPORTED, "wavefront pipeline: depth > 4095 (use the megakernel)");
    // the light-biased integrators trace their shadow rays from inside the shading step: megakernel only
    if ((o->flags & RTW_FLAG_WAVEFRONT) && light_integrator(P.integrator)) return fail(RTW_E_UNSUPPORTED, "light-biased integrators run in the megakernel only");
    s->last_wavefront = (o->flags & RTW_FLAG_WAVEFRONT) ||
                        (!(o->flags & RTW_FLAG_MEGAKERNEL) && P.depth <= WF_MAX_DEPTH && !light_integrator(P.integrator) &&
                         wavefront_pays(s, n_paths));
    if (s->last_wavefront) {
        s->launches = 0;
        if ((rc = enqueue_wavefront(s, P, variant, d_out, st))) return rc;
    } else {
        CK(variant == 1   ? launch_mega_strict(s->d_cur, P, d_out, s->d_ctr, s->d_work, L, st)
           : variant == 2 ? launch_mega_count(s->d_cur, P, d_out, s->d_ctr, s->d_work, L, st)
                          : launch_mega_fast(s->d_cur, P, d_out, s->d_ctr, s->d_work, L, st));
        s->launches = 1;
    }
    CK(cudaEventRecord(s->ev1, st));
    if (o->output == RTW_OUT_RESOLVED) {
        resolve_kernel<<<grid_for(n_values, 256), 256, 0, st>>>(d_out, n_values, (float)total, 1.0f / o->gamma);
        CK(cudaGetLastError());
        s->launches += 1;
    }
    CK(cudaMemcpyAsync(s->h_ctr, s->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(s->ev2, st));
    s->have_stats = true;
    return RTW_OK;
}

int ensure_img(rtw_scene* s, size_t floats) {
    if (s->d_img_floats >= floats) return RTW_OK;
    if (s->d_img) { CK(cudaStreamSynchronize(s->stream)); cached_free(s->d_img); s->d_img = nullptr; s->d_img_floats = 0; }
    CK(cached_alloc_t(&s->d_img, floats * sizeof(float)));
    s->d_img_floats = floats;
    return RTW_OK;
}

}  // namespace

extern "C" {

int rtw_abi_version(void) { return RTW_ABI_VERSION; }
const char* rtw_last_error(void) { return g_err.c_str(); }

int rtw_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int rtw_camera_init(rtw_camera* cam, uint32_t width, float aspect_ratio, float vfov_deg, const float origin[3],
                    const float direction[3], const float vup[3], float lens_radius) {
    if (!cam || !origin || !direction || !vup || width == 0 || !(aspect_ratio > 0.0f)) return fail(RTW_E_INVALID, "bad camera arguments");
    camera_init(cam, width, aspect_ratio, vfov_deg, origin, direction, vup, lens_radius);
    return RTW_OK;
}
int rtw_camera_init_res(rtw_camera* cam, uint32_t width, uint32_t height, float vfov_deg, const float origin[3],
                        const float direction[3], const float vup[3], float lens_radius) {
    if (height == 0) return fail(RTW_E_INVALID, "bad camera arguments");
    return rtw_camera_init(cam, width, (float)width / (float)height, vfov_deg, origin, direction, vup, lens_radius);   // viewport.rs:402-428
}

void rtw_render_opts_default(rtw_render_opts* o) {
    if (!o) return;
    std::memset(o, 0, sizeof *o);
    o->samples = 1; o->depth = 10; o->gamma = 2.0f;
    o->integrator = RTW_INT_GRADIENT; o->mode = RTW_MODE_RANDOM;
    o->shutter_speed = 0.0f; o->fps = 30.0f; o->frame = 0;   // viewport.rs:394-398
    o->seed = 1; o->output = RTW_OUT_RESOLVED;
}

int rtw_scene_create(const rtw_scene_desc* desc, int device, rtw_scene** out) {
    if (!out) return fail(RTW_E_INVALID, "null out pointer");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(RTW_E_CUDA, std::string("no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count 0") +
                                    "); this library has no CPU path");
    }
    if (device < 0) { CK(cudaGetDevice(&device)); }
    if (device >= ndev) return fail(RTW_E_INVALID, "device ordinal out of range");
    HostScene hs; std::string err;
    int rc = build_host_scene(desc, hs, err);
    if (rc) return fail(rc, err);
    DeviceGuard guard;
    CK(cudaSetDevice(device));
    rtw_scene* s = new (std::nothrow) rtw_scene();
    if (!s) return fail(RTW_E_NOMEM, "out of host memory");
    s->device = device;
    s->info = hs.info;
    s->info.device_bytes = 0; s->info.h2d_bytes = 0;
    auto bail = [&](int code) { delete s; return code; };
    if ((rc = upload_scene(s, hs))) return bail(rc);
    cudaError_t ce;
    if ((ce = cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) return bail(cuda_fail(ce, "sm count"));
    s->stack_words[1] = hs.any_levels ? stack_words(hs.top_depth, hs.inner_depth) : stack_words_flat(hs.top_depth);
    s->stack_words[0] = hs.fast_levels ? stack_words(hs.fast_top_depth, hs.fast_inner_depth) : stack_words_flat(hs.fast_top_depth);
    if (s->stack_words[0] > s->stack_words[1]) s->stack_words[0] = s->stack_words[1];
    for (int k = 0; k < 2; ++k) { s->sah_stack_words[k] = s->stack_words[k]; s->smem_bytes[k] = (size_t)s->stack_words[k] * RTW_MEGA_BLOCK * sizeof(uint32_t); }
    if ((rc = refresh_mega_occupancy(s))) return bail(rc);
    if ((ce = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(cuda_fail(ce, "stream"));
    if ((ce = cudaEventCreate(&s->ev0)) != cudaSuccess || (ce = cudaEventCreate(&s->ev1)) != cudaSuccess ||
        (ce = cudaEventCreate(&s->ev2)) != cudaSuccess) return bail(cuda_fail(ce, "events"));
    if ((ce = cached_alloc_t(&s->d_ctr, sizeof(Counters))) != cudaSuccess) return bail(cuda_fail(ce, "counters"));
    if ((ce = cached_alloc_t(&s->d_work, sizeof(uint32_t))) != cudaSuccess) return bail(cuda_fail(ce, "work counter"));
    if ((ce = cached_alloc_t(&s->h_ctr, sizeof(Counters), true)) != cudaSuccess) return bail(cuda_fail(ce, "pinned counters"));
    std::memset(s->h_ctr, 0, sizeof(Counters));
    *out = s;
    return RTW_OK;
}

void rtw_scene_destroy(rtw_scene* scene) { delete scene; }

int rtw_render_device(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts, float* d_out_rgb, void* stream) {
    if (!scene || !cam || !opts || !d_out_rgb) return fail(RTW_E_INVALID, "null argument");
    DeviceGuard guard;
    CK(cudaSetDevice(scene->device));
    return enqueue_render(scene, cam, opts, d_out_rgb, static_cast<cudaStream_t>(stream));
}

int rtw_render(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts, float* out_rgb) {
    if (!scene || !cam || !opts || !out_rgb) return fail(RTW_E_INVALID, "null argument");
    DeviceGuard guard;
    CK(cudaSetDevice(scene->device));
    const uint32_t rows = opts->row_count ? opts->row_count : cam->height;
    const size_t n_values = (size_t)rows * cam->width * 3u;
    int rc = ensure_img(scene, n_values);
    if (rc) return rc;
    if ((rc = enqueue_render(scene, cam, opts, scene->d_img, scene->stream))) return rc;
    CK(cudaMemcpyAsync(out_rgb, scene->d_img, n_values * sizeof(float), cudaMemcpyDeviceToHost, scene->stream));
    CK(cudaEventRecord(scene->ev2, scene->stream));
    CK(cudaStreamSynchronize(scene->stream));
    return RTW_OK;
}

int rtw_render_frames(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts, uint32_t first_frame,
                      uint32_t n_frames, float* out_rgb) {
    if (!scene || !cam || !opts || !out_rgb) return fail(RTW_E_INVALID, "null argument");
    const uint32_t rows = opts->row_count ? opts->row_count : cam->height;
    const size_t n_values = (size_t)rows * cam->width * 3u;
    rtw_render_opts o = *opts;
    for (uint32_t k = 0; k < n_frames; ++k) {   // render_multi (viewport.rs:256-267): only `frame` changes
        o.frame = first_frame + k;
        int rc = rtw_render(scene, cam, &o, out_rgb + (size_t)k * n_values);
        if (rc) return rc;
    }
    return RTW_OK;
}

int rtw_resolve_device(float* d_rgb, uint64_t n_pixels, uint32_t total_samples, float gamma, int device, void* stream) {
    if (!d_rgb || total_samples == 0 || !(gamma > 0.0f)) return fail(RTW_E_INVALID, "bad resolve arguments");
    DeviceGuard guard;
    if (device >= 0) CK(cudaSetDevice(device));
    const unsigned long long n = n_pixels * 3ull;
    resolve_kernel<<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_rgb, n, (float)total_samples, 1.0f / gamma);
    CK(cudaGetLastError());
    return RTW_OK;
}

int rtw_quantize_device(const float* d_rgb, uint8_t* d_rgb8, uint64_t n_values, int device, void* stream) {
    if (!d_rgb || !d_rgb8) return fail(RTW_E_INVALID, "null argument");
    DeviceGuard guard;
    if (device >= 0) CK(cudaSetDevice(device));
    quantize_kernel<<<grid_for(n_values, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_rgb, d_rgb8, n_values);
    CK(cudaGetLastError());
    return RTW_OK;
}

namespace {
// stack depth, shared-memory size and the occupancies that depend on it, per kernel family as at scene creation:
// depth[0] = {top, inner} of what the fast kernels walk, depth[1] = of the reference's structure
int refresh_launch_limits(rtw_scene* s, const uint32_t depth[2][2]) {
    uint32_t words[2];
    words[1] = s->any_levels ? stack_words(depth[1][0], depth[1][1]) : stack_words_flat(depth[1][0]);
    words[0] = s->fast_levels ? stack_words(depth[0][0], depth[0][1]) : stack_words_flat(depth[0][0]);
    if (words[0] > words[1]) words[0] = words[1];
    if (words[1] > 160u) return fail(RTW_E_UNSUPPORTED, "BVH deeper than the traversal stack");
    for (int k = 0; k < 2; ++k) { s->stack_words[k] = words[k]; s->smem_bytes[k] = (size_t)words[k] * RTW_MEGA_BLOCK * sizeof(uint32_t); }
    int rc = refresh_mega_occupancy(s);
    if (rc) return rc;
    if (s->d_wc && (rc = refresh_wf_occupancy(s))) return rc;
    return RTW_OK;
}
int set_instance_level(rtw_scene* s, uint32_t k, const Level& lv) {
    Instance* base = const_cast<Instance*>(s->d.instances);
    CK(cudaMemcpyAsync(reinterpret_cast<char*>(base + k) + offsetof(Instance, level), &lv, sizeof(Level),
                       cudaMemcpyHostToDevice, s->stream));
    return RTW_OK;
}
}  // namespace

int rtw_scene_rebuild_lbvh(rtw_scene* s, float* out_build_ms) {
    if (!s) return fail(RTW_E_INVALID, "null argument");
    DeviceGuard guard;
    CK(cudaSetDevice(s->device));
    uint32_t extra = 0;
    for (const auto& lv : s->lb_levels) if (lv.count >= 2u) extra += lv.count - 1u;
    if (out_build_ms) *out_build_ms = 0.f;
    if (extra == 0u) return RTW_OK;                       // nothing but one-primitive levels
    const size_t n_total = (size_t)s->n_sah_nodes + extra;
    if (!s->lb_nodes) {
        // the node arrays grow once: SAH nodes stay in front (their references remain valid), the
        // rebuilt levels go behind them
        // all four blocks or none: a later call must not find lb_nodes set and the others missing
        cudaError_t ae = cached_alloc_t(&s->lb_nodes, n_total * sizeof(Node));
        if (ae == cudaSuccess) ae = cached_alloc_t(&s->lb_nodes16, n_total * sizeof(Node16));
        if (ae == cudaSuccess) ae = cached_alloc_t(&s->lb_leaves_d, s->lb_leaves_h.size() * sizeof(LbvhLeaf));
        if (ae == cudaSuccess) ae = cached_alloc_t(&s->lb_out_d, s->lb_levels.size() * sizeof(LbvhLevelOut));
        if (ae != cudaSuccess) {
            cached_free(s->lb_nodes); cached_free(s->lb_nodes16); cached_free(s->lb_leaves_d); cached_free(s->lb_out_d);
            s->lb_nodes = nullptr; s->lb_nodes16 = nullptr; s->lb_leaves_d = nullptr; s->lb_out_d = nullptr;
            return cuda_fail(ae, "LBVH node arrays");
        }
        CK(cudaMemcpy(s->lb_nodes, s->d.nodes, (size_t)s->n_sah_nodes * sizeof(Node), cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(s->lb_nodes16, s->d.nodes16, (size_t)s->n_sah_nodes * sizeof(Node16), cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(s->lb_leaves_d, s->lb_leaves_h.data(), s->lb_leaves_h.size() * sizeof(LbvhLeaf), cudaMemcpyHostToDevice));
        s->info.device_bytes += n_total * (sizeof(Node) + sizeof(Node16)) + s->lb_leaves_h.size() * sizeof(LbvhLeaf);
        s->d.nodes = s->lb_nodes; s->d.nodes16 = s->lb_nodes16;
    }
    CK(cudaStreamSynchronize(s->stream));
    CK(cudaEventRecord(s->ev0, s->stream));
    uint32_t base = s->n_sah_nodes;
    std::vector<uint32_t> level_base(s->lb_levels.size(), 0u);
    if (s->lb_scratch.size() < s->lb_levels.size()) s->lb_scratch.resize(s->lb_levels.size());
    if (!s->lb_fork) CK(cudaEventCreateWithFlags(&s->lb_fork, cudaEventDisableTiming));
    CK(cudaEventRecord(s->lb_fork, s->stream));
    int lanes_used = 0, n_built = 0;
    for (size_t i = 0; i < s->lb_levels.size(); ++i) {
        const LbvhLevelDesc& lv = s->lb_levels[i];
        if (lv.count < 2u) continue;
        level_base[i] = base;
        const int lane = n_built++ % rtw_scene::kLbLanes;
        if (!s->lb_stream[lane]) {
            CK(cudaStreamCreateWithFlags(&s->lb_stream[lane], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&s->lb_join[lane], cudaEventDisableTiming));
        }
        if (lane >= lanes_used) { CK(cudaStreamWaitEvent(s->lb_stream[lane], s->lb_fork, 0)); lanes_used = lane + 1; }
        cudaError_t e = lbvh_build_level(s->lb_leaves_d + lv.first, lv.count, base, s->lb_nodes, s->lb_nodes16,
                                         s->lb_scratch[i], s->lb_out_d + i, s->lb_stream[lane]);
        if (e != cudaSuccess) return cuda_fail(e, "lbvh_build_level");
        base += lv.count - 1u;
    }
    for (int lane = 0; lane < lanes_used; ++lane) {
        CK(cudaEventRecord(s->lb_join[lane], s->lb_stream[lane]));
        CK(cudaStreamWaitEvent(s->stream, s->lb_join[lane], 0));
    }
    CK(cudaEventRecord(s->ev1, s->stream));
    std::vector<LbvhLevelOut> out(s->lb_levels.size());
    CK(cudaMemcpyAsync(out.data(), s->lb_out_d, out.size() * sizeof(LbvhLevelOut), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    if (out_build_ms) CK(cudaEventElapsedTime(out_build_ms, s->ev0, s->ev1));
    // stack bound first: if the rebuilt trees are too deep the scene keeps walking the SAH levels. Per kernel family,
    // like the SAH levels' bound: the fast kernels walk the baked top levels (-3 / -4) when there are any and only the
    // instances that were not baked — one bound for both families made C5's fast kernels carry 66-word stack columns
    // instead of 34 through the rebuilt levels, and cost them 8 % (profiles/r2r_lbvh_vs_sah.txt against r2s).
    uint32_t depth[2][2] = {{s->sah_fast_top_depth, s->sah_fast_inner_depth}, {s->sah_top_depth, s->sah_inner_depth}};
    for (size_t i = 0; i < s->lb_levels.size(); ++i) {
        if (s->lb_levels[i].count < 2u) continue;
        const uint32_t pushes = lbvh_pushes(out[i].max_depth);
        const int32_t owner = s->lb_levels[i].owner;
        if (owner < 0) {
            const bool baked_top = owner <= -3;
            depth[1][0] = std::max(depth[1][0], pushes);          // the strict kernels' bound covers every top level
            if (baked_top == s->has_baked) depth[0][0] = std::max(depth[0][0], pushes);
        } else {
            depth[1][1] = std::max(depth[1][1], pushes);
            if ((size_t)owner >= s->inst_baked.size() || !s->inst_baked[owner]) depth[0][1] = std::max(depth[0][1], pushes);
        }
    }
    int rc = refresh_launch_limits(s, depth);
    if (rc) return rc;
    for (size_t i = 0; i < s->lb_levels.size(); ++i) {
        const LbvhLevelDesc& lv = s->lb_levels[i];
        if (lv.count < 2u) continue;
        Level L = lv.owner < 0 ? s->sah_tops[-1 - lv.owner] : s->sah_inst_levels[lv.owner];
        L.root = out[i].root;
        for (int a = 0; a < 3; ++a) { L.lo[a] = out[i].lo[a]; L.hi[a] = out[i].hi[a]; }
        if (lv.owner < 0) { s->tops[-1 - lv.owner] = L; if (lv.owner == -1) s->d.top = L; }
        else if ((rc = set_instance_level(s, (uint32_t)lv.owner, L))) return rc;
    }
    CK(cudaStreamSynchronize(s->stream));
    return RTW_OK;
}

int rtw_scene_use_sah(rtw_scene* s) {
    if (!s) return fail(RTW_E_INVALID, "null argument");
    DeviceGuard guard;
    CK(cudaSetDevice(s->device));
    CK(cudaStreamSynchronize(s->stream));
    for (int k = 0; k < 4; ++k) s->tops[k] = s->sah_tops[k];
    s->d.top = s->sah_tops[0];
    int rc;
    for (const auto& lv : s->lb_levels)
        if (lv.owner >= 0 && (rc = set_instance_level(s, (uint32_t)lv.owner, s->sah_inst_levels[lv.owner]))) return rc;
    CK(cudaStreamSynchronize(s->stream));
    for (int k = 0; k < 2; ++k) { s->stack_words[k] = s->sah_stack_words[k]; s->smem_bytes[k] = (size_t)s->stack_words[k] * RTW_MEGA_BLOCK * sizeof(uint32_t); }
    int rc2 = refresh_mega_occupancy(s);
    if (rc2) return rc2;
    return s->d_wc ? refresh_wf_occupancy(s) : (int)RTW_OK;
}

int rtw_quantize(const float* rgb, uint8_t* rgb8, uint64_t n_values, int device) {
    if (!rgb || !rgb8) return fail(RTW_E_INVALID, "null argument");
    if (n_values == 0) return RTW_OK;
    DeviceGuard guard;
    if (device >= 0) CK(cudaSetDevice(device));
    float* d_in = nullptr; uint8_t* d_out = nullptr;
    cudaError_t e = cached_alloc_t(&d_in, n_values * sizeof(float));
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = cached_alloc_t(&d_out, n_values)) != cudaSuccess) { cached_free(d_in); return cuda_fail(e, "cudaMalloc"); }
    e = cudaMemcpy(d_in, rgb, n_values * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        quantize_kernel<<<grid_for(n_values, 256), 256>>>(d_in, d_out, n_values);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(rgb8, d_out, n_values, cudaMemcpyDeviceToHost);
    else cudaDeviceSynchronize();
    cached_free(d_in); cached_free(d_out);
    if (e != cudaSuccess) return cuda_fail(e, "rtw_quantize");
    return RTW_OK;
}

// Rust2/src/postprocessing.rs:70-135 on the device (rtw_post.cu)
int rtw_bilateral_filter_device(const uint8_t* d_rgb8, uint32_t width, uint32_t height, uint32_t size, int proximity,
                                uint8_t* d_out_rgb8, float* d_params, int device, void* stream) {
    if (!d_rgb8 || !d_out_rgb8) return fail(RTW_E_INVALID, "null argument");
    if (d_rgb8 == d_out_rgb8) return fail(RTW_E_INVALID, "bilateral filter cannot run in place");
    if (width < 2u || height < 2u || width > 32768u || height > 32768u) return fail(RTW_E_INVALID, "bilateral filter: image must be 2..32768 pixels a side");
    if (proximity != RTW_PROX_SQUARE && proximity != RTW_PROX_EDGES) return fail(RTW_E_INVALID, "unknown proximity type");
    if (size > 4096u) return fail(RTW_E_INVALID, "bilateral filter: window size > 4096");
    DeviceGuard guard;
    if (device >= 0) CK(cudaSetDevice(device));
    int launches = 0;
    cudaError_t e = bilateral_filter_enqueue(d_rgb8, width, height, size, proximity, d_out_rgb8, d_params,
                                             static_cast<cudaStream_t>(stream), &launches);
    if (e != cudaSuccess) return cuda_fail(e, "rtw_bilateral_filter_device");
    return RTW_OK;
}

int rtw_bilateral_filter(const uint8_t* rgb8, uint32_t width, uint32_t height, uint32_t size, int proximity,
                         uint8_t* out_rgb8, float* out_params, float* out_kernel_ms, int device) {
    if (!rgb8 || !out_rgb8) return fail(RTW_E_INVALID, "null argument");
    if (width < 2u || height < 2u || width > 32768u || height > 32768u) return fail(RTW_E_INVALID, "bilateral filter: image must be 2..32768 pixels a side");
    DeviceGuard guard;
    if (device >= 0) CK(cudaSetDevice(device));
    const size_t n = (size_t)width * height * 3u;
    uint8_t* d_buf = nullptr; float* d_par = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t e = cached_alloc_t(&d_buf, 2u * n + 256u);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    uint8_t* d_out = d_buf + (n + 255u) / 256u * 256u;
    int rc = RTW_OK;
    auto done = [&](int code) {
        cudaStreamSynchronize(nullptr);
        cached_free(d_buf); cached_free(d_par);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        return code;
    };
    if ((e = cached_alloc_t(&d_par, 4 * sizeof(float))) != cudaSuccess) return done(cuda_fail(e, "cudaMalloc"));
    if ((e = cudaEventCreate(&e0)) != cudaSuccess || (e = cudaEventCreate(&e1)) != cudaSuccess) return done(cuda_fail(e, "cudaEventCreate"));
    if ((e = cudaMemcpy(d_buf, rgb8, n, cudaMemcpyHostToDevice)) != cudaSuccess) return done(cuda_fail(e, "cudaMemcpy"));
    cudaEventRecord(e0, 0);
    rc = rtw_bilateral_filter_device(d_buf, width, height, size, proximity, d_out, d_par, -1, nullptr);
    if (rc) return done(rc);
    cudaEventRecord(e1, 0);
    if ((e = cudaMemcpy(out_rgb8, d_out, n, cudaMemcpyDeviceToHost)) != cudaSuccess) return done(cuda_fail(e, "rtw_bilateral_filter"));
    if (out_params && (e = cudaMemcpy(out_params, d_par, 4 * sizeof(float), cudaMemcpyDeviceToHost)) != cudaSuccess) return done(cuda_fail(e, "cudaMemcpy"));
    if (out_kernel_ms) { *out_kernel_ms = 0.f; cudaEventSynchronize(e1); cudaEventElapsedTime(out_kernel_ms, e0, e1); }
    return done(RTW_OK);
}

int rtw_render_rgb8(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts, uint8_t* out_rgb8) {
    if (!scene || !cam || !opts || !out_rgb8) return fail(RTW_E_INVALID, "null argument");
    if (opts->output != RTW_OUT_RESOLVED) return fail(RTW_E_INVALID, "rtw_render_rgb8 needs RTW_OUT_RESOLVED");
    DeviceGuard guard;
    CK(cudaSetDevice(scene->device));
    const uint32_t rows = opts->row_count ? opts->row_count : cam->height;
    const size_t n_values = (size_t)rows * cam->width * 3u;
    // the f32 image, then its 8-bit copy behind it in the same allocation
    int rc = ensure_img(scene, n_values + (n_values + 3u) / 4u);
    if (rc) return rc;
    if ((rc = enqueue_render(scene, cam, opts, scene->d_img, scene->stream))) return rc;
    uint8_t* d8 = reinterpret_cast<uint8_t*>(scene->d_img + n_values);
    quantize_kernel<<<grid_for(n_values, 256), 256, 0, scene->stream>>>(scene->d_img, d8, n_values);
    CK(cudaGetLastError());
    scene->launches += 1;
    CK(cudaMemcpyAsync(out_rgb8, d8, n_values, cudaMemcpyDeviceToHost, scene->stream));
    CK(cudaEventRecord(scene->ev2, scene->stream));
    CK(cudaStreamSynchronize(scene->stream));
    return RTW_OK;
}

int rtw_trim_memory(int device) {
    MemCache& c = mem_cache();
    std::lock_guard<std::mutex> lk(c.m);
    trim_locked(c, device);
    return RTW_OK;
}

int rtw_get_stats(rtw_scene* scene, rtw_stats* out) {
    if (!scene || !out) return fail(RTW_E_INVALID, "null argument");
    std::memset(out, 0, sizeof *out);
    if (!scene->have_stats) return RTW_OK;
    DeviceGuard guard;
    CK(cudaSetDevice(scene->device));
    CK(cudaEventSynchronize(scene->ev2));
    out->paths = scene->last_wavefront ? scene->wf_paths : scene->h_ctr->paths; out->rays = scene->h_ctr->rays; out->nan_paths = scene->h_ctr->nan_paths;
    out->gpu_launches = scene->launches;
    out->nodes = scene->h_ctr->nodes; out->sphere_tests = scene->h_ctr->spheres; out->quad_tests = scene->h_ctr->quads;
    out->instance_tests = scene->h_ctr->instances; out->hits = scene->h_ctr->hits;
    if (scene->last_wavefront) {
        out->iterations = scene->wf_iterations;
        // the samples still in the ring (the last kTimed at most) join the sums harvested on the way
        double se = scene->wf_extend_sum, ss = scene->wf_shade_sum;
        for (int k = scene->wf_harvested; k < scene->wf_timed; ++k) {
            float a = 0.f, b = 0.f;
            cudaEvent_t* te = scene->wf_t[k % rtw_scene::kTimed];
            CK(cudaEventElapsedTime(&a, te[0], te[1]));
            CK(cudaEventElapsedTime(&b, te[1], te[2]));
            se += a; ss += b;
        }
        const int n = scene->wf_timed;
        if (n > 0) { out->extend_ms = (float)(se / n); out->shade_ms = (float)(ss / n); }
    }
    CK(cudaEventElapsedTime(&out->kernel_ms, scene->ev0, scene->ev1));
    CK(cudaEventElapsedTime(&out->total_ms, scene->ev0, scene->ev2));
    return RTW_OK;
}

int rtw_get_sched_stats(rtw_scene* scene, uint64_t out[14]) {
    if (!scene || !out) return fail(RTW_E_INVALID, "null argument");
    std::memset(out, 0, 14 * sizeof(uint64_t));
    if (!scene->have_stats) return RTW_OK;
    DeviceGuard guard;
    CK(cudaSetDevice(scene->device));
    CK(cudaEventSynchronize(scene->ev2));
    out[0] = scene->h_ctr->sched_iters;
    for (int k = 0; k < 4; ++k) { out[1 + k] = scene->h_ctr->sched_steps[k]; out[5 + k] = scene->h_ctr->sched_lanes[k]; }
    for (int k = 0; k < 5; ++k) out[9 + k] = scene->h_ctr->sched_cycles[k];
    return RTW_OK;
}

int rtw_microbench(int device, int which, int reps, double* out_value) {
    if (!out_value || reps < 1) return fail(RTW_E_INVALID, "bad microbench arguments");
    DeviceGuard guard;
    if (device >= 0) CK(cudaSetDevice(device));
    int sms = 0, dev = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    double best = 0.0;
    if (which == RTW_BENCH_FP32) {
        float* d = nullptr; CK(cudaMalloc(&d, 16));
        const int iters = 1 << 16, block = 256, grid = sms * 8;
        for (int r = 0; r < reps + 1; ++r) {
            CK(cudaEventRecord(e0));
            bench_fp32_kernel<<<grid, block>>>(d, iters, 1.0000001f, 1e-7f);
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
            double tf = 2.0 * 8.0 * iters * (double)grid * block / (ms * 1e-3) / 1e12;
            if (r > 0 && tf > best) best = tf;
        }
        cudaFree(d);
    } else if (which == RTW_BENCH_L2 || which == RTW_BENCH_HBM) {
        const size_t bytes = which == RTW_BENCH_L2 ? (32ull << 20) : (2ull << 30);
        const int passes = which == RTW_BENCH_L2 ? 64 : 1;
        uint4* d = nullptr; unsigned* o = nullptr;
        CK(cudaMalloc(&d, bytes)); CK(cudaMalloc(&o, 16));
        CK(cudaMemset(d, 1, bytes));
        for (int r = 0; r < reps + 1; ++r) {
            CK(cudaEventRecord(e0));
            bench_read_kernel<<<sms * 8, 512>>>(d, bytes / 16, passes, o);
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
            double gbs = (double)bytes * passes / (ms * 1e-3) / 1e9;
            if (r > 0 && gbs > best) best = gbs;
        }
        cudaFree(d); cudaFree(o);
    } else {
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        return fail(RTW_E_INVALID, "unknown microbench");
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    CK(cudaGetLastError());
    *out_value = best;
    return RTW_OK;
}

int rtw_scene_get_info(rtw_scene* scene, rtw_scene_info* out) {
    if (!scene || !out) return fail(RTW_E_INVALID, "null argument");
    *out = scene->info;
    return RTW_OK;
}

}  // extern "C"
