// rtw_kernels.cu — the render hot path on sm_100a: camera ray generation, the ray_color recursion
// as an iterative loop, BVH traversal, material scatter, texture sampling and accumulation.
// Replaces render_row / async_render (Rust/src/viewport.rs:215-305) and everything below them.
//
// Persistent-lane megakernel: a warp owns an 8x4 pixel tile x a chunk of samples (32*chunk paths).
// Every lane carries one path in registers; the warp alternates between three warp-synchronous
// phases so that the 32 lanes always execute the same code:
//   refill  lanes without a path take the item's next (pixel, sample) and generate its camera ray
//   extend  all lanes walk the two-level BVH in one loop (rtw_trace.h): lanes on a node step
//           together, then leaves / instance entry / exit; the loop is left early when fewer than
//           `refill_below` lanes still walk and others wait to be shaded (their state stays put)
//   shade   lanes whose search ended evaluate the Hit, scatter (next segment) or finish the path
// The traversal stack is a per-thread column in shared memory; scene data (BVH nodes 64 B,
// primitives 16-64 B) is read through the read-only path (LDG.128, L1/L2 resident: the whole C5
// scene is < 0.5 MB); Perlin tables are staged in shared memory; image textures go through
// texture objects. Finished paths add into the warp's tile sums in shared memory; the tile leaves
// the SM as three f32 REDs per pixel per item. No tensor cores: nothing here is a contraction.
//
// Compiled three times (see rtw_kernels.h): RTW_VARIANT = fast | strict | count.
#include <mutex>
#include "rtw_kernels.h"

#ifndef RTW_VARIANT
#define RTW_VARIANT fast
#endif
#define RTW_CAT2(a, b) a##b
#define RTW_CAT(a, b) RTW_CAT2(a, b)

namespace rtw {
namespace RTW_CAT(variant_, RTW_VARIANT) {

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
    return v;
}

// kLights: the instance that runs Rust2's light-biased integrators (shadow rays from inside segment_end: a second,
// non-inlined traversal). It is a kernel of its own so that the other integrators keep the register allocation
// they had without it (no stack frame, no spills). The plain instance is compiled for five resident blocks per SM: 96
// registers, still without a spill, against the 128 it takes when left alone — 20 warps per SM instead of 16 for a
// kernel that has no warp eligible on 42 % of its cycles (profiles/r2r_mega_kernel_C3_details.txt): C1 +7 %, C3 +9 %,
// C2 +2.6 %, C5 at 41 M paths +13 % (profiles/r2s_ab_mega_resident_blocks.txt; six blocks = 80 registers with
// spills and eight = 64 are slower again). The light instance spills more at 96 registers than at 128 and stays as it was.
template <bool kLights>
__device__ __forceinline__ void mega_body(const Scene& sc, const RenderParams& P, float* __restrict__ sum, Counters* __restrict__ ctr,
                                          uint32_t* __restrict__ work, const MegaLaunch& L) {
    extern __shared__ uint32_t s_stack[];                 // [stack_words][RTW_MEGA_BLOCK]
    __shared__ Noise s_noise[RTW_SMEM_NOISES];
    __shared__ float s_acc[RTW_MEGA_BLOCK / 32][3][32];   // per warp: the tile's 32 pixel sums
    // ---- stage the Perlin tables (texture.rs:63-69) in shared memory
    const Noise* nt = sc.noises;
    if (sc.n_noises > 0 && sc.n_noises <= RTW_SMEM_NOISES) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(sc.noises);
        uint32_t* dst = reinterpret_cast<uint32_t*>(s_noise);
        const uint32_t nwords = sc.n_noises * (uint32_t)(sizeof(Noise) / 4);
        for (uint32_t w = threadIdx.x; w < nwords; w += blockDim.x) dst[w] = src[w];
        nt = s_noise;
    }
    __syncthreads();

    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t W = P.cam.width;
    Stack S; S.base = s_stack + threadIdx.x; S.stride = RTW_MEGA_BLOCK;
    uint32_t n_paths = 0, n_rays = 0, n_nan = 0;
#if defined(RTW_COUNT_WORK)
    Work wk = {0u, 0u, 0u, 0u, 0u};
    Work* const wp = &wk;
#else
    Work* const wp = nullptr;
#endif

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(work, 1u);
        item = __shfl_sync(FULL, item, 0);
        if (item >= L.n_items) break;
        const uint32_t chunk = item % L.n_chunks, tile = item / L.n_chunks;
        const uint32_t tx = tile % L.tiles_x, ty = tile / L.tiles_x;
        const uint32_t s0 = P.sample_begin + chunk * L.chunk;
        const uint32_t ns = min(L.chunk, P.sample_begin + P.sample_count - s0);
        const uint32_t n_total = ns * 32u;                // paths of this item, handed out in order
        uint32_t next = 0;                                // warp-uniform
        s_acc[warp][0][lane] = 0.f; s_acc[warp][1][lane] = 0.f; s_acc[warp][2][lane] = 0.f;
        __syncwarp();

        Path path;
        Trav T;
        T.cur = CUR_DONE;
        TraceCtx cx = path_ctx(sc, P, path, wp);
        uint32_t pix = 0;
        bool alive = false;
        for (;;) {
            // ---- refill: every free lane takes the item's next path (generate, viewport.rs:287-298)
            const uint32_t want = __ballot_sync(FULL, !alive);
            if (want != 0u && next < n_total) {
                const uint32_t n = next + __popc(want & lt_mask);
                next += __popc(want);
                if (!alive && n < n_total) {
                    pix = n & 31u;
                    const uint32_t i = tx * 8u + (pix & 7u), jr = ty * 4u + (pix >> 3);
                    if (i < W && jr < P.row_count) {
                        generate(P, i, image_row(P, jr), s0 + (n >> 5), path);
                        alive = true;
                        cx = path_ctx(sc, P, path, wp);
                        if (!segment_begin(cx, P, path, T, S)) T.cur = CUR_DONE;   // depth 0: nothing to trace
                        else ++n_rays;
                    }
                }
            }
            if (!__any_sync(FULL, alive)) {
                if (next >= n_total) break;
                continue;
            }
            // ---- extend: warp-synchronous walk; all lanes on a node step together, then the rest
            for (;;) {
                const uint32_t walking = __ballot_sync(FULL, T.cur != CUR_DONE);
                if (walking == 0u) break;
                if ((uint32_t)__popc(walking) < L.refill_below && next < n_total &&
                    __any_sync(FULL, alive && T.cur == CUR_DONE)) break;
                if (T.cur < NODE_LIMIT) trav_step_node<true>(cx, T, S);
                if (T.cur >= NODE_LIMIT && T.cur != CUR_DONE) trav_step_other(cx, path.ray.time, T, S);
            }
            // ---- shade / scatter every lane whose search is over
            if (alive && T.cur == CUR_DONE) {
                bool finished = true;
                if (first_hit_integrator(P.integrator) || path.depth_left >= 1u) finished = segment_end<false>(cx, nt, P, path, T, kLights ? &S : nullptr, &n_rays);
                if (!finished) {
                    cx.bounce = P.depth - path.depth_left;
                    if (segment_begin(cx, P, path, T, S)) ++n_rays; else finished = true;
                }
                if (finished) {
                    V3 c = path_result(path);
                    if (!finite3(c)) ++n_nan;
                    atomicAdd(&s_acc[warp][0][pix], c.x);
                    atomicAdd(&s_acc[warp][1][pix], c.y);
                    atomicAdd(&s_acc[warp][2][pix], c.z);
                    ++n_paths;
                    alive = false;
                    T.cur = CUR_DONE;
                }
            }
        }
        __syncwarp();
        {
            const uint32_t i = tx * 8u + (lane & 7u), jr = ty * 4u + (lane >> 3);
            if (i < W && jr < P.row_count) {
                float* px = sum + ((size_t)jr * W + i) * 3u;
                atomicAdd(px + 0, s_acc[warp][0][lane]);
                atomicAdd(px + 1, s_acc[warp][1][lane]);
                atomicAdd(px + 2, s_acc[warp][2][lane]);
            }
        }
        __syncwarp();
    }
    unsigned long long p = warp_sum_u64(n_paths), r = warp_sum_u64(n_rays), n = warp_sum_u64(n_nan);
    if (lane == 0) {
        atomicAdd(&ctr->paths, p);
        atomicAdd(&ctr->rays, r);
        if (n) atomicAdd(&ctr->nan_paths, n);
    }
#if defined(RTW_COUNT_WORK)
    unsigned long long w0 = warp_sum_u64(wk.nodes), w1 = warp_sum_u64(wk.spheres), w2 = warp_sum_u64(wk.quads),
                       w3 = warp_sum_u64(wk.instances), w4 = warp_sum_u64(wk.hits);
    if (lane == 0) {
        atomicAdd(&ctr->nodes, w0); atomicAdd(&ctr->spheres, w1); atomicAdd(&ctr->quads, w2);
        atomicAdd(&ctr->instances, w3); atomicAdd(&ctr->hits, w4);
    }
#endif
}

__global__ void __launch_bounds__(RTW_MEGA_BLOCK, RTW_MEGA_MINB)
mega_kernel(const Scene sc, const RenderParams P, float* __restrict__ sum, Counters* __restrict__ ctr,
            uint32_t* __restrict__ work, const MegaLaunch L) {
    mega_body<false>(sc, P, sum, ctr, work, L);
}
__global__ void __launch_bounds__(RTW_MEGA_BLOCK)
mega_kernel_lights(const Scene sc, const RenderParams P, float* __restrict__ sum, Counters* __restrict__ ctr,
                   uint32_t* __restrict__ work, const MegaLaunch L) {
    mega_body<true>(sc, P, sum, ctr, work, L);
}

// =====================================================================================
// Wavefront pipeline (see rtw_kernels.h): extend + shade over a pool of path slots in HBM.
// =====================================================================================

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// Path records are read once and written once per iteration: with the streaming (evict-first) cache operators
// (ld.global.cs / st.global.cs) they leave the BVH nodes and primitives alone in L1 and L2. RTW_WF_STREAM: bit 0 =
// extend, bit 1 = shade; both: C5 at 625 spp +1.4 %, at 2 500 spp +1.1 %, C4 +0.6 % (each kernel's launches 1 % shorter;
// profiles/r3a_ab_streaming_records.txt).
#ifndef RTW_WF_STREAM
#define RTW_WF_STREAM 3
#endif
#if RTW_WF_STREAM & 1
#define EXT_LD(p) __ldcs(p)
#define EXT_ST(p, v) __stcs(p, v)
#else
#define EXT_LD(p) (*(p))
#define EXT_ST(p, v) (*(p) = (v))
#endif
#if RTW_WF_STREAM & 2
#define SHD_LD(p) __ldcs(p)
#define SHD_ST(p, v) __stcs(p, v)
#else
#define SHD_LD(p) (*(p))
#define SHD_ST(p, v) (*(p) = (v))
#endif

// extend: Scene::collision_normal for every ray of queue[cur]. Persistent warps. A warp reserves
// `chunk` consecutive queue positions with one atomic; whenever `fetch_min` of its lanes are idle
// they take the next positions of the chunk (already prefetched into L1) and join the walk loop,
// so the loop stays populated while the rays differ wildly in length. A lane whose search is over
// stores (t, primitive) at its ray's position.
__global__ void __launch_bounds__(RTW_WF_BLOCK, RTW_WF_EXTEND_MINB)
wf_extend_kernel(const Scene sc, const RenderParams P, const WfPool pool, WfCounters* __restrict__ wc,
                 Counters* __restrict__ ctr, const WfLaunch L) {
    extern __shared__ uint32_t s_stack[];                 // [stack_words][RTW_WF_BLOCK]
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    Stack S; S.base = s_stack + threadIdx.x; S.stride = RTW_WF_BLOCK;
    const uint32_t n = pool.capacity;
    const WfQueue q = pool.q;
    uint32_t n_rays = 0u;
#if defined(RTW_COUNT_WORK)
    Work wk = {0u, 0u, 0u, 0u, 0u};
    Work* const wp = &wk;
#else
    Work* const wp = nullptr;
#endif
    TraceCtx cx; cx.sc = &sc; cx.k0 = P.k0; cx.k1 = P.k1; cx.pixel = 0u; cx.sample = 0u; cx.bounce = 0u;
    cx.mint = P.mint; cx.maxt = P.maxt; cx.work = wp;
#if defined(RTW_COUNT_WORK)
    uint32_t sc_iters = 0u, sc_steps[4] = {0u, 0u, 0u, 0u}, sc_lanes[4] = {0u, 0u, 0u, 0u};
#define SCHED_COUNT(what, k, n) SCHED_COUNT_##what(k, n)
#define SCHED_COUNT_iters(k, n) (++sc_iters)
#define SCHED_COUNT_step(k, n) (++sc_steps[k], sc_lanes[k] += (uint32_t)(n))
    unsigned long long sc_cyc[5] = {0ull, 0ull, 0ull, 0ull, 0ull};
    long long sc_t0 = clock64(), sc_t1;
    int sc_kind = 4;
#define SCHED_TIME_BEGIN(k) (sc_t1 = clock64(), sc_cyc[4] += (unsigned long long)(sc_t1 - sc_t0), sc_t0 = sc_t1, sc_kind = (k))
#define SCHED_TIME_END() (sc_t1 = clock64(), sc_cyc[sc_kind] += (unsigned long long)(sc_t1 - sc_t0), sc_t0 = sc_t1, sc_kind = 4)
#else
#define SCHED_COUNT(what, k, n) ((void)0)
#define SCHED_TIME_BEGIN(k) ((void)0)
#define SCHED_TIME_END() ((void)0)
#endif
    Trav T; T.cur = CUR_DONE;
#if RTW_WF_PENDING
    uint32_t pend = REF_NONE;                             // the lane's parked leaf test
#endif
    uint32_t pos = 0u;
    float time = 0.f;
    RTW_WF_FLAG_T busy = false, more = n > 0u;         // (bool: the compiler packs them into bytes of one register and PRMTs them apart; A/B with uint32_t)
    uint32_t c_pos = 0u, c_end = 0u;                      // the warp's reserved chunk (uniform)
#if RTW_WF_WINDOW
    __shared__ RTW_WF_ORDER_T s_order[RTW_WF_BLOCK / 32][RTW_WF_WINDOW];
    __shared__ uint32_t s_bins[RTW_WF_BLOCK / 32][10];
    const uint32_t wid = threadIdx.x >> 5;
    uint32_t c_base = 0u;
#endif
    for (;;) {
#if RTW_WF_REDUX
        // One REDUX.SUM gives the lane count of every state at once (a byte each: node | leaf test << 8 | event << 16 |
        // idle << 24; a lane whose search has just ended counts nowhere) instead of a ballot + popc per state.
        uint32_t counts = __reduce_add_sync(FULL, !busy ? 0x01000000u : T.cur == CUR_DONE ? 0u : T.cur < NODE_LIMIT ? 1u : ref_is_event(T.cur) ? 0x10000u : 0x100u);
        if ((counts >> 24) == 32u && !more) break;
        if (more && (counts >> 24) >= L.fetch_min) {
            const uint32_t idle = __ballot_sync(FULL, !busy);
#else
        const uint32_t idle = __ballot_sync(FULL, !busy);
        if (idle == FULL && !more) break;
        if (more && (uint32_t)__popc(idle) >= L.fetch_min) {
#endif
#if RTW_WF_WINDOW
            // Window sort (A/B): the warp reserves a window of RTW_WF_WINDOW consecutive slots, reads their directions once
            // and takes them in the order camera rays / the eight direction octants of the bounced rays (dead slots
            // dropped), so that the rays that join the walk together start alike. c_pos / c_end index s_order.
            if (c_pos >= c_end && more) {
                uint32_t base = 0u;
                if (lane == 0u) base = atomicAdd(&wc->head, (uint32_t)RTW_WF_WINDOW);
                base = __shfl_sync(FULL, base, 0);
                c_pos = 0u; c_end = 0u; c_base = base;
                if (base >= n) more = false;
                else {
                    uint32_t live_groups = 0xFFFFFFFFu;
                    if (L.drain) {
                        const uint32_t f = lane < RTW_WF_WINDOW / 32 ? __ldg(&pool.warp_live[(base >> 5) + lane]) : 0u;
                        live_groups = __ballot_sync(FULL, f != 0u);
                    }
                    if (lane < 10u) s_bins[wid][lane] = 0u;
                    __syncwarp();
                    unsigned long long keys = 0ull;                            // 4 bits per round
#pragma unroll
                    for (int j = 0; j < RTW_WF_WINDOW / 32; ++j) {
                        uint32_t key = 9u;
                        if ((live_groups >> j) & 1u) {
#if RTW_WF_WINDOW_CS
                            const float4 rd = __ldcs(&q.d[base + 32u * j + lane]);     // streaming: the key pass must not evict BVH nodes from L1
#else
                            const float4 rd = q.d[base + 32u * j + lane];
#endif
                            const uint32_t bn = f2u(rd.w);
                            if (bn != WF_DEAD) key = bn == 0u ? 0u : 1u + ((rd.x < 0.f ? 1u : 0u) | (rd.y < 0.f ? 2u : 0u) | (rd.z < 0.f ? 4u : 0u));
                        }
                        keys |= (unsigned long long)key << (4 * j);
                        const uint32_t same = __match_any_sync(FULL, key);
                        if (lane == (uint32_t)(__ffs(same) - 1)) atomicAdd(&s_bins[wid][key], (uint32_t)__popc(same));
                    }
                    __syncwarp();
                    // exclusive scan of the ten bins (lanes 0..9)
                    uint32_t cnt = lane < 10u ? s_bins[wid][lane] : 0u, inc = cnt;
#pragma unroll
                    for (int o = 1; o < 16; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= (uint32_t)o) inc += t; }
                    __syncwarp();
                    if (lane < 10u) s_bins[wid][lane] = inc - cnt;
                    c_end = __shfl_sync(FULL, inc, 8);                          // everything but the dead bin
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < RTW_WF_WINDOW / 32; ++j) {
                        const uint32_t key = (uint32_t)(keys >> (4 * j)) & 15u;
                        const uint32_t same = __match_any_sync(FULL, key);
                        const uint32_t leader = (uint32_t)(__ffs(same) - 1);
                        uint32_t at = 0u;
                        if (lane == leader) at = atomicAdd(&s_bins[wid][key], (uint32_t)__popc(same));
                        at = __shfl_sync(FULL, at, leader) + __popc(same & lt_mask);
                        if (key != 9u) s_order[wid][at] = (RTW_WF_ORDER_T)(32u * j + lane);
                    }
                    __syncwarp();
                }
            }
            const uint32_t take = min((uint32_t)__popc(idle), c_end - c_pos);
            SCHED_COUNT(step, 3, take);
            SCHED_TIME_BEGIN(3);
            const uint32_t rank = __popc(idle & lt_mask);
            if (!busy && rank < take) {
                pos = c_base + s_order[wid][c_pos + rank];
                const float4 rd = q.d[pos];
                const float4 ro = q.o[pos];
                const uint2 id = q.id[pos];
                Ray r; r.o = v3(ro.x, ro.y, ro.z); r.d = v3(rd.x, rd.y, rd.z); r.time = ro.w;
                time = ro.w;
                cx.pixel = id.x; cx.sample = id.y; cx.bounce = f2u(rd.w);
                trav_begin(cx, r, T, S);
                busy = true;
                ++n_rays;
            }
            c_pos += take;
            if (c_pos + lane < c_end) {                   // the next rays of the window -> L1
                const uint32_t nx = c_base + s_order[wid][c_pos + lane];
                prefetch_l1(&q.o[nx]); prefetch_l1(&q.id[nx]);
            }
#else
            if (c_pos >= c_end) {
                uint32_t base = 0u;
                if (lane == 0u) base = atomicAdd(&wc->head, L.chunk);
                base = __shfl_sync(FULL, base, 0);
                c_pos = base; c_end = min(base + L.chunk, n);
                if (base >= n) { more = false; c_end = c_pos; }
            }
            if (L.drain) {
                // skip whole groups of 32 slots whose shade warp has gone dead for good: 32 flags per load
                while (c_pos < c_end && (c_pos & 31u) == 0u) {
                    const uint32_t ng = min(32u, (c_end - c_pos) >> 5);
                    if (ng == 0u) break;
                    const uint32_t f = lane < ng ? __ldg(&pool.warp_live[(c_pos >> 5) + lane]) : 0u;
                    const uint32_t lm = __ballot_sync(FULL, f != 0u);
                    if (lm == 0u) { c_pos += 32u * ng; continue; }
                    c_pos += 32u * (uint32_t)(__ffs(lm) - 1);
                    break;
                }
            }
            const uint32_t take = min((uint32_t)__popc(idle), c_end - c_pos);
            SCHED_COUNT(step, 3, take);
            SCHED_TIME_BEGIN(3);
            const uint32_t rank = __popc(idle & lt_mask);
            if (!busy && rank < take) {
                const float4 rd = EXT_LD(&q.d[c_pos + rank]);
                if (f2u(rd.w) != WF_DEAD) {               // dead slots only exist once the sample supply has run out
                    pos = c_pos + rank;
                    const float4 ro = EXT_LD(&q.o[pos]);
                    const uint2 id = EXT_LD(&q.id[pos]);
                    Ray r; r.o = v3(ro.x, ro.y, ro.z); r.d = v3(rd.x, rd.y, rd.z); r.time = ro.w;
                    time = ro.w;
                    cx.pixel = id.x; cx.sample = id.y; cx.bounce = f2u(rd.w);
                    trav_begin(cx, r, T, S);
                    busy = true;
                    ++n_rays;
                }
            }
            c_pos += take;
            if (c_pos + lane < c_end) {                   // the next positions of the chunk -> L1
                prefetch_l1(&q.o[c_pos + lane]); prefetch_l1(&q.d[c_pos + lane]);
                if ((lane & 1u) == 0u) prefetch_l1(&q.id[c_pos + lane]);
            }
#endif
            __syncwarp();
            SCHED_TIME_END();
#if RTW_WF_REDUX
            counts = __reduce_add_sync(FULL, !busy ? 0x01000000u : T.cur == CUR_DONE ? 0u : T.cur < NODE_LIMIT ? 1u : ref_is_event(T.cur) ? 0x10000u : 0x100u);
#endif
        }
        // ---- one step of ONE kind per iteration: the kind most lanes are waiting for. Lanes in
        // other states sit the iteration out, so that the rarer, dearer steps (leaf tests, instance
        // entry / exit) run with many lanes at once instead of after every node step. (Voting sphere
        // tests, quad tests and one-sphere instances as separate kinds starves the rarer ones: -11 %
        // on C5, profiles/r1h_ab_split_fastprim_volskip.txt; the tests are branch-free instead.)
#if RTW_WF_PENDING
        // A lane that pops a leaf test does not wait for a leaf step: it parks the test in `pend`
        // and carries on with the next reference, so it keeps taking node steps; the parked test
        // runs in the next leaf step, which every lane with a parked test joins. Only a second leaf,
        // an instance entry, a level exit or the end of the search force the parked test first. The
        // closest hit does not depend on the order of the tests (closer() breaks ties by reference).
        if (pend == REF_NONE && ref_is_leaf_test(T.cur)) { pend = T.cur; T.cur = S.pop(T.sp); }
        const uint32_t cur = T.cur;
        const bool has_p = pend != REF_NONE;
        const bool on_node = cur < NODE_LIMIT;
        const bool must_p = has_p && !on_node;
        const bool on_event = !has_p && ref_is_event(cur);
        const int cn = __popc(__ballot_sync(FULL, on_node)), cm = __popc(__ballot_sync(FULL, must_p)),
                  ce = __popc(__ballot_sync(FULL, on_event));
        SCHED_COUNT(iters, 0, 0);
        if (cn >= cm && cn >= ce) {
            SCHED_COUNT(step, 0, cn);
            SCHED_TIME_BEGIN(0);
            if (on_node) trav_step_node<false>(cx, T, S);
            SCHED_TIME_END();
        } else if (cm >= ce) {
            SCHED_COUNT(step, 1, __popc(__ballot_sync(FULL, has_p)));
            SCHED_TIME_BEGIN(1);
            if (has_p) { trav_test_parked(cx, pend, T, S); pend = REF_NONE; }
            SCHED_TIME_END();
        } else {
            SCHED_COUNT(step, 2, ce);
            SCHED_TIME_BEGIN(2);
            if (on_event) trav_step_event(cx, time, T, S);
            SCHED_TIME_END();
        }
        if (busy && T.cur == CUR_DONE && pend == REF_NONE) {
#else
        const uint32_t cur = T.cur;
        const bool on_node = cur < NODE_LIMIT, on_event = ref_is_event(cur), on_prim = !on_node && !on_event && cur != CUR_DONE;
#if RTW_WF_REDUX
        const int cn = (int)(counts & 0xFFu);
        int cp = (int)((counts >> 8) & 0xFFu), ce = (int)((counts >> 16) & 0xFFu);
        bool node_step = cn * 4 >= cp * (int)L.prim_w && cn * 4 >= ce * (int)L.event_w;
#else
        const int cn = __popc(__ballot_sync(FULL, on_node));
        int cp = 0, ce = 0;
        bool node_step = true;
#endif
#if RTW_WF_FASTVOTE && !RTW_WF_REDUX
        // With `node_sure` or more lanes on a node the node step wins whatever the other 32 - cn lanes wait for
        // (cn * 4 >= (32 - cn) * max(prim_w, event_w)): the two other ballots are only taken below that.
        if (cn < (int)L.node_sure)
#endif
#if !RTW_WF_REDUX
        {
            cp = __popc(__ballot_sync(FULL, on_prim)); ce = __popc(__ballot_sync(FULL, on_event));
            node_step = cn * 4 >= cp * (int)L.prim_w && cn * 4 >= ce * (int)L.event_w;
        }
#endif
        SCHED_COUNT(iters, 0, 0);
        if (node_step) {
            SCHED_COUNT(step, 0, cn);
            SCHED_TIME_BEGIN(0);
            if (on_node) {
                trav_step_node<false>(cx, T, S);
#if RTW_WF_PREFETCH
                if (T.cur < NODE_LIMIT) prefetch_l1(&sc.nodes16[T.cur]);
                else if ((T.cur >> 30) == KIND_SPHERE) prefetch_l1(&sc.spheres[ref_index(T.cur)]);
                else if ((T.cur >> 30) == KIND_QUAD) prefetch_l1(&sc.quads[ref_index(T.cur)]);
#endif
            }
#if RTW_WF_NODE_BURST > 1
#pragma unroll 1
            for (int burst = 1; burst < RTW_WF_NODE_BURST; ++burst) {
                const bool again = T.cur < NODE_LIMIT;
                if (__popc(__ballot_sync(FULL, again)) < cn - (cn >> 2)) break;   // most of the node lanes are still on nodes
                SCHED_COUNT(step, 0, cn);
                if (again) trav_step_node<false>(cx, T, S);
            }
#endif
            SCHED_TIME_END();
        } else if (cp >= ce) {
            SCHED_COUNT(step, 1, cp);
            SCHED_TIME_BEGIN(1);
            if (on_prim) trav_step_prim(cx, T, S);
            SCHED_TIME_END();
        } else {
            SCHED_COUNT(step, 2, ce);
            SCHED_TIME_BEGIN(2);
            if (on_event) trav_step_event(cx, time, T, S);
            SCHED_TIME_END();
        }
        if (busy && T.cur == CUR_DONE) {
#endif
            EXT_ST(&pool.hit[pos], make_uint4(f2u(T.t), T.ref, T.inner, 0u));
            busy = false;
        }
    }
    {
        const unsigned long long r = warp_sum_u64(n_rays);
        if (lane == 0u && r != 0ull) atomicAdd(&ctr->rays, r);
    }
#if defined(RTW_COUNT_WORK)
    unsigned long long w0 = warp_sum_u64(wk.nodes), w1 = warp_sum_u64(wk.spheres), w2 = warp_sum_u64(wk.quads),
                       w3 = warp_sum_u64(wk.instances);
    if (lane == 0) {
        atomicAdd(&ctr->nodes, w0); atomicAdd(&ctr->spheres, w1); atomicAdd(&ctr->quads, w2); atomicAdd(&ctr->instances, w3);
        atomicAdd(&ctr->sched_iters, (unsigned long long)sc_iters);
        for (int k = 0; k < 4; ++k) {
            atomicAdd(&ctr->sched_steps[k], (unsigned long long)sc_steps[k]);
            atomicAdd(&ctr->sched_lanes[k], (unsigned long long)sc_lanes[k]);
        }
        sc_cyc[4] += (unsigned long long)(clock64() - sc_t0);
        for (int k = 0; k < 5; ++k) atomicAdd(&ctr->sched_cycles[k], sc_cyc[k]);
    }
#endif
#undef SCHED_COUNT
#undef SCHED_TIME_BEGIN
#undef SCHED_TIME_END
}

// shade: one thread per pool slot, record rewritten in place. init != 0: first fill.
__global__ void __launch_bounds__(RTW_WF_SHADE_BLOCK, RTW_WF_SHADE_MINB)
wf_shade_kernel(const Scene sc, const RenderParams P, const WfPool pool, WfCounters* __restrict__ wc,
                Counters* __restrict__ ctr, float* __restrict__ sum, const uint32_t mark, const int init, const WfLaunch L) {
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e == 0u) {
        wc->head = 0u;                                      // this iteration's extend is over
        wc->iterations += 1ull;
    }
    const Noise* nt = sc.noises;
#if RTW_WF_SHADE_SMEM_NOISE
    // A/B of north_star's "Perlin tables staged in shared memory" for THIS kernel (the persistent megakernel stages
    // them once per block lifetime, above). Here a block lives for 128 slots: every block of every launch copies the
    // 4.9 KB table (38 B per slot, a third of the slot's own record) to serve the few slots that hit a noise texture,
    // and 10 resident blocks hold 49-98 KB of shared memory that L1 loses. Measured, not kept
    // (profiles/r2*_ab_shade_smem_noise.txt); the tables stay in global memory, where the lanes that need them
    // find them in L1 / L2.
    __shared__ Noise s_noise[RTW_SMEM_NOISES];
    if (sc.n_noises > 0 && sc.n_noises <= RTW_SMEM_NOISES) {
        const uint4* src = reinterpret_cast<const uint4*>(sc.noises);
        uint4* dst = reinterpret_cast<uint4*>(s_noise);
        const uint32_t nvec = sc.n_noises * (uint32_t)(sizeof(Noise) / 16);
        for (uint32_t w = threadIdx.x; w < nvec; w += blockDim.x) dst[w] = __ldg(&src[w]);
        nt = s_noise;
        __syncthreads();
    }
#endif
    const WfQueue q = pool.q;                               // grid * block == capacity (a multiple of 256)
    const uint32_t W = P.cam.width;
    const unsigned long long total = L.total_paths;        // a launch parameter: reading it from wc put an L2 round trip at the head of every warp
    // live_mark is read here, long before it is needed: once the first warps of the launch have set
    // it, every later warp sees `mark` already and leaves without waiting for an L2 round trip (that
    // wait was 21 % of the kernel's memory stalls, profiles/r1i_wf_shade_by_line.txt). Warps that
    // saw the old value look again before storing: thousands of stores to one word cost 3x the
    // whole kernel (profiles/r1j_ab_shade_live_mark.txt).
    const uint32_t live_mark_seen = __ldcg(&wc->live_mark);
#if defined(RTW_COUNT_WORK)
    Work wk = {0u, 0u, 0u, 0u, 0u};
    Work* const wp = &wk;
#else
    Work* const wp = nullptr;
#endif
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    Path path;
    RTW_WF_FLAG_T alive = false, was_live = false, fresh = false;
    if (L.drain && !init && __ldg(&pool.warp_live[e >> 5]) == 0u) return;   // (whole warp) dead for good: WfLaunch::drain
    if (!init) {
        const float4 rd = SHD_LD(&q.d[e]);
        was_live = f2u(rd.w) != WF_DEAD;
        if (was_live) {
            const float4 ro = SHD_LD(&q.o[e]);
            const uint2 id = SHD_LD(&q.id[e]);
            const uint4 hit = SHD_LD(&pool.hit[e]);
            const float4 thr = SHD_LD(&q.thr[e]);
            path.ray.time = ro.w;
            path.thr = v3(thr.x, thr.y, thr.z); path.rad = v3(0.f, 0.f, 0.f);   // this segment's contribution only
            path.pixel = id.x; path.sample = id.y;
            const uint32_t st = f2u(thr.w);
            path.block = st & WF_BLOCK_MASK; path.depth_left = (st >> WF_DEPTH_SHIFT) & WF_DEPTH_MASK; path.poison = (st >> 31) != 0u;
            Trav T;
            T.o = v3(ro.x, ro.y, ro.z); T.d = v3(rd.x, rd.y, rd.z);
            T.t = u2f(hit.x); T.ref = hit.y; T.inner = hit.z;
            TraceCtx cx = path_ctx(sc, P, path, wp);
            bool finished = segment_end<true>(cx, nt, P, path, T);
            if (!finished && path.depth_left < 1u) finished = true;   // ray_color.rs:14-16: the next call returns black
            // L = emitted + col_mod * L(next): what this segment adds (thr * emitted, thr *
            // background) goes straight into the frame, so no radiance travels with the path. A path
            // that is or becomes NaN (N8) adds NaN once, when it ends, as the reference's sum would be.
            V3 c = path.rad;
            if (!finite3(c)) path.poison = true;
            float* px = sum + (size_t)output_pixel(P, path.pixel) * 3u;
            if (finished && path.poison) {
                atomicAdd(&ctr->nan_paths, 1ull);
                const float qn = u2f(0x7FC00000u);
                c = v3(qn, qn, qn);
            }
            if ((finished || !path.poison) && (c.x != 0.f || c.y != 0.f || c.z != 0.f)) {
                atomicAdd(px + 0, c.x); atomicAdd(px + 1, c.y); atomicAdd(px + 2, c.z);
            }
            alive = !finished;
        }
    }
    // ---- regeneration: free slots take the next camera samples of the warp's reserved range.
    // 32 consecutive indices = one 8x4 tile x one sample, so primary rays that start together are coherent.
    const bool need = !alive && !(!first_hit_integrator(P.integrator) && P.depth < 1u);
    const uint32_t m = __ballot_sync(FULL, need);
    if (m != 0u) {
        unsigned long long* sup = pool.supply + 2u * (e >> 5);
        unsigned long long nx = sup[0], en = sup[1];                  // same address for the whole warp
        const uint32_t cnt = __popc(m), rank = __popc(m & ((1u << lane) - 1u));
        const unsigned long long avail = en - nx;
        unsigned long long base = 0ull, bend = 0ull;
        if (avail < cnt) {                                            // refill: one atomic per warp per supply_chunk samples
            if (lane == 0u) {                                         // (no atomic once the cursor has passed the end)
                base = *reinterpret_cast<volatile unsigned long long*>(&wc->next_path);
                if (base < total) base = atomicAdd(&wc->next_path, (unsigned long long)L.supply_chunk);
            }
            base = __shfl_sync(FULL, base, 0);
            if (base > total) base = total;
            bend = base + L.supply_chunk < total ? base + L.supply_chunk : total;
        }
        unsigned long long k = ~0ull;
        if (need) {
            if (rank < avail) k = nx + rank;
            else if (base + (rank - avail) < bend) k = base + (rank - avail);
        }
        if (avail < cnt) { nx = base + (cnt - avail) < bend ? base + (cnt - avail) : bend; en = bend; }
        else nx += cnt;
        if (lane == 0u) { sup[0] = nx; sup[1] = en; }
        if (k != ~0ull) {
            const uint32_t pix = (uint32_t)(k & 31ull);
            const unsigned long long g = k >> 5;
            uint32_t tile, s;
            if (L.sample_major) {                                     // the whole frame once per sample
                const uint32_t nt = L.tiles_x * L.tiles_y;
                if (L.index32) { const uint32_t g32 = (uint32_t)g; s = g32 / nt; tile = g32 - s * nt; }
                else { s = (uint32_t)(g / nt); tile = (uint32_t)(g % nt); }
                s += P.sample_begin;
            } else {                                                  // every sample of a tile, then the next tile
                if (L.index32) { const uint32_t g32 = (uint32_t)g; tile = g32 / P.sample_count; s = g32 - tile * P.sample_count; }
                else { tile = (uint32_t)(g / P.sample_count); s = (uint32_t)(g % P.sample_count); }
                s += P.sample_begin;
            }
            const uint32_t tx = tile % L.tiles_x, ty = tile / L.tiles_x;
            const uint32_t i = tx * 8u + (pix & 7u), jr = ty * 4u + (pix >> 3);
            if (i < W && jr < P.row_count) {                          // edge tiles are padded; such an index is dropped
                generate(P, i, image_row(P, jr), s, path);
                alive = true;
                fresh = true;
            }
        }
    }
    {
        // A warp that ends a launch without a live slot found the sample supply exhausted (every lane asked and none
        // got a sample): it stays dead for the rest of the frame, and says so for the drain (WfLaunch::drain).
        const uint32_t alive_mask = __ballot_sync(FULL, alive);
        if (lane == 0u && (init || alive_mask == 0u)) pool.warp_live[e >> 5] = alive_mask != 0u ? 1u : 0u;
    }
    if (!alive && !was_live && !init) return;                         // dead before, dead now: nothing to write
    if (alive) {
        SHD_ST(&q.o[e], make_float4(path.ray.o.x, path.ray.o.y, path.ray.o.z, path.ray.time));
        SHD_ST(&q.d[e], make_float4(path.ray.d.x, path.ray.d.y, path.ray.d.z, u2f(P.depth - path.depth_left)));
        if (fresh) SHD_ST(&q.id[e], make_uint2(path.pixel, path.sample));     // a continuing path keeps its identity
        SHD_ST(&q.thr[e], make_float4(path.thr.x, path.thr.y, path.thr.z,
                               u2f((path.block & WF_BLOCK_MASK) | (path.depth_left << WF_DEPTH_SHIFT) | (path.poison ? 0x80000000u : 0u))));
        const uint32_t am = __activemask();
        // same value from every warp; read first (L2) so that only the first few warps store
        if (lane == (uint32_t)(__ffs(am) - 1) && live_mark_seen != mark && __ldcg(&wc->live_mark) != mark) wc->live_mark = mark;
    } else {
        q.d[e] = make_float4(0.f, 0.f, 0.f, u2f(WF_DEAD));
    }
#if defined(RTW_COUNT_WORK)
    if (wk.hits) atomicAdd(&ctr->hits, (unsigned long long)wk.hits);
#endif
}

}  // namespace variant

#define RTW_FN(name) RTW_CAT(RTW_CAT(name, _), RTW_VARIANT)

cudaError_t RTW_FN(launch_mega)(const Scene& sc, const RenderParams& P, float* d_sum, Counters* d_ctr,
                                uint32_t* d_work, const MegaLaunch& L, cudaStream_t st) {
    if (light_integrator(P.integrator))
        RTW_CAT(variant_, RTW_VARIANT)::mega_kernel_lights<<<L.grid, L.block, L.smem_bytes, st>>>(sc, P, d_sum, d_ctr, d_work, L);
    else
        RTW_CAT(variant_, RTW_VARIANT)::mega_kernel<<<L.grid, L.block, L.smem_bytes, st>>>(sc, P, d_sum, d_ctr, d_work, L);
    return cudaGetLastError();
}

cudaError_t RTW_FN(launch_wf_extend)(const Scene& sc, const RenderParams& P, const WfPool& pool, WfCounters* d_wc,
                                     Counters* d_ctr, const WfLaunch& L, cudaStream_t st) {
    RTW_CAT(variant_, RTW_VARIANT)::wf_extend_kernel<<<L.extend_grid, RTW_WF_BLOCK, L.smem_bytes, st>>>(sc, P, pool, d_wc, d_ctr, L);
    return cudaGetLastError();
}

cudaError_t RTW_FN(launch_wf_shade)(const Scene& sc, const RenderParams& P, const WfPool& pool, WfCounters* d_wc,
                                    Counters* d_ctr, float* d_sum, uint32_t mark, int init, const WfLaunch& L, cudaStream_t st) {
    RTW_CAT(variant_, RTW_VARIANT)::wf_shade_kernel<<<L.shade_grid, RTW_WF_SHADE_BLOCK, 0, st>>>(sc, P, pool, d_wc, d_ctr, d_sum, mark, init, L);
    return cudaGetLastError();
}

// The opt-in limit for dynamic shared memory only ever grows (per kernel and device, rounded up to 8 KB): scenes
// with different stack depths share these kernels, also across host threads (rtw_multi.cu), and a limit that followed
// the last query made a concurrent launch of a deeper scene fail with "invalid argument".
static cudaError_t raise_smem_limit(const void* kernel, size_t smem_bytes, int which) {
    static std::mutex m;
    static int limit[3][64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const int want = (int)((smem_bytes + 8191u) & ~(size_t)8191u);
    std::lock_guard<std::mutex> lk(m);
    if (dev >= 0 && dev < 64 && limit[which][dev] >= want) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, want);
    if (e == cudaSuccess && dev >= 0 && dev < 64) limit[which][dev] = want;
    return e;
}
cudaError_t RTW_FN(wf_extend_occupancy)(int* blocks_per_sm, size_t smem_bytes) {
    cudaError_t e = raise_smem_limit((const void*)RTW_CAT(variant_, RTW_VARIANT)::wf_extend_kernel, smem_bytes, 0);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, RTW_CAT(variant_, RTW_VARIANT)::wf_extend_kernel,
                                                         RTW_WF_BLOCK, smem_bytes);
}

cudaError_t RTW_FN(mega_occupancy)(int* blocks_per_sm, size_t smem_bytes) {
    // blocks_per_sm[0]: the plain instance, [1]: the light-biased one (their register counts differ)
    cudaError_t e = raise_smem_limit((const void*)RTW_CAT(variant_, RTW_VARIANT)::mega_kernel, smem_bytes, 1);
    if (e == cudaSuccess) e = raise_smem_limit((const void*)RTW_CAT(variant_, RTW_VARIANT)::mega_kernel_lights, smem_bytes, 2);
    if (e != cudaSuccess) return e;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm[0], RTW_CAT(variant_, RTW_VARIANT)::mega_kernel, RTW_MEGA_BLOCK, smem_bytes);
    if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm[1], RTW_CAT(variant_, RTW_VARIANT)::mega_kernel_lights, RTW_MEGA_BLOCK, smem_bytes);
    return e;
}

}  // namespace rtw
