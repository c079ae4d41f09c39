// rtw_kernels.cu — the render hot path on sm_100a: camera ray generation, the ray_color recursion
// as an iterative loop, BVH traversal, material scatter, texture sampling and accumulation.
// Replaces render_row / async_render (Rust/src/viewport.rs:215-305) and everything below them.
//
// Megakernel with path regeneration: one warp owns an 8x4 pixel tile x a chunk of samples; every
// lane carries one path in registers and, the moment its path ends (miss, depth exhausted), starts
// its pixel's next sample, so all lanes stay in the extend/shade loop until the item is done.
// Scene data (BVH nodes 64 B, primitives 16-64 B) is read through the read-only path (LDG.128,
// L1/L2 resident: the whole C5 scene is < 0.5 MB); Perlin tables are staged in shared memory;
// image textures go through texture objects. Per-pixel sums leave the SM as three f32 REDs per
// lane per item. No tensor cores: nothing here is a contraction.
//
// Compiled three times (see rtw_kernels.h): RTW_VARIANT = fast | strict | count.
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

__global__ void __launch_bounds__(RTW_MEGA_BLOCK)
mega_kernel(const Scene sc, const RenderParams P, float* __restrict__ sum, Counters* __restrict__ ctr,
            uint32_t* __restrict__ work, const MegaLaunch L) {
    // ---- stage the Perlin tables (texture.rs:63-69) in shared memory
    __shared__ Noise s_noise[RTW_SMEM_NOISES];
    const Noise* nt = sc.noises;
    if (sc.n_noises > 0 && sc.n_noises <= RTW_SMEM_NOISES) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(sc.noises);
        uint32_t* dst = reinterpret_cast<uint32_t*>(s_noise);
        const uint32_t nwords = sc.n_noises * (uint32_t)(sizeof(Noise) / 4);
        for (uint32_t w = threadIdx.x; w < nwords; w += blockDim.x) dst[w] = src[w];
        nt = s_noise;
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t W = P.cam.width;
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
        item = __shfl_sync(0xFFFFFFFFu, item, 0);
        if (item >= L.n_items) break;
        const uint32_t chunk = item % L.n_chunks, tile = item / L.n_chunks;
        const uint32_t tx = tile % L.tiles_x, ty = tile / L.tiles_x;
        const uint32_t i = tx * 8u + (lane & 7u), jr = ty * 4u + (lane >> 3);
        const bool valid = i < W && jr < P.row_count;
        const uint32_t j = P.row_begin + jr;
        uint32_t s = P.sample_begin + chunk * L.chunk;
        const uint32_t s_end = min(s + L.chunk, P.sample_begin + P.sample_count);
        V3 acc = v3(0.f, 0.f, 0.f);
        if (valid) {
            Path path;
            bool alive = false;
            for (;;) {
                if (!alive) {
                    if (s >= s_end) break;
                    generate(P, i, j, s, path);
                    ++s;
                    alive = true;
                }
                if (path_step(sc, nt, P, path, n_rays, wp)) {
                    V3 c = path_result(path);
                    if (!finite3(c)) ++n_nan;
                    acc = acc + c;
                    ++n_paths;
                    alive = false;
                }
            }
            float* px = sum + ((size_t)jr * W + i) * 3u;
            atomicAdd(px + 0, acc.x);
            atomicAdd(px + 1, acc.y);
            atomicAdd(px + 2, acc.z);
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

}  // namespace variant

#define RTW_FN(name) RTW_CAT(RTW_CAT(name, _), RTW_VARIANT)

cudaError_t RTW_FN(launch_mega)(const Scene& sc, const RenderParams& P, float* d_sum, Counters* d_ctr,
                                uint32_t* d_work, const MegaLaunch& L, cudaStream_t st) {
    RTW_CAT(variant_, RTW_VARIANT)::mega_kernel<<<L.grid, L.block, 0, st>>>(sc, P, d_sum, d_ctr, d_work, L);
    return cudaGetLastError();
}

cudaError_t RTW_FN(mega_occupancy)(int* blocks_per_sm) {
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, RTW_CAT(variant_, RTW_VARIANT)::mega_kernel,
                                                         RTW_MEGA_BLOCK, 0);
}

}  // namespace rtw
