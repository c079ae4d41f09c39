// rtw_kernels.h — launch interface of rtw_kernels.cu. That file is compiled twice:
//   fast   : nvcc default (a*b+c contracts to FFMA)             -> launch_*_fast
//   strict : -fmad=false -DRTW_STRICT_FP (rustc-like arithmetic) -> launch_*_strict
//   count  : fast + -DRTW_COUNT_WORK (node/primitive test counters) -> launch_*_count
// selected per render call by RTW_FLAG_STRICT_FP.
#pragma once
#include <cuda_runtime.h>
#include "rtw_shade.h"

namespace rtw {

struct Counters {
    unsigned long long paths, rays, nan_paths;
    unsigned long long nodes, spheres, quads, instances, hits;   // only from the RTW_COUNT_WORK build
};

// Work decomposition of the path-regenerating megakernel: an item is one 8x4 pixel tile x one
// chunk of consecutive samples; warps pull items from a global counter.
struct MegaLaunch {
    uint32_t chunk;        // samples per item
    uint32_t n_chunks;
    uint32_t tiles_x, tiles_y;
    uint32_t n_items;
    int grid, block;
};

#define RTW_MEGA_BLOCK 128
#define RTW_SMEM_NOISES 2

#define RTW_DECLARE_VARIANT(SUFFIX)                                                                          \
    cudaError_t launch_mega_##SUFFIX(const Scene& sc, const RenderParams& P, float* d_sum, Counters* d_ctr,  \
                                     uint32_t* d_work, const MegaLaunch& L, cudaStream_t st);                \
    cudaError_t mega_occupancy_##SUFFIX(int* blocks_per_sm);

RTW_DECLARE_VARIANT(fast)
RTW_DECLARE_VARIANT(strict)
RTW_DECLARE_VARIANT(count)    // fast + algorithmic work counters (RTW_FLAG_COUNT_WORK)

}  // namespace rtw
