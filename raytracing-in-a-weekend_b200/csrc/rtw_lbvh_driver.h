// rtw_lbvh_driver.h — host-callable interface of rtw_lbvh.cu (the on-device LBVH rebuild).
#pragma once
#include <cuda_runtime.h>
#include <cstring>
#include "rtw_lbvh.h"

namespace rtw {

struct LbvhLevelOut {      // what a level's build reports back (device memory, read after the stream syncs)
    float lo[3]; uint32_t root;
    float hi[3]; uint32_t max_depth;     // wide nodes on the longest way down (lbvh_pushes takes it)
};

struct LbvhScratch {       // device scratch, grown to the largest level seen and reused across rebuilds
    uint32_t *codes = nullptr, *codes_sorted = nullptr, *order_in = nullptr, *order = nullptr;
    int *left = nullptr, *right = nullptr, *parent_i = nullptr, *parent_l = nullptr;
    LbvhBox* ibox = nullptr;
    uint32_t* flag = nullptr;
    int* queue = nullptr;
    float* sah_cost = nullptr;
    float* dp_cut = nullptr;           // 4 per internal node
    uint32_t* dp_choice = nullptr;
    float* cb = nullptr;
    void* cub_tmp = nullptr;
    size_t n_codes = 0, cub_bytes = 0;
    LbvhScratch() = default;
    LbvhScratch(const LbvhScratch&) = delete;
    LbvhScratch& operator=(const LbvhScratch&) = delete;
    LbvhScratch(LbvhScratch&& o) noexcept { std::memcpy(this, &o, sizeof *this); std::memset(&o, 0, sizeof o); }   // (owns raw device pointers)
    ~LbvhScratch();
};

// Builds the level over d_leaves[0..n) (n >= 2) into d_nodes / d_nodes16 starting at node_base (at most
// n - 1 wide nodes) and writes the level's root and box to *d_out. Everything is enqueued on `st`.
cudaError_t lbvh_build_level(const LbvhLeaf* d_leaves, uint32_t n, uint32_t node_base, Node* d_nodes, Node16* d_nodes16,
                             LbvhScratch& w, LbvhLevelOut* d_out, cudaStream_t st);

}  // namespace rtw
