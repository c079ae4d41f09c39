// rtw_lbvh.cu — device driver of the LBVH rebuild (rtw_lbvh.h has the per-element steps and the
// references to the builders it replaces). One level at a time on the caller's stream; the sort
// and the scan come from CUB (a plain library sort is all this off-path step needs).
#include <cub/cub.cuh>
#include "rtw_lbvh.h"
#include "rtw_lbvh_driver.h"

namespace rtw {

namespace {

__device__ __forceinline__ uint32_t ordered_bits(float f) {          // monotone float -> uint
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_float(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

// centroid bounds of the level -> clo[3], cinv[3] (one block; levels are small)
__global__ void lbvh_bounds_kernel(const LbvhLeaf* __restrict__ leaves, uint32_t n, float* __restrict__ cb) {
    __shared__ uint32_t s_lo[3], s_hi[3];
    if (threadIdx.x < 3) { s_lo[threadIdx.x] = 0xFFFFFFFFu; s_hi[threadIdx.x] = 0u; }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        for (int a = 0; a < 3; ++a) {
            const float c = 0.5f * leaves[i].lo[a] + 0.5f * leaves[i].hi[a];
            if (c == c) { atomicMin(&s_lo[a], ordered_bits(c)); atomicMax(&s_hi[a], ordered_bits(c)); }
        }
    __syncthreads();
    if (threadIdx.x < 3) {
        const float lo = ordered_float(s_lo[threadIdx.x]), hi = ordered_float(s_hi[threadIdx.x]);
        cb[threadIdx.x] = lo;
        cb[3 + threadIdx.x] = hi > lo ? 1.0f / (hi - lo) : 0.0f;
    }
}

__global__ void lbvh_morton_kernel(const LbvhLeaf* __restrict__ leaves, uint32_t n, const float* __restrict__ cb,
                                   uint32_t* __restrict__ codes, uint32_t* __restrict__ order) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float clo[3] = {cb[0], cb[1], cb[2]}, cinv[3] = {cb[3], cb[4], cb[5]};
    codes[i] = lbvh_morton(leaves[i], clo, cinv);
    order[i] = i;
}

__global__ void lbvh_karras_kernel(const uint32_t* __restrict__ codes, int n, int* left, int* right, int* parent_i,
                                   int* parent_l, uint32_t* flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    flag[i] = 0u;
    lbvh_karras(codes, n, i, left, right, parent_i, parent_l);
}

__device__ __forceinline__ LbvhBox load_box_cg(const LbvhBox* p) {   // written by another thread: bypass L1
    LbvhBox b;
    const float* f = reinterpret_cast<const float*>(p);
    b.lo[0] = __ldcg(f + 0); b.lo[1] = __ldcg(f + 1); b.lo[2] = __ldcg(f + 2);
    b.hi[0] = __ldcg(f + 3); b.hi[1] = __ldcg(f + 4); b.hi[2] = __ldcg(f + 5);
    return b;
}

// bottom-up fit: every leaf walks up; the second thread to reach a node computes its box
__global__ void lbvh_fit_kernel(const LbvhLeaf* __restrict__ leaves, const uint32_t* __restrict__ order, int n,
                                const int* __restrict__ left, const int* __restrict__ right,
                                const int* __restrict__ parent_i, const int* __restrict__ parent_l, LbvhBox* ibox,
                                uint32_t* flag) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int cur = parent_l[p];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&flag[cur], 1u) == 0u) return;
        __threadfence();
        const int l = left[cur], r = right[cur];
        const LbvhBox a = l < 0 ? lbvh_leaf_box(leaves[order[~l]]) : load_box_cg(&ibox[l]);
        const LbvhBox b = r < 0 ? lbvh_leaf_box(leaves[order[~r]]) : load_box_cg(&ibox[r]);
        ibox[cur] = lbvh_union(a, b);
        cur = parent_i[cur];
    }
}

__global__ void lbvh_depth_kernel(const int* __restrict__ parent_i, int n, uint32_t* __restrict__ wide,
                                  uint32_t* __restrict__ max_depth) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const uint32_t d = lbvh_depth(parent_i, i);
    wide[i] = (d & 1u) ? 0u : 1u;
    atomicMax(max_depth, d + 1u);          // its leaf children sit one level below
}

__global__ void lbvh_emit_kernel(LbvhTree t, int n, const uint32_t* __restrict__ wide, Node* nodes, Node16* nodes16) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || !wide[i]) return;
    lbvh_emit(t, i, nodes, nodes16);
}

__global__ void lbvh_finish_kernel(const LbvhBox* __restrict__ ibox, uint32_t node_base, const uint32_t* __restrict__ max_depth,
                                   LbvhLevelOut* out) {
    for (int a = 0; a < 3; ++a) { out->lo[a] = ibox[0].lo[a]; out->hi[a] = ibox[0].hi[a]; }
    out->root = make_ref(KIND_NODE, node_base);      // the binary root is the level's first wide node
    out->max_depth = *max_depth;
}

template <class T>
cudaError_t grow(T** p, size_t* have, size_t need) {
    if (*have >= need) return cudaSuccess;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(p), need * sizeof(T));
    if (e == cudaSuccess) *have = need;
    return e;
}

}  // namespace

LbvhScratch::~LbvhScratch() {
    cudaFree(codes); cudaFree(codes_sorted); cudaFree(order_in); cudaFree(order); cudaFree(left); cudaFree(right);
    cudaFree(parent_i); cudaFree(parent_l); cudaFree(ibox); cudaFree(flag); cudaFree(wide); cudaFree(wide_index);
    cudaFree(cb); cudaFree(max_depth); cudaFree(cub_tmp);
}

#define LB(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return e__; } while (0)

cudaError_t lbvh_build_level(const LbvhLeaf* d_leaves, uint32_t n, uint32_t node_base, Node* d_nodes, Node16* d_nodes16,
                             LbvhScratch& w, LbvhLevelOut* d_out, cudaStream_t st) {
    if (n < 2u) return cudaErrorInvalidValue;
    size_t h;
    h = w.n_codes; LB(grow(&w.codes, &h, n)); h = w.n_codes; LB(grow(&w.codes_sorted, &h, n));
    h = w.n_codes; LB(grow(&w.order_in, &h, n)); h = w.n_codes; LB(grow(&w.order, &h, n));
    h = w.n_codes; LB(grow(&w.parent_l, &h, n)); h = w.n_codes; LB(grow(&w.left, &h, n));
    h = w.n_codes; LB(grow(&w.right, &h, n)); h = w.n_codes; LB(grow(&w.parent_i, &h, n));
    h = w.n_codes; LB(grow(&w.ibox, &h, n)); h = w.n_codes; LB(grow(&w.flag, &h, n));
    h = w.n_codes; LB(grow(&w.wide, &h, n)); h = w.n_codes; LB(grow(&w.wide_index, &h, n));
    if (w.n_codes < n) w.n_codes = n;
    if (!w.cb) LB(cudaMalloc(reinterpret_cast<void**>(&w.cb), 6 * sizeof(float)));
    if (!w.max_depth) LB(cudaMalloc(reinterpret_cast<void**>(&w.max_depth), sizeof(uint32_t)));
    size_t sort_bytes = 0, scan_bytes = 0;
    LB(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, w.codes, w.codes_sorted, w.order_in, w.order, (int)n, 0, 30, st));
    LB(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, w.wide, w.wide_index, (int)(n - 1u), st));
    const size_t tmp = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    if (w.cub_bytes < tmp) {
        if (w.cub_tmp) cudaFree(w.cub_tmp);
        w.cub_tmp = nullptr; w.cub_bytes = 0;
        LB(cudaMalloc(&w.cub_tmp, tmp));
        w.cub_bytes = tmp;
    }
    const int B = 128;
    const int g_leaf = (int)((n + B - 1) / B), g_int = (int)((n - 1 + B - 1) / B);
    LB(cudaMemsetAsync(w.max_depth, 0, sizeof(uint32_t), st));
    lbvh_bounds_kernel<<<1, 256, 0, st>>>(d_leaves, n, w.cb);
    lbvh_morton_kernel<<<g_leaf, B, 0, st>>>(d_leaves, n, w.cb, w.codes, w.order_in);
    size_t sb = w.cub_bytes;
    LB(cub::DeviceRadixSort::SortPairs(w.cub_tmp, sb, w.codes, w.codes_sorted, w.order_in, w.order, (int)n, 0, 30, st));
    lbvh_karras_kernel<<<g_int, B, 0, st>>>(w.codes_sorted, (int)n, w.left, w.right, w.parent_i, w.parent_l, w.flag);
    lbvh_fit_kernel<<<g_leaf, B, 0, st>>>(d_leaves, w.order, (int)n, w.left, w.right, w.parent_i, w.parent_l, w.ibox, w.flag);
    lbvh_depth_kernel<<<g_int, B, 0, st>>>(w.parent_i, (int)n, w.wide, w.max_depth);
    sb = w.cub_bytes;
    LB(cub::DeviceScan::ExclusiveSum(w.cub_tmp, sb, w.wide, w.wide_index, (int)(n - 1u), st));
    LbvhTree t;
    t.leaves = d_leaves; t.order = w.order; t.left = w.left; t.right = w.right; t.ibox = w.ibox;
    t.wide_index = w.wide_index; t.node_base = node_base;
    lbvh_emit_kernel<<<g_int, B, 0, st>>>(t, (int)n, w.wide, d_nodes, d_nodes16);
    lbvh_finish_kernel<<<1, 1, 0, st>>>(w.ibox, node_base, w.max_depth, d_out);
    return cudaGetLastError();
}

}  // namespace rtw
