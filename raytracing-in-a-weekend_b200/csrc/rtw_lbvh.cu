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

// bottom-up fit: every leaf walks up; the second thread to reach a node owns its subtree from then on: it restructures
// the node's treelet, computes its box, its subtree cost and its collapse table (lbvh_fit_node). What other threads of
// the launch have written (links, boxes, costs, tables) is read with ld.global.cg behind the fence / atomic handshake;
// a restructuring only re-wires nodes inside the owned subtree, so every node still sees exactly two arrivals.
__global__ void lbvh_fit_kernel(LbvhBuild b, int n, uint32_t* flag) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int cur = b.parent_l[p];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&flag[cur], 1u) == 0u) return;
        __threadfence();
        lbvh_fit_node(b, cur);
        cur = b.parent_i[cur];
    }
}

// Breadth-first collapse by ONE block (a level of a scene is a few thousand leaves; the loop is O(n / 1024) steps of
// lbvh_pick + lbvh_emit per thread for any n). queue[k] = the binary node behind wide node k; the children of a
// tree level's wide nodes are appended in thread order through a block scan, so the node order is the emulator's
// sequential one: the binary root is wide node 0 and every tree level is contiguous (the top of the tree stays hot).
constexpr int LBVH_COLLAPSE_BLOCK = 1024;
__global__ void __launch_bounds__(LBVH_COLLAPSE_BLOCK)
lbvh_collapse_kernel(LbvhTree t, int* queue, Node* nodes, Node16* nodes16, LbvhLevelOut* out) {
    typedef cub::BlockScan<uint32_t, LBVH_COLLAPSE_BLOCK> Scan;
    __shared__ typename Scan::TempStorage tmp;
    uint32_t head = 0u, tail = 1u, depth = 0u;           // the same in every thread
    if (threadIdx.x == 0) queue[0] = 0;
    __syncthreads();
    while (head < tail) {
        ++depth;
        const uint32_t level_end = tail;
        for (uint32_t base = head; base < level_end; base += LBVH_COLLAPSE_BLOCK) {
            const uint32_t me = base + threadIdx.x;
            int cand[4]; int nc = 0; uint32_t k = 0u;
            if (me < level_end) {
                nc = lbvh_pick(t, queue[me], cand);
                for (int j = 0; j < nc; ++j) k += cand[j] >= 0 ? 1u : 0u;
            }
            uint32_t off, total;
            Scan(tmp).ExclusiveSum(k, off, total);
            if (me < level_end) {
                uint32_t pos[4] = {0u, 0u, 0u, 0u};
                uint32_t w = tail + off;
                for (int j = 0; j < nc; ++j) if (cand[j] >= 0) { pos[j] = w; queue[w] = cand[j]; ++w; }
                lbvh_emit(t, cand, nc, pos, me, nodes, nodes16);
            }
            tail += total;
            __syncthreads();                             // the scan's storage is reused; queue entries become visible
        }
        head = level_end;
    }
    if (threadIdx.x == 0) {
        for (int a = 0; a < 3; ++a) { out->lo[a] = t.ibox[0].lo[a]; out->hi[a] = t.ibox[0].hi[a]; }
        out->root = make_ref(KIND_NODE, t.node_base);    // the binary root is the level's first wide node
        out->max_depth = depth;
    }
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
    cudaFree(parent_i); cudaFree(parent_l); cudaFree(ibox); cudaFree(flag); cudaFree(queue); cudaFree(dp_cut); cudaFree(dp_choice); cudaFree(sah_cost);
    cudaFree(cb); cudaFree(cub_tmp);
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
    h = w.n_codes; LB(grow(&w.queue, &h, n));
    h = w.n_codes; LB(grow(&w.dp_choice, &h, n));
    h = w.n_codes; LB(grow(&w.sah_cost, &h, n));
    h = 4 * w.n_codes; LB(grow(&w.dp_cut, &h, 4 * (size_t)n));
    if (w.n_codes < n) w.n_codes = n;
    if (!w.cb) LB(cudaMalloc(reinterpret_cast<void**>(&w.cb), 6 * sizeof(float)));
    size_t tmp = 0;
    LB(cub::DeviceRadixSort::SortPairs(nullptr, tmp, w.codes, w.codes_sorted, w.order_in, w.order, (int)n, 0, 30, st));
    if (w.cub_bytes < tmp) {
        if (w.cub_tmp) cudaFree(w.cub_tmp);
        w.cub_tmp = nullptr; w.cub_bytes = 0;
        LB(cudaMalloc(&w.cub_tmp, tmp));
        w.cub_bytes = tmp;
    }
    const int B = 128;
    const int g_leaf = (int)((n + B - 1) / B), g_int = (int)((n - 1 + B - 1) / B);
    lbvh_bounds_kernel<<<1, 256, 0, st>>>(d_leaves, n, w.cb);
    lbvh_morton_kernel<<<g_leaf, B, 0, st>>>(d_leaves, n, w.cb, w.codes, w.order_in);
    size_t sb = w.cub_bytes;
    LB(cub::DeviceRadixSort::SortPairs(w.cub_tmp, sb, w.codes, w.codes_sorted, w.order_in, w.order, (int)n, 0, 30, st));
    lbvh_karras_kernel<<<g_int, B, 0, st>>>(w.codes_sorted, (int)n, w.left, w.right, w.parent_i, w.parent_l, w.flag);
    LbvhBuild b;
    b.leaves = d_leaves; b.order = w.order; b.left = w.left; b.right = w.right; b.parent_i = w.parent_i; b.parent_l = w.parent_l;
    b.ibox = w.ibox; b.dp_cut = w.dp_cut; b.dp_choice = w.dp_choice; b.sah_cost = w.sah_cost;
    for (int pass = 0; pass < LBVH_FIT_PASSES; ++pass) {
        if (pass > 0) LB(cudaMemsetAsync(w.flag, 0, (size_t)(n - 1u) * sizeof(uint32_t), st));
        lbvh_fit_kernel<<<g_leaf, B, 0, st>>>(b, (int)n, w.flag);
    }
    LbvhTree t;
    t.leaves = d_leaves; t.order = w.order; t.left = w.left; t.right = w.right; t.ibox = w.ibox; t.dp_choice = w.dp_choice; t.node_base = node_base;
    lbvh_collapse_kernel<<<1, LBVH_COLLAPSE_BLOCK, 0, st>>>(t, w.queue, d_nodes, d_nodes16, d_out);
    return cudaGetLastError();
}

}  // namespace rtw
