// rtw_kernels.h — launch interface of rtw_kernels.cu. That file is compiled three times:
//   fast   : nvcc default (a*b+c contracts to FFMA)             -> launch_*_fast
//   strict : -fmad=false -DRTW_STRICT_FP (rustc-like arithmetic) -> launch_*_strict
//   count  : fast + -DRTW_COUNT_WORK (node/primitive test counters) -> launch_*_count
// selected per render call by RTW_FLAG_STRICT_FP / RTW_FLAG_COUNT_WORK.
#pragma once
#include <cuda_runtime.h>
#include "rtw_shade.h"

namespace rtw {

struct Counters {
    unsigned long long paths, rays, nan_paths;
    unsigned long long nodes, spheres, quads, instances, hits;   // only from the RTW_COUNT_WORK build
    // warp scheduling of wf_extend (RTW_COUNT_WORK build): iterations, and per step kind how often
    // it ran and with how many lanes in total. [0] node [1] prim [2] event [3] fetch
    unsigned long long sched_iters, sched_steps[4], sched_lanes[4];
    unsigned long long sched_cycles[5];   // warp cycles inside node / prim / event / fetch steps, and in the rest of the loop
};

// Work decomposition of the persistent-lane megakernel: an item is one 8x4 pixel tile x one chunk
// of consecutive samples (32 * chunk paths); warps pull items from a global counter and hand the
// item's paths to whichever lane is free.
struct MegaLaunch {
    uint32_t chunk;        // samples per item
    uint32_t n_chunks;
    uint32_t tiles_x, tiles_y;
    uint32_t n_items;
    uint32_t stack_words;  // per-thread traversal stack depth (stack_words(), rtw_trace.h)
    uint32_t refill_below; // leave the traversal loop to shade/refill when fewer lanes than this still traverse
    int grid, block;
    size_t smem_bytes;     // dynamic shared memory: stack_words * block * 4
};

// ---- wavefront pipeline (RTW_FLAG_WAVEFRONT): the paths in flight live in HBM as a pool of
// `capacity` path records (SoA), one fixed slot per shade thread. One iteration =
//   extend  persistent warps pull consecutive slots (chunk reservation + L1 prefetch), walk the BVH
//           for the live ones and store the closest hit in the slot;
//   shade   one thread per slot: miss / Hit / scatter, record rewritten in place. A finished path
//           adds its radiance to the frame and its thread starts the next camera sample of the
//           frame's tile-major sample order. Samples are handed out from a per-warp range of
//           `supply_chunk` indices (ballot ranks inside the warp; one atomic on the global cursor
//           per refill), so nothing has to be compacted while the supply lasts: every slot stays
//           live. A slot that finds the supply exhausted is marked dead.
struct WfQueue {
    float4* o;           // origin.xyz, time
    float4* d;           // direction.xyz, bits(bounce) (WF_DEAD: empty slot)
    uint2*  id;          // pixel index, sample index (Philox counter words 0/1)
    float4* thr;         // throughput.xyz, bits(next Philox block | depth_left << 19 | NaN-poison << 31)
};
// No radiance travels with a path: each segment's contribution (thr * emitted, thr * background) is
// added to the frame when it arises (rtw_kernels.cu, wf_shade_kernel).
#define WF_BLOCK_MASK 0x7FFFFu      // 19 bits of Philox block counter
#define WF_DEPTH_SHIFT 19
#define WF_DEPTH_MASK 0xFFFu        // depth_left <= 4095: deeper renders take the megakernel (rtw_api.cu)
#define WF_MAX_DEPTH 4095u
#define WF_DEAD 0xFFFFFFFFu
struct WfPool {
    WfQueue q;
    uint4*  hit;         // bits(t), ref, inner, 0   (extend -> shade)
    unsigned long long* supply;   // per shade warp: [next, end) of its reserved sample indices
    uint32_t* warp_live;          // per shade warp (32 slots): 0 once none of its slots can ever be live again (drain, below)
    uint32_t capacity;
};
struct WfCounters {
    uint32_t head;                   // extend's reservation cursor into the pool
    uint32_t live_mark;              // = `mark` of the last shade launch that left a live slot
    unsigned long long next_path;    // global cursor into the tile-major sample order
    unsigned long long total_paths;  // 32 * tiles * samples of this render (edge-tile padding included)
    unsigned long long iterations;
};
struct WfLaunch {
    uint32_t tiles_x, tiles_y;
    uint32_t stack_words;
    uint32_t fetch_min;              // extend fetches new rays when at least this many lanes are idle
    uint32_t chunk;                  // pool slots an extend warp reserves per atomic
    uint32_t supply_chunk;           // sample indices a shade warp reserves per atomic
    uint32_t sample_major;           // order of the sample supply: 0 = tile-major (all samples of a tile, then the next tile), 1 = sample-major
    uint32_t index32;                // tiles * samples < 2^32: 32-bit index arithmetic
    unsigned long long total_paths;  // = WfCounters::total_paths
    // extend's vote: node steps run while node lanes * 4 >= leaf lanes * prim_w and >= event lanes * event_w.
    // 4 / 4 is the plain majority. Where instance events are rare (C5: three instance leaves among 2 409)
    // the few lanes that wait for a leaf test or a level change should not wait for a node majority to
    // dissolve: 8 / 16 is +3.6 % there; where they are common (C4: two of eight leaves) any bias costs
    // 1-6 % (profiles/r1j_ab_vote_weights.txt). rtw_api.cu picks by the share of instance leaves.
    uint32_t prim_w, event_w;
    // Drain: once the frame's sample supply is exhausted (the host sees WfCounters::next_path pass the end, two batches
    // late) slots die for good, and for the last ~depth iterations both kernels used to scan a pool that is almost or
    // entirely dead: ~0.2 ms per iteration over 6 Mi slots, ~13 ms per frame, a fixed cost that does not shrink with
    // the share of the frame a GPU renders (the 1 -> 8 GPU efficiency loss of round 1). In drain mode a shade warp
    // whose `warp_live` word is 0 leaves after that one load, and extend reads 32 of those words at a time and skips
    // dead groups of 32 slots without touching their records (and reserves `chunk` = 2048 slots per atomic).
    uint32_t drain;
    uint32_t node_sure;              // lanes on a node from which the node step wins for certain: ceil(32 w / (4 + w)), w = max(prim_w, event_w)
    int extend_grid, shade_grid;
    size_t smem_bytes;
};
#define RTW_MAX_STACK_WORDS 160     // deepest traversal stack column any scene may ask for (scene_build.cpp / rtw_api.cu refuse more)
#define RTW_WF_BLOCK 128
#ifndef RTW_WF_EXTEND_MINB
#define RTW_WF_EXTEND_MINB 7      // min resident blocks per SM asked of the compiler for wf_extend (71 registers, no spills since the lane flags are
                                  // words and the stack pointer an address; 6 blocks / 80 registers until round 2: 7 blocks then cost -3.2 %, 8 now -1.3 %)
#endif
#ifndef RTW_WF_PENDING
#define RTW_WF_PENDING 0          // 1: wf_extend parks a popped leaf test and keeps walking nodes (see the kernel); A/B: leaf steps
                                  // run with 15.7 instead of 13.2 lanes, node steps unchanged (16.0 vs 15.6), +2 % node visits, -4 % overall
#endif
#ifndef RTW_WF_PREFETCH
#define RTW_WF_PREFETCH 0         // 1: the node step prefetches the record it will visit next into L1
#endif
#ifndef RTW_WF_NODE_BURST
#define RTW_WF_NODE_BURST 3       // node steps per vote while the node lanes keep their majority
#endif
#ifndef RTW_WF_WINDOW
#define RTW_WF_WINDOW 0           // 256 (or 128 / 512 / 1024): wf_extend takes the slots of a window of that many in the order camera rays, then
                                  // the direction octants of the bounced rays (A/B, see the kernel); 0: slot order
#endif
#ifndef RTW_WF_WINDOW_CS
#define RTW_WF_WINDOW_CS 0
#endif
#if RTW_WF_WINDOW > 256
#define RTW_WF_ORDER_T uint16_t
#else
#define RTW_WF_ORDER_T uint8_t
#endif
#ifndef RTW_WF_FLAG_T
#define RTW_WF_FLAG_T uint32_t     // wf_extend's `busy` / `more`: as bool the compiler packed them into bytes of one register and PRMT-ed them
                                  // apart every iteration; as words: 73 instead of 79 registers, 2 % fewer instructions, +1.2 % (C5), and with
                                  // that room 7 resident blocks: +3.4 % together (profiles/r2l_ab_u32_flags_7_blocks.txt)
#endif
#ifndef RTW_WF_REDUX
#define RTW_WF_REDUX 0            // 1: one REDUX.SUM of packed byte counters per loop iteration instead of a ballot + popc per lane state (A/B)
#endif
#ifndef RTW_WF_FASTVOTE
#define RTW_WF_FASTVOTE 1         // 1: skip the leaf / event ballots of wf_extend's vote when the node lanes win anyway (same decisions)
#endif
#ifndef RTW_WF_SHADE_SMEM_NOISE
#define RTW_WF_SHADE_SMEM_NOISE 0   // 1: wf_shade stages the Perlin tables in shared memory per block (A/B only, see the kernel)
#endif
#ifndef RTW_WF_SHADE_MINB
#define RTW_WF_SHADE_MINB 10      // min resident blocks per SM asked of the compiler for wf_shade
#endif
#define RTW_WF_SHADE_BLOCK 128

#define RTW_MEGA_BLOCK 128
#ifndef RTW_MEGA_MINB
#define RTW_MEGA_MINB 5        // resident blocks per SM the plain megakernel's register allocation aims at (96 registers)
#endif
#define RTW_SMEM_NOISES 2

#define RTW_DECLARE_VARIANT(SUFFIX)                                                                          \
    cudaError_t launch_mega_##SUFFIX(const Scene& sc, const RenderParams& P, float* d_sum, Counters* d_ctr,  \
                                     uint32_t* d_work, const MegaLaunch& L, cudaStream_t st);                \
    cudaError_t mega_occupancy_##SUFFIX(int* blocks_per_sm /* [2]: plain, light-biased */, size_t smem_bytes); \
    cudaError_t launch_wf_extend_##SUFFIX(const Scene& sc, const RenderParams& P, const WfPool& pool, WfCounters* d_wc,   \
                                          Counters* d_ctr, const WfLaunch& L, cudaStream_t st);                          \
    cudaError_t launch_wf_shade_##SUFFIX(const Scene& sc, const RenderParams& P, const WfPool& pool, WfCounters* d_wc,    \
                                         Counters* d_ctr, float* d_sum, uint32_t mark, int init, const WfLaunch& L,      \
                                         cudaStream_t st);                                                               \
    cudaError_t wf_extend_occupancy_##SUFFIX(int* blocks_per_sm, size_t smem_bytes);

RTW_DECLARE_VARIANT(fast)
RTW_DECLARE_VARIANT(strict)
RTW_DECLARE_VARIANT(count)    // fast + algorithmic work counters (RTW_FLAG_COUNT_WORK)

}  // namespace rtw
