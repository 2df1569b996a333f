// vkr_scene_kernels.cuh — the kernels of update_scene on the device and the record of a scene handed over with
// vkr_scene_create (included by vkr_api.cu between VkrDevice and submit_frame; bodies in vkr_scene.cuh).
//
// Per frame, on the upload stream, in front of prepare_draws_kernel / expand_chunks_kernel (which read what these write):
//   [scene_nodes_kernel x levels]   only after a transform changed: worldTransform of every node, a level of the tree per launch
//   scene_draws_kernel              one thread per draw: is_visible, MVP, light MVPs, order keys -> DObject[n], shadowMvp, keys
//   scene_rank_kernel               one thread per (pass, draw): its place in the pass's front-to-back order (counting rank);
//                                   above SCENE_RANK_MAX draws instead scene_bucket_count / _scan / _place_kernel: a stable
//                                   counting sort on SCENE_BUCKETS slices of the pass's key range
//   scene_scan_kernel               one CTA: dense primitive ids of the visible draws, the passes' first work-item slots, the counts
// The host's share of a frame no longer grows with the number of draws: it uploads nothing and launches a fixed list of kernels.
#pragma once
#include "vkr_scene.cuh"

__global__ void scene_nodes_kernel(const float* __restrict__ local, const int32_t* __restrict__ parent, float* world,
                                   const uint32_t* __restrict__ levelOrder, uint32_t first, uint32_t count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) scene_node_world(local, parent, world, levelOrder[first + i]);
}

__global__ void __launch_bounds__(128) scene_draws_kernel(const __grid_constant__ SceneFrame f, const SceneDraw* __restrict__ draws, const float* __restrict__ world,
                                                          const DObject* __restrict__ tmpl, DObject* __restrict__ objects, float* __restrict__ shadowMvp,
                                                          float* __restrict__ keys, uint32_t* __restrict__ visTris, uint32_t* keyRange) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < f.nDraws;
    if (live) scene_draw_body(f, draws[i], world, tmpl[i], i, objects[i], shadowMvp, keys, visTris);
    if (keyRange == nullptr) return;
    // many draws (bucket order): the range of every pass's keys over the visible draws, one pair of atomics per warp and pass
    const float inf = __uint_as_float(0x7F800000u);
    for (uint32_t p = 0; p <= f.shadowLayers; ++p) {
        const bool takes = live && (p > 0 || i < f.opaqueCount);
        const float k = takes ? keys[(size_t)p * f.nDraws + i] : inf;
        const uint32_t o = k == inf ? 0u : scene_ordered(k), no = k == inf ? 0u : ~scene_ordered(k);
        const uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, o), mn = __reduce_max_sync(0xFFFFFFFFu, no);
        if ((threadIdx.x & 31u) == 0u && mx != 0u) { atomicMax(keyRange + 2 * p, mx); atomicMax(keyRange + 2 * p + 1, mn); }
    }
}

// order[pass][rank] = draw.  The main pass (blockIdx.y = 0) orders the opaque prefix only; the shadow layers order every draw.
// The keys of 256 draws at a time go through shared memory; every thread counts the draws in front of its own (scene_rank).
__global__ void __launch_bounds__(256) scene_rank_kernel(const float* __restrict__ keys, uint32_t nDraws, uint32_t opaqueCount, uint32_t* __restrict__ order) {
    __shared__ float s_keys[256];
    const uint32_t p = blockIdx.y, count = p == 0 ? opaqueCount : nDraws;
    if (blockIdx.x * 256u >= count) return;
    const float* __restrict__ k = keys + (size_t)p * nDraws;
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    const float ki = i < count ? k[i] : 0.0f;
    uint32_t rank = 0;
    for (uint32_t base = 0; base < count; base += 256u) {
        __syncthreads();
        s_keys[threadIdx.x] = base + threadIdx.x < count ? k[base + threadIdx.x] : 0.0f;
        __syncthreads();
        const uint32_t m = min(256u, count - base);
        if (i < count)
            for (uint32_t j = 0; j < m; ++j) rank += scene_key_less(s_keys[j], base + j, ki, i) ? 1u : 0u;
    }
    if (i < count) order[(size_t)p * nDraws + rank] = i;
}

// ---- many draws: the bucket order (vkr_scene.cuh).  A CTA owns SCENE_BLOCK consecutive draws of a pass.
// hist[pass][bucket][block] = draws of the block in the bucket; bucketOf[pass][draw] = its bucket.
__global__ void __launch_bounds__(256) scene_bucket_count_kernel(const float* __restrict__ keys, uint32_t nDraws, uint32_t opaqueCount, uint32_t nBlocks,
                                                                  const uint32_t* __restrict__ keyRange, uint16_t* __restrict__ bucketOf, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[SCENE_BUCKETS];
    const uint32_t p = blockIdx.y, count = p == 0 ? opaqueCount : nDraws, block = blockIdx.x;
    for (uint32_t k = threadIdx.x; k < SCENE_BUCKETS; k += 256u) s_hist[k] = 0u;
    __syncthreads();
    const uint32_t rmax = keyRange[2 * p], rneg = keyRange[2 * p + 1];
    for (uint32_t e = threadIdx.x; e < SCENE_BLOCK; e += 256u) {
        const uint32_t i = block * SCENE_BLOCK + e;
        if (i >= count) break;
        const uint32_t b = scene_bucket(keys[(size_t)p * nDraws + i], rmax, rneg);
        bucketOf[(size_t)p * nDraws + i] = (uint16_t)b;
        atomicAdd(&s_hist[b], 1u);
    }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < SCENE_BUCKETS; k += 256u) hist[((size_t)p * SCENE_BUCKETS + k) * nBlocks + block] = s_hist[k];
}
// order[pass][hist-prefix of (bucket, block) + draws of the same bucket earlier in the block] = draw  (stable)
__global__ void __launch_bounds__(256) scene_bucket_place_kernel(uint32_t nDraws, uint32_t opaqueCount, uint32_t nBlocks, const uint16_t* __restrict__ bucketOf,
                                                                  const uint32_t* __restrict__ histPrefix, uint32_t* __restrict__ order) {
    __shared__ uint16_t s_b[SCENE_BLOCK];
    const uint32_t p = blockIdx.y, count = p == 0 ? opaqueCount : nDraws, block = blockIdx.x;
    const uint32_t base = block * SCENE_BLOCK, m = base < count ? min(SCENE_BLOCK, count - base) : 0u;
    for (uint32_t e = threadIdx.x; e < m; e += 256u) s_b[e] = bucketOf[(size_t)p * nDraws + base + e];
    __syncthreads();
    for (uint32_t e = threadIdx.x; e < m; e += 256u) {
        const uint32_t b = s_b[e];
        uint32_t before = 0;
        for (uint32_t j = 0; j < e; ++j) before += s_b[j] == b ? 1u : 0u;
        order[(size_t)p * nDraws + histPrefix[((size_t)p * SCENE_BUCKETS + b) * nBlocks + block] + before] = base + e;
    }
}

// Exclusive prefix sums of n values by ONE CTA of 1024 threads; thread t owns a contiguous run of elements.  put(k, prefix).
template <class Get, class Put>
__device__ __forceinline__ uint32_t scene_block_scan(uint32_t n, Get get, Put put, uint32_t* s_warp /* [33] */) {
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5, per = (n + 1023u) / 1024u;
    const uint32_t b = min(n, t * per), e = min(n, b + per);
    uint32_t sum = 0;
    for (uint32_t k = b; k < e; ++k) sum += get(k);
    uint32_t incl = sum;
    for (uint32_t o = 1; o < 32u; o <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= o) incl += v; }
    if (lane == 31u) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_warp[lane];
        uint32_t wi = w;
        for (uint32_t o = 1; o < 32u; o <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, wi, o); if (lane >= o) wi += v; }
        s_warp[lane] = wi - w;
        if (lane == 31u) s_warp[32] = wi;
    }
    __syncthreads();
    uint32_t run = s_warp[warp] + incl - sum;
    for (uint32_t k = b; k < e; ++k) { const uint32_t v = get(k); put(k, run); run += v; }
    const uint32_t total = s_warp[32];
    __syncthreads();                 // s_warp is free for the next scan
    return total;
}

// hist[pass][bucket][block] -> its exclusive prefix over (bucket, block), one CTA per pass
__global__ void __launch_bounds__(1024) scene_bucket_scan_kernel(uint32_t* hist, uint32_t nBlocks) {
    __shared__ uint32_t s_warp[33];
    uint32_t* h = hist + (size_t)blockIdx.x * SCENE_BUCKETS * nBlocks;
    scene_block_scan(SCENE_BUCKETS * nBlocks, [&](uint32_t k) { return h[k]; }, [&](uint32_t k, uint32_t pre) { h[k] = pre; }, s_warp);
}

// counts[0] = draws is_visible kept, counts[1] = their triangles (pinned host words: EngineStats::drawcall_count / triangle_count,
// src/vk_engine_run.cpp:309-310).  slotBase[pass][draw] = first work-item slot of the draw in the pass (0xFFFFFFFF: not part of it).
__global__ void __launch_bounds__(1024) scene_scan_kernel(const __grid_constant__ SceneFrame f, const SceneDraw* __restrict__ draws, DObject* __restrict__ objects,
                                                           const uint32_t* __restrict__ visTris, const uint32_t* __restrict__ order,
                                                           uint32_t* __restrict__ slotBase, uint32_t* counts) {
    __shared__ uint32_t s_warp[33];
    const uint32_t n = f.nDraws;
    const uint32_t tris = scene_block_scan(n, [&](uint32_t k) { return visTris[k]; }, [&](uint32_t k, uint32_t pre) { objects[k].primBase = pre; }, s_warp);
    const uint32_t shown = scene_block_scan(n, [&](uint32_t k) { return (objects[k].flags & DRAW_HIDDEN) ? 0u : 1u; }, [](uint32_t, uint32_t) {}, s_warp);
    if (threadIdx.x == 0 && counts != nullptr) { counts[0] = shown; counts[1] = tris; }
    for (uint32_t p = 0; p <= f.shadowLayers; ++p) {
        const uint32_t count = p == 0 ? f.opaqueCount : n;
        const uint32_t* __restrict__ ord = order + (size_t)p * n;
        uint32_t* __restrict__ sb = slotBase + (size_t)p * n;
        for (uint32_t k = count + threadIdx.x; k < n; k += blockDim.x) sb[k] = 0xFFFFFFFFu;      // Transparent draws: not in the main pass's list
        scene_block_scan(count, [&](uint32_t k) { return draws[ord[k]].nClusters; }, [&](uint32_t k, uint32_t pre) { sb[ord[k]] = pre; }, s_warp);
    }
}

// A scene graph handed over with vkr_scene_create: static per-draw data on the device, the host copies it was built from.
struct SceneRec {
    uint32_t nNodes = 0, nDraws = 0, opaqueCount = 0;
    uint32_t nClusters = 0, opaqueClusters = 0, nSeqSlots = 0, totalTris = 0;
    bool anyCutout = false;
    bool worldDirty = true;                 // some localTransform changed since the world matrices were computed
    uint64_t flagsGen = ~0ull;              // VkrDevice::bindGen the draws' flags were derived at (bindless slots decide bit1 / bit2)
    std::vector<uint32_t> levelFirst;       // levelFirst[l] .. levelFirst[l+1]: the nodes of depth l in dLevelOrder
    float* dLocal = nullptr; int32_t* dParent = nullptr; float* dWorld = nullptr; uint32_t* dLevelOrder = nullptr;
    SceneDraw* dDraws = nullptr; DObject* dTmpl = nullptr;
    uint32_t* dClusterBase = nullptr; uint32_t* dSeqSlotBase = nullptr;      // nDraws + 1 each: static, every draw keeps its slots
    float* dKeys = nullptr; uint32_t* dOrder = nullptr; uint32_t* dVisTris = nullptr;   // per-frame scratch (upload stream only)
    // more than SCENE_RANK_MAX draws: the bucket order's scratch (key ranges, per-draw bucket, per-block histograms)
    uint32_t* dKeyRange = nullptr; uint16_t* dBucketOf = nullptr; uint32_t* dHist = nullptr; uint32_t nBlocks = 0;
    std::vector<VkrSceneSurface> surfaces;  // draw order: the traversal's, MainColor first (stable)
    std::vector<DObject> tmplHost;
};
