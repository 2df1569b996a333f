// vkr_api.cu — the C-ABI (include/vkr.h) and the kernels behind it.  sm_100a only.
//
// Frame = [upload draw list] -> [shadow: clear, geometry, large-triangle raster] x layers
//       -> [main: clear visibility buffer, geometry, large-triangle raster] -> [shade+resolve].
// All work is enqueued on one stream; vkr_draw never blocks on the device.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <nccl.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "vkr_internal.h"
#include "vkr_raster.cuh"
#include "vkr_shade.cuh"
#include "vkr_frameprep.h"

using namespace vkr;

#include "vkr_scene_kernels.cuh"      // update_scene on the device: kernels + SceneRec

// ======================================================================= kernels

// ---- fills / format helpers
__global__ void fill_u64_kernel(uint64_t* __restrict__ dst, size_t n, uint64_t v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    ulonglong2* d2 = reinterpret_cast<ulonglong2*>(dst);
    const size_t n2 = n >> 1;
    for (size_t k = i; k < n2; k += stride) d2[k] = make_ulonglong2(v, v);
    if (i == 0 && (n & 1)) dst[n - 1] = v;
}
__global__ void fill_u32_kernel(uint32_t* __restrict__ dst, size_t n, uint32_t v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    const size_t n4 = n >> 2;
    for (size_t k = i; k < n4; k += stride) d4[k] = make_uint4(v, v, v, v);
    if (i == 0) for (size_t k = n4 << 2; k < n; ++k) dst[k] = v;
}
// VKR_TARGET_RGB8: four RGBA8 pixels in (16 B), twelve bytes out; the tail of an image whose pixel count is not a multiple
// of four is packed bytewise by the first thread
__global__ void pack_rgb8_kernel(const uint32_t* __restrict__ rgba, uint32_t* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t n4 = n >> 2;
    if (i < n4) {
        const uint4 p = reinterpret_cast<const uint4*>(rgba)[i];
        const uint32_t a = p.x & 0xFFFFFFu, b = p.y & 0xFFFFFFu, c = p.z & 0xFFFFFFu, d = p.w & 0xFFFFFFu;
        out[3 * i] = a | (b << 24);
        out[3 * i + 1] = (b >> 8) | (c << 16);
        out[3 * i + 2] = (c >> 16) | (d << 8);
    }
    if (i == 0) {
        uint8_t* o = reinterpret_cast<uint8_t*>(out);
        for (size_t k = n4 << 2; k < n; ++k) { const uint32_t v = rgba[k]; o[3 * k] = (uint8_t)v; o[3 * k + 1] = (uint8_t)(v >> 8); o[3 * k + 2] = (uint8_t)(v >> 16); }
    }
}
// Cube-map texels decoded once, at upload, to the four floats the shading kernel's fetch computes from them (sRGB table
// lookups / 255ths): one 16-byte load per texel afterwards instead of a 4-byte load, three table lookups and four conversions.
__global__ void decode_texels_kernel(const uchar4* __restrict__ src, float4* __restrict__ dst, size_t n, const float* __restrict__ lut, int srgb) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uchar4 p = src[i];
    dst[i] = srgb ? make_float4(lut[p.x], lut[p.y], lut[p.z], p.w * (1.0f / 255.0f))
                  : make_float4(p.x * (1.0f / 255.0f), p.y * (1.0f / 255.0f), p.z * (1.0f / 255.0f), p.w * (1.0f / 255.0f));
}
__global__ void extract_vis_kernel(const uint64_t* __restrict__ vis, uint32_t* __restrict__ out, size_t n, int what,
                                   const uint2* __restrict__ chunkObj, const DObject* __restrict__ objects) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t k = vis[i];
    if (what == 0) { out[i] = (uint32_t)(k >> 32); return; }                     // depth bits
    uint32_t lo = (uint32_t)k;
    if (lo == VIS_EMPTY_LO) { out[i] = 0xFFFFFFFFu; return; }
    const uint32_t seq = 0xFFFFFFFEu - lo;                                        // slot*256 + triangle within the slot
    const uint2 ch = chunkObj[seq >> 8];
    out[i] = objects[ch.x].primBase + ch.y + (seq & 255u);                        // dense submission-order id
}

// ---- sort-first exchange: ranks own row bands of different heights (NCCL has no all-gather-v), so the exchange
// is one grouped send/recv: every rank sends its band to every peer and receives theirs in place.
static ncclResult_t exchange_bands(const uint32_t* src, uint32_t* dst, const std::vector<int32_t>& band, int lo, int hi, size_t rowTexels,
                                   int rank, int nranks, ncclComm_t comm, cudaStream_t stream) {
    auto clampRow = [&](int32_t y) { return std::min(std::max(y, lo), hi); };
    ncclResult_t rc = ncclGroupStart();
    const int32_t m0 = clampRow(band[rank]), m1 = clampRow(band[rank + 1]);
    for (int peer = 0; peer < nranks && rc == ncclSuccess; ++peer) {
        const int32_t p0 = clampRow(band[peer]), p1 = clampRow(band[peer + 1]);
        if (peer == rank && src == dst) continue;      // in place: the own band is already there
        if (m1 > m0) rc = ncclSend(src + (size_t)m0 * rowTexels, (size_t)(m1 - m0) * rowTexels, ncclUint32, peer, comm, stream);
        if (rc == ncclSuccess && p1 > p0) rc = ncclRecv(dst + (size_t)p0 * rowTexels, (size_t)(p1 - p0) * rowTexels, ncclUint32, peer, comm, stream);
    }
    if (rc == ncclSuccess) rc = ncclGroupEnd();
    return rc;
}

// ---- sort-first exchange over PEER MEMORY (NVLink / NVSwitch; one process per GPU, buffers shared through CUDA IPC):
// instead of two NCCL groups per frame, the ranks store their rows straight into their peers' shadow maps (a copy kernel) and
// into rank 0's presented image (the shading kernel's own stores), and tell each other with flags.  Every flag is a frame
// number that only grows, lives in the memory of the rank that WAITS on it and is written by a peer with a system-scope
// release store after a system-scope fence; waiters spin with acquire loads, bounded by a timeout that raises an error word in
// pinned host memory instead of hanging the device.
//   shadowCleared[s][p] = f+1 : rank p has cleared the shadow map of frame slot s (= f mod 2) for frame f: rows may be stored
//                              into it.  Normally raised a frame AHEAD (the clear of frame f+1's map is issued in front of frame f's
//                              main pass), so that nobody ever waits for it
//   shadowReady[p]   = f+1 : rank p's rows of frame f have landed in my shadow map
//   frameReady[p]    = f+1 : rank p's rows of frame f have landed in my (rank 0's) presented image
//   gatheredFree     = f+1 : rank 0 has finished with the image that frame f overwrites
static const int MAX_RANKS = 16;
struct ShardFlags {
    uint32_t shadowCleared[2][MAX_RANKS], shadowReady[MAX_RANKS], frameReady[MAX_RANKS];
    uint32_t gatheredFree, pad[15];
};
static_assert(VKR_FRAME_OVERLAP == 2, "ShardFlags::shadowCleared has one row per frame slot");
struct PeerList { uint32_t* p[MAX_RANKS]; int n; };
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
// Everything the stream has done before this kernel is visible to a peer that reads the new flag value.
__global__ void shard_signal_kernel(const __grid_constant__ PeerList dst, uint32_t value) {
    if ((int)threadIdx.x < dst.n) { __threadfence_system(); st_release_sys(dst.p[threadIdx.x], value); }
}
// Lane i waits until flags[i] >= value, for every i in `mask`.  ~4 s of spinning raises *err (pinned host memory) and goes on.
__global__ void shard_wait_kernel(const uint32_t* flags, uint32_t mask, uint32_t value, uint32_t* err, uint32_t code) {
    const uint32_t i = threadIdx.x;
    if (i < (uint32_t)MAX_RANKS && ((mask >> i) & 1u)) {
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(flags + i) - value) < 0) {
            if (clock64() - t0 > 8000000000ll) { *err = code | (i << 8); break; }
            __nanosleep(64);
        }
    }
}
__device__ __forceinline__ void spin_until(const uint32_t* flag, uint32_t value, uint32_t* err, uint32_t code) {
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flag) - value) < 0) {
        if (clock64() - t0 > 8000000000ll) { *err = code; break; }
        __nanosleep(64);
    }
}
// Tell the peers (dst), then wait for theirs (flags[i] >= value for i in mask): the two ends of one exchange in one launch.
__global__ void shard_signal_wait_kernel(const __grid_constant__ PeerList dst, const uint32_t* flags, uint32_t mask, uint32_t value, uint32_t* err, uint32_t code) {
    if ((int)threadIdx.x < dst.n) { __threadfence_system(); st_release_sys(dst.p[threadIdx.x], value); }
    const uint32_t i = threadIdx.x;
    if (i < (uint32_t)MAX_RANKS && ((mask >> i) & 1u)) spin_until(flags + i, value, err, code | (i << 8));
}
// My rows of one shadow layer into the same rows of every peer's map (texels [first, first + count) of the layer), once every
// peer has cleared its map for this frame (waitFlags[i] >= value for i in mask; each CTA checks before its first store — the
// peers' clears do not wait for anything on this GPU).
__global__ void shard_push_rows_kernel(const uint32_t* __restrict__ src, const __grid_constant__ PeerList dst, size_t first, size_t count,
                                       const uint32_t* waitFlags, uint32_t mask, uint32_t value, uint32_t* err) {
    if (threadIdx.x < (unsigned)MAX_RANKS && ((mask >> threadIdx.x) & 1u)) spin_until(waitFlags + threadIdx.x, value, err, 1u | (threadIdx.x << 8));
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x, tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (((first | count) & 3u) == 0u) {
        const uint4* s4 = reinterpret_cast<const uint4*>(src + first);
        for (size_t i = tid; i < (count >> 2); i += stride) {
            const uint4 v = __ldcg(s4 + i);                    // written by this frame's raster atomics: read at L2
            for (int k = 0; k < dst.n; ++k) reinterpret_cast<uint4*>(dst.p[k] + first)[i] = v;
        }
    } else {
        for (size_t i = tid; i < count; i += stride) {
            const uint32_t v = __ldcg(src + first + i);
            for (int k = 0; k < dst.n; ++k) dst.p[k][first + i] = v;
        }
    }
}

// ---- upload_mesh: AoS Vertex (48 B) -> three float4 streams
__global__ void repack_vertices_kernel(const VkrVertex* __restrict__ v, size_t n, float4* __restrict__ pos,
                                       float4* __restrict__ attrA, float4* __restrict__ attrB) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4* src = reinterpret_cast<const float4*>(v + i);   // 48 B = 3 x 16 B, 16-aligned
    float4 a = src[0], b = src[1], c = src[2];
    // a = (px,py,pz,uv_x)  b = (nx,ny,nz,uv_y)  c = color
    pos[i] = make_float4(a.x, a.y, a.z, 1.0f);
    attrA[i] = make_float4(b.x, b.y, b.z, a.w);
    attrB[i] = make_float4(c.x, c.y, c.z, b.w);
}

// ---- generate_mipmaps (src/vk_images.cpp:64-157): LINEAR blit level i -> i+1
__device__ __forceinline__ uint32_t linear_to_srgb8_dev(float l) {
    if (!(l > 0.0f)) return 0u;
    if (l >= 1.0f) return 255u;
    float s = l <= 0.0031308f ? l * 12.92f : 1.055f * powf(l, 1.0f / 2.4f) - 0.055f;
    return unorm8(s);
}
__global__ void mip_blit_kernel(const uchar4* __restrict__ src, int sw, int sh, uchar4* __restrict__ dst, int dw, int dh,
                                int srgb, const float* __restrict__ lut) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= dw || y >= dh) return;
    float scx = (float)sw / (float)dw, scy = (float)sh / (float)dh;
    float u = __fsub_rn(__fmul_rn((float)x + 0.5f, scx), 0.5f), v = __fsub_rn(__fmul_rn((float)y + 0.5f, scy), 0.5f);
    float fu = floorf(u), fv = floorf(v);
    float a = u - fu, b = v - fv;
    int x0 = min(max((int)fu, 0), sw - 1), x1 = min(max((int)fu + 1, 0), sw - 1);
    int y0 = min(max((int)fv, 0), sh - 1), y1 = min(max((int)fv + 1, 0), sh - 1);
    uchar4 q[4] = {src[(size_t)y0 * sw + x0], src[(size_t)y0 * sw + x1], src[(size_t)y1 * sw + x0], src[(size_t)y1 * sw + x1]};
    float wgt[4] = {__fmul_rn(1.0f - a, 1.0f - b), __fmul_rn(a, 1.0f - b), __fmul_rn(1.0f - a, b), __fmul_rn(a, b)};
    float c[4] = {0, 0, 0, 0};
    for (int k = 0; k < 4; ++k) {
        float r, g, bl;
        if (srgb) { r = lut[q[k].x]; g = lut[q[k].y]; bl = lut[q[k].z]; }
        else { r = __fdiv_rn((float)q[k].x, 255.0f); g = __fdiv_rn((float)q[k].y, 255.0f); bl = __fdiv_rn((float)q[k].z, 255.0f); }
        float al = __fdiv_rn((float)q[k].w, 255.0f);
        c[0] = __fadd_rn(c[0], __fmul_rn(r, wgt[k])); c[1] = __fadd_rn(c[1], __fmul_rn(g, wgt[k]));
        c[2] = __fadd_rn(c[2], __fmul_rn(bl, wgt[k])); c[3] = __fadd_rn(c[3], __fmul_rn(al, wgt[k]));
    }
    uchar4 o;
    if (srgb) { o.x = linear_to_srgb8_dev(c[0]); o.y = linear_to_srgb8_dev(c[1]); o.z = linear_to_srgb8_dev(c[2]); }
    else { o.x = unorm8(c[0]); o.y = unorm8(c[1]); o.z = unorm8(c[2]); }
    o.w = unorm8(c[3]);
    dst[(size_t)y * dw + x] = o;
}

// ---- per-frame tables, written on the device from the per-draw prefixes (one CTA per draw):
// for every pass the work items in processing order (geometry warp slot -> cluster, front to back by draw), each with the
// outcome of its cluster test (one thread per cluster projects the bounding box's eight corners through the pass's matrix:
// frustum planes, pixel rectangle, nearest depth — CullRec); one more list in submission order with the main pass's matrix
// (the transparent pass); the 48-byte cluster records the cut-out kernel walks; and the (draw, first triangle) pairs the
// shading pass looks up through seq >> 8.
// Per draw (one thread each): the eye points of its matrices for the facing cull of whole clusters ([pass][draw][8]: pass 0 the
// main matrix, 1.. the shadow layers) and its sequence slots.
__global__ void prepare_draws_kernel(const DObject* __restrict__ objects, const uint32_t* __restrict__ seqSlotBase, const float* __restrict__ shadowMvp,
                                     uint32_t nobj, uint32_t passes, float* __restrict__ eyes, uint2* __restrict__ chunkObj) {
    const uint32_t obj = blockIdx.x * blockDim.x + threadIdx.x;
    if (obj >= nobj) return;
    for (uint32_t p = 0; p < passes; ++p)
        eye_point(p == 0 ? objects[obj].mvp : shadowMvp + ((size_t)(p - 1) * nobj + obj) * 16, eyes + ((size_t)p * nobj + obj) * 8);
    const uint32_t sbase = seqSlotBase[obj], ns = seqSlotBase[obj + 1] - sbase;
    for (uint32_t k = 0; k < ns; ++k) chunkObj[sbase + k] = make_uint2(obj, k * SEQ_SLOT);
}
// Per cluster of the frame (one thread each; the draw is found by bisection of the cluster prefix).
__global__ void __launch_bounds__(128) expand_chunks_kernel(const DObject* __restrict__ objects, const uint32_t* __restrict__ clusterBase,
                                     const uint32_t* __restrict__ seqSlotBase, const uint32_t* __restrict__ slotBase,
                                     const float* __restrict__ shadowMvp, const float* __restrict__ eyes, uint32_t nobj, uint32_t passes, uint32_t listStride,
                                     const __grid_constant__ RasterTarget rtMain, const __grid_constant__ RasterTarget rtShadow,
                                     DChunk* __restrict__ chunks, WorkItem* __restrict__ work) {
    const uint32_t ci = blockIdx.x * blockDim.x + threadIdx.x;
    if (ci >= listStride) return;
    uint32_t lo = 0, hi = nobj;                       // the last draw with clusterBase[draw] <= ci (draws without clusters share a base: skipped)
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(clusterBase + mid) <= ci) lo = mid; else hi = mid; }
    const uint32_t obj = lo, c = ci - __ldg(clusterBase + obj), sbase = __ldg(seqSlotBase + obj);
    const DObject& o = objects[obj];
    chunks[ci] = make_chunk(o, obj, c, sbase);
    const float4 bmin = o.clBounds[2 * c], bmax = o.clBounds[2 * c + 1];
    const ClusterCone cone{o.clCone[3 * c], o.clCone[3 * c + 1], o.clCone[3 * c + 2]};
    // a draw the device's is_visible rejected (vkr_draw_scene, DRAW_HIDDEN) keeps its slots in every list, as culled clusters
    const bool hidden = (o.flags & DRAW_HIDDEN) != 0u;
    const CullRec none{0, 0, 0, 0, 0u, CLUSTER_CULLED};
    const CullRec mainRec = hidden ? none : make_cull_rec(o.mvp, bmin, bmax, rtMain, &cone, eyes + (size_t)obj * 8);
    work[(size_t)passes * listStride + ci] = make_work_item(o, obj, c, sbase, mainRec);      // submission order
    for (uint32_t p = 0; p < passes; ++p) {
        const uint32_t sb = slotBase[(size_t)p * nobj + obj];
        if (sb == 0xFFFFFFFFu) continue;
        work[(size_t)p * listStride + sb + c] = make_work_item(o, obj, c, sbase, (p == 0 || hidden) ? mainRec :
            make_cull_rec(shadowMvp + ((size_t)(p - 1) * nobj + obj) * 16, bmin, bmax, rtShadow, &cone, eyes + ((size_t)p * nobj + obj) * 8));
    }
}

// ---- geometry: one WARP per cluster (<= 128 triangles, <= 128 distinct vertices of one draw); the warps of a
// CTA never synchronise with each other, so every warp of the SM is an independent latency-hiding stream.
// The warp reads its 32-byte work item, and unless expand_chunks_kernel's box test already dropped the cluster, one lane
// starts ONE bulk copy (cp.async.bulk: the TMA unit, completion on the warp's own mbarrier) of the cluster's block — packed
// triangles and vertex positions, 1-3 KB — into the warp's shared memory.  While it flies the warp walks the cluster's
// pixel rectangle in the live depth / visibility buffer (phase 0: hidden -> dropped).  Phase 1 transforms each DISTINCT
// vertex once, in place, into snapped window coordinates; phase 2 runs the triangles, 32 at a time, on those.
// PASS: 0 main (visibility buffer), 1 shadow (depth only), 2 transparent (shade + additive blend, no depth write).
enum { PASS_MAIN = 0, PASS_SHADOW = 1, PASS_TRANSPARENT = 2 };

// VKR_STAGE_TRIS 1: the whole block (packed triangles + positions) is staged; 0 (default): only the positions, the packed
// triangles are read from L2 one iteration ahead (1 KB less shared memory per warp: 12 instead of 10 CTAs per SM, and it is
// resident warps, not instructions, that this kernel is short of — profiles/r02d_ab.md)
#ifndef VKR_STAGE_TRIS
#define VKR_STAGE_TRIS 0
#endif
template <int WARPS>
struct alignas(16) GeomSmem {              // the shared memory of one cluster's warp (WARPS = 1) or CTA (WARPS = 4): one base register, constant offsets
#if VKR_STAGE_TRIS
    uint2 tris[CL_TRIS];                   // bulk-copy destination, part 1 ...
#endif
    uint4 vtx[CL_VERTS];                   // ... part 2: positions, overwritten in place by the transformed vertices (CachedVtx)
    ParkedTri park[32 * WARPS];
#if defined(VKR_RASTER_CANDIDATES)
    uint2 map[32];
#endif
    float mvp[16];
    unsigned long long bar;                // the mbarrier of the bulk copy
    unsigned long long pad;
};
typedef GeomSmem<1> WarpSmem;
static_assert(sizeof(WarpSmem) % 16 == 0 && offsetof(WarpSmem, vtx) % 16 == 0, "bulk copies land on 16-byte boundaries");

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int WARPS>
__device__ __forceinline__ void bulk_load_start(GeomSmem<WARPS>* ws, const void* src, uint32_t bytes) {      // one lane
#if VKR_STAGE_TRIS
    const uint32_t bar = smem_addr(&ws->bar), dst = smem_addr(ws->tris);
#else
    const uint32_t bar = smem_addr(&ws->bar), dst = smem_addr(ws->vtx);
#endif
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");          // the async proxy must see the initialised barrier
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
template <int WARPS>
__device__ __forceinline__ bool bulk_load_wait(GeomSmem<WARPS>* ws) {           // every lane; false: gave up (never seen; keeps a fault from hanging the device)
    const uint32_t bar = smem_addr(&ws->bar);
    for (int spin = 0; spin < (1 << 24); ++spin) {
        uint32_t done;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar) : "memory");
        if (done) return true;
    }
    return false;
}

// WARPS = 1: the warp owns the cluster (the throughput form: no barrier between clusters).  WARPS = 4: the four warps of a CTA
// share one cluster — a quarter of the vertices, one 32-triangle iteration and every fourth occlusion footprint each, two
// CTA barriers — which cuts the LATENCY of a cluster to a third: the form for frames with fewer clusters than the GPU has
// warp slots (a rank's band of a sharded frame), where the kernel's duration is the duration of its slowest cluster.
template <int WARPS> __device__ __forceinline__ void geom_sync() { if (WARPS == 1) __syncwarp(); else __syncthreads(); }
template <int PASS, int WARPS, class MakeSink>
__device__ __forceinline__ void geometry_warp(const WorkItem* __restrict__ item, const float* __restrict__ mvps, uint32_t mvpStride,
                                              const RasterTarget& rt, const Queues& q, GeomSmem<WARPS>* ws, const MakeSink& makeSink) {
    const int lane = threadIdx.x & 31;
    const int wi = WARPS == 1 ? 0 : (int)(threadIdx.x >> 5);          // this warp's number among the cluster's warps
    const int tid = WARPS == 1 ? lane : (int)threadIdx.x;             // this thread's number among the cluster's threads
    const uint4 w0 = __ldg(reinterpret_cast<const uint4*>(item)), w1 = __ldg(reinterpret_cast<const uint4*>(item) + 1);
    const uint32_t info = w1.x, flags = (info >> 16) & 0xFFu;
    const int32_t verdict = (int32_t)(info >> 24);
    if (PASS == PASS_MAIN && (flags & 2u)) return;     // alpha-tested draw: geometry_cutout_kernel owns this cluster
    if (verdict == CLUSTER_CULLED) { if (tid == 0) atomicAdd(q.counters + 3, 1u); return; }
    const uint4* __restrict__ blob = reinterpret_cast<const uint4*>(((uint64_t)w0.y << 32) | w0.x);
    const uint32_t obj = w0.z, seqBase = w0.w;
    const uint32_t nVerts = info & 0xFFu, nTris = (info >> 8) & 0xFFu;
#if VKR_STAGE_TRIS
    if (tid == 0) bulk_load_start(ws, blob, CL_BLOB_TRI_BYTES + nVerts * 16u);
#else
    if (tid == 0) bulk_load_start(ws, blob + CL_BLOB_TRI_BYTES / 16u, nVerts * 16u);
    const uint2* __restrict__ gtris = reinterpret_cast<const uint2*>(blob);
    uint2 tpNext = (uint32_t)(wi * 32 + lane) < nTris ? __ldg(gtris + wi * 32 + lane) : make_uint2(0u, 0u);      // the first iteration's triangles, in flight during the cull
#endif
    if (tid < 16) ws->mvp[tid] = __ldg(mvps + (size_t)obj * mvpStride + tid);
    geom_sync<WARPS>();
    const float* mvp = ws->mvp;

    // ---- phase 0: the occlusion test against the LIVE depth / visibility buffer (the box test — frustum planes, pixel
    // rectangle, nearest depth — was done once per frame by expand_chunks_kernel)
    if (verdict == CLUSTER_TEST) {
        const int32_t rx0 = (int32_t)(w1.y & 0xFFFFu), ry0 = (int32_t)(w1.y >> 16), rx1 = (int32_t)(w1.z & 0xFFFFu), ry1 = (int32_t)(w1.z >> 16);
        const uint32_t zminBits = w1.w;
        // the warp walks the rectangle as 8x16 pixel footprints (four independent loads per lane in flight) and stops
        // at the first footprint with a visible pixel
        bool hidden = true;
        const int32_t lx = lane & 7, ly = lane >> 3;
        if (WARPS > 1) {
            // the footprints dealt to the warps (warp w takes every WARPS-th), a warp stops at its first visible pixel
            const int32_t fw = (rx1 - rx0 + 8) >> 3, fh = (ry1 - ry0 + 16) >> 4;
            for (int32_t f = wi; f < fw * fh && hidden; f += WARPS) {
                const int32_t fy = f / fw, bx = rx0 + (f - fy * fw) * 8, by = ry0 + fy * 16;
                const int32_t px = bx + lx;
                uint32_t stored[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int32_t py = by + ly + 4 * k;
                    stored[k] = 0u;
                    if (px <= rx1 && py <= ry1) {
                        const uint32_t pix = (uint32_t)(py * rt.width + px);
                        if (PASS == PASS_SHADOW) stored[k] = __ldcg(rt.depthBits + pix);
                        else stored[k] = __ldcg(reinterpret_cast<const uint32_t*>(rt.vis + pix) + 1);
                    }
                }
                const bool h = depth_hides(stored[0], zminBits) && depth_hides(stored[1], zminBits) &&
                               depth_hides(stored[2], zminBits) && depth_hides(stored[3], zminBits);
                hidden = __all_sync(0xFFFFFFFFu, h);
            }
            hidden = __syncthreads_and(hidden ? 1 : 0) != 0;
        } else
#if defined(VKR_OCC_WIDE)
        // A/B: 16x16 footprints (eight independent loads per lane in flight): half the dependent round trips of a culled cluster
        for (int32_t by = ry0; by <= ry1 && hidden; by += 16) {
            for (int32_t bx = rx0; bx <= rx1 && hidden; bx += 16) {
                uint32_t stored[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int32_t px = bx + lx + 8 * (k >> 2), py = by + ly + 4 * (k & 3);
                    stored[k] = 0u;
                    if (px <= rx1 && py <= ry1) {
                        const uint32_t pix = (uint32_t)(py * rt.width + px);
                        if (PASS == PASS_SHADOW) stored[k] = __ldcg(rt.depthBits + pix);
                        else stored[k] = __ldcg(reinterpret_cast<const uint32_t*>(rt.vis + pix) + 1);
                    }
                }
                bool h = true;
#pragma unroll
                for (int k = 0; k < 8; ++k) h = h && depth_hides(stored[k], zminBits);
                hidden = __all_sync(0xFFFFFFFFu, h);
            }
        }
#else
        for (int32_t by = ry0; by <= ry1 && hidden; by += 16) {
            for (int32_t bx = rx0; bx <= rx1 && hidden; bx += 8) {
                const int32_t px = bx + lx;
                uint32_t stored[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int32_t py = by + ly + 4 * k;
                    stored[k] = 0u;                                     // outside the rectangle: counts as hidden
                    if (px <= rx1 && py <= ry1) {
                        // .cg: read at L2, where the RED.MIN of the other warps land (stale L1 lines would only lose culls)
                        const uint32_t pix = (uint32_t)(py * rt.width + px);
                        if (PASS == PASS_SHADOW) stored[k] = __ldcg(rt.depthBits + pix);
                        else stored[k] = __ldcg(reinterpret_cast<const uint32_t*>(rt.vis + pix) + 1);
                    }
                }
                const bool h = depth_hides(stored[0], zminBits) && depth_hides(stored[1], zminBits) &&
                               depth_hides(stored[2], zminBits) && depth_hides(stored[3], zminBits);
                hidden = __all_sync(0xFFFFFFFFu, h);
            }
        }
#endif
        if (hidden) {
            bulk_load_wait(ws);                                         // the copy must have landed before this shared memory is given to another CTA
            if (tid == 0) atomicAdd(q.counters + 3, 1u);
            return;
        }
    }
    if (!bulk_load_wait(ws)) { if (tid == 0) q.counters[2] = 2u; return; }
    uint4* s_vtx = ws->vtx;
    ParkedTri* s_park = ws->park + wi * 32;

    // ---- phase 1: distinct vertices, in place (each thread reads and rewrites its own slots)
#pragma unroll
    for (int k = 0; k < CL_VERTS / (32 * WARPS); ++k)
        if (tid + 32u * WARPS * k < nVerts) {
            const uint4 pv = s_vtx[tid + 32u * WARPS * k];
            const CachedVtx cv = cluster_vertex(mvp, make_float4(__uint_as_float(pv.x), __uint_as_float(pv.y), __uint_as_float(pv.z), 1.0f), rt);
            s_vtx[tid + 32u * WARPS * k] = make_uint4((uint32_t)cv.X, (uint32_t)cv.Y, cv.zbits, cv.flag);
        }
    geom_sync<WARPS>();

    // ---- phase 2: triangles, 32 at a time.  Every lane sets one up and parks it in its shared-memory slot; then the warp
    // works through what was parked: the candidate samples of all the small ones dealt evenly to the lanes (most triangles
    // of a dense mesh cover one or two samples, a few cover dozens: a lane never waits for its neighbour's bounding box),
    // the medium and large ones queued for large_raster_kernel (vkr_raster.cuh, "warp-cooperative handling")
    for (uint32_t base = (uint32_t)wi * 32u; base < nTris; base += 32u * WARPS) {
        const uint32_t t = base + lane;
        int kind = 0;
        uint32_t count = 0, seqOfLane = 0;
#if !VKR_STAGE_TRIS
        const uint2 tpCur = tpNext;
        if (t + 32u * WARPS < nTris) tpNext = __ldg(gtris + t + 32u * WARPS);      // the next iteration's, one iteration ahead
#endif
        if (t < nTris) {
#if VKR_STAGE_TRIS
            const uint2 tp = ws->tris[t];
#else
            const uint2 tp = tpCur;
#endif
            const uint32_t l0 = tp.x & 127u, l1 = (tp.x >> 7) & 127u, l2 = (tp.x >> 14) & 127u;
            const uint4 a = s_vtx[l0], b = s_vtx[l1], c = s_vtx[l2];
            const CachedVtx va{(int32_t)a.x, (int32_t)a.y, a.z, a.w}, vb{(int32_t)b.x, (int32_t)b.y, b.z, b.w}, vc{(int32_t)c.x, (int32_t)c.y, c.z, c.w};
            // submission-order sequence number, whatever the processing order (the shadow pass has no use for it)
            const uint32_t seq = PASS == PASS_SHADOW ? 0u : seqBase + tp.y;
            seqOfLane = seq;
            // (a triangle that needs clipping re-reads its ORIGINAL positions from the cluster's block in global memory)
            kind = cluster_triangle_deferred(va, vb, vc, mvp, reinterpret_cast<const float4*>(blob + CL_BLOB_TRI_BYTES / 16u), l0, l1, l2, seq, rt, q,
                                             makeSink(seq), s_park[lane], count);
            if (kind != 3) count = 0;
        }
        const uint32_t medium = __ballot_sync(0xFFFFFFFFu, kind == 1);
        uint32_t large = __ballot_sync(0xFFFFFFFFu, kind == 2);
#if defined(VKR_RASTER_CANDIDATES)
        __syncwarp();
        {   // inclusive prefix sum of the candidate counts
            uint32_t incl = count;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += v; }
            const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
            if (total) raster_candidates_warp(s_park, ws->map, incl - count, count, seqOfLane, total, lane, makeSink);
        }
#else
        (void)count; (void)seqOfLane;
#endif
#if defined(VKR_RASTER_CANDIDATES)
        if ((medium | large) == 0u) { __syncwarp(); continue; }
#else
        if ((medium | large) == 0u) continue;        // (nobody reads a parked slot: nothing to order)
        __syncwarp();                                // the parked triangles are visible to the whole warp
#endif
        if (medium) {
            // medium triangles (extent < 64 px): one queue item each, no tile enumeration; the slots of the whole warp are
            // claimed with ONE 64-bit atomic over the two adjacent counters (triangles | items << 32).  large_raster_kernel
            // rasterises them with 32-bit edge functions.  (Claiming room for a whole cluster at once and padding the rest with
            // skip items was measured and lost: profiles/r02_ab.md, "queue room per cluster".)
            unsigned long long claimed = 0;
            if (lane == 0) claimed = atomicAdd(reinterpret_cast<unsigned long long*>(q.counters), (unsigned long long)__popc(medium) * 0x100000001ull);
            claimed = __shfl_sync(0xFFFFFFFFu, claimed, 0);
            const uint32_t firstTri = (uint32_t)claimed, firstItem = (uint32_t)(claimed >> 32);
            if (kind == 1) {
                const uint32_t k = (uint32_t)__popc(medium & ((1u << lane) - 1u));
                const bool fits = firstTri + k < q.triCap && firstItem + k < q.itemCap;
                if (fits) q.tris[firstTri + k] = s_park[lane].lt;
                else q.counters[2] = 1;
                if (firstItem + k < q.itemCap) {      // every claimed item slot below the capacity is written: a skip marker if the triangle did not fit
                    LargeItem it; it.tri = fits ? firstTri + k : ITEM_SKIP; it.tx = ITEM_WHOLE; it.ty = ITEM_WHOLE;
                    q.items[firstItem + k] = it;
                }
            }
        }
        __syncwarp();
        while (large) {
            const int src = __ffs((int)large) - 1;
            large &= large - 1;
            enqueue_large_coop(s_park[src].lt, q, lane);
        }
        __syncwarp();
    }
}

struct MakeVisSink { uint64_t* vis; int32_t width; __device__ VisSink operator()(uint32_t seq) const { return VisSink{vis, width, 0xFFFFFFFEu - seq}; } };
struct MakeDepthSink { uint32_t* bits; int32_t width; __device__ DepthSink operator()(uint32_t) const { return DepthSink{bits, width}; } };
struct MakeBlendSink { const ShadeConst* sc; const ShadeBuffers* sb; __device__ BlendSink operator()(uint32_t seq) const { return BlendSink{sc, sb, seq}; } };

template <bool SHADOW>
__global__ void __launch_bounds__(GEOM_WARPS * 32, VKR_GEOM_MIN_CTAS) geometry_kernel(const WorkItem* __restrict__ work, uint32_t nClusters,
                                                                     const float* __restrict__ mvps, uint32_t mvpStride,
                                                                     const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q) {
    __shared__ WarpSmem s_warp[GEOM_WARPS];
    const uint32_t slot = blockIdx.x * GEOM_WARPS + (threadIdx.x >> 5);
    if (slot >= nClusters) return;
    WarpSmem* ws = &s_warp[threadIdx.x >> 5];
    if (SHADOW) geometry_warp<PASS_SHADOW, 1>(work + slot, mvps, mvpStride, rt, q, ws, MakeDepthSink{rt.depthBits, rt.width});
    else geometry_warp<PASS_MAIN, 1>(work + slot, mvps, mvpStride, rt, q, ws, MakeVisSink{rt.vis, rt.width});
}
#if !VKR_STAGE_TRIS && !defined(VKR_RASTER_CANDIDATES)
// The latency form: one CTA of four warps per cluster (geometry_warp<.., 4>).  submit_frame picks it when a pass has fewer
// clusters than VKR_CTA_CLUSTERS (default 32768: on one GPU-eighth bands of config 4 the two forms cross between 26 k and 45 k clusters, profiles/r02_ab.md).
__global__ void __launch_bounds__(128, VKR_GEOM_MIN_CTAS) geometry_cta_kernel_main(const WorkItem* __restrict__ work, uint32_t nClusters,
                                                                     const float* __restrict__ mvps, uint32_t mvpStride,
                                                                     const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q) {
    __shared__ GeomSmem<4> s_cta;
    if (blockIdx.x >= nClusters) return;
    geometry_warp<PASS_MAIN, 4>(work + blockIdx.x, mvps, mvpStride, rt, q, &s_cta, MakeVisSink{rt.vis, rt.width});
}
__global__ void __launch_bounds__(128, VKR_GEOM_MIN_CTAS) geometry_cta_kernel_shadow(const WorkItem* __restrict__ work, uint32_t nClusters,
                                                                     const float* __restrict__ mvps, uint32_t mvpStride,
                                                                     const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q) {
    __shared__ GeomSmem<4> s_cta;
    if (blockIdx.x >= nClusters) return;
    geometry_warp<PASS_SHADOW, 4>(work + blockIdx.x, mvps, mvpStride, rt, q, &s_cta, MakeDepthSink{rt.depthBits, rt.width});
}
#define VKR_HAVE_CTA_GEOMETRY 1
#endif

// ---- "Transparent" draws (the tail of the draw list), after every opaque fragment is in the visibility buffer
__global__ void __launch_bounds__(GEOM_WARPS * 32) geometry_transparent_kernel(const WorkItem* __restrict__ work, uint32_t nClusters,
                                                                              const float* __restrict__ mvps, uint32_t mvpStride,
                                                                              const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q,
                                                                              const __grid_constant__ ShadeConst sc, const __grid_constant__ ShadeBuffers sbIn) {
    __shared__ WarpSmem s_warp[GEOM_WARPS];
    __shared__ float s_lut[256];
    for (uint32_t i = threadIdx.x; i < 256u; i += GEOM_WARPS * 32) s_lut[i] = sbIn.srgbLut[i];
    __syncthreads();
    __shared__ ShadeBuffers s_sb;
    if (threadIdx.x == 0) { s_sb = sbIn; s_sb.srgbLut = s_lut; }
    __syncthreads();
    const uint32_t slot = blockIdx.x * GEOM_WARPS + (threadIdx.x >> 5);
    if (slot >= nClusters) return;
    geometry_warp<PASS_TRANSPARENT, 1>(work + slot, mvps, mvpStride, rt, q, &s_warp[threadIdx.x >> 5], MakeBlendSink{&sc, &s_sb});
}

// ---- main pass, alpha-tested draws only (launched when the frame has any): one thread per triangle,
// straight from the original vertices, coverage gated by the albedo alpha (vkr_raster.cuh alpha_passes)
__global__ void __launch_bounds__(CL_TRIS) geometry_cutout_kernel(const DObject* __restrict__ objects, const DChunk* __restrict__ chunks,
                                                                  const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q,
                                                                  const __grid_constant__ AlphaCtx ax) {
    const DChunk& ch = chunks[blockIdx.x];
    if (!(ch.flags & 2u)) return;
    const uint32_t nTris = ch.counts >> 16;
    if (threadIdx.x >= nTris) return;
    const DObject& o = objects[ch.obj];
    const uint32_t packed = ch.clTris[threadIdx.x];
    const uint32_t g0 = ch.clVerts[packed & 127u], g1 = ch.clVerts[(packed >> 7) & 127u], g2 = ch.clVerts[(packed >> 14) & 127u];
    const uint32_t seq = ch.seqBase + ch.clOrig[threadIdx.x];
    VisSink sink{rt.vis, rt.width, 0xFFFFFFFEu - seq};
    cutout_triangle(o, o.mvp, g0, g1, g2, seq, rt, q, ax, sink);
}

// ---- large triangles: one warp (or, on a short queue, up to eight) per (triangle, TILE x TILE tile) or whole-box item, taken from a shared cursor
// CTAs of large_raster_kernel per SM (a register cap: 1 -> 125 registers, 2 resident; 3 -> 85; 4 -> 64 with spills), which is also
// the size of its persistent grid.  Configs 1 / 2 / 3, frames/s: 2 resident 19930 / 8195 / 891, 3 -> 20090 / 8282 / 899, 4 -> 20590 / 8333 / 885.
#ifndef VKR_LARGE_MIN_CTAS
#define VKR_LARGE_MIN_CTAS 3
#endif
#ifndef VKR_ITEM_CHUNK_MAX
#define VKR_ITEM_CHUNK_MAX 4u      // items a warp of large_raster_kernel takes per trip to the cursor (at most 32)
#endif
template <bool SHADOW>
__global__ void __launch_bounds__(256, VKR_LARGE_MIN_CTAS) large_raster_kernel(const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q,
                                                           const __grid_constant__ AlphaCtx ax, int anyCutout) {
    __shared__ uint32_t s_queue[SHADOW ? 1 : 8][SHADOW ? 1 : RECT_QUEUE];      // cut-out items: compaction queue per warp
    const uint32_t nItems = min(q.counters[1], q.itemCap), nTris = min(q.counters[0], q.triCap);
    const uint32_t warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    // Items differ in cost by orders of magnitude (an alpha-tested tile vs. an opaque sliver), so warps take them from a shared
    // cursor — but an item is also two DEPENDENT loads away (item -> triangle record), about a microsecond at this kernel's
    // occupancy, and a medium triangle's raster work is shorter than that: with one item per trip a frame with a million of
    // them (config 3) ran at the latency of those loads, 60 % of its stall samples on the cursor atomic in front of them.
    // A warp therefore takes `cs` consecutive items per trip — cs = 1 when there are fewer than 4 items per warp (a few hundred
    // heavy tile items must spread over all warps: a fixed chunk of 8 cost config 1 two thirds of its frame rate), up to
    // VKR_ITEM_CHUNK_MAX = 4 (config 3, frames/s: 1 -> 691, 4 -> 718, 32 -> 661: long chunks of heavy cut-out items make a tail)
    // — its lanes fetch the chunk's items and triangle records side by side into shared memory, and the cursor is asked for
    // the next chunk while this one is rasterised.  No atomic at all when the queue is shorter than the grid.
    // profiles/r02_ab.md, "item cursor".
    __shared__ LargeItem s_item[8][VKR_ITEM_CHUNK_MAX];
    __shared__ uint4 s_tri[8][VKR_ITEM_CHUNK_MAX][3];
    static_assert(sizeof(LargeTri) == 48, "three 16-byte loads per triangle record");
    const int wslot = threadIdx.x >> 5;
    const uint32_t cs = min((uint32_t)VKR_ITEM_CHUNK_MAX, max(1u, nItems / (warpsPerGrid * 4u)));
    // The other end of the scale: a queue much SHORTER than the grid (a few hundred tile items, 64 x 64 samples each at the time: config 1's
    // ground plane) leaves most warps idle while each busy one walks 4096 samples alone — the kernel then lasts as long as
    // one tile (71 us of config 1's 147 us frame).  With at most a quarter / half / as many items as warps, 8 / 4 / 2
    // warps share an item, each taking a band of its rows ("virtual items": item * wpi + part).  Config 1: 9550 -> 15350 frames/s.
    const uint32_t wsh = nItems * 4u <= warpsPerGrid ? 3u : nItems * 2u <= warpsPerGrid ? 2u : nItems <= warpsPerGrid ? 1u : 0u;
    const uint32_t nVirtual = nItems << wsh;
    uint32_t chunk = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * cs, nextChunk = 0xFFFFFFFFu;
    while (chunk < nVirtual) {
        if (lane == 0) nextChunk = (warpsPerGrid + atomicAdd(&q.counters[4], 1u)) * cs;
        const uint32_t n = min(cs, nVirtual - chunk);
        if ((uint32_t)lane < n) {
            LargeItem mine = q.items[(chunk + lane) >> wsh];
            if (mine.tri >= nTris) mine.tri = ITEM_SKIP;     // a skip marker of an overflowed frame (the host grows the queues and redraws)
            else {
                const uint4* src = reinterpret_cast<const uint4*>(q.tris + mine.tri);
                s_tri[wslot][lane][0] = src[0]; s_tri[wslot][lane][1] = src[1]; s_tri[wslot][lane][2] = src[2];
            }
            s_item[wslot][lane] = mine;
        }
        __syncwarp();
        for (uint32_t k = 0; k < n; ++k) {
            const LargeItem it = s_item[wslot][k];
            if (it.tri == ITEM_SKIP) continue;
            LargeTri lt;
            {
                uint4* dst = reinterpret_cast<uint4*>(&lt);
                dst[0] = s_tri[wslot][k][0]; dst[1] = s_tri[wslot][k][1]; dst[2] = s_tri[wslot][k][2];
            }
            const bool whole = it.tx == ITEM_WHOLE;          // medium triangle: its whole bounding box, 32-bit edge functions
            const uint32_t part = (chunk + k) & ((1u << wsh) - 1u);
            const bool cutoutItem = !SHADOW && anyCutout && (ax.objects[ax.chunkObj[lt.seq >> 8].x].flags & 2u);
            if (whole && !cutoutItem) {
                if (SHADOW) { DepthSink sink{rt.depthBits, rt.width}; raster_medium_coop(lt, sink, lane, part, wsh); }
                else { VisSink sink{rt.vis, rt.width, 0xFFFFFFFEu - lt.seq}; raster_medium_coop(lt, sink, lane, part, wsh); }
                continue;
            }
            int32_t x0, y0, x1, y1;
            item_rect(lt, it, x0, y0, x1, y1);
            if (wsh) {                                       // this warp's band of the item's rows
                const int32_t per = ((y1 - y0 + 1) + (int32_t)(1u << wsh) - 1) >> wsh;
                y0 += (int32_t)part * per;
                y1 = min(y1, y0 + per - 1);
                if (y0 > y1) continue;
            }
            if (SHADOW) {
                DepthSink sink{rt.depthBits, rt.width};
                raster_rect64(lt.X, lt.Y, lt.z0, lt.dzdx, lt.dzdy, x0, y0, x1, y1, sink, lane, 32);
            } else {
                VisSink sink{rt.vis, rt.width, 0xFFFFFFFEu - lt.seq};
                if (cutoutItem) {
                    const uint2 co = ax.chunkObj[lt.seq >> 8];
                    const DObject& o = ax.objects[co.x];
                    const uint32_t tri = co.y + (lt.seq & 255u);
                    AlphaTri at;
                    make_alpha_tri(o, o.mvp, o.idx[3 * tri], o.idx[3 * tri + 1], o.idx[3 * tri + 2], ax, rt, at);
                    CutoutSink<VisSink> cs2{sink, &at};
                    raster_rect64_compact(lt.X, lt.Y, lt.z0, lt.dzdx, lt.dzdy, x0, y0, x1, y1, cs2, lane, s_queue[SHADOW ? 0 : wslot]);
                } else raster_rect64(lt.X, lt.Y, lt.z0, lt.dzdx, lt.dzdy, x0, y0, x1, y1, sink, lane, 32);
            }
        }
        __syncwarp();                                        // the chunk's slots are free again
        chunk = __shfl_sync(0xFFFFFFFFu, nextChunk, 0);
    }
}
__global__ void __launch_bounds__(256) large_transparent_kernel(const __grid_constant__ RasterTarget rt, const __grid_constant__ Queues q,
                                                                const __grid_constant__ ShadeConst sc, const __grid_constant__ ShadeBuffers sbIn) {
    __shared__ float s_lut[256];
    __shared__ uint32_t s_queue[8][RECT_QUEUE];
    s_lut[threadIdx.x] = sbIn.srgbLut[threadIdx.x];
    __syncthreads();
    ShadeBuffers sb = sbIn;
    sb.srgbLut = s_lut;
    const uint32_t nItems = min(q.counters[1], q.itemCap), nTris = min(q.counters[0], q.triCap);
    const uint32_t warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < nItems;) {
        const uint32_t cur = i;
        if (lane == 0) i = warpsPerGrid + atomicAdd(&q.counters[4], 1u);
        i = __shfl_sync(0xFFFFFFFFu, i, 0);
        const LargeItem it = q.items[cur];
        if (it.tri >= nTris) continue;
        const LargeTri lt = q.tris[it.tri];
        BlendSink sink{&sc, &sb, lt.seq};
        raster_item_compact(lt, it, sink, lane, s_queue[threadIdx.x >> 5]);       // shading at raster time: keep the warp converged
    }
}

// ---- shade + resolve: one thread per pixel of the scissor rectangle; a CTA is a 16x8 pixel block,
// each warp an 8x4 footprint (triangle / texel / shadow-map locality).  The sRGB decode table is read
// through L1 (see below).
__global__ void __launch_bounds__(128, VKR_SHADE_MIN_CTAS) shade_kernel(const __grid_constant__ ShadeConst sc, const __grid_constant__ ShadeBuffers sbIn) {
#if defined(VKR_SHADE_LUT_SHARED)
    __shared__ float s_lut[256];
    s_lut[threadIdx.x] = sbIn.srgbLut[threadIdx.x];
    s_lut[threadIdx.x + 128] = sbIn.srgbLut[threadIdx.x + 128];
#endif
    // sharded over peer memory: rank 0 must have released the image this frame's pixels are also stored into (nothing on this
    // GPU stands between rank 0 and that release)
    if (threadIdx.x == 0 && sbIn.peerFree != nullptr) spin_until(sbIn.peerFree, sbIn.peerFreeValue, sbIn.countersHost + 12, 3u);
#if !defined(VKR_SHADE_LUT_SHARED)
    // the sRGB decode table (1 KB) is read from global memory and lives in L1: staging it in shared memory cost every 128-pixel
    // CTA 256 loads, 256 stores and a barrier, +2.8 % frame rate without (profiles/r02_ab.md, "decode table")
    if (sbIn.peerFree != nullptr) __syncthreads();
    const ShadeBuffers& sb = sbIn;
#else
    __syncthreads();
    ShadeBuffers sb = sbIn;
    sb.srgbLut = s_lut;
#endif
    if ((blockIdx.x | blockIdx.y) == 0 && threadIdx.x < 12 && sbIn.countersHost != nullptr) {
        // every raster pass of the frame has ended (stream order / the shadow join): publish their counters to the host
        const uint32_t* c = sbIn.counters[threadIdx.x >> 2];
        sbIn.countersHost[threadIdx.x] = c ? c[threadIdx.x & 3] : 0u;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lx = (lane & 7) + ((warp & 1) << 3), ly = (lane >> 3) + ((warp >> 1) << 2);
    const int px = sc.sx0 + blockIdx.x * 16 + lx, py = sc.sy0 + blockIdx.y * 8 + ly;
    if (px >= sc.sx1 || py >= sc.sy1) return;
    shade_pixel(sc, sb, px, py);
}

// ======================================================================= host side

#define CK(call)                                                                      \
    do {                                                                              \
        cudaError_t e_ = (call);                                                      \
        if (e_ != cudaSuccess) { return dev->fail(VKR_ERROR_CUDA, #call, e_); }       \
    } while (0)

// after a kernel launch: a launch that the driver refused is reported with the kernel's name, on the thread that issued it
#define LAUNCHED(name)                                                                \
    do {                                                                              \
        cudaError_t e_ = cudaGetLastError();                                          \
        if (e_ != cudaSuccess) { return dev->fail(VKR_ERROR_CUDA, "launch of " name, e_); } \
    } while (0)

struct ImageRec {
    uint32_t texel0 = 0;
    uchar4* data = nullptr;
    uint32_t w = 0, h = 0, levels = 1;
    uint32_t mipOffset[16] = {0};
    size_t texels = 0;
    int format = 0;
    bool hasCutout = false;
};

struct FrameSlot {
    void* pinned = nullptr; size_t pinnedCap = 0;
    void* devbuf = nullptr; size_t devCap = 0;
    cudaEvent_t uploaded = nullptr;
    cudaEvent_t rendered = nullptr;
    cudaEvent_t copied = nullptr;       // last asynchronous read-back of this slot's presented image (copy stream)
    bool copyPending = false;
    uint64_t copiedFrame = ~0ull;       // dev->frameIndex when that read-back was issued (= the frame it reads, plus one)
    // per-frame tables written on the device by expand_chunks_kernel (one set per frame in flight: the next frame's
    // tables are built on the upload stream while this frame's kernels still read theirs)
    DChunk* chunkBuf = nullptr; size_t chunkCap = 0;       // cluster table
    uint2* chunkObjBuf = nullptr; size_t seqSlotCap = 0;   // seq >> 8 -> (draw, first triangle)
    float* eyeBuf = nullptr; size_t eyeCap = 0;             // (1 + shadow layers) x draws x 8 floats: eye points (prepare_draws_kernel)
    WorkItem* workBuf = nullptr; size_t workCap = 0;       // (2 + shadow layers) x clusters: every pass's work items in processing order, then one list in submission order
};

struct VkrDevice {
    int device = 0;
    cudaStream_t stream = nullptr, copyStream = nullptr;   // frame stream; device->host copies
    cudaStream_t shadowStream = nullptr, upStream = nullptr; // shadow pass runs beside the main pass; host->device uploads
    cudaEvent_t evShadow = nullptr, packedEv = nullptr;
    // sharded over peer memory: the shadow map of slot s was cleared ahead of time for the frame whose flag value is
    // preclearedFor[s], preclearedLayers[s] layers of it; the clear is complete at preclearEv[s] (recorded on the frame stream)
    cudaEvent_t preclearEv[VKR_FRAME_OVERLAP] = {};
    uint32_t preclearedFor[VKR_FRAME_OVERLAP] = {}, preclearedLayers[VKR_FRAME_OVERLAP] = {};
    bool ownStream = false;
    uint32_t W = 0, H = 0, S = 2048, flags = 0;
    int smCount = 148;

    uint64_t* vis = nullptr;
    bool visClean = false;                 // VKR_CONFIG_TRANSIENT_DEPTH: the last frame's shading pass left `vis` cleared
    uint32_t* shadowBits = nullptr;        // the shadow map of the frame being (or last) drawn: one of shadowBuf
    uint32_t* shadowBuf[VKR_FRAME_OVERLAP] = {};   // per frame in flight: frame i+1's shadow pass overlaps frame i's shading
    uint16_t* hdr = nullptr;
    uint32_t* rgba8[VKR_FRAME_OVERLAP] = {nullptr, nullptr};
    uint32_t* scratch = nullptr;        // W*H*4, lazily allocated for depth / primid readback
    uint32_t* rgb8[VKR_FRAME_OVERLAP] = {nullptr, nullptr};   // W*H*3 (+pad), lazily allocated for VKR_TARGET_RGB8
    float* accum = nullptr;             // W*H*16, lazily allocated: additive sums of the transparent pass
    float* lut = nullptr;

    std::vector<DMesh> meshes;
    std::vector<void*> allocations;     // freed LIFO at destroy (DeletionQueue, src/vk_types.h:34-48)
    std::vector<ImageRec> images;
    std::vector<DTexture> texHost;
    DTexture* texDev = nullptr;
    bool texDirty = true;
    DTexture* cubeDev = nullptr;
    bool haveCube = false;
    std::vector<DMaterial> materials;
    DMaterial* matDev = nullptr; size_t matCap = 0; bool matDirty = true;

    struct RangeKey { uint32_t mesh, first, count; bool operator<(const RangeKey& o) const {
        return mesh != o.mesh ? mesh < o.mesh : (first != o.first ? first < o.first : count < o.count); } };
    std::map<RangeKey, ClusterRange> ranges;   // cluster data per (mesh, firstIndex, indexCount), built at first use
    std::vector<std::vector<uint32_t>> meshIndexHost;   // host copy of each mesh's indices (cluster building)
    std::vector<std::vector<float4>> meshPosHost;       // ... and positions (cluster bounds, Morton order)
    size_t uploadBytes = 0;
    // VKR_HOST_PROF=1: mean host time of vkr_draw's sections, printed at vkr_destroy
    bool hostProf = false; double hostProfUs[10] = {0}; uint64_t hostProfFrames = 0;
    // a pass with fewer clusters than this runs one CTA (four warps) per cluster instead of one warp: the latency form
    // (geometry_cta_kernel_*); VKR_CTA_CLUSTERS in the environment overrides it (0: never)
    uint32_t ctaClusters = 32768;


    Queues q{};                         // large-triangle queue of the main pass
    Queues qShadow{};                   // ... of the shadow pass (the two passes overlap)
    uint32_t* transCounters = nullptr;  // counters of the transparent pass (it reuses the main pass's queues)
    uint32_t* countersHost = nullptr;   // pinned mirror of the three passes' counters after the last frame (written by shade_kernel)

    FramePlan* lastPlan = nullptr; size_t lastN = 0;   // the frame drawn last (vkr_sync redraws it after a queue overflow)
    bool overflowReported = false;
    FrameSlot slots[VKR_FRAME_OVERLAP];
    uint64_t frameIndex = 0;
    int lastSlot = 0;
    const uint2* lastChunkObj = nullptr; const DObject* lastObjects = nullptr;
    FramePlan plans[2]; int planIdx = 0;      // two: with a submit thread one is being submitted while the next is planned
    // VKR_CONFIG_SUBMIT_THREAD: vkr_draw plans the frame on the caller's thread and hands the CUDA/NCCL calls to this thread
    // and asynchronous read-backs (vkr_readback_async) follow their frame through the same queue, so the caller never waits
    // for a frame to be issued before it can plan the next one
    std::thread worker;
    struct Job { int kind; FramePlan* plan; size_t n; VkrTarget what; void* dst; size_t dstBytes; };   // kind 0: frame, 1: read-back, 2: read-back of this rank's rows
    static const uint32_t JOBS = 4;
    Job jobs[JOBS];
    std::atomic<uint32_t> jobHead{0}, jobTail{0};   // single producer (the caller's thread), single consumer; head advances when a job is DONE
    std::atomic<bool> workerStop{false};
    std::mutex workerMutex; std::condition_variable workerCv;
    int asyncRc = 0;                          // first failure of an asynchronously submitted frame, reported by the next call

    cudaEvent_t tev[15] = {nullptr};  // 13, 14: around the last read-back transfer (copy stream)
    bool rbTimed = false;
     // ... 9 transparent done; 10 shadow raster done (before the exchange); 11, 12 frame gather (gather stream)   // 0 start, 1 uploaded, 3 cleared, 4 geometry kernel, 5 large kernel, 6 shaded; 7,8 shadow pass (shadow stream)
    bool tevValid = false, shadowTimed = false;
    float timings[VKR_T_COUNT] = {0};
    static const int KRING = 128;          // VKR_CONFIG_TIME_GEOM_KERNEL: event pairs of the last KRING frames
    cudaEvent_t kev[KRING][2] = {};
    uint64_t kevRecorded = 0, kevRead = 0;

    VkrShadowExchangeFn shadowExchange = nullptr; void* shadowExchangeUser = nullptr;
    // native sort-first sharding (vkr_shard_init)
    int shardRank = 0, shardRanks = 1;
    ncclComm_t commShadow = nullptr, commFrame = nullptr;
    cudaStream_t gatherStream = nullptr;
    uint32_t* gathered[VKR_FRAME_OVERLAP] = {nullptr, nullptr};
    cudaEvent_t gatheredEv[VKR_FRAME_OVERLAP] = {nullptr, nullptr};
    bool gatheredValid[VKR_FRAME_OVERLAP] = {false, false};
    // ... over peer memory (CUDA IPC): see "sort-first exchange over PEER MEMORY" above
    bool p2p = false;
    uint64_t shardFrame = 0;                       // frames drawn since vkr_shard_init (the same number on every rank)
    ShardFlags* myFlags = nullptr;                 // in my memory (tail of gathered[0]'s allocation)
    ShardFlags* peerFlags[MAX_RANKS] = {};         // [p]: rank p's flags, mapped here ([rank] = myFlags)
    uint32_t* peerShadow[MAX_RANKS][VKR_FRAME_OVERLAP] = {};   // [p][slot]: rank p's shadow maps ([rank] = my own)
    uint32_t* rootGathered[VKR_FRAME_OVERLAP] = {};            // rank 0's presented (assembled) images
    std::vector<void*> ipcOpened;
    int32_t lastBand[2] = {0, 0};                  // rows [y0, y1) of the last frame this rank rendered
    uint32_t lastGatherF1 = 0;                     // the last frame's flag value if it was assembled on rank 0 over peer memory, else 0
    VkrEngineStats stats{};
    std::vector<SceneRec*> scenes;                 // vkr_scene_create
    uint64_t bindGen = 0;                          // counts the changes of bindless slots (a scene's draw flags depend on them)
    uint32_t lastOpaqueClusters = 0, lastShadowClusters = 0;
    int launches = 0;
    std::string err;

    int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
        err = what;
        if (e != cudaSuccess) { err += ": "; err += cudaGetErrorString(e); }
        return code;
    }
};

static float g_srgb_lut[256];
static void init_lut() {
    for (int i = 0; i < 256; ++i) {
        double c = i / 255.0;
        g_srgb_lut[i] = (float)(c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4));
    }
}

static int dev_alloc(VkrDevice* dev, void** p, size_t bytes);
static int drain(VkrDevice* dev);
#define DRAIN(dev) do { if ((dev)->worker.joinable()) { int arc_ = drain(dev); if (arc_ != VKR_SUCCESS) return arc_; } } while (0)

struct DeviceClusterProvider : ClusterProvider {
    VkrDevice* dev;
    explicit DeviceClusterProvider(VkrDevice* d) : dev(d) {}
    bool get(uint32_t mesh, uint32_t firstIndex, uint32_t indexCount, ClusterRange& out) override {
        VkrDevice::RangeKey key{mesh, firstIndex, indexCount};
        auto it = dev->ranges.find(key);
        if (it != dev->ranges.end()) { out = it->second; return true; }
        { const int arc = drain(dev); if (arc != VKR_SUCCESS) dev->asyncRc = arc; }     // building touches state the submit thread uses
        ClusterData cd;
        build_clusters(dev->meshIndexHost[mesh].data() + firstIndex, indexCount / 3, dev->meshes[mesh].nv, dev->meshPosHost[mesh].data(), cd);
        uint32_t *dv = nullptr, *dt = nullptr, *dorig = nullptr; uint2* di = nullptr; float4* db = nullptr; uint4* dblob = nullptr; float4* dcone = nullptr;
        if (dev_alloc(dev, (void**)&dv, cd.verts.size() * 4) || dev_alloc(dev, (void**)&dt, cd.tris.size() * 4) ||
            dev_alloc(dev, (void**)&dorig, cd.orig.size() * 4) || dev_alloc(dev, (void**)&di, cd.info.size() * sizeof(uint2)) ||
            dev_alloc(dev, (void**)&db, cd.bounds.size() * sizeof(float4)) || dev_alloc(dev, (void**)&dblob, cd.blob.size() * sizeof(uint4)) ||
            dev_alloc(dev, (void**)&dcone, cd.cone.size() * sizeof(float4)))
            return false;
        bool ok = cudaMemcpyAsync(dv, cd.verts.data(), cd.verts.size() * 4, cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaMemcpyAsync(dt, cd.tris.data(), cd.tris.size() * 4, cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaMemcpyAsync(dorig, cd.orig.data(), cd.orig.size() * 4, cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaMemcpyAsync(di, cd.info.data(), cd.info.size() * sizeof(uint2), cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaMemcpyAsync(db, cd.bounds.data(), cd.bounds.size() * sizeof(float4), cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaMemcpyAsync(dblob, cd.blob.data(), cd.blob.size() * sizeof(uint4), cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaMemcpyAsync(dcone, cd.cone.data(), cd.cone.size() * sizeof(float4), cudaMemcpyHostToDevice, dev->stream) == cudaSuccess &&
                  cudaStreamSynchronize(dev->stream) == cudaSuccess;   // cd is a local
        if (!ok) return false;
        out = ClusterRange{dv, dt, dorig, di, db, dblob, dcone, 0u, {0, 0, 0}, 0};
        range_sphere(cd, out);
        dev->ranges[key] = out;
        return true;
    }
};

static int dev_alloc(VkrDevice* dev, void** p, size_t bytes) {
    cudaError_t e = cudaMalloc(p, bytes ? bytes : 16);
    if (e != cudaSuccess) return dev->fail(e == cudaErrorMemoryAllocation ? VKR_ERROR_OUT_OF_MEMORY : VKR_ERROR_CUDA, "cudaMalloc", e);
    dev->allocations.push_back(*p);
    return VKR_SUCCESS;
}
// Release one tracked allocation now (the caller has made sure the device no longer uses it).
static void dev_free(VkrDevice* dev, void* p) {
    if (!p) return;
    for (auto it = dev->allocations.begin(); it != dev->allocations.end(); ++it)
        if (*it == p) { dev->allocations.erase(it); break; }
    cudaFree(p);
}
#define ALLOC(ptr, bytes)                                                    \
    do { int rc_ = dev_alloc(dev, (void**)&(ptr), (bytes)); if (rc_) return rc_; } while (0)

static inline unsigned grid_for(size_t n, unsigned block, unsigned maxBlocks) {
    size_t g = (n + block - 1) / block;
    return (unsigned)std::max<size_t>(1, std::min<size_t>(g, maxBlocks));
}

extern "C" {

const char* vkr_version(void) { return "vkr 0.1 (sm_100a)"; }

const char* vkr_last_error(VkrDevice* dev) { return dev ? dev->err.c_str() : "null device"; }

int vkr_create(const VkrConfig* cfg, VkrDevice** out) {
    if (!cfg || !out || cfg->width == 0 || cfg->height == 0 || cfg->width > 16384 || cfg->height > 16384)
        return VKR_ERROR_INVALID_ARGUMENT;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || cfg->device < 0 || cfg->device >= count) {
        cudaGetLastError();
        return VKR_ERROR_NO_DEVICE;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess || prop.major < 10) return VKR_ERROR_NO_DEVICE;
    VkrDevice* dev = new VkrDevice();
    dev->device = cfg->device;
    dev->W = cfg->width; dev->H = cfg->height;
    dev->S = cfg->shadowExtent ? cfg->shadowExtent : 2048;
    dev->flags = cfg->flags;
    dev->smCount = prop.multiProcessorCount;
    dev->hostProf = std::getenv("VKR_HOST_PROF") != nullptr;
    if (const char* e = std::getenv("VKR_CTA_CLUSTERS")) dev->ctaClusters = (uint32_t)std::strtoul(e, nullptr, 10);
    *out = dev;
    auto body = [&]() -> int {
        CK(cudaSetDevice(dev->device));
        if (cfg->stream) dev->stream = (cudaStream_t)cfg->stream;
        else { CK(cudaStreamCreateWithFlags(&dev->stream, cudaStreamNonBlocking)); dev->ownStream = true; }
        // the shadow stream carries the shadow exchange of a sharded frame: highest priority, so that the collective's
        // CTAs are dispatched ahead of the main pass's thousands of queued geometry CTAs; the copy stream's pack kernel likewise
        int prLo = 0, prHi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&prLo, &prHi));
        CK(cudaStreamCreateWithPriority(&dev->copyStream, cudaStreamNonBlocking, prHi));
        CK(cudaEventCreateWithFlags(&dev->packedEv, cudaEventDisableTiming));
        CK(cudaStreamCreateWithPriority(&dev->shadowStream, cudaStreamNonBlocking, prHi));
        CK(cudaStreamCreateWithFlags(&dev->upStream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&dev->evShadow, cudaEventDisableTiming));
        for (auto& e : dev->preclearEv) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        const size_t px = (size_t)dev->W * dev->H;
        ALLOC(dev->vis, px * 8);
        ALLOC(dev->hdr, px * 8);
        for (int i = 0; i < VKR_FRAME_OVERLAP; ++i) ALLOC(dev->rgba8[i], px * 4);
        for (auto& b : dev->shadowBuf) ALLOC(b, (size_t)dev->S * dev->S * 4 * VKR_NUM_CASCADES);
        dev->shadowBits = dev->shadowBuf[0];
        ALLOC(dev->lut, 256 * 4);
        init_lut();
        CK(cudaMemcpyAsync(dev->lut, g_srgb_lut, sizeof(g_srgb_lut), cudaMemcpyHostToDevice, dev->stream));
        ALLOC(dev->texDev, sizeof(DTexture) * VKR_MAX_BINDLESS);
        ALLOC(dev->cubeDev, sizeof(DTexture) * 6);
        dev->texHost.assign(VKR_MAX_BINDLESS, DTexture{});
        dev->q.triCap = 1u << 20; dev->q.itemCap = 1u << 22;
        dev->qShadow.triCap = 1u << 18; dev->qShadow.itemCap = 1u << 20;
        if (const char* caps = std::getenv("VKR_QUEUE_CAPS")) {     // "tris,items": starting capacities of both queues (tests of the overflow path)
            unsigned a = 0, b = 0;
            if (std::sscanf(caps, "%u,%u", &a, &b) == 2 && a && b) { dev->q.triCap = dev->qShadow.triCap = a; dev->q.itemCap = dev->qShadow.itemCap = b; }
        }
        ALLOC(dev->q.tris, sizeof(LargeTri) * dev->q.triCap);
        ALLOC(dev->q.items, sizeof(LargeItem) * dev->q.itemCap);
        ALLOC(dev->q.counters, 32);          // [0..3] see Queues; [4] work cursor of the large-triangle kernels
#if defined(VKR_BIN_PROBE)
        {   // A/B probe only (vkr_raster.cuh): bins for the append half of a tile-binned rasteriser
            const size_t tiles = (size_t)((dev->W + 63) / 64) * ((dev->H + 63) / 64);
            uint4* pool = nullptr; uint32_t* cnt = nullptr;
            ALLOC(pool, tiles * VKR_BIN_CAP * 48);
            ALLOC(cnt, tiles * 4);
            CK(cudaMemset(cnt, 0, tiles * 4));
            CK(cudaMemcpyToSymbol(g_binPool, &pool, sizeof pool));
            CK(cudaMemcpyToSymbol(g_binCount, &cnt, sizeof cnt));
        }
#endif
        ALLOC(dev->qShadow.tris, sizeof(LargeTri) * dev->qShadow.triCap);
        ALLOC(dev->qShadow.items, sizeof(LargeItem) * dev->qShadow.itemCap);
        ALLOC(dev->qShadow.counters, 32 * VKR_FRAME_OVERLAP);     // per frame in flight: frame i+1's shadow pass runs beside frame i's shading, which publishes frame i's
        ALLOC(dev->transCounters, 32);
        CK(cudaMallocHost(&dev->countersHost, 64));     // (mapped: unified addressing makes every pinned allocation device-visible)
        std::memset(dev->countersHost, 0, 64);          // [0..11] three passes' counters, [12] peer-exchange error word, [13..14] vkr_draw_scene: visible draws, their triangles
        for (auto& s : dev->slots) {
            CK(cudaEventCreateWithFlags(&s.uploaded, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&s.rendered, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&s.copied, cudaEventDisableTiming));
        }
        for (auto& e : dev->tev) CK(cudaEventCreate(&e));
        // shadow image starts cleared to 1.0 so an un-rendered map reads "unshadowed"
        for (uint32_t* b : dev->shadowBuf)
            fill_u32_kernel<<<grid_for((size_t)dev->S * dev->S * VKR_NUM_CASCADES / 4, 256, 4096), 256, 0, dev->stream>>>(
                b, (size_t)dev->S * dev->S * VKR_NUM_CASCADES, DEPTH_ONE_BITS);
        fill_u64_kernel<<<grid_for(px / 2, 256, 4096), 256, 0, dev->stream>>>(dev->vis, px, VIS_CLEAR);
        CK(cudaGetLastError());
        // init_default_data: bindless slot 0 = default normal texture, NEAREST (src/vk_engine_init.cpp:351-355)
        uint32_t normalDefault = 0xFFFF8080u;
        VkrImage img;
        int rc = vkr_create_image(dev, &normalDefault, 1, 1, VKR_FORMAT_R8G8B8A8_UNORM, 0, &img);
        if (rc) return rc;
        rc = vkr_update_bindless_texture(dev, 0, img, VKR_SAMPLER_NEAREST);
        if (rc) return rc;
        CK(cudaStreamSynchronize(dev->stream));
        return VKR_SUCCESS;
    };
    int rc = body();
    if (rc != VKR_SUCCESS) {
        std::fprintf(stderr, "vkr_create: %s\n", dev->err.c_str());
        vkr_destroy(dev);
        *out = nullptr;
    }
    return rc;
}

void vkr_destroy(VkrDevice* dev) {
    if (!dev) return;
    if (dev->worker.joinable()) {
        (void)drain(dev);
        dev->workerStop.store(true);
        dev->workerCv.notify_all();
        dev->worker.join();
    }
    cudaSetDevice(dev->device);
    cudaDeviceSynchronize();
    if (dev->hostProf && dev->hostProfFrames) {
        const double f = 1.0 / (double)dev->hostProfFrames;
        std::fprintf(stderr, "[vkr host rank %d] vkr_draw us/frame over %llu frames: plan %.1f | upload+expand %.1f | shadow (+exchange) %.1f | "
                     "main raster %.1f | transparent+shade %.1f | gather %.1f | async read-backs %.1f | caller waiting for the submit thread %.1f | submit thread waiting for the device %.1f\n",
                     dev->shardRank, (unsigned long long)dev->hostProfFrames,
                     dev->hostProfUs[0] * f, dev->hostProfUs[1] * f, dev->hostProfUs[2] * f, dev->hostProfUs[3] * f, dev->hostProfUs[4] * f, dev->hostProfUs[5] * f,
                     dev->hostProfUs[6] * f, dev->hostProfUs[7] * f, dev->hostProfUs[8] * f);
    }
    for (SceneRec* sr : dev->scenes) delete sr;            // (their device arrays are among the tracked allocations)
    for (auto it = dev->allocations.rbegin(); it != dev->allocations.rend(); ++it) cudaFree(*it);
    for (auto& s : dev->slots) {
        if (s.pinned) cudaFreeHost(s.pinned);
        if (s.devbuf) cudaFree(s.devbuf);
        if (s.uploaded) cudaEventDestroy(s.uploaded);
        if (s.rendered) cudaEventDestroy(s.rendered);
        if (s.copied) cudaEventDestroy(s.copied);
    }
    for (auto& e : dev->tev) if (e) cudaEventDestroy(e);
    if (dev->countersHost) cudaFreeHost(dev->countersHost);
    for (void* p : dev->ipcOpened) cudaIpcCloseMemHandle(p);
    if (dev->commShadow) ncclCommDestroy(dev->commShadow);
    if (dev->commFrame) ncclCommDestroy(dev->commFrame);
    if (dev->gatherStream) cudaStreamDestroy(dev->gatherStream);
    for (auto& e : dev->gatheredEv) if (e) cudaEventDestroy(e);
    for (auto& pr : dev->kev) { if (pr[0]) cudaEventDestroy(pr[0]); if (pr[1]) cudaEventDestroy(pr[1]); }
    if (dev->copyStream) cudaStreamDestroy(dev->copyStream);
    if (dev->shadowStream) cudaStreamDestroy(dev->shadowStream);
    if (dev->upStream) cudaStreamDestroy(dev->upStream);
    if (dev->evShadow) cudaEventDestroy(dev->evShadow);
    for (auto& e : dev->preclearEv) if (e) cudaEventDestroy(e);
    if (dev->packedEv) cudaEventDestroy(dev->packedEv);
    if (dev->ownStream && dev->stream) cudaStreamDestroy(dev->stream);
    delete dev;
}

int vkr_upload_mesh(VkrDevice* dev, const uint32_t* indices, size_t ni, const VkrVertex* vertices, size_t nv, VkrMesh* out) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!out || (ni && !indices) || (nv && !vertices) || nv >= 0xFFFFFFFFull || ni >= 0xFFFFFFFFull)
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_upload_mesh: bad arguments");
    for (size_t i = 0; i < ni; ++i)
        if (indices[i] >= nv) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_upload_mesh: index out of range");
    CK(cudaSetDevice(dev->device));
    DMesh m{};
    m.nv = (uint32_t)nv; m.ni = (uint32_t)ni;
    float4 *pos, *aA, *aB; uint32_t* idx; VkrVertex* staging = nullptr;
    ALLOC(pos, nv * 16); ALLOC(aA, nv * 16); ALLOC(aB, nv * 16); ALLOC(idx, ni * 4);
    if (nv) {
        cudaError_t e = cudaMalloc(&staging, nv * sizeof(VkrVertex));
        if (e != cudaSuccess) return dev->fail(VKR_ERROR_OUT_OF_MEMORY, "cudaMalloc(staging)", e);
        CK(cudaMemcpyAsync(staging, vertices, nv * sizeof(VkrVertex), cudaMemcpyHostToDevice, dev->stream));
        repack_vertices_kernel<<<(unsigned)((nv + 255) / 256), 256, 0, dev->stream>>>(staging, nv, pos, aA, aB);
        CK(cudaGetLastError());
    }
    if (ni) CK(cudaMemcpyAsync(idx, indices, ni * 4, cudaMemcpyHostToDevice, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));     // caller keeps ownership of the host arrays (immediate_submit semantics)
    if (staging) cudaFree(staging);
    m.pos = pos; m.attrA = aA; m.attrB = aB; m.idx = idx;
    dev->meshes.push_back(m);
    dev->meshIndexHost.emplace_back(indices, indices + ni);
    {
        std::vector<float4> hp(nv);
        for (size_t i = 0; i < nv; ++i) hp[i] = make_float4(vertices[i].position[0], vertices[i].position[1], vertices[i].position[2], 1.0f);
        dev->meshPosHost.push_back(std::move(hp));
    }
    *out = (VkrMesh)(dev->meshes.size() - 1);
    return VKR_SUCCESS;
}

static int build_image(VkrDevice* dev, const void* rgba8, uint32_t w, uint32_t h, VkrFormat fmt, int mipmapped, ImageRec& im) {
    im.w = w; im.h = h; im.format = fmt;
    im.levels = mipmapped ? (uint32_t)std::floor(std::log2((double)std::max(w, h))) + 1 : 1;
    if (im.levels > 16) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "image too large");
    size_t total = 0;
    for (uint32_t l = 0; l < im.levels; ++l) {
        im.mipOffset[l] = (uint32_t)total;
        total += (size_t)std::max(1u, w >> l) * std::max(1u, h >> l);
    }
    im.texels = total;
    ALLOC(im.data, total * 4);
    CK(cudaMemcpyAsync(im.data, rgba8, (size_t)w * h * 4, cudaMemcpyHostToDevice, dev->stream));
    for (uint32_t l = 0; l + 1 < im.levels; ++l) {
        int sw = std::max(1u, w >> l), sh = std::max(1u, h >> l), dw = std::max(1u, w >> (l + 1)), dh = std::max(1u, h >> (l + 1));
        dim3 b(16, 16), g((dw + 15) / 16, (dh + 15) / 16);
        mip_blit_kernel<<<g, b, 0, dev->stream>>>(im.data + im.mipOffset[l], sw, sh, im.data + im.mipOffset[l + 1], dw, dh,
                                                  fmt == VKR_FORMAT_R8G8B8A8_SRGB, dev->lut);
    }
    CK(cudaGetLastError());
    const uint8_t* p = (const uint8_t*)rgba8;
    std::memcpy(&im.texel0, rgba8, 4);
    im.hasCutout = false;
    for (size_t i = 0; i < (size_t)w * h; ++i) if (p[i * 4 + 3] < 128) { im.hasCutout = true; break; }
    CK(cudaStreamSynchronize(dev->stream));
    return VKR_SUCCESS;
}

static DTexture make_dtexture(const ImageRec& im, int sampler) {
    DTexture t{};
    t.data = im.data; t.w = im.w; t.h = im.h; t.levels = im.levels;
    t.flags = (im.format == VKR_FORMAT_R8G8B8A8_SRGB ? 1u : 0u) | (sampler == VKR_SAMPLER_LINEAR ? 2u : 0u) | (im.hasCutout ? 4u : 0u);
    std::memcpy(t.mipOffset, im.mipOffset, sizeof(t.mipOffset));
    t.texel0 = im.texel0;
    return t;
}

// The sky is sampled at every background pixel of every frame: its six faces keep decoded RGBA32F texels (DTexture.flags bit 4)
static int predecode_cube_face(VkrDevice* dev, DTexture& d) {
    if (d.flags & (8u | 16u)) return VKR_SUCCESS;                     // f16 texels, or decoded already
    const uint32_t last = d.levels - 1;
    const size_t total = (size_t)d.mipOffset[last] + (size_t)std::max(1u, d.w >> last) * std::max(1u, d.h >> last);
    float4* out = nullptr;
    ALLOC(out, total * sizeof(float4));
    decode_texels_kernel<<<(unsigned)((total + 255) / 256), 256, 0, dev->stream>>>(d.data, out, total, dev->lut, (int)(d.flags & 1u));
    LAUNCHED("decode_texels_kernel");
    d.data = reinterpret_cast<const uchar4*>(out);
    d.flags |= 16u;
    return VKR_SUCCESS;
}

int vkr_create_image(VkrDevice* dev, const void* rgba8, uint32_t w, uint32_t h, VkrFormat fmt, int mipmapped, VkrImage* out) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!rgba8 || !out || w == 0 || h == 0 || (fmt != VKR_FORMAT_R8G8B8A8_UNORM && fmt != VKR_FORMAT_R8G8B8A8_SRGB))
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_create_image: bad arguments");
    CK(cudaSetDevice(dev->device));
    ImageRec im;
    int rc = build_image(dev, rgba8, w, h, fmt, mipmapped, im);
    if (rc) return rc;
    dev->images.push_back(im);
    *out = (VkrImage)(dev->images.size() - 1);
    return VKR_SUCCESS;
}

int vkr_update_bindless_texture(VkrDevice* dev, int index, VkrImage image, VkrSampler sampler) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (index < 0 || index >= VKR_MAX_BINDLESS || image >= dev->images.size())
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_update_bindless_texture: bad index or image");
    dev->texHost[index] = make_dtexture(dev->images[image], sampler);
    dev->texDirty = true;
    dev->bindGen++;
    return VKR_SUCCESS;
}

int vkr_set_cubemap(VkrDevice* dev, const void* const faces[6], uint32_t w, uint32_t h, VkrFormat fmt, int mipmapped) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!faces || w == 0 || h == 0 || (fmt != VKR_FORMAT_R8G8B8A8_UNORM && fmt != VKR_FORMAT_R8G8B8A8_SRGB))
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_set_cubemap: bad arguments");
    CK(cudaSetDevice(dev->device));
    DTexture t[6];
    for (int f = 0; f < 6; ++f) {
        if (!faces[f]) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_set_cubemap: null face");
        ImageRec im;
        int rc = build_image(dev, faces[f], w, h, fmt, mipmapped, im);
        if (rc) return rc;
        t[f] = make_dtexture(im, VKR_SAMPLER_LINEAR);
        rc = predecode_cube_face(dev, t[f]);
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(dev->cubeDev, t, sizeof(t), cudaMemcpyHostToDevice, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));
    dev->haveCube = true;
    return VKR_SUCCESS;
}

int vkr_set_cubemap_levels(VkrDevice* dev, const void* const* levels, uint32_t w, uint32_t h, uint32_t levelCount, VkrFormat fmt) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!levels || w == 0 || h == 0 || w > 16384 || h > 16384 || levelCount == 0 || levelCount > 16 ||
        (fmt != VKR_FORMAT_R8G8B8A8_UNORM && fmt != VKR_FORMAT_R8G8B8A8_SRGB && fmt != VKR_FORMAT_R16G16B16A16_SFLOAT))
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_set_cubemap_levels: bad arguments");
    for (uint32_t i = 0; i < levelCount * 6; ++i)
        if (!levels[i]) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_set_cubemap_levels: null level");
    CK(cudaSetDevice(dev->device));
    const size_t bpt = fmt == VKR_FORMAT_R16G16B16A16_SFLOAT ? 8 : 4;
    DTexture t[6];
    for (int f = 0; f < 6; ++f) {
        DTexture& d = t[f];
        d = DTexture{};
        size_t total = 0;
        for (uint32_t l = 0; l < levelCount; ++l) {
            d.mipOffset[l] = (uint32_t)total;
            total += (size_t)std::max(1u, w >> l) * std::max(1u, h >> l);
        }
        char* data = nullptr;
        ALLOC(data, total * bpt);
        for (uint32_t l = 0; l < levelCount; ++l)
            CK(cudaMemcpyAsync(data + (size_t)d.mipOffset[l] * bpt, levels[(size_t)l * 6 + f],
                               (size_t)std::max(1u, w >> l) * std::max(1u, h >> l) * bpt, cudaMemcpyHostToDevice, dev->stream));
        d.data = reinterpret_cast<const uchar4*>(data);
        d.w = w; d.h = h; d.levels = levelCount;
        d.flags = (fmt == VKR_FORMAT_R8G8B8A8_SRGB ? 1u : 0u) | 2u | (fmt == VKR_FORMAT_R16G16B16A16_SFLOAT ? 8u : 0u);
        const int rc = predecode_cube_face(dev, d);
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(dev->cubeDev, t, sizeof(t), cudaMemcpyHostToDevice, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));      // the caller's arrays and `t` may go away
    dev->haveCube = true;
    return VKR_SUCCESS;
}

int vkr_build_material(VkrDevice* dev, const VkrMaterialConstants* p, VkrMaterialPass pass, VkrMaterial* out) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!p || !out) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_build_material: bad arguments");
    DMaterial m;
    std::memcpy(m.colorFactors, p->colorFactors, 16);
    std::memcpy(m.metalRough, p->metal_rough_factors, 16);
    m.albedoID = p->albedoID; m.normalID = p->normalID; m.metalRoughID = p->metalRoughID; m.pass = (int32_t)pass;
    m.constFlags = 0u;               // (filled at the next vkr_draw: material_constants)
    dev->materials.push_back(m);
    dev->matDirty = true;
    *out = (VkrMaterial)(dev->materials.size() - 1);
    return VKR_SUCCESS;
}

static int ensure_slot(VkrDevice* dev, FrameSlot& s, size_t bytes) {
    if (s.pinnedCap < bytes) {
        if (s.pinned) cudaFreeHost(s.pinned);
        if (s.devbuf) cudaFree(s.devbuf);
        size_t cap = std::max<size_t>(bytes * 3 / 2, 1 << 16);
        CK(cudaMallocHost(&s.pinned, cap));
        CK(cudaMalloc(&s.devbuf, cap));
        s.pinnedCap = s.devCap = cap;
    }
    return VKR_SUCCESS;
}

}  // extern "C"

// ---- submit thread (VKR_CONFIG_SUBMIT_THREAD) -------------------------------------------------------------------
static int submit_frame(VkrDevice* dev, FramePlan& plan, size_t n);

extern "C" {
static int readback_issue(VkrDevice* dev, VkrTarget what, void* dst, size_t dstBytes);
static int readback_band_issue(VkrDevice* dev, VkrTarget what, void* frameBase, size_t frameBytes);
}

static void worker_main(VkrDevice* dev) {
    cudaSetDevice(dev->device);
    for (;;) {
        int spins = 0;
        uint32_t head = dev->jobHead.load(std::memory_order_relaxed);
        while (dev->jobTail.load(std::memory_order_acquire) == head && !dev->workerStop.load(std::memory_order_acquire)) {
            if (++spins < 20000) { __builtin_ia32_pause(); continue; }          // ~0.2 ms of spinning covers the gap between frames
            std::unique_lock<std::mutex> lk(dev->workerMutex);
            dev->workerCv.wait_for(lk, std::chrono::milliseconds(2), [&] { return dev->jobTail.load() != head || dev->workerStop.load(); });
            spins = 0;
        }
        if (dev->jobTail.load(std::memory_order_acquire) == head) return;        // stop requested and nothing left to issue
        const VkrDevice::Job job = dev->jobs[head % VkrDevice::JOBS];
        const auto jobT = std::chrono::steady_clock::now();
        const int rc = job.kind == 0 ? submit_frame(dev, *job.plan, job.n)
                     : job.kind == 1 ? readback_issue(dev, job.what, job.dst, job.dstBytes) : readback_band_issue(dev, job.what, job.dst, job.dstBytes);
        if (job.kind != 0 && dev->hostProf && dev->frameIndex >= 16)
            dev->hostProfUs[6] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - jobT).count();
        if (rc != VKR_SUCCESS && dev->asyncRc == VKR_SUCCESS) dev->asyncRc = rc;
        dev->jobHead.store(head + 1, std::memory_order_release);
    }
}
// Hand a job to the submit thread (blocks only while all JOBS slots are taken).
static void post_job(VkrDevice* dev, const VkrDevice::Job& job) {
    const uint32_t tail = dev->jobTail.load(std::memory_order_relaxed);
    while (tail - dev->jobHead.load(std::memory_order_acquire) >= VkrDevice::JOBS) __builtin_ia32_pause();
    dev->jobs[tail % VkrDevice::JOBS] = job;
    dev->jobTail.store(tail + 1, std::memory_order_release);
    dev->workerCv.notify_one();
}
// Wait until the submit thread has issued everything it was given; returns (and clears) a deferred failure.
static int drain(VkrDevice* dev) {
    if (!dev->worker.joinable()) return VKR_SUCCESS;
    int spins = 0;
    const uint32_t tail = dev->jobTail.load(std::memory_order_relaxed);
    while (dev->jobHead.load(std::memory_order_acquire) != tail) { if (++spins < 4000) __builtin_ia32_pause(); else std::this_thread::yield(); }
    const int rc = dev->asyncRc;
    dev->asyncRc = VKR_SUCCESS;
    return rc;
}

extern "C" int vkr_draw(VkrDevice* dev, const VkrGPUSceneData* scene, const VkrRenderObject* objs, size_t n, const VkrFrameParams* fp) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    if (!scene || !fp || (n && !objs)) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_draw: null argument");
    CK(cudaSetDevice(dev->device));
    const bool async = (dev->flags & VKR_CONFIG_SUBMIT_THREAD) && !dev->shadowExchange;     // a user hook runs on the caller's thread
    if (!async) DRAIN(dev);
    if (async && !dev->worker.joinable()) dev->worker = std::thread(worker_main, dev);
    const auto hostT = std::chrono::steady_clock::now();
    FramePlan& plan = dev->plans[dev->planIdx];        // the other one may still be in the submit thread's hands
    DeviceClusterProvider provider(dev);
    std::string why = plan_frame(scene, objs, n, fp, dev->meshes, dev->materials, dev->texHost, provider, dev->W, dev->H, dev->S, plan,
                                 dev->shardRank, dev->shardRanks);     // sharded: also splits the rows into cost-balanced bands
    if (!why.empty()) { DRAIN(dev); return dev->fail(VKR_ERROR_INVALID_ARGUMENT, why.c_str()); }
    if (dev->hostProf && dev->frameIndex >= 16)
        dev->hostProfUs[0] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - hostT).count();
    if (!async) return submit_frame(dev, plan, n);
    const auto waitT = std::chrono::steady_clock::now();
    DRAIN(dev);                                        // the previous frame is fully issued (and its failure, if any, reported here)
    if (dev->hostProf && dev->frameIndex >= 16)
        dev->hostProfUs[7] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - waitT).count();
    dev->planIdx ^= 1;
    post_job(dev, VkrDevice::Job{0, &plan, n, VKR_TARGET_RGBA8, nullptr, 0});
    return VKR_SUCCESS;
}

static int submit_frame(VkrDevice* dev, FramePlan& plan, size_t n) {
    auto hostT = std::chrono::steady_clock::now();
    auto mark = [&](int k) {
        if (!dev->hostProf || dev->frameIndex < 16) return;      // skip allocation / NCCL warm-up frames
        const auto now = std::chrono::steady_clock::now();
        dev->hostProfUs[k] += std::chrono::duration<double, std::micro>(now - hostT).count();
        hostT = now;
    };
    cudaStream_t st = dev->stream;
    dev->launches = 0;
    const bool timing = dev->flags & VKR_CONFIG_TIMINGS;
    const bool timeKernel = !timing && (dev->flags & VKR_CONFIG_TIME_GEOM_KERNEL);
    if (timing) CK(cudaEventRecord(dev->tev[0], st));

    // ---- descriptor-ish state
    if (dev->texDirty) {
        CK(cudaMemcpyAsync(dev->texDev, dev->texHost.data(), sizeof(DTexture) * VKR_MAX_BINDLESS, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));   // texHost may change right after return
        dev->texDirty = false;
        dev->matDirty = true;            // the materials' uv-independent texels follow the slots
    }
    if (dev->matDirty) {
        for (DMaterial& m : dev->materials) material_constants(m, dev->texHost.data(), g_srgb_lut);
        if (dev->matCap < dev->materials.size()) {
            dev->matCap = std::max<size_t>(dev->materials.size() * 2, 64);
            ALLOC(dev->matDev, dev->matCap * sizeof(DMaterial));
        }
        CK(cudaMemcpyAsync(dev->matDev, dev->materials.data(), dev->materials.size() * sizeof(DMaterial), cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        dev->matDirty = false;
    }

    // ---- upload the draw list (replaces per-object push constants + the scene UBO)
    const int slotIdx = (int)(dev->frameIndex % VKR_FRAME_OVERLAP);
    FrameSlot& slot = dev->slots[slotIdx];
    dev->uploadBytes = 0;
    // vkr_draw_scene: the per-draw arrays are written by the scene kernels (same places in the slot's device buffer), nothing is uploaded
    SceneRec* scn = static_cast<SceneRec*>(plan.scene);
    const size_t szObj = sizeof(DObject) * n, szBase = 4 * (n + 1);
    const unsigned nChunks = scn ? scn->nClusters : plan.clusterBase[n], nSeqSlots = scn ? scn->nSeqSlots : plan.seqSlotBase[n];
    const uint32_t passes = 1u + (uint32_t)plan.shadowLayers;
    const size_t szShadow = (size_t)plan.shadowLayers * n * 64, szSlot = (size_t)passes * n * 4;
    auto align16 = [](size_t v) { return (v + 15) & ~(size_t)15; };
    const size_t offObj = 0, offBase = align16(offObj + szObj), offSeq = align16(offBase + szBase), offShadow = align16(offSeq + szBase),
                 offSlot = align16(offShadow + szShadow);
    const size_t total = align16(offSlot + szSlot) + 16;
    if (slot.chunkCap < nChunks || slot.seqSlotCap < nSeqSlots || slot.workCap < (size_t)(passes + 1) * nChunks || slot.eyeCap < (size_t)passes * n) {
        // a table has to grow: the frame that last used this slot must have finished with the old one before it is released
        if (dev->frameIndex >= VKR_FRAME_OVERLAP) CK(cudaEventSynchronize(slot.rendered));
        if (slot.chunkCap < nChunks) {
            dev_free(dev, slot.chunkBuf); slot.chunkBuf = nullptr;
            slot.chunkCap = std::max<size_t>((size_t)nChunks * 3 / 2, 1024);
            ALLOC(slot.chunkBuf, slot.chunkCap * sizeof(DChunk));
        }
        if (slot.seqSlotCap < nSeqSlots) {
            dev_free(dev, slot.chunkObjBuf); slot.chunkObjBuf = nullptr;
            slot.seqSlotCap = std::max<size_t>((size_t)nSeqSlots * 3 / 2, 1024);
            ALLOC(slot.chunkObjBuf, slot.seqSlotCap * sizeof(uint2));
        }
        if (slot.eyeCap < (size_t)passes * n) {
            dev_free(dev, slot.eyeBuf); slot.eyeBuf = nullptr;
            slot.eyeCap = std::max<size_t>((size_t)passes * n * 3 / 2, 256);
            ALLOC(slot.eyeBuf, slot.eyeCap * 8 * sizeof(float));
        }
        if (slot.workCap < (size_t)(passes + 1) * nChunks) {
            dev_free(dev, slot.workBuf); slot.workBuf = nullptr;
            slot.workCap = std::max<size_t>((size_t)(passes + 1) * nChunks * 3 / 2, 1024);
            ALLOC(slot.workBuf, slot.workCap * sizeof(WorkItem));
        }
    }
    if (dev->frameIndex >= VKR_FRAME_OVERLAP) {      // pinned buffer reuse: this is where a submit thread that runs ahead waits for the device
        const auto w0 = std::chrono::steady_clock::now();
        CK(cudaEventSynchronize(slot.uploaded));
        if (dev->hostProf && dev->frameIndex >= 16) {
            const auto w1 = std::chrono::steady_clock::now();
            dev->hostProfUs[8] += std::chrono::duration<double, std::micro>(w1 - w0).count();
            hostT += w1 - w0;                        // not counted as issue time of the section
        }
    }
    int rc = ensure_slot(dev, slot, total);
    if (rc) return rc;
    char* hp = (char*)slot.pinned;
    char* dp = (char*)slot.devbuf;
    // the upload rides its own stream so it overlaps the previous frame's kernels; the slot's device
    // buffer is free once the frame that last used it (FRAME_OVERLAP frames ago) has rendered
    CK(cudaStreamWaitEvent(dev->upStream, slot.rendered, 0));
    const uint32_t* dClusterBase = (const uint32_t*)(dp + offBase);
    const uint32_t* dSeqSlotBase = (const uint32_t*)(dp + offSeq);
    if (!scn) {
        if (n) std::memcpy(hp + offObj, plan.objects.data(), szObj);
        std::memcpy(hp + offBase, plan.clusterBase.data(), szBase);
        std::memcpy(hp + offSeq, plan.seqSlotBase.data(), szBase);
        if (szShadow) std::memcpy(hp + offShadow, plan.shadowMvp.data(), szShadow);
        if (szSlot) std::memcpy(hp + offSlot, plan.slotBase.data(), szSlot);
        CK(cudaMemcpyAsync(slot.devbuf, slot.pinned, total, cudaMemcpyHostToDevice, dev->upStream));
        dev->uploadBytes = total;
    } else if (n) {
        // update_scene on the device (vkr_scene_kernels.cuh): world matrices if a transform changed, then per draw is_visible,
        // MVPs and order keys, the passes' processing order, and the prefix sums — into the places the upload would have filled
        cudaStream_t up = dev->upStream;
        if (scn->worldDirty) {
            for (size_t l = 0; l + 1 < scn->levelFirst.size(); ++l) {
                const uint32_t first = scn->levelFirst[l], count = scn->levelFirst[l + 1] - first;
                if (!count) continue;
                scene_nodes_kernel<<<(count + 127) / 128, 128, 0, up>>>(scn->dLocal, scn->dParent, scn->dWorld, scn->dLevelOrder, first, count);
                dev->launches += 1;
            }
            scn->worldDirty = false;
        }
        const bool buckets = scn->nBlocks != 0;          // many draws: bucket order instead of the counting rank
        if (buckets) CK(cudaMemsetAsync(scn->dKeyRange, 0, 2 * 4 * (1 + VKR_NUM_CASCADES), up));
        scene_draws_kernel<<<(unsigned)((n + 127) / 128), 128, 0, up>>>(plan.sceneFrame, scn->dDraws, scn->dWorld, scn->dTmpl, (DObject*)(dp + offObj),
                                                                        (float*)(dp + offShadow), scn->dKeys, scn->dVisTris, buckets ? scn->dKeyRange : nullptr);
        if (buckets) {
            scene_bucket_count_kernel<<<dim3(scn->nBlocks, passes), 256, 0, up>>>(scn->dKeys, (uint32_t)n, scn->opaqueCount, scn->nBlocks, scn->dKeyRange, scn->dBucketOf, scn->dHist);
            scene_bucket_scan_kernel<<<passes, 1024, 0, up>>>(scn->dHist, scn->nBlocks);
            scene_bucket_place_kernel<<<dim3(scn->nBlocks, passes), 256, 0, up>>>((uint32_t)n, scn->opaqueCount, scn->nBlocks, scn->dBucketOf, scn->dHist, scn->dOrder);
            dev->launches += 2;
        } else
            scene_rank_kernel<<<dim3((unsigned)((n + 255) / 256), passes), 256, 0, up>>>(scn->dKeys, (uint32_t)n, scn->opaqueCount, scn->dOrder);
        scene_scan_kernel<<<1, 1024, 0, up>>>(plan.sceneFrame, scn->dDraws, (DObject*)(dp + offObj), scn->dVisTris, scn->dOrder, (uint32_t*)(dp + offSlot),
                                              dev->countersHost + 13);
        LAUNCHED("scene kernels");
        dev->launches += 3;
        dClusterBase = scn->dClusterBase; dSeqSlotBase = scn->dSeqSlotBase;
    } else {                                     // a scene without a single surface: nothing runs that would count
        CK(cudaStreamSynchronize(dev->upStream));
        dev->countersHost[13] = dev->countersHost[14] = 0u;
    }
    const DObject* dObjects = (const DObject*)(dp + offObj);
    const DChunk* dChunks = slot.chunkBuf;
    const float* dShadowMvp = (const float*)(dp + offShadow);
    const WorkItem* dWork = slot.workBuf;
    if (n) {   // the frame's tables, on the upload stream too: neither the main nor the shadow stream has to wait for the other
        prepare_draws_kernel<<<(unsigned)((n + 127) / 128), 128, 0, dev->upStream>>>(dObjects, dSeqSlotBase, dShadowMvp, (uint32_t)n, passes,
                                                                                     slot.eyeBuf, slot.chunkObjBuf);
        if (nChunks)
            expand_chunks_kernel<<<(nChunks + 127) / 128, 128, 0, dev->upStream>>>(dObjects, dClusterBase, dSeqSlotBase, (const uint32_t*)(dp + offSlot),
                                                                                   dShadowMvp, slot.eyeBuf, (uint32_t)n, passes, nChunks, plan.mainTarget, plan.shadowTarget,
                                                                                   slot.chunkBuf, slot.workBuf);
        dev->launches += 2;
    }
    CK(cudaEventRecord(slot.uploaded, dev->upStream));
    CK(cudaStreamWaitEvent(st, slot.uploaded, 0));
    if (timing) CK(cudaEventRecord(dev->tev[1], st));
    mark(1);

    const unsigned largeGrid = (unsigned)dev->smCount * VKR_LARGE_MIN_CTAS;      // every CTA resident: the cursor logic counts the grid's warps
    // sort-first over peer memory (see "sort-first exchange over PEER MEMORY"): this frame's number + 1 is what its flags carry
    const bool sharded = dev->shardRanks > 1;
    const bool p2p = sharded && dev->p2p && !dev->shadowExchange;
    const bool gatherFrame = sharded && !(dev->flags & VKR_CONFIG_LOCAL_PRESENT);
    const uint32_t F1 = (uint32_t)dev->shardFrame + 1u;
    const uint32_t allMask = sharded ? (uint32_t)((1ull << dev->shardRanks) - 1ull) : 0u, peersMask = allMask & ~(1u << dev->shardRank);
    if (p2p && gatherFrame && dev->shardRank == 0) {
        // rank 0 presents: once the image that this frame overwrites has been read back (if anybody asked), every rank may store
        // its rows of this frame into it
        if (slot.copyPending) CK(cudaStreamWaitEvent(dev->gatherStream, slot.copied, 0));
        PeerList pl{}; pl.n = 0;
        for (int r = 0; r < dev->shardRanks; ++r) pl.p[pl.n++] = &dev->peerFlags[r]->gatheredFree;
        shard_signal_kernel<<<1, 32, 0, dev->gatherStream>>>(pl, F1);
        LAUNCHED("shard_signal_kernel");
    }

    // ---- draw_shadow (src/vk_engine_run.cpp:334-382) on the shadow stream, concurrently with the main
    // pass's rasterisation: both are latency-bound, together they fill the SMs' issue slots
    cudaStream_t ss = dev->shadowStream;
    if (plan.shadowLayers) {
        // the shadow pass depends on this frame's upload only (and, through it, on the frame that last used this slot's shadow
        // map and tables): it may run ahead of the main stream, which is still shading the previous frame
        CK(cudaStreamWaitEvent(ss, slot.uploaded, 0));
        dev->shadowBits = dev->shadowBuf[slotIdx];
        if (timing) CK(cudaEventRecord(dev->tev[7], ss));
        Queues qs = dev->qShadow;
        qs.counters = dev->qShadow.counters + 8 * slotIdx;
        const size_t sn = (size_t)dev->S * dev->S;
        // every layer is cleared before anything else: a sharded device then tells its peers that their rows may land.  Over peer
        // memory that was normally done a frame ago (below, in front of the previous frame's main pass): the peers' pushes, which
        // have to wait for the clear, then never find it missing
        if (p2p && dev->preclearedFor[slotIdx] == F1 && dev->preclearedLayers[slotIdx] >= (uint32_t)plan.shadowLayers) {
            CK(cudaStreamWaitEvent(ss, dev->preclearEv[slotIdx], 0));
        } else {
            fill_u32_kernel<<<grid_for(sn * plan.shadowLayers / 4, 256, 2048), 256, 0, ss>>>(dev->shadowBits, sn * plan.shadowLayers, DEPTH_ONE_BITS);
            dev->launches += 1;
            if (p2p) {
                PeerList pl{}; pl.n = 0;
                for (int r = 0; r < dev->shardRanks; ++r) if (r != dev->shardRank) pl.p[pl.n++] = &dev->peerFlags[r]->shadowCleared[slotIdx][dev->shardRank];
                shard_signal_kernel<<<1, 32, 0, ss>>>(pl, F1);
                LAUNCHED("shard_signal_kernel");
            }
        }
        for (int layer = 0; layer < plan.shadowLayers; ++layer) {
            RasterTarget rt = plan.shadowTarget;
            rt.depthBits = dev->shadowBits + (size_t)layer * sn;
            if (layer == 0) CK(cudaMemsetAsync(qs.counters, 0, 32, ss));
            else {        // the queue and its cursor start over; the overflow flag and the culled count are the frame's
                CK(cudaMemsetAsync(qs.counters, 0, 8, ss));
                CK(cudaMemsetAsync(qs.counters + 4, 0, 4, ss));
            }
            const unsigned shadowSlots = plan.passSlots[1 + layer];
            if (shadowSlots) {
#if defined(VKR_HAVE_CTA_GEOMETRY)
                if (shadowSlots < dev->ctaClusters)
                    geometry_cta_kernel_shadow<<<shadowSlots, 128, 0, ss>>>(dWork + (size_t)(1 + layer) * nChunks, shadowSlots, dShadowMvp + (size_t)layer * n * 16, 16u, rt, qs);
                else
#endif
                geometry_kernel<true><<<(shadowSlots + GEOM_WARPS - 1) / GEOM_WARPS, GEOM_WARPS * 32, 0, ss>>>(dWork + (size_t)(1 + layer) * nChunks, shadowSlots,
                                                                                                      dShadowMvp + (size_t)layer * n * 16, 16u, rt, qs);
                large_raster_kernel<true><<<largeGrid, 256, 0, ss>>>(rt, qs, AlphaCtx{}, 0);
                dev->launches += 2;
            }
        }
        if (timing) CK(cudaEventRecord(dev->tev[10], ss));
        if (dev->shardRanks > 1) {      // complete every layer from the other ranks' row bands (only rows a caster can reach)
            auto clampRow = [](int32_t y, int lo, int hi) { return std::min(std::max(y, lo), hi); };
            if (p2p) {
                // peer memory: wait until every peer has cleared its map for this frame, store my rows into all of them, tell
                // them, and wait for theirs
                PeerList maps{}; maps.n = 0;
                for (int r = 0; r < dev->shardRanks; ++r) if (r != dev->shardRank) maps.p[maps.n++] = dev->peerShadow[r][slotIdx];
                for (int layer = 0; layer < plan.shadowLayers; ++layer) {
                    const int lo = layer == 0 ? plan.activeShadow[0] : 0, hi = layer == 0 ? plan.activeShadow[1] : (int)dev->S;
                    const int32_t m0 = clampRow(plan.bandShadow[dev->shardRank], lo, hi), m1 = clampRow(plan.bandShadow[dev->shardRank + 1], lo, hi);
                    if (m1 <= m0) continue;
                    const size_t first = (size_t)layer * sn + (size_t)m0 * dev->S, count = (size_t)(m1 - m0) * dev->S;
                    shard_push_rows_kernel<<<grid_for(count / 4, 256, (unsigned)dev->smCount * 4), 256, 0, ss>>>(dev->shadowBits, maps, first, count,
                                                                                                                 dev->myFlags->shadowCleared[slotIdx], peersMask, F1, dev->countersHost + 12);
                    LAUNCHED("shard_push_rows_kernel");
                    dev->launches += 1;
                }
                PeerList pl{}; pl.n = 0;
                for (int r = 0; r < dev->shardRanks; ++r) if (r != dev->shardRank) pl.p[pl.n++] = &dev->peerFlags[r]->shadowReady[dev->shardRank];
                shard_signal_wait_kernel<<<1, 32, 0, ss>>>(pl, dev->myFlags->shadowReady, peersMask, F1, dev->countersHost + 12, 2u);
                LAUNCHED("shard_signal_wait_kernel");
            } else {
                for (int layer = 0; layer < plan.shadowLayers; ++layer) {
                    uint32_t* base = dev->shadowBits + (size_t)layer * sn;
                    // layer 0's casters define the active rows; other cascades see the scene at another scale: exchange them whole
                    const int lo = layer == 0 ? plan.activeShadow[0] : 0, hi = layer == 0 ? plan.activeShadow[1] : (int)dev->S;
                    if (exchange_bands(base, base, plan.bandShadow, lo, hi, dev->S, dev->shardRank, dev->shardRanks, dev->commShadow, ss) != ncclSuccess)
                        return dev->fail(VKR_ERROR_CUDA, "NCCL shadow band exchange");
                }
            }
        }
        if (dev->shadowExchange) {
            if (dev->shadowExchange(dev->shadowExchangeUser, dev->shadowBits, (size_t)dev->S * dev->S * 4, plan.shadowLayers, (void*)ss) != 0)
                return dev->fail(VKR_ERROR_CUDA, "shadow exchange hook failed");
        }
        if (timing) CK(cudaEventRecord(dev->tev[8], ss));
        CK(cudaEventRecord(dev->evShadow, ss));
        if (p2p) {
            // The NEXT frame's shadow map (the other slot's) is cleared here, a frame ahead, behind this frame's shadow work, and
            // the peers are told: its last reader, the previous frame's shading pass, has ended (`rendered`), and the peers'
            // pushes of the next frame, which must not land before the clear, are a frame away — so they never wait for it
            // (cleared at the start of its own frame, every rank's push waited for the slowest rank's clear: a second
            // rendezvous per frame).
            const int other = slotIdx ^ 1;
            const size_t texels = (size_t)dev->S * dev->S * plan.shadowLayers;
            if (dev->frameIndex > 0) CK(cudaStreamWaitEvent(ss, dev->slots[other].rendered, 0));
            fill_u32_kernel<<<grid_for(texels / 4, 256, 2048), 256, 0, ss>>>(dev->shadowBuf[other], texels, DEPTH_ONE_BITS);
            dev->launches += 1;
            PeerList pl{}; pl.n = 0;
            for (int r = 0; r < dev->shardRanks; ++r) if (r != dev->shardRank) pl.p[pl.n++] = &dev->peerFlags[r]->shadowCleared[other][dev->shardRank];
            shard_signal_kernel<<<1, 32, 0, ss>>>(pl, F1 + 1u);
            LAUNCHED("shard_signal_kernel (shadow map cleared ahead)");
            CK(cudaEventRecord(dev->preclearEv[other], ss));
            dev->preclearedFor[other] = F1 + 1u;
            dev->preclearedLayers[other] = (uint32_t)plan.shadowLayers;
        }
    }

    mark(2);
    // ---- draw_geometry, raster half (src/vk_engine_run.cpp:244-311)
    {
        RasterTarget rt = plan.mainTarget;
        rt.vis = dev->vis;
        const size_t px = (size_t)dev->W * dev->H;
        // a transient attachment was left cleared by the previous frame's shading pass (every rasterised pixel lies in the
        // scissor rectangle that pass walks); otherwise loadOp CLEAR is a pass of its own
        if (!dev->visClean) {
            fill_u64_kernel<<<grid_for(px / 2, 256, 4096), 256, 0, st>>>(dev->vis, px, VIS_CLEAR);
            dev->launches += 1;
        }
        CK(cudaMemsetAsync(dev->q.counters, 0, 32, st));
        if (timing) CK(cudaEventRecord(dev->tev[3], st));
        cudaEvent_t* kpair = nullptr;
        if (timeKernel) {
            kpair = dev->kev[dev->kevRecorded % VkrDevice::KRING];
            if (!kpair[0]) { CK(cudaEventCreate(&kpair[0])); CK(cudaEventCreate(&kpair[1])); }
            CK(cudaEventRecord(kpair[0], st));
        }
        // only the opaque prefix writes depth; transparent draws are blended in a later pass
        const unsigned opaqueChunks = plan.opaqueClusters, mainSlots = plan.passSlots[0];
        if (opaqueChunks) {
            const AlphaCtx ax{dObjects, slot.chunkObjBuf, dev->matDev, dev->texDev};
#if !defined(VKR_CUTOUT_LAST)
            // alpha-tested clusters first, so that their items — the heaviest — are at the FRONT of the large-triangle queue (a dynamic
            // cursor wants the heaviest items first: config 3 801 -> 848 frames/s; -DVKR_CUTOUT_LAST restores round 1 order)
            if (plan.anyCutout) {
                geometry_cutout_kernel<<<opaqueChunks, CL_TRIS, 0, st>>>(dObjects, dChunks, rt, dev->q, ax);
                dev->launches += 1;
            }
#endif
            if (mainSlots) {
#if defined(VKR_HAVE_CTA_GEOMETRY)
                if (mainSlots < dev->ctaClusters)
                    geometry_cta_kernel_main<<<mainSlots, 128, 0, st>>>(dWork, mainSlots, (const float*)dObjects, (uint32_t)(sizeof(DObject) / 4), rt, dev->q);
                else
#endif
                geometry_kernel<false><<<(mainSlots + GEOM_WARPS - 1) / GEOM_WARPS, GEOM_WARPS * 32, 0, st>>>(dWork, mainSlots, (const float*)dObjects,
                                                                                                            (uint32_t)(sizeof(DObject) / 4), rt, dev->q);
            }
            if (timing) CK(cudaEventRecord(dev->tev[4], st));
            if (kpair) { CK(cudaEventRecord(kpair[1], st)); dev->kevRecorded++; }
#if defined(VKR_CUTOUT_LAST)
            if (plan.anyCutout) {
                geometry_cutout_kernel<<<opaqueChunks, CL_TRIS, 0, st>>>(dObjects, dChunks, rt, dev->q, ax);
                dev->launches += 1;
            }
#endif
            large_raster_kernel<false><<<largeGrid, 256, 0, st>>>(rt, dev->q, ax, plan.anyCutout ? 1 : 0);
            dev->launches += 2;
        } else if (timing) CK(cudaEventRecord(dev->tev[4], st));
    }
    if (timing) CK(cudaEventRecord(dev->tev[5], st));
    mark(3);

    if (plan.shadowLayers) CK(cudaStreamWaitEvent(st, dev->evShadow, 0));   // join: the shaders sample the shadow map
    ShadeBuffers sb{};
    sb.vis = dev->vis; sb.objects = dObjects; sb.chunkObj = slot.chunkObjBuf; sb.materials = dev->matDev;
    sb.textures = dev->texDev; sb.cube = dev->haveCube ? dev->cubeDev : nullptr;
    sb.shadow = reinterpret_cast<const float*>(dev->shadowBits);
    sb.hdr = dev->hdr; sb.rgba8 = dev->rgba8[slotIdx]; sb.srgbLut = dev->lut;
    sb.accum = nullptr;
    const bool transientDepth = (dev->flags & VKR_CONFIG_TRANSIENT_DEPTH) != 0;
    plan.sc.clearVis = transientDepth ? 1 : 0;
    plan.sc.storeHdr = (dev->flags & VKR_CONFIG_TRANSIENT_HDR) ? 0 : 1;
    const ShadeConst& sc = plan.sc;

    // ---- "Transparent" draws: depth-tested against the finished opaque depth, shaded and summed (BlendSink)
    const unsigned transChunks = nChunks - plan.opaqueClusters;
    if (transChunks) {
        const size_t px = (size_t)dev->W * dev->H;
        if (!dev->accum) ALLOC(dev->accum, px * 16);
        sb.accum = dev->accum;
        CK(cudaMemsetAsync(dev->accum, 0, px * 16, st));
        Queues qt = dev->q;                          // the main pass's queues (drained by now), this pass's own counters
        qt.counters = dev->transCounters;
        CK(cudaMemsetAsync(qt.counters, 0, 32, st));
        RasterTarget rt = plan.mainTarget;
        rt.vis = dev->vis;
        geometry_transparent_kernel<<<(transChunks + GEOM_WARPS - 1) / GEOM_WARPS, GEOM_WARPS * 32, 0, st>>>(
            dWork + (size_t)passes * nChunks + plan.opaqueClusters, transChunks, (const float*)dObjects, (uint32_t)(sizeof(DObject) / 4), rt, qt, sc, sb);
        large_transparent_kernel<<<largeGrid, 256, 0, st>>>(rt, qt, sc, sb);
        dev->launches += 2;
    }
    sb.counters[0] = dev->q.counters;
    sb.counters[1] = plan.shadowLayers ? dev->qShadow.counters + 8 * slotIdx : nullptr;
    sb.counters[2] = transChunks ? dev->transCounters : nullptr;
    sb.countersHost = dev->countersHost;
    if (timing) CK(cudaEventRecord(dev->tev[9], st));

    // ---- shading half of draw_geometry + draw_skybox + draw_background + draw_postprocess + present blit
    if (sharded && !p2p && dev->gatheredValid[slotIdx]) CK(cudaStreamWaitEvent(st, dev->gatheredEv[slotIdx], 0));   // rgba8[slot] still being gathered
    if (slot.copyPending) CK(cudaStreamWaitEvent(st, slot.copied, 0));      // ... or still being read back (FRAME_OVERLAP frames ago)
    if (p2p && gatherFrame) {
        // the shading kernel stores this rank's rows into rank 0's image as well: wait until rank 0 has released it
        sb.rgba8Peer = dev->rootGathered[slotIdx];
        sb.peerFree = &dev->myFlags->gatheredFree; sb.peerFreeValue = F1;        // (every CTA of the shading kernel checks it before its first store)
    }
    {
        dim3 grid((sc.sx1 - sc.sx0 + 15) / 16, (sc.sy1 - sc.sy0 + 7) / 8);
        if (grid.x && grid.y) { shade_kernel<<<grid, 128, 0, st>>>(sc, sb); dev->launches += 1; }
        else {                                       // nothing to shade (empty rectangle): the counters still reach the host
            for (int k = 0; k < 3; ++k)
                if (sb.counters[k]) CK(cudaMemcpyAsync(dev->countersHost + 4 * k, sb.counters[k], 16, cudaMemcpyDeviceToHost, st));
                else std::memset(dev->countersHost + 4 * k, 0, 16);
        }
        dev->visClean = transientDepth;
    }
    if (timing) { CK(cudaEventRecord(dev->tev[6], st)); dev->tevValid = true; dev->shadowTimed = plan.shadowLayers > 0; }
    if (p2p && gatherFrame) {           // my rows of this frame are in rank 0's image
        PeerList pl{}; pl.n = 1; pl.p[0] = &dev->peerFlags[0]->frameReady[dev->shardRank];
        shard_signal_kernel<<<1, 32, 0, st>>>(pl, F1);
        LAUNCHED("shard_signal_kernel");
        dev->lastGatherF1 = F1;
    } else if (sharded) dev->lastGatherF1 = 0;
    CK(cudaGetLastError());
    CK(cudaEventRecord(slot.rendered, st));
    mark(4);
    if (sharded && !p2p && gatherFrame) {          // assemble the presented image; overlaps the next frame's rendering
        CK(cudaStreamWaitEvent(dev->gatherStream, slot.rendered, 0));
        if (slot.copyPending) CK(cudaStreamWaitEvent(dev->gatherStream, slot.copied, 0));   // the landing buffer may still be being read back
        if (timing) CK(cudaEventRecord(dev->tev[11], dev->gatherStream));
        if (exchange_bands(dev->rgba8[slotIdx], dev->gathered[slotIdx], plan.bandMain, 0, (int)dev->H, dev->W, dev->shardRank, dev->shardRanks,
                           dev->commFrame, dev->gatherStream) != ncclSuccess)
            return dev->fail(VKR_ERROR_CUDA, "NCCL frame band gather");
        if (timing) CK(cudaEventRecord(dev->tev[12], dev->gatherStream));
        CK(cudaEventRecord(dev->gatheredEv[slotIdx], dev->gatherStream));
        dev->gatheredValid[slotIdx] = true;
    }
    mark(5);
    if (dev->hostProf && dev->frameIndex >= 16) dev->hostProfFrames++;
    dev->lastSlot = slotIdx;
    if (sharded) { dev->shardFrame++; dev->lastBand[0] = plan.bandMain[dev->shardRank]; dev->lastBand[1] = plan.bandMain[dev->shardRank + 1]; }
    else { dev->lastBand[0] = 0; dev->lastBand[1] = (int32_t)dev->H; }
    dev->lastPlan = &plan; dev->lastN = n;
    dev->lastChunkObj = slot.chunkObjBuf; dev->lastObjects = dObjects;
    dev->frameIndex++;
    dev->stats.drawcall_count = (int32_t)n;
    dev->stats.triangle_count = (int32_t)plan.totalTris;
    dev->lastOpaqueClusters = plan.passSlots[0];
    dev->lastShadowClusters = 0;                    // over all layers, like the culled count
    for (int layer = 0; layer < plan.shadowLayers; ++layer) dev->lastShadowClusters += plan.passSlots[1 + layer];
    return VKR_SUCCESS;
}

extern "C" {

int vkr_sync(VkrDevice* dev) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    CK(cudaSetDevice(dev->device));
    CK(cudaStreamSynchronize(dev->stream));
    CK(cudaStreamSynchronize(dev->copyStream));
    CK(cudaStreamSynchronize(dev->shadowStream));
    if (dev->gatherStream) CK(cudaStreamSynchronize(dev->gatherStream));
    if (dev->p2p && dev->shardRank == 0 && dev->lastGatherF1) {      // "everything issued" includes the peers' rows of the last frame
        shard_wait_kernel<<<1, 32, 0, dev->copyStream>>>(dev->myFlags->frameReady, (uint32_t)((1ull << dev->shardRanks) - 1ull), dev->lastGatherF1, dev->countersHost + 12, 4u);
        LAUNCHED("shard_wait_kernel (vkr_sync)");
        CK(cudaStreamSynchronize(dev->copyStream));
    }
    if (dev->countersHost[12]) {
        char msg[128];
        std::snprintf(msg, sizeof msg, "peer exchange timed out (wait %u on rank %u)", dev->countersHost[12] & 0xFFu, dev->countersHost[12] >> 8);
        dev->countersHost[12] = 0;
        return dev->fail(VKR_ERROR_CUDA, msg);
    }
    // A large-triangle queue overflowed in the last frame (its image misses those triangles): the counters kept counting, so
    // they say how much room the frame needs.  Grow the queues and draw the frame again before returning; a read-back that
    // follows sees the complete frame.  Only an asynchronous read-back ALREADY issued for the incomplete frame cannot be
    // repaired: that is reported once as VKR_ERROR_CAPACITY (the queues have grown: drawing again succeeds).
    for (int attempt = 0; attempt < 4 && (dev->countersHost[2] || dev->countersHost[6] || dev->countersHost[10]); ++attempt) {
        if (dev->shardRanks > 1 || !dev->lastPlan) return dev->fail(VKR_ERROR_CAPACITY, "large-triangle queue overflow");   // (a sharded redraw would need every rank to take part)
        const bool stale = dev->slots[dev->lastSlot].copiedFrame == dev->frameIndex;   // (synchronous read-backs sync first: see do_readback)
        // a pass whose flag is set gets at least twice the room, and what its counters ask for plus a half (the shadow pass's
        // counters are those of its last layer only, hence the doubling)
        auto grow = [&](Queues& qq, uint32_t needTris, uint32_t needItems) -> int {
            dev_free(dev, qq.tris); qq.tris = nullptr;
            qq.triCap = std::max(needTris + needTris / 2, 2 * qq.triCap);
            ALLOC(qq.tris, sizeof(LargeTri) * (size_t)qq.triCap);
            dev_free(dev, qq.items); qq.items = nullptr;
            qq.itemCap = std::max(needItems + needItems / 2, 2 * qq.itemCap);
            ALLOC(qq.items, sizeof(LargeItem) * (size_t)qq.itemCap);
            return VKR_SUCCESS;
        };
        int rc = VKR_SUCCESS;
        if (dev->countersHost[2] || dev->countersHost[10])      // main and transparent passes share dev->q
            rc = grow(dev->q, std::max(dev->countersHost[0], dev->countersHost[8]), std::max(dev->countersHost[1], dev->countersHost[9]));
        if (rc == VKR_SUCCESS && dev->countersHost[6]) rc = grow(dev->qShadow, dev->countersHost[4], dev->countersHost[5]);
        if (rc != VKR_SUCCESS) return rc;
        std::memset(dev->countersHost, 0, 48);
        rc = submit_frame(dev, *dev->lastPlan, dev->lastN);
        if (rc != VKR_SUCCESS) return rc;
        CK(cudaStreamSynchronize(dev->stream));
        CK(cudaStreamSynchronize(dev->copyStream));
        CK(cudaStreamSynchronize(dev->shadowStream));
        if (stale && !dev->overflowReported) {
            dev->overflowReported = true;
            return dev->fail(VKR_ERROR_CAPACITY, "large-triangle queue overflow: an asynchronous read-back saw the incomplete frame (queues grown, frame redrawn)");
        }
    }
    if (dev->countersHost[2] || dev->countersHost[6] || dev->countersHost[10]) return dev->fail(VKR_ERROR_CAPACITY, "large-triangle queue overflow");
    if (dev->kevRecorded > dev->kevRead) {      // mean main-pass geometry kernel time over the frames since the last sync
        const uint64_t first = std::max(dev->kevRead, dev->kevRecorded > (uint64_t)VkrDevice::KRING ? dev->kevRecorded - VkrDevice::KRING : 0);
        double sum = 0.0;
        for (uint64_t k = first; k < dev->kevRecorded; ++k) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, dev->kev[k % VkrDevice::KRING][0], dev->kev[k % VkrDevice::KRING][1]) != cudaSuccess) { cudaGetLastError(); ms = 0; }
            sum += ms;
        }
        dev->timings[VKR_T_GEOM_KERNEL] = (float)(sum / (double)(dev->kevRecorded - first));
        dev->kevRead = dev->kevRecorded;
    }
    if (dev->tevValid) {
        // (an event pair that this frame's configuration never recorded reads as 0 and must not leave an error behind)
        auto el = [&](int a, int b) { float ms = 0; if (cudaEventElapsedTime(&ms, dev->tev[a], dev->tev[b]) != cudaSuccess) { cudaGetLastError(); ms = 0; } return ms; };
        dev->timings[VKR_T_UPLOAD] = el(0, 1);
        dev->timings[VKR_T_SHADOW] = dev->shadowTimed ? el(7, 8) : 0.0f;   // overlaps the geometry pass
        dev->timings[VKR_T_GEOMETRY] = el(1, 5);
        dev->timings[VKR_T_CLEAR] = el(1, 3);
        dev->timings[VKR_T_GEOM_KERNEL] = el(3, 4);
        dev->timings[VKR_T_LARGE_KERNEL] = el(4, 5);
        dev->timings[VKR_T_SHADE] = el(9, 6);
        dev->timings[VKR_T_TRANSPARENT] = el(5, 9); dev->timings[VKR_T_POST] = 0;   // tonemap + present are fused into the shade kernel
        dev->timings[VKR_T_TOTAL] = el(0, 6);
        dev->timings[VKR_T_SHADOW_EXCHANGE] = dev->shadowTimed ? el(10, 8) : 0.0f;
        dev->timings[VKR_T_FRAME_GATHER] = (dev->shardRanks > 1 && !dev->p2p) ? el(11, 12) : 0.0f;     // (over peer memory the shading kernel's stores are the gather)
        dev->timings[VKR_T_READBACK] = dev->rbTimed ? el(13, 14) : 0.0f;
        dev->stats.frametime = dev->timings[VKR_T_TOTAL];
        dev->stats.mesh_draw_time = dev->timings[VKR_T_SHADOW] + dev->timings[VKR_T_GEOMETRY];
    }
    return VKR_SUCCESS;
}

static int target_info(VkrDevice* dev, VkrTarget what, const void** src, size_t* bytes) {
    const size_t px = (size_t)dev->W * dev->H;
    switch (what) {
        case VKR_TARGET_RGBA8: *src = dev->rgba8[dev->lastSlot]; *bytes = px * 4; return 0;
        case VKR_TARGET_RGBA8_GATHERED: case VKR_TARGET_RGB8:
            // over peer memory only rank 0 holds the assembled frame, and only if the last frame was assembled at all
            if (dev->shardRanks > 1 && dev->p2p && (dev->shardRank != 0 || dev->lastGatherF1 == 0)) return -1;
            *src = dev->shardRanks <= 1 ? dev->rgba8[dev->lastSlot] : dev->gathered[dev->lastSlot];
            *bytes = what == VKR_TARGET_RGB8 ? px * 3 : px * 4;
            return 0;
        case VKR_TARGET_HDR16F: *src = dev->hdr; *bytes = px * 8; return 0;
        case VKR_TARGET_SHADOW: *src = dev->shadowBits; *bytes = (size_t)dev->S * dev->S * 4 * VKR_NUM_CASCADES; return 0;
        case VKR_TARGET_DEPTH: case VKR_TARGET_PRIMID: *src = nullptr; *bytes = px * 4; return 0;
        default: return -1;
    }
}

// The copy-stream half of a read-back (colour targets and the shadow map): runs on whichever thread issues CUDA work for the
// device, after the frame it reads (dev->lastSlot is that frame's slot).
static int readback_issue(VkrDevice* dev, VkrTarget what, void* dst, size_t dstBytes) {
    const void* src; size_t bytes;
    if (target_info(dev, what, &src, &bytes) || dstBytes < bytes) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback: bad target");
    // copy engine work goes to the copy stream so the next frame's kernels overlap it
    CK(cudaStreamWaitEvent(dev->copyStream, dev->slots[dev->lastSlot].rendered, 0));
    if ((what == VKR_TARGET_RGBA8_GATHERED || what == VKR_TARGET_RGB8) && dev->shardRanks > 1) {
        if (dev->p2p) {    // every rank's rows of that frame have landed in my image
            shard_wait_kernel<<<1, 32, 0, dev->copyStream>>>(dev->myFlags->frameReady, (uint32_t)((1ull << dev->shardRanks) - 1ull), dev->lastGatherF1, dev->countersHost + 12, 4u);
            LAUNCHED("shard_wait_kernel (read-back)");
        } else CK(cudaStreamWaitEvent(dev->copyStream, dev->gatheredEv[dev->lastSlot], 0));
    }
    if (what == VKR_TARGET_RGB8) {
        const size_t px = (size_t)dev->W * dev->H;
        uint32_t*& packed = dev->rgb8[dev->lastSlot];
        if (!packed) ALLOC(packed, px * 3 + 16);
        // Unsharded: packed on the frame stream, straight behind the frame's shading pass (whose wait on this slot's previous
        // copy also protects `packed`), so the transfer starts as soon as the frame is done.  Sharded: the source is the
        // gathered image, which the frame stream must not wait for, so the pack stays on the (high-priority) copy stream.
        cudaStream_t ps = dev->shardRanks > 1 ? dev->copyStream : dev->stream;
        pack_rgb8_kernel<<<(unsigned)(((px >> 2) + 256) / 256), 256, 0, ps>>>((const uint32_t*)src, packed, px);
        CK(cudaGetLastError());
        if (ps != dev->copyStream) {
            CK(cudaEventRecord(dev->packedEv, ps));
            CK(cudaStreamWaitEvent(dev->copyStream, dev->packedEv, 0));
        }
        src = packed;
    }
    const bool timing = (dev->flags & VKR_CONFIG_TIMINGS) != 0;
    if (timing) CK(cudaEventRecord(dev->tev[13], dev->copyStream));
    CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, dev->copyStream));
    if (timing) { CK(cudaEventRecord(dev->tev[14], dev->copyStream)); dev->rbTimed = true; }
    if (what == VKR_TARGET_RGBA8 || what == VKR_TARGET_RGBA8_GATHERED || what == VKR_TARGET_RGB8) {
        // the frame FRAME_OVERLAP draws from now writes this image again: it must wait for this copy
        FrameSlot& s = dev->slots[dev->lastSlot];
        CK(cudaEventRecord(s.copied, dev->copyStream));
        s.copyPending = true;
        s.copiedFrame = dev->frameIndex;
    }
    return VKR_SUCCESS;
}

static int do_readback(VkrDevice* dev, VkrTarget what, void* dst, size_t dstBytes, bool sync) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    const void* src; size_t bytes;
    const bool queued = !sync && dev->worker.joinable() && (dev->flags & VKR_CONFIG_SUBMIT_THREAD);
    // (a request that is queued behind its frame is checked again when the submit thread reaches it: what the last frame left
    // behind — e.g. whether it was assembled on this rank — is only known then; everything else waits for the frame to be issued)
    if (!queued) DRAIN(dev);
    if (!dst || (target_info(dev, what, &src, &bytes) && !(queued && (what == VKR_TARGET_RGBA8_GATHERED || what == VKR_TARGET_RGB8)))) {
        DRAIN(dev);
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback: bad target");
    }
    if (what == VKR_TARGET_RGBA8_GATHERED || what == VKR_TARGET_RGB8) bytes = (size_t)dev->W * dev->H * (what == VKR_TARGET_RGB8 ? 3 : 4);
    if (dstBytes < bytes) { DRAIN(dev); return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback: destination too small"); }
    const bool visTarget = what == VKR_TARGET_DEPTH || what == VKR_TARGET_PRIMID;
    // only the presented image is double-buffered and fenced against the frames that follow: the draw image is one buffer and
    // the shadow map is rewritten FRAME_OVERLAP frames later without waiting for a copy
    if (!sync && (what == VKR_TARGET_HDR16F || what == VKR_TARGET_SHADOW)) { DRAIN(dev); return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback_async: only the presented image (RGBA8 / RGBA8_GATHERED / RGB8) can be read back asynchronously"); }
    if (what == VKR_TARGET_HDR16F && (dev->flags & VKR_CONFIG_TRANSIENT_HDR)) { DRAIN(dev); return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback: the draw image is transient (VKR_CONFIG_TRANSIENT_HDR)"); }
    if (visTarget && (dev->flags & VKR_CONFIG_TRANSIENT_DEPTH)) { DRAIN(dev); return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback: the depth attachment is transient (VKR_CONFIG_TRANSIENT_DEPTH)"); }
    if (!sync && !visTarget && dev->worker.joinable() && (dev->flags & VKR_CONFIG_SUBMIT_THREAD)) {
        // behind the frame it reads, on the submit thread: the caller goes straight on to planning its next frame
        post_job(dev, VkrDevice::Job{1, nullptr, 0, what, dst, dstBytes});
        return VKR_SUCCESS;
    }
    DRAIN(dev);
    CK(cudaSetDevice(dev->device));
    if (sync) {     // the frame is complete (and, had a queue overflowed, redrawn) before it is copied
        const int rc0 = vkr_sync(dev);
        if (rc0 != VKR_SUCCESS) return rc0;
    }
    if (visTarget) {
        const size_t px = (size_t)dev->W * dev->H;
        if (!dev->scratch) ALLOC(dev->scratch, px * 4);
        extract_vis_kernel<<<(unsigned)((px + 255) / 256), 256, 0, dev->stream>>>(dev->vis, dev->scratch, px, what == VKR_TARGET_DEPTH ? 0 : 1,
                                                                                    dev->lastChunkObj, dev->lastObjects);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(dst, dev->scratch, bytes, cudaMemcpyDeviceToHost, dev->stream));
        if (sync) return vkr_sync(dev);
        return VKR_SUCCESS;
    }
    const int rc = readback_issue(dev, what, dst, dstBytes);
    if (rc != VKR_SUCCESS) return rc;
    if (sync) return vkr_sync(dev);
    return VKR_SUCCESS;
}

// This rank's rows of the presented image of the last frame, into their place in a whole-frame buffer at frameBase.
static int readback_band_issue(VkrDevice* dev, VkrTarget what, void* frameBase, size_t frameBytes) {
    const size_t bpp = what == VKR_TARGET_RGB8 ? 3 : 4, rowBytes = (size_t)dev->W * bpp;
    if ((what != VKR_TARGET_RGB8 && what != VKR_TARGET_RGBA8) || !frameBase || frameBytes < rowBytes * dev->H)
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_readback_band_async: RGBA8 or RGB8 into a whole-frame buffer");
    const int32_t y0 = dev->lastBand[0], y1 = dev->lastBand[1];
    FrameSlot& sl = dev->slots[dev->lastSlot];
    const uint32_t* src = dev->rgba8[dev->lastSlot] + (size_t)y0 * dev->W;
    const size_t px = (size_t)(y1 - y0) * dev->W;
    CK(cudaStreamWaitEvent(dev->copyStream, sl.rendered, 0));
    const void* from = src;
    if (what == VKR_TARGET_RGB8 && px) {
        const size_t all = (size_t)dev->W * dev->H;
        uint32_t*& packed = dev->rgb8[dev->lastSlot];
        if (!packed) ALLOC(packed, all * 3 + 16);
        // (the band starts at a multiple of W pixels: 4-pixel groups stay aligned when W is a multiple of 4, else pack the whole image)
        if ((dev->W & 3u) == 0u) {
            uint32_t* out = packed + (size_t)y0 * dev->W * 3 / 4;
            pack_rgb8_kernel<<<(unsigned)(((px >> 2) + 256) / 256), 256, 0, dev->stream>>>(src, out, px);
            from = out;
        } else {
            pack_rgb8_kernel<<<(unsigned)(((all >> 2) + 256) / 256), 256, 0, dev->stream>>>(dev->rgba8[dev->lastSlot], packed, all);
            from = reinterpret_cast<const uint8_t*>(packed) + (size_t)y0 * rowBytes;
        }
        CK(cudaGetLastError());
        CK(cudaEventRecord(dev->packedEv, dev->stream));
        CK(cudaStreamWaitEvent(dev->copyStream, dev->packedEv, 0));
    }
    if (px) CK(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(frameBase) + (size_t)y0 * rowBytes, from, px * bpp, cudaMemcpyDeviceToHost, dev->copyStream));
    CK(cudaEventRecord(sl.copied, dev->copyStream));
    sl.copyPending = true;
    sl.copiedFrame = dev->frameIndex;
    return VKR_SUCCESS;
}

int vkr_readback_band_async(VkrDevice* dev, VkrTarget what, void* frameBase, size_t frameBytes) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    if (dev->worker.joinable() && (dev->flags & VKR_CONFIG_SUBMIT_THREAD)) {
        post_job(dev, VkrDevice::Job{2, nullptr, 0, what, frameBase, frameBytes});
        return VKR_SUCCESS;
    }
    DRAIN(dev);
    CK(cudaSetDevice(dev->device));
    return readback_band_issue(dev, what, frameBase, frameBytes);
}

int vkr_readback(VkrDevice* dev, VkrTarget what, void* dst, size_t dstBytes) { return do_readback(dev, what, dst, dstBytes, true); }
int vkr_readback_async(VkrDevice* dev, VkrTarget what, void* dst, size_t dstBytes) { return do_readback(dev, what, dst, dstBytes, false); }

int vkr_stats(VkrDevice* dev, VkrEngineStats* out) {
    if (!dev || !out) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (dev->lastPlan && dev->lastPlan->scene) {       // vkr_draw_scene: the device counted what is_visible kept
        CK(cudaSetDevice(dev->device));
        CK(cudaStreamSynchronize(dev->upStream));
        dev->stats.drawcall_count = (int32_t)dev->countersHost[13];
        dev->stats.triangle_count = (int32_t)dev->countersHost[14];
    }
    *out = dev->stats;
    return VKR_SUCCESS;
}

int vkr_set_flags(VkrDevice* dev, uint32_t flags) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!(flags & VKR_CONFIG_TIMINGS)) dev->tevValid = false;
    dev->flags = flags;      // (visClean stays as the last frame left it: a frame drawn without the flag resets it)
    return VKR_SUCCESS;
}

int vkr_get_timings(VkrDevice* dev, float* msOut, int count) {
    if (!dev || !msOut || count < 0) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    for (int i = 0; i < count && i < VKR_T_COUNT; ++i) msOut[i] = dev->timings[i];
    return VKR_SUCCESS;
}

// Mean host time per frame (microseconds) of vkr_draw's sections since the last reset: [0] plan_frame on the caller's
// thread, [1] upload + expand, [2] shadow pass (+ exchange), [3] main raster, [4] transparent + shading, [5] gather,
// [6] asynchronous read-backs, [7] caller waiting for the submit thread, [8] submit thread waiting for the device (pinned
// buffer reuse).  [1..6] are the submit thread's issue time.  reset != 0 starts a new measurement (and switches it on).
int vkr_host_times(VkrDevice* dev, float* usOut, int count, int reset) {
    if (!dev || (count > 0 && !usOut) || count < 0) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    const double f = dev->hostProfFrames ? 1.0 / (double)dev->hostProfFrames : 0.0;
    for (int i = 0; i < count && i < 9; ++i) usOut[i] = (float)(dev->hostProfUs[i] * f);
    if (reset) {
        for (double& v : dev->hostProfUs) v = 0.0;
        dev->hostProfFrames = 0;
        dev->hostProf = true;
    }
    return VKR_SUCCESS;
}

int vkr_device_ptr(VkrDevice* dev, VkrTarget what, void** ptr, size_t* bytes) {
    if (!dev || !ptr || !bytes) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    const void* src; size_t n;
    if (target_info(dev, what, &src, &n) || !src || (what == VKR_TARGET_HDR16F && (dev->flags & VKR_CONFIG_TRANSIENT_HDR)))
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_device_ptr: target has no linear buffer");
    *ptr = const_cast<void*>(src); *bytes = n;
    return VKR_SUCCESS;
}

int vkr_comm_unique_id(void* out, size_t bytes) {
    static_assert(sizeof(ncclUniqueId) == VKR_COMM_ID_BYTES, "ncclUniqueId size");
    if (!out || bytes < VKR_COMM_ID_BYTES) return VKR_ERROR_INVALID_ARGUMENT;
    ncclUniqueId id;
    if (ncclGetUniqueId(&id) != ncclSuccess) return VKR_ERROR_CUDA;
    std::memcpy(out, &id, VKR_COMM_ID_BYTES);
    return VKR_SUCCESS;
}

int vkr_shard_init(VkrDevice* dev, int rank, int nranks, const void* twoIds, size_t bytes) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!twoIds || bytes < 2 * VKR_COMM_ID_BYTES || nranks < 1 || nranks > MAX_RANKS || rank < 0 || rank >= nranks)
        return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_shard_init: bad arguments");
    if ((int)dev->H < nranks || (int)dev->S < nranks) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_shard_init: more ranks than rows");
    if (dev->commShadow || dev->commFrame) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_shard_init: already initialised");
    if (dev->frameIndex != 0) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_shard_init: must precede the first vkr_draw (the ranks count frames together)");
    CK(cudaSetDevice(dev->device));
    if (nranks == 1) return VKR_SUCCESS;
    ncclUniqueId ids[2];
    std::memcpy(ids, twoIds, sizeof(ids));
    ncclResult_t rc = ncclCommInitRank(&dev->commShadow, nranks, ids[0], rank);
    if (rc == ncclSuccess) rc = ncclCommInitRank(&dev->commFrame, nranks, ids[1], rank);
    if (rc != ncclSuccess) { dev->err = std::string("ncclCommInitRank: ") + ncclGetErrorString(rc); return VKR_ERROR_CUDA; }
    int prLo = 0, prHi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prLo, &prHi));
    CK(cudaStreamCreateWithPriority(&dev->gatherStream, cudaStreamNonBlocking, prHi));
    const size_t px = (size_t)dev->W * dev->H;
    for (int i = 0; i < VKR_FRAME_OVERLAP; ++i) {
        ALLOC(dev->gathered[i], px * 4 + (i == 0 ? 4096 : 0));      // + the flags (one allocation = one IPC handle)
        CK(cudaEventCreateWithFlags(&dev->gatheredEv[i], cudaEventDisableTiming));
    }
    dev->shardRank = rank; dev->shardRanks = nranks;

    // ---- peer memory: every rank publishes IPC handles of its two shadow maps and its two assembled images (the flags sit
    // behind the first image), the handles travel through the communicator that was just made, every rank maps its peers'.
    // VKR_EXCHANGE=nccl keeps the NCCL send/recv exchange (also the fallback when the mapping fails).
    const char* mode = std::getenv("VKR_EXCHANGE");
    if (mode && std::strcmp(mode, "nccl") == 0) return VKR_SUCCESS;
    struct Pack { cudaIpcMemHandle_t shadow[VKR_FRAME_OVERLAP], gathered[VKR_FRAME_OVERLAP]; };
    std::vector<Pack> packs((size_t)nranks);
    Pack& mine = packs[(size_t)rank];
    bool ok = true;
    for (int i = 0; i < VKR_FRAME_OVERLAP && ok; ++i)
        ok = cudaIpcGetMemHandle(&mine.shadow[i], dev->shadowBuf[i]) == cudaSuccess && cudaIpcGetMemHandle(&mine.gathered[i], dev->gathered[i]) == cudaSuccess;
    Pack* dpacks = nullptr;
    ALLOC(dpacks, sizeof(Pack) * (size_t)nranks);
    // (every rank takes part in the all-gather even if its own export failed: `ok` travels in the first byte being zeroed)
    if (!ok) std::memset(&mine, 0, sizeof(Pack));
    CK(cudaMemcpyAsync(dpacks + rank, &mine, sizeof(Pack), cudaMemcpyHostToDevice, dev->stream));
    CK(cudaMemsetAsync(reinterpret_cast<char*>(dev->gathered[0]) + px * 4, 0, 4096, dev->stream));
    if (ncclAllGather(dpacks + rank, dpacks, sizeof(Pack), ncclChar, dev->commShadow, dev->stream) != ncclSuccess)
        return dev->fail(VKR_ERROR_CUDA, "vkr_shard_init: exchanging the IPC handles");
    CK(cudaMemcpyAsync(packs.data(), dpacks, sizeof(Pack) * (size_t)nranks, cudaMemcpyDeviceToHost, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));
    dev_free(dev, dpacks);
    const Pack zero{};
    for (int p = 0; p < nranks; ++p) if (std::memcmp(&packs[(size_t)p], &zero, sizeof(Pack)) == 0) ok = false;
    dev->myFlags = reinterpret_cast<ShardFlags*>(reinterpret_cast<char*>(dev->gathered[0]) + px * 4);
    for (int p = 0; p < nranks && ok; ++p) {
        if (p == rank) {
            for (int i = 0; i < VKR_FRAME_OVERLAP; ++i) dev->peerShadow[p][i] = dev->shadowBuf[i];
            dev->peerFlags[p] = dev->myFlags;
            if (p == 0) for (int i = 0; i < VKR_FRAME_OVERLAP; ++i) dev->rootGathered[i] = dev->gathered[i];
            continue;
        }
        for (int i = 0; i < VKR_FRAME_OVERLAP && ok; ++i) {
            void* q = nullptr;
            ok = cudaIpcOpenMemHandle(&q, packs[(size_t)p].shadow[i], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            if (ok) { dev->ipcOpened.push_back(q); dev->peerShadow[p][i] = (uint32_t*)q; }
        }
        // the flags of every peer (behind its first image); the images themselves only of rank 0, which presents
        void* g0 = nullptr;
        if (ok) ok = cudaIpcOpenMemHandle(&g0, packs[(size_t)p].gathered[0], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
        if (ok) { dev->ipcOpened.push_back(g0); dev->peerFlags[p] = reinterpret_cast<ShardFlags*>(reinterpret_cast<char*>(g0) + px * 4); }
        if (ok && p == 0) {
            void* g1 = nullptr;
            ok = cudaIpcOpenMemHandle(&g1, packs[0].gathered[1], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            if (ok) { dev->ipcOpened.push_back(g1); dev->rootGathered[0] = (uint32_t*)g0; dev->rootGathered[1] = (uint32_t*)g1; }
        }
    }
    if (!ok) {
        cudaGetLastError();
        for (void* q : dev->ipcOpened) cudaIpcCloseMemHandle(q);
        dev->ipcOpened.clear();
    }
    // all or nothing: the ranks agree on the exchange they use (min over the ranks of `ok`)
    int* dflag = nullptr;
    ALLOC(dflag, sizeof(int));
    const int mineOk = ok ? 1 : 0;
    CK(cudaMemcpyAsync(dflag, &mineOk, sizeof(int), cudaMemcpyHostToDevice, dev->stream));
    if (ncclAllReduce(dflag, dflag, 1, ncclInt, ncclMin, dev->commShadow, dev->stream) != ncclSuccess)
        return dev->fail(VKR_ERROR_CUDA, "vkr_shard_init: agreeing on the exchange");
    int allOk = 0;
    CK(cudaMemcpyAsync(&allOk, dflag, sizeof(int), cudaMemcpyDeviceToHost, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));
    dev_free(dev, dflag);
    dev->p2p = allOk == 1;
    if (!dev->p2p) dev->err = "vkr_shard_init: peer memory unavailable, exchanging over NCCL";
    return VKR_SUCCESS;
}

int vkr_shard_info(VkrDevice* dev, int32_t out[4]) {
    if (!dev || !out) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    out[0] = dev->lastBand[0]; out[1] = dev->lastBand[1]; out[2] = dev->p2p ? 1 : 0; out[3] = dev->shardRanks;
    return VKR_SUCCESS;
}

int vkr_set_shadow_exchange(VkrDevice* dev, VkrShadowExchangeFn fn, void* user) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    dev->shadowExchange = fn; dev->shadowExchangeUser = user;
    return VKR_SUCCESS;
}

int vkr_cull_stats(VkrDevice* dev, uint32_t out[4]) {
    if (!dev || !out) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    out[0] = dev->lastOpaqueClusters; out[1] = dev->countersHost[3];
    out[2] = dev->lastShadowClusters; out[3] = dev->countersHost[7];
    return VKR_SUCCESS;
}

int vkr_launch_count(VkrDevice* dev) { if (!dev) return 0; (void)drain(dev); return dev->launches; }
size_t vkr_upload_bytes(VkrDevice* dev) { if (!dev) return 0; (void)drain(dev); return dev->uploadBytes; }

}  // extern "C"

// ======================================================================= allocation-free RGB8 pack (vkr_pack_rgb8_static)
// No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops, harnesses that forbid it).  It is pack_rgb8_kernel — the kernel behind
// vkr_readback(VKR_TARGET_RGB8), whose packed image is device scratch memory before it is copied out — with that scratch in a
// file-scope device array instead of a cudaMalloc'd buffer.  Every GPU that loads the library holds the array until exit.
//
// In HOST code `g_packStatic` is only a stand-in for the symbol, not a device address (a kernel handed it would fault, or, where
// the GPU can reach host memory, quietly use the wrong memory): the address the kernel gets comes from cudaGetSymbolAddress,
// asked once per device with that device current.
static __device__ __align__(256) uint32_t g_packStatic[VKR_PACK_STATIC_MAX_BYTES / 4];
static std::mutex g_packStaticMutex;
static uint32_t* g_packStaticPtr[64] = {};       // per device ordinal: the device address of g_packStatic on that device

extern "C" int vkr_pack_rgb8_static(VkrDevice* dev, const void* rgba8Device, uint32_t width, uint32_t height, void* rgb8Out, size_t rgb8Bytes,
                                    void* cudaStream, int wait) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    if (!rgba8Device || !rgb8Out || width == 0 || height == 0) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_pack_rgb8_static: null or empty argument");
    if ((reinterpret_cast<uintptr_t>(rgba8Device) & 15u) != 0) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_pack_rgb8_static: the RGBA8 image must be 16-byte aligned");
    const size_t px = (size_t)width * height, bytes = px * 3;
    if (bytes > (size_t)VKR_PACK_STATIC_MAX_BYTES) {
        char msg[192];
        std::snprintf(msg, sizeof msg, "vkr_pack_rgb8_static: %ux%u needs %zu bytes of scratch, the static array holds %u (VKR_PACK_STATIC_MAX_BYTES); use vkr_readback",
                      width, height, bytes, (unsigned)VKR_PACK_STATIC_MAX_BYTES);
        return dev->fail(VKR_ERROR_CAPACITY, msg);
    }
    if (rgb8Bytes < bytes) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_pack_rgb8_static: destination too small");
    if (dev->device < 0 || dev->device >= 64) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_pack_rgb8_static: device ordinal out of range");
    CK(cudaSetDevice(dev->device));
    uint32_t* scratch;
    {
        std::lock_guard<std::mutex> lock(g_packStaticMutex);
        if (!g_packStaticPtr[dev->device]) {         // set-up, once per device, that device current
            void* p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_packStatic);
            if (e != cudaSuccess || !p) { cudaGetLastError(); return dev->fail(VKR_ERROR_CUDA, "vkr_pack_rgb8_static: cudaGetSymbolAddress", e); }
            g_packStaticPtr[dev->device] = static_cast<uint32_t*>(p);
        }
        scratch = g_packStaticPtr[dev->device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cudaStream);
    // (nothing in the array has to start at zero: the kernel writes every one of the 3 * width * height bytes that are copied out)
    pack_rgb8_kernel<<<(unsigned)(((px >> 2) + 256) / 256), 256, 0, st>>>(static_cast<const uint32_t*>(rgba8Device), scratch, px);
    LAUNCHED("pack_rgb8_kernel (vkr_pack_rgb8_static)");
    CK(cudaMemcpyAsync(rgb8Out, scratch, bytes, cudaMemcpyDefault, st));
    if (wait) CK(cudaStreamSynchronize(st));
    return VKR_SUCCESS;
}

#include "vkr_scene_api.cuh"        // vkr_scene_create / vkr_scene_set_transforms / vkr_draw_scene / vkr_scene_draw_list
