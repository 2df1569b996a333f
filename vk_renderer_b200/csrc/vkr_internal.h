// vkr_internal.h — device-side data layout shared by the kernels and the C-ABI host code.
// Everything here is POD; matrices are column-major (m[col*4+row]) like glm/GLSL.
#pragma once
#include <stdint.h>
#include "../../include/vkr.h"

#if defined(__CUDACC__)
#include <vector_types.h>
#define VKR_HD __host__ __device__ __forceinline__
#define VKR_HDN __host__ __device__ __noinline__
#else
#define VKR_HD inline
#define VKR_HDN inline
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(8) float2 { float x, y; };
struct alignas(4) uchar4 { unsigned char x, y, z, w; };
struct alignas(8) uint2 { unsigned int x, y; };
struct alignas(16) uint4 { unsigned int x, y, z, w; };
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
#endif

namespace vkr {

// ---- visibility-buffer encoding -------------------------------------------------
// key = (depth bits << 32) | (0xFFFFFFFE - primSeq).  64-bit atomicMin keeps the nearest
// fragment; among equal depths the LARGER primSeq wins, which is LESS_OR_EQUAL with
// depth writes in submission order (src/vk_engine_init.cpp:539).  Depth is clamped to
// [0,1] so the IEEE bit pattern orders like the value.
static const uint64_t VIS_CLEAR = 0x3F800000FFFFFFFFull;   // depth 1.0, no primitive
static const uint32_t VIS_EMPTY_LO = 0xFFFFFFFFu;
static const uint32_t DEPTH_ONE_BITS = 0x3F800000u;

// Large triangles are queued as (triangle, TILE x TILE tile) items, one warp each.  Measured (profiles/r02_ab.md, "tile edge";
// frames/s of configs 1 / 2 / 3 / 4): 64 -> 14960 / 8180 / 843 / 2586, 32 -> 20200 / 8190 / 887 / 2576, 16 -> 19790 / 8190 / 866.
#ifndef VKR_TILE
#define VKR_TILE 32
#endif
static const int TILE = VKR_TILE;          // large-triangle tile edge in pixels
// In-thread raster when the bounding box is <= SMALL_MAX x SMALL_MAX samples; bigger triangles go to large_raster_kernel, where
// a whole warp works on one.  Measured (profiles/r02_ab.md, "in-lane threshold"; frames/s of configs 1 / 2 / 3 / 4):
// 16 -> 7447 / 6360 / 714 / 2492, 12 -> 7230 / 6646 / 697, 10 -> 7597 / 6874 / 685 / 2496, 8 -> 8081 / 7244 / 681 / 2489,
// 6 -> 8580 / 7160 / 555: a lane alone on a 16 x 16 box holds its 31 neighbours for 256 iterations.
#ifndef VKR_SMALL_MAX
#define VKR_SMALL_MAX 8
#endif
static const int SMALL_MAX = VKR_SMALL_MAX;
static const int CUTOUT_SMALL_MAX = 4;     // ... for alpha-tested triangles (each fragment costs a texture evaluation)
static const int CL_TRIS = 128;            // a cluster (one warp's unit of work) holds at most this many triangles ...
static const int CL_VERTS = 128;           // ... and this many distinct vertices (7-bit local ids, 2 KB of shared memory)
static const int SEQ_SLOT = 256;           // primitive sequence numbers are handed out to draws in slots of 256
#ifndef VKR_GEOM_WARPS
#define VKR_GEOM_WARPS 4
#endif
static const int GEOM_WARPS = VKR_GEOM_WARPS;   // warps (clusters) per geometry CTA
#ifndef VKR_SHADE_MIN_CTAS
#define VKR_SHADE_MIN_CTAS 12              // shading CTAs (128 threads) resident per SM the register allocation must allow
#endif
#ifndef VKR_GEOM_MIN_CTAS
#define VKR_GEOM_MIN_CTAS 12               // geometry CTAs resident per SM the register allocation must allow (48 warps, 40 registers; 15 KB of shared memory per CTA)
#endif
static const float GUARD_BAND = 64.0f;     // clip x,y only beyond +-GUARD_BAND*w

// One mesh in HBM: SoA, 3 x 16-byte streams + u32 indices.
//   pos   = (x, y, z, 1)               read by every geometry pass (128-bit loads)
//   attrA = (nx, ny, nz, uv_x)         read only by the shading pass
//   attrB = (r, g, b, uv_y)            read only by the shading pass (Vertex.color.w is never used)
struct DMesh {
    const float4* pos;
    const float4* attrA;
    const float4* attrB;
    const uint32_t* idx;
    uint32_t nv, ni;
};

// Per-draw record, rebuilt every frame from VkrRenderObject (replaces GPUDrawPushConstants,
// src/vk_types.h:72-76, plus the index-buffer binding).
struct DObject {
    float mvp[16];            // viewproj * world, glm association, computed on the host
    float world[16];
    const float4* pos;
    const float4* attrA;
    const float4* attrB;
    const uint32_t* idx;      // already offset by firstIndex
    const uint32_t* clVerts;  // cluster data of this draw's index range (see ClusterRange)
    const uint32_t* clTris;
    const uint32_t* clOrig;
    const uint2* clInfo;
    const float4* clBounds;
    uint32_t triCount;
    uint32_t primBase;        // dense submission-order id of this draw's first triangle (readback only)
    uint32_t material;
    uint32_t flags;           // bit0: transparent pass, bit1: alpha-tested (albedo texture has alpha < 0.5), bit2: a sampled texture has mips,
                              // bit7: DRAW_HIDDEN (set on the device by vkr_draw_scene's is_visible test)
    const uint4* clBlob;      // the same clusters as self-contained CL_BLOB_BYTES blocks (what the geometry kernels stage with one bulk copy)
    const float4* clCone;     // three float4 per cluster: normal cone and bounding sphere (ClusterCone, vkr_raster.cuh)
    uint64_t pad;
};
static_assert(sizeof(DObject) % 16 == 0, "DObject array is read with 16-byte loads");

// ---- update_scene on the device (vkr_scene_create / vkr_draw_scene; bodies in vkr_scene.cuh) ----------------------------
static const uint32_t DRAW_HIDDEN = 0x80u;   // DObject::flags bit7: is_visible said no — the draw takes part in no pass of this frame

// One RenderObject the traversal emits (a surface of a node's mesh), static part.  48 bytes.
struct SceneDraw {
    uint32_t node;                     // the node whose worldTransform the draw takes
    uint32_t nClusters;                // clusters of its index range (0 for a draw without triangles)
    float origin[3], sphereRadius;     // MeshAsset::bounds (src/vk_types.h:112-116): what is_visible tests
    float center[3], radius;           // bounding sphere of the range's clusters (ClusterRange): the processing-order key
    uint32_t pad[2];
};
static_assert(sizeof(SceneDraw) == 48, "SceneDraw is read as three 16-byte words");

// Per frame, passed by value.
struct SceneFrame {
    float viewproj[16];
    float lightViewproj[VKR_NUM_CASCADES][16];
    float planes[24];                  // extract_frustum_planes (src/vk_engine_run.cpp:420-433): 6 x (n.xyz, d), normalised
    uint32_t nDraws, opaqueCount, shadowLayers, pad;
};


// Cluster data, built once per (mesh, firstIndex, indexCount) at first use (build_clusters):
// the draw's triangles are sorted along a Morton curve of their centroids and cut into clusters of
// <= CL_TRIS triangles and <= CL_VERTS distinct vertices, so a cluster is a compact surface patch
// with a tight bounding box.
//   clInfo[c]              = (offset of the cluster's first entry in clVerts, vertexCount | triCount << 16)
//   clVerts[...]           = the cluster's distinct mesh vertices
//   clTris[c*CL_TRIS + t]  = l0 | l1 << 7 | l2 << 14   (local vertex ids)
//   clOrig[c*CL_TRIS + t]  = the triangle's index in the draw's index range (submission order)
//   clBounds[2c, 2c+1]     = object-space AABB (min.xyz, max.xyz) of the cluster's vertices
//
// Primitive sequence numbers stay in SUBMISSION order whatever the processing order: a draw owns
// ceil(triCount / 256) slots of 256 numbers starting at seqBase, triangle t of the draw is seqBase + t,
// so seq >> 8 indexes the frame's (draw, first triangle) table and later primitives win equal-depth ties.
// A cluster as ONE contiguous block (CL_BLOB_BYTES apart), what a geometry warp brings into its shared memory with a single
// bulk copy (cp.async.bulk, the TMA unit) before it knows whether the cluster survives the occlusion test:
//   [0, 1024)     uint2[CL_TRIS]   (l0 | l1 << 7 | l2 << 14, the triangle's number in the draw)   — clTris and clOrig interleaved
//   [1024, 3072)  float4[CL_VERTS] the cluster's distinct vertices' positions (x, y, z, 1)         — pos[clVerts[..]] gathered
// Only the first 1024 + 16 * vertexCount bytes are copied.
static const uint32_t CL_BLOB_TRI_BYTES = CL_TRIS * 8u, CL_BLOB_BYTES = CL_BLOB_TRI_BYTES + CL_VERTS * 16u;

// One geometry warp's work: cluster `c` of a draw in one pass, with the outcome of its cluster test (CullRec), written per
// frame by expand_chunks_kernel in the pass's PROCESSING order (front to back by draw) — the warp reads 32 bytes at its slot,
// no indirection.
struct WorkItem {
    const uint4* blob;        // the cluster's block
    uint32_t obj, seqBase;    // draw; first sequence number of the draw
    uint32_t info;            // vertexCount | triCount << 8 | draw flags << 16 (bit0 transparent, bit1 alpha-tested) | verdict << 24
    uint32_t rect0, rect1;    // CullRec: x0 | y0 << 16, x1 | y1 << 16
    uint32_t zminBits;
};
static_assert(sizeof(WorkItem) == 32, "WorkItem is read as two 16-byte words");

struct DChunk {               // a cluster as the cut-out kernel and the read-back paths see it, written on the device by expand_chunks_kernel
    const float4* pos;        // mesh positions
    const uint32_t* clVerts;  // this cluster's distinct-vertex list
    const uint32_t* clTris;   // this cluster's packed triangles
    const uint32_t* clOrig;   // ... and their submission-order triangle numbers
    uint32_t obj, seqBase;    // draw; first sequence number of the draw
    uint32_t counts;          // vertexCount | triCount << 16
    uint32_t flags;           // draw flags (bit0 transparent, bit1 alpha-tested)
};
static_assert(sizeof(DChunk) == 48, "DChunk is read as three 16-byte words");
// The outcome of the cluster test of one (pass, cluster), computed once per frame by expand_chunks_kernel from the cluster's
// bounding box (one THREAD per cluster there; in the geometry kernel it would be one warp per cluster): the warp reads 16 bytes
// instead of projecting eight corners.
struct CullRec {
    int16_t x0, y0, x1, y1;   // pixel rectangle the cluster can touch (valid for verdict CLUSTER_TEST)
    uint32_t zminBits;        // nearest depth it can write, as float bits
    int32_t verdict;          // CLUSTER_CULLED / CLUSTER_DRAW / CLUSTER_TEST (vkr_raster.cuh)
};
static_assert(sizeof(CullRec) == 16, "CullRec is read as one 16-byte word");
// Cluster c of draw `obj` whose first sequence slot is `seqSlot`.
VKR_HD DChunk make_chunk(const DObject& o, uint32_t obj, uint32_t c, uint32_t seqSlot) {
    const uint2 info = o.clInfo[c];
    DChunk ch;
    ch.pos = o.pos; ch.clVerts = o.clVerts + info.x; ch.clTris = o.clTris + (size_t)c * CL_TRIS;
    ch.clOrig = o.clOrig + (size_t)c * CL_TRIS;
    ch.obj = obj; ch.seqBase = seqSlot * SEQ_SLOT; ch.counts = info.y;
    ch.flags = (o.flags & DRAW_HIDDEN) ? DRAW_HIDDEN : o.flags;      // a hidden draw is nobody's work (the cut-out kernel selects on bit1)
    return ch;
}

// The work item of cluster c of draw `obj` (first sequence slot `seqSlot`) with the outcome `cr` of its cluster test.
VKR_HD WorkItem make_work_item(const DObject& o, uint32_t obj, uint32_t c, uint32_t seqSlot, const CullRec& cr) {
    const uint2 info = o.clInfo[c];
    WorkItem w;
    w.blob = o.clBlob + (size_t)c * (CL_BLOB_BYTES / 16u);
    w.obj = obj; w.seqBase = seqSlot * SEQ_SLOT;
    w.info = (info.y & 0xFFu) | ((info.y >> 16) << 8) | ((o.flags & 0xFFu) << 16) | ((uint32_t)cr.verdict << 24);
    w.rect0 = (uint32_t)(uint16_t)cr.x0 | ((uint32_t)(uint16_t)cr.y0 << 16);
    w.rect1 = (uint32_t)(uint16_t)cr.x1 | ((uint32_t)(uint16_t)cr.y1 << 16);
    w.zminBits = cr.zminBits;
    return w;
}

struct DMaterial {
    float colorFactors[4];
    float metalRough[4];
    int32_t albedoID, normalID, metalRoughID, pass;
    // what texture() returns for the albedo / metal-rough slot when that does not depend on uv — a 1x1 image (the engine's default
    // textures, i.e. every untextured material) or an unbound slot; filled by material_constants() whenever materials or bindless
    // slots change, so the shading pass does not walk a descriptor to fetch the same texel for every pixel
    float albedoTexel[4], mrTexel[4];
    uint32_t constFlags;      // bit0: albedoTexel valid, bit1: mrTexel valid
    uint32_t pad[3];
};
static_assert(sizeof(DMaterial) == 96, "DMaterial");

struct DTexture {             // one bindless slot (combined image sampler)
    const uchar4* data;       // mip chain, level l at data + mipOffset[l]
    uint32_t w, h, levels;
    uint32_t flags;           // bit0: sRGB, bit1: LINEAR sampler (else NEAREST), bit2: has alpha < 0.5 somewhere,
                              // bit3: texels are 4 x f16 (8 bytes; cube maps only), `data` then points at uint2 texels,
                              // bit4: texels are 4 x f32 (8-bit cube-map texels decoded at upload), `data` points at float4 texels
    uint32_t mipOffset[16];
    uint32_t texel0, pad;     // the first texel of level 0 as uploaded (host-side use: material_constants)
};

// A projected, snapped, oriented triangle too large for the in-thread path (48 B).
struct LargeTri {
    int32_t X[3], Y[3];       // 24.8 fixed point, doubled area > 0
    float z0, dzdx, dzdy;
    int16_t bx0, by0, bx1, by1;
    uint32_t seq;
};
struct LargeItem { uint32_t tri; uint16_t tx, ty; };   // (triangle, TILE x TILE tile); tx == ITEM_WHOLE: the triangle's whole bounding box
static const uint16_t ITEM_WHOLE = 0xFFFFu;
static const uint32_t ITEM_SKIP = 0xFFFFFFFFu;    // LargeItem::tri of a slot claimed by a triangle that did not fit the queue

struct Queues {
    LargeTri* tris;
    LargeItem* items;
    uint32_t* counters;       // [0] tri count, [1] item count, [2] overflow flag, [3] clusters culled
    uint32_t triCap, itemCap;
};

// Raster target of one geometry pass.
struct RasterTarget {
    uint64_t* vis;            // main pass: W*H keys
    uint32_t* depthBits;      // shadow pass: S*S float bits
    int32_t width, height;
    int32_t sx0, sy0, sx1, sy1;   // scissor [sx0,sx1) x [sy0,sy1)
    float halfW, halfH;
    float halfW256, halfH256; // 128*W, 128*H: viewport scale fused with the 8-bit sub-pixel snap
    int32_t cullFront;        // 1: cull front faces (shadow), 0: cull back faces (main)
};

// Per-frame constants for the shading pass.
struct ShadeConst {
    VkrGPUSceneData scene;
    float camPos[3];
    float lightVec[3];                 // normalize(-sunlightDirection.xyz)
    float shadowMat[VKR_NUM_CASCADES][16];   // biasMat * lightViewproj[i]
    float bgTop[4], bgBottom[4];
    int32_t width, height, shadowExtent;
    int32_t sx0, sy0, sx1, sy1;
    int32_t enableBackground, enablePostprocess, enableSkybox, fragShader;
    int32_t shadowMode, enableShadows;
    int32_t clearVis;                  // VKR_CONFIG_TRANSIENT_DEPTH: every shaded pixel leaves VIS_CLEAR behind
    int32_t storeHdr;                  // 0 under VKR_CONFIG_TRANSIENT_HDR: the RGBA16F draw image is not written (both roundings still applied)
    uint32_t nobj;
    float halfW, halfH;
    float invHalfW, invHalfH;
    // skybox ray through window position (x,y): dir = skyO + x*skyX + y*skyY
    // (= transpose(mat3(view)) * unproject, y negated: shaders/skybox.vert:10-15)
    float skyO[3], skyX[3], skyY[3];
};

struct ShadeBuffers {
    uint64_t* vis;                     // read; written only to clear it (ShadeConst::clearVis)
    const DObject* objects;
    const uint2* chunkObj;             // sequence slots: seq >> 8 -> (draw, first triangle of the slot)
    const DMaterial* materials;
    const DTexture* textures;          // VKR_MAX_BINDLESS slots
    const DTexture* cube;              // 6 faces (nullptr when no cubemap)
    const float* shadow;               // S*S*NUM_CASCADES
    float* accum;                      // RGBA32F sums of the transparent pass (nullptr: the frame has none)
    uint16_t* hdr;                     // RGBA16F
    uint32_t* rgba8;
    uint32_t* rgba8Peer;               // sharded over peer memory: the presenting rank's image (mapped here), stored to as well; else nullptr
    const uint32_t* peerFree; uint32_t peerFreeValue;   // ... and the flag (>= value) that says it may be stored to
    const float* srgbLut;              // 256 floats
    // the frame's queue counters (main, shadow, transparent pass; 4 words each, nullptr: pass not run) and their pinned host
    // mirror, which the shading kernel fills through its mapped address: a 16-byte cudaMemcpyAsync on a render stream would
    // queue behind the previous frame's 25 MB read-back on the device-to-host copy engine and stall the stream for it
    const uint32_t* counters[3];
    uint32_t* countersHost;
};

}  // namespace vkr
