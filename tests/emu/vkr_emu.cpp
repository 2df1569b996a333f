// vkr_emu.cpp — TEST-ONLY host emulation of the CUDA kernels' per-thread bodies.
//
// There is no GPU in the development container, so this harness compiles the very same
// __host__ __device__ functions the kernels call (vkr_raster.cuh, vkr_shade.cuh,
// vkr_frameprep.h) with g++ and runs "one thread per triangle / per pixel" as plain loops.
// It lets the CPU test-suite check the kernels' logic against the oracle before a GPU run.
// It is NOT part of the product: libvkr.so does not contain it and nothing in
// vk_renderer_b200/ references it.  Warp-cooperative code paths (lane-strided loops) run
// with nlanes = 32 lanes executed sequentially.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fPIC -shared (tests/conftest.py does it).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../vk_renderer_b200/csrc/vkr_internal.h"
#include "../../vk_renderer_b200/csrc/vkr_raster.cuh"
#include "../../vk_renderer_b200/csrc/vkr_shade.cuh"
#include "../../vk_renderer_b200/csrc/vkr_frameprep.h"
#include "../../vk_renderer_b200/csrc/vkr_scene.cuh"

using namespace vkr;

struct EmuImage {
    std::vector<uchar4> data; uint32_t w, h, levels; uint32_t mipOffset[16]; int format; bool cutout;
};
struct EmuMesh { std::vector<float4> pos, a, b; std::vector<uint32_t> idx; };

struct EmuDevice;
struct EmuClusters : ClusterProvider {
    EmuDevice* d;
    std::vector<ClusterData*> store;
    std::vector<std::pair<std::vector<uint32_t>, ClusterRange>> cache;   // key (mesh, first, count)
    bool get(uint32_t mesh, uint32_t firstIndex, uint32_t indexCount, ClusterRange& out) override;
    ~EmuClusters() { for (auto* c : store) delete c; }
};

// vkr_scene_create's record (the device library's SceneRec, host arrays instead of device ones)
struct EmuScene {
    std::vector<float> local, world;
    std::vector<int32_t> parent;
    std::vector<VkrSceneSurface> surfaces;      // draw order: emission order, MainColor first
    std::vector<SceneDraw> draws;
    std::vector<DObject> tmpl;
    std::vector<uint32_t> clusterBase, seqSlotBase;
    uint32_t opaqueCount = 0, opaqueClusters = 0, totalTris = 0;
};

struct EmuDevice {
    uint32_t W, H, S;
    std::vector<EmuScene*> scenes;
    EmuScene* lastScene = nullptr;
    EmuClusters clusters;
    std::vector<EmuMesh*> meshStore;
    std::vector<DMesh> meshes;
    std::vector<EmuImage*> images;
    std::vector<DTexture> tex;
    DTexture cube[6]; bool haveCube = false;
    std::vector<DMaterial> materials;
    std::vector<uint64_t> vis;
    std::vector<uint32_t> shadowBits;
    std::vector<uint16_t> hdr;
    std::vector<uint32_t> rgba8;
    std::vector<float> accum;
    std::vector<LargeTri> ltris; std::vector<LargeItem> litems; uint32_t counters[4];
    float lut[256];
    FramePlan plan;
    std::vector<DChunk> chunks;
    std::vector<uint2> chunkObj;
    std::vector<uint32_t> perm;          // passes x clusters
    uint32_t culled[2] = {0, 0};         // clusters dropped by the cluster test: [0] main, [1] shadow
    VkrEngineStats stats{};
    VkrShadowExchangeFn shadowExchange = nullptr; void* shadowExchangeUser = nullptr;
    int shardRank = 0, shardRanks = 1;      // vkr_shard_init's band split, without the exchange (the test supplies it)
    std::string err;
};

bool EmuClusters::get(uint32_t mesh, uint32_t firstIndex, uint32_t indexCount, ClusterRange& out) {
    for (auto& e : cache) if (e.first[0] == mesh && e.first[1] == firstIndex && e.first[2] == indexCount) { out = e.second; return true; }
    ClusterData* cd = new ClusterData();
    build_clusters(d->meshes[mesh].idx + firstIndex, indexCount / 3, d->meshes[mesh].nv, d->meshes[mesh].pos, *cd);
    store.push_back(cd);
    out = ClusterRange{cd->verts.data(), cd->tris.data(), cd->orig.data(), cd->info.data(), cd->bounds.data(), cd->blob.data(), cd->cone.data(), 0u, {0, 0, 0}, 0};
    range_sphere(*cd, out);
    cache.push_back({{mesh, firstIndex, indexCount}, out});
    return true;
}

static uint32_t srgb8(float l) {
    if (!(l > 0.0f)) return 0u;
    if (l >= 1.0f) return 255u;
    float s = l <= 0.0031308f ? l * 12.92f : 1.055f * powf(l, 1.0f / 2.4f) - 0.055f;
    return unorm8(s);
}

static EmuImage* build_image(EmuDevice* d, const void* rgba8, uint32_t w, uint32_t h, int fmt, int mipmapped) {
    EmuImage* im = new EmuImage();
    im->w = w; im->h = h; im->format = fmt;
    im->levels = mipmapped ? (uint32_t)std::floor(std::log2((double)std::max(w, h))) + 1 : 1;
    size_t total = 0;
    for (uint32_t l = 0; l < im->levels; ++l) { im->mipOffset[l] = (uint32_t)total; total += (size_t)std::max(1u, w >> l) * std::max(1u, h >> l); }
    im->data.resize(total);
    std::memcpy(im->data.data(), rgba8, (size_t)w * h * 4);
    im->cutout = false;
    for (size_t i = 0; i < (size_t)w * h; ++i) if (im->data[i].w < 128) { im->cutout = true; break; }
    for (uint32_t l = 0; l + 1 < im->levels; ++l) {   // same arithmetic as mip_blit_kernel
        int sw = std::max(1u, w >> l), sh = std::max(1u, h >> l), dw = std::max(1u, w >> (l + 1)), dh = std::max(1u, h >> (l + 1));
        const uchar4* src = im->data.data() + im->mipOffset[l];
        uchar4* dst = im->data.data() + im->mipOffset[l + 1];
        float scx = (float)sw / (float)dw, scy = (float)sh / (float)dh;
        for (int y = 0; y < dh; ++y) for (int x = 0; x < dw; ++x) {
            float u = ((float)x + 0.5f) * scx - 0.5f, v = ((float)y + 0.5f) * scy - 0.5f;
            float fu = floorf(u), fv = floorf(v), a = u - fu, b = v - fv;
            int x0 = std::min(std::max((int)fu, 0), sw - 1), x1 = std::min(std::max((int)fu + 1, 0), sw - 1);
            int y0 = std::min(std::max((int)fv, 0), sh - 1), y1 = std::min(std::max((int)fv + 1, 0), sh - 1);
            uchar4 q[4] = {src[(size_t)y0 * sw + x0], src[(size_t)y0 * sw + x1], src[(size_t)y1 * sw + x0], src[(size_t)y1 * sw + x1]};
            float wgt[4] = {(1.0f - a) * (1.0f - b), a * (1.0f - b), (1.0f - a) * b, a * b};
            float c[4] = {0, 0, 0, 0};
            for (int k = 0; k < 4; ++k) {
                float r, g, bl;
                if (fmt == 1) { r = d->lut[q[k].x]; g = d->lut[q[k].y]; bl = d->lut[q[k].z]; }
                else { r = q[k].x / 255.0f; g = q[k].y / 255.0f; bl = q[k].z / 255.0f; }
                float al = q[k].w / 255.0f;
                c[0] = c[0] + r * wgt[k]; c[1] = c[1] + g * wgt[k]; c[2] = c[2] + bl * wgt[k]; c[3] = c[3] + al * wgt[k];
            }
            uchar4 o;
            if (fmt == 1) { o.x = srgb8(c[0]); o.y = srgb8(c[1]); o.z = srgb8(c[2]); }
            else { o.x = unorm8(c[0]); o.y = unorm8(c[1]); o.z = unorm8(c[2]); }
            o.w = unorm8(c[3]);
            dst[(size_t)y * dw + x] = o;
        }
    }
    return im;
}
static DTexture make_tex(const EmuImage* im, int sampler) {
    DTexture t{};
    t.data = im->data.data(); t.w = im->w; t.h = im->h; t.levels = im->levels;
    t.flags = (im->format == 1 ? 1u : 0u) | (sampler == 1 ? 2u : 0u) | (im->cutout ? 4u : 0u);
    std::memcpy(t.mipOffset, im->mipOffset, sizeof(t.mipOffset));
    std::memcpy(&t.texel0, im->data.data(), 4);
    return t;
}

extern "C" {

int emu_create_image(EmuDevice* d, const void* rgba8, uint32_t w, uint32_t h, int fmt, int mipmapped, uint32_t* out);
int emu_update_bindless_texture(EmuDevice* d, int index, uint32_t image, int sampler);

int emu_create(const VkrConfig* cfg, EmuDevice** out) {
    EmuDevice* d = new EmuDevice();
    d->clusters.d = d;
    d->W = cfg->width; d->H = cfg->height; d->S = cfg->shadowExtent ? cfg->shadowExtent : 2048;
    size_t px = (size_t)d->W * d->H;
    d->vis.assign(px, VIS_CLEAR);
    d->hdr.assign(px * 4, 0); d->rgba8.assign(px, 0);
    d->shadowBits.assign((size_t)d->S * d->S * VKR_NUM_CASCADES, DEPTH_ONE_BITS);
    d->tex.assign(VKR_MAX_BINDLESS, DTexture{});
    d->ltris.resize(1 << 16); d->litems.resize(1 << 20);
    for (int i = 0; i < 256; ++i) { double c = i / 255.0; d->lut[i] = (float)(c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4)); }
    uint32_t nd = 0xFFFF8080u, img;
    emu_create_image(d, &nd, 1, 1, 0, 0, &img);
    emu_update_bindless_texture(d, 0, img, 0);
    *out = d;
    return 0;
}
void emu_destroy(EmuDevice* d) {
    if (!d) return;
    for (auto* m : d->meshStore) delete m;
    for (auto* i : d->images) delete i;
    for (auto* sc : d->scenes) delete sc;
    delete d;
}
int emu_upload_mesh(EmuDevice* d, const uint32_t* idx, size_t ni, const VkrVertex* v, size_t nv, uint32_t* out) {
    EmuMesh* m = new EmuMesh();
    m->idx.assign(idx, idx + ni);
    m->pos.resize(nv); m->a.resize(nv); m->b.resize(nv);
    for (size_t i = 0; i < nv; ++i) {   // repack_vertices_kernel
        m->pos[i] = make_float4(v[i].position[0], v[i].position[1], v[i].position[2], 1.0f);
        m->a[i] = make_float4(v[i].normal[0], v[i].normal[1], v[i].normal[2], v[i].uv_x);
        m->b[i] = make_float4(v[i].color[0], v[i].color[1], v[i].color[2], v[i].uv_y);
    }
    d->meshStore.push_back(m);
    DMesh dm{m->pos.data(), m->a.data(), m->b.data(), m->idx.data(), (uint32_t)nv, (uint32_t)ni};
    d->meshes.push_back(dm);
    *out = (uint32_t)d->meshes.size() - 1;
    return 0;
}
int emu_create_image(EmuDevice* d, const void* rgba8, uint32_t w, uint32_t h, int fmt, int mipmapped, uint32_t* out) {
    d->images.push_back(build_image(d, rgba8, w, h, fmt, mipmapped));
    *out = (uint32_t)d->images.size() - 1;
    return 0;
}
int emu_update_bindless_texture(EmuDevice* d, int index, uint32_t image, int sampler) {
    d->tex[index] = make_tex(d->images[image], sampler);
    return 0;
}
int emu_set_cubemap(EmuDevice* d, const void* const faces[6], uint32_t w, uint32_t h, int fmt, int mipmapped) {
    for (int f = 0; f < 6; ++f) {
        EmuImage* im = build_image(d, faces[f], w, h, fmt, mipmapped);
        d->images.push_back(im);
        d->cube[f] = make_tex(im, 1);
    }
    d->haveCube = true;
    return 0;
}
int emu_set_cubemap_levels(EmuDevice* d, const void* const* levels, uint32_t w, uint32_t h, uint32_t levelCount, int fmt) {
    const size_t bpt = fmt == 2 ? 8 : 4;          // vkr_set_cubemap_levels: the file's own levels, 4- or 8-byte texels
    for (int f = 0; f < 6; ++f) {
        EmuImage* im = new EmuImage();
        im->w = w; im->h = h; im->levels = levelCount; im->format = fmt; im->cutout = false;
        size_t total = 0;
        for (uint32_t l = 0; l < levelCount; ++l) { im->mipOffset[l] = (uint32_t)total; total += (size_t)std::max(1u, w >> l) * std::max(1u, h >> l); }
        im->data.resize(total * bpt / 4);
        for (uint32_t l = 0; l < levelCount; ++l)
            std::memcpy((char*)im->data.data() + (size_t)im->mipOffset[l] * bpt, levels[(size_t)l * 6 + f],
                        (size_t)std::max(1u, w >> l) * std::max(1u, h >> l) * bpt);
        d->images.push_back(im);
        d->cube[f] = make_tex(im, 1);
        if (fmt == 2) d->cube[f].flags |= 8u;
    }
    d->haveCube = true;
    return 0;
}
int emu_build_material(EmuDevice* d, const VkrMaterialConstants* p, int pass, uint32_t* out) {
    DMaterial m;
    std::memcpy(m.colorFactors, p->colorFactors, 16);
    std::memcpy(m.metalRough, p->metal_rough_factors, 16);
    m.albedoID = p->albedoID; m.normalID = p->normalID; m.metalRoughID = p->metalRoughID; m.pass = pass;
    m.constFlags = 0u;
    d->materials.push_back(m);
    *out = (uint32_t)d->materials.size() - 1;
    return 0;
}

}  // extern "C"

// geometry_kernel phase 0 (cluster culling): the box test is expand_chunks_kernel's make_cull_rec, the occlusion test walks the
// rectangle
template <bool SHADOW>
static bool cluster_survives(const CullRec& c, const RasterTarget& rt) {
    if (c.verdict == CLUSTER_CULLED) return false;
    if (c.verdict == CLUSTER_DRAW) return true;
    for (int32_t py = c.y0; py <= c.y1; ++py)
        for (int32_t px = c.x0; px <= c.x1; ++px) {
            const uint32_t stored = SHADOW ? rt.depthBits[(size_t)py * rt.width + px] : (uint32_t)(rt.vis[(size_t)py * rt.width + px] >> 32);
            if (!depth_hides(stored, c.zminBits)) return true;
        }
    return false;
}

// PASS: 0 main, 1 shadow, 2 transparent (clusters [first, first + nChunks) in natural order, BlendSink)
template <int PASS>
static void run_pass(EmuDevice* d, const FramePlan& plan, RasterTarget rt, const float* mvps, uint32_t stride, unsigned nChunks,
                     const uint32_t* perm, unsigned first = 0, const ShadeBuffers* sbT = nullptr, unsigned cutoutChunks = 0) {
    const bool SHADOW = PASS == 1;
    Queues q{d->ltris.data(), d->litems.data(), d->counters, (uint32_t)d->ltris.size(), (uint32_t)d->litems.size()};
    d->counters[0] = d->counters[1] = d->counters[2] = 0;
    std::vector<CachedVtx> svtx(CL_VERTS);
    const AlphaCtx ax{plan.objects.data(), d->chunkObj.data(), d->materials.data(), d->tex.data()};
    if (PASS == 0)
        for (unsigned c = 0; c < cutoutChunks; ++c) {        // geometry_cutout_kernel: natural order, no culling
            const DChunk ch = d->chunks[c];
            if (!(ch.flags & 2u)) continue;
            const DObject& ob = plan.objects[ch.obj];
            for (uint32_t t = 0; t < (ch.counts >> 16); ++t) {
                const uint32_t packed = ch.clTris[t];
                const uint32_t seq = ch.seqBase + ch.clOrig[t];
                VisSink s{rt.vis, rt.width, 0xFFFFFFFEu - seq};
                cutout_triangle(ob, ob.mvp, ch.clVerts[packed & 127u], ch.clVerts[(packed >> 7) & 127u], ch.clVerts[(packed >> 14) & 127u], seq, rt, q, ax, s);
            }
        }
    for (unsigned slot = 0; slot < nChunks; ++slot) {        // geometry_kernel: one CTA per cluster, in processing order
        const uint32_t ci = perm ? perm[slot] : first + slot;
        const DChunk ch = d->chunks[ci];
        if (ch.flags & DRAW_HIDDEN) continue;                  // expand_chunks_kernel: a hidden draw's clusters are culled in every list
        if (PASS == 0 && (ch.flags & 2u)) continue;
        const float* mvp = mvps + (size_t)ch.obj * stride;
        const DObject& ob = plan.objects[ch.obj];
        const uint32_t lc = ci - plan.clusterBase[ch.obj];
        const ClusterCone cone{ob.clCone[3 * lc], ob.clCone[3 * lc + 1], ob.clCone[3 * lc + 2]};
        float eye[8];
        eye_point(mvp, eye);
        const CullRec cr = make_cull_rec(mvp, ob.clBounds[2 * lc], ob.clBounds[2 * lc + 1], rt, &cone, eye);
        if (!cluster_survives<PASS == 1>(cr, rt)) { d->culled[SHADOW ? 1 : 0]++; continue; }
        const uint32_t nVerts = ch.counts & 0xFFFFu, nTris = ch.counts >> 16;
        const uint32_t* clVerts = ch.clVerts;
        for (uint32_t v = 0; v < nVerts; ++v) svtx[v] = cluster_vertex(mvp, ch.pos[clVerts[v]], rt);     // phase 1
        for (uint32_t t = 0; t < nTris; ++t) {                                                           // phase 2
            const uint32_t packed = ch.clTris[t];
            const uint32_t l0 = packed & 127u, l1 = (packed >> 7) & 127u, l2 = (packed >> 14) & 127u;
            const uint32_t seq = ch.seqBase + ch.clOrig[t];
            if (PASS == 1) { DepthSink s{rt.depthBits, rt.width}; cluster_triangle(svtx[l0], svtx[l1], svtx[l2], mvp, ch.pos, clVerts[l0], clVerts[l1], clVerts[l2], 0u, rt, q, s); }
            else if (PASS == 2) { BlendSink s{&plan.sc, sbT, seq}; cluster_triangle(svtx[l0], svtx[l1], svtx[l2], mvp, ch.pos, clVerts[l0], clVerts[l1], clVerts[l2], seq, rt, q, s); }
            else { VisSink s{rt.vis, rt.width, 0xFFFFFFFEu - seq}; cluster_triangle(svtx[l0], svtx[l1], svtx[l2], mvp, ch.pos, clVerts[l0], clVerts[l1], clVerts[l2], seq, rt, q, s); }
        }
    }
    for (uint32_t i = 0; i < d->counters[1]; ++i) {          // large_raster_kernel
        const LargeItem it = q.items[i];
        const LargeTri lt = q.tris[it.tri];
        for (int lane = 0; lane < 32; ++lane) {
            if (PASS == 1) { DepthSink s{rt.depthBits, rt.width}; raster_item(lt, it, s, lane, 32); }
            else if (PASS == 2) { BlendSink s{&plan.sc, sbT, lt.seq}; raster_item(lt, it, s, lane, 32); }
            else {
                VisSink s{rt.vis, rt.width, 0xFFFFFFFEu - lt.seq};
                const uint2 co = d->chunkObj[lt.seq >> 8];
                if (plan.objects[co.x].flags & 2u) {
                    const DObject& ob = plan.objects[co.x];
                    const uint32_t tri = co.y + (lt.seq & 255u);
                    AlphaTri at;
                    make_alpha_tri(ob, ob.mvp, ob.idx[3 * tri], ob.idx[3 * tri + 1], ob.idx[3 * tri + 2], ax, rt, at);
                    CutoutSink<VisSink> cs{s, &at};
                    raster_item(lt, it, cs, lane, 32);
                } else raster_item(lt, it, s, lane, 32);
            }
        }
    }
}

extern "C" {

static int emu_execute(EmuDevice* d, size_t n);

int emu_draw(EmuDevice* d, const VkrGPUSceneData* scene, const VkrRenderObject* objs, size_t n, const VkrFrameParams* fp) {
    FramePlan& plan = d->plan;
    for (DMaterial& m : d->materials) material_constants(m, d->tex.data(), d->lut);
    std::string why = plan_frame(scene, objs, n, fp, d->meshes, d->materials, d->tex, d->clusters, d->W, d->H, d->S, plan, d->shardRank, d->shardRanks);
    if (!why.empty()) { d->err = why; return VKR_ERROR_INVALID_ARGUMENT; }
    d->lastScene = nullptr;
    const int rc = emu_execute(d, n);
    d->stats.drawcall_count = (int)n; d->stats.triangle_count = (int)plan.totalTris;
    return rc;
}

// Everything behind the plan: expand_chunks_kernel, the raster passes, the shading pass.
static int emu_execute(EmuDevice* d, size_t n) {
    FramePlan& plan = d->plan;
    const unsigned nChunks = plan.clusterBase[n];
    d->chunks.resize(nChunks);
    d->chunkObj.resize(plan.seqSlotBase[n]);
    const uint32_t passes = 1u + (uint32_t)plan.shadowLayers;
    d->perm.assign((size_t)passes * nChunks, 0u);
    for (size_t o = 0; o < n; ++o)                            // expand_chunks_kernel
        for (uint32_t c = plan.clusterBase[o]; c < plan.clusterBase[o + 1]; ++c) {
            const uint32_t lc = c - plan.clusterBase[o];
            d->chunks[c] = make_chunk(plan.objects[o], (uint32_t)o, lc, plan.seqSlotBase[o]);
            for (uint32_t p = 0; p < passes; ++p) {
                const uint32_t sb = plan.slotBase[(size_t)p * n + o];
                if (sb != 0xFFFFFFFFu) d->perm[(size_t)p * nChunks + sb + lc] = c;
            }
        }
    for (size_t o = 0; o < n; ++o)
        for (uint32_t k = plan.seqSlotBase[o]; k < plan.seqSlotBase[o + 1]; ++k) d->chunkObj[k] = uint2{(uint32_t)o, (k - plan.seqSlotBase[o]) * SEQ_SLOT};
    d->culled[0] = d->culled[1] = 0;
    for (int layer = 0; layer < plan.shadowLayers; ++layer) {
        RasterTarget rt = plan.shadowTarget;
        rt.depthBits = d->shadowBits.data() + (size_t)layer * d->S * d->S;
        std::fill(rt.depthBits, rt.depthBits + (size_t)d->S * d->S, DEPTH_ONE_BITS);
        run_pass<1>(d, plan, rt, plan.shadowMvp.data() + (size_t)layer * n * 16, 16, plan.passSlots[1 + layer], d->perm.data() + (size_t)(1 + layer) * nChunks);
    }
    if (plan.shadowLayers && d->shadowExchange &&
        d->shadowExchange(d->shadowExchangeUser, d->shadowBits.data(), (size_t)d->S * d->S * 4, plan.shadowLayers, nullptr) != 0) {
        d->err = "shadow exchange hook failed"; return VKR_ERROR_CUDA;
    }
    RasterTarget rt = plan.mainTarget;
    rt.vis = d->vis.data();
    std::fill(d->vis.begin(), d->vis.end(), VIS_CLEAR);
    const unsigned opaqueChunks = plan.opaqueClusters;
    run_pass<0>(d, plan, rt, (const float*)plan.objects.data(), sizeof(DObject) / 4, plan.passSlots[0], d->perm.data(), 0, nullptr, opaqueChunks);
    if (d->counters[2]) { d->err = "queue overflow"; return VKR_ERROR_CAPACITY; }
    ShadeBuffers sb{};
    sb.accum = nullptr;
    sb.vis = d->vis.data(); sb.objects = plan.objects.data(); sb.chunkObj = d->chunkObj.data();
    sb.materials = d->materials.data(); sb.textures = d->tex.data(); sb.cube = d->haveCube ? d->cube : nullptr;
    sb.shadow = reinterpret_cast<const float*>(d->shadowBits.data());
    sb.hdr = d->hdr.data(); sb.rgba8 = d->rgba8.data(); sb.srgbLut = d->lut;
    if (nChunks > opaqueChunks) {                             // geometry_transparent_kernel + large_transparent_kernel
        d->accum.assign((size_t)d->W * d->H * 4, 0.0f);
        sb.accum = d->accum.data();
        run_pass<2>(d, plan, rt, (const float*)plan.objects.data(), sizeof(DObject) / 4, nChunks - opaqueChunks, nullptr, opaqueChunks, &sb);
        if (d->counters[2]) { d->err = "queue overflow"; return VKR_ERROR_CAPACITY; }
    }
    for (int py = plan.sc.sy0; py < plan.sc.sy1; ++py)
        for (int px = plan.sc.sx0; px < plan.sc.sx1; ++px) shade_pixel(plan.sc, sb, px, py);
    return 0;
}

// ---- update_scene on the device (vkr_scene_create / vkr_draw_scene): the same per-thread bodies (vkr_scene.cuh) as loops
int emu_scene_create(EmuDevice* d, const VkrSceneNode* nodes, size_t nn, const VkrSceneSurface* surfaces, size_t ns, uint32_t* out) {
    if (!nodes || !nn || !out) return VKR_ERROR_INVALID_ARGUMENT;
    EmuScene* s = new EmuScene();
    s->local.resize(nn * 16); s->world.assign(nn * 16, 0.0f); s->parent.resize(nn);
    for (size_t i = 0; i < nn; ++i) {
        if (nodes[i].parent < -1 || nodes[i].parent >= (int32_t)i) { delete s; return VKR_ERROR_INVALID_ARGUMENT; }
        std::memcpy(&s->local[i * 16], nodes[i].localTransform, 64); s->parent[i] = nodes[i].parent;
    }
    for (size_t i = 0; i < ns; ++i)
        if (surfaces[i].node >= nn || surfaces[i].mesh >= d->meshes.size() || surfaces[i].material >= d->materials.size() ||
            (uint64_t)surfaces[i].firstIndex + surfaces[i].indexCount > d->meshes[surfaces[i].mesh].ni) { delete s; return VKR_ERROR_INVALID_ARGUMENT; }
    for (int pass = 0; pass < 2; ++pass)
        for (size_t i = 0; i < ns; ++i)
            if ((d->materials[surfaces[i].material].pass == VKR_PASS_TRANSPARENT) == (pass == 1)) s->surfaces.push_back(surfaces[i]);
    while (s->opaqueCount < ns && d->materials[s->surfaces[s->opaqueCount].material].pass != VKR_PASS_TRANSPARENT) s->opaqueCount++;
    s->draws.resize(ns); s->tmpl.resize(ns); s->clusterBase.resize(ns + 1); s->seqSlotBase.resize(ns + 1);
    uint32_t ncl = 0, nslot = 0, prim = 0;
    for (size_t i = 0; i < ns; ++i) {
        const VkrSceneSurface& sf = s->surfaces[i];
        const DMesh& m = d->meshes[sf.mesh];
        DObject& o = s->tmpl[i];
        std::memset(&o, 0, sizeof(o));
        o.pos = m.pos; o.attrA = m.attrA; o.attrB = m.attrB; o.idx = m.idx + sf.firstIndex; o.triCount = sf.indexCount / 3;
        ClusterRange cr{};
        if (o.triCount) d->clusters.get(sf.mesh, sf.firstIndex, sf.indexCount, cr);
        o.clVerts = cr.verts; o.clTris = cr.tris; o.clOrig = cr.orig; o.clInfo = cr.info; o.clBounds = cr.bounds; o.clBlob = cr.blob; o.clCone = cr.cone;
        o.material = sf.material;
        SceneDraw& sd = s->draws[i];
        std::memset(&sd, 0, sizeof(sd));
        sd.node = sf.node; sd.nClusters = o.triCount ? cr.nClusters : 0u;
        std::memcpy(sd.origin, sf.boundsOrigin, 12); sd.sphereRadius = sf.boundsSphereRadius;
        std::memcpy(sd.center, cr.center, 12); sd.radius = cr.radius;
        s->clusterBase[i] = ncl; s->seqSlotBase[i] = nslot;
        ncl += sd.nClusters; nslot += (o.triCount + SEQ_SLOT - 1) / SEQ_SLOT; prim += o.triCount;
        if (i + 1 == s->opaqueCount) s->opaqueClusters = ncl;
    }
    s->clusterBase[ns] = ncl; s->seqSlotBase[ns] = nslot; s->totalTris = prim;
    d->scenes.push_back(s);
    *out = (uint32_t)d->scenes.size() - 1;
    return 0;
}
// Test probes of the kernels' texture unit (vkr_shade.cuh texture2d / texture_cube, the mip chain of create_image), with the
// signatures of the oracle's ork_image_level / ork_cb_tex2d / ork_cb_cube: tests/test_rules_on_kernels.py asks them the
// known answers of the Vulkan texture rules.
struct EmuTexCtx { EmuDevice* dev; float duvdx[2], duvdy[2]; };
int emu_image_level(EmuDevice* d, uint32_t image, int level, void* dst, size_t bytes, uint32_t* w, uint32_t* h) {
    if (!d || image >= d->images.size()) return VKR_ERROR_INVALID_ARGUMENT;
    const EmuImage& im = *d->images[image];
    if (level < 0) return (int)im.levels;
    if (level >= (int)im.levels) return VKR_ERROR_INVALID_ARGUMENT;
    const uint32_t lw = std::max(1u, im.w >> level), lh = std::max(1u, im.h >> level);
    if (w) *w = lw;
    if (h) *h = lh;
    if (dst) {
        if (bytes < (size_t)lw * lh * 4) return VKR_ERROR_INVALID_ARGUMENT;
        std::memcpy(dst, im.data.data() + im.mipOffset[level], (size_t)lw * lh * 4);
    }
    return 0;
}
void emu_cb_tex2d(void* ctx, int id, float u, float v, float out[4]) {
    const EmuTexCtx* c = (const EmuTexCtx*)ctx;
    ShadeBuffers sb{};
    sb.textures = c->dev->tex.data(); sb.srgbLut = c->dev->lut;
    const float4 t = texture2d(sb, id, u, v, c->duvdx[0], c->duvdx[1], c->duvdy[0], c->duvdy[1]);
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
}
void emu_cb_cube(void* ctx, const float dir[3], float out[4]) {
    const EmuTexCtx* c = (const EmuTexCtx*)ctx;
    ShadeBuffers sb{};
    sb.textures = c->dev->tex.data(); sb.srgbLut = c->dev->lut; sb.cube = c->dev->cube;
    const f3 r = {dir[0], dir[1], dir[2]};
    const float4 t = texture_cube(sb, r, r, r);
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
}

// Test hooks for the bucket order's helpers (vkr_scene.cuh): the order-preserving float -> uint32 map and the slice of a key
uint32_t emu_scene_ordered(float f) { return scene_ordered(f); }
float emu_scene_unordered(uint32_t u) { return scene_unordered(u); }
void emu_scene_buckets(const float* keys, uint32_t n, uint32_t* out) {
    const float inf = bits2f(0x7F800000u);
    uint32_t rmax = 0, rneg = 0;
    for (uint32_t i = 0; i < n; ++i) if (keys[i] != inf) { rmax = std::max(rmax, scene_ordered(keys[i])); rneg = std::max(rneg, ~scene_ordered(keys[i])); }
    for (uint32_t i = 0; i < n; ++i) out[i] = scene_bucket(keys[i], rmax, rneg);
}
int emu_scene_destroy(EmuDevice* d, uint32_t scene) {
    if (scene >= d->scenes.size() || !d->scenes[scene]) return VKR_ERROR_INVALID_ARGUMENT;
    if (d->lastScene == d->scenes[scene]) d->lastScene = nullptr;
    delete d->scenes[scene];
    d->scenes[scene] = nullptr;
    return 0;
}
int emu_scene_set_transforms(EmuDevice* d, uint32_t scene, uint32_t first, uint32_t count, const float* local) {
    if (scene >= d->scenes.size() || !d->scenes[scene] || (uint64_t)first + count > d->scenes[scene]->parent.size()) return VKR_ERROR_INVALID_ARGUMENT;
    if (count) std::memcpy(&d->scenes[scene]->local[(size_t)first * 16], local, (size_t)count * 64);
    return 0;
}
int emu_draw_scene(EmuDevice* d, uint32_t scene, const VkrGPUSceneData* sceneData, const float* planes, const VkrFrameParams* fp) {
    if (scene >= d->scenes.size() || !d->scenes[scene]) return VKR_ERROR_INVALID_ARGUMENT;
    EmuScene& s = *d->scenes[scene];
    FramePlan& plan = d->plan;
    for (DMaterial& m : d->materials) material_constants(m, d->tex.data(), d->lut);
    const size_t n = s.draws.size();
    plan.shadowLayers = fp->enableShadows ? (fp->shadowMode < 3 ? 1 : VKR_NUM_CASCADES) : 0;
    const uint32_t passes = 1u + (uint32_t)plan.shadowLayers;
    SceneFrame f;
    std::memcpy(f.viewproj, sceneData->viewproj, 64);
    std::memcpy(f.lightViewproj, sceneData->lightViewproj, sizeof(f.lightViewproj));
    std::memcpy(f.planes, planes, 96);
    f.nDraws = (uint32_t)n; f.opaqueCount = s.opaqueCount; f.shadowLayers = (uint32_t)plan.shadowLayers; f.pad = 0;
    for (size_t i = 0; i < s.parent.size(); ++i) scene_node_world(s.local.data(), s.parent.data(), s.world.data(), (uint32_t)i);   // scene_nodes_kernel (any parents-first order)
    plan.objects.resize(n); plan.shadowMvp.assign((size_t)plan.shadowLayers * n * 16, 0.0f);
    std::vector<float> keys((size_t)passes * n);
    std::vector<uint32_t> visTris(n), order((size_t)passes * n, 0u);
    for (size_t i = 0; i < n; ++i) {                           // scene_draws_kernel
        s.tmpl[i].flags = draw_flags(d->materials[s.surfaces[i].material], d->tex);
        scene_draw_body(f, s.draws[i], s.world.data(), s.tmpl[i], (uint32_t)i, plan.objects[i], plan.shadowMvp.data(), keys.data(), visTris.data());
    }
    plan.slotBase.assign((size_t)passes * n, 0xFFFFFFFFu);
    plan.passSlots.assign(passes, 0u);
    uint32_t shown = 0, tris = 0;
    plan.anyCutout = false;
    for (size_t i = 0; i < n; ++i) {                           // scene_scan_kernel, first part
        plan.objects[i].primBase = tris; tris += visTris[i];
        if (!(plan.objects[i].flags & DRAW_HIDDEN)) shown++;
        if (plan.objects[i].flags & 2u) plan.anyCutout = true;
    }
    for (uint32_t p = 0; p < passes; ++p) {                    // scene_rank_kernel, then scene_scan_kernel's per-pass part
        const uint32_t count = p == 0 ? s.opaqueCount : (uint32_t)n;
        const float* kp = keys.data() + (size_t)p * n;
        if (n <= SCENE_RANK_MAX) {
            for (uint32_t i = 0; i < count; ++i) order[(size_t)p * n + scene_rank(kp, count, i)] = i;
        } else {                                               // many draws: scene_draws_kernel's key range, then the three bucket kernels
            const float inf = bits2f(0x7F800000u);
            uint32_t rmax = 0, rneg = 0;
            for (uint32_t i = 0; i < count; ++i)
                if (kp[i] != inf) { rmax = std::max(rmax, scene_ordered(kp[i])); rneg = std::max(rneg, ~scene_ordered(kp[i])); }
            std::vector<uint32_t> start(SCENE_BUCKETS + 1, 0u);
            for (uint32_t i = 0; i < count; ++i) start[scene_bucket(kp[i], rmax, rneg) + 1]++;
            for (uint32_t b = 0; b < SCENE_BUCKETS; ++b) start[b + 1] += start[b];
            for (uint32_t i = 0; i < count; ++i) order[(size_t)p * n + start[scene_bucket(kp[i], rmax, rneg)]++] = i;
            // (front to back up to the width of a bucket: a draw never precedes one whose key is more than a bucket nearer)
            const float width = (scene_unordered(rmax) - scene_unordered(~rneg)) / (float)(SCENE_BUCKETS - 1u);
            float nearest = -inf;
            for (uint32_t k = 0; k < count; ++k) {
                const float key = kp[order[(size_t)p * n + k]];
                if (key == inf) continue;
                if (key < nearest - 1.001f * width - 1e-6f * std::fabs(nearest)) { d->err = "bucket order is not front to back"; return VKR_ERROR_INVALID_ARGUMENT; }
                nearest = std::max(nearest, key);
            }
        }
        {   // the order must be a permutation of the pass's draws: every draw gets its slots exactly once
            std::vector<uint8_t> seen(count, 0);
            for (uint32_t k = 0; k < count; ++k) { const uint32_t o = order[(size_t)p * n + k]; if (o >= count || seen[o]) { d->err = "processing order is not a permutation"; return VKR_ERROR_INVALID_ARGUMENT; } seen[o] = 1; }
        }
        uint32_t slot = 0;
        for (uint32_t k = 0; k < count; ++k) { const uint32_t o = order[(size_t)p * n + k]; plan.slotBase[(size_t)p * n + o] = slot; slot += s.draws[o].nClusters; }
        plan.passSlots[p] = slot;
    }
    plan.clusterBase = s.clusterBase; plan.seqSlotBase = s.seqSlotBase;
    plan.primBase.assign(n + 1, 0u);
    plan.opaqueCount = s.opaqueCount; plan.opaqueClusters = s.opaqueClusters; plan.totalTris = s.totalTris;
    plan.bandMain.assign(1, 0); plan.bandShadow.assign(1, 0);
    plan_constants(sceneData, fp, d->W, d->H, d->S, n, plan);
    d->lastScene = &s;
    const int rc = emu_execute(d, n);
    d->stats.drawcall_count = (int)shown; d->stats.triangle_count = (int)tris;
    return rc;
}
int emu_scene_draw_list(EmuDevice* d, uint32_t scene, VkrRenderObject* out, size_t capacity, size_t* count) {
    if (scene >= d->scenes.size() || !d->scenes[scene] || d->lastScene != d->scenes[scene] || !count) return VKR_ERROR_INVALID_ARGUMENT;
    const EmuScene& s = *d->scenes[scene];
    size_t k = 0;
    for (size_t i = 0; i < s.draws.size(); ++i) {
        if (d->plan.objects[i].flags & DRAW_HIDDEN) continue;
        if (k < capacity) {
            out[k].mesh = s.surfaces[i].mesh; out[k].material = s.surfaces[i].material;
            std::memcpy(out[k].transform, d->plan.objects[i].world, 64);
            out[k].indexCount = s.surfaces[i].indexCount; out[k].firstIndex = s.surfaces[i].firstIndex;
        }
        ++k;
    }
    *count = k;
    return 0;
}
int emu_readback(EmuDevice* d, int what, void* dst, size_t bytes) {
    size_t px = (size_t)d->W * d->H;
    switch (what) {
        case VKR_TARGET_RGBA8: if (bytes < px * 4) return -1; std::memcpy(dst, d->rgba8.data(), px * 4); return 0;
        case VKR_TARGET_RGB8:
            if (bytes < px * 3) return -1;
            for (size_t i = 0; i < px; ++i) std::memcpy((uint8_t*)dst + 3 * i, (const uint8_t*)d->rgba8.data() + 4 * i, 3);
            return 0;
        case VKR_TARGET_HDR16F: if (bytes < px * 8) return -1; std::memcpy(dst, d->hdr.data(), px * 8); return 0;
        case VKR_TARGET_SHADOW: std::memcpy(dst, d->shadowBits.data(), d->shadowBits.size() * 4); return 0;
        case VKR_TARGET_DEPTH: for (size_t i = 0; i < px; ++i) ((uint32_t*)dst)[i] = (uint32_t)(d->vis[i] >> 32); return 0;
        case VKR_TARGET_PRIMID:
            for (size_t i = 0; i < px; ++i) {
                uint32_t lo = (uint32_t)d->vis[i];
                if (lo == VIS_EMPTY_LO) { ((uint32_t*)dst)[i] = 0xFFFFFFFFu; continue; }
                uint32_t seq = 0xFFFFFFFEu - lo;
                const uint2 ch = d->chunkObj[seq >> 8];
                ((uint32_t*)dst)[i] = d->plan.objects[ch.x].primBase + ch.y + (seq & 255u);
            }
            return 0;
    }
    return -1;
}
int emu_stats(EmuDevice* d, VkrEngineStats* out) { *out = d->stats; return 0; }
int emu_set_shadow_exchange(EmuDevice* d, VkrShadowExchangeFn fn, void* user) { d->shadowExchange = fn; d->shadowExchangeUser = user; return 0; }
int emu_set_shard(EmuDevice* d, int rank, int nranks) { d->shardRank = rank; d->shardRanks = nranks; return 0; }
// row split points of the last frame: out[0..nranks] frame rows, out[nranks+1 .. 2*nranks+1] shadow rows
int emu_bands(EmuDevice* d, int32_t* out) {
    for (size_t i = 0; i < d->plan.bandMain.size(); ++i) out[i] = d->plan.bandMain[i];
    for (size_t i = 0; i < d->plan.bandShadow.size(); ++i) out[d->plan.bandMain.size() + i] = d->plan.bandShadow[i];
    return (int)d->plan.bandMain.size();
}
void* emu_host_ptr(EmuDevice* d, int what) { return what == VKR_TARGET_RGBA8 ? (void*)d->rgba8.data() : nullptr; }
// Test hook: build_clusters (vkr_frameprep.h) on a bare index / position array.  Returns the number of clusters; when the
// output arrays are large enough (capacities in elements) the cluster data is copied out: verts (concatenated distinct-vertex
// lists), tris / orig (clusters * CL_TRIS), info (2 u32 per cluster), bounds (8 floats per cluster).
int emu_build_clusters(const uint32_t* idx, uint32_t triCount, uint32_t nv, const float* pos4, uint32_t* verts, size_t vertsCap,
                       uint32_t* tris, uint32_t* orig, size_t trisCap, uint32_t* info, float* bounds, size_t clustersCap, size_t* vertsUsed) {
    ClusterData cd;
    build_clusters(idx, triCount, nv, reinterpret_cast<const float4*>(pos4), cd);
    const size_t nc = cd.info.size();
    if (vertsUsed) *vertsUsed = cd.verts.size();
    if (cd.verts.size() <= vertsCap && cd.tris.size() <= trisCap && nc <= clustersCap) {
        std::memcpy(verts, cd.verts.data(), cd.verts.size() * 4);
        std::memcpy(tris, cd.tris.data(), cd.tris.size() * 4);
        std::memcpy(orig, cd.orig.data(), cd.orig.size() * 4);
        std::memcpy(info, cd.info.data(), nc * 8);
        std::memcpy(bounds, cd.bounds.data(), nc * 32);
    }
    return (int)nc;
}
void emu_culled(EmuDevice* d, uint32_t out[2]) { out[0] = d->culled[0]; out[1] = d->culled[1]; }
const char* emu_last_error(EmuDevice* d) { return d->err.c_str(); }

}  // extern "C"

// Test hook for the facing cull of whole clusters (vkr_raster.cuh cone_culls): clusters are built from the index / position
// arrays, every cluster is run through make_cull_rec with its cone, and for every cluster the cone dropped, each of its
// triangles goes through the kernels' own vertex and set-up path WITHOUT the cone.  Returns the number of clusters the cone
// dropped; *violations counts triangles of such clusters that the per-triangle path would have kept (must be 0).
extern "C" int emu_cone_check(const uint32_t* idx, uint32_t triCount, uint32_t nv, const float* pos4, const float* mvp, int width, int height,
                              int cullFront, uint32_t* violations, uint32_t* clusters) {
    ClusterData cd;
    build_clusters(idx, triCount, nv, reinterpret_cast<const float4*>(pos4), cd);
    RasterTarget rt{};
    rt.width = width; rt.height = height; rt.sx0 = 0; rt.sy0 = 0; rt.sx1 = width; rt.sy1 = height;
    rt.halfW = 0.5f * width; rt.halfH = 0.5f * height; rt.halfW256 = 128.0f * width; rt.halfH256 = 128.0f * height;
    rt.cullFront = cullFront;
    float eye[8];
    eye_point(mvp, eye);
    int dropped = 0;
    *violations = 0;
    *clusters = (uint32_t)cd.info.size();
    const float4* P = reinterpret_cast<const float4*>(pos4);
    for (size_t c = 0; c < cd.info.size(); ++c) {
        const ClusterCone cone{cd.cone[3 * c], cd.cone[3 * c + 1], cd.cone[3 * c + 2]};
        const CullRec with = make_cull_rec(mvp, cd.bounds[2 * c], cd.bounds[2 * c + 1], rt, &cone, eye);
        const CullRec without = make_cull_rec(mvp, cd.bounds[2 * c], cd.bounds[2 * c + 1], rt);
        if (!(with.verdict == CLUSTER_CULLED && without.verdict != CLUSTER_CULLED)) continue;
        ++dropped;
        const uint32_t off = cd.info[c].x, nt = cd.info[c].y >> 16;
        for (uint32_t t = 0; t < nt; ++t) {
            const uint32_t pk = cd.tris[c * CL_TRIS + t];
            const uint32_t g[3] = {cd.verts[off + (pk & 127u)], cd.verts[off + ((pk >> 7) & 127u)], cd.verts[off + ((pk >> 14) & 127u)]};
            CachedVtx v[3];
            for (int k = 0; k < 3; ++k) v[k] = cluster_vertex(mvp, P[g[k]], rt);
            if (v[0].flag | v[1].flag | v[2].flag) { ++*violations; continue; }      // (allFront clusters have no clipped vertices near w = 0; count any)
            // the per-triangle facing decision alone (setup_snapped's first lines), without the scissor / sample tests
            const int64_t d1x = v[1].X - v[0].X, d1y = v[1].Y - v[0].Y, d2x = v[2].X - v[0].X, d2y = v[2].Y - v[0].Y;
            const int64_t A2 = d1x * d2y - d2x * d1y;
            if (A2 == 0) continue;
            const bool front = A2 < 0;
            if (!(cullFront ? front : !front)) ++*violations;
        }
    }
    return dropped;
}
