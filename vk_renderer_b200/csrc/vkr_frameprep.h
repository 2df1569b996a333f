// vkr_frameprep.h — host-side per-frame preparation (pure C++, no CUDA calls).
//
// Turns the reference's per-object command recording (src/vk_engine_run.cpp:284-311 for the
// geometry pass, :367-378 for the shadow pass: bind index buffer, push 80 bytes of constants,
// vkCmdDrawIndexed) into flat arrays the kernels consume: one DObject per draw, a prefix sum
// of triangle counts (submission-order primitive ids), GEOM_BLOCK-triangle work chunks and
// the per-cascade light MVPs.
//
// Must be compiled without FMA contraction (-ffp-contract=off): the matrices computed here
// feed the bit-exact position path.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "vkr_internal.h"

namespace vkr {

// glm mat4*mat4 association: ((A0*b0 + A1*b1) + A2*b2) + A3*b3, every product and sum rounded (no FMA).
// Column j of the result is a 4-wide expression of A's columns: written with SSE on x86 (mulps/addps round exactly
// like the scalar operations), scalar elsewhere.
#if defined(__SSE__)
}  // namespace vkr
#include <xmmintrin.h>
namespace vkr {
inline void mat_mul(const float* a, const float* b, float* r) {
    const __m128 a0 = _mm_loadu_ps(a), a1 = _mm_loadu_ps(a + 4), a2 = _mm_loadu_ps(a + 8), a3 = _mm_loadu_ps(a + 12);
    for (int j = 0; j < 4; ++j) {
        const __m128 t = _mm_add_ps(_mm_mul_ps(a0, _mm_set1_ps(b[j * 4 + 0])), _mm_mul_ps(a1, _mm_set1_ps(b[j * 4 + 1])));
        const __m128 u = _mm_add_ps(t, _mm_mul_ps(a2, _mm_set1_ps(b[j * 4 + 2])));
        _mm_storeu_ps(r + j * 4, _mm_add_ps(u, _mm_mul_ps(a3, _mm_set1_ps(b[j * 4 + 3]))));
    }
}
#else
inline void mat_mul(const float* a, const float* b, float* r) {
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i)
            r[j * 4 + i] = ((a[0 * 4 + i] * b[j * 4 + 0] + a[1 * 4 + i] * b[j * 4 + 1]) + a[2 * 4 + i] * b[j * 4 + 2]) +
                           a[3 * 4 + i] * b[j * 4 + 3];
}
#endif

// inverse(view)[3].xyz by cofactors (shaders/mesh_pbr.frag:187 evaluates this per fragment).
inline void camera_position(const float* m, float out[3]) {
    double inv12 = -(double)m[4] * m[9] * m[14] + (double)m[4] * m[10] * m[13] + (double)m[8] * m[5] * m[14] -
                   (double)m[8] * m[6] * m[13] - (double)m[12] * m[5] * m[10] + (double)m[12] * m[6] * m[9];
    double inv13 = (double)m[0] * m[9] * m[14] - (double)m[0] * m[10] * m[13] - (double)m[8] * m[1] * m[14] +
                   (double)m[8] * m[2] * m[13] + (double)m[12] * m[1] * m[10] - (double)m[12] * m[2] * m[9];
    double inv14 = -(double)m[0] * m[5] * m[14] + (double)m[0] * m[6] * m[13] + (double)m[4] * m[1] * m[14] -
                   (double)m[4] * m[2] * m[13] - (double)m[12] * m[1] * m[6] + (double)m[12] * m[2] * m[5];
    double inv15 = (double)m[0] * m[5] * m[10] - (double)m[0] * m[6] * m[9] - (double)m[4] * m[1] * m[10] +
                   (double)m[4] * m[2] * m[9] + (double)m[8] * m[1] * m[6] - (double)m[8] * m[2] * m[5];
    // column 3 of the inverse = (inv[12], inv[13], inv[14], inv[15]) / det; det*inv[15]... use w-normalised form
    double inv0 = (double)m[5] * m[10] * m[15] - (double)m[5] * m[11] * m[14] - (double)m[9] * m[6] * m[15] +
                  (double)m[9] * m[7] * m[14] + (double)m[13] * m[6] * m[11] - (double)m[13] * m[7] * m[10];
    double inv4 = -(double)m[4] * m[10] * m[15] + (double)m[4] * m[11] * m[14] + (double)m[8] * m[6] * m[15] -
                  (double)m[8] * m[7] * m[14] - (double)m[12] * m[6] * m[11] + (double)m[12] * m[7] * m[10];
    double inv8 = (double)m[4] * m[9] * m[15] - (double)m[4] * m[11] * m[13] - (double)m[8] * m[5] * m[15] +
                  (double)m[8] * m[7] * m[13] + (double)m[12] * m[5] * m[11] - (double)m[12] * m[7] * m[9];
    double det = (double)m[0] * inv0 + (double)m[1] * inv4 + (double)m[2] * inv8 + (double)m[3] * inv12;
    (void)inv15;
    out[0] = (float)(inv12 / det);
    out[1] = (float)(inv13 / det);
    out[2] = (float)(inv14 / det);
}

// Host-side builder of the per-range cluster data (DObject::clVerts / clTris / clOrig / clInfo / clBounds).
struct ClusterData {
    std::vector<uint32_t> verts;     // concatenated distinct-vertex lists
    std::vector<uint32_t> tris;      // nClusters * CL_TRIS packed local ids
    std::vector<uint32_t> orig;      // nClusters * CL_TRIS submission-order triangle numbers
    std::vector<uint2> info;         // per cluster
    std::vector<float4> bounds;      // per cluster: min, max
    std::vector<uint4> blob;         // per cluster CL_BLOB_BYTES: interleaved (tris, orig), then the gathered positions
    std::vector<float4> cone;        // per cluster: (axis, cos alpha), (centre, radius), (Nmin, sin alpha, 0, 0) — ClusterCone
};

// Spatial order of a draw's triangles for the greedy cut below: a k-d partition of the centroids whose split points are
// multiples of CL_TRIS.  A range that needs k = ceil(n / CL_TRIS) clusters is split along the longest axis of its centroid
// bounds into the floor(k / 2) * CL_TRIS lowest triangles and the rest, recursively, so every leaf is one full cluster (the
// last one takes the remainder) and a compact patch of surface.  (A Morton curve, used before, jumps: on BASELINE config 4's
// sphere half of the clusters were two patches in one box, median box diagonal 0.93 of the radius against 0.5 now, and a
// box that is too big on screen is never tried against the depth buffer, vkr_raster.cuh OCC_MAX_PIXELS.)
// Deterministic: ties broken by the triangle number, leaves listed in submission order.
inline void spatial_order(const uint32_t* idx, uint32_t triCount, const float4* pos, std::vector<uint32_t>& order) {
    order.resize(triCount);
    for (uint32_t t = 0; t < triCount; ++t) order[t] = t;
    if (triCount <= (uint32_t)CL_TRIS) return;
    std::vector<float> cen((size_t)triCount * 3);
    for (uint32_t t = 0; t < triCount; ++t) {
        const float4 a = pos[idx[3 * (size_t)t]], b = pos[idx[3 * (size_t)t + 1]], c = pos[idx[3 * (size_t)t + 2]];
        const float v[3] = {(a.x + b.x + c.x) * (1.0f / 3.0f), (a.y + b.y + c.y) * (1.0f / 3.0f), (a.z + b.z + c.z) * (1.0f / 3.0f)};
        for (int k = 0; k < 3; ++k) cen[(size_t)t * 3 + k] = std::isfinite(v[k]) ? v[k] : 0.0f;     // a strict weak order needs no NaN
    }
    struct Span { uint32_t begin, end; };
    std::vector<Span> stack{{0u, triCount}};
    while (!stack.empty()) {
        const Span sp = stack.back();
        stack.pop_back();
        const uint32_t n = sp.end - sp.begin;
        if (n <= (uint32_t)CL_TRIS) { std::sort(order.begin() + sp.begin, order.begin() + sp.end); continue; }
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = sp.begin; i < sp.end; ++i)
            for (int k = 0; k < 3; ++k) { const float v = cen[(size_t)order[i] * 3 + k]; lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v); }
        int axis = 0;
        if (hi[1] - lo[1] > hi[axis] - lo[axis]) axis = 1;
        if (hi[2] - lo[2] > hi[axis] - lo[axis]) axis = 2;
        const uint32_t clusters = (n + (uint32_t)CL_TRIS - 1) / (uint32_t)CL_TRIS;
        const uint32_t mid = sp.begin + (clusters / 2) * (uint32_t)CL_TRIS;
        std::nth_element(order.begin() + sp.begin, order.begin() + mid, order.begin() + sp.end, [&](uint32_t x, uint32_t y) {
            const float cx = cen[(size_t)x * 3 + axis], cy = cen[(size_t)y * 3 + axis];
            return cx != cy ? cx < cy : x < y;
        });
        stack.push_back({sp.begin, mid});
        stack.push_back({mid, sp.end});
    }
}

// `pos` are the mesh's positions (x,y,z,1).  Triangles are put in spatial_order, then cut greedily: a cluster
// closes when it holds CL_TRIS triangles or the next triangle would take it past CL_VERTS vertices.
inline void build_clusters(const uint32_t* idx, uint32_t triCount, uint32_t nv, const float4* pos, ClusterData& out) {
    out.verts.clear(); out.tris.clear(); out.orig.clear(); out.info.clear(); out.bounds.clear(); out.blob.clear(); out.cone.clear();
    std::vector<uint32_t> order;
    spatial_order(idx, triCount, pos, order);
    std::vector<uint32_t> stamp(nv, 0xFFFFFFFFu), local(nv, 0u);
    uint32_t s = 0, c = 0;
    while (s < triCount) {
        const uint32_t off = (uint32_t)out.verts.size();
        uint32_t count = 0, nt = 0;
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        out.tris.resize((size_t)(c + 1) * CL_TRIS, 0u);
        out.orig.resize((size_t)(c + 1) * CL_TRIS, 0u);
        while (s < triCount && nt < (uint32_t)CL_TRIS) {
            const uint32_t t = order[s];
            uint32_t fresh = 0;
            for (int k = 0; k < 3; ++k) {
                const uint32_t v = idx[3 * (size_t)t + k];
                bool seen = stamp[v] == c;
                for (int j = 0; j < k && !seen; ++j) seen = idx[3 * (size_t)t + j] == v;
                if (!seen) ++fresh;
            }
            if (count + fresh > (uint32_t)CL_VERTS) break;
            uint32_t l[3];
            for (int k = 0; k < 3; ++k) {
                const uint32_t v = idx[3 * (size_t)t + k];
                if (stamp[v] != c) {
                    stamp[v] = c; local[v] = count++; out.verts.push_back(v);
                    const float4 p = pos[v];
                    const float pc[3] = {p.x, p.y, p.z};
                    for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], pc[a]); hi[a] = std::max(hi[a], pc[a]); }
                }
                l[k] = local[v];
            }
            out.tris[(size_t)c * CL_TRIS + nt] = l[0] | (l[1] << 7) | (l[2] << 14);
            out.orig[(size_t)c * CL_TRIS + nt] = t;
            ++nt; ++s;
        }
        out.info.push_back(uint2{off, count | (nt << 16)});
        // (positions are required to be finite; min/max skip NaNs)
        out.bounds.push_back(make_float4(lo[0], lo[1], lo[2], 0.0f));
        out.bounds.push_back(make_float4(hi[0], hi[1], hi[2], 0.0f));
        {   // normal cone + bounding sphere (ClusterCone), rounded to the safe side; a degenerate triangle or a cone wider than
            // ~80 degrees switches the facing cull off for the cluster (Nmin = 0)
            double ax = 0, ay = 0, az = 0, nmin = INFINITY, lmax = 0;
            std::vector<double> nn((size_t)nt * 3);
            for (uint32_t k = 0; k < nt; ++k) {
                const uint32_t pk = out.tris[(size_t)c * CL_TRIS + k];
                const float4 q0 = pos[out.verts[off + (pk & 127u)]], q1 = pos[out.verts[off + ((pk >> 7) & 127u)]], q2 = pos[out.verts[off + ((pk >> 14) & 127u)]];
                const double e1[3] = {(double)q1.x - q0.x, (double)q1.y - q0.y, (double)q1.z - q0.z}, e2[3] = {(double)q2.x - q0.x, (double)q2.y - q0.y, (double)q2.z - q0.z};
                const double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
                const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
                nmin = std::min(nmin, len);
                const double e3[3] = {(double)q2.x - q1.x, (double)q2.y - q1.y, (double)q2.z - q1.z};
                lmax = std::max({lmax, std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]), std::sqrt(e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2]),
                                 std::sqrt(e3[0] * e3[0] + e3[1] * e3[1] + e3[2] * e3[2])});
                for (int a = 0; a < 3; ++a) nn[(size_t)k * 3 + a] = len > 0 ? n[a] / len : 0.0;
                ax += nn[(size_t)k * 3]; ay += nn[(size_t)k * 3 + 1]; az += nn[(size_t)k * 3 + 2];
            }
            const double al = std::sqrt(ax * ax + ay * ay + az * az);
            double cosA = -1.0;
            if (al > 1e-12 && nmin > 0.0 && std::isfinite(nmin)) {
                ax /= al; ay /= al; az /= al;
                cosA = 1.0;
                for (uint32_t k = 0; k < nt; ++k) cosA = std::min(cosA, nn[(size_t)k * 3] * ax + nn[(size_t)k * 3 + 1] * ay + nn[(size_t)k * 3 + 2] * az);
            }
            const double cx = 0.5 * ((double)lo[0] + hi[0]), cy = 0.5 * ((double)lo[1] + hi[1]), cz = 0.5 * ((double)lo[2] + hi[2]);
            double rad = 0.0;
            for (uint32_t k = 0; k < count; ++k) {
                const float4 q = pos[out.verts[off + k]];
                rad = std::max(rad, std::sqrt((q.x - cx) * (q.x - cx) + (q.y - cy) * (q.y - cy) + (q.z - cz) * (q.z - cz)));
            }
            const bool usable = cosA > 0.17 && std::isfinite(rad) && std::isfinite(cx + cy + cz);
            const double cosSafe = usable ? cosA * (1.0 - 1e-5) - 1e-6 : 0.0, sinSafe = usable ? std::min(1.0, std::sqrt(std::max(0.0, 1.0 - cosSafe * cosSafe)) * (1.0 + 1e-5) + 1e-6) : 1.0;
            out.cone.push_back(make_float4((float)ax, (float)ay, (float)az, (float)cosSafe));
            out.cone.push_back(make_float4((float)cx, (float)cy, (float)cz, (float)(rad * (1.0 + 1e-5) + 1e-30)));
            out.cone.push_back(make_float4(usable && std::isfinite(lmax) ? (float)(nmin * (1.0 - 1e-5)) : 0.0f, (float)sinSafe, (float)(lmax * (1.0 + 1e-5)), 0.0f));
        }
        {   // the cluster's self-contained block
            const size_t b0 = out.blob.size();
            out.blob.resize(b0 + CL_BLOB_BYTES / 16, uint4{0u, 0u, 0u, 0u});
            uint32_t* tw = reinterpret_cast<uint32_t*>(out.blob.data() + b0);
            for (uint32_t k = 0; k < nt; ++k) { tw[2 * k] = out.tris[(size_t)c * CL_TRIS + k]; tw[2 * k + 1] = out.orig[(size_t)c * CL_TRIS + k]; }
            float4* pw = reinterpret_cast<float4*>(out.blob.data() + b0 + CL_BLOB_TRI_BYTES / 16);
            for (uint32_t k = 0; k < count; ++k) { const float4 q = pos[out.verts[off + k]]; pw[k] = make_float4(q.x, q.y, q.z, 1.0f); }
        }
        ++c;
    }
}

struct ClusterRange {
    const uint32_t* verts; const uint32_t* tris; const uint32_t* orig; const uint2* info; const float4* bounds; const uint4* blob; const float4* cone;
    uint32_t nClusters;
    float center[3], radius;          // bounding sphere of the whole range (object space), for the processing order
};
// Implemented by the owner of the device (or emulated) memory: returns the cached range, building it on first use.
struct ClusterProvider {
    virtual ~ClusterProvider() {}
    virtual bool get(uint32_t mesh, uint32_t firstIndex, uint32_t indexCount, ClusterRange& out) = 0;
};
inline void range_sphere(const ClusterData& cd, ClusterRange& r) {
    r.nClusters = (uint32_t)cd.info.size();
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (size_t c = 0; c + 1 < cd.bounds.size(); c += 2) {
        const float4 a = cd.bounds[c], b = cd.bounds[c + 1];
        lo[0] = std::min(lo[0], a.x); lo[1] = std::min(lo[1], a.y); lo[2] = std::min(lo[2], a.z);
        hi[0] = std::max(hi[0], b.x); hi[1] = std::max(hi[1], b.y); hi[2] = std::max(hi[2], b.z);
    }
    float r2 = 0.0f;
    for (int k = 0; k < 3; ++k) {
        r.center[k] = std::isfinite(lo[k]) && std::isfinite(hi[k]) ? 0.5f * (lo[k] + hi[k]) : 0.0f;
        const float e = std::isfinite(lo[k]) && std::isfinite(hi[k]) ? 0.5f * (hi[k] - lo[k]) : 0.0f;
        r2 += e * e;
    }
    r.radius = std::sqrt(r2);
}

// DObject::flags of a draw with this material: bit0 Transparent pass, bit1 alpha-tested (its albedo texture has texels with
// alpha < 0.5, DTexture flag bit2), bit2 some texture the fragment shaders sample has a mip chain, i.e. the shading pass needs
// uv derivatives for this draw (a 1x1 or single-level image is sampled without them, vkr_shade.cuh texture2d)
inline uint32_t draw_flags(const DMaterial& mat, const std::vector<DTexture>& textures) {
    uint32_t flags = mat.pass == VKR_PASS_TRANSPARENT ? 1u : 0u;
    const int aid = mat.albedoID;
    if (aid >= 0 && aid < (int)textures.size() && textures[aid].data != nullptr && (textures[aid].flags & 4u)) flags |= 2u;
    const int ids[2] = {aid, mat.metalRoughID};
    for (int id : ids)
        if (id >= 0 && id < (int)textures.size() && textures[id].data != nullptr && textures[id].levels > 1u && (textures[id].w | textures[id].h) != 1u) flags |= 4u;
    return flags;
}

struct FramePlan {
    std::vector<DObject> objects;
    std::vector<uint32_t> primBase;     // nobj + 1 (dense ids, for stats / readback)
    std::vector<uint32_t> clusterBase;  // nobj + 1: first frame-cluster index of every draw
    std::vector<uint32_t> seqSlotBase;  // nobj + 1: first 256-number sequence slot of every draw
    uint32_t opaqueClusters = 0;        // clusters of the opaque prefix
    bool anyCutout = false;             // some draw needs the raster-stage alpha test
    std::vector<float> shadowMvp;       // layers * nobj * 16
    // Processing order (front to back, so the cluster occlusion test finds the occluders already drawn):
    // slotBase[p*nobj + i] = first geometry-CTA slot of draw i in pass p (0 = main, 1.. = shadow layers),
    // 0xFFFFFFFF when the draw is not part of the pass.  Primitive ids do not depend on it.
    std::vector<uint32_t> slotBase;
    std::vector<int32_t> bandMain, bandShadow;   // sort-first: row split points (nranks + 1) of the frame / the shadow layers
    std::vector<double> rowCost;
    int32_t activeMain[2] = {0, 0}, activeShadow[2] = {0, 0};   // rows reachable by some draw's bounding sphere
    std::vector<uint32_t> passSlots;    // per pass: geometry warp slots in use (= clusters of the draws taking part)
    std::vector<std::pair<float, uint32_t>> sortScratch;
    // the last frame planned into this object, per pass: the draws that took part (submission order) and their processing order
    // (positions into that list) — the next frame's keys are usually in nearly the same order, which an insertion sort finishes in O(n)
    std::vector<std::vector<uint32_t>> prevCand, prevOrder;
    std::vector<float> sphereScratch;
    int shadowLayers = 0;
    uint32_t totalTris = 0;
    uint32_t opaqueCount = 0;           // objects [0, opaqueCount) are MainColor, the rest Transparent
    ShadeConst sc;
    RasterTarget mainTarget;
    RasterTarget shadowTarget;
    // vkr_draw_scene (update_scene on the device): `scene` is the device library's scene record and the per-draw arrays above
    // (objects, clusterBase, seqSlotBase, shadowMvp, slotBase) stay empty — the device writes them from the record and
    // `sceneFrame`; the counts a launch needs (passSlots, opaqueClusters, ...) are the scene's static totals.
    void* scene = nullptr;
    SceneFrame sceneFrame;
};

// The frame's constants: what the shading pass and the two raster passes are told (O(1) in the draw count).
inline void plan_constants(const VkrGPUSceneData* scene, const VkrFrameParams* fp, uint32_t W, uint32_t H, uint32_t S, size_t n, FramePlan& plan) {
    ShadeConst& sc = plan.sc;
    sc.scene = *scene;
    camera_position(scene->view, sc.camPos);
    {   // outLightVec = normalize(-sunlightDirection.xyz) (shaders/mesh.vert:21)
        float x = -scene->sunlightDirection[0], y = -scene->sunlightDirection[1], z = -scene->sunlightDirection[2];
        float inv = 1.0f / std::sqrt(x * x + y * y + z * z);
        sc.lightVec[0] = x * inv; sc.lightVec[1] = y * inv; sc.lightVec[2] = z * inv;
    }
    const float bias[16] = {0.5f, 0, 0, 0, 0, 0.5f, 0, 0, 0, 0, 1.0f, 0, 0.5f, 0.5f, 0, 1.0f};   // mesh_pbr.frag:31-35
    for (int l = 0; l < VKR_NUM_CASCADES; ++l) mat_mul(bias, scene->lightViewproj[l], sc.shadowMat[l]);
    std::memcpy(sc.bgTop, fp->background.data1, 16);
    std::memcpy(sc.bgBottom, fp->background.data2, 16);
    sc.width = (int32_t)W; sc.height = (int32_t)H; sc.shadowExtent = (int32_t)S;
    sc.sx0 = 0; sc.sy0 = 0; sc.sx1 = (int32_t)W; sc.sy1 = (int32_t)H;
    if (fp->scissor[2] > 0 && fp->scissor[3] > 0) {
        sc.sx0 = std::max(0, fp->scissor[0]); sc.sy0 = std::max(0, fp->scissor[1]);
        sc.sx1 = std::min((int32_t)W, fp->scissor[0] + fp->scissor[2]);
        sc.sy1 = std::min((int32_t)H, fp->scissor[1] + fp->scissor[3]);
    }
    sc.enableBackground = fp->enableBackground; sc.enablePostprocess = fp->enablePostprocess;
    sc.enableSkybox = fp->enableSkybox; sc.fragShader = fp->fragShader;
    // the shaders read the flags from the UBO .w components (src/vk_engine_run.cpp:124-125)
    sc.shadowMode = (int32_t)scene->sunlightDirection[3];
    sc.enableShadows = scene->sunlightColor[3] > 0.5f ? 1 : 0;
    sc.nobj = (uint32_t)n;
    sc.clearVis = 0;                   // set by the device library under VKR_CONFIG_TRANSIENT_DEPTH
    sc.storeHdr = 1;                   // cleared by the device library under VKR_CONFIG_TRANSIENT_HDR
    sc.halfW = (float)W * 0.5f; sc.halfH = (float)H * 0.5f;
    sc.invHalfW = 1.0f / sc.halfW; sc.invHalfH = 1.0f / sc.halfH;
    {
        const float* P = scene->proj; const float* V = scene->view;
        for (int j = 0; j < 3; ++j) {
            double sx = (double)V[4 * j] / ((double)sc.halfW * P[0]), sy = (double)V[4 * j + 1] / ((double)sc.halfH * P[5]);
            double so = (double)V[4 * j] * ((double)P[8] - 1.0) / P[0] + (double)V[4 * j + 1] * ((double)P[9] - 1.0) / P[5] - (double)V[4 * j + 2];
            const double sgn = j == 1 ? -1.0 : 1.0;
            sc.skyX[j] = (float)(sgn * sx); sc.skyY[j] = (float)(sgn * sy); sc.skyO[j] = (float)(sgn * so);
        }
    }

    RasterTarget& mt = plan.mainTarget;
    mt.vis = nullptr; mt.depthBits = nullptr;
    mt.width = (int32_t)W; mt.height = (int32_t)H;
    mt.sx0 = sc.sx0; mt.sy0 = sc.sy0; mt.sx1 = sc.sx1; mt.sy1 = sc.sy1;
    mt.halfW = sc.halfW; mt.halfH = sc.halfH;
    mt.halfW256 = sc.halfW * 256.0f; mt.halfH256 = sc.halfH * 256.0f;
    mt.cullFront = 0;                                   // VK_CULL_MODE_BACK_BIT (src/vk_engine_init.cpp:536)
    RasterTarget& st = plan.shadowTarget;
    st = mt;
    st.width = st.height = (int32_t)S;
    st.sx0 = st.sy0 = 0; st.sx1 = st.sy1 = (int32_t)S;
    st.halfW = st.halfH = (float)S * 0.5f;
    st.halfW256 = st.halfH256 = (float)S * 128.0f;
    if (fp->shadowScissor[2] > 0 && fp->shadowScissor[3] > 0) {
        st.sx0 = std::max(0, fp->shadowScissor[0]); st.sy0 = std::max(0, fp->shadowScissor[1]);
        st.sx1 = std::min((int32_t)S, fp->shadowScissor[0] + fp->shadowScissor[2]);
        st.sy1 = std::min((int32_t)S, fp->shadowScissor[1] + fp->shadowScissor[3]);
    }
    st.cullFront = 1;                                   // VK_CULL_MODE_FRONT_BIT (src/vk_engine_init.cpp:441)
}

// Returns an empty string on success, else the reason the draw list is invalid.
inline std::string plan_frame(const VkrGPUSceneData* scene, const VkrRenderObject* objs, size_t n, const VkrFrameParams* fp,
                              const std::vector<DMesh>& meshes, const std::vector<DMaterial>& materials,
                              const std::vector<DTexture>& textures, ClusterProvider& clusters, uint32_t W, uint32_t H, uint32_t S, FramePlan& plan,
                              int shardRank = 0, int shardRanks = 1) {
    VkrFrameParams fpLocal = *fp;       // sort-first sharding rewrites the two scissors below
    fp = &fpLocal;
    plan.objects.resize(n);
    plan.primBase.resize(n + 1);
    plan.clusterBase.resize(n + 1);
    plan.seqSlotBase.resize(n + 1);
    uint64_t ncl = 0, nslot = 0;
    plan.shadowLayers = fp->enableShadows ? (fp->shadowMode < 3 ? 1 : VKR_NUM_CASCADES) : 0;
    plan.shadowMvp.resize((size_t)plan.shadowLayers * n * 16);
    uint64_t prim = 0;
    bool seenTransparent = false;
    std::vector<float>& sphere = plan.sphereScratch;
    sphere.resize(4 * n + 4);
    plan.opaqueCount = (uint32_t)n;
    plan.opaqueClusters = 0;
    plan.anyCutout = false;
    bool haveLast = false;
    uint32_t lastMesh = 0, lastFirst = 0, lastCount = 0;
    ClusterRange lastRange{};
    for (size_t i = 0; i < n; ++i) {
        const VkrRenderObject& ro = objs[i];
        if (ro.mesh >= meshes.size()) return "render object references an unknown mesh";
        if (ro.material >= materials.size()) return "render object references an unknown material";
        const DMesh& m = meshes[ro.mesh];
        if ((uint64_t)ro.firstIndex + ro.indexCount > m.ni) return "render object index range exceeds the mesh";
        const bool transparent = materials[ro.material].pass == VKR_PASS_TRANSPARENT;
        if (transparent && !seenTransparent) { seenTransparent = true; plan.opaqueCount = (uint32_t)i; }
        if (!transparent && seenTransparent)
            return "draw list must be sorted opaque-first as update_scene does (src/vk_engine_run.cpp:454-458)";
        DObject& o = plan.objects[i];
        mat_mul(scene->viewproj, ro.transform, o.mvp);
        std::memcpy(o.world, ro.transform, 64);
        o.pos = m.pos; o.attrA = m.attrA; o.attrB = m.attrB;
        o.idx = m.idx + ro.firstIndex;
        o.triCount = ro.indexCount / 3;
        // (instances of one range follow each other in most draw lists: the provider is asked once per run)
        if (o.triCount && !(haveLast && lastMesh == ro.mesh && lastFirst == ro.firstIndex && lastCount == ro.indexCount)) {
            if (!clusters.get(ro.mesh, ro.firstIndex, ro.indexCount, lastRange)) return "out of memory building clusters";
            haveLast = true; lastMesh = ro.mesh; lastFirst = ro.firstIndex; lastCount = ro.indexCount;
        }
        const ClusterRange cr = o.triCount ? lastRange : ClusterRange{};
        o.clVerts = cr.verts; o.clTris = cr.tris; o.clOrig = cr.orig; o.clInfo = cr.info; o.clBounds = cr.bounds; o.clBlob = cr.blob; o.clCone = cr.cone; o.pad = 0;
        // bounding sphere of the draw's range (object space): sort key of the processing order below
        sphere[4 * i] = cr.center[0]; sphere[4 * i + 1] = cr.center[1]; sphere[4 * i + 2] = cr.center[2]; sphere[4 * i + 3] = cr.radius;
        o.primBase = (uint32_t)prim;
        o.material = ro.material;
        o.flags = draw_flags(materials[ro.material], textures);
        if (o.flags & 2u) plan.anyCutout = true;
        plan.primBase[i] = (uint32_t)prim;
        plan.clusterBase[i] = (uint32_t)ncl;
        plan.seqSlotBase[i] = (uint32_t)nslot;
        ncl += o.triCount ? cr.nClusters : 0u;
        nslot += (o.triCount + SEQ_SLOT - 1) / SEQ_SLOT;
        if (!transparent) plan.opaqueClusters = (uint32_t)ncl;
        prim += o.triCount;
        if (prim >= 0xFFFFFFF0ull || nslot >= 0xFFFFFFull || ncl >= 0x7FFFFFFFull) return "too many primitives in one frame";
        for (int l = 0; l < plan.shadowLayers; ++l)
            mat_mul(scene->lightViewproj[l], ro.transform, &plan.shadowMvp[((size_t)l * n + i) * 16]);
    }
    plan.primBase[n] = (uint32_t)prim;
    plan.clusterBase[n] = (uint32_t)ncl;
    plan.seqSlotBase[n] = (uint32_t)nslot;
    plan.totalTris = (uint32_t)prim;

    // ---- sort-first sharding: split the rows of the frame and of the shadow map into one contiguous band per rank
    // with equal estimated cost (triangles whose draw reaches the row + pixels to shade), from the draws' bounding
    // spheres.  Every rank computes the same split from the same draw list.
    plan.bandMain.assign(1, 0); plan.bandShadow.assign(1, 0);
    if (shardRanks > 1) {
        for (int p = 0; p <= (plan.shadowLayers ? 1 : 0); ++p) {
            const uint32_t rows = p == 0 ? H : S;
            std::vector<double>& cost = plan.rowCost;
            cost.assign(rows + 1, 0.0);
            const size_t cnt = p == 0 ? plan.opaqueCount : n;
            for (size_t i = 0; i < cnt; ++i) {
                const float* m = p == 0 ? plan.objects[i].mvp : &plan.shadowMvp[i * 16];      // shadow: layer 0 decides for all layers
                const float cx = sphere[4 * i], cy = sphere[4 * i + 1], cz = sphere[4 * i + 2], rad = sphere[4 * i + 3];
                const float yc = m[1] * cx + m[5] * cy + m[9] * cz + m[13], wc = m[3] * cx + m[7] * cy + m[11] * cz + m[15];
                const float ry = rad * std::sqrt(m[1] * m[1] + m[5] * m[5] + m[9] * m[9]), rw = rad * std::sqrt(m[3] * m[3] + m[7] * m[7] + m[11] * m[11]);
                float r0 = 0.0f, r1 = (float)rows;
                if (wc - rw > 1e-6f) {       // wholly in front: interval bound of y/w
                    const float ylo = yc - ry, yhi = yc + ry, ilo = 1.0f / (wc - rw), ihi = 1.0f / (wc + rw);
                    const float nlo = std::min(ylo * ilo, ylo * ihi), nhi = std::max(yhi * ilo, yhi * ihi);
                    r0 = (nlo + 1.0f) * 0.5f * (float)rows - 1.0f; r1 = (nhi + 1.0f) * 0.5f * (float)rows + 1.0f;   // +-1: the reciprocals' rounding
                }
                if (!(r0 == r0) || !(r1 == r1)) { r0 = 0.0f; r1 = (float)rows; }
                if (r1 < 0.0f || r0 > (float)rows) continue;
                const uint32_t a = (uint32_t)std::max(0.0f, r0), b = (uint32_t)std::min((float)rows - 1.0f, std::max(0.0f, r1));   // >= 0: the cast floors
                const double per = (double)plan.objects[i].triCount / (double)(b - a + 1);
                cost[a] += per; cost[b + 1] -= per;            // difference array
            }
            const double perRow = p == 0 ? (double)W : 0.25 * (double)S;     // a shaded pixel costs about what a triangle costs; clearing a texel less
            double total = 0.0;
            {   // one sweep: rows any draw can reach (the rest of a shadow layer stays cleared on every rank and is never
                // exchanged), and cost[y] := cumulative cost through row y
                double run = 0.0;
                int32_t a0 = (int32_t)rows, a1 = 0;
                for (uint32_t y = 0; y < rows; ++y) {
                    run += cost[y];
                    if (run > 1e-9) { if (a0 > (int32_t)y) a0 = (int32_t)y; a1 = (int32_t)y + 1; }
                    total += run + perRow;
                    cost[y] = total;
                }
                if (a0 > a1) a0 = a1 = 0;
                int32_t* act = p == 0 ? plan.activeMain : plan.activeShadow;
                act[0] = std::max(0, a0 - 1); act[1] = std::min((int32_t)rows, a1 + 1);
            }
            std::vector<int32_t>& band = p == 0 ? plan.bandMain : plan.bandShadow;
            band.assign((size_t)shardRanks + 1, 0);
            uint32_t y = 0;
            for (int k = 1; k < shardRanks; ++k) {
                const double target = total * (double)k / (double)shardRanks;
                while (y + 1 < rows && cost[y] < target) ++y;
                // split after row y; keep every band non-empty
                const int32_t lo = band[k - 1] + 1, hi = (int32_t)rows - (shardRanks - k);
                band[k] = std::min(hi, std::max(lo, (int32_t)y + 1));
            }
            band[shardRanks] = (int32_t)rows;
            int32_t* sci = p == 0 ? fpLocal.scissor : fpLocal.shadowScissor;
            sci[0] = 0; sci[1] = band[shardRank]; sci[2] = (int32_t)(p == 0 ? W : S); sci[3] = band[shardRank + 1] - band[shardRank];
        }
    }

    // ---- processing order per pass; draws whose bounding sphere misses the pass's scissor rows are left out
    // (sort-first sharding: a rank only touches the draws that reach its band)
    plan.slotBase.assign((size_t)(1 + plan.shadowLayers) * n, 0xFFFFFFFFu);
    plan.passSlots.assign((size_t)(1 + plan.shadowLayers), 0u);
    for (int p = 0; p <= plan.shadowLayers; ++p) {
        auto& keys = plan.sortScratch;
        keys.clear();
        const size_t cnt = p == 0 ? plan.opaqueCount : n;
        // scissor rows of the pass as an NDC interval, widened by a pixel (vertex snapping, rounding)
        const int32_t* sci = p == 0 ? fp->scissor : fp->shadowScissor;
        const float extent = p == 0 ? (float)H : (float)S;
        const bool banded = sci[2] > 0 && sci[3] > 0;
        const float b0 = banded ? 2.0f * ((float)sci[1] - 1.0f) / extent - 1.0f : -2.0f;
        const float b1 = banded ? 2.0f * ((float)(sci[1] + sci[3]) + 1.0f) / extent - 1.0f : 2.0f;
        for (size_t i = 0; i < cnt; ++i) {
            const DObject& o = plan.objects[i];
            const float cx = sphere[4 * i], cy = sphere[4 * i + 1], cz = sphere[4 * i + 2], rad = sphere[4 * i + 3];
            const float* m = p == 0 ? o.mvp : &plan.shadowMvp[((size_t)(p - 1) * n + i) * 16];
            if (banded) {
                // every point of the sphere satisfies y - b*w > 0 (resp. < 0): nothing of it can land on a row <= b1 (>= b0);
                // y - b*w is linear in the object-space position, its range over the sphere is centre value +- r*|gradient|
                const float yc = m[1] * cx + m[5] * cy + m[9] * cz + m[13], wc = m[3] * cx + m[7] * cy + m[11] * cz + m[15];
                const float r1 = rad * 1.001f + 1e-6f;
                auto gradLen = [&](float b) {
                    const float gx = m[1] - b * m[3], gy = m[5] - b * m[7], gz = m[9] - b * m[11];
                    return std::sqrt(gx * gx + gy * gy + gz * gz);
                };
                const float slack = 1e-5f * (std::fabs(yc) + std::fabs(wc));
                if ((yc - b1 * wc) - r1 * gradLen(b1) > slack) continue;      // entirely past the last row
                if ((yc - b0 * wc) + r1 * gradLen(b0) < -slack) continue;     // entirely before the first row
            }
            // main pass: clip w (view depth); shadow pass (orthographic light): clip z
            const int r = p == 0 ? 3 : 2;
            const float d = m[r] * cx + m[4 + r] * cy + m[8 + r] * cz + m[12 + r];
            const float ext = rad * std::sqrt(m[r] * m[r] + m[4 + r] * m[4 + r] + m[8 + r] * m[8 + r]);
            float key = d - ext;
            if (!(key == key)) key = 0.0f;
            keys.emplace_back(key, (uint32_t)i);
        }
        // (key, submission index) pairs: the pair order is the stable order.  The unique sorted order is reached from the previous
        // frame's order when the same draws take part (an insertion sort, linear for a camera that moves a little), else by std::sort
        if ((int)plan.prevCand.size() <= p) { plan.prevCand.resize((size_t)p + 1); plan.prevOrder.resize((size_t)p + 1); }
        std::vector<uint32_t>& pc = plan.prevCand[(size_t)p];
        std::vector<uint32_t>& po = plan.prevOrder[(size_t)p];
        bool same = pc.size() == keys.size();
        for (size_t k = 0; same && k < keys.size(); ++k) same = pc[k] == keys[k].second;
        if (same) {
            size_t moves = 0;
            const size_t budget = 16 * po.size() + 64;          // an order that changed a lot (a cut to another camera) goes to std::sort
            for (size_t k = 1; k < po.size() && moves <= budget; ++k) {
                const uint32_t v = po[k];
                size_t j = k;
                while (j > 0 && keys[v] < keys[po[j - 1]]) { po[j] = po[j - 1]; --j; ++moves; }
                po[j] = v;
            }
            if (moves > budget) std::sort(po.begin(), po.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
        } else {
            pc.resize(keys.size()); po.resize(keys.size());
            for (size_t k = 0; k < keys.size(); ++k) { pc[k] = keys[k].second; po[k] = (uint32_t)k; }
            std::sort(po.begin(), po.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
        }
        uint32_t slot = 0;
        for (const uint32_t k : po) {
            const uint32_t obj = keys[k].second;
            plan.slotBase[(size_t)p * n + obj] = slot;
            slot += plan.clusterBase[obj + 1] - plan.clusterBase[obj];
        }
        plan.passSlots[p] = slot;
    }

    plan.scene = nullptr;
    plan_constants(scene, fp, W, H, S, n, plan);
    return std::string();
}

}  // namespace vkr
