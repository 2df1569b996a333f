// vkr_scene.cuh — update_scene on the device: per-thread bodies (host + device; the kernels around them are in
// vkr_scene_impl.cuh, the test-only host emulation runs the same bodies as loops).
//
// Replaces, for a scene graph handed over once (vkr_scene_create):
//   Node::refreshTransform                      src/vk_types.h:148-163       -> scene_node_world   (one thread per node, level by level)
//   VulkanEngine::is_visible                    src/vk_engine_run.cpp:461-480 -> scene_is_visible   (one thread per draw)
//   update_scene's cull loop + opaque-first sort src/vk_engine_run.cpp:447-458 -> the draws are kept in the reference's order
//                                                  (traversal order, MainColor first, stable) and a culled draw is only MARKED
//                                                  (DRAW_HIDDEN): every later stage skips it, nothing is compacted, so no count
//                                                  ever has to travel back to the host
//   plan_frame's per-draw work (vkr_frameprep.h) -> scene_draw_body: MVP and light MVPs with glm's association, the
//                                                  front-to-back key of every pass; scene_key_less + the scans in the kernels give
//                                                  the processing order and the dense primitive ids
//
// Arithmetic: the world matrices and the MVPs feed the bit-exact position path and is_visible decides which draws exist, so
// both are evaluated with explicitly rounded operations in exactly the order glm (and csrc/host/vk_math.h) uses:
//   mat4 * mat4   ((A0*b0 + A1*b1) + A2*b2) + A3*b3        detail/type_mat4x4.inl:606-625
//   mat4 * vec4   (m0*x + m1*y) + (m2*z + m3*w)            detail/type_mat4x4.inl:561-572
//   dot(vec3)     (a.x*b.x + a.y*b.y) + a.z*b.z ; length = sqrt(dot)
#pragma once
#include "vkr_raster.cuh"

namespace vkr {

#if defined(__CUDA_ARCH__)
VKR_HD float xsqrt(float a) { return __fsqrt_rn(a); }
#else
VKR_HD float xsqrt(float a) { return sqrtf(a); }
#endif

// r = a * b, glm association.  r must not alias a or b.
VKR_HD void scene_mat_mul(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ r) {
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i)
            r[j * 4 + i] = xadd(xadd(xadd(xmul(a[i], b[j * 4]), xmul(a[4 + i], b[j * 4 + 1])), xmul(a[8 + i], b[j * 4 + 2])), xmul(a[12 + i], b[j * 4 + 3]));
}

// worldTransform = parentMatrix * localTransform (src/vk_types.h:150); the root is refreshed with mat4(1) (src/vk_engine_run.cpp:441),
// and that product is evaluated too (it is the identity only for finite matrices).  The parent's world matrix must be final.
VKR_HD void scene_node_world(const float* __restrict__ local, const int32_t* __restrict__ parent, float* world, uint32_t node) {
    const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float a[16], out[16];
    const int32_t p = parent[node];
    for (int k = 0; k < 16; ++k) a[k] = p < 0 ? ident[k] : world[(size_t)p * 16 + k];
    scene_mat_mul(a, local + (size_t)node * 16, out);
    for (int k = 0; k < 16; ++k) world[(size_t)node * 16 + k] = out[k];
}

VKR_HD float scene_std_max(float a, float b) { return a < b ? b : a; }     // std::max: the first argument unless it is smaller

// VulkanEngine::is_visible (src/vk_engine_run.cpp:461-480), operation for operation.
VKR_HD bool scene_is_visible(const float* __restrict__ T, const float* __restrict__ origin, float sphereRadius, const float* __restrict__ planes) {
    float c[3];
    for (int i = 0; i < 3; ++i)
        c[i] = xadd(xadd(xmul(T[i], origin[0]), xmul(T[4 + i], origin[1])), xadd(xmul(T[8 + i], origin[2]), xmul(T[12 + i], 1.0f)));
    float s[3];
    for (int k = 0; k < 3; ++k) {
        const float x = T[4 * k], y = T[4 * k + 1], z = T[4 * k + 2];
        s[k] = xsqrt(xadd(xadd(xmul(x, x), xmul(y, y)), xmul(z, z)));
    }
    const float maxScale = scene_std_max(scene_std_max(s[0], s[1]), s[2]);
    const float globalRadius = xmul(sphereRadius, maxScale);
    for (int i = 0; i < 6; ++i) {
        const float* pl = planes + 4 * i;
        const float dist = xadd(xadd(xadd(xmul(pl[0], c[0]), xmul(pl[1], c[1])), xmul(pl[2], c[2])), pl[3]);
        if (dist < -globalRadius) return false;
    }
    return true;
}

// Front-to-back key of a draw in a pass (plan_frame's: nearest point of the bounding sphere along the view depth — clip w — for
// the main pass, along clip z for the orthographic light).  The processing order only decides how much the occlusion test of
// the geometry kernels finds already drawn: any order gives the same image.
VKR_HD float scene_order_key(const float* __restrict__ m, const float* __restrict__ center, float radius, bool mainPass) {
    const int r = mainPass ? 3 : 2;
    const float d = m[r] * center[0] + m[4 + r] * center[1] + m[8 + r] * center[2] + m[12 + r];
    const float ext = radius * sqrtf(m[r] * m[r] + m[4 + r] * m[4 + r] + m[8 + r] * m[8 + r]);
    const float key = d - ext;
    return key == key ? key : 0.0f;
}
// strict total order on (key, draw): no NaN keys (scene_order_key), ties by submission order
VKR_HD bool scene_key_less(float ka, uint32_t a, float kb, uint32_t b) { return ka < kb || (ka == kb && a < b); }

// ---- processing order of many draws: buckets.  Counting every draw in front of one's own (scene_rank) is n^2 compares per
// pass — 1 M for 1000 draws, 256 M for 16,000.  Above SCENE_RANK_MAX draws the order is instead: SCENE_BUCKETS equal slices of
// the pass's key range, front to back, draws of one slice in submission order (a stable counting sort on the slice number).
static const uint32_t SCENE_RANK_MAX = 2048u, SCENE_BUCKETS = 2048u, SCENE_BLOCK = 1024u;
// monotone map float -> uint32 (so that atomicMax finds the range of the keys)
VKR_HD uint32_t scene_ordered(float f) { const uint32_t b = f2bits(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
VKR_HD float scene_unordered(uint32_t u) { return bits2f((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }
// range[0] = max of ordered(key), range[1] = max of ~ordered(key) over the pass's visible draws (both start at 0: no draw)
VKR_HD uint32_t scene_bucket(float key, uint32_t rangeMax, uint32_t rangeNegMin) {
    if (key == bits2f(0x7F800000u) || rangeMax == 0u) return SCENE_BUCKETS - 1u;          // hidden draws last
    const float kmin = scene_unordered(~rangeNegMin), kmax = scene_unordered(rangeMax);
    const float span = kmax - kmin;
    if (!(span > 0.0f)) return 0u;
    const float x = (key - kmin) * ((float)(SCENE_BUCKETS - 1u) / span);                   // monotone in key
    const uint32_t b = x > 0.0f ? (uint32_t)x : 0u;
    return b < SCENE_BUCKETS - 2u ? b : SCENE_BUCKETS - 2u;
}

// One thread per draw: the frame's DObject from the static template, the light MVPs, the keys of every pass, and what the
// scans need (triangles if visible).  keys: [pass][nDraws]; shadowMvp: [layer][nDraws][16].
VKR_HD void scene_draw_body(const SceneFrame& f, const SceneDraw& sd, const float* __restrict__ nodeWorld, const DObject& tmpl, uint32_t i,
                            DObject& out, float* __restrict__ shadowMvp, float* __restrict__ keys, uint32_t* __restrict__ visTris) {
    float world[16];
    for (int k = 0; k < 16; ++k) world[k] = nodeWorld[(size_t)sd.node * 16 + k];
    const bool visible = scene_is_visible(world, sd.origin, sd.sphereRadius, f.planes);
    out = tmpl;
    scene_mat_mul(f.viewproj, world, out.mvp);
    for (int k = 0; k < 16; ++k) out.world[k] = world[k];
    out.flags = tmpl.flags | (visible ? 0u : DRAW_HIDDEN);
    const float inf = bits2f(0x7F800000u);
    keys[i] = visible ? scene_order_key(out.mvp, sd.center, sd.radius, true) : inf;
    for (uint32_t l = 0; l < f.shadowLayers; ++l) {
        float* sm = shadowMvp + ((size_t)l * f.nDraws + i) * 16;
        float m[16];
        scene_mat_mul(f.lightViewproj[l], world, m);
        for (int k = 0; k < 16; ++k) sm[k] = m[k];
        keys[(size_t)(1 + l) * f.nDraws + i] = visible ? scene_order_key(m, sd.center, sd.radius, false) : inf;
    }
    visTris[i] = visible ? tmpl.triCount : 0u;
}

// One thread per (pass, draw taking part): its place in the pass's processing order, by counting the draws in front of it.
VKR_HD uint32_t scene_rank(const float* __restrict__ keys, uint32_t count, uint32_t i) {
    const float ki = keys[i];
    uint32_t rank = 0;
    for (uint32_t j = 0; j < count; ++j) rank += scene_key_less(keys[j], j, ki, i) ? 1u : 0u;
    return rank;
}

}  // namespace vkr
