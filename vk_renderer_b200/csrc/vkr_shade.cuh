// vkr_shade.cuh — deferred shading of the visibility buffer.
//
// Replaces mesh.vert's varyings + mesh_pbr.frag / mesh.frag (shaders/mesh.vert:14-24,
// shaders/mesh_pbr.frag:185-226, shaders/mesh.frag:124-182), skybox.vert/.frag
// (shaders/skybox.vert:8-17, skybox.frag:9-12), gradient_color.comp (:16-31), tonemap.comp
// (:9-22), the RGBA16F colour-attachment store and the present blit to UNORM8
// (src/vk_images.cpp:30-62) — one pass over the pixels instead of the reference's five.
//
// Floating point here only has to land within 1/255 of the oracle on the 8-bit image, so
// unlike vkr_raster.cuh it is written for speed (FMA contraction allowed).
#pragma once
#include "vkr_internal.h"
#include <math.h>

namespace vkr {

struct f3 { float x, y, z; };
VKR_HD f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
VKR_HD f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
VKR_HD f3 operator*(f3 a, f3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
VKR_HD f3 operator*(f3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
VKR_HD f3 operator*(float s, f3 a) { return {a.x * s, a.y * s, a.z * s}; }
VKR_HD float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
VKR_HD f3 norm3(f3 a);
VKR_HD float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }

#if defined(__CUDA_ARCH__)
#define VKR_LDG(p) __ldg(p)
// Colour-only arithmetic (tolerance 1/255 on the 8-bit image): the approximate, flush-to-zero SFU forms, one instruction each —
// the non-ftz intrinsics wrap every MUFU in a denormal range fix-up.
VKR_HD float frcp(float a) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
VKR_HD float fdiv(float a, float b) { return a * frcp(b); }
VKR_HD float frsqrt(float a) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
VKR_HD float flog2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
VKR_HD float fexp2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
VKR_HD float fpow(float x, float e) { return fexp2(e * flog2(x)); }      // x >= 0; pow(0, e > 0) = ex2(-inf) = 0
#else
#define VKR_LDG(p) (*(p))
VKR_HD float fdiv(float a, float b) { return a / b; }
VKR_HD float frcp(float a) { return 1.0f / a; }
VKR_HD float frsqrt(float a) { return 1.0f / sqrtf(a); }
VKR_HD float fpow(float x, float e) { return powf(x, e); }
VKR_HD float flog2(float x) { return log2f(x); }
#endif

VKR_HD uint16_t f2h_bits(float f) {
#if defined(__CUDA_ARCH__)
    return __half_as_ushort(__float2half_rn(f));
#else
    uint32_t x; __builtin_memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u, absx = x & 0x7FFFFFFFu;
    if (absx > 0x7F800000u) return (uint16_t)(sign | 0x7E00u);
    if (absx >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);
    if (absx < 0x38800000u) {
        if (absx < 0x33000000u) return (uint16_t)sign;
        uint32_t e = absx >> 23, mant = (absx & 0x7FFFFFu) | 0x800000u, shift = 126 - e;
        uint32_t h = mant >> shift, rem = mant & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) h++;
        return (uint16_t)(sign | h);
    }
    uint32_t h = (absx - 0x38000000u) >> 13, rem = absx & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) h++;
    return (uint16_t)(sign | h);
#endif
}
VKR_HD float h2f_bits(uint16_t h) {
#if defined(__CUDA_ARCH__)
    return __half2float(__ushort_as_half(h));
#else
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu, x;
    if (e == 0) {
        if (m == 0) x = sign;
        else { int s = 0; while (!(m & 0x400u)) { m <<= 1; ++s; } m &= 0x3FFu; x = sign | ((uint32_t)(113 - s) << 23) | (m << 13); }
    } else if (e == 31) x = sign | 0x7F800000u | (m << 13);
    else x = sign | ((e + 112) << 23) | (m << 13);
    float f; __builtin_memcpy(&f, &x, 4); return f;
#endif
}
VKR_HD float round_f16(float f) { return h2f_bits(f2h_bits(f)); }
// Presentation only (colour, +-1 LSB tolerance): saturate + round-to-nearest-even; differs from unorm8() below on exact .5 ties.
// Texture creation (mip chain; its alpha decides cut-out coverage) must use unorm8().
VKR_HD uint32_t unorm8_fast(float x) {
#if defined(__CUDA_ARCH__)
    return __float2uint_rn(__saturatef(x) * 255.0f);     // NaN -> 0
#else
    if (!(x > 0.0f)) return 0u;
    if (x >= 1.0f) return 255u;
    return (uint32_t)floorf(x * 255.0f + 0.5f);
#endif
}
VKR_HD uint32_t unorm8(float x) {
    if (!(x > 0.0f)) return 0u;
    if (x >= 1.0f) return 255u;
    return (uint32_t)floorf(x * 255.0f + 0.5f);
}

VKR_HD f3 norm3(f3 a) { return a * frsqrt(dot3(a, a)); }

// ---- texture units in software ------------------------------------------------------
VKR_HD float4 fetch_texel(const DTexture& t, const float* __restrict__ lut, int level, int x, int y, int w) {
    uchar4 p = VKR_LDG(t.data + t.mipOffset[level] + (size_t)y * w + x);
    if (t.flags & 1u) return make_float4(lut[p.x], lut[p.y], lut[p.z], p.w * (1.0f / 255.0f));   // lut: shared memory on the device
    return make_float4(p.x * (1.0f / 255.0f), p.y * (1.0f / 255.0f), p.z * (1.0f / 255.0f), p.w * (1.0f / 255.0f));
}
VKR_HD int wrap_repeat(int i, int n) {
    if ((n & (n - 1)) == 0) return i & (n - 1);       // power-of-two extent (the usual case): two's-complement AND is the floor-mod
    int r = i % n; return r < 0 ? r + n : r;
}
VKR_HD int level_dim(uint32_t d, int level) { int v = (int)(d >> level); return v < 1 ? 1 : v; }

VKR_HD float4 sample_level(const DTexture& t, const float* lut, int level, float u, float v) {
    const int w = level_dim(t.w, level), h = level_dim(t.h, level);
    if (!(t.flags & 2u)) {
        int x = wrap_repeat((int)floorf(u * (float)w), w), y = wrap_repeat((int)floorf(v * (float)h), h);
        return fetch_texel(t, lut, level, x, y, w);
    }
    float fu = u * (float)w - 0.5f, fv = v * (float)h - 0.5f;
    float iu = floorf(fu), iv = floorf(fv);
    float a = fu - iu, b = fv - iv;
    int x0 = wrap_repeat((int)iu, w), x1 = wrap_repeat((int)iu + 1, w);
    int y0 = wrap_repeat((int)iv, h), y1 = wrap_repeat((int)iv + 1, h);
    float4 t00 = fetch_texel(t, lut, level, x0, y0, w), t10 = fetch_texel(t, lut, level, x1, y0, w);
    float4 t01 = fetch_texel(t, lut, level, x0, y1, w), t11 = fetch_texel(t, lut, level, x1, y1, w);
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return make_float4(t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11,
                       t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11,
                       t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11,
                       t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11);
}

// lod from the SQUARE of the footprint's longer axis: log2(sqrt(x)) = log2(x) / 2, no square roots
VKR_HD float lod_from_rho2(float rho2, int levels) {
    float lambda = rho2 > 0.0f ? 0.5f * flog2(rho2) : -126.0f;
    return fminf(fmaxf(lambda, 0.0f), (float)(levels - 1));
}

VKR_HD float4 texture2d(const ShadeBuffers& sb, int id, float u, float v, float dudx, float dvdx, float dudy, float dvdy) {
    if (id < 0 || id >= VKR_MAX_BINDLESS) return make_float4(0, 0, 0, 1);
    const DTexture& t = sb.textures[id];
    if (t.data == nullptr) return make_float4(0, 0, 0, 1);
    // 1x1 textures — the engine's default white/grey/black/normal images (src/vk_engine_init.cpp:320-333), i.e. every
    // untextured material: any filter returns that texel
    if ((t.w | t.h) == 1u) return fetch_texel(t, sb.srgbLut, 0, 0, 0, 1);
    const int levels = (int)t.levels;
    if (levels == 1) return sample_level(t, sb.srgbLut, 0, u, v);
    float w = (float)t.w, h = (float)t.h;
    float rx2 = (dudx * w) * (dudx * w) + (dvdx * h) * (dvdx * h);
    float ry2 = (dudy * w) * (dudy * w) + (dvdy * h) * (dvdy * h);
    float lod = lod_from_rho2(fmaxf(rx2, ry2), levels);
    int l0 = (int)floorf(lod);
    int l1 = l0 + 1 < levels ? l0 + 1 : levels - 1;
    float f = lod - (float)l0;
    float4 c0 = sample_level(t, sb.srgbLut, l0, u, v);
    if (f == 0.0f) return c0;
    float4 c1 = sample_level(t, sb.srgbLut, l1, u, v);
    return make_float4(c0.x * (1.0f - f) + c1.x * f, c0.y * (1.0f - f) + c1.y * f, c0.z * (1.0f - f) + c1.z * f,
                       c0.w * (1.0f - f) + c1.w * f);
}

// Host side: the uv-independent texture results of a material (DMaterial::albedoTexel / mrTexel), with fetch_texel's arithmetic.
// `textures` are the VKR_MAX_BINDLESS slots as the device will see them, `lut` the 256-entry sRGB decode table.
inline void material_constants(DMaterial& m, const DTexture* textures, const float* lut) {
    m.constFlags = 0u;
    const int ids[2] = {m.albedoID, m.metalRoughID};
    float* dst[2] = {m.albedoTexel, m.mrTexel};
    for (int k = 0; k < 2; ++k) {
        const bool unbound = ids[k] < 0 || ids[k] >= VKR_MAX_BINDLESS || textures[ids[k]].data == nullptr;
        if (unbound) { dst[k][0] = dst[k][1] = dst[k][2] = 0.0f; dst[k][3] = 1.0f; m.constFlags |= 1u << k; continue; }     // texture2d: (0,0,0,1)
        const DTexture& t = textures[ids[k]];
        if ((t.w | t.h) != 1u) continue;
        const uint32_t p = t.texel0;
        const uint32_t c[4] = {p & 255u, (p >> 8) & 255u, (p >> 16) & 255u, p >> 24};
        for (int i = 0; i < 3; ++i) dst[k][i] = (t.flags & 1u) ? lut[c[i]] : (float)c[i] * (1.0f / 255.0f);
        dst[k][3] = (float)c[3] * (1.0f / 255.0f);
        m.constFlags |= 1u << k;
    }
}

// _shadowSampler: LINEAR, CLAMP_TO_BORDER, border 1.0, no compare (src/vk_engine_init.cpp:392-394)
VKR_HD float shadow_fetch(const float* __restrict__ sm, int S, int x, int y) {
    if ((unsigned)x >= (unsigned)S || (unsigned)y >= (unsigned)S) return 1.0f;
    return VKR_LDG(sm + (size_t)y * S + x);
}
VKR_HD float texture_shadow(const float* __restrict__ sm, int S, float u, float v) {
    float fu = u * (float)S - 0.5f, fv = v * (float)S - 0.5f;
    float iu = floorf(fu), iv = floorf(fv);
    float a = fu - iu, b = fv - iv;
    iu = fminf(fmaxf(iu, -2.0f), (float)S + 1.0f);
    iv = fminf(fmaxf(iv, -2.0f), (float)S + 1.0f);
    int x0 = (int)iu, y0 = (int)iv;
    float t00 = shadow_fetch(sm, S, x0, y0), t10 = shadow_fetch(sm, S, x0 + 1, y0);
    float t01 = shadow_fetch(sm, S, x0, y0 + 1), t11 = shadow_fetch(sm, S, x0 + 1, y0 + 1);
    return t00 * ((1.0f - a) * (1.0f - b)) + t10 * (a * (1.0f - b)) + t01 * ((1.0f - a) * b) + t11 * (a * b);
}

// ---- shadow modes: shaders/mesh_pbr.frag:39-156 ------------------------------------------
#define VKR_SHADOW_BIAS 0.0005f
#define VKR_NEAR_PLANE 0.1f
#define VKR_LIGHT_SIZE_UV (2.0f / 200.0f)

// poissonDisk — shaders/mesh_pbr.frag:68-85
#define VKR_POISSON_TABLE                                                                          \
    {{-0.94201624f, -0.39906216f}, {0.94558609f, -0.76890725f}, {-0.094184101f, -0.92938870f},      \
     {0.34495938f, 0.29387760f},   {-0.91588581f, 0.45771432f}, {-0.81544232f, -0.87912464f},       \
     {-0.38277543f, 0.27676845f},  {0.97484398f, 0.75648379f},  {0.44323325f, -0.97511554f},        \
     {0.53742981f, -0.47373420f},  {-0.26496911f, -0.41893023f}, {0.79197514f, 0.19090188f},        \
     {-0.24188840f, 0.99706507f},  {-0.81409955f, 0.91437590f}, {0.19984126f, 0.78641367f},         \
     {0.14383161f, -0.14100790f}}

VKR_HD float shadow_pcss(const float* sm, int S, float su, float sv, float sz) {
    const float c_poisson[16][2] = VKR_POISSON_TABLE;
    float blockerSum = 0.0f;
    int numBlockers = 0;
    float searchWidth = VKR_LIGHT_SIZE_UV * (sz - VKR_NEAR_PLANE) / sz;
    for (int i = 0; i < 16; ++i) {
        float z = texture_shadow(sm, S, su + c_poisson[i][0] * searchWidth, sv + c_poisson[i][1] * searchWidth);
        if (z + VKR_SHADOW_BIAS < sz) { blockerSum += z; numBlockers++; }
    }
    if (numBlockers == 0) return 0.0f;
    float zBlocker = blockerSum / (float)numBlockers;
    if (zBlocker < 0.0f) return 0.0f;
    float penumbraRatio = (sz - zBlocker) / zBlocker;
    float r = penumbraRatio * VKR_LIGHT_SIZE_UV * VKR_NEAR_PLANE / sz;
    float sum = 0.0f;
    for (int i = 0; i < 16; ++i) {
        float pd = texture_shadow(sm, S, su + c_poisson[i][0] * r, sv + c_poisson[i][1] * r);
        sum += (sz - VKR_SHADOW_BIAS > pd) ? 1.0f : 0.0f;
    }
    return sum / 16.0f;
}

// PCF (shaders/mesh_pbr.frag:48-59): nine bilinear taps one texel apart share a 4x4 texel block and
// the same fractional weights, so fetch 16 texels once and filter separably (12 + 9 lerps).
VKR_HD float pcf3x3(const float* __restrict__ sm, int S, float su, float sv, float sz) {
    const float fu = su * (float)S - 0.5f, fv = sv * (float)S - 0.5f;
    float iu = floorf(fu), iv = floorf(fv);
    const float a = fu - iu, b = fv - iv;
    iu = fminf(fmaxf(iu, -4.0f), (float)S + 3.0f);
    iv = fminf(fmaxf(iv, -4.0f), (float)S + 3.0f);
    const int x0 = (int)iu - 1, y0 = (int)iv - 1;
    float h[4][3];
    const bool interior = x0 >= 0 && y0 >= 0 && x0 + 3 < S && y0 + 3 < S;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float t0, t1, t2, t3;
        if (interior) {
            const float* row = sm + (size_t)(y0 + r) * S + x0;
            t0 = VKR_LDG(row); t1 = VKR_LDG(row + 1); t2 = VKR_LDG(row + 2); t3 = VKR_LDG(row + 3);
        } else {
            t0 = shadow_fetch(sm, S, x0, y0 + r); t1 = shadow_fetch(sm, S, x0 + 1, y0 + r);
            t2 = shadow_fetch(sm, S, x0 + 2, y0 + r); t3 = shadow_fetch(sm, S, x0 + 3, y0 + r);
        }
        h[r][0] = t0 + a * (t1 - t0); h[r][1] = t1 + a * (t2 - t1); h[r][2] = t2 + a * (t3 - t2);
    }
    const float ref = sz - VKR_SHADOW_BIAS;
    float shadow = 0.0f;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float d = h[r][c] + b * (h[r + 1][c] - h[r][c]);
            shadow += (ref > d) ? 1.0f : 0.0f;
        }
    return shadow * (1.0f / 9.0f);
}

VKR_HD float calc_shadow(const ShadeConst& sc, const ShadeBuffers& sb, f3 pw, float viewZ) {
    if (!sc.enableShadows) return 0.0f;
    int layer = 0;
    if (sc.shadowMode >= 3) {
        float viewDepth = fabsf(viewZ);
        layer = VKR_NUM_CASCADES - 1;
        for (int i = 0; i < VKR_NUM_CASCADES; ++i)
            if (viewDepth < sc.scene.cascadeDistances[i]) { layer = i; break; }
    }
    const float* m = sc.shadowMat[layer];
    // glm association (m0*x + m1*y) + (m2*z + m3*w), w = 1
    float su = (m[0] * pw.x + m[4] * pw.y) + (m[8] * pw.z + m[12]);
    float sv = (m[1] * pw.x + m[5] * pw.y) + (m[9] * pw.z + m[13]);
    float sz = (m[2] * pw.x + m[6] * pw.y) + (m[10] * pw.z + m[14]);
    const int S = sc.shadowExtent;
    const float* sm = sb.shadow + (size_t)layer * S * S;
    if (sc.shadowMode == 0) {
        float d = texture_shadow(sm, S, su, sv);
        return (d + VKR_SHADOW_BIAS < sz) ? 1.0f : 0.0f;
    }
    if (sc.shadowMode == 1) return pcf3x3(sm, S, su, sv, sz);
    return shadow_pcss(sm, S, su, sv, sz);
}

// ---- Cook-Torrance: shaders/mesh_pbr.frag:159-183 ---------------------------------------
#define VKR_PI 3.14159265359f
VKR_HD float distribution_ggx(f3 N, f3 H, float roughness) {
    float a = roughness * roughness, a2 = a * a;
    float NdotH = fmaxf(dot3(N, H), 0.0f);
    float denom = NdotH * NdotH * (a2 - 1.0f) + 1.0f;
    return fdiv(a2, VKR_PI * denom * denom);
}
VKR_HD float geometry_schlick_ggx(float NdotV, float roughness) {
    float r = roughness + 1.0f, k = r * r / 8.0f;
    return fdiv(NdotV, NdotV * (1.0f - k) + k);
}
VKR_HD float pow5(float x) { float x2 = x * x; return x2 * x2 * x; }

struct Fragment {
    f3 normal, color, posWorld;
    float u, v, viewZ;
    float dudx, dvdx, dudy, dvdy;
};

// The shaders' `discard` (albedo alpha < 0.5) was already applied in the raster stage (vkr_raster.cuh
// alpha_passes), which owns that decision: a fragment that reached the visibility buffer is shaded.
VKR_HD bool shade_fragment(const ShadeConst& sc, const ShadeBuffers& sb, const DMaterial& m, const Fragment& fr, f3& out, bool alphaTest = false) {
    float4 albedoTex = (m.constFlags & 1u) ? make_float4(m.albedoTexel[0], m.albedoTexel[1], m.albedoTexel[2], m.albedoTexel[3])
                                           : texture2d(sb, m.albedoID, fr.u, fr.v, fr.dudx, fr.dvdx, fr.dudy, fr.dvdy);
    if (alphaTest && albedoTex.w < 0.5f) return false;      // `discard` of a fragment that writes no depth (transparent pass)
    f3 base = f3{m.colorFactors[0], m.colorFactors[1], m.colorFactors[2]} * f3{albedoTex.x, albedoTex.y, albedoTex.z} * fr.color;
    f3 N = norm3(fr.normal);
    f3 L = norm3(f3{sc.lightVec[0], sc.lightVec[1], sc.lightVec[2]});
    f3 sun = {sc.scene.sunlightColor[0], sc.scene.sunlightColor[1], sc.scene.sunlightColor[2]};
    f3 amb = {sc.scene.ambientColor[0], sc.scene.ambientColor[1], sc.scene.ambientColor[2]};
    // A fragment facing away from the light gets no direct light: both shaders multiply the direct term by max(N.L, 0), which
    // is then exactly 0, and what is left is ambient * base, bit for bit (x + 0 = x).  The shadow-map lookup — 16 texels for
    // PCF —, the metallic-roughness fetch and the BRDF are skipped for it.  (N.L <= 0 is false for a NaN normal, which takes
    // the full path like before; the one case this changes is a direct term that is 0 * NaN in the shader — roughness 0 with
    // the half vector exactly on the normal.)
    if (dot3(N, L) <= 0.0f) { out = amb * base; return true; }
    const float shadow = calc_shadow(sc, sb, fr.posWorld, fr.viewZ);
    if (sc.fragShader == VKR_FRAG_MESH_LAMBERT) {          // shaders/mesh.frag:171-181
        f3 diffuse = fmaxf(0.0f, dot3(N, L)) * base * sun;
        out = diffuse * (1.0f - shadow) + amb * base;
        return true;
    }
    f3 V = norm3(f3{sc.camPos[0], sc.camPos[1], sc.camPos[2]} - fr.posWorld);
    f3 H = norm3(V + L);
    float4 mrTex = (m.constFlags & 2u) ? make_float4(m.mrTexel[0], m.mrTexel[1], m.mrTexel[2], m.mrTexel[3])
                                       : texture2d(sb, m.metalRoughID, fr.u, fr.v, fr.dudx, fr.dvdx, fr.dudy, fr.dvdy);
    float metallic = m.metalRough[0] * mrTex.z;
    float roughness = m.metalRough[1] * mrTex.y;
    f3 F0 = f3{0.04f, 0.04f, 0.04f} * (1.0f - metallic) + base * metallic;
    float NDF = distribution_ggx(N, H, roughness);
    float NdotV = fmaxf(dot3(N, V), 0.0f), NdotL = fmaxf(dot3(N, L), 0.0f);
    float G = geometry_schlick_ggx(NdotV, roughness) * geometry_schlick_ggx(NdotL, roughness);
    float p5 = pow5(clamp01(1.0f - fmaxf(dot3(H, V), 0.0f)));
    f3 F = F0 + (f3{1.0f, 1.0f, 1.0f} - F0) * p5;
    f3 kD = (f3{1.0f, 1.0f, 1.0f} - F) * (1.0f - metallic);
    float denominator = 4.0f * NdotV * NdotL + 0.0001f;
    f3 specular = (NDF * G * F) * frcp(denominator);
    f3 Lo = (kD * base * (1.0f / VKR_PI) + specular) * NdotL * sun;
    out = amb * base + Lo * (1.0f - shadow);
    return true;
}

// ---- cube map (Vulkan face selection table) -------------------------------------------
VKR_HD void cube_project(f3 r, int face, float& s, float& t) {
    float sc, tc, ma;
    switch (face) {
        case 0: sc = -r.z; tc = -r.y; ma = fabsf(r.x); break;
        case 1: sc = r.z; tc = -r.y; ma = fabsf(r.x); break;
        case 2: sc = r.x; tc = r.z; ma = fabsf(r.y); break;
        case 3: sc = r.x; tc = -r.z; ma = fabsf(r.y); break;
        case 4: sc = r.x; tc = -r.y; ma = fabsf(r.z); break;
        default: sc = -r.x; tc = -r.y; ma = fabsf(r.z); break;
    }
    const float rma = frcp(ma);
    s = 0.5f * (sc * rma + 1.0f);
    t = 0.5f * (tc * rma + 1.0f);
}
VKR_HD int cube_select(f3 r) {
    float ax = fabsf(r.x), ay = fabsf(r.y), az = fabsf(r.z);
    if (az >= ax && az >= ay) return r.z >= 0 ? 4 : 5;
    if (ay >= ax) return r.y >= 0 ? 2 : 3;
    return r.x >= 0 ? 0 : 1;
}
VKR_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
// A cube-map texel: RGBA8 (UNORM / sRGB) or RGBA16F (the engine's skybox asset, assets/pisa_cube.ktx, is an HDR image).
// The format is a template parameter chosen once per lookup (texture_cube): as a per-texel test the compiler predicates
// both fetch paths into every texel, +8 % instructions for the whole shading kernel (profiles/r01z).
// FMT: 0 RGBA8 (UNORM / sRGB), 1 RGBA16F, 2 RGBA32F (8-bit texels decoded at upload: DTexture.flags bit 4, device only)
template <int FMT>
VKR_HD float4 fetch_cube_texel(const DTexture& t, const float* __restrict__ lut, int level, int x, int y, int w) {
    if (FMT == 2) return VKR_LDG(reinterpret_cast<const float4*>(t.data) + t.mipOffset[level] + (size_t)y * w + x);
    if (FMT == 1) {
        const uint2 p = VKR_LDG(reinterpret_cast<const uint2*>(t.data) + t.mipOffset[level] + (size_t)y * w + x);
        return make_float4(h2f_bits((uint16_t)(p.x & 0xFFFFu)), h2f_bits((uint16_t)(p.x >> 16)),
                           h2f_bits((uint16_t)(p.y & 0xFFFFu)), h2f_bits((uint16_t)(p.y >> 16)));
    }
    return fetch_texel(t, lut, level, x, y, w);
}
// A bilinear tap ONE texel beyond an edge of `face` (Vulkan filters cube maps seamlessly): the texel of the face across that
// edge.  The two faces unfolded flat about the edge, the tap lies as far inside the neighbour as it lay outside this face:
// with the face written as  major + u * sAxis + v * tAxis  (u = 2s - 1, v = 2t - 1; the selection table read backwards),
// the tap at |u| = 1 + d is the direction  (1 - d) * major + sign(u) * sAxis + v * tAxis, which is then selected and projected
// like any other.  False: the tap is beyond two edges — a cube corner, where no texel exists.
template <int FMT>
VKR_HD bool cube_edge_tap(const DTexture* cube, const float* lut, int face, int level, int x, int y, int w, int h, float4& out) {
    const bool xo = (unsigned)x >= (unsigned)w, yo = (unsigned)y >= (unsigned)h;
    if (xo && yo) return false;
    const float u = (float)(2 * x + 1) * frcp((float)w) - 1.0f, v = (float)(2 * y + 1) * frcp((float)h) - 1.0f;
    const float along = xo ? v : u, beyond = xo ? u : v;
    const float sg = beyond > 0.0f ? 1.0f : -1.0f, depth = 2.0f - fabsf(beyond);          // 1 - d
    // (major, sAxis, tAxis) per face as signed unit axes: +X (x, -z, -y), -X (-x, z, -y), +Y (y, x, z), -Y (-y, x, -z), +Z (z, x, -y), -Z (-z, -x, -y)
    const float a0 = depth, a1 = xo ? sg : along, a2 = xo ? along : sg;                      // coefficients of major, sAxis, tAxis
    f3 r;
    switch (face) {
        case 0: r = {a0, -a2, -a1}; break;
        case 1: r = {-a0, -a2, a1}; break;
        case 2: r = {a1, a0, a2}; break;
        case 3: r = {a1, -a0, -a2}; break;
        case 4: r = {a1, -a2, a0}; break;
        default: r = {-a1, -a2, -a0}; break;
    }
    const int nf = cube_select(r);
    float ns, nt;
    cube_project(r, nf, ns, nt);
    const DTexture& t = cube[nf];
    const int nw = level_dim(t.w, level), nh = level_dim(t.h, level);
    out = fetch_cube_texel<FMT>(t, lut, level, clampi((int)floorf(ns * (float)nw), 0, nw - 1), clampi((int)floorf(nt * (float)nh), 0, nh - 1), nw);
    return true;
}
template <int FMT>
VKR_HD float4 sample_face_level(const DTexture* cube, int face, const float* lut, int level, float s, float tt) {
    const DTexture& t = cube[face];
    const int w = level_dim(t.w, level), h = level_dim(t.h, level);
    float fu = s * (float)w - 0.5f, fv = tt * (float)h - 0.5f;
    float iu = floorf(fu), iv = floorf(fv);
    float a = fu - iu, b = fv - iv;
    const int x0 = clampi((int)iu, -1, w - 1), y0 = clampi((int)iv, -1, h - 1);
    float4 t00, t10, t01, t11;
    if (x0 >= 0 && y0 >= 0 && x0 + 1 < w && y0 + 1 < h) {          // the whole footprint on this face: every pixel but those on a seam
        t00 = fetch_cube_texel<FMT>(t, lut, level, x0, y0, w); t10 = fetch_cube_texel<FMT>(t, lut, level, x0 + 1, y0, w);
        t01 = fetch_cube_texel<FMT>(t, lut, level, x0, y0 + 1, w); t11 = fetch_cube_texel<FMT>(t, lut, level, x0 + 1, y0 + 1, w);
    } else {
        float4 tap[4];
        bool have[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = x0 + (k & 1), y = y0 + (k >> 1);
            if ((unsigned)x < (unsigned)w && (unsigned)y < (unsigned)h) { tap[k] = fetch_cube_texel<FMT>(t, lut, level, x, y, w); have[k] = true; }
            else have[k] = cube_edge_tap<FMT>(cube, lut, face, level, x, y, w, h, tap[k]);
        }
        // a footprint at a cube corner lacks one tap: the average of the other three (Vulkan "Cube Map Corner Handling")
        float4 sum = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float n = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (have[k]) { sum.x += tap[k].x; sum.y += tap[k].y; sum.z += tap[k].z; sum.w += tap[k].w; n += 1.0f; }
        const float inv = n > 0.0f ? 1.0f / n : 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (!have[k]) tap[k] = make_float4(sum.x * inv, sum.y * inv, sum.z * inv, sum.w * inv);
        t00 = tap[0]; t10 = tap[1]; t01 = tap[2]; t11 = tap[3];
    }
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return make_float4(t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11,
                       t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11,
                       t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11,
                       t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11);
}
template <int FMT>
VKR_HD float4 texture_cube_fmt(const ShadeBuffers& sb, f3 r, f3 rdx, f3 rdy, int face, float s, float t) {
    const DTexture& tex = sb.cube[face];
    const int levels = (int)tex.levels;
    if (levels == 1) return sample_face_level<FMT>(sb.cube, face, sb.srgbLut, 0, s, t);
    float sx, tx, sy, ty;
    cube_project(rdx, face, sx, tx);
    cube_project(rdy, face, sy, ty);
    float w = (float)tex.w, h = (float)tex.h;
    float ax = (sx - s) * w, bx = (tx - t) * h, ay = (sy - s) * w, by = (ty - t) * h;
    float lod = lod_from_rho2(fmaxf(ax * ax + bx * bx, ay * ay + by * by), levels);
    int l0 = (int)floorf(lod);
    int l1 = l0 + 1 < levels ? l0 + 1 : levels - 1;
    float f = lod - (float)l0;
    float4 c0 = sample_face_level<FMT>(sb.cube, face, sb.srgbLut, l0, s, t);
    if (f == 0.0f) return c0;
    float4 c1 = sample_face_level<FMT>(sb.cube, face, sb.srgbLut, l1, s, t);
    return make_float4(c0.x * (1.0f - f) + c1.x * f, c0.y * (1.0f - f) + c1.y * f, c0.z * (1.0f - f) + c1.z * f,
                       c0.w * (1.0f - f) + c1.w * f);
}
VKR_HD float4 texture_cube(const ShadeBuffers& sb, f3 r, f3 rdx, f3 rdy) {
    int face = cube_select(r);
    float s, t;
    cube_project(r, face, s, t);
    const uint32_t fl = sb.cube[face].flags;                         // all six faces share the format
    if (fl & 16u) return texture_cube_fmt<2>(sb, r, rdx, rdy, face, s, t);
    if (fl & 8u) return texture_cube_fmt<1>(sb, r, rdx, rdy, face, s, t);
    return texture_cube_fmt<0>(sb, r, rdx, rdy, face, s, t);
}
VKR_HD f3 sky_dir(const ShadeConst& sc, float sx, float sy) {
    return {sc.skyO[0] + sx * sc.skyX[0] + sy * sc.skyY[0], sc.skyO[1] + sx * sc.skyX[1] + sy * sc.skyY[1],
            sc.skyO[2] + sx * sc.skyX[2] + sy * sc.skyY[2]};
}

// ---- perspective-correct barycentrics of the original triangle (2-D homogeneous) --------
VKR_HD void bary_at(const float4 c[3], float nx, float ny, float b[3]) {
    float q0x = c[0].x - nx * c[0].w, q0y = c[0].y - ny * c[0].w;
    float q1x = c[1].x - nx * c[1].w, q1y = c[1].y - ny * c[1].w;
    float q2x = c[2].x - nx * c[2].w, q2y = c[2].y - ny * c[2].w;
    float e0 = q1x * q2y - q2x * q1y;
    float e1 = q2x * q0y - q0x * q2y;
    float e2 = q0x * q1y - q1x * q0y;
    float s = e0 + e1 + e2;
    if (s == 0.0f) { b[0] = 1.0f; b[1] = 0.0f; b[2] = 0.0f; return; }
    float inv = frcp(s);
    b[0] = e0 * inv; b[1] = e1 * inv; b[2] = e2 * inv;
}

VKR_HD float4 xform(const float* m, float4 p) {
    return make_float4((m[0] * p.x + m[4] * p.y) + (m[8] * p.z + m[12] * p.w), (m[1] * p.x + m[5] * p.y) + (m[9] * p.z + m[13] * p.w),
                       (m[2] * p.x + m[6] * p.y) + (m[10] * p.z + m[14] * p.w), (m[3] * p.x + m[7] * p.y) + (m[11] * p.z + m[15] * p.w));
}

// Rebuild the fragment of primitive `seq` at pixel (px,py) from the SoA vertex streams.
VKR_HD void fetch_fragment(const ShadeConst& sc, const ShadeBuffers& sb, uint32_t seq, int px, int py,
                           Fragment& fr, uint32_t& material) {
    // cluster slots: seq >> 8 -> (draw, first triangle of the slot); seq is in submission order
    const uint2 ch = VKR_LDG(sb.chunkObj + (seq >> 8));
    const DObject& o = sb.objects[ch.x];
    const uint32_t tri = ch.y + (seq & 255u);
    material = o.material;
    const uint32_t i0 = VKR_LDG(o.idx + 3 * tri), i1 = VKR_LDG(o.idx + 3 * tri + 1), i2 = VKR_LDG(o.idx + 3 * tri + 2);
    const float4 p0 = VKR_LDG(o.pos + i0), p1 = VKR_LDG(o.pos + i1), p2 = VKR_LDG(o.pos + i2);
    float4 c[3] = {xform(o.mvp, p0), xform(o.mvp, p1), xform(o.mvp, p2)};
    const float cx = (float)px + 0.5f, cy = (float)py + 0.5f;
    const float ihw = sc.invHalfW, ihh = sc.invHalfH;
    float b[3], bx[3], by[3];
    const float ox = (px & 1) ? -1.0f : 1.0f, oy = (py & 1) ? -1.0f : 1.0f;
    bary_at(c, (cx - sc.halfW) * ihw, (cy - sc.halfH) * ihh, b);
    // uv derivatives (the quad neighbours' barycentrics) only for draws with a mip-mapped texture: plan_frame sets bit 2 of
    // the draw's flags, which sit in the record this function reads anyway
#if defined(__CUDA_ARCH__)
    // (warp-uniform, so that the compiler branches around the two evaluations instead of predicating ~80 instructions)
    const bool derivs = __any_sync(__activemask(), (o.flags & 4u) != 0u);
#else
    const bool derivs = (o.flags & 4u) != 0u;
#endif
    if (derivs) {
        bary_at(c, (cx + ox - sc.halfW) * ihw, (cy - sc.halfH) * ihh, bx);
        bary_at(c, (cx - sc.halfW) * ihw, (cy + oy - sc.halfH) * ihh, by);
    }
    const float4 a0 = VKR_LDG(o.attrA + i0), a1 = VKR_LDG(o.attrA + i1), a2 = VKR_LDG(o.attrA + i2);
    const float4 c0 = VKR_LDG(o.attrB + i0), c1 = VKR_LDG(o.attrB + i1), c2 = VKR_LDG(o.attrB + i2);
    f3 nobj = {b[0] * a0.x + b[1] * a1.x + b[2] * a2.x, b[0] * a0.y + b[1] * a1.y + b[2] * a2.y, b[0] * a0.z + b[1] * a1.z + b[2] * a2.z};
    const float* wm = o.world;
    fr.normal = {wm[0] * nobj.x + wm[4] * nobj.y + wm[8] * nobj.z, wm[1] * nobj.x + wm[5] * nobj.y + wm[9] * nobj.z,
                 wm[2] * nobj.x + wm[6] * nobj.y + wm[10] * nobj.z};
    fr.color = {b[0] * c0.x + b[1] * c1.x + b[2] * c2.x, b[0] * c0.y + b[1] * c1.y + b[2] * c2.y, b[0] * c0.z + b[1] * c1.z + b[2] * c2.z};
    fr.u = b[0] * a0.w + b[1] * a1.w + b[2] * a2.w;
    fr.v = b[0] * c0.w + b[1] * c1.w + b[2] * c2.w;
    fr.dudx = fr.dvdx = fr.dudy = fr.dvdy = 0.0f;
    if (derivs) {
        float ux = bx[0] * a0.w + bx[1] * a1.w + bx[2] * a2.w, vx = bx[0] * c0.w + bx[1] * c1.w + bx[2] * c2.w;
        float uy = by[0] * a0.w + by[1] * a1.w + by[2] * a2.w, vy = by[0] * c0.w + by[1] * c1.w + by[2] * c2.w;
        fr.dudx = (ux - fr.u) * ox; fr.dvdx = (vx - fr.v) * ox;
        fr.dudy = (uy - fr.u) * oy; fr.dvdy = (vy - fr.v) * oy;
    }
    float4 pobj = make_float4(b[0] * p0.x + b[1] * p1.x + b[2] * p2.x, b[0] * p0.y + b[1] * p1.y + b[2] * p2.y,
                              b[0] * p0.z + b[1] * p1.z + b[2] * p2.z, 1.0f);
    float4 pw = xform(wm, pobj);
    fr.posWorld = {pw.x, pw.y, pw.z};
    const float* vm = sc.scene.view;
    fr.viewZ = (vm[2] * pw.x + vm[6] * pw.y) + (vm[10] * pw.z + vm[14] * pw.w);
}

// ---- "Transparent" material pass (src/vk_engine_init.cpp:545-547): depth test LESS_OR_EQUAL against the opaque
// depth, no depth write, blend rgb = src.rgb * src.a + dst.rgb, a = src.a (src/vk_pipelines.cpp:108-118).
// The blend is a sum, so fragments are shaded where they are rasterised and added (f32 atomics) into an
// accumulation image that shade_pixel folds into the draw image before tone mapping; the reference rounds the
// destination to f16 after every fragment, this path once — far inside the 1/255 tolerance.
struct BlendSink {
    const ShadeConst* sc; const ShadeBuffers* sb; uint32_t seq;
    VKR_HD void put(int32_t px, int32_t py, float z) const {
        const size_t pix = (size_t)py * sc->width + px;
        const uint32_t stored = (uint32_t)(sb->vis[pix] >> 32);
        if (!(f2bits(z) <= stored)) return;
        // draw_skybox runs after the geometry (src/vk_engine_run.cpp:313) and overwrites every texel still at depth 1.0
        if (sc->enableSkybox && sb->cube != nullptr && stored >= DEPTH_ONE_BITS) return;
        Fragment fr;
        uint32_t material;
        fetch_fragment(*sc, *sb, seq, px, py, fr, material);
        f3 col;
        if (!shade_fragment(*sc, *sb, sb->materials[material], fr, col, true)) return;
        float* acc = sb->accum + pix * 4;     // src alpha is 1.0 (outFragColor.a, shaders/mesh_pbr.frag:225)
#if defined(__CUDA_ARCH__)
        atomicAdd(acc, col.x); atomicAdd(acc + 1, col.y); atomicAdd(acc + 2, col.z); atomicAdd(acc + 3, 1.0f);
#else
        acc[0] += col.x; acc[1] += col.y; acc[2] += col.z; acc[3] += 1.0f;
#endif
    }
};

VKR_HD f3 tonemap_reinhard(f3 c) {      // shaders/tonemap.comp:13-19
    f3 m = {fdiv(c.x, c.x + 1.0f), fdiv(c.y, c.y + 1.0f), fdiv(c.z, c.z + 1.0f)};
    const float e = 1.0f / 2.2f;
    return {fpow(m.x, e), fpow(m.y, e), fpow(m.z, e)};
}

// Resolve one pixel: visibility key -> draw-image texel (RGBA16F) -> presented texel (RGBA8).
VKR_HD void shade_pixel(const ShadeConst& sc, const ShadeBuffers& sb, int px, int py) {
    const size_t pix = (size_t)py * sc.width + px;
    const uint64_t key = sb.vis[pix];
    const uint32_t lo = (uint32_t)key;
    float r, g, b, a = 1.0f;
    bool have = false;
    const bool inScissor = px >= sc.sx0 && px < sc.sx1 && py >= sc.sy0 && py < sc.sy1;
    if (lo != VIS_EMPTY_LO) {
        Fragment fr;
        uint32_t material;
        fetch_fragment(sc, sb, 0xFFFFFFFEu - lo, px, py, fr, material);
        f3 col;
        // alpha-tested fragments were already rejected in the raster stage for cut-out
        // materials; a discard here can only be a texel/LOD rounding disagreement.
        if (shade_fragment(sc, sb, sb.materials[material], fr, col)) { r = col.x; g = col.y; b = col.z; have = true; }
    }
    if (!have) {
        if (sc.enableSkybox && sb.cube != nullptr && inScissor && (uint32_t)(key >> 32) >= DEPTH_ONE_BITS) {
            const float cx = (float)px + 0.5f, cy = (float)py + 0.5f;
            const float ox = (px & 1) ? -1.0f : 1.0f, oy = (py & 1) ? -1.0f : 1.0f;
            const f3 d0 = sky_dir(sc, cx, cy);
            const f3 dx = {d0.x + ox * sc.skyX[0], d0.y + ox * sc.skyX[1], d0.z + ox * sc.skyX[2]};
            const f3 dy = {d0.x + oy * sc.skyY[0], d0.y + oy * sc.skyY[1], d0.z + oy * sc.skyY[2]};
            float4 c = texture_cube(sb, d0, dx, dy);
            r = c.x; g = c.y; b = c.z; a = c.w;
        } else if (sc.enableBackground) {                   // gradient_color.comp:27-29
            float blend = (float)py / (float)sc.height;
            r = sc.bgTop[0] * (1.0f - blend) + sc.bgBottom[0] * blend;
            g = sc.bgTop[1] * (1.0f - blend) + sc.bgBottom[1] * blend;
            b = sc.bgTop[2] * (1.0f - blend) + sc.bgBottom[2] * blend;
            a = sc.bgTop[3] * (1.0f - blend) + sc.bgBottom[3] * blend;
        } else { r = 0.0f; g = 0.0f; b = 0.0f; a = 1.0f; }   // clear (0,0,0,1), src/vk_engine_run.cpp:247
    }
    if (sb.accum != nullptr) {      // "Transparent" draws: additive blend accumulated by BlendSink
        const float4 t = reinterpret_cast<const float4*>(sb.accum)[pix];
        if (t.w > 0.0f) { r += t.x; g += t.y; b += t.z; a = 1.0f; }
    }
    // colour attachment store (RGBA16F), then tonemap.comp reads that image
    r = round_f16(r); g = round_f16(g); b = round_f16(b); a = round_f16(a);
    if (sc.enablePostprocess) {
        f3 m = tonemap_reinhard({r, g, b});
        r = round_f16(m.x); g = round_f16(m.y); b = round_f16(m.z); a = 1.0f;
    }
    if (sc.storeHdr) {
        uint16_t* h = sb.hdr + pix * 4;
#if defined(__CUDA_ARCH__)
        uint2 packed;
        packed.x = (uint32_t)f2h_bits(r) | ((uint32_t)f2h_bits(g) << 16);
        packed.y = (uint32_t)f2h_bits(b) | ((uint32_t)f2h_bits(a) << 16);
        *reinterpret_cast<uint2*>(h) = packed;
#else
        h[0] = f2h_bits(r); h[1] = f2h_bits(g); h[2] = f2h_bits(b); h[3] = f2h_bits(a);
#endif
    }
    const uint32_t texel = unorm8_fast(r) | (unorm8_fast(g) << 8) | (unorm8_fast(b) << 16) | (unorm8_fast(a) << 24);
    sb.rgba8[pix] = texel;
    if (sb.rgba8Peer != nullptr) sb.rgba8Peer[pix] = texel;      // the presenting rank's image, over NVLink
    // a transient attachment: this is the next frame's clear.  Issued with the other stores, long after the key arrived — a
    // store right behind the load of the same address waits for that load (a DRAM round trip) with the warp's later loads
    // queued behind it, which made the kernel 3.6x slower (gpurun_out ab_z2)
    if (sc.clearVis) sb.vis[pix] = VIS_CLEAR;
}

}  // namespace vkr
