// ork.cpp — CPU ORACLE (test infrastructure only; see ork.h).
//
// A plain scalar restatement of the reference's per-frame path: the GLSL shaders are
// restated function by function (citations on each), and the Vulkan fixed-function
// stages they sit in (clip, viewport, snapping, top-left rule, depth test, blending,
// texture filtering, format conversion) are restated from the Vulkan 1.3 rules
// summarised in SURVEY.md Appendix B.  It renders FORWARD, in submission order, like
// the reference (src/vk_engine_run.cpp:140-161), which is deliberately a different
// algorithm from the product's visibility-buffer pipeline.
//
// PARITY UNPINNED for fixed function (no Vulkan driver can run here; the reference
// has no golden images).  Shader math is pinned by tests/golden (see ork.h).
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fPIC -shared (see oracle/Makefile).
// -ffp-contract=off matters: the position path must round exactly as written.

#include "ork.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

namespace {

// ------------------------------------------------------------------ small math
struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct mat4 { float m[4][4]; };  // m[col][row], column-major like glm / GLSL

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 normalize(vec3 a) {
    float inv = 1.0f / std::sqrt(dot(a, a));  // GLSL normalize = v * inversesqrt(dot(v,v))
    return a * inv;
}
inline vec3 V3(float s) { return {s, s, s}; }
inline vec3 xyz(vec4 v) { return {v.x, v.y, v.z}; }
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline float clampf(float x, float lo, float hi) { return std::min(std::max(x, lo), hi); }

inline mat4 load_mat(const float* p) { mat4 r; std::memcpy(r.m, p, 64); return r; }

// glm 0.9.9.7 mat4*vec4 association: (m0*x + m1*y) + (m2*z + m3*w)
// (external/glm/glm/detail/type_mat4x4.inl:561-572).
inline vec4 mul(const mat4& m, vec4 v) {
    vec4 r;
    float* o = &r.x;
    for (int i = 0; i < 4; ++i)
        o[i] = (m.m[0][i] * v.x + m.m[1][i] * v.y) + (m.m[2][i] * v.z + m.m[3][i] * v.w);
    return r;
}
// Position path (spec): clip = M*(p,1) as three nested FMAs per component.
inline vec4 mul_point_fma(const mat4& m, float x, float y, float z) {
    vec4 r;
    float* o = &r.x;
    for (int i = 0; i < 4; ++i) o[i] = std::fmaf(m.m[0][i], x, std::fmaf(m.m[1][i], y, std::fmaf(m.m[2][i], z, m.m[3][i])));
    return r;
}
// glm mat4*mat4: column j = ((A0*b0 + A1*b1) + A2*b2) + A3*b3 (type_mat4x4.inl:606-625).
inline mat4 mul(const mat4& a, const mat4& b) {
    mat4 r;
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i)
            r.m[j][i] = ((a.m[0][i] * b.m[j][0] + a.m[1][i] * b.m[j][1]) + a.m[2][i] * b.m[j][2]) +
                        a.m[3][i] * b.m[j][3];
    return r;
}
inline vec3 mul3(const mat4& m, vec3 v) {  // mat3(m) * v
    return {m.m[0][0] * v.x + m.m[1][0] * v.y + m.m[2][0] * v.z,
            m.m[0][1] * v.x + m.m[1][1] * v.y + m.m[2][1] * v.z,
            m.m[0][2] * v.x + m.m[1][2] * v.y + m.m[2][2] * v.z};
}
// General 4x4 inverse by cofactors (what GLSL inverse() / glm::inverse compute).
mat4 inverse(const mat4& in) {
    const float* m = &in.m[0][0];
    float inv[16];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    float id = 1.0f / det;
    mat4 r;
    for (int i = 0; i < 16; ++i) (&r.m[0][0])[i] = inv[i] * id;
    return r;
}

// ------------------------------------------------------------- format helpers
// f32 -> f16 round-to-nearest-even (colour attachment / image store conversion).
uint16_t f32_to_f16(float f) {
    uint32_t x;
    std::memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t absx = x & 0x7FFFFFFFu;
    if (absx > 0x7F800000u) return (uint16_t)(sign | 0x7E00u);             // NaN
    if (absx >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);            // >= 65520 -> inf
    if (absx < 0x38800000u) {                                              // subnormal half / zero
        if (absx < 0x33000000u) return (uint16_t)sign;                     // < 2^-25 -> 0
        uint32_t e = absx >> 23;
        uint32_t mant = (absx & 0x7FFFFFu) | 0x800000u;
        uint32_t shift = 126 - e;                                          // 14..24
        uint32_t h = mant >> shift;
        uint32_t rem = mant & ((1u << shift) - 1);
        uint32_t half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) h++;
        return (uint16_t)(sign | h);
    }
    uint32_t h = ((absx - 0x38000000u) >> 13);
    uint32_t rem = absx & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) h++;
    return (uint16_t)(sign | h);
}
float f16_to_f32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t e = (h >> 10) & 0x1Fu, m = h & 0x3FFu;
    uint32_t x;
    if (e == 0) {
        if (m == 0) x = sign;
        else {
            int s = 0;
            while (!(m & 0x400u)) { m <<= 1; ++s; }
            m &= 0x3FFu;
            x = sign | ((uint32_t)(113 - s) << 23) | (m << 13);
        }
    } else if (e == 31) x = sign | 0x7F800000u | (m << 13);
    else x = sign | ((e + 112) << 23) | (m << 13);
    float f;
    std::memcpy(&f, &x, 4);
    return f;
}
inline float roundtrip_f16(float f) { return f16_to_f32(f32_to_f16(f)); }

// UNORM8 conversion of the present blit (src/vk_images.cpp:30-62): clamp, *255, round.
inline uint8_t to_unorm8(float x) {
    if (!(x > 0.0f)) return 0;
    if (x >= 1.0f) return 255;
    return (uint8_t)std::floor(x * 255.0f + 0.5f);
}

float g_srgb_lut[256];
bool g_lut_ready = false;
void init_lut() {
    if (g_lut_ready) return;
    for (int i = 0; i < 256; ++i) {
        double c = i / 255.0;
        double l = c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4);
        g_srgb_lut[i] = (float)l;
    }
    g_lut_ready = true;
}
inline uint8_t linear_to_srgb8(float l) {
    if (!(l > 0.0f)) return 0;
    if (l >= 1.0f) return 255;
    float s = l <= 0.0031308f ? l * 12.92f : 1.055f * std::pow(l, 1.0f / 2.4f) - 0.055f;
    return to_unorm8(s);
}

// -------------------------------------------------------------------- objects
struct Image {
    int format = 0;
    std::vector<int> w, h;
    std::vector<std::vector<uint8_t>> mips;
    int levels() const { return (int)mips.size(); }
};
struct Slot { int image = -1; int sampler = 0; };
struct Mesh { std::vector<VkrVertex> v; std::vector<uint32_t> idx; };
struct Material { VkrMaterialConstants c; int pass; };

inline vec4 fetch_texel(const Image& im, int level, int x, int y) {
    if (im.format == VKR_FORMAT_R16G16B16A16_SFLOAT) {          // cube maps loaded from an RGBA16F KTX file
        const uint8_t* q = &im.mips[level][((size_t)y * im.w[level] + x) * 8];
        uint16_t hv[4];
        std::memcpy(hv, q, 8);
        return {f16_to_f32(hv[0]), f16_to_f32(hv[1]), f16_to_f32(hv[2]), f16_to_f32(hv[3])};
    }
    const uint8_t* p = &im.mips[level][((size_t)y * im.w[level] + x) * 4];
    if (im.format == VKR_FORMAT_R8G8B8A8_SRGB)
        return {g_srgb_lut[p[0]], g_srgb_lut[p[1]], g_srgb_lut[p[2]], p[3] / 255.0f};
    return {p[0] / 255.0f, p[1] / 255.0f, p[2] / 255.0f, p[3] / 255.0f};
}

// generate_mipmaps (src/vk_images.cpp:64-157): level i+1 = LINEAR blit of level i to
// max(1, w>>(i+1)) x max(1, h>>(i+1)); for sRGB formats filtering is done on decoded
// values and the result re-encoded.
void build_mips(Image& im) {
    int W = im.w[0], H = im.h[0];
    int levels = (int)std::floor(std::log2((double)std::max(W, H))) + 1;
    for (int i = 0; i < levels - 1; ++i) {
        int sw = std::max(1, W >> i), sh = std::max(1, H >> i);   // blit src extent (>>i, see :94-95)
        // NB: the reference passes (W>>i) unclamped as the src extent; a 0 extent is an empty
        // region in Vulkan. We clamp to 1 (the image's real level size).
        int dw = std::max(1, W >> (i + 1)), dh = std::max(1, H >> (i + 1));
        std::vector<uint8_t> dst((size_t)dw * dh * 4);
        float scx = (float)sw / (float)dw, scy = (float)sh / (float)dh;
        for (int y = 0; y < dh; ++y)
            for (int x = 0; x < dw; ++x) {
                // vkCmdBlitImage LINEAR: src coord of dst texel centre, unnormalised, minus 0.5
                float u = ((float)x + 0.5f) * scx - 0.5f;
                float v = ((float)y + 0.5f) * scy - 0.5f;
                float fu = std::floor(u), fv = std::floor(v);
                float a = u - fu, b = v - fv;
                int x0 = std::clamp((int)fu, 0, sw - 1), x1 = std::clamp((int)fu + 1, 0, sw - 1);
                int y0 = std::clamp((int)fv, 0, sh - 1), y1 = std::clamp((int)fv + 1, 0, sh - 1);
                vec4 t00 = fetch_texel(im, i, x0, y0), t10 = fetch_texel(im, i, x1, y0);
                vec4 t01 = fetch_texel(im, i, x0, y1), t11 = fetch_texel(im, i, x1, y1);
                float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
                float c[4] = {t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11,
                              t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11,
                              t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11,
                              t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11};
                uint8_t* o = &dst[((size_t)y * dw + x) * 4];
                if (im.format == VKR_FORMAT_R8G8B8A8_SRGB) {
                    o[0] = linear_to_srgb8(c[0]); o[1] = linear_to_srgb8(c[1]); o[2] = linear_to_srgb8(c[2]);
                } else {
                    o[0] = to_unorm8(c[0]); o[1] = to_unorm8(c[1]); o[2] = to_unorm8(c[2]);
                }
                o[3] = to_unorm8(c[3]);
            }
        im.w.push_back(dw); im.h.push_back(dh);
        im.mips.push_back(std::move(dst));
    }
}

}  // namespace

struct OrkDevice {
    int W = 0, H = 0, S = 2048;
    int nthreads = 0;
    std::vector<Mesh> meshes;
    std::vector<Image> images;
    std::vector<Material> materials;
    Slot slots[VKR_MAX_BINDLESS];
    Image cube[6];
    bool haveCube = false;
    std::vector<float> depth, shadow;
    std::vector<uint16_t> hdr;      // RGBA16F draw image
    std::vector<uint32_t> primid;
    std::vector<uint8_t> rgba8;
    VkrEngineStats stats{};
    std::string err;
};

namespace {

// ------------------------------------------------------- texture sampling
inline int wrap_repeat(int i, int n) { int r = i % n; return r < 0 ? r + n : r; }

// One level, sampler filter NEAREST or LINEAR, address REPEAT
// (vkinit::sampler_create_info, src/vk_initializers.cpp:228-248; src/vk_engine_init.cpp:343-348).
vec4 sample_level(const Image& im, int level, int sampler, float u, float v) {
    int w = im.w[level], h = im.h[level];
    if (sampler == VKR_SAMPLER_NEAREST) {
        int x = wrap_repeat((int)std::floor(u * (float)w), w);
        int y = wrap_repeat((int)std::floor(v * (float)h), h);
        return fetch_texel(im, level, x, y);
    }
    float fu = u * (float)w - 0.5f, fv = v * (float)h - 0.5f;
    float iu = std::floor(fu), iv = std::floor(fv);
    float a = fu - iu, b = fv - iv;
    int x0 = wrap_repeat((int)iu, w), x1 = wrap_repeat((int)iu + 1, w);
    int y0 = wrap_repeat((int)iv, h), y1 = wrap_repeat((int)iv + 1, h);
    vec4 t00 = fetch_texel(im, level, x0, y0), t10 = fetch_texel(im, level, x1, y0);
    vec4 t01 = fetch_texel(im, level, x0, y1), t11 = fetch_texel(im, level, x1, y1);
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return {t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11,
            t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11,
            t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11,
            t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11};
}

// Implicit-LOD texture(): rho = max(|d(uv*size)/dx|, |d(uv*size)/dy|) (Euclidean lengths),
// lambda = log2(rho), mipmapMode LINEAR between floor(lambda) and +1, LOD clamp [0, levels-1].
vec4 texture2D(const OrkDevice* d, int id, vec2 uv, vec2 ddx, vec2 ddy) {
    if (id < 0 || id >= VKR_MAX_BINDLESS || d->slots[id].image < 0) return {0, 0, 0, 1};
    const Slot& s = d->slots[id];
    const Image& im = d->images[s.image];
    int levels = im.levels();
    if (levels == 1) return sample_level(im, 0, s.sampler, uv.x, uv.y);
    float w = (float)im.w[0], h = (float)im.h[0];
    float rx = std::sqrt((ddx.x * w) * (ddx.x * w) + (ddx.y * h) * (ddx.y * h));
    float ry = std::sqrt((ddy.x * w) * (ddy.x * w) + (ddy.y * h) * (ddy.y * h));
    float rho = std::max(rx, ry);
    float lambda = rho > 0.0f ? std::log2(rho) : -126.0f;
    float lod = clampf(lambda, 0.0f, (float)(levels - 1));
    int l0 = (int)std::floor(lod);
    int l1 = std::min(l0 + 1, levels - 1);
    float f = lod - (float)l0;
    vec4 c0 = sample_level(im, l0, s.sampler, uv.x, uv.y);
    if (f == 0.0f) return c0;
    vec4 c1 = sample_level(im, l1, s.sampler, uv.x, uv.y);
    return {c0.x * (1.0f - f) + c1.x * f, c0.y * (1.0f - f) + c1.y * f, c0.z * (1.0f - f) + c1.z * f,
            c0.w * (1.0f - f) + c1.w * f};
}

// texture(shadowMap, vec3(uv, layer)).r with _shadowSampler: LINEAR, CLAMP_TO_BORDER,
// border FLOAT_OPAQUE_WHITE, compare disabled (src/vk_engine_init.cpp:392-394).
float shadow_fetch(const OrkDevice* d, int layer, int x, int y) {
    if (x < 0 || y < 0 || x >= d->S || y >= d->S) return 1.0f;
    return d->shadow[((size_t)layer * d->S + y) * d->S + x];
}
float texture_shadow(const OrkDevice* d, float u, float v, int layer) {
    float fu = u * (float)d->S - 0.5f, fv = v * (float)d->S - 0.5f;
    float iu = std::floor(fu), iv = std::floor(fv);
    float a = fu - iu, b = fv - iv;
    // keep the int conversion defined for wild coordinates
    iu = clampf(iu, -2.0f, (float)d->S + 1.0f);
    iv = clampf(iv, -2.0f, (float)d->S + 1.0f);
    int x0 = (int)iu, y0 = (int)iv;
    float t00 = shadow_fetch(d, layer, x0, y0), t10 = shadow_fetch(d, layer, x0 + 1, y0);
    float t01 = shadow_fetch(d, layer, x0, y0 + 1), t11 = shadow_fetch(d, layer, x0 + 1, y0 + 1);
    return t00 * ((1.0f - a) * (1.0f - b)) + t10 * (a * (1.0f - b)) + t01 * ((1.0f - a) * b) + t11 * (a * b);
}

// Cube map face selection per the Vulkan spec table ("Cube Map Face Selection").
void cube_face(vec3 r, int& face, float& s, float& t) {
    float ax = std::fabs(r.x), ay = std::fabs(r.y), az = std::fabs(r.z);
    float sc, tc, ma;
    if (az >= ax && az >= ay) {          // spec: z wins ties, then y, then x
        if (r.z >= 0) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; }
        ma = az;
    } else if (ay >= ax) {
        if (r.y >= 0) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; }
        ma = ay;
    } else {
        if (r.x >= 0) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; }
        ma = ax;
    }
    s = 0.5f * (sc / ma + 1.0f);
    t = 0.5f * (tc / ma + 1.0f);
}
void cube_face_fixed(vec3 r, int face, float& s, float& t) {  // project onto a given face
    float sc, tc, ma;
    switch (face) {
        case 0: sc = -r.z; tc = -r.y; ma = std::fabs(r.x); break;
        case 1: sc = r.z; tc = -r.y; ma = std::fabs(r.x); break;
        case 2: sc = r.x; tc = r.z; ma = std::fabs(r.y); break;
        case 3: sc = r.x; tc = -r.z; ma = std::fabs(r.y); break;
        case 4: sc = r.x; tc = -r.y; ma = std::fabs(r.z); break;
        default: sc = -r.x; tc = -r.y; ma = std::fabs(r.z); break;
    }
    s = 0.5f * (sc / ma + 1.0f);
    t = 0.5f * (tc / ma + 1.0f);
}
// A texel of a cube face addressed possibly ONE texel beyond the face (a bilinear tap at a seam).  Vulkan filters cube maps
// seamlessly ("Cube Map Edge Handling"): a tap beyond one edge comes from the face that shares the edge.  The faces are
// unfolded about that edge: the tap's centre, d beyond the edge on this face's plane, is the point d INSIDE the neighbour's
// plane.  With the face-selection table read backwards, a face is  M + u S + v T  (M the major axis, u = 2s - 1, v = 2t - 1):
//   +X: M = (1,0,0), S = (0,0,-1), T = (0,-1,0)   -X: M = (-1,0,0), S = (0,0,1), T = (0,-1,0)
//   +Y: M = (0,1,0), S = (1,0,0),  T = (0,0,1)    -Y: M = (0,-1,0), S = (1,0,0), T = (0,0,-1)
//   +Z: M = (0,0,1), S = (1,0,0),  T = (0,-1,0)   -Z: M = (0,0,-1), S = (-1,0,0), T = (0,-1,0)
// Returns false for a tap beyond TWO edges (a cube corner: no such texel exists).
static const float CUBE_M[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
static const float CUBE_S[6][3] = {{0, 0, -1}, {0, 0, 1}, {1, 0, 0}, {1, 0, 0}, {1, 0, 0}, {-1, 0, 0}};
static const float CUBE_T[6][3] = {{0, -1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}, {0, -1, 0}, {0, -1, 0}};
bool cube_tap(const OrkDevice* d, int face, int level, int x, int y, vec4& out) {
    const Image& im = d->cube[face];
    const int w = im.w[level], h = im.h[level];
    const bool xo = x < 0 || x >= w, yo = y < 0 || y >= h;
    if (!xo && !yo) { out = fetch_texel(im, level, x, y); return true; }
    if (xo && yo) return false;
    const float u = (float)(2 * x + 1) / (float)w - 1.0f, v = (float)(2 * y + 1) / (float)h - 1.0f;
    const float* M = CUBE_M[face]; const float* S = CUBE_S[face]; const float* T = CUBE_T[face];
    float P[3];
    if (xo) {
        const float sg = u > 0 ? 1.0f : -1.0f, dd = std::fabs(u) - 1.0f;
        for (int k = 0; k < 3; ++k) P[k] = M[k] * (1.0f - dd) + S[k] * sg + T[k] * v;
    } else {
        const float sg = v > 0 ? 1.0f : -1.0f, dd = std::fabs(v) - 1.0f;
        for (int k = 0; k < 3; ++k) P[k] = M[k] * (1.0f - dd) + S[k] * u + T[k] * sg;
    }
    int nf; float ns, nt;
    cube_face(vec3{P[0], P[1], P[2]}, nf, ns, nt);
    const Image& nim = d->cube[nf];
    const int nw = nim.w[level], nh = nim.h[level];
    const int nx = std::clamp((int)std::floor(ns * (float)nw), 0, nw - 1), ny = std::clamp((int)std::floor(nt * (float)nh), 0, nh - 1);
    out = fetch_texel(nim, level, nx, ny);
    return true;
}
// LINEAR on a cube face, seamless: taps beyond an edge come from the neighbouring face; the one tap a footprint at a cube
// corner lacks is the average of the other three ("Cube Map Corner Handling": the texels of the three faces averaged).
vec4 sample_face_level(const OrkDevice* d, int face, int level, float s, float t) {
    const Image& im = d->cube[face];
    int w = im.w[level], h = im.h[level];
    float fu = s * (float)w - 0.5f, fv = t * (float)h - 0.5f;
    float iu = std::floor(fu), iv = std::floor(fv);
    float a = fu - iu, b = fv - iv;
    const int x0 = std::clamp((int)iu, -1, w - 1), y0 = std::clamp((int)iv, -1, h - 1);       // s, t in [0, 1]: at most one texel beyond
    vec4 tap[4];
    bool have[4];
    have[0] = cube_tap(d, face, level, x0, y0, tap[0]);
    have[1] = cube_tap(d, face, level, x0 + 1, y0, tap[1]);
    have[2] = cube_tap(d, face, level, x0, y0 + 1, tap[2]);
    have[3] = cube_tap(d, face, level, x0 + 1, y0 + 1, tap[3]);
    for (int k = 0; k < 4; ++k)
        if (!have[k]) {
            vec4 sum{0, 0, 0, 0};
            int n = 0;
            for (int j = 0; j < 4; ++j) if (have[j]) { sum.x += tap[j].x; sum.y += tap[j].y; sum.z += tap[j].z; sum.w += tap[j].w; ++n; }
            const float inv = n ? 1.0f / (float)n : 0.0f;
            tap[k] = {sum.x * inv, sum.y * inv, sum.z * inv, sum.w * inv};
        }
    const vec4 &t00 = tap[0], &t10 = tap[1], &t01 = tap[2], &t11 = tap[3];
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return {t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11,
            t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11,
            t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11,
            t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11};
}
// texture(samplerCube, dir) with derivatives taken in the centre fragment's face plane.
vec4 texture_cube(const OrkDevice* d, vec3 r, vec3 rdx, vec3 rdy) {
    int face; float s, t;
    cube_face(r, face, s, t);
    const Image& im = d->cube[face];
    int levels = im.levels();
    if (levels == 1) return sample_face_level(d, face, 0, s, t);
    float sx, tx, sy, ty;
    cube_face_fixed(rdx, face, sx, tx);
    cube_face_fixed(rdy, face, sy, ty);
    float w = (float)im.w[0], h = (float)im.h[0];
    float ax = (sx - s) * w, bx = (tx - t) * h, ay = (sy - s) * w, by = (ty - t) * h;
    float rho = std::max(std::sqrt(ax * ax + bx * bx), std::sqrt(ay * ay + by * by));
    float lambda = rho > 0.0f ? std::log2(rho) : -126.0f;
    float lod = clampf(lambda, 0.0f, (float)(levels - 1));
    int l0 = (int)std::floor(lod), l1 = std::min(l0 + 1, levels - 1);
    float f = lod - (float)l0;
    vec4 c0 = sample_face_level(d, face, l0, s, t);
    if (f == 0.0f) return c0;
    vec4 c1 = sample_face_level(d, face, l1, s, t);
    return {c0.x * (1.0f - f) + c1.x * f, c0.y * (1.0f - f) + c1.y * f, c0.z * (1.0f - f) + c1.z * f,
            c0.w * (1.0f - f) + c1.w * f};
}

// ------------------------------------------------------------------ shaders
const float PI = 3.14159265359f;   // shaders/mesh_pbr.frag:28
const float SHADOW_BIAS = 0.0005f; // :38

const float kPoisson[16][2] = {     // :68-85
    {-0.94201624f, -0.39906216f}, {0.94558609f, -0.76890725f}, {-0.094184101f, -0.92938870f},
    {0.34495938f, 0.29387760f},   {-0.91588581f, 0.45771432f}, {-0.81544232f, -0.87912464f},
    {-0.38277543f, 0.27676845f},  {0.97484398f, 0.75648379f},  {0.44323325f, -0.97511554f},
    {0.53742981f, -0.47373420f},  {-0.26496911f, -0.41893023f}, {0.79197514f, 0.19090188f},
    {-0.24188840f, 0.99706507f},  {-0.81409955f, 0.91437590f}, {0.19984126f, 0.78641367f},
    {0.14383161f, -0.14100790f}};

struct ShaderCtx {
    const OrkDevice* d;
    const VkrGPUSceneData* scene;
};

// compute_shadow — shaders/mesh_pbr.frag:39-46
float compute_shadow(const ShaderCtx& c, vec4 sc, int layer) {
    float shadow = 0.0f;
    float depth = texture_shadow(c.d, sc.x, sc.y, layer);
    float cur = sc.z;
    if (depth + SHADOW_BIAS < cur) shadow = 1.0f;
    return shadow;
}
// PCF — :48-59
float PCF(const ShaderCtx& c, vec4 sc, int layer) {
    float shadow = 0.0f;
    float texSize = 1.0f / (float)c.d->S;
    float cur = sc.z;
    for (int i = -1; i <= 1; ++i)
        for (int j = -1; j <= 1; ++j) {
            float pd = texture_shadow(c.d, sc.x + (float)i * texSize, sc.y + (float)j * texSize, layer);
            shadow += (cur - SHADOW_BIAS > pd) ? 1.0f : 0.0f;
        }
    return shadow / 9.0f;
}
const float NEAR_PLANE = 0.1f;                    // :63
const float LIGHT_SIZE_UV = 2.0f / 200.0f;        // :64-66
// findBlocker — :87-100
float findBlocker(const ShaderCtx& c, vec4 sc, int layer) {
    float zReceiver = sc.z;
    float blockerSum = 0;
    int numBlockers = 0;
    float searchWidth = LIGHT_SIZE_UV * (zReceiver - NEAR_PLANE) / zReceiver;
    for (int i = 0; i < 16; ++i) {
        float z = texture_shadow(c.d, sc.x + kPoisson[i][0] * searchWidth, sc.y + kPoisson[i][1] * searchWidth, layer);
        if (z + SHADOW_BIAS < zReceiver) { blockerSum += z; numBlockers++; }
    }
    return numBlockers != 0 ? blockerSum / (float)numBlockers : -1.0f;
}
// PCF_filter — :102-110
float PCF_filter(const ShaderCtx& c, vec4 sc, float r, int layer) {
    float cur = sc.z, sum = 0.0f;
    for (int i = 0; i < 16; ++i) {
        float pd = texture_shadow(c.d, sc.x + kPoisson[i][0] * r, sc.y + kPoisson[i][1] * r, layer);
        sum += (cur - SHADOW_BIAS > pd) ? 1.0f : 0.0f;
    }
    return sum / 16.0f;
}
// PCSS — :112-121 ; CSM — :123-125
float PCSS(const ShaderCtx& c, vec4 sc, int layer) {
    float zBlocker = findBlocker(c, sc, layer);
    if (zBlocker < 0) return 0.0f;
    float penumbraRatio = (sc.z - zBlocker) / zBlocker;
    float filterRadiusUV = penumbraRatio * LIGHT_SIZE_UV * NEAR_PLANE / sc.z;
    return PCF_filter(c, sc, filterRadiusUV, layer);
}
// calcShadow — :127-156 (identical block in shaders/mesh.frag:125-150)
float calcShadow(const ShaderCtx& c, vec4 posWorld, vec4 posView) {
    const VkrGPUSceneData& sd = *c.scene;
    float viewDepth = std::fabs(posView.z);
    int layer = VKR_NUM_CASCADES - 1;
    for (int i = 0; i < VKR_NUM_CASCADES; ++i)
        if (viewDepth < sd.cascadeDistances[i]) { layer = i; break; }
    int shadowMode = (int)sd.sunlightDirection[3];
    if (shadowMode < 3) layer = 0;
    // shadowCoord = biasMat * lightViewproj[layer] * inFragPosWorld  (left-assoc: (bias*L)*p)
    mat4 bias = {{{0.5f, 0, 0, 0}, {0, 0.5f, 0, 0}, {0, 0, 1.0f, 0}, {0.5f, 0.5f, 0, 1.0f}}};
    mat4 L = load_mat(sd.lightViewproj[layer]);
    vec4 sc = mul(mul(bias, L), posWorld);
    float shadow = 0.0f;
    bool enableShadow = sd.sunlightColor[3] > 0.5f;
    if (enableShadow) {
        if (shadowMode == 0) shadow = compute_shadow(c, sc, layer);
        else if (shadowMode == 1) shadow = PCF(c, sc, layer);
        else if (shadowMode == 2) shadow = PCSS(c, sc, layer);
        else if (shadowMode == 3) shadow = PCSS(c, sc, layer);
    }
    return shadow;
}

// Cook-Torrance helpers — shaders/mesh_pbr.frag:159-183
float DistributionGGX(vec3 N, vec3 H, float roughness) {
    float a = roughness * roughness;
    float a2 = a * a;
    float NdotH = std::max(dot(N, H), 0.0f);
    float denom = NdotH * NdotH * (a2 - 1.0f) + 1.0f;
    return a2 / (PI * denom * denom);
}
float GeometrySchlickGGX(float NdotV, float roughness) {
    float r = roughness + 1.0f;
    float k = r * r / 8.0f;
    return NdotV / (NdotV * (1.0f - k) + k);
}
float GeometrySmith(vec3 N, vec3 V, vec3 L, float roughness) {
    float NdotV = std::max(dot(N, V), 0.0f);
    float NdotL = std::max(dot(N, L), 0.0f);
    return GeometrySchlickGGX(NdotV, roughness) * GeometrySchlickGGX(NdotL, roughness);
}
vec3 fresnelSchlick(float cosTheta, vec3 F0) {
    float p = std::pow(clampf(1.0f - cosTheta, 0.0f, 1.0f), 5.0f);
    return F0 + (V3(1.0f) - F0) * p;
}

struct FragIn {
    vec3 normal, color; vec2 uv; vec3 lightVec; vec4 posWorld, posView;
    vec2 duvdx, duvdy;
};

// mesh_pbr.frag main() — shaders/mesh_pbr.frag:185-226. Returns false on discard.
bool frag_pbr(const ShaderCtx& c, const VkrMaterialConstants& m, vec3 camPos, const FragIn& in, vec4& out) {
    const VkrGPUSceneData& sd = *c.scene;
    vec3 N = normalize(in.normal);
    vec3 V = normalize(camPos - xyz(in.posWorld));
    vec3 L = normalize(in.lightVec);
    vec3 H = normalize(V + L);

    vec4 albedoTex = texture2D(c.d, m.albedoID, in.uv, in.duvdx, in.duvdy);
    if (albedoTex.w < 0.5f) return false;
    vec3 albedo = vec3{m.colorFactors[0], m.colorFactors[1], m.colorFactors[2]} * xyz(albedoTex) * in.color;
    vec4 mrTex = texture2D(c.d, m.metalRoughID, in.uv, in.duvdx, in.duvdy);
    float metallic = m.metal_rough_factors[0] * mrTex.z;
    float roughness = m.metal_rough_factors[1] * mrTex.y;

    vec3 F0 = V3(0.04f);
    F0 = mix(F0, albedo, metallic);

    vec3 Lo = V3(0.0f);
    vec3 radiance = {sd.sunlightColor[0], sd.sunlightColor[1], sd.sunlightColor[2]};

    float NDF = DistributionGGX(N, H, roughness);
    float G = GeometrySmith(N, V, L, roughness);
    vec3 F = fresnelSchlick(std::max(dot(H, V), 0.0f), F0);

    vec3 kS = F;
    vec3 kD = V3(1.0f) - kS;
    kD = kD * (1.0f - metallic);

    vec3 numerator = NDF * G * F;
    float denominator = 4.0f * std::max(dot(N, V), 0.0f) * std::max(dot(N, L), 0.0f) + 0.0001f;
    vec3 specular = numerator / denominator;

    float NdotL = std::max(dot(N, L), 0.0f);
    Lo = Lo + (kD * albedo / PI + specular) * NdotL * radiance;

    vec3 ambient = vec3{sd.ambientColor[0], sd.ambientColor[1], sd.ambientColor[2]} * albedo;
    float shadow = calcShadow(c, in.posWorld, in.posView);
    vec3 col = ambient + Lo * (1.0f - shadow);
    out = {col.x, col.y, col.z, 1.0f};
    return true;
}

// mesh.frag main() — shaders/mesh.frag:124-182 (Lambert; shadow evaluated before the discard).
bool frag_lambert(const ShaderCtx& c, const VkrMaterialConstants& m, const FragIn& in, vec4& out) {
    const VkrGPUSceneData& sd = *c.scene;
    float shadow = calcShadow(c, in.posWorld, in.posView);
    vec4 texColor = texture2D(c.d, m.albedoID, in.uv, in.duvdx, in.duvdy);
    vec3 color = in.color * xyz(texColor) * vec3{m.colorFactors[0], m.colorFactors[1], m.colorFactors[2]};
    if (texColor.w < 0.5f) return false;
    vec3 N = normalize(in.normal);
    vec3 L = normalize(in.lightVec);
    vec3 diffuse = std::max(0.0f, dot(N, L)) * color * vec3{sd.sunlightColor[0], sd.sunlightColor[1], sd.sunlightColor[2]};
    vec3 ambient = vec3{sd.ambientColor[0], sd.ambientColor[1], sd.ambientColor[2]} * color;
    vec3 col = diffuse * (1.0f - shadow) + ambient;
    out = {col.x, col.y, col.z, 1.0f};
    return true;
}

// tonemap.comp main() — shaders/tonemap.comp:9-22
vec3 tonemap(vec3 color) {
    vec3 mapped = color / (color + V3(1.0f));
    const float gamma = 2.2f;
    float e = 1.0f / gamma;
    return {std::pow(mapped.x, e), std::pow(mapped.y, e), std::pow(mapped.z, e)};
}

// ------------------------------------------------- fixed function: geometry
struct VSOut {            // mesh.vert outputs — shaders/mesh.vert:14-24
    vec4 clip;
    vec3 normal, color; vec2 uv; vec3 lightVec; vec4 posWorld, posView;
};

const float GUARD = 64.0f;  // guard band: clip x,y only beyond +-GUARD*w

struct RTri {               // a screen-space triangle ready for rasterisation
    int32_t X[3], Y[3];     // 24.8 fixed point, oriented so the doubled area is > 0
    float z0, dzdx, dzdy;   // depth plane about vertex 0 (pixels)
    int32_t bx0, by0, bx1, by1;  // inclusive pixel bounds (already scissored)
    uint32_t seq;           // submission-order primitive id
    uint32_t obj, tri;      // source object / triangle
};

inline vec4 clip_lerp(vec4 in, vec4 out, float din, float dout) {
    float t = din / (din - dout);
    return {in.x + t * (out.x - in.x), in.y + t * (out.y - in.y), in.z + t * (out.z - in.z), in.w + t * (out.w - in.w)};
}
inline float plane_dist(int p, vec4 v) {
    switch (p) {
        case 0: return v.z;                     // near: z >= 0
        case 1: return v.w - v.z;               // far:  z <= w
        case 2: return v.x + GUARD * v.w;       // guard band
        case 3: return GUARD * v.w - v.x;
        case 4: return v.y + GUARD * v.w;
        default: return GUARD * v.w - v.y;
    }
}
// Sutherland-Hodgman against plane p; intersection always computed from the inside vertex
// towards the outside vertex so both triangles sharing an edge get the same point.
int clip_poly(int p, const vec4* in, int n, vec4* out) {
    int m = 0;
    for (int i = 0; i < n; ++i) {
        vec4 a = in[i], b = in[(i + 1) % n];
        float da = plane_dist(p, a), db = plane_dist(p, b);
        bool ia = da >= 0.0f, ib = db >= 0.0f;
        if (ia) out[m++] = a;
        if (ia != ib) out[m++] = ia ? clip_lerp(a, b, da, db) : clip_lerp(b, a, db, da);
    }
    return m;
}

struct Viewport { float halfW, halfH; int sx0, sy0, sx1, sy1; };  // scissor: [sx0,sx1) x [sy0,sy1)

enum CullMode { CULL_BACK = 0, CULL_FRONT = 1 };

// Project, snap, cull and set up one clip-space triangle (all vertices inside the clip volume).
bool setup_tri(const vec4 c[3], const Viewport& vp, CullMode cull, RTri& t) {
    int32_t X[3], Y[3];
    float z[3];
    for (int i = 0; i < 3; ++i) {
        float rw = 1.0f / c[i].w;
        float xn = c[i].x * rw, yn = c[i].y * rw;
        z[i] = c[i].z * rw;
        // viewport transform (x=0,y=0,minDepth=0,maxDepth=1) fused with the snap to 8 sub-pixel bits
        // (subPixelPrecisionBits = 8): X = rint(xn * 128W + 128W), round-half-even
        X[i] = (int32_t)std::nearbyint(std::fmaf(xn, vp.halfW * 256.0f, vp.halfW * 256.0f));
        Y[i] = (int32_t)std::nearbyint(std::fmaf(yn, vp.halfH * 256.0f, vp.halfH * 256.0f));
    }
    int64_t A2 = (int64_t)(X[1] - X[0]) * (Y[2] - Y[0]) - (int64_t)(X[2] - X[0]) * (Y[1] - Y[0]);
    // Vulkan: a = -1/2 * sum(x_i*y_j - x_j*y_i) in y-down framebuffer coords; frontFace CCW => a>0 front.
    // a = -A2/2, so front-facing <=> A2 < 0.  (src/vk_engine_init.cpp:441,536)
    if (A2 == 0) return false;
    bool front = A2 < 0;
    if (cull == CULL_BACK && !front) return false;
    if (cull == CULL_FRONT && front) return false;
    if (A2 < 0) { std::swap(X[1], X[2]); std::swap(Y[1], Y[2]); std::swap(z[1], z[2]); A2 = -A2; }
    int32_t minX = std::min(X[0], std::min(X[1], X[2])), maxX = std::max(X[0], std::max(X[1], X[2]));
    int32_t minY = std::min(Y[0], std::min(Y[1], Y[2])), maxY = std::max(Y[0], std::max(Y[1], Y[2]));
    // pixel centres at (i*256+128)
    int32_t bx0 = (minX + 127) >> 8, bx1 = (maxX - 128) >> 8;
    int32_t by0 = (minY + 127) >> 8, by1 = (maxY - 128) >> 8;
    bx0 = std::max(bx0, vp.sx0); by0 = std::max(by0, vp.sy0);
    bx1 = std::min(bx1, vp.sx1 - 1); by1 = std::min(by1, vp.sy1 - 1);
    if (bx0 > bx1 || by0 > by1) return false;
    for (int i = 0; i < 3; ++i) { t.X[i] = X[i]; t.Y[i] = Y[i]; }
    t.bx0 = bx0; t.by0 = by0; t.bx1 = bx1; t.by1 = by1;
    // depth plane: z(x,y) = z0 + dzdx*(x-x0) + dzdy*(y-y0), x,y in pixels
    float d1x = (float)(X[1] - X[0]), d1y = (float)(Y[1] - Y[0]);
    float d2x = (float)(X[2] - X[0]), d2y = (float)(Y[2] - Y[0]);
    float d1z = z[1] - z[0], d2z = z[2] - z[0];
    float fA = (float)A2;
    t.z0 = z[0];
    float rA = 1.0f / fA;
    t.dzdx = ((d1z * d2y - d2z * d1y) * rA) * 256.0f;
    t.dzdy = ((d2z * d1x - d1z * d2x) * rA) * 256.0f;
    return true;
}

// Edge function data for coverage with the top-left rule.
struct Edges { int64_t a[3], b[3], c[3]; };  // E_i(P) = a*Px + b*Py + c, inside iff E_i >= 0 after bias
inline void make_edges(const RTri& t, Edges& e) {
    for (int i = 0; i < 3; ++i) {
        int j = (i + 1) % 3, k = (i + 2) % 3;          // edge j->k is opposite vertex i
        int64_t dx = t.X[k] - t.X[j], dy = t.Y[k] - t.Y[j];
        // cross((k-j),(P-j)) = dx*(Py-Yj) - dy*(Px-Xj)
        e.a[i] = -dy; e.b[i] = dx;
        e.c[i] = dy * t.X[j] - dx * t.Y[j];
        bool topleft = (dy < 0) || (dy == 0 && dx > 0);  // left edge: dy<0 ; top edge: dy==0 && dx>0 (y down, area>0)
        if (!topleft) e.c[i] -= 1;
    }
}
inline bool covered(const Edges& e, int px, int py) {
    int64_t Px = (int64_t)px * 256 + 128, Py = (int64_t)py * 256 + 128;
    for (int i = 0; i < 3; ++i)
        if (e.a[i] * Px + e.b[i] * Py + e.c[i] < 0) return false;
    return true;
}
inline float depth_at(const RTri& t, int px, int py) {
    float fx = (float)(px * 256 + 128 - t.X[0]) * (1.0f / 256.0f);
    float fy = (float)(py * 256 + 128 - t.Y[0]) * (1.0f / 256.0f);
    float z = std::fmaf(t.dzdy, fy, std::fmaf(t.dzdx, fx, t.z0));
    if (!(z > 0.0f)) z = 0.0f;       // depth clamp to [0,1] for the D32 attachment
    if (z > 1.0f) z = 1.0f;
    return z;
}

// Transform one triangle to clip space and emit its raster triangles (0..7 after clipping).
template <class Emit>
void process_triangle(const vec4 cin[3], const Viewport& vp, CullMode cull, uint32_t seq, uint32_t obj,
                      uint32_t tri, Emit&& emit) {
    // trivial reject against the view volume
    bool rej = (cin[0].z < 0 && cin[1].z < 0 && cin[2].z < 0) ||
               (cin[0].z > cin[0].w && cin[1].z > cin[1].w && cin[2].z > cin[2].w) ||
               (cin[0].x < -cin[0].w && cin[1].x < -cin[1].w && cin[2].x < -cin[2].w) ||
               (cin[0].x > cin[0].w && cin[1].x > cin[1].w && cin[2].x > cin[2].w) ||
               (cin[0].y < -cin[0].w && cin[1].y < -cin[1].w && cin[2].y < -cin[2].w) ||
               (cin[0].y > cin[0].w && cin[1].y > cin[1].w && cin[2].y > cin[2].w);
    if (rej) return;
    bool needClip = false;
    for (int i = 0; i < 3; ++i)
        for (int p = 0; p < 6; ++p)
            if (!(plane_dist(p, cin[i]) >= 0.0f)) needClip = true;
    RTri t;
    t.seq = seq; t.obj = obj; t.tri = tri;
    if (!needClip) {
        if (setup_tri(cin, vp, cull, t)) emit(t);
        return;
    }
    vec4 bufA[12], bufB[12];
    int n = 3;
    bufA[0] = cin[0]; bufA[1] = cin[1]; bufA[2] = cin[2];
    vec4 *src = bufA, *dst = bufB;
    for (int p = 0; p < 6 && n >= 3; ++p) {
        n = clip_poly(p, src, n, dst);
        std::swap(src, dst);
    }
    if (n < 3) return;
    for (int i = 1; i + 1 < n; ++i) {
        vec4 c[3] = {src[0], src[i], src[i + 1]};
        if (!(c[0].w > 0.0f && c[1].w > 0.0f && c[2].w > 0.0f)) continue;
        if (setup_tri(c, vp, cull, t)) emit(t);
    }
}

// Perspective-correct barycentrics of the ORIGINAL (unclipped) triangle at a window position,
// by 2-D homogeneous edge functions on clip coordinates.
inline void bary_at(const vec4 c[3], float sx, float sy, const Viewport& vp, float b[3]) {
    float nx = (sx - vp.halfW) / vp.halfW, ny = (sy - vp.halfH) / vp.halfH;
    float qx[3], qy[3];
    for (int i = 0; i < 3; ++i) { qx[i] = c[i].x - nx * c[i].w; qy[i] = c[i].y - ny * c[i].w; }
    float e0 = qx[1] * qy[2] - qx[2] * qy[1];
    float e1 = qx[2] * qy[0] - qx[0] * qy[2];
    float e2 = qx[0] * qy[1] - qx[1] * qy[0];
    float s = e0 + e1 + e2;
    if (s == 0.0f) { b[0] = 1; b[1] = 0; b[2] = 0; return; }
    b[0] = e0 / s; b[1] = e1 / s; b[2] = e2 / s;
}
template <class T> inline T lerp3(const T& a0, const T& a1, const T& a2, const float b[3]);
template <> inline float lerp3(const float& a0, const float& a1, const float& a2, const float b[3]) {
    return b[0] * a0 + b[1] * a1 + b[2] * a2;
}
inline vec3 lerp3v(vec3 a0, vec3 a1, vec3 a2, const float b[3]) {
    return {lerp3(a0.x, a1.x, a2.x, b), lerp3(a0.y, a1.y, a2.y, b), lerp3(a0.z, a1.z, a2.z, b)};
}
inline vec4 lerp3v(vec4 a0, vec4 a1, vec4 a2, const float b[3]) {
    return {lerp3(a0.x, a1.x, a2.x, b), lerp3(a0.y, a1.y, a2.y, b), lerp3(a0.z, a1.z, a2.z, b), lerp3(a0.w, a1.w, a2.w, b)};
}

void parallel_for(int nthreads, int n, const std::function<void(int)>& fn) {
    if (nthreads <= 1 || n <= 1) { for (int i = 0; i < n; ++i) fn(i); return; }
    std::atomic<int> next{0};
    std::vector<std::thread> th;
    int nt = std::min(nthreads, n);
    for (int t = 0; t < nt; ++t)
        th.emplace_back([&] { for (;;) { int i = next.fetch_add(1); if (i >= n) break; fn(i); } });
    for (auto& t : th) t.join();
}

struct DrawCtx {
    OrkDevice* d;
    const VkrGPUSceneData* scene;
    const VkrRenderObject* objs;
    size_t nobj;
    const VkrFrameParams* fp;
};

// Geometry front end for one pass: returns raster triangles in submission order.
void build_rtris(const DrawCtx& dc, const mat4& vpMat, const Viewport& vp, CullMode cull, bool countSeq,
                 std::vector<RTri>& out) {
    OrkDevice* d = dc.d;
    // chunk objects' triangles for parallel setup, keep order
    struct Chunk { uint32_t obj, t0, t1, seq0; };
    std::vector<Chunk> chunks;
    uint32_t seq = 0;
    for (size_t o = 0; o < dc.nobj; ++o) {
        uint32_t nt = dc.objs[o].indexCount / 3;
        for (uint32_t t0 = 0; t0 < nt; t0 += 4096)
            chunks.push_back({(uint32_t)o, t0, std::min(nt, t0 + 4096), seq + t0});
        seq += nt;
    }
    (void)countSeq;
    std::vector<std::vector<RTri>> parts(chunks.size());
    parallel_for(d->nthreads, (int)chunks.size(), [&](int ci) {
        const Chunk& ch = chunks[ci];
        const VkrRenderObject& ro = dc.objs[ch.obj];
        const Mesh& mesh = d->meshes[ro.mesh];
        mat4 world = load_mat(ro.transform);
        mat4 M = mul(vpMat, world);   // viewproj * worldMatrix (left-assoc GLSL, shaders/mesh.vert:16)
        auto& outp = parts[ci];
        for (uint32_t t = ch.t0; t < ch.t1; ++t) {
            vec4 c[3];
            for (int k = 0; k < 3; ++k) {
                uint32_t vi = mesh.idx[ro.firstIndex + 3 * t + k];
                const VkrVertex& v = mesh.v[vi];
                c[k] = mul_point_fma(M, v.position[0], v.position[1], v.position[2]);
            }
            process_triangle(c, vp, cull, ch.seq0 + (t - ch.t0), ch.obj, t, [&](const RTri& rt) { outp.push_back(rt); });
        }
    });
    size_t total = 0;
    for (auto& p : parts) total += p.size();
    out.clear();
    out.reserve(total);
    for (auto& p : parts) out.insert(out.end(), p.begin(), p.end());
}

// draw_shadow — src/vk_engine_run.cpp:334-382 with shadow.vert (shaders/shadow.vert:6-9):
// depth-only, cull FRONT, LEQUAL + write, clear 1.0, one layer (mode<3) or NUM_CASCADES.
void pass_shadow(const DrawCtx& dc) {
    OrkDevice* d = dc.d;
    int layers = dc.fp->shadowMode < 3 ? 1 : VKR_NUM_CASCADES;
    int S = d->S;
    Viewport vp{(float)S * 0.5f, (float)S * 0.5f, 0, 0, S, S};
    for (int layer = 0; layer < layers; ++layer) {
        float* sm = &d->shadow[(size_t)layer * S * S];
        std::fill(sm, sm + (size_t)S * S, 1.0f);
        std::vector<RTri> tris;
        build_rtris(dc, load_mat(dc.scene->lightViewproj[layer]), vp, CULL_FRONT, false, tris);
        const int BAND = 64;
        int nb = (S + BAND - 1) / BAND;
        parallel_for(d->nthreads, nb, [&](int b) {
            int y0 = b * BAND, y1 = std::min(S, y0 + BAND) - 1;
            for (const RTri& t : tris) {
                if (t.by1 < y0 || t.by0 > y1) continue;
                Edges e; make_edges(t, e);
                for (int py = std::max(t.by0, y0); py <= std::min(t.by1, y1); ++py)
                    for (int px = t.bx0; px <= t.bx1; ++px) {
                        if (!covered(e, px, py)) continue;
                        float z = depth_at(t, px, py);
                        float& dst = sm[(size_t)py * S + px];
                        if (z <= dst) dst = z;
                    }
            }
        });
    }
}

// mesh.vert main() — shaders/mesh.vert:14-24
VSOut vertex_main(const VkrVertex& v, const mat4& world, const mat4& M, const VkrGPUSceneData& sd) {
    VSOut o;
    vec4 p = {v.position[0], v.position[1], v.position[2], 1.0f};
    o.clip = mul_point_fma(M, v.position[0], v.position[1], v.position[2]);
    o.normal = mul3(world, vec3{v.normal[0], v.normal[1], v.normal[2]});
    o.color = {v.color[0], v.color[1], v.color[2]};
    o.uv = {v.uv_x, v.uv_y};
    o.lightVec = normalize(-vec3{sd.sunlightDirection[0], sd.sunlightDirection[1], sd.sunlightDirection[2]});
    o.posWorld = mul(world, p);
    o.posView = mul(load_mat(sd.view), o.posWorld);
    return o;
}

void write_hdr(OrkDevice* d, size_t pix, vec4 c) {
    uint16_t* p = &d->hdr[pix * 4];
    p[0] = f32_to_f16(c.x); p[1] = f32_to_f16(c.y); p[2] = f32_to_f16(c.z); p[3] = f32_to_f16(c.w);
}
vec4 read_hdr(const OrkDevice* d, size_t pix) {
    const uint16_t* p = &d->hdr[pix * 4];
    return {f16_to_f32(p[0]), f16_to_f32(p[1]), f16_to_f32(p[2]), f16_to_f32(p[3])};
}

// gradient_color.comp main() — shaders/gradient_color.comp:16-31: mix(top, bottom, y / height)
vec4 gradient_color(const float* top, const float* bot, int y, int height) {
    float blend = (float)y / (float)height;
    return {top[0] * (1.0f - blend) + bot[0] * blend, top[1] * (1.0f - blend) + bot[1] * blend,
            top[2] * (1.0f - blend) + bot[2] * blend, top[3] * (1.0f - blend) + bot[3] * blend};
}

// Per-pixel form of skybox.vert (shaders/skybox.vert:8-17): the cube position whose projection
// proj * mat4(mat3(view)) * pos lands on window position (sx, sy), with outUVW.y negated.
vec3 skybox_dir(const mat4& view, const mat4& proj, float halfW, float halfH, float sx, float sy) {
    float nx = (sx - halfW) / halfW, ny = (sy - halfH) / halfH;
    // view-space ray: proj * (xv, yv, -1, 1) = (nx, ny, ., 1) * w  for a perspective proj
    vec3 dv = {(nx + proj.m[2][0]) / proj.m[0][0], (ny + proj.m[2][1]) / proj.m[1][1], -1.0f};
    // pos = transpose(mat3(view)) * dv   (inverse of the rotation skybox.vert applies, skybox.vert:14-15)
    vec3 p = {view.m[0][0] * dv.x + view.m[0][1] * dv.y + view.m[0][2] * dv.z,
              view.m[1][0] * dv.x + view.m[1][1] * dv.y + view.m[1][2] * dv.z,
              view.m[2][0] * dv.x + view.m[2][1] * dv.y + view.m[2][2] * dv.z};
    p.y *= -1.0f;   // outUVW.y *= -1.0 (skybox.vert:11)
    return p;
}

// draw_background — src/vk_engine_run.cpp:234-242 + gradient_color.comp:16-31
void pass_background(const DrawCtx& dc) {
    OrkDevice* d = dc.d;
    const float* top = dc.fp->background.data1;
    const float* bot = dc.fp->background.data2;
    parallel_for(d->nthreads, d->H, [&](int y) {
        vec4 c = gradient_color(top, bot, y, d->H);
        for (int x = 0; x < d->W; ++x) write_hdr(d, (size_t)y * d->W + x, c);
    });
}

// draw_geometry — src/vk_engine_run.cpp:244-316 (+ draw_skybox :318-332)
void pass_geometry(const DrawCtx& dc) {
    OrkDevice* d = dc.d;
    const VkrGPUSceneData& sd = *dc.scene;
    int W = d->W, H = d->H;
    Viewport vp{(float)W * 0.5f, (float)H * 0.5f, 0, 0, W, H};
    if (dc.fp->scissor[2] > 0 && dc.fp->scissor[3] > 0) {
        vp.sx0 = std::max(0, dc.fp->scissor[0]); vp.sy0 = std::max(0, dc.fp->scissor[1]);
        vp.sx1 = std::min(W, dc.fp->scissor[0] + dc.fp->scissor[2]);
        vp.sy1 = std::min(H, dc.fp->scissor[1] + dc.fp->scissor[3]);
    }
    // attachment clears are limited to the render area = full extent (vkinit::rendering_info)
    std::fill(d->depth.begin(), d->depth.end(), 1.0f);
    std::fill(d->primid.begin(), d->primid.end(), 0xFFFFFFFFu);
    if (!dc.fp->enableBackground)
        parallel_for(d->nthreads, H, [&](int y) { for (int x = 0; x < W; ++x) write_hdr(d, (size_t)y * W + x, {0, 0, 0, 1}); });

    mat4 viewproj = load_mat(sd.viewproj);
    std::vector<RTri> tris;
    build_rtris(dc, viewproj, vp, CULL_BACK, true, tris);

    mat4 invView = inverse(load_mat(sd.view));
    vec3 camPos = {invView.m[3][0], invView.m[3][1], invView.m[3][2]};   // vec3(inverse(view)[3])
    ShaderCtx sc{d, dc.scene};

    const int BAND = 16;
    int nb = (H + BAND - 1) / BAND;
    parallel_for(d->nthreads, nb, [&](int b) {
        int y0 = b * BAND, y1 = std::min(H, y0 + BAND) - 1;
        uint32_t cachedSeq = 0xFFFFFFFFu;
        VSOut vs[3];
        vec4 clip[3];
        for (const RTri& t : tris) {
            if (t.by1 < y0 || t.by0 > y1) continue;
            const VkrRenderObject& ro = dc.objs[t.obj];
            const Material& mat = d->materials[ro.material];
            bool transparent = mat.pass == VKR_PASS_TRANSPARENT;
            Edges e; make_edges(t, e);
            for (int py = std::max(t.by0, y0); py <= std::min(t.by1, y1); ++py)
                for (int px = t.bx0; px <= t.bx1; ++px) {
                    if (!covered(e, px, py)) continue;
                    float z = depth_at(t, px, py);
                    size_t pix = (size_t)py * W + px;
                    if (!(z <= d->depth[pix])) continue;       // VK_COMPARE_OP_LESS_OR_EQUAL
                    if (cachedSeq != t.seq) {                 // run the vertex shader for this primitive
                        const Mesh& mesh = d->meshes[ro.mesh];
                        mat4 world = load_mat(ro.transform);
                        mat4 M = mul(viewproj, world);
                        for (int k = 0; k < 3; ++k) {
                            vs[k] = vertex_main(mesh.v[mesh.idx[ro.firstIndex + 3 * t.tri + k]], world, M, sd);
                            clip[k] = vs[k].clip;
                        }
                        cachedSeq = t.seq;
                    }
                    // interpolate varyings; derivatives are quad finite differences on the primitive's plane
                    float bc[3], bx[3], by[3];
                    float cx = (float)px + 0.5f, cy = (float)py + 0.5f;
                    float ox = (px & 1) ? -1.0f : 1.0f, oy = (py & 1) ? -1.0f : 1.0f;
                    bary_at(clip, cx, cy, vp, bc);
                    bary_at(clip, cx + ox, cy, vp, bx);
                    bary_at(clip, cx, cy + oy, vp, by);
                    FragIn in;
                    in.normal = lerp3v(vs[0].normal, vs[1].normal, vs[2].normal, bc);
                    in.color = lerp3v(vs[0].color, vs[1].color, vs[2].color, bc);
                    in.uv = {lerp3(vs[0].uv.x, vs[1].uv.x, vs[2].uv.x, bc), lerp3(vs[0].uv.y, vs[1].uv.y, vs[2].uv.y, bc)};
                    in.lightVec = lerp3v(vs[0].lightVec, vs[1].lightVec, vs[2].lightVec, bc);
                    in.posWorld = lerp3v(vs[0].posWorld, vs[1].posWorld, vs[2].posWorld, bc);
                    in.posView = lerp3v(vs[0].posView, vs[1].posView, vs[2].posView, bc);
                    vec2 uvx = {lerp3(vs[0].uv.x, vs[1].uv.x, vs[2].uv.x, bx), lerp3(vs[0].uv.y, vs[1].uv.y, vs[2].uv.y, bx)};
                    vec2 uvy = {lerp3(vs[0].uv.x, vs[1].uv.x, vs[2].uv.x, by), lerp3(vs[0].uv.y, vs[1].uv.y, vs[2].uv.y, by)};
                    in.duvdx = {(uvx.x - in.uv.x) * ox, (uvx.y - in.uv.y) * ox};
                    in.duvdy = {(uvy.x - in.uv.x) * oy, (uvy.y - in.uv.y) * oy};
                    vec4 col;
                    bool keep = dc.fp->fragShader == VKR_FRAG_MESH_LAMBERT ? frag_lambert(sc, mat.c, in, col)
                                                                            : frag_pbr(sc, mat.c, camPos, in, col);
                    if (!keep) continue;                       // discard: no depth, no colour
                    if (transparent) {
                        // SRC_ALPHA, ONE, ADD ; alpha ONE, ZERO (src/vk_pipelines.cpp:108-118); no depth write
                        vec4 dst = read_hdr(d, pix);
                        write_hdr(d, pix, {col.x * col.w + dst.x, col.y * col.w + dst.y, col.z * col.w + dst.z, col.w});
                    } else {
                        d->depth[pix] = z;
                        d->primid[pix] = t.seq;
                        write_hdr(d, pix, col);
                    }
                }
        }
    });

    // draw_skybox: cull FRONT cube at depth 1.0, LEQUAL, no depth write -> every pixel whose depth is still 1.0
    if (dc.fp->enableSkybox && d->haveCube) {
        mat4 view = load_mat(sd.view), proj = load_mat(sd.proj);
        auto dir_at = [&](float sx, float sy) { return skybox_dir(view, proj, vp.halfW, vp.halfH, sx, sy); };
        parallel_for(d->nthreads, vp.sy1 - vp.sy0, [&](int yy) {
            int py = vp.sy0 + yy;
            for (int px = vp.sx0; px < vp.sx1; ++px) {
                size_t pix = (size_t)py * W + px;
                if (!(1.0f <= d->depth[pix])) continue;
                float cx = (float)px + 0.5f, cy = (float)py + 0.5f;
                float ox = (px & 1) ? -1.0f : 1.0f, oy = (py & 1) ? -1.0f : 1.0f;
                vec4 c = texture_cube(d, dir_at(cx, cy), dir_at(cx + ox, cy), dir_at(cx, cy + oy));
                write_hdr(d, pix, c);   // skybox.frag:11
            }
        });
    }
}

// draw_postprocess — src/vk_engine_run.cpp:384-408 + tonemap.comp; result copied back to the draw image
void pass_post(const DrawCtx& dc) {
    OrkDevice* d = dc.d;
    parallel_for(d->nthreads, d->H, [&](int y) {
        for (int x = 0; x < d->W; ++x) {
            size_t pix = (size_t)y * d->W + x;
            vec4 c = read_hdr(d, pix);
            vec3 m = tonemap({c.x, c.y, c.z});
            write_hdr(d, pix, {m.x, m.y, m.z, 1.0f});
        }
    });
}

// copy_image_to_image draw -> swapchain (src/vk_engine_run.cpp:159-161): 1:1 LINEAR blit to UNORM8
void pass_present(const DrawCtx& dc) {
    OrkDevice* d = dc.d;
    parallel_for(d->nthreads, d->H, [&](int y) {
        for (int x = 0; x < d->W; ++x) {
            size_t pix = (size_t)y * d->W + x;
            vec4 c = read_hdr(d, pix);
            uint8_t* o = &d->rgba8[pix * 4];
            o[0] = to_unorm8(c.x); o[1] = to_unorm8(c.y); o[2] = to_unorm8(c.z); o[3] = to_unorm8(c.w);
        }
    });
}

}  // namespace

// ----------------------------------------------------------------- C API
extern "C" {

int ork_create(const VkrConfig* cfg, OrkDevice** out) {
    if (!cfg || !out || cfg->width == 0 || cfg->height == 0) return VKR_ERROR_INVALID_ARGUMENT;
    init_lut();
    OrkDevice* d = new OrkDevice();
    d->W = (int)cfg->width; d->H = (int)cfg->height;
    d->S = cfg->shadowExtent ? (int)cfg->shadowExtent : 2048;
    d->nthreads = (int)std::thread::hardware_concurrency();
    size_t n = (size_t)d->W * d->H;
    d->depth.assign(n, 1.0f);
    d->primid.assign(n, 0xFFFFFFFFu);
    d->hdr.assign(n * 4, 0);
    d->rgba8.assign(n * 4, 0);
    d->shadow.assign((size_t)d->S * d->S * VKR_NUM_CASCADES, 1.0f);
    // init_default_data (src/vk_engine_init.cpp:318-355): slot 0 ends up holding the default NORMAL texture
    uint32_t normalDefault = 0xFFFF8080u;
    VkrImage img;
    ork_create_image(d, &normalDefault, 1, 1, VKR_FORMAT_R8G8B8A8_UNORM, 0, &img);
    ork_update_bindless_texture(d, 0, img, VKR_SAMPLER_NEAREST);
    *out = d;
    return VKR_SUCCESS;
}
void ork_destroy(OrkDevice* d) { delete d; }
int ork_set_threads(OrkDevice* d, int n) {
    if (!d) return VKR_ERROR_INVALID_ARGUMENT;
    d->nthreads = n > 0 ? n : (int)std::thread::hardware_concurrency();
    return VKR_SUCCESS;
}
int ork_upload_mesh(OrkDevice* d, const uint32_t* indices, size_t ni, const VkrVertex* v, size_t nv, VkrMesh* out) {
    if (!d || !out || (ni && !indices) || (nv && !v)) return VKR_ERROR_INVALID_ARGUMENT;
    for (size_t i = 0; i < ni; ++i) if (indices[i] >= nv) { d->err = "index out of range"; return VKR_ERROR_INVALID_ARGUMENT; }
    Mesh m;
    m.idx.assign(indices, indices + ni);
    m.v.assign(v, v + nv);
    d->meshes.push_back(std::move(m));
    *out = (VkrMesh)(d->meshes.size() - 1);
    return VKR_SUCCESS;
}
int ork_create_image(OrkDevice* d, const void* rgba8, uint32_t w, uint32_t h, VkrFormat fmt, int mipmapped, VkrImage* out) {
    if (!d || !rgba8 || !out || w == 0 || h == 0) return VKR_ERROR_INVALID_ARGUMENT;
    Image im;
    im.format = fmt;
    im.w.push_back((int)w); im.h.push_back((int)h);
    im.mips.emplace_back((const uint8_t*)rgba8, (const uint8_t*)rgba8 + (size_t)w * h * 4);
    if (mipmapped) build_mips(im);
    d->images.push_back(std::move(im));
    *out = (VkrImage)(d->images.size() - 1);
    return VKR_SUCCESS;
}
int ork_update_bindless_texture(OrkDevice* d, int index, VkrImage image, VkrSampler sampler) {
    if (!d || index < 0 || index >= VKR_MAX_BINDLESS || image >= d->images.size()) return VKR_ERROR_INVALID_ARGUMENT;
    d->slots[index].image = (int)image;
    d->slots[index].sampler = sampler;
    return VKR_SUCCESS;
}
int ork_set_cubemap(OrkDevice* d, const void* const faces[6], uint32_t w, uint32_t h, VkrFormat fmt, int mipmapped) {
    if (!d || !faces || w == 0 || h == 0) return VKR_ERROR_INVALID_ARGUMENT;
    for (int f = 0; f < 6; ++f) {
        if (!faces[f]) return VKR_ERROR_INVALID_ARGUMENT;
        Image im;
        im.format = fmt;
        im.w.push_back((int)w); im.h.push_back((int)h);
        im.mips.emplace_back((const uint8_t*)faces[f], (const uint8_t*)faces[f] + (size_t)w * h * 4);
        if (mipmapped) build_mips(im);
        d->cube[f] = std::move(im);
    }
    d->haveCube = true;
    return VKR_SUCCESS;
}
// load_cubemap: every level comes from the KTX file (ktxTexture_VkUploadEx, src/vk_loader.cpp:533-536); no mip generation.
int ork_set_cubemap_levels(OrkDevice* d, const void* const* levels, uint32_t w, uint32_t h, uint32_t levelCount, VkrFormat fmt) {
    if (!d || !levels || w == 0 || h == 0 || levelCount == 0 || levelCount > 16) return VKR_ERROR_INVALID_ARGUMENT;
    const size_t bpt = fmt == VKR_FORMAT_R16G16B16A16_SFLOAT ? 8 : 4;
    for (int f = 0; f < 6; ++f) {
        Image im;
        im.format = fmt;
        for (uint32_t l = 0; l < levelCount; ++l) {
            const int lw = (int)std::max(1u, w >> l), lh = (int)std::max(1u, h >> l);
            const uint8_t* src = (const uint8_t*)levels[(size_t)l * 6 + f];
            if (!src) return VKR_ERROR_INVALID_ARGUMENT;
            im.w.push_back(lw); im.h.push_back(lh);
            im.mips.emplace_back(src, src + (size_t)lw * lh * bpt);
        }
        d->cube[f] = std::move(im);
    }
    d->haveCube = true;
    return VKR_SUCCESS;
}
int ork_build_material(OrkDevice* d, const VkrMaterialConstants* p, VkrMaterialPass pass, VkrMaterial* out) {
    if (!d || !p || !out) return VKR_ERROR_INVALID_ARGUMENT;
    d->materials.push_back({*p, (int)pass});
    *out = (VkrMaterial)(d->materials.size() - 1);
    return VKR_SUCCESS;
}
int ork_draw(OrkDevice* d, const VkrGPUSceneData* scene, const VkrRenderObject* objs, size_t n, const VkrFrameParams* fp) {
    if (!d || !scene || !fp || (n && !objs)) return VKR_ERROR_INVALID_ARGUMENT;
    d->stats.triangle_count = 0;
    d->stats.drawcall_count = 0;
    for (size_t i = 0; i < n; ++i) {
        const VkrRenderObject& o = objs[i];
        if (o.mesh >= d->meshes.size() || o.material >= d->materials.size()) { d->err = "bad handle"; return VKR_ERROR_INVALID_ARGUMENT; }
        if ((size_t)o.firstIndex + o.indexCount > d->meshes[o.mesh].idx.size()) { d->err = "index range"; return VKR_ERROR_INVALID_ARGUMENT; }
        d->stats.drawcall_count++;
        d->stats.triangle_count += (int)(o.indexCount / 3);
    }
    DrawCtx dc{d, scene, objs, n, fp};
    if (fp->enableShadows) pass_shadow(dc);
    if (fp->enableBackground) pass_background(dc);
    pass_geometry(dc);
    if (fp->enablePostprocess) pass_post(dc);
    pass_present(dc);
    return VKR_SUCCESS;
}
int ork_readback(OrkDevice* d, VkrTarget what, void* dst, size_t bytes) {
    if (!d || !dst) return VKR_ERROR_INVALID_ARGUMENT;
    const void* src; size_t n;
    switch (what) {
        case VKR_TARGET_RGBA8: src = d->rgba8.data(); n = d->rgba8.size(); break;
        case VKR_TARGET_HDR16F: src = d->hdr.data(); n = d->hdr.size() * 2; break;
        case VKR_TARGET_DEPTH: src = d->depth.data(); n = d->depth.size() * 4; break;
        case VKR_TARGET_SHADOW: src = d->shadow.data(); n = d->shadow.size() * 4; break;
        case VKR_TARGET_PRIMID: src = d->primid.data(); n = d->primid.size() * 4; break;
        case VKR_TARGET_RGB8: {
            const size_t px = d->rgba8.size() / 4;
            if (bytes < px * 3) return VKR_ERROR_INVALID_ARGUMENT;
            uint8_t* o = (uint8_t*)dst;
            for (size_t i = 0; i < px; ++i) { o[3 * i] = d->rgba8[4 * i]; o[3 * i + 1] = d->rgba8[4 * i + 1]; o[3 * i + 2] = d->rgba8[4 * i + 2]; }
            return VKR_SUCCESS;
        }
        default: return VKR_ERROR_INVALID_ARGUMENT;
    }
    if (bytes < n) return VKR_ERROR_INVALID_ARGUMENT;
    std::memcpy(dst, src, n);
    return VKR_SUCCESS;
}
int ork_stats(OrkDevice* d, VkrEngineStats* out) {
    if (!d || !out) return VKR_ERROR_INVALID_ARGUMENT;
    *out = d->stats;
    return VKR_SUCCESS;
}
int ork_image_level(OrkDevice* d, VkrImage image, int level, void* dst, size_t bytes, uint32_t* w, uint32_t* h) {
    if (!d || image >= d->images.size()) return VKR_ERROR_INVALID_ARGUMENT;
    const Image& im = d->images[image];
    if (level < 0) return im.levels();
    if (level >= im.levels()) return VKR_ERROR_INVALID_ARGUMENT;
    if (w) *w = (uint32_t)im.w[level];
    if (h) *h = (uint32_t)im.h[level];
    if (dst) {
        if (bytes < im.mips[level].size()) return VKR_ERROR_INVALID_ARGUMENT;
        std::memcpy(dst, im.mips[level].data(), im.mips[level].size());
    }
    return VKR_SUCCESS;
}
const char* ork_last_error(OrkDevice* d) { return d ? d->err.c_str() : "null device"; }

int ork_shade_fragment(OrkDevice* d, const VkrGPUSceneData* scene, const VkrMaterialConstants* mat, int fragShader,
                       const OrkFragIn* in, float outColor[4]) {
    if (!d || !scene || !mat || !in || !outColor) return VKR_ERROR_INVALID_ARGUMENT;
    ShaderCtx sc{d, scene};
    FragIn f;
    f.normal = {in->normal[0], in->normal[1], in->normal[2]};
    f.color = {in->color[0], in->color[1], in->color[2]};
    f.uv = {in->uv[0], in->uv[1]};
    f.lightVec = {in->lightVec[0], in->lightVec[1], in->lightVec[2]};
    f.posWorld = {in->posWorld[0], in->posWorld[1], in->posWorld[2], in->posWorld[3]};
    f.posView = {in->posView[0], in->posView[1], in->posView[2], in->posView[3]};
    f.duvdx = {in->duvdx[0], in->duvdx[1]};
    f.duvdy = {in->duvdy[0], in->duvdy[1]};
    mat4 invView = inverse(load_mat(scene->view));
    vec3 camPos = {invView.m[3][0], invView.m[3][1], invView.m[3][2]};
    vec4 col{0, 0, 0, 0};
    bool keep = fragShader == VKR_FRAG_MESH_LAMBERT ? frag_lambert(sc, *mat, f, col) : frag_pbr(sc, *mat, camPos, f, col);
    outColor[0] = col.x; outColor[1] = col.y; outColor[2] = col.z; outColor[3] = col.w;
    return keep ? 0 : 1;
}
int ork_load_shadow(OrkDevice* d, const float* depth, size_t count) {
    if (!d || !depth || count > d->shadow.size()) return VKR_ERROR_INVALID_ARGUMENT;
    std::memcpy(d->shadow.data(), depth, count * 4);
    return VKR_SUCCESS;
}
void ork_brdf(const float N[3], const float V[3], const float L[3], float roughness, const float F0[3], float* D, float* G, float F[3]) {
    vec3 n{N[0], N[1], N[2]}, v{V[0], V[1], V[2]}, l{L[0], L[1], L[2]};
    vec3 h = normalize(v + l);
    *D = DistributionGGX(n, h, roughness);
    *G = GeometrySmith(n, v, l, roughness);
    vec3 f = fresnelSchlick(std::max(dot(h, v), 0.0f), vec3{F0[0], F0[1], F0[2]});
    F[0] = f.x; F[1] = f.y; F[2] = f.z;
}
void ork_tonemap(const float in[3], float out[3]) {
    vec3 m = tonemap({in[0], in[1], in[2]});
    out[0] = m.x; out[1] = m.y; out[2] = m.z;
}
void ork_gradient(const VkrComputePushConstants* push, int y, int height, float out[4]) {
    vec4 c = gradient_color(push->data1, push->data2, y, height);
    out[0] = c.x; out[1] = c.y; out[2] = c.z; out[3] = c.w;
}
void ork_vertex(const VkrGPUSceneData* scene, const float world[16], const VkrVertex* v, float out[23]) {
    mat4 w = load_mat(world);
    VSOut o = vertex_main(*v, w, mul(load_mat(scene->viewproj), w), *scene);
    float r[23] = {o.clip.x, o.clip.y, o.clip.z, o.clip.w, o.normal.x, o.normal.y, o.normal.z, o.color.x, o.color.y, o.color.z,
                   o.uv.x, o.uv.y, o.lightVec.x, o.lightVec.y, o.lightVec.z, o.posWorld.x, o.posWorld.y, o.posWorld.z, o.posWorld.w,
                   o.posView.x, o.posView.y, o.posView.z, o.posView.w};
    std::memcpy(out, r, sizeof r);
}
void ork_shadow_vertex(const VkrGPUSceneData* scene, const float world[16], int cascade, const VkrVertex* v, float out[4]) {
    mat4 M = mul(load_mat(scene->lightViewproj[cascade]), load_mat(world));
    vec4 c = mul_point_fma(M, v->position[0], v->position[1], v->position[2]);
    out[0] = c.x; out[1] = c.y; out[2] = c.z; out[3] = c.w;
}
void ork_skybox_dir(const VkrGPUSceneData* scene, int width, int height, float sx, float sy, float out[3]) {
    vec3 p = skybox_dir(load_mat(scene->view), load_mat(scene->proj), (float)width * 0.5f, (float)height * 0.5f, sx, sy);
    out[0] = p.x; out[1] = p.y; out[2] = p.z;
}
void ork_cb_tex2d(void* ctx, int id, float u, float v, float out[4]) {
    const OrkTexCtx* c = (const OrkTexCtx*)ctx;
    vec4 t = texture2D(c->dev, id, {u, v}, {c->duvdx[0], c->duvdx[1]}, {c->duvdy[0], c->duvdy[1]});
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
}
float ork_cb_shadow(void* ctx, float u, float v, int layer) {
    return texture_shadow(((const OrkTexCtx*)ctx)->dev, u, v, layer);
}
void ork_cb_cube(void* ctx, const float dir[3], float out[4]) {
    vec3 r = {dir[0], dir[1], dir[2]};
    vec4 t = texture_cube(((const OrkTexCtx*)ctx)->dev, r, r, r);
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
}

}  // extern "C"
