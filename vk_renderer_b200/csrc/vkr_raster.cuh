// vkr_raster.cuh — vertex transform, clipping, triangle setup, coverage and depth.
//
// Replaces (for the geometry half of the path): shadow.vert + mesh.vert's gl_Position
// (shaders/shadow.vert:6-9, shaders/mesh.vert:16) and the Vulkan fixed-function clip /
// viewport / rasterisation / depth-test stages configured at
// src/vk_engine_init.cpp:439-447,532-547 and src/vk_engine_run.cpp:257-269,348-360.
//
// Arithmetic contract (DESIGN.md "Position path"): every float operation on the position
// path is an explicitly rounded single operation (no implicit FMA contraction), in a fixed
// association order; snapped coordinates are 24.8 fixed point; coverage is decided by
// exact 64-bit integer edge functions with the top-left rule; depth is a plane evaluated
// with two explicit FMAs.  This is what makes coverage and depth-test outcomes bit-exact
// against the CPU oracle.
#pragma once
#include "vkr_internal.h"
#include <math.h>

namespace vkr {

// ---- explicitly rounded float ops (device: intrinsics that the compiler may not contract;
//      host emulation build: plain ops, the TU is compiled with -ffp-contract=off) ----------
#if defined(__CUDA_ARCH__)
VKR_HD float xmul(float a, float b) { return __fmul_rn(a, b); }
VKR_HD float xadd(float a, float b) { return __fadd_rn(a, b); }
VKR_HD float xsub(float a, float b) { return __fsub_rn(a, b); }
VKR_HD float xdiv(float a, float b) { return __fdiv_rn(a, b); }
VKR_HD float xfma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
VKR_HD float xi2f(int32_t i) { return __int2float_rn(i); }
VKR_HD float xll2f(int64_t i) { return __ll2float_rn(i); }
#else
VKR_HD float xmul(float a, float b) { return a * b; }
VKR_HD float xadd(float a, float b) { return a + b; }
VKR_HD float xsub(float a, float b) { return a - b; }
VKR_HD float xdiv(float a, float b) { return a / b; }
VKR_HD float xfma(float a, float b, float c) { return fmaf(a, b, c); }
VKR_HD float xi2f(int32_t i) { return (float)i; }
VKR_HD float xll2f(int64_t i) { return (float)i; }
#endif

VKR_HD int32_t imin(int32_t a, int32_t b) { return a < b ? a : b; }
VKR_HD int32_t imax(int32_t a, int32_t b) { return a > b ? a : b; }

// clip = M * (p,1) with glm's association (m0*x + m1*y) + (m2*z + m3)
VKR_HD float4 transform_point(const float* __restrict__ m, float4 p) {
    float4 r;
    r.x = xadd(xadd(xmul(m[0], p.x), xmul(m[4], p.y)), xadd(xmul(m[8], p.z), m[12]));
    r.y = xadd(xadd(xmul(m[1], p.x), xmul(m[5], p.y)), xadd(xmul(m[9], p.z), m[13]));
    r.z = xadd(xadd(xmul(m[2], p.x), xmul(m[6], p.y)), xadd(xmul(m[10], p.z), m[14]));
    r.w = xadd(xadd(xmul(m[3], p.x), xmul(m[7], p.y)), xadd(xmul(m[11], p.z), m[15]));
    return r;
}

struct SetupTri {
    int32_t X[3], Y[3];
    float z0, dzdx, dzdy;
    int32_t bx0, by0, bx1, by1;
};

// Project + snap + cull + set up one clip-space triangle whose vertices are all inside the
// clip volume.  Returns false if it produces no fragments.
VKR_HD bool setup_triangle(const float4 c[3], const RasterTarget& rt, SetupTri& t) {
    int32_t X[3], Y[3];
    float z[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float xn = xdiv(c[i].x, c[i].w), yn = xdiv(c[i].y, c[i].w);
        z[i] = xdiv(c[i].z, c[i].w);
        float xf = xadd(xmul(xn, rt.halfW), rt.halfW);
        float yf = xadd(xmul(yn, rt.halfH), rt.halfH);
        X[i] = (int32_t)floorf(xadd(xmul(xf, 256.0f), 0.5f));
        Y[i] = (int32_t)floorf(xadd(xmul(yf, 256.0f), 0.5f));
    }
    int64_t A2 = (int64_t)(X[1] - X[0]) * (Y[2] - Y[0]) - (int64_t)(X[2] - X[0]) * (Y[1] - Y[0]);
    if (A2 == 0) return false;
    bool front = A2 < 0;                      // CCW front in Vulkan's y-down framebuffer
    if (rt.cullFront ? front : !front) return false;
    if (A2 < 0) {
        int32_t ti = X[1]; X[1] = X[2]; X[2] = ti;
        ti = Y[1]; Y[1] = Y[2]; Y[2] = ti;
        float tf = z[1]; z[1] = z[2]; z[2] = tf;
        A2 = -A2;
    }
    int32_t minX = imin(X[0], imin(X[1], X[2])), maxX = imax(X[0], imax(X[1], X[2]));
    int32_t minY = imin(Y[0], imin(Y[1], Y[2])), maxY = imax(Y[0], imax(Y[1], Y[2]));
    int32_t bx0 = imax((minX + 127) >> 8, rt.sx0), bx1 = imin((maxX - 128) >> 8, rt.sx1 - 1);
    int32_t by0 = imax((minY + 127) >> 8, rt.sy0), by1 = imin((maxY - 128) >> 8, rt.sy1 - 1);
    if (bx0 > bx1 || by0 > by1) return false;
#pragma unroll
    for (int i = 0; i < 3; ++i) { t.X[i] = X[i]; t.Y[i] = Y[i]; }
    t.bx0 = bx0; t.by0 = by0; t.bx1 = bx1; t.by1 = by1;
    float d1x = xi2f(X[1] - X[0]), d1y = xi2f(Y[1] - Y[0]);
    float d2x = xi2f(X[2] - X[0]), d2y = xi2f(Y[2] - Y[0]);
    float d1z = xsub(z[1], z[0]), d2z = xsub(z[2], z[0]);
    float fA = xll2f(A2);
    t.z0 = z[0];
    t.dzdx = xmul(xdiv(xsub(xmul(d1z, d2y), xmul(d2z, d1y)), fA), 256.0f);
    t.dzdy = xmul(xdiv(xsub(xmul(d2z, d1x), xmul(d1z, d2x)), fA), 256.0f);
    return true;
}

// Edge functions E_i(P) = a_i*Px + b_i*Py + c_i in sub-pixel units; inside iff all >= 0
// (the top-left rule is folded into c_i).
struct EdgeEq { int64_t a[3], b[3], c[3]; };

VKR_HD void make_edges(const int32_t X[3], const int32_t Y[3], EdgeEq& e) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3, k = (i + 2) % 3;
        int64_t dx = (int64_t)X[k] - X[j], dy = (int64_t)Y[k] - Y[j];
        e.a[i] = -dy;
        e.b[i] = dx;
        int64_t c = dy * X[j] - dx * Y[j];
        bool topleft = (dy < 0) || (dy == 0 && dx > 0);
        e.c[i] = topleft ? c : c - 1;
    }
}

VKR_HD float depth_at(float z0, float dzdx, float dzdy, int32_t X0, int32_t Y0, int32_t px, int32_t py) {
    float fx = xmul(xi2f(px * 256 + 128 - X0), 1.0f / 256.0f);
    float fy = xmul(xi2f(py * 256 + 128 - Y0), 1.0f / 256.0f);
    float z = xfma(dzdy, fy, xfma(dzdx, fx, z0));
    if (!(z > 0.0f)) z = 0.0f;
    if (z > 1.0f) z = 1.0f;
    return z;
}

// ---- fragment sinks: the only place the two geometry passes differ ---------------------
struct VisSink {           // main pass: 64-bit key into the visibility buffer
    uint64_t* vis; int32_t width; uint32_t lo;
    VKR_HD void put(int32_t px, int32_t py, float z) const {
        uint64_t key = ((uint64_t)__float_as_uint_hd(z) << 32) | lo;
        uint64_t* p = vis + (size_t)py * width + px;
#if defined(__CUDA_ARCH__)
        if (key < __ldcg((const unsigned long long*)p)) atomicMin((unsigned long long*)p, (unsigned long long)key);
#else
        if (key < *p) *p = key;
#endif
    }
    static VKR_HD uint32_t __float_as_uint_hd(float f) {
#if defined(__CUDA_ARCH__)
        return __float_as_uint(f);
#else
        uint32_t u; __builtin_memcpy(&u, &f, 4); return u;
#endif
    }
};
struct DepthSink {         // shadow pass: float bits, LESS_OR_EQUAL + write == min
    uint32_t* bits; int32_t width;
    VKR_HD void put(int32_t px, int32_t py, float z) const {
        uint32_t v = VisSink::__float_as_uint_hd(z);
        uint32_t* p = bits + (size_t)py * width + px;
#if defined(__CUDA_ARCH__)
        if (v < __ldcg(p)) atomicMin(p, v);
#else
        if (v < *p) *p = v;
#endif
    }
};

// Rasterise the pixel rectangle [x0,x1]x[y0,y1] of a set-up triangle; `lane`/`nlanes` split the
// rectangle's pixels between cooperating threads (1 thread: lane 0 of 1).
template <class Sink>
VKR_HD void raster_rect(const int32_t X[3], const int32_t Y[3], float z0, float dzdx, float dzdy,
                        int32_t x0, int32_t y0, int32_t x1, int32_t y1, const Sink& sink,
                        int lane = 0, int nlanes = 1) {
    EdgeEq e;
    make_edges(X, Y, e);
    const int32_t w = x1 - x0 + 1, h = y1 - y0 + 1;
    if (nlanes == 1) {
        // incremental evaluation along the row
        int64_t Px = (int64_t)x0 * 256 + 128, Py = (int64_t)y0 * 256 + 128;
        int64_t r0 = e.a[0] * Px + e.b[0] * Py + e.c[0];
        int64_t r1 = e.a[1] * Px + e.b[1] * Py + e.c[1];
        int64_t r2 = e.a[2] * Px + e.b[2] * Py + e.c[2];
        const int64_t ax0 = e.a[0] * 256, ax1 = e.a[1] * 256, ax2 = e.a[2] * 256;
        const int64_t by0 = e.b[0] * 256, by1 = e.b[1] * 256, by2 = e.b[2] * 256;
        for (int32_t py = y0; py <= y1; ++py) {
            int64_t f0 = r0, f1 = r1, f2 = r2;
            for (int32_t px = x0; px <= x1; ++px) {
                if ((f0 | f1 | f2) >= 0) sink.put(px, py, depth_at(z0, dzdx, dzdy, X[0], Y[0], px, py));
                f0 += ax0; f1 += ax1; f2 += ax2;
            }
            r0 += by0; r1 += by1; r2 += by2;
        }
    } else {
        const int32_t n = w * h;
        for (int32_t i = lane; i < n; i += nlanes) {
            int32_t py = y0 + i / w, px = x0 + i % w;
            int64_t Px = (int64_t)px * 256 + 128, Py = (int64_t)py * 256 + 128;
            int64_t f0 = e.a[0] * Px + e.b[0] * Py + e.c[0];
            int64_t f1 = e.a[1] * Px + e.b[1] * Py + e.c[1];
            int64_t f2 = e.a[2] * Px + e.b[2] * Py + e.c[2];
            if ((f0 | f1 | f2) >= 0) sink.put(px, py, depth_at(z0, dzdx, dzdy, X[0], Y[0], px, py));
        }
    }
}

// Conservative tile-vs-triangle test: false if the tile's pixel centres are all outside one edge.
VKR_HD bool tile_overlaps(const EdgeEq& e, int32_t x0, int32_t y0, int32_t x1, int32_t y1) {
    int64_t Px0 = (int64_t)x0 * 256 + 128, Px1 = (int64_t)x1 * 256 + 128;
    int64_t Py0 = (int64_t)y0 * 256 + 128, Py1 = (int64_t)y1 * 256 + 128;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        int64_t v = e.a[i] * (e.a[i] >= 0 ? Px1 : Px0) + e.b[i] * (e.b[i] >= 0 ? Py1 : Py0) + e.c[i];
        if (v < 0) return false;
    }
    return true;
}

// ---- queue push (large triangles) ---------------------------------------------------
VKR_HD uint32_t queue_alloc(uint32_t* counter, uint32_t n) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(counter, n);
#else
    uint32_t r = *counter; *counter += n; return r;
#endif
}

template <class Sink>
VKR_HD void emit_setup(const SetupTri& t, uint32_t seq, const RasterTarget& rt, const Queues& q, const Sink& sink) {
    const int32_t w = t.bx1 - t.bx0 + 1, h = t.by1 - t.by0 + 1;
    if (w <= SMALL_MAX && h <= SMALL_MAX) {
        raster_rect(t.X, t.Y, t.z0, t.dzdx, t.dzdy, t.bx0, t.by0, t.bx1, t.by1, sink);
        return;
    }
    uint32_t slot = queue_alloc(&q.counters[0], 1);
    if (slot >= q.triCap) { q.counters[2] = 1; return; }
    LargeTri lt;
#pragma unroll
    for (int i = 0; i < 3; ++i) { lt.X[i] = t.X[i]; lt.Y[i] = t.Y[i]; }
    lt.z0 = t.z0; lt.dzdx = t.dzdx; lt.dzdy = t.dzdy;
    lt.bx0 = (int16_t)t.bx0; lt.by0 = (int16_t)t.by0; lt.bx1 = (int16_t)t.bx1; lt.by1 = (int16_t)t.by1;
    lt.seq = seq;
    q.tris[slot] = lt;
    EdgeEq e;
    make_edges(t.X, t.Y, e);
    const int32_t tx0 = t.bx0 / TILE, tx1 = t.bx1 / TILE, ty0 = t.by0 / TILE, ty1 = t.by1 / TILE;
    for (int32_t ty = ty0; ty <= ty1; ++ty)
        for (int32_t tx = tx0; tx <= tx1; ++tx) {
            int32_t x0 = imax(tx * TILE, t.bx0), x1 = imin(tx * TILE + TILE - 1, t.bx1);
            int32_t y0 = imax(ty * TILE, t.by0), y1 = imin(ty * TILE + TILE - 1, t.by1);
            if (!tile_overlaps(e, x0, y0, x1, y1)) continue;
            uint32_t is = queue_alloc(&q.counters[1], 1);
            if (is >= q.itemCap) { q.counters[2] = 1; return; }
            LargeItem it; it.tri = slot; it.tx = (uint16_t)tx; it.ty = (uint16_t)ty;
            q.items[is] = it;
        }
}

// ---- clipping (slow path) -------------------------------------------------------------
VKR_HD float plane_dist(int p, float4 v) {
    switch (p) {
        case 0: return v.z;
        case 1: return xsub(v.w, v.z);
        case 2: return xadd(v.x, xmul(GUARD_BAND, v.w));
        case 3: return xsub(xmul(GUARD_BAND, v.w), v.x);
        case 4: return xadd(v.y, xmul(GUARD_BAND, v.w));
        default: return xsub(xmul(GUARD_BAND, v.w), v.y);
    }
}
VKR_HD float4 clip_lerp(float4 in, float4 out, float din, float dout) {
    float t = xdiv(din, xsub(din, dout));
    float4 r;
    r.x = xadd(in.x, xmul(t, xsub(out.x, in.x)));
    r.y = xadd(in.y, xmul(t, xsub(out.y, in.y)));
    r.z = xadd(in.z, xmul(t, xsub(out.z, in.z)));
    r.w = xadd(in.w, xmul(t, xsub(out.w, in.w)));
    return r;
}

template <class Sink>
VKR_HDN void clip_and_emit(const float4 cin[3], uint32_t seq, const RasterTarget& rt, const Queues& q, const Sink& sink) {
    float4 bufA[12], bufB[12];
    int n = 3;
    bufA[0] = cin[0]; bufA[1] = cin[1]; bufA[2] = cin[2];
    float4 *src = bufA, *dst = bufB;
    for (int p = 0; p < 6 && n >= 3; ++p) {
        int m = 0;
        for (int i = 0; i < n; ++i) {
            float4 a = src[i], b = src[(i + 1) % n];
            float da = plane_dist(p, a), db = plane_dist(p, b);
            bool ia = da >= 0.0f, ib = db >= 0.0f;
            if (ia) dst[m++] = a;
            if (ia != ib) dst[m++] = ia ? clip_lerp(a, b, da, db) : clip_lerp(b, a, db, da);
        }
        n = m;
        float4* tmp = src; src = dst; dst = tmp;
    }
    if (n < 3) return;
    for (int i = 1; i + 1 < n; ++i) {
        float4 c[3] = {src[0], src[i], src[i + 1]};
        if (!(c[0].w > 0.0f && c[1].w > 0.0f && c[2].w > 0.0f)) continue;
        SetupTri t;
        if (setup_triangle(c, rt, t)) emit_setup(t, seq, rt, q, sink);
    }
}

// One triangle of one draw: fetch, transform, reject / clip, set up, rasterise or enqueue.
template <class Sink>
VKR_HD void geometry_triangle(const float* __restrict__ mvp, const float4* __restrict__ pos,
                              const uint32_t* __restrict__ idx, uint32_t tri, uint32_t seq,
                              const RasterTarget& rt, const Queues& q, const Sink& sink) {
    float4 c[3];
#if defined(__CUDA_ARCH__)
    const uint32_t i0 = __ldg(idx + 3 * tri), i1 = __ldg(idx + 3 * tri + 1), i2 = __ldg(idx + 3 * tri + 2);
    c[0] = transform_point(mvp, __ldg(pos + i0));
    c[1] = transform_point(mvp, __ldg(pos + i1));
    c[2] = transform_point(mvp, __ldg(pos + i2));
#else
    c[0] = transform_point(mvp, pos[idx[3 * tri]]);
    c[1] = transform_point(mvp, pos[idx[3 * tri + 1]]);
    c[2] = transform_point(mvp, pos[idx[3 * tri + 2]]);
#endif
    // trivial reject against the view volume
    if ((c[0].z < 0 && c[1].z < 0 && c[2].z < 0) ||
        (c[0].z > c[0].w && c[1].z > c[1].w && c[2].z > c[2].w) ||
        (c[0].x < -c[0].w && c[1].x < -c[1].w && c[2].x < -c[2].w) ||
        (c[0].x > c[0].w && c[1].x > c[1].w && c[2].x > c[2].w) ||
        (c[0].y < -c[0].w && c[1].y < -c[1].w && c[2].y < -c[2].w) ||
        (c[0].y > c[0].w && c[1].y > c[1].w && c[2].y > c[2].w))
        return;
    bool needClip = false;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float gw = xmul(GUARD_BAND, c[i].w);
        bool inside = (c[i].z >= 0.0f) && (xsub(c[i].w, c[i].z) >= 0.0f) && (xadd(c[i].x, gw) >= 0.0f) &&
                      (xsub(gw, c[i].x) >= 0.0f) && (xadd(c[i].y, gw) >= 0.0f) && (xsub(gw, c[i].y) >= 0.0f);
        needClip |= !inside;
    }
    if (needClip) { clip_and_emit(c, seq, rt, q, sink); return; }
    SetupTri t;
    if (setup_triangle(c, rt, t)) emit_setup(t, seq, rt, q, sink);
}

// One (large triangle, tile) work item; lanes of a warp split the tile's pixels.
template <class Sink>
VKR_HD void raster_item(const LargeTri& lt, const LargeItem& it, const Sink& sink, int lane, int nlanes) {
    int32_t x0 = imax((int32_t)it.tx * TILE, (int32_t)lt.bx0), x1 = imin((int32_t)it.tx * TILE + TILE - 1, (int32_t)lt.bx1);
    int32_t y0 = imax((int32_t)it.ty * TILE, (int32_t)lt.by0), y1 = imin((int32_t)it.ty * TILE + TILE - 1, (int32_t)lt.by1);
    raster_rect(lt.X, lt.Y, lt.z0, lt.dzdx, lt.dzdy, x0, y0, x1, y1, sink, lane, nlanes);
}

}  // namespace vkr
