// vkr_raster.cuh — vertex transform, clipping, triangle setup, coverage and depth.
//
// Replaces (for the geometry half of the path): shadow.vert + mesh.vert's gl_Position
// (shaders/shadow.vert:6-9, shaders/mesh.vert:16) and the Vulkan fixed-function clip /
// viewport / rasterisation / depth-test stages configured at
// src/vk_engine_init.cpp:439-447,532-547 and src/vk_engine_run.cpp:257-269,348-360.
//
// Arithmetic contract (DESIGN.md "Position path"): every float operation on the position
// path is an explicitly rounded single operation (mul, add, sub, rcp or FMA where the contract
// says FMA — never an implicit contraction), in a fixed association order; snapped
// coordinates are 24.8 fixed point; coverage is decided by exact integer edge functions with
// the top-left rule; depth is a plane evaluated with two explicit FMAs.  This is what makes
// coverage and depth-test outcomes bit-exact against the CPU oracle.
#pragma once
#include "vkr_internal.h"
#include <math.h>

namespace vkr {

// ---- explicitly rounded float ops (device: intrinsics the compiler may not contract;
//      host emulation build: plain ops, the TU is compiled with -ffp-contract=off) ----------
#if defined(__CUDA_ARCH__)
VKR_HD float xmul(float a, float b) { return __fmul_rn(a, b); }
VKR_HD float xadd(float a, float b) { return __fadd_rn(a, b); }
VKR_HD float xsub(float a, float b) { return __fsub_rn(a, b); }
VKR_HD float xrcp(float a) { return __frcp_rn(a); }
VKR_HD float xdiv(float a, float b) { return __fdiv_rn(a, b); }
VKR_HD float xfma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
VKR_HD float xi2f(int32_t i) { return __int2float_rn(i); }
VKR_HD float xll2f(int64_t i) { return __ll2float_rn(i); }
VKR_HD int32_t xf2i_rn(float f) { return __float2int_rn(f); }
VKR_HD uint32_t f2bits(float f) { return __float_as_uint(f); }
VKR_HD float bits2f(uint32_t u) { return __uint_as_float(u); }
#else
VKR_HD float xmul(float a, float b) { return a * b; }
VKR_HD float xadd(float a, float b) { return a + b; }
VKR_HD float xsub(float a, float b) { return a - b; }
VKR_HD float xrcp(float a) { return 1.0f / a; }
VKR_HD float xdiv(float a, float b) { return a / b; }
VKR_HD float xfma(float a, float b, float c) { return fmaf(a, b, c); }
VKR_HD float xi2f(int32_t i) { return (float)i; }
VKR_HD float xll2f(int64_t i) { return (float)i; }
VKR_HD int32_t xf2i_rn(float f) { return (int32_t)nearbyintf(f); }
VKR_HD uint32_t f2bits(float f) { uint32_t u; __builtin_memcpy(&u, &f, 4); return u; }
VKR_HD float bits2f(uint32_t u) { float f; __builtin_memcpy(&f, &u, 4); return f; }
#endif

VKR_HD int32_t imin(int32_t a, int32_t b) { return a < b ? a : b; }
VKR_HD int32_t imax(int32_t a, int32_t b) { return a > b ? a : b; }
VKR_HD int32_t iabs(int32_t a) { return a < 0 ? -a : a; }

// clip = M * (p,1): per component fma(m0,x, fma(m1,y, fma(m2,z, m3)))
VKR_HD float4 transform_point(const float* __restrict__ m, float x, float y, float z) {
    float4 r;
    r.x = xfma(m[0], x, xfma(m[4], y, xfma(m[8], z, m[12])));
    r.y = xfma(m[1], x, xfma(m[5], y, xfma(m[9], z, m[13])));
    r.z = xfma(m[2], x, xfma(m[6], y, xfma(m[10], z, m[14])));
    r.w = xfma(m[3], x, xfma(m[7], y, xfma(m[11], z, m[15])));
    return r;
}

// Inside the clip volume used for rasterisation: 0 <= z <= w and |x|,|y| <= GUARD_BAND*w.
VKR_HD bool inside_clip_volume(float4 c) {
    const float gw = xmul(GUARD_BAND, c.w);
    return (c.z >= 0.0f) && (xsub(c.w, c.z) >= 0.0f) && (xadd(c.x, gw) >= 0.0f) && (xsub(gw, c.x) >= 0.0f) &&
           (xadd(c.y, gw) >= 0.0f) && (xsub(gw, c.y) >= 0.0f);
}

// Perspective divide + viewport + snap to 24.8 (round-half-even).
struct SnapVtx { int32_t X, Y; float z; };
VKR_HD SnapVtx project_vertex(float4 c, const RasterTarget& rt) {
    SnapVtx v;
    const float rw = xrcp(c.w);
    const float xn = xmul(c.x, rw), yn = xmul(c.y, rw);
    v.z = xmul(c.z, rw);
    v.X = xf2i_rn(xfma(xn, rt.halfW256, rt.halfW256));
    v.Y = xf2i_rn(xfma(yn, rt.halfH256, rt.halfH256));
    return v;
}

struct SetupTri {
    int32_t X[3], Y[3];
    float z0, dzdx, dzdy;
    int32_t bx0, by0, bx1, by1;
    int32_t ext;              // max extent of the (unclipped-by-scissor) triangle in sub-pixels
};

// Cull + orient + bound + depth plane of three snapped vertices.  False: no fragments.
VKR_HD bool setup_snapped(SnapVtx v0, SnapVtx v1, SnapVtx v2, const RasterTarget& rt, SetupTri& t) {
    const int32_t d1x = v1.X - v0.X, d1y = v1.Y - v0.Y, d2x = v2.X - v0.X, d2y = v2.Y - v0.Y;
    int64_t A2;
    const bool narrow = (iabs(d1x) | iabs(d1y) | iabs(d2x) | iabs(d2y)) < 32768;      // the doubled area fits int32 (every small triangle)
    if (narrow) A2 = (int64_t)(d1x * d2y - d2x * d1y);
    else A2 = (int64_t)d1x * d2y - (int64_t)d2x * d1y;
    if (A2 == 0) return false;
    const bool front = A2 < 0;                 // CCW front in Vulkan's y-down framebuffer (a = -A2/2 > 0)
    if (rt.cullFront ? front : !front) return false;
    if (A2 < 0) { SnapVtx tmp = v1; v1 = v2; v2 = tmp; A2 = -A2; }
    const int32_t minX = imin(v0.X, imin(v1.X, v2.X)), maxX = imax(v0.X, imax(v1.X, v2.X));
    const int32_t minY = imin(v0.Y, imin(v1.Y, v2.Y)), maxY = imax(v0.Y, imax(v1.Y, v2.Y));
    // pixel centres sit at i*256+128
    const int32_t bx0 = imax((minX + 127) >> 8, rt.sx0), bx1 = imin((maxX - 128) >> 8, rt.sx1 - 1);
    const int32_t by0 = imax((minY + 127) >> 8, rt.sy0), by1 = imin((maxY - 128) >> 8, rt.sy1 - 1);
    if (bx0 > bx1 || by0 > by1) return false;
    t.X[0] = v0.X; t.X[1] = v1.X; t.X[2] = v2.X;
    t.Y[0] = v0.Y; t.Y[1] = v1.Y; t.Y[2] = v2.Y;
    t.bx0 = bx0; t.by0 = by0; t.bx1 = bx1; t.by1 = by1;
    t.ext = imax(maxX - minX, maxY - minY);
    // depth plane about vertex 0: z(x,y) = z0 + dzdx*(x-x0) + dzdy*(y-y0), x,y in pixels
    const float e1x = xi2f(v1.X - v0.X), e1y = xi2f(v1.Y - v0.Y), e2x = xi2f(v2.X - v0.X), e2y = xi2f(v2.Y - v0.Y);
    const float d1z = xsub(v1.z, v0.z), d2z = xsub(v2.z, v0.z);
    const float rA = xrcp(narrow ? xi2f((int32_t)A2) : xll2f(A2));      // (the same value: an int32 converts to the same float either way)
    t.z0 = v0.z;
    t.dzdx = xmul(xmul(xsub(xmul(d1z, e2y), xmul(d2z, e1y)), rA), 256.0f);
    t.dzdy = xmul(xmul(xsub(xmul(d2z, e1x), xmul(d1z, e2x)), rA), 256.0f);
    return true;
}

VKR_HD float depth_at(float z0, float dzdx, float dzdy, int32_t X0, int32_t Y0, int32_t px, int32_t py) {
    const float fx = xmul(xi2f(px * 256 + 128 - X0), 1.0f / 256.0f);
    const float fy = xmul(xi2f(py * 256 + 128 - Y0), 1.0f / 256.0f);
    float z = xfma(dzdy, fy, xfma(dzdx, fx, z0));
    if (!(z > 0.0f)) z = 0.0f;      // depth is clamped to [0,1] (also keeps the bit pattern ordered)
    if (z > 1.0f) z = 1.0f;
    return z;
}

// ---- fragment sinks: the only place the two geometry passes differ ---------------------
struct VisSink {           // main pass: 64-bit key into the visibility buffer
    uint64_t* vis; int32_t width; uint32_t lo;
    // Early depth test for fragments that are expensive to produce (alpha-tested): true iff the stored depth is already
    // nearer than z.  Stored depths only decrease, so such a fragment can never win (LEQUAL: equal depth still can).
    VKR_HD bool hidden(int32_t px, int32_t py, float z) const {
        const uint32_t* hi = reinterpret_cast<const uint32_t*>(vis + (uint32_t)(py * width + px)) + 1;     // (W * H < 2^32)
#if defined(__CUDA_ARCH__)
        return __ldcg(hi) < f2bits(z);
#else
        return *hi < f2bits(z);
#endif
    }
    VKR_HD void put(int32_t px, int32_t py, float z) const {
        const uint64_t key = ((uint64_t)f2bits(z) << 32) | lo;
        uint64_t* p = vis + (uint32_t)(py * width + px);
#if defined(__CUDA_ARCH__)
        atomicMin((unsigned long long*)p, (unsigned long long)key);   // result unused -> RED.MIN.64: fire and forget, no L2 round trip
#else
        if (key < *p) *p = key;
#endif
    }
};
struct DepthSink {         // shadow pass: float bits, LESS_OR_EQUAL + write == min
    uint32_t* bits; int32_t width;
    VKR_HD void put(int32_t px, int32_t py, float z) const {
        const uint32_t v = f2bits(z);
        uint32_t* p = bits + (uint32_t)(py * width + px);
#if defined(__CUDA_ARCH__)
        atomicMin(p, v);                                                // RED.MIN.32
#else
        if (v < *p) *p = v;
#endif
    }
};

// ---- coverage ------------------------------------------------------------------------------
// Edge i runs from vertex j=(i+1)%3 to k=(i+2)%3; with doubled area > 0 the interior has
// cross(k-j, P-j) > 0.  Top-left rule (y down): an edge owns its boundary samples iff it is a
// left edge (dy < 0) or a top edge (dy == 0 && dx > 0); otherwise the edge value must be >= 1.

// Small triangles (extent < 2^14 sub-pixels, bounding box <= SMALL_MAX x SMALL_MAX samples): everything fits 32-bit integers.
// A set-up small triangle is a 64-byte record; its bounding box holds w*h CANDIDATE samples, numbered row by row, and
// raster_candidate() decides and emits one of them.  The geometry kernel parks the records of a warp's 32 triangles in
// shared memory and deals the candidates of all of them to the 32 lanes (geometry_warp: prefix sum of the counts, owner
// search), so a lane with a 1-sample triangle does not wait for its neighbour's 4x4 box; the host emulation walks them in
// order.  Both produce the values of the plain per-triangle loop: the edge functions at sample (ix, iy) of the box are
// e + ix*ax + iy*by exactly (32-bit, no overflow: |e| < 2^30), and the depth is depth_at() of the same integers.
// w | ceil(2^16 / w) << 8 for the box widths 1..SMALL_MAX: k / w == (k * (that >> 8)) >> 16 for k < 256 (a table: the integer
// division costs the set-up ~25 instructions)
struct WdivTable {
    uint32_t v[SMALL_MAX + 1];
    constexpr WdivTable() : v() { for (uint32_t w = 1; w <= (uint32_t)SMALL_MAX; ++w) v[w] = w | (((65536u + w - 1u) / w) << 8); }
};
#if defined(__CUDACC__)
__device__ constexpr WdivTable kWdivDev{};
#endif
constexpr WdivTable kWdivHost{};

struct alignas(16) FragRec {
    int32_t e0, e1, e2;           // edge functions at the box's first sample, top-left bias folded in (covered iff all >= 0)
    int32_t ax0, ax1, ax2;        // their steps per pixel in x ...
    int32_t by0, by1, by2;        // ... and in y
    float z0, dzdx, dzdy;         // depth plane about vertex 0
    float fx0, fy0;               // (first sample - vertex 0) in pixels: a multiple of 1/256 below 2^7, exact in f32
    uint32_t origin;              // first sample of the box: x | y << 16
    uint32_t wdiv;                // box width w | ceil(2^16 / w) << 8   (k / w == (k * that) >> 16 for k < 256)
};
static_assert(sizeof(FragRec) == 64, "FragRec is read as four 16-byte words");

// The record of a small triangle; returns its candidate count (w * h <= SMALL_MAX^2).
VKR_HD uint32_t make_frag_rec(const SetupTri& t, FragRec& r) {
    const int32_t Px = t.bx0 * 256 + 128, Py = t.by0 * 256 + 128;
    int32_t e[3], ax[3], by[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3, k = (i + 2) % 3;
        const int32_t dx = t.X[k] - t.X[j], dy = t.Y[k] - t.Y[j];
        const bool topleft = (dy < 0) || (dy == 0 && dx > 0);
        e[i] = dx * (Py - t.Y[j]) - dy * (Px - t.X[j]) - (topleft ? 0 : 1);
        ax[i] = -dy * 256;
        by[i] = dx * 256;
    }
    r.e0 = e[0]; r.e1 = e[1]; r.e2 = e[2];
    r.ax0 = ax[0]; r.ax1 = ax[1]; r.ax2 = ax[2];
    r.by0 = by[0]; r.by1 = by[1]; r.by2 = by[2];
    const uint32_t w = (uint32_t)(t.bx1 - t.bx0 + 1), h = (uint32_t)(t.by1 - t.by0 + 1);
    r.origin = (uint32_t)t.bx0 | ((uint32_t)t.by0 << 16);
#if defined(__CUDA_ARCH__)
    r.wdiv = kWdivDev.v[w];
#else
    r.wdiv = kWdivHost.v[w];
#endif
    r.z0 = t.z0; r.dzdx = t.dzdx; r.dzdy = t.dzdy;
    // depth_at() divides (sample - vertex 0), an integer number of sub-pixels below 2^15, by 256: exact, and so is adding the
    // sample's integer offset inside the box afterwards (the sum still has at most 15 + 8 significant bits)
    r.fx0 = xmul(xi2f(Px - t.X[0]), 1.0f / 256.0f);
    r.fy0 = xmul(xi2f(Py - t.Y[0]), 1.0f / 256.0f);
    return w * h;
}

// Candidate k of the record: coverage, then depth and the sink.
template <class Sink>
VKR_HD void raster_candidate(const FragRec& r, uint32_t k, const Sink& sink) {
    const uint32_t w = r.wdiv & 0xFFu;
    const int32_t iy = (int32_t)((k * (r.wdiv >> 8)) >> 16), ix = (int32_t)k - iy * (int32_t)w;
    const int32_t f0 = r.e0 + ix * r.ax0 + iy * r.by0, f1 = r.e1 + ix * r.ax1 + iy * r.by1, f2 = r.e2 + ix * r.ax2 + iy * r.by2;
    if ((f0 | f1 | f2) < 0) return;
    // depth_at(): fx = (sample - vertex 0) / 256 = fx0 + ix, both exact (see make_frag_rec)
    const float fx = xadd(r.fx0, xi2f(ix)), fy = xadd(r.fy0, xi2f(iy));
    float z = xfma(r.dzdy, fy, xfma(r.dzdx, fx, r.z0));
    if (!(z > 0.0f)) z = 0.0f;
    if (z > 1.0f) z = 1.0f;
    sink.put((int32_t)(r.origin & 0xFFFFu) + ix, (int32_t)(r.origin >> 16) + iy, z);
}

// The plain per-triangle double loop (incremental edge functions): every lane rasterises its own triangle and waits for the
// largest box of the warp.  This is what the geometry kernels run.  The candidate stage below (-DVKR_RASTER_CANDIDATES: the
// samples of all 32 boxes dealt evenly to the lanes through a prefix sum) keeps every lane busy but pays ~90 instructions
// per round of 32 samples against ~12 per sample here, and measured slower on BASELINE config 4 (profiles/r02d_ab.md).
template <class Sink>
VKR_HD void raster_small_loop(const SetupTri& t, const Sink& sink) {
    FragRec r;
    make_frag_rec(t, r);
    int32_t e0 = r.e0, e1 = r.e1, e2 = r.e2;
    for (int32_t py = t.by0; py <= t.by1; ++py) {
        int32_t f0 = e0, f1 = e1, f2 = e2;
        for (int32_t px = t.bx0; px <= t.bx1; ++px) {
            if ((f0 | f1 | f2) >= 0) sink.put(px, py, depth_at(t.z0, t.dzdx, t.dzdy, t.X[0], t.Y[0], px, py));
            f0 += r.ax0; f1 += r.ax1; f2 += r.ax2;
        }
        e0 += r.by0; e1 += r.by1; e2 += r.by2;
    }
}

// One lane (or the host emulation) walking every candidate of its own triangle.
template <class Sink>
VKR_HD void raster_small(const SetupTri& t, const Sink& sink) {
    FragRec r;
    const uint32_t n = make_frag_rec(t, r);
    for (uint32_t k = 0; k < n; ++k) raster_candidate(r, k, sink);
}

// General case: 64-bit edge functions E_i(P) = a_i*Px + b_i*Py + c_i.
struct EdgeEq { int64_t a[3], b[3], c[3]; };
VKR_HD void make_edges(const int32_t X[3], const int32_t Y[3], EdgeEq& e) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3, k = (i + 2) % 3;
        const int64_t dx = (int64_t)X[k] - X[j], dy = (int64_t)Y[k] - Y[j];
        e.a[i] = -dy;
        e.b[i] = dx;
        const int64_t c = dy * X[j] - dx * Y[j];
        const bool topleft = (dy < 0) || (dy == 0 && dx > 0);
        e.c[i] = topleft ? c : c - 1;
    }
}

// Coverage of the strip pxa..pxb (at most 8 pixels) of row py.  The edge functions are linear along the strip, so their
// values at its two ends bound them everywhere in between: both ends inside every edge -> the whole strip is covered, both
// ends outside some edge -> none of it is; only a strip an edge actually crosses is walked pixel by pixel.  Exact (the same
// integers, the same >= 0 test); interior strips of a big triangle — most of a tile item — cost three multiply-adds instead
// of eight 64-bit steps.  Measured neutral to slightly slower (configs 1 / 2 / 3 / 4: 20 140 / 8202 / 949 / 2532 frames/s with,
// 20 345 / 8271 / 956 / 2529 without; profiles/r02_ab.md): the large-triangle kernel does not wait for its coverage
// arithmetic.  Off; -DVKR_STRIP_BOUNDS switches it on.
VKR_HD uint32_t strip_coverage(const EdgeEq& e, int64_t ax0, int64_t ax1, int64_t ax2, int32_t pxa, int32_t pxb, int32_t py) {
    const int64_t Px = (int64_t)pxa * 256 + 128, Py = (int64_t)py * 256 + 128;
    int64_t f0 = e.a[0] * Px + e.b[0] * Py + e.c[0];
    int64_t f1 = e.a[1] * Px + e.b[1] * Py + e.c[1];
    int64_t f2 = e.a[2] * Px + e.b[2] * Py + e.c[2];
#if defined(VKR_STRIP_BOUNDS)
    const int32_t last = pxb - pxa;                       // 0..7
    const int64_t g0 = f0 + ax0 * last, g1 = f1 + ax1 * last, g2 = f2 + ax2 * last;
    if ((f0 | g0 | f1 | g1 | f2 | g2) >= 0) return (2u << last) - 1u;
    if (((f0 & g0) | (f1 & g1) | (f2 & g2)) < 0) return 0u;
#endif
    uint32_t mask = 0;
#pragma unroll
    for (int32_t k = 0; k < 8; ++k) {
        if (pxa + k <= pxb && (f0 | f1 | f2) >= 0) mask |= 1u << k;
        f0 += ax0; f1 += ax1; f2 += ax2;
    }
    return mask;
}

// Rectangle [x0,x1]x[y0,y1] of a large triangle, pixels dealt to `nlanes` cooperating lanes in
// 8-pixel-wide strips (lane -> (row, 8-px column group)), incremental along the strip.
template <class Sink>
VKR_HD void raster_rect64(const int32_t X[3], const int32_t Y[3], float z0, float dzdx, float dzdy,
                          int32_t x0, int32_t y0, int32_t x1, int32_t y1, const Sink& sink, int lane, int nlanes) {
    EdgeEq e;
    make_edges(X, Y, e);
    const int32_t w = x1 - x0 + 1, h = y1 - y0 + 1;
    const int32_t strips = (w + 7) >> 3;
    const int32_t n = strips * h;
    const int64_t ax0 = e.a[0] * 256, ax1 = e.a[1] * 256, ax2 = e.a[2] * 256;
    for (int32_t i = lane; i < n; i += nlanes) {
        const int32_t row = i / strips, sx = i - row * strips;
        const int32_t py = y0 + row, pxa = x0 + sx * 8, pxb = imin(pxa + 7, x1);
        // coverage of the strip first (cheap, every lane in step), then one sink call per covered pixel: lanes meet at
        // their k-th covered pixel, so an expensive sink (alpha test, blending) runs converged even on thin slivers
        uint32_t mask = strip_coverage(e, ax0, ax1, ax2, pxa, pxb, py);
        while (mask) {
#if defined(__CUDA_ARCH__)
            const int32_t k = __ffs((int)mask) - 1;
#else
            const int32_t k = __builtin_ctz(mask);
#endif
            mask &= mask - 1;
            sink.put(pxa + k, py, depth_at(z0, dzdx, dzdy, X[0], Y[0], pxa + k, py));
        }
    }
}

#if defined(__CUDACC__)
// The same rectangle for an EXPENSIVE sink (alpha test, shade-and-blend): one warp, covered pixels are first compacted
// into a per-warp queue in shared memory (`wq`, RECT_QUEUE entries) and the sink then runs on 32 queued pixels at a time,
// fully converged — thin slivers cover one or two pixels per row, which would otherwise leave a few lanes doing all the
// sink work while the others wait.
static const int RECT_QUEUE = 32 + 32 * 8;
template <class Sink>
__device__ __forceinline__ void raster_rect64_compact(const int32_t X[3], const int32_t Y[3], float z0, float dzdx, float dzdy,
                                                      int32_t x0, int32_t y0, int32_t x1, int32_t y1, const Sink& sink, int lane, uint32_t* wq) {
    EdgeEq e;
    make_edges(X, Y, e);
    const int32_t w = x1 - x0 + 1, h = y1 - y0 + 1;
    const int32_t strips = (w + 7) >> 3;
    const int32_t n = strips * h;
    const int64_t ax0 = e.a[0] * 256, ax1 = e.a[1] * 256, ax2 = e.a[2] * 256;
    uint32_t count = 0;                                   // queued pixels (warp-uniform)
    for (int32_t base = 0; base < n; base += 32) {
        const int32_t i = base + lane;
        uint32_t mask = 0;
        int32_t py = 0, pxa = 0;
        if (i < n) {
            const int32_t row = i / strips, sx = i - row * strips;
            py = y0 + row; pxa = x0 + sx * 8;
            mask = strip_coverage(e, ax0, ax1, ax2, pxa, imin(pxa + 7, x1), py);
        }
        if (!__any_sync(0xFFFFFFFFu, mask != 0)) continue;
        const uint32_t c = (uint32_t)__popc(mask);
        uint32_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        uint32_t off = count + incl - c;
        while (mask) {
            const int32_t k = __ffs((int)mask) - 1;
            mask &= mask - 1;
            wq[off++] = (uint32_t)(pxa + k) | ((uint32_t)py << 16);
        }
        __syncwarp();
        count += total;
        uint32_t head = 0;
        while (count - head >= 32) {
            const uint32_t q = wq[head + lane];
            const int32_t px = (int32_t)(q & 0xFFFFu), qy = (int32_t)(q >> 16);
            sink.put(px, qy, depth_at(z0, dzdx, dzdy, X[0], Y[0], px, qy));
            head += 32;
        }
        const uint32_t rem = count - head;                // < 32: move the tail to the front
        __syncwarp();
        const uint32_t v = lane < (int)rem ? wq[head + lane] : 0u;
        __syncwarp();
        if (lane < (int)rem) wq[lane] = v;
        __syncwarp();
        count = rem;
    }
    if (lane < (int)count) {
        const uint32_t q = wq[lane];
        const int32_t px = (int32_t)(q & 0xFFFFu), qy = (int32_t)(q >> 16);
        sink.put(px, qy, depth_at(z0, dzdx, dzdy, X[0], Y[0], px, qy));
    }
    __syncwarp();
}
#endif

// Conservative tile-vs-triangle test: false if the tile's pixel centres are all outside one edge.
VKR_HD bool tile_overlaps(const EdgeEq& e, int32_t x0, int32_t y0, int32_t x1, int32_t y1) {
    const int64_t Px0 = (int64_t)x0 * 256 + 128, Px1 = (int64_t)x1 * 256 + 128;
    const int64_t Py0 = (int64_t)y0 * 256 + 128, Py1 = (int64_t)y1 * 256 + 128;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int64_t v = e.a[i] * (e.a[i] >= 0 ? Px1 : Px0) + e.b[i] * (e.b[i] >= 0 ? Py1 : Py0) + e.c[i];
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

// strips: (alpha-tested triangles, whose every fragment costs a texture evaluation) items of TILE x 16 samples instead of TILE x TILE or
// the whole box — `ty` then holds ITEM_STRIP | row / 16 — so that no warp of large_raster_kernel is left alone with 4096 alpha
// tests while the others have run out of items (config 3: 15 % of that kernel's warp slots active, profiles/r02_ab.md)
#ifndef VKR_CUTOUT_STRIPS
#define VKR_CUTOUT_STRIPS 0      // measured neutral on config 3 (848 vs 849 frames/s, profiles/r02_ab.md): off
#endif
static const uint16_t ITEM_STRIP = 0x8000u;
static const int STRIP_ROWS = 16;
VKR_HDN void enqueue_large(const SetupTri& t, uint32_t seq, const Queues& q, bool strips = false) {
    const uint32_t slot = queue_alloc(&q.counters[0], 1);
    if (slot >= q.triCap) { q.counters[2] = 1; return; }
    LargeTri lt;
#pragma unroll
    for (int i = 0; i < 3; ++i) { lt.X[i] = t.X[i]; lt.Y[i] = t.Y[i]; }
    lt.z0 = t.z0; lt.dzdx = t.dzdx; lt.dzdy = t.dzdy;
    lt.bx0 = (int16_t)t.bx0; lt.by0 = (int16_t)t.by0; lt.bx1 = (int16_t)t.bx1; lt.by1 = (int16_t)t.by1;
    lt.seq = seq;
    q.tris[slot] = lt;
    if (strips) {
        EdgeEq e;
        make_edges(t.X, t.Y, e);
        const int32_t tx0 = t.bx0 / TILE, tx1 = t.bx1 / TILE, sy0 = t.by0 / STRIP_ROWS, sy1 = t.by1 / STRIP_ROWS;
        for (int32_t sy = sy0; sy <= sy1; ++sy)
            for (int32_t tx = tx0; tx <= tx1; ++tx) {
                const int32_t x0 = imax(tx * TILE, t.bx0), x1 = imin(tx * TILE + TILE - 1, t.bx1);
                const int32_t y0 = imax(sy * STRIP_ROWS, t.by0), y1 = imin(sy * STRIP_ROWS + STRIP_ROWS - 1, t.by1);
                if (!tile_overlaps(e, x0, y0, x1, y1)) continue;
                const uint32_t is = queue_alloc(&q.counters[1], 1);
                if (is >= q.itemCap) { q.counters[2] = 1; return; }
                LargeItem it; it.tri = slot; it.tx = (uint16_t)tx; it.ty = (uint16_t)(ITEM_STRIP | (uint32_t)sy);
                q.items[is] = it;
            }
        return;
    }
    if (t.ext < 16384) {          // medium: one item for the whole bounding box (at most 64 x 64 pixels), 32-bit edge functions
        const uint32_t is = queue_alloc(&q.counters[1], 1);
        if (is >= q.itemCap) { q.counters[2] = 1; return; }
        LargeItem it; it.tri = slot; it.tx = ITEM_WHOLE; it.ty = ITEM_WHOLE;
        q.items[is] = it;
        return;
    }
    EdgeEq e;
    make_edges(t.X, t.Y, e);
    const int32_t tx0 = t.bx0 / TILE, tx1 = t.bx1 / TILE, ty0 = t.by0 / TILE, ty1 = t.by1 / TILE;
    for (int32_t ty = ty0; ty <= ty1; ++ty)
        for (int32_t tx = tx0; tx <= tx1; ++tx) {
            const int32_t x0 = imax(tx * TILE, t.bx0), x1 = imin(tx * TILE + TILE - 1, t.bx1);
            const int32_t y0 = imax(ty * TILE, t.by0), y1 = imin(ty * TILE + TILE - 1, t.by1);
            if (!tile_overlaps(e, x0, y0, x1, y1)) continue;
            const uint32_t is = queue_alloc(&q.counters[1], 1);
            if (is >= q.itemCap) { q.counters[2] = 1; return; }
            LargeItem it; it.tri = slot; it.tx = (uint16_t)tx; it.ty = (uint16_t)ty;
            q.items[is] = it;
        }
}

// smallMax: largest bounding box (pixels per side) the calling lane rasterises itself; expensive sinks pass a small one so
// that their fragments are produced by the large-triangle kernel, a warp per triangle with converged sink calls
template <class Sink>
VKR_HD void emit_setup(const SetupTri& t, uint32_t seq, const Queues& q, const Sink& sink, int32_t smallMax = SMALL_MAX, bool strips = false) {
    const int32_t w = t.bx1 - t.bx0 + 1, h = t.by1 - t.by0 + 1;
    if (w <= smallMax && h <= smallMax && t.ext < 16384) raster_small(t, sink);
    else enqueue_large(t, seq, q, strips);
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
    const float t = xdiv(din, xsub(din, dout));
    float4 r;
    r.x = xadd(in.x, xmul(t, xsub(out.x, in.x)));
    r.y = xadd(in.y, xmul(t, xsub(out.y, in.y)));
    r.z = xadd(in.z, xmul(t, xsub(out.z, in.z)));
    r.w = xadd(in.w, xmul(t, xsub(out.w, in.w)));
    return r;
}

// A triangle with at least one vertex outside the clip volume: trivial reject, else
// Sutherland-Hodgman (intersection always from the inside vertex towards the outside one) + fan.
template <class Sink>
VKR_HDN void clip_and_emit(float4 c0, float4 c1, float4 c2, uint32_t seq, const RasterTarget& rt, const Queues& q, const Sink& sink) {
    if ((c0.z < 0 && c1.z < 0 && c2.z < 0) || (c0.z > c0.w && c1.z > c1.w && c2.z > c2.w) ||
        (c0.x < -c0.w && c1.x < -c1.w && c2.x < -c2.w) || (c0.x > c0.w && c1.x > c1.w && c2.x > c2.w) ||
        (c0.y < -c0.w && c1.y < -c1.w && c2.y < -c2.w) || (c0.y > c0.w && c1.y > c1.w && c2.y > c2.w))
        return;
    float4 bufA[12], bufB[12];
    int n = 3;
    bufA[0] = c0; bufA[1] = c1; bufA[2] = c2;
    float4 *src = bufA, *dst = bufB;
    for (int p = 0; p < 6 && n >= 3; ++p) {
        int m = 0;
        for (int i = 0; i < n; ++i) {
            const float4 a = src[i], b = src[(i + 1) % n];
            const float da = plane_dist(p, a), db = plane_dist(p, b);
            const bool ia = da >= 0.0f, ib = db >= 0.0f;
            if (ia) dst[m++] = a;
            if (ia != ib) dst[m++] = ia ? clip_lerp(a, b, da, db) : clip_lerp(b, a, db, da);
        }
        n = m;
        float4* tmp = src; src = dst; dst = tmp;
    }
    if (n < 3) return;
    for (int i = 1; i + 1 < n; ++i) {
        if (!(src[0].w > 0.0f && src[i].w > 0.0f && src[i + 1].w > 0.0f)) continue;
        SetupTri t;
        if (setup_snapped(project_vertex(src[0], rt), project_vertex(src[i], rt), project_vertex(src[i + 1], rt), rt, t))
            emit_setup(t, seq, q, sink);
    }
}

// One (large triangle, tile) work item; lanes of a warp split the tile's pixels.
// The sample rectangle of an item: the triangle's whole box, a TILE x TILE tile of it, or a TILE x 16 strip.
VKR_HD void item_rect(const LargeTri& lt, const LargeItem& it, int32_t& x0, int32_t& y0, int32_t& x1, int32_t& y1) {
    if (it.tx == ITEM_WHOLE) { x0 = lt.bx0; y0 = lt.by0; x1 = lt.bx1; y1 = lt.by1; return; }
    x0 = imax((int32_t)it.tx * TILE, (int32_t)lt.bx0); x1 = imin((int32_t)it.tx * TILE + TILE - 1, (int32_t)lt.bx1);
    if (it.ty & ITEM_STRIP) {
        const int32_t sy = (int32_t)(it.ty & (uint16_t)~ITEM_STRIP);
        y0 = imax(sy * STRIP_ROWS, (int32_t)lt.by0); y1 = imin(sy * STRIP_ROWS + STRIP_ROWS - 1, (int32_t)lt.by1);
    } else { y0 = imax((int32_t)it.ty * TILE, (int32_t)lt.by0); y1 = imin((int32_t)it.ty * TILE + TILE - 1, (int32_t)lt.by1); }
}
template <class Sink>
VKR_HD void raster_item(const LargeTri& lt, const LargeItem& it, const Sink& sink, int lane, int nlanes) {
    int32_t x0, y0, x1, y1;
    item_rect(lt, it, x0, y0, x1, y1);
    raster_rect64(lt.X, lt.Y, lt.z0, lt.dzdx, lt.dzdy, x0, y0, x1, y1, sink, lane, nlanes);
}

#if defined(__CUDACC__)
template <class Sink>
__device__ __forceinline__ void raster_item_compact(const LargeTri& lt, const LargeItem& it, const Sink& sink, int lane, uint32_t* wq) {
    int32_t x0, y0, x1, y1;
    item_rect(lt, it, x0, y0, x1, y1);
    raster_rect64_compact(lt.X, lt.Y, lt.z0, lt.dzdx, lt.dzdy, x0, y0, x1, y1, sink, lane, wq);
}
#endif

// ---- the two phases of the geometry kernel (one CTA = one cluster of <= 256 triangles) -----
// Phase 1, one unique vertex: transform, classify, project, snap.  w == 1 flags "outside the clip
// volume": triangles touching such a vertex take the clipping path from the original positions.
struct CachedVtx { int32_t X, Y; uint32_t zbits; uint32_t flag; };   // 16 B, one LDS.128

VKR_HD CachedVtx cluster_vertex(const float* __restrict__ mvp, float4 p, const RasterTarget& rt) {
    const float4 c = transform_point(mvp, p.x, p.y, p.z);
    CachedVtx o;
    if (!inside_clip_volume(c)) { o.X = 0; o.Y = 0; o.zbits = 0; o.flag = 1u; return o; }
    const SnapVtx s = project_vertex(c, rt);
    o.X = s.X; o.Y = s.Y; o.zbits = f2bits(s.z); o.flag = 0u;
    return o;
}

// Phase 2, one triangle.  `gid` maps a local vertex id back to the mesh vertex for the slow path.
template <class Sink>
VKR_HD void cluster_triangle(const CachedVtx& a, const CachedVtx& b, const CachedVtx& c, const float* __restrict__ mvp,
                             const float4* __restrict__ pos, uint32_t g0, uint32_t g1, uint32_t g2, uint32_t seq,
                             const RasterTarget& rt, const Queues& q, const Sink& sink) {
    if ((a.flag | b.flag | c.flag) != 0u) {
#if defined(__CUDA_ARCH__)
        const float4 p0 = __ldg(pos + g0), p1 = __ldg(pos + g1), p2 = __ldg(pos + g2);
#else
        const float4 p0 = pos[g0], p1 = pos[g1], p2 = pos[g2];
#endif
        clip_and_emit(transform_point(mvp, p0.x, p0.y, p0.z), transform_point(mvp, p1.x, p1.y, p1.z),
                      transform_point(mvp, p2.x, p2.y, p2.z), seq, rt, q, sink);
        return;
    }
    SetupTri t;
    const SnapVtx v0{a.X, a.Y, bits2f(a.zbits)}, v1{b.X, b.Y, bits2f(b.zbits)}, v2{c.X, c.Y, bits2f(c.zbits)};
    if (setup_snapped(v0, v1, v2, rt, t)) emit_setup(t, seq, q, sink);
}

#if defined(__CUDACC__)
#if defined(VKR_BIN_PROBE)
static const uint32_t VKR_BIN_CAP = 16384u;          // records per tile bin (the probe wraps around: nothing reads the bins)
__device__ uint4* g_binPool = nullptr;
__device__ uint32_t* g_binCount = nullptr;
#endif
// ---- warp-cooperative handling: candidates of small triangles dealt to the lanes; triangles too big for that ------------
// cluster_triangle_deferred sets one triangle up and says who rasterises it: 0 nobody (culled, or clipped and emitted on the
// spot), 1 medium (extent < 64 px: the warp queues it as one item for large_raster_kernel's 32-bit path), 2 large (the warp
// enumerates its TILE x TILE tiles and queues them), 3 small: its record is in `out.rec` and `count` candidates wait for the warp's
// candidate stage (geometry_warp).  `out` is the lane's slot of the warp's shared memory.
#if defined(VKR_RASTER_CANDIDATES)
union ParkedTri { FragRec rec; LargeTri lt; };
static_assert(sizeof(ParkedTri) == 64, "one 64-byte shared-memory slot per lane");
#else
struct ParkedTri { LargeTri lt; };           // only medium and large triangles are parked
#endif
template <class Sink>
__device__ __forceinline__ int cluster_triangle_deferred(const CachedVtx& a, const CachedVtx& b, const CachedVtx& c, const float* __restrict__ mvp,
                                                         const float4* __restrict__ pos, uint32_t g0, uint32_t g1, uint32_t g2, uint32_t seq,
                                                         const RasterTarget& rt, const Queues& q, const Sink& sink, ParkedTri& out, uint32_t& count) {
    if ((a.flag | b.flag | c.flag) != 0u) {
        const float4 p0 = __ldg(pos + g0), p1 = __ldg(pos + g1), p2 = __ldg(pos + g2);
        clip_and_emit(transform_point(mvp, p0.x, p0.y, p0.z), transform_point(mvp, p1.x, p1.y, p1.z),
                      transform_point(mvp, p2.x, p2.y, p2.z), seq, rt, q, sink);
        return 0;
    }
    SetupTri t;
    const SnapVtx v0{a.X, a.Y, bits2f(a.zbits)}, v1{b.X, b.Y, bits2f(b.zbits)}, v2{c.X, c.Y, bits2f(c.zbits)};
    if (!setup_snapped(v0, v1, v2, rt, t)) return 0;
    const int32_t w = t.bx1 - t.bx0 + 1, h = t.by1 - t.by0 + 1;
#if defined(VKR_BIN_PROBE)
    // A/B probe (profiles/r02_ab.md, "tile bins"): what the APPEND half of a tile-binned rasteriser would add to this kernel — the
    // set-up triangle (48 B) stored into the bin of the 64x64 tile of its first corner, its slot taken from that tile's counter.
    // Nothing reads the bins: the image is unchanged, only the cost of writing them is measured.
    if (g_binPool != nullptr) {
        const uint32_t tile = (uint32_t)(t.by0 >> 6) * (uint32_t)((rt.width + 63) >> 6) + (uint32_t)(t.bx0 >> 6);
        const uint32_t slot = atomicAdd(g_binCount + tile, 1u) & (VKR_BIN_CAP - 1u);
        uint4* rec = g_binPool + ((size_t)tile * VKR_BIN_CAP + slot) * 3u;
        rec[0] = make_uint4((uint32_t)t.X[0], (uint32_t)t.X[1], (uint32_t)t.X[2], (uint32_t)t.Y[0]);
        rec[1] = make_uint4((uint32_t)t.Y[1], (uint32_t)t.Y[2], f2bits(t.z0), f2bits(t.dzdx));
        rec[2] = make_uint4(f2bits(t.dzdy), (uint32_t)t.bx0 | ((uint32_t)t.by0 << 16), (uint32_t)t.bx1 | ((uint32_t)t.by1 << 16), seq);
    }
#endif
#if defined(VKR_RASTER_CANDIDATES)
    if (w <= SMALL_MAX && h <= SMALL_MAX && t.ext < 16384) { count = make_frag_rec(t, out.rec); return 3; }
#else
    if (w <= SMALL_MAX && h <= SMALL_MAX && t.ext < 16384) { raster_small_loop(t, sink); return 0; }
#endif
    LargeTri& lt = out.lt;
#pragma unroll
    for (int i = 0; i < 3; ++i) { lt.X[i] = t.X[i]; lt.Y[i] = t.Y[i]; }
    lt.z0 = t.z0; lt.dzdx = t.dzdx; lt.dzdy = t.dzdy;
    lt.bx0 = (int16_t)t.bx0; lt.by0 = (int16_t)t.by0; lt.bx1 = (int16_t)t.bx1; lt.by1 = (int16_t)t.by1;
    lt.seq = seq;
    return t.ext < 16384 ? 1 : 2;
}

// The candidate stage: the 32 lanes of a warp hold `count` candidates each (records in their `slots`); candidate j of the
// concatenation goes to lane j mod 32, 32 candidates per round.  `excl` is the exclusive prefix sum of the counts over the
// lanes, `total` the sum.  The non-empty triangles are numbered densely, `map[d]` = (lane | first candidate << 8, seq) of the
// d-th one.  In the round that starts at candidate jb, the triangles whose first candidate falls inside the round set a bit at
// that position (one warp-wide OR, REDUX); the owner of the candidate on lane L is then the triangle number
// (triangles that start before the round) + (bits at positions <= L) - 1: no search.
#if defined(VKR_RASTER_CANDIDATES)
template <class MakeSink>
__device__ __forceinline__ void raster_candidates_warp(const ParkedTri* slots, uint2* map, uint32_t excl, uint32_t count, uint32_t seq, uint32_t total,
                                                       int lane, const MakeSink& makeSink) {
    const uint32_t nonEmpty = __ballot_sync(0xFFFFFFFFu, count != 0u);
    if (count) map[__popc(nonEmpty & ((1u << lane) - 1u))] = make_uint2((uint32_t)lane | (excl << 8), seq);
    __syncwarp();
    const uint32_t below = 0xFFFFFFFFu >> (31 - lane);                  // positions <= lane
    for (uint32_t jb = 0; jb < total; jb += 32) {
        const uint32_t pos = excl - jb;                                  // (wraps for triangles that started earlier: no bit)
        const uint32_t heads = __reduce_or_sync(0xFFFFFFFFu, (count != 0u && pos < 32u) ? 1u << pos : 0u);
        const uint32_t before = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, count != 0u && excl < jb));
        const uint32_t j = jb + lane;
        if (j < total) {
            const uint2 m = map[before + (uint32_t)__popc(heads & below) - 1u];
            const uint4* rp = reinterpret_cast<const uint4*>(&slots[m.x & 31u].rec);
            const uint4 r0 = rp[0], r1 = rp[1], r2 = rp[2], r3 = rp[3];
            FragRec r;
            r.e0 = (int32_t)r0.x; r.e1 = (int32_t)r0.y; r.e2 = (int32_t)r0.z; r.ax0 = (int32_t)r0.w;
            r.ax1 = (int32_t)r1.x; r.ax2 = (int32_t)r1.y; r.by0 = (int32_t)r1.z; r.by1 = (int32_t)r1.w;
            r.by2 = (int32_t)r2.x; r.z0 = __uint_as_float(r2.y); r.dzdx = __uint_as_float(r2.z); r.dzdy = __uint_as_float(r2.w);
            r.fx0 = __uint_as_float(r3.x); r.fy0 = __uint_as_float(r3.y); r.origin = r3.z; r.wdiv = r3.w;
            raster_candidate(r, j - (m.x >> 8), makeSink(m.y));
        }
    }
}

#endif

// All 32 lanes rasterise one medium triangle (warp-uniform `lt`): 8-pixel strips of its bounding box dealt to the lanes,
// 32-bit edge functions as in raster_small.
template <class Sink>
__device__ __forceinline__ void raster_medium_coop(const LargeTri& lt, const Sink& sink, int lane, uint32_t part = 0u, uint32_t partShift = 0u) {
    int32_t e[3], ax[3], by[3];
    const int32_t Px = (int32_t)lt.bx0 * 256 + 128, Py = (int32_t)lt.by0 * 256 + 128;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3, k = (i + 2) % 3;
        const int32_t dx = lt.X[k] - lt.X[j], dy = lt.Y[k] - lt.Y[j];
        const bool topleft = (dy < 0) || (dy == 0 && dx > 0);
        e[i] = dx * (Py - lt.Y[j]) - dy * (Px - lt.X[j]) - (topleft ? 0 : 1);
        ax[i] = -dy * 256;
        by[i] = dx * 256;
    }
    const int32_t w = lt.bx1 - lt.bx0 + 1, h = lt.by1 - lt.by0 + 1;
    // (part, partShift): this warp's band of the rows when 2^partShift warps share the triangle (large_raster_kernel on a short queue)
    const int32_t per = (h + (int32_t)(1u << partShift) - 1) >> partShift, rowBegin = (int32_t)part * per, rowEnd = imin(h, rowBegin + per);
    const int32_t strips = (w + 7) >> 3, n = strips * rowEnd;
    for (int32_t i = strips * rowBegin + lane; i < n; i += 32) {
        const int32_t row = i / strips, sx = i - row * strips;
        const int32_t py = lt.by0 + row, pxa = lt.bx0 + sx * 8;
        int32_t f0 = e[0] + row * by[0] + sx * 8 * ax[0], f1 = e[1] + row * by[1] + sx * 8 * ax[1], f2 = e[2] + row * by[2] + sx * 8 * ax[2];
#if defined(VKR_STRIP_BOUNDS)
        {   // both ends of the strip outside one edge: none of it is covered (strip_coverage; half of a bounding box is such strips)
            const int32_t last = imin(7, (int32_t)lt.bx1 - pxa);
            const int32_t g0 = f0 + last * ax[0], g1 = f1 + last * ax[1], g2 = f2 + last * ax[2];
            if (((f0 & g0) | (f1 & g1) | (f2 & g2)) < 0) continue;
        }
#endif
#pragma unroll
        for (int32_t k = 0; k < 8; ++k) {
            if (pxa + k <= lt.bx1 && (f0 | f1 | f2) >= 0) sink.put(pxa + k, py, depth_at(lt.z0, lt.dzdx, lt.dzdy, lt.X[0], lt.Y[0], pxa + k, py));
            f0 += ax[0]; f1 += ax[1]; f2 += ax[2];
        }
    }
}

// All 32 lanes enumerate the tiles of one large triangle (warp-uniform `lt`) and queue the overlapping ones.
__device__ __forceinline__ void enqueue_large_coop(const LargeTri& lt, const Queues& q, int lane) {
    uint32_t slot = 0;
    if (lane == 0) slot = atomicAdd(&q.counters[0], 1u);
    slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
    if (slot >= q.triCap) { if (lane == 0) q.counters[2] = 1; return; }
    if (lane == 0) q.tris[slot] = lt;
    EdgeEq e;
    make_edges(lt.X, lt.Y, e);
    const int32_t tx0 = lt.bx0 / TILE, tx1 = lt.bx1 / TILE, ty0 = lt.by0 / TILE, ty1 = lt.by1 / TILE;
    const int32_t ntx = tx1 - tx0 + 1, n = ntx * (ty1 - ty0 + 1);
    for (int32_t base = 0; base < n; base += 32) {
        const int32_t i = base + lane;
        bool hit = false;
        int32_t tx = 0, ty = 0;
        if (i < n) {
            ty = ty0 + i / ntx; tx = tx0 + i % ntx;
            const int32_t x0 = imax(tx * TILE, (int32_t)lt.bx0), x1 = imin(tx * TILE + TILE - 1, (int32_t)lt.bx1);
            const int32_t y0 = imax(ty * TILE, (int32_t)lt.by0), y1 = imin(ty * TILE + TILE - 1, (int32_t)lt.by1);
            hit = tile_overlaps(e, x0, y0, x1, y1);
        }
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, hit);
        if (!m) continue;
        uint32_t first = 0;
        if (lane == 0) first = atomicAdd(&q.counters[1], (uint32_t)__popc(m));
        first = __shfl_sync(0xFFFFFFFFu, first, 0);
        if (hit) {
            const uint32_t is = first + (uint32_t)__popc(m & ((1u << lane) - 1u));
            if (is >= q.itemCap) q.counters[2] = 1;
            else { LargeItem it; it.tri = slot; it.tx = (uint16_t)tx; it.ty = (uint16_t)ty; q.items[is] = it; }
        }
    }
}
#endif

// ---- cluster culling --------------------------------------------------------------------------
// Before a geometry CTA transforms anything it projects the 8 corners of its cluster's object-space
// bounding box and drops the cluster when no fragment of it can survive:
//   (1) every corner outside one frustum plane  -> all triangles are clipped away or scissored;
//   (2) the projected rectangle holds no pixel centre inside the scissor;
//   (3) every pixel of that rectangle already stores a depth nearer than the nearest the cluster can
//       produce (occlusion against the live depth / visibility buffer; draws are processed front to back).
// All three are conservative, so the image is identical with and without them.  Rigour: x, y, z and w
// are affine in the object-space position, so over the box they are bounded by their corner values,
// and x/w, y/w, z/w (linear-fractional, w > 0) too.  The corners are evaluated in floating point like
// the vertices; both are within 3 roundings of the exact affine value, |err| <= 3*2^-24*mag with
// mag = sum |m_i * p_i|; the margins below use 2^-20*mag (>5x that bound, both sides) plus the
// 1/512 px vertex snap.  Depth written by the rasteriser is the triangle's plane evaluated at a covered
// sample (inside the triangle), so it lies in the vertex depth range up to the rounding of the plane
// evaluation; a quarter of the cluster's own depth range plus 2^-19 is subtracted for that.
struct CornerAcc {
    float xmin, xmax, ymin, ymax;   // NDC bounds, already widened by the margins (valid when allFront)
    float zmin, zmax;               // depth bounds (valid when allFront)
    uint32_t outAll;                // bit p set: every corner is outside frustum plane p by more than the margin
    uint32_t allFront;              // every corner has w safely > 0
    float wmax, wmin;               // bounds of w over the box (valid when allFront; wmin > 0)
    float mgx, mgy;                 // largest rounding margin of x/w and y/w over the corners, NDC units (valid when allFront)
};
static const float CULL_EPS = 9.5367431640625e-07f;   // 2^-20

// One clip-space component (0 = x, 1 = y, 2 = z, 3 = w) of one box corner and its rounding-error bound.
VKR_HD void corner_component(const float* __restrict__ m, int comp, float px, float py, float pz, float& v, float& e) {
    v = xfma(m[comp], px, xfma(m[4 + comp], py, xfma(m[8 + comp], pz, m[12 + comp])));
    e = CULL_EPS * (fabsf(m[comp]) * fabsf(px) + fabsf(m[4 + comp]) * fabsf(py) + fabsf(m[8 + comp]) * fabsf(pz) + fabsf(m[12 + comp]));
}
// Outside tests of a component against the planes it takes part in: x, y against +-w; z against 0 and w.
VKR_HD void corner_outside(int comp, float v, float e, float w, float ew, bool& outLo, bool& outHi) {
    outLo = comp < 2 ? (v + w < -(e + ew)) : (v < -e);
    outHi = v - w > (e + ew);
}
VKR_HD bool corner_front(float w, float ew) { return w > 2.0f * ew && w > 1e-30f; }
// Widened NDC interval of a component of a corner with w safely > 0.
VKR_HD float corner_ndc(float v, float e, float w, float ew, float& lo, float& hi) {
    const float rw = 1.0f / (w - ew);             // >= 1/w_exact
    const float n = v / w;
    const float mg = (e + fabsf(n) * ew) * rw * 2.0f;
    lo = n - mg; hi = n + mg;
    return mg;
}

VKR_HD CornerAcc corner_eval(const float* __restrict__ m, float4 bmin, float4 bmax, int corner) {
    const float px = (corner & 1) ? bmax.x : bmin.x, py = (corner & 2) ? bmax.y : bmin.y, pz = (corner & 4) ? bmax.z : bmin.z;
    float v[4], e[4];
    for (int c = 0; c < 4; ++c) corner_component(m, c, px, py, pz, v[c], e[c]);
    CornerAcc a;
    uint32_t out = 0;
    for (int c = 0; c < 3; ++c) {
        bool lo, hi;
        corner_outside(c, v[c], e[c], v[3], e[3], lo, hi);
        if (lo) out |= 1u << (2 * c);
        if (hi) out |= 2u << (2 * c);
    }
    a.outAll = out;
    const bool front = corner_front(v[3], e[3]);
    a.allFront = front ? 1u : 0u;
    if (front) {
        a.mgx = corner_ndc(v[0], e[0], v[3], e[3], a.xmin, a.xmax);
        a.mgy = corner_ndc(v[1], e[1], v[3], e[3], a.ymin, a.ymax);
        corner_ndc(v[2], e[2], v[3], e[3], a.zmin, a.zmax);
        a.wmax = v[3] + e[3]; a.wmin = v[3] - e[3];
    } else {
        a.xmin = a.ymin = a.zmin = 0.0f; a.xmax = a.ymax = a.zmax = 0.0f;
        a.wmax = 0.0f; a.wmin = 0.0f; a.mgx = a.mgy = 0.0f;
    }
    return a;
}
VKR_HD CornerAcc corner_merge(const CornerAcc& a, const CornerAcc& b) {
    CornerAcc r;
    r.xmin = fminf(a.xmin, b.xmin); r.xmax = fmaxf(a.xmax, b.xmax);
    r.ymin = fminf(a.ymin, b.ymin); r.ymax = fmaxf(a.ymax, b.ymax);
    r.zmin = fminf(a.zmin, b.zmin); r.zmax = fmaxf(a.zmax, b.zmax);
    r.outAll = a.outAll & b.outAll;
    r.allFront = a.allFront & b.allFront;
    r.wmax = fmaxf(a.wmax, b.wmax); r.wmin = fminf(a.wmin, b.wmin); r.mgx = fmaxf(a.mgx, b.mgx); r.mgy = fmaxf(a.mgy, b.mgy);
    return r;
}

static const int CLUSTER_DRAW = 1, CLUSTER_CULLED = 0, CLUSTER_TEST = 2;
#ifndef VKR_OCC_MAX_PIXELS
#define VKR_OCC_MAX_PIXELS 1024
#endif
static const int OCC_MAX_PIXELS = VKR_OCC_MAX_PIXELS;      // occlusion test only when the rectangle is at most this many pixels
struct ClusterRect { int32_t x0, y0, x1, y1; uint32_t zminBits; int32_t verdict; };

VKR_HD ClusterRect cluster_finalize(const CornerAcc& a, const RasterTarget& rt) {
    ClusterRect r;
    r.x0 = rt.sx0; r.y0 = rt.sy0; r.x1 = rt.sx1 - 1; r.y1 = rt.sy1 - 1; r.zminBits = 0u;
    if (a.outAll) { r.verdict = CLUSTER_CULLED; return r; }
    if (!a.allFront) { r.verdict = CLUSTER_DRAW; return r; }
    // pixel i is sampled at i + 0.5; a covered sample lies within the widened bounds (+ the snap of 1/512 px)
    const float snap = 0.004f;
    const float fx0 = fmaf(a.xmin, rt.halfW, rt.halfW) - snap - 0.5f, fx1 = fmaf(a.xmax, rt.halfW, rt.halfW) + snap - 0.5f;
    const float fy0 = fmaf(a.ymin, rt.halfH, rt.halfH) - snap - 0.5f, fy1 = fmaf(a.ymax, rt.halfH, rt.halfH) + snap - 0.5f;
    // NaN-safe clamps: fmaxf/fminf return the non-NaN operand, which is the scissor bound (= widest rectangle)
    const float cx0 = fminf(fmaxf(ceilf(fx0), (float)rt.sx0), (float)rt.sx1), cx1 = fmaxf(fminf(floorf(fx1), (float)(rt.sx1 - 1)), (float)rt.sx0 - 1.0f);
    const float cy0 = fminf(fmaxf(ceilf(fy0), (float)rt.sy0), (float)rt.sy1), cy1 = fmaxf(fminf(floorf(fy1), (float)(rt.sy1 - 1)), (float)rt.sy0 - 1.0f);
    r.x0 = (int32_t)cx0; r.x1 = (int32_t)cx1; r.y0 = (int32_t)cy0; r.y1 = (int32_t)cy1;
    if (r.x0 > r.x1 || r.y0 > r.y1) { r.verdict = CLUSTER_CULLED; return r; }
    if ((int64_t)(r.x1 - r.x0 + 1) * (r.y1 - r.y0 + 1) > OCC_MAX_PIXELS) { r.verdict = CLUSTER_DRAW; return r; }
    float z = a.zmin - (0.25f * (a.zmax - a.zmin) + 1.9073486328125e-06f);
    if (!(z > 0.0f)) { r.verdict = CLUSTER_DRAW; return r; }        // reaches depth 0 (or NaN): nothing can hide it
    if (z > 1.0f) z = 1.0f;
    r.zminBits = f2bits(z);
    r.verdict = CLUSTER_TEST;
    return r;
}
// ---- facing cull of a whole cluster ---------------------------------------------------------------
// The face cull is decided per triangle on SNAPPED window coordinates (setup_snapped: sign of the doubled area A2).  A cluster
// can be dropped before any vertex work when every triangle of it would be culled that way, and that can be argued exactly:
//   * with clip coordinates (x, y, w) = M' (p, 1), the determinant D of a triangle's three (x, y, w) rows equals pi . E, where
//     pi = (N, -N . p0), N = (p1 - p0) x (p2 - p0), is the triangle's plane and E (4 numbers per matrix, `eye`, computed on the
//     host in double precision from the very matrix the kernel uses) is the homogeneous point M' maps to (0, 0, 0): the eye.
//     With every w > 0 the doubled area in sub-pixels is A2 = (128 W)(128 H) D / (w0 w1 w2), so its sign is D's;
//   * every unit normal of the cluster lies within the angle alpha of the cone axis a (ClusterCone, built with the clusters),
//     its vertices within r of c.  With v = E.xyz - E.w c and beta the angle between a and +-v:
//     |D| >= Nmin (|v| cos(alpha + beta) - |E.w| r)  whenever that is positive (Nmin: the smallest |N| of the cluster);
//   * the kernel's vertices differ from the exact ones by at most delta = 1 sub-pixel (rounding to the grid: 0.5, the divide
//     and the viewport FMA: < 0.1 at 2^20 sub-pixels) plus the rounding of the matrix product, a quarter of the margin the box
//     test carries (mgx, mgy: 2^-20 |terms| where three roundings make 3 * 2^-24); moving every vertex by delta changes A2 by at
//     most 8 delta R + 8 delta^2, R the largest extent of a TRIANGLE in sub-pixels: an object-space edge of length L projects to
//     at most L (|row x| + max|x/w| |row w|) / wmin in NDC (x1/w1 - x0/w0 = (dx - (x0/w0) dw) / w1), Lmax being stored with the cone.
// If the lower bound of |A2| exceeds that with a factor to spare, no triangle changes sign (or vanishes) under snapping: all of
// them are culled by the per-triangle test, and dropping the cluster changes nothing.
// The point E with M' E = 0 for the 3 x 4 matrix M' of the x, y and w rows of `mvp` (column-major): E_i = +-minor_i.  The sign is
// fixed so that det[(x, y, w) rows of a triangle] == plane(p0, p1, p2) . E, checked on one triangle in double precision.
VKR_HD void eye_point(const float* m, float out[8]) {
    const double r[3][4] = {{m[0], m[4], m[8], m[12]}, {m[1], m[5], m[9], m[13]}, {m[3], m[7], m[11], m[15]}};
    double E[4];
    for (int skip = 0; skip < 4; ++skip) {
        const int c0 = skip == 0 ? 1 : 0, c1 = skip <= 1 ? 2 : 1, c2 = skip <= 2 ? 3 : 2;
        const double mn = r[0][c0] * (r[1][c1] * r[2][c2] - r[1][c2] * r[2][c1]) - r[0][c1] * (r[1][c0] * r[2][c2] - r[1][c2] * r[2][c0]) +
                          r[0][c2] * (r[1][c0] * r[2][c1] - r[1][c1] * r[2][c0]);
        E[skip] = (skip & 1) ? -mn : mn;
    }
    // reference triangle (0,0,0) (1,0,0) (0,1,0): plane N = (0,0,1), pi = (0,0,1,0) -> pi . E = E[2]
    const double P[3][3] = {{r[0][3], r[1][3], r[2][3]}, {r[0][0] + r[0][3], r[1][0] + r[1][3], r[2][0] + r[2][3]}, {r[0][1] + r[0][3], r[1][1] + r[1][3], r[2][1] + r[2][3]}};
    const double D = P[0][0] * (P[1][1] * P[2][2] - P[1][2] * P[2][1]) - P[0][1] * (P[1][0] * P[2][2] - P[1][2] * P[2][0]) + P[0][2] * (P[1][0] * P[2][1] - P[1][1] * P[2][0]);
    if ((D < 0.0) != (E[2] < 0.0)) for (int i = 0; i < 4; ++i) E[i] = -E[i];
    // scale into float range (only ratios and signs matter to the test, but a2min uses the magnitude: keep it)
    for (int i = 0; i < 4; ++i) out[i] = (float)E[i];
    // Lipschitz constants of x, y and w in the object-space position (rounded up), for the triangle-extent bound of cone_culls
    for (int k = 0; k < 3; ++k) out[4 + k] = (float)(sqrt(r[k][0] * r[k][0] + r[k][1] * r[k][1] + r[k][2] * r[k][2]) * (1.0 + 1e-6));
    out[7] = 0.0f;
}

struct ClusterCone { float4 axisCos; float4 centerRadius; float4 nminSin; };      // (a, cos alpha) (c, r) (Nmin, sin alpha, Lmax, -)

VKR_HD bool cone_culls(const ClusterCone& k, const float* __restrict__ eye, const CornerAcc& a, const RasterTarget& rt) {
    if (!a.allFront || !(k.nminSin.x > 0.0f) || !(a.wmin > 0.0f)) return false;
    const float sgn = rt.cullFront ? -1.0f : 1.0f;            // main pass: culled when D > 0; shadow pass (cull FRONT): when D < 0
    const float vx = eye[0] - eye[3] * k.centerRadius.x, vy = eye[1] - eye[3] * k.centerRadius.y, vz = eye[2] - eye[3] * k.centerRadius.z;
    const float lv = sqrtf(vx * vx + vy * vy + vz * vz);
    if (!(lv > 0.0f)) return false;
    const float cosb = sgn * (k.axisCos.x * vx + k.axisCos.y * vy + k.axisCos.z * vz) / lv;
    if (!(cosb > 0.0f)) return false;
    const float sinb = sqrtf(fmaxf(0.0f, 1.0f - cosb * cosb));
    const float cab = k.axisCos.w * cosb - k.nminSin.y * sinb;          // cos(alpha + beta)
    if (!(cab > 0.0f)) return false;
    const float lower = k.nminSin.x * (lv * cab - fabsf(eye[3]) * k.centerRadius.w);
    if (!(lower > 0.0f)) return false;
    const float a2min = (rt.halfW256 * rt.halfH256) * lower / (a.wmax * a.wmax * a.wmax);
    // eye[4..6]: |row x|, |row y|, |row w| of the matrix (xyz parts); the largest triangle extent, sub-pixels
    const float ndx = fmaxf(fabsf(a.xmin), fabsf(a.xmax)), ndy = fmaxf(fabsf(a.ymin), fabsf(a.ymax));
    const float ex = k.nminSin.z * (eye[4] + ndx * eye[6]) / a.wmin * rt.halfW256, ey = k.nminSin.z * (eye[5] + ndy * eye[6]) / a.wmin * rt.halfH256;
    const float R = fmaxf(ex, ey) * 1.001f + 2.0f;
    const float delta = 1.0f + 0.25f * fmaxf(a.mgx * rt.halfW256, a.mgy * rt.halfH256);
    const float moved = 8.0f * delta * R + 8.0f * delta * delta;
    return a2min * 0.9f > moved;                               // (the bounds are exact; 10 % for the float evaluation of this test.  NaN anywhere: false)
}

// The whole test of one cluster for one pass, by one thread (host emulation; expand_chunks_kernel spreads the corners over
// eight lanes and calls cluster_verdict on the merged result).
VKR_HD CullRec cluster_verdict(const CornerAcc& a, const ClusterCone* cone, const float* __restrict__ eye, const RasterTarget& rt) {
    ClusterRect r = cluster_finalize(a, rt);
#if !defined(VKR_NO_CONE)      // (-DVKR_NO_CONE: the A/B of the facing cull)
    if (r.verdict != CLUSTER_CULLED && cone != nullptr && cone_culls(*cone, eye, a, rt)) r.verdict = CLUSTER_CULLED;
#endif
    CullRec c;
    c.x0 = (int16_t)r.x0; c.y0 = (int16_t)r.y0; c.x1 = (int16_t)r.x1; c.y1 = (int16_t)r.y1;     // extents are <= 16384
    c.zminBits = r.zminBits; c.verdict = r.verdict;
    return c;
}
VKR_HD CullRec make_cull_rec(const float* __restrict__ mvp, float4 bmin, float4 bmax, const RasterTarget& rt,
                             const ClusterCone* cone = nullptr, const float* eye = nullptr) {
    CornerAcc a = corner_eval(mvp, bmin, bmax, 0);
    for (int k = 1; k < 8; ++k) a = corner_merge(a, corner_eval(mvp, bmin, bmax, k));
    return cluster_verdict(a, cone, eye, rt);
}
// A stored depth hides the cluster at that pixel iff it is strictly nearer than anything the cluster can write.
VKR_HD bool depth_hides(uint32_t storedBits, uint32_t zminBits) { return storedBits < zminBits; }

// ---- alpha test in the raster stage ---------------------------------------------------------
// mesh_pbr.frag / mesh.frag `discard` fragments whose albedo alpha < 0.5 (shaders/mesh_pbr.frag:193):
// such fragments must not write depth, so for draws whose albedo texture has any texel alpha < 0.5 the
// test runs here, before the visibility-buffer atomic.  It decides coverage, so it is evaluated with
// explicitly rounded operations in the oracle's order (texture2D + bary_at in oracle/ork.cpp).
struct AlphaCtx {              // what the alpha test needs to find a primitive's texture and vertices
    const DObject* objects;
    const uint2* chunkObj;       // seq >> 8 -> (draw, first triangle of the slot)
    const DMaterial* materials;
    const DTexture* textures;
};
struct AlphaTri {
    float cx[3], cy[3], cw[3];   // clip x, y, w of the ORIGINAL triangle's vertices
    float u[3], v[3];
    const DTexture* tex;
    float halfW, halfH;
};

// a / 255.0f, correctly rounded, for a = 0..255: the division the oracle does per texel, from a table (bit-identical)
struct Unorm8Table {
    float v[256];
    constexpr Unorm8Table() : v() { for (int i = 0; i < 256; ++i) v[i] = (float)i / 255.0f; }
};
#if defined(__CUDACC__)
__device__ constexpr Unorm8Table kUnorm8Dev{};
#endif
constexpr Unorm8Table kUnorm8Host{};
VKR_HD float alpha_texel(const DTexture& t, int level, int x, int y, int w) {
    const uint8_t a = t.data[t.mipOffset[level] + (size_t)y * w + x].w;
#if defined(__CUDA_ARCH__)
    return kUnorm8Dev.v[a];
#else
    return kUnorm8Host.v[a];
#endif
}
VKR_HD int alpha_wrap(int i, int n) {
    if ((n & (n - 1)) == 0) return i & (n - 1);       // power-of-two extent: the AND is the floor-mod
    int r = i % n; return r < 0 ? r + n : r;
}
VKR_HD float alpha_sample_level(const DTexture& t, int level, float u, float v) {
    int w = (int)(t.w >> level), h = (int)(t.h >> level);
    if (w < 1) w = 1;
    if (h < 1) h = 1;
    if (!(t.flags & 2u)) {
        const int x = alpha_wrap((int)floorf(xmul(u, (float)w)), w), y = alpha_wrap((int)floorf(xmul(v, (float)h)), h);
        return alpha_texel(t, level, x, y, w);
    }
    const float fu = xsub(xmul(u, (float)w), 0.5f), fv = xsub(xmul(v, (float)h), 0.5f);
    const float iu = floorf(fu), iv = floorf(fv);
    const float a = xsub(fu, iu), b = xsub(fv, iv);
    const int x0 = alpha_wrap((int)iu, w), x1 = alpha_wrap((int)iu + 1, w);
    const int y0 = alpha_wrap((int)iv, h), y1 = alpha_wrap((int)iv + 1, h);
    const float t00 = alpha_texel(t, level, x0, y0, w), t10 = alpha_texel(t, level, x1, y0, w);
    const float t01 = alpha_texel(t, level, x0, y1, w), t11 = alpha_texel(t, level, x1, y1, w);
    const float w00 = xmul(xsub(1.0f, a), xsub(1.0f, b)), w10 = xmul(a, xsub(1.0f, b)), w01 = xmul(xsub(1.0f, a), b), w11 = xmul(a, b);
    return xadd(xadd(xadd(xmul(t00, w00), xmul(t10, w10)), xmul(t01, w01)), xmul(t11, w11));
}
VKR_HD void alpha_bary(const AlphaTri& a, float sx, float sy, float b[3]) {
    const float nx = xdiv(xsub(sx, a.halfW), a.halfW), ny = xdiv(xsub(sy, a.halfH), a.halfH);
    float qx[3], qy[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) { qx[i] = xsub(a.cx[i], xmul(nx, a.cw[i])); qy[i] = xsub(a.cy[i], xmul(ny, a.cw[i])); }
    const float e0 = xsub(xmul(qx[1], qy[2]), xmul(qx[2], qy[1]));
    const float e1 = xsub(xmul(qx[2], qy[0]), xmul(qx[0], qy[2]));
    const float e2 = xsub(xmul(qx[0], qy[1]), xmul(qx[1], qy[0]));
    const float s = xadd(xadd(e0, e1), e2);
    if (s == 0.0f) { b[0] = 1.0f; b[1] = 0.0f; b[2] = 0.0f; return; }
    b[0] = xdiv(e0, s); b[1] = xdiv(e1, s); b[2] = xdiv(e2, s);
}
VKR_HD float alpha_lerp(const float a[3], const float b[3]) { return xadd(xadd(xmul(b[0], a[0]), xmul(b[1], a[1])), xmul(b[2], a[2])); }

VKR_HDN bool alpha_passes(const AlphaTri& a, int px, int py) {
    const DTexture& t = *a.tex;
    if (t.data == nullptr) return true;                       // unbound slot samples (0,0,0,1)
    const float cx = xadd((float)px, 0.5f), cy = xadd((float)py, 0.5f);
    float bc[3];
    alpha_bary(a, cx, cy, bc);
    const float u = alpha_lerp(a.u, bc), v = alpha_lerp(a.v, bc);
    const int levels = (int)t.levels;
    float alpha;
    if (levels == 1) alpha = alpha_sample_level(t, 0, u, v);
    else {
        const float ox = (px & 1) ? -1.0f : 1.0f, oy = (py & 1) ? -1.0f : 1.0f;
        float bx[3], by[3];
        alpha_bary(a, xadd(cx, ox), cy, bx);
        alpha_bary(a, cx, xadd(cy, oy), by);
        const float dudx = xmul(xsub(alpha_lerp(a.u, bx), u), ox), dvdx = xmul(xsub(alpha_lerp(a.v, bx), v), ox);
        const float dudy = xmul(xsub(alpha_lerp(a.u, by), u), oy), dvdy = xmul(xsub(alpha_lerp(a.v, by), v), oy);
        const float w = (float)t.w, h = (float)t.h;
        const float ax = xmul(dudx, w), bxv = xmul(dvdx, h), ay = xmul(dudy, w), byv = xmul(dvdy, h);
        const float rx = sqrtf(xadd(xmul(ax, ax), xmul(bxv, bxv))), ry = sqrtf(xadd(xmul(ay, ay), xmul(byv, byv)));
        const float rho = fmaxf(rx, ry);
        const float lambda = rho > 0.0f ? log2f(rho) : -126.0f;
        const float lod = fminf(fmaxf(lambda, 0.0f), (float)(levels - 1));
        const int l0 = (int)floorf(lod), l1 = l0 + 1 < levels ? l0 + 1 : levels - 1;
        const float f = xsub(lod, (float)l0);
        const float c0 = alpha_sample_level(t, l0, u, v);
        if (f == 0.0f) alpha = c0;
        else alpha = xadd(xmul(c0, xsub(1.0f, f)), xmul(alpha_sample_level(t, l1, u, v), f));
    }
    return !(alpha < 0.5f);
}

// Gather the alpha-test inputs of primitive `tri` of draw `o` (positions re-transformed exactly as the
// geometry pass does, uv from the attribute streams).
VKR_HD void make_alpha_tri(const DObject& o, const float* __restrict__ mvp, uint32_t g0, uint32_t g1, uint32_t g2,
                           const AlphaCtx& ax, const RasterTarget& rt, AlphaTri& a) {
    const uint32_t g[3] = {g0, g1, g2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 p = o.pos[g[k]];
        const float4 c = transform_point(mvp, p.x, p.y, p.z);
        a.cx[k] = c.x; a.cy[k] = c.y; a.cw[k] = c.w;
        a.u[k] = o.attrA[g[k]].w; a.v[k] = o.attrB[g[k]].w;
    }
    const int id = ax.materials[o.material].albedoID;
    a.tex = ax.textures + ((id >= 0 && id < VKR_MAX_BINDLESS) ? id : 0);
    a.halfW = rt.halfW; a.halfH = rt.halfH;
}

template <class Base>
struct CutoutSink {
    Base base;
    const AlphaTri* a;
    VKR_HD void put(int32_t px, int32_t py, float z) const {
        if (base.hidden(px, py, z)) return;              // depth first: the alpha test is ~100x the cost of the probe
        if (alpha_passes(*a, px, py)) base.put(px, py, z);
    }
};

// Phase 2 for a triangle of an alpha-tested draw (always from the original vertices).
template <class Sink>
VKR_HDN void cutout_triangle(const DObject& o, const float* __restrict__ mvp, uint32_t g0, uint32_t g1, uint32_t g2, uint32_t seq,
                             const RasterTarget& rt, const Queues& q, const AlphaCtx& ax, const Sink& base) {
    AlphaTri a;
    make_alpha_tri(o, mvp, g0, g1, g2, ax, rt, a);
    CutoutSink<Sink> sink{base, &a};
    const float4 p0 = o.pos[g0], p1 = o.pos[g1], p2 = o.pos[g2];
    const float4 c0 = transform_point(mvp, p0.x, p0.y, p0.z), c1 = transform_point(mvp, p1.x, p1.y, p1.z), c2 = transform_point(mvp, p2.x, p2.y, p2.z);
    if (inside_clip_volume(c0) && inside_clip_volume(c1) && inside_clip_volume(c2)) {
        SetupTri t;
        if (setup_snapped(project_vertex(c0, rt), project_vertex(c1, rt), project_vertex(c2, rt), rt, t)) emit_setup(t, seq, q, sink, CUTOUT_SMALL_MAX, VKR_CUTOUT_STRIPS != 0);
    } else clip_and_emit(c0, c1, c2, seq, rt, q, sink);
}

}  // namespace vkr
