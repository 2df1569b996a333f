// vk_jpeg.cpp — JPEG (ITU-T T.81 / JFIF) -> RGBA8 for glTF images: what stbi_load does for the reference's loader
// (src/vk_loader.cpp:280,293: every glTF image goes through stb_image, and the engine's own scene, Khronos Sponza, ships JPEG
// textures).  stb_image is not available here, so this file follows the standard itself:
//   * baseline and extended sequential DCT (SOF0 / SOF1) and progressive DCT (SOF2: spectral selection and successive
//     approximation, T.81 Annex G), Huffman coding, 8-bit samples, restart intervals;
//   * 1 component (grey) or 3 components (YCbCr per JFIF, or RGB when an Adobe APP14 marker says transform 0 or the component
//     ids are 'R','G','B'), any integer sampling ratio; 2x chroma is upsampled with the usual triangle filter ("fancy
//     upsampling": 3/4 nearest + 1/4 next-nearest sample), other ratios by replication;
//   * the inverse DCT is evaluated in double precision and rounded once, so every sample is within one level of any
//     conforming decoder's (stb_image's fixed-point IDCT included).
// Arithmetic coding, lossless / hierarchical modes, 12-bit samples and CMYK are refused (decode_jpeg returns false and the
// image takes the error checkerboard, like any image stbi_load fails on).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "vk_gltf.h"

namespace {

const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Huffman {                 // canonical code (T.81 Annex C) with an 8-bit look-ahead table
    bool present = false;
    uint8_t vals[256];
    int32_t maxcode[18];         // largest code of each length, -1 when none
    int32_t valptr[17], mincode[17];
    uint16_t look[256];          // (length << 8 | value) for codes of <= 8 bits, 0 otherwise

    bool build(const uint8_t counts[16], const uint8_t* symbols, int n) {
        std::memcpy(vals, symbols, (size_t)n);
        std::memset(look, 0, sizeof(look));
        int32_t code = 0, k = 0;
        for (int len = 1; len <= 16; ++len) {
            valptr[len] = k; mincode[len] = code;
            for (int i = 0; i < counts[len - 1]; ++i, ++k, ++code) {
                if (len <= 8) {
                    const int first = code << (8 - len), span = 1 << (8 - len);
                    for (int j = 0; j < span; ++j) look[first + j] = (uint16_t)((len << 8) | vals[k]);
                }
            }
            maxcode[len] = counts[len - 1] ? code - 1 : -1;
            if (code > (1 << len)) return false;
            code <<= 1;
        }
        maxcode[17] = 0x7FFFFFFF;
        present = true;
        return k == n;
    }
};

struct BitReader {               // entropy-coded segment: 0xFF00 is a data byte 0xFF, any other marker ends the segment
    const uint8_t* p; const uint8_t* end;
    uint32_t acc = 0; int cnt = 0;
    int marker = 0;              // the marker that stopped the segment (0: none yet)
    void fill() {
        while (cnt <= 24) {
            int b = 0;
            if (!marker && p < end) {
                b = *p++;
                if (b == 0xFF) {
                    int m = p < end ? *p : 0xD9;
                    while (m == 0xFF && p + 1 < end) { ++p; m = *p; }       // fill bytes
                    if (m == 0) ++p;
                    else { marker = m; ++p; b = 0; }
                }
            }
            acc |= (uint32_t)b << (24 - cnt);
            cnt += 8;
        }
    }
    int bits(int n) {            // n <= 16
        if (n == 0) return 0;
        if (cnt < n) fill();
        const int v = (int)(acc >> (32 - n));
        acc <<= n; cnt -= n;
        return v;
    }
    int bit() { return bits(1); }
    int peek8() { if (cnt < 8) fill(); return (int)(acc >> 24); }
    void reset() { acc = 0; cnt = 0; marker = 0; }
};

inline int extend(int v, int n) { return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v; }     // T.81 F.2.2.1

int decode_symbol(BitReader& br, const Huffman& h) {
    const uint16_t e = h.look[br.peek8()];
    if (e) { br.bits(e >> 8); return e & 0xFF; }
    int code = br.bits(8);
    for (int len = 9; len <= 16; ++len) {
        code = (code << 1) | br.bit();
        if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len]) return h.vals[h.valptr[len] + code - h.mincode[len]];
    }
    return -1;
}

struct Component {
    int id = 0, h = 1, v = 1, tq = 0;
    int blocksW = 0, blocksH = 0;     // allocated grid (whole MCUs)
    int ownW = 0, ownH = 0;           // blocks a non-interleaved scan of this component walks: ceil(samples / 8)
    std::vector<int16_t> coef;        // blocksW * blocksH * 64, natural (row-major) order inside a block
    int pred = 0, td = 0, ta = 0;
    std::vector<uint8_t> plane;       // blocksW*8 x blocksH*8 samples after the IDCT
};

struct Decoder {
    const uint8_t* d; size_t n; size_t pos = 0;
    uint16_t qt[4][64]; bool haveQt[4] = {false, false, false, false};
    Huffman dc[4], ac[4];
    int W = 0, H = 0, ncomp = 0, hmax = 1, vmax = 1, mcusX = 0, mcusY = 0;
    bool progressive = false, haveFrame = false;
    Component comp[3];
    int restartInterval = 0;
    int adobeTransform = -1;
    int eobrun = 0;

    int u8() { return pos < n ? d[pos++] : -1; }
    int u16() { const int a = u8(), b = u8(); return a < 0 || b < 0 ? -1 : (a << 8) | b; }

    bool dqt(size_t endPos) {
        while (pos < endPos) {
            const int pq = u8();
            if (pq < 0) return false;
            const int prec = pq >> 4, t = pq & 15;
            if (t > 3 || prec > 1) return false;
            for (int i = 0; i < 64; ++i) {
                const int v = prec ? u16() : u8();
                if (v < 0) return false;
                qt[t][kZigzag[i]] = (uint16_t)v;          // stored in natural order
            }
            haveQt[t] = true;
        }
        return true;
    }
    bool dht(size_t endPos) {
        while (pos < endPos) {
            const int tc = u8();
            if (tc < 0 || (tc >> 4) > 1 || (tc & 15) > 3) return false;
            uint8_t counts[16]; int total = 0;
            for (int i = 0; i < 16; ++i) { const int c = u8(); if (c < 0) return false; counts[i] = (uint8_t)c; total += c; }
            if (total > 256 || pos + (size_t)total > endPos) return false;
            Huffman& h = (tc >> 4) ? ac[tc & 15] : dc[tc & 15];
            if (!h.build(counts, d + pos, total)) return false;
            pos += (size_t)total;
        }
        return true;
    }
    bool sof(int marker) {
        if (haveFrame) return false;
        progressive = marker == 0xC2;
        const int prec = u8();
        H = u16(); W = u16(); ncomp = u8();
        if (prec != 8 || W <= 0 || H <= 0 || (ncomp != 1 && ncomp != 3) || W > 32768 || H > 32768) return false;
        for (int i = 0; i < ncomp; ++i) {
            Component& c = comp[i];
            c.id = u8();
            const int hv = u8();
            c.tq = u8();
            if (hv < 0 || c.tq < 0 || c.tq > 3) return false;
            c.h = hv >> 4; c.v = hv & 15;
            if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4) return false;
            hmax = std::max(hmax, c.h); vmax = std::max(vmax, c.v);
        }
        if (ncomp == 1) { comp[0].h = comp[0].v = 1; hmax = vmax = 1; }      // a single component is never subsampled against itself
        for (int i = 0; i < ncomp; ++i) if (hmax % comp[i].h || vmax % comp[i].v) return false;      // integer ratios only
        mcusX = (W + 8 * hmax - 1) / (8 * hmax); mcusY = (H + 8 * vmax - 1) / (8 * vmax);
        for (int i = 0; i < ncomp; ++i) {
            Component& c = comp[i];
            c.blocksW = mcusX * c.h; c.blocksH = mcusY * c.v;
            const int sw = (W * c.h + hmax - 1) / hmax, sh = (H * c.v + vmax - 1) / vmax;
            c.ownW = (sw + 7) / 8; c.ownH = (sh + 7) / 8;
            c.coef.assign((size_t)c.blocksW * c.blocksH * 64, 0);
        }
        haveFrame = true;
        return true;
    }

    // ---- one block of one scan
    bool block_baseline(BitReader& br, Component& c, int16_t* b) {
        const int t = decode_symbol(br, dc[c.td]);
        if (t < 0 || t > 11) return false;
        c.pred += t ? extend(br.bits(t), t) : 0;
        b[0] = (int16_t)c.pred;
        for (int k = 1; k < 64;) {
            const int rs = decode_symbol(br, ac[c.ta]);
            if (rs < 0) return false;
            const int r = rs >> 4, s = rs & 15;
            if (s == 0) { if (r == 15) { k += 16; continue; } break; }
            k += r;
            if (k > 63) return false;
            b[kZigzag[k]] = (int16_t)extend(br.bits(s), s);
            ++k;
        }
        return true;
    }
    bool block_dc_progressive(BitReader& br, Component& c, int16_t* b, int ah, int al) {
        if (ah == 0) {
            const int t = decode_symbol(br, dc[c.td]);
            if (t < 0 || t > 11) return false;
            c.pred += t ? extend(br.bits(t), t) : 0;
            b[0] = (int16_t)(c.pred * (1 << al));
        } else if (br.bit()) b[0] = (int16_t)(b[0] | (1 << al));
        return true;
    }
    bool block_ac_first(BitReader& br, Component& c, int16_t* b, int ss, int se, int al) {
        if (eobrun > 0) { --eobrun; return true; }
        for (int k = ss; k <= se;) {
            const int rs = decode_symbol(br, ac[c.ta]);
            if (rs < 0) return false;
            const int r = rs >> 4, s = rs & 15;
            if (s == 0) {
                if (r < 15) { eobrun = (1 << r) - 1; if (r) eobrun += br.bits(r); break; }
                k += 16;
                continue;
            }
            k += r;
            if (k > se) return false;
            b[kZigzag[k]] = (int16_t)(extend(br.bits(s), s) * (1 << al));
            ++k;
        }
        return true;
    }
    bool block_ac_refine(BitReader& br, Component& c, int16_t* b, int ss, int se, int al) {
        const int p1 = 1 << al, m1 = -(1 << al);
        auto refine = [&](int16_t& v) {          // one correction bit for a coefficient that is already non-zero
            if (br.bit() && (v & p1) == 0) v = (int16_t)(v + (v >= 0 ? p1 : m1));
        };
        int k = ss;
        if (eobrun == 0) {
            for (; k <= se; ++k) {
                const int rs = decode_symbol(br, ac[c.ta]);
                if (rs < 0) return false;
                int r = rs >> 4;
                const int s = rs & 15;
                int val = 0;
                if (s) {
                    if (s != 1) return false;
                    val = br.bit() ? p1 : m1;
                } else if (r != 15) {
                    eobrun = 1 << r;
                    if (r) eobrun += br.bits(r);
                    break;
                }
                // skip r zero-history coefficients (refining the non-zero ones on the way); the new coefficient goes to the next zero one
                for (; k <= se; ++k) {
                    int16_t& v = b[kZigzag[k]];
                    if (v != 0) refine(v);
                    else if (--r < 0) break;
                }
                if (s && k <= se) b[kZigzag[k]] = (int16_t)val;
            }
        }
        if (eobrun > 0) {
            for (; k <= se; ++k) { int16_t& v = b[kZigzag[k]]; if (v != 0) refine(v); }
            --eobrun;
        }
        return true;
    }

    bool sos() {
        if (!haveFrame) return false;
        const int len = u16();
        const size_t endHdr = pos - 2 + (size_t)len;
        const int ns = u8();
        if (ns < 1 || ns > ncomp) return false;
        Component* sc[3];
        for (int i = 0; i < ns; ++i) {
            const int cs = u8(), t = u8();
            if (t < 0) return false;
            sc[i] = nullptr;
            for (int j = 0; j < ncomp; ++j) if (comp[j].id == cs) sc[i] = &comp[j];
            if (!sc[i]) return false;
            sc[i]->td = t >> 4; sc[i]->ta = t & 15;
            if (sc[i]->td > 3 || sc[i]->ta > 3) return false;
        }
        const int ss = u8(), se = u8(), a = u8();
        if (a < 0 || pos != endHdr) return false;
        const int ah = a >> 4, al = a & 15;
        if (progressive) {
            if (ss > se || se > 63 || (ss == 0 && se != 0) || (ss > 0 && ns != 1) || al > 13) return false;
        } else if (ss != 0 || se != 63 || a != 0) return false;
        for (int i = 0; i < ns; ++i) {
            const bool needDc = !progressive || (ss == 0 && ah == 0), needAc = !progressive || ss > 0;
            if ((needDc && !dc[sc[i]->td].present) || (needAc && !ac[sc[i]->ta].present)) return false;
            sc[i]->pred = 0;
        }
        eobrun = 0;
        BitReader br{d + pos, d + n};
        auto one = [&](Component& c, int bx, int by) -> bool {
            int16_t* b = c.coef.data() + ((size_t)by * c.blocksW + bx) * 64;
            if (!progressive) return block_baseline(br, c, b);
            if (ss == 0) return block_dc_progressive(br, c, b, ah, al);
            return ah == 0 ? block_ac_first(br, c, b, ss, se, al) : block_ac_refine(br, c, b, ss, se, al);
        };
        int untilRestart = restartInterval, nextRst = 0;
        auto restart = [&]() -> bool {        // between two restart intervals: byte-align, expect RSTn, reset the predictors
            if (!restartInterval || --untilRestart > 0) return true;
            br.cnt = 0; br.acc = 0;
            if (!br.marker) {                 // the marker has not been reached by the bit reader yet: find it
                while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) {
                    if (br.p[0] == 0xFF && br.p[1] != 0 && br.p[1] != 0xFF) return true;     // some other marker: the scan ends here
                    ++br.p;
                }
                if (br.p + 1 >= br.end) return true;
                br.marker = br.p[1]; br.p += 2;
            }
            if (br.marker != 0xD0 + nextRst) return br.marker >= 0xD0 && br.marker <= 0xD7 ? (br.reset(), true) : true;
            nextRst = (nextRst + 1) & 7;
            br.reset();
            for (int i = 0; i < ns; ++i) sc[i]->pred = 0;
            eobrun = 0;
            untilRestart = restartInterval;
            return true;
        };
        if (ns == 1) {        // non-interleaved: the component's own block grid, row by row
            Component& c = *sc[0];
            for (int by = 0; by < c.ownH; ++by)
                for (int bx = 0; bx < c.ownW; ++bx) {
                    if (!one(c, bx, by)) return false;
                    if (!(by == c.ownH - 1 && bx == c.ownW - 1)) restart();
                }
        } else {
            for (int my = 0; my < mcusY; ++my)
                for (int mx = 0; mx < mcusX; ++mx) {
                    for (int i = 0; i < ns; ++i)
                        for (int v = 0; v < sc[i]->v; ++v)
                            for (int h = 0; h < sc[i]->h; ++h)
                                if (!one(*sc[i], mx * sc[i]->h + h, my * sc[i]->v + v)) return false;
                    if (!(my == mcusY - 1 && mx == mcusX - 1)) restart();
                }
        }
        // continue after the entropy-coded data: at the marker the reader stopped at, or scan forward to the next one
        if (br.marker) pos = (size_t)(br.p - d) - 2;
        else {
            const uint8_t* q = br.p;
            while (q + 1 < d + n && !(q[0] == 0xFF && q[1] != 0 && q[1] != 0xFF && !(q[1] >= 0xD0 && q[1] <= 0xD7))) ++q;
            pos = (size_t)(q - d);
        }
        return true;
    }

    // ---- dequantisation + inverse DCT (T.81 A.3.3), double precision, one rounding
    void idct_all() {
        double cs[8][8];
        for (int x = 0; x < 8; ++x)
            for (int u = 0; u < 8; ++u) cs[x][u] = (u == 0 ? std::sqrt(0.125) : 0.5) * std::cos((2 * x + 1) * u * 3.14159265358979323846 / 16.0);
        for (int ci = 0; ci < ncomp; ++ci) {
            Component& c = comp[ci];
            const uint16_t* q = qt[c.tq];
            const int pw = c.blocksW * 8;
            c.plane.assign((size_t)pw * c.blocksH * 8, 0);
            for (int by = 0; by < c.blocksH; ++by)
                for (int bx = 0; bx < c.blocksW; ++bx) {
                    const int16_t* b = c.coef.data() + ((size_t)by * c.blocksW + bx) * 64;
                    double f[64], t[64];
                    bool acZero = true;
                    for (int i = 0; i < 64; ++i) { f[i] = (double)b[i] * q[i]; if (i && b[i]) acZero = false; }
                    uint8_t* out = c.plane.data() + (size_t)by * 8 * pw + bx * 8;
                    if (acZero) {
                        const int v = (int)std::floor(f[0] / 8.0 + 128.0 + 0.5);
                        const uint8_t s = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
                        for (int y = 0; y < 8; ++y) std::memset(out + (size_t)y * pw, s, 8);
                        continue;
                    }
                    for (int v = 0; v < 8; ++v)          // rows: t[v][x] = sum_u cs[x][u] f[v][u]
                        for (int x = 0; x < 8; ++x) {
                            double s = 0;
                            for (int u = 0; u < 8; ++u) s += cs[x][u] * f[v * 8 + u];
                            t[v * 8 + x] = s;
                        }
                    for (int y = 0; y < 8; ++y)
                        for (int x = 0; x < 8; ++x) {
                            double s = 0;
                            for (int v = 0; v < 8; ++v) s += cs[y][v] * t[v * 8 + x];
                            const int r = (int)std::floor(s + 128.0 + 0.5);
                            out[(size_t)y * pw + x] = (uint8_t)(r < 0 ? 0 : r > 255 ? 255 : r);
                        }
                }
        }
    }

    // component ci at full resolution, W x H
    void upsample(int ci, std::vector<uint8_t>& out) {
        const Component& c = comp[ci];
        const int hs = hmax / c.h, vs = vmax / c.v, pw = c.blocksW * 8;
        const int sw = (W * c.h + hmax - 1) / hmax, sh = (H * c.v + vmax - 1) / vmax;      // the component's own sample counts
        out.resize((size_t)W * H);
        auto at = [&](int x, int y) -> int {
            x = x < 0 ? 0 : x >= sw ? sw - 1 : x; y = y < 0 ? 0 : y >= sh ? sh - 1 : y;
            return c.plane[(size_t)y * pw + x];
        };
        if (hs == 1 && vs == 1) {
            for (int y = 0; y < H; ++y) std::memcpy(out.data() + (size_t)y * W, c.plane.data() + (size_t)y * pw, (size_t)W);
            return;
        }
        const bool triH = hs == 2, triV = vs == 2;
        for (int y = 0; y < H; ++y) {
            // vertical: nearest row (weight 3) and the next-nearest (weight 1) for a 2x ratio, else replication
            const int y0 = y / vs, y1 = triV ? ((y & 1) ? y0 + 1 : y0 - 1) : y0;
            const int wy0 = triV ? 3 : 1, wy1 = triV ? 1 : 0;
            for (int x = 0; x < W; ++x) {
                const int x0 = x / hs, x1 = triH ? ((x & 1) ? x0 + 1 : x0 - 1) : x0;
                const int wx0 = triH ? 3 : 1, wx1 = triH ? 1 : 0;
                const int sum = wy0 * (wx0 * at(x0, y0) + wx1 * at(x1, y0)) + wy1 * (wx0 * at(x0, y1) + wx1 * at(x1, y1));
                const int den = (wy0 + wy1) * (wx0 + wx1);
                out[(size_t)y * W + x] = (uint8_t)((sum + den / 2) / den);
            }
        }
    }

    bool run(std::vector<unsigned char>& rgba, uint32_t& outW, uint32_t& outH) {
        if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return false;
        pos = 2;
        bool sawScan = false;
        for (;;) {
            int m = u8();
            if (m < 0) break;
            if (m != 0xFF) continue;
            do { m = u8(); } while (m == 0xFF);
            if (m < 0) break;
            if (m == 0xD9) break;                                   // EOI
            if (m == 0 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
            if (m == 0xDA) { if (!sos()) return false; sawScan = true; continue; }
            const int len = u16();
            if (len < 2 || pos - 2 + (size_t)len > n) return false;
            const size_t endPos = pos - 2 + (size_t)len;
            if (m == 0xDB) { if (!dqt(endPos)) return false; }
            else if (m == 0xC4) { if (!dht(endPos)) return false; }
            else if (m == 0xC0 || m == 0xC1 || m == 0xC2) { if (!sof(m)) return false; }
            else if (m == 0xC3 || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC)) return false;      // lossless, hierarchical, arithmetic
            else if (m == 0xCC) return false;                       // arithmetic conditioning
            else if (m == 0xDD) { restartInterval = u16(); if (restartInterval < 0) return false; }
            else if (m == 0xEE && len >= 14 && std::memcmp(d + pos, "Adobe", 5) == 0) adobeTransform = d[pos + 11];
            pos = endPos;
        }
        if (!haveFrame || !sawScan) return false;
        for (int i = 0; i < ncomp; ++i) if (!haveQt[comp[i].tq]) return false;
        idct_all();
        rgba.resize((size_t)W * H * 4);
        outW = (uint32_t)W; outH = (uint32_t)H;
        std::vector<uint8_t> p0, p1, p2;
        upsample(0, p0);
        if (ncomp == 1) {
            for (size_t i = 0; i < (size_t)W * H; ++i) { rgba[4 * i] = rgba[4 * i + 1] = rgba[4 * i + 2] = p0[i]; rgba[4 * i + 3] = 255; }
            return true;
        }
        upsample(1, p1); upsample(2, p2);
        const bool isRgb = adobeTransform == 0 || (adobeTransform < 0 && comp[0].id == 'R' && comp[1].id == 'G' && comp[2].id == 'B');
        auto clamp8 = [](double v) { const int r = (int)std::floor(v + 0.5); return (unsigned char)(r < 0 ? 0 : r > 255 ? 255 : r); };
        for (size_t i = 0; i < (size_t)W * H; ++i) {
            if (isRgb) { rgba[4 * i] = p0[i]; rgba[4 * i + 1] = p1[i]; rgba[4 * i + 2] = p2[i]; }
            else {      // JFIF: full-range BT.601
                const double y = p0[i], cb = p1[i] - 128.0, cr = p2[i] - 128.0;
                rgba[4 * i] = clamp8(y + 1.402 * cr);
                rgba[4 * i + 1] = clamp8(y - 0.344136 * cb - 0.714136 * cr);
                rgba[4 * i + 2] = clamp8(y + 1.772 * cb);
            }
            rgba[4 * i + 3] = 255;
        }
        return true;
    }
};

}  // namespace

bool decode_jpeg(const unsigned char* data, size_t size, std::vector<unsigned char>& rgba, uint32_t& width, uint32_t& height) {
    if (!data || size < 4) return false;
    Decoder dec{data, size};
    return dec.run(rgba, width, height);
}
