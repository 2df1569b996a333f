// vk_gltf.cpp — see vk_gltf.h.  Own JSON reader, base64, GLB container, PNG decoder (zlib), accessor conversion (JPEG: vk_jpeg.cpp).
#include "vk_gltf.h"

#include <zlib.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <utility>

using namespace vkm;

namespace {

// ------------------------------------------------------------------------------------------------- JSON
struct Json {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    const Json* find(const char* key) const {
        if (type != Object) return nullptr;
        for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool has(const char* key) const { return find(key) != nullptr; }
    double number(const char* key, double def) const { const Json* j = find(key); return j && j->type == Number ? j->num : def; }
    long integer(const char* key, long def) const { const Json* j = find(key); return j && j->type == Number ? (long)j->num : def; }
    std::string string(const char* key, const std::string& def = "") const { const Json* j = find(key); return j && j->type == String ? j->str : def; }
    const std::vector<Json>& array(const char* key) const {
        static const std::vector<Json> empty;
        const Json* j = find(key);
        return j && j->type == Array ? j->arr : empty;
    }
};

struct JsonParser {
    const char* p; const char* end; bool ok = true;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
    bool lit(const char* s) { size_t n = std::strlen(s); if ((size_t)(end - p) >= n && std::memcmp(p, s, n) == 0) { p += n; return true; } return false; }
    static void utf8(std::string& out, unsigned cp) {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 63)); }
        else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 63)); out += (char)(0x80 | (cp & 63)); }
        else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 63)); out += (char)(0x80 | ((cp >> 6) & 63)); out += (char)(0x80 | (cp & 63)); }
    }
    unsigned hex4() {
        unsigned v = 0;
        for (int i = 0; i < 4; ++i) {
            if (p >= end) { ok = false; return 0; }
            char c = *p++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= c - '0';
            else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
            else { ok = false; return 0; }
        }
        return v;
    }
    std::string parseString() {
        std::string s;
        ++p;   // opening quote
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) break;
                char c = *p++;
                switch (c) {
                    case 'n': s += '\n'; break; case 't': s += '\t'; break; case 'r': s += '\r'; break;
                    case 'b': s += '\b'; break; case 'f': s += '\f'; break;
                    case 'u': {
                        unsigned cp = hex4();
                        if (cp >= 0xD800 && cp < 0xDC00 && lit("\\u")) { unsigned lo = hex4(); cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); }
                        utf8(s, cp);
                        break;
                    }
                    default: s += c;   // \" \\ \/
                }
            } else s += *p++;
        }
        if (p >= end) ok = false; else ++p;
        return s;
    }
    Json parse(int depth = 0) {
        Json j;
        ws();
        if (p >= end || depth > 64) { ok = false; return j; }
        if (*p == '{') {
            j.type = Json::Object; ++p; ws();
            if (p < end && *p == '}') { ++p; return j; }
            while (ok) {
                ws();
                if (p >= end || *p != '"') { ok = false; break; }
                std::string key = parseString();
                ws();
                if (p >= end || *p != ':') { ok = false; break; }
                ++p;
                j.obj.emplace_back(std::move(key), parse(depth + 1));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; break; }
                ok = false;
            }
        } else if (*p == '[') {
            j.type = Json::Array; ++p; ws();
            if (p < end && *p == ']') { ++p; return j; }
            while (ok) {
                j.arr.push_back(parse(depth + 1));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; break; }
                ok = false;
            }
        } else if (*p == '"') { j.type = Json::String; j.str = parseString(); }
        else if (lit("true")) { j.type = Json::Bool; j.b = true; }
        else if (lit("false")) { j.type = Json::Bool; j.b = false; }
        else if (lit("null")) { j.type = Json::Null; }
        else {
            char* e = nullptr;
            std::string tmp(p, (size_t)std::min<ptrdiff_t>(end - p, 64));
            j.num = std::strtod(tmp.c_str(), &e);
            if (e == tmp.c_str()) { ok = false; return j; }
            p += e - tmp.c_str();
            j.type = Json::Number;
        }
        return j;
    }
};

// ------------------------------------------------------------------------------------------------- bytes
bool read_file(const std::string& path, std::vector<unsigned char>& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return false; }
    out.resize((size_t)n);
    size_t got = n ? std::fread(out.data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == (size_t)n;
}

bool base64_decode(const char* s, size_t n, std::vector<unsigned char>& out) {
    static int8_t T[256];
    static bool init = false;
    if (!init) {
        std::memset(T, -1, sizeof(T));
        const char* A = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
        for (int i = 0; i < 64; ++i) T[(unsigned char)A[i]] = (int8_t)i;
        init = true;
    }
    out.clear();
    out.reserve(n * 3 / 4);
    uint32_t acc = 0; int bits = 0;
    for (size_t i = 0; i < n; ++i) {
        unsigned char c = (unsigned char)s[i];
        if (c == '=' || c == '\n' || c == '\r') continue;
        if (T[c] < 0) return false;
        acc = (acc << 6) | (uint32_t)T[c]; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((unsigned char)((acc >> bits) & 0xFF)); }
    }
    return true;
}

std::string percent_decode(const std::string& s) {
    std::string o;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] == '%' && i + 2 < s.size() + 0 && std::isxdigit((unsigned char)s[i + 1]) && std::isxdigit((unsigned char)s[i + 2])) {
            o += (char)std::strtol(s.substr(i + 1, 2).c_str(), nullptr, 16);
            i += 2;
        } else o += s[i];
    }
    return o;
}

std::string parent_path(const std::string& file) {
    size_t k = file.find_last_of("/\\");
    return k == std::string::npos ? std::string(".") : file.substr(0, k);
}

// A URI as glTF uses it: data:...;base64,xxxx or a path relative to the .gltf.
bool load_uri(const std::string& uri, const std::string& baseDir, std::vector<unsigned char>& out) {
    if (uri.rfind("data:", 0) == 0) {
        size_t k = uri.find(";base64,");
        if (k == std::string::npos) return false;
        return base64_decode(uri.data() + k + 8, uri.size() - k - 8, out);
    }
    return read_file(baseDir + "/" + percent_decode(uri), out);
}

}  // namespace

// ------------------------------------------------------------------------------------------------- PNG
bool decode_png(const unsigned char* d, size_t n, std::vector<unsigned char>& rgba, uint32_t& W, uint32_t& H) {
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (n < 8 + 25 || std::memcmp(d, sig, 8) != 0) return false;
    auto be32 = [](const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; };
    size_t pos = 8;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    bool haveHdr = false;
    while (pos + 12 <= n) {
        uint32_t len = be32(d + pos);
        const unsigned char* tag = d + pos + 4;
        const unsigned char* body = d + pos + 8;
        if (pos + 12 + (size_t)len > n) return false;
        if (!std::memcmp(tag, "IHDR", 4) && len >= 13) {
            W = be32(body); H = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
            haveHdr = true;
        } else if (!std::memcmp(tag, "PLTE", 4)) plte.assign(body, body + len);
        else if (!std::memcmp(tag, "tRNS", 4)) trns.assign(body, body + len);
        else if (!std::memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!std::memcmp(tag, "IEND", 4)) break;
        pos += 12 + (size_t)len;
    }
    if (!haveHdr || W == 0 || H == 0 || W > 16384 || H > 16384 || interlace > 1) return false;
    if (depth != 8 && depth != 16 && !(ctype == 3 || ctype == 0)) return false;
    int channels;
    switch (ctype) { case 0: channels = 1; break; case 2: channels = 3; break; case 3: channels = 1; break; case 4: channels = 2; break; case 6: channels = 4; break; default: return false; }
    if ((ctype == 3 && (depth > 8 || plte.empty())) || (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16)) return false;
    const size_t bpp = std::max<size_t>(1, (size_t)channels * depth / 8);        // filter unit
    // The image is one pass, or (interlace method 1, Adam7 — stb_image reads those too) seven reduced images stored one after
    // the other: pass p holds the pixels (x0 + i * dx, y0 + j * dy), each with its own scan lines and filters.
    struct Pass { uint32_t x0, y0, dx, dy; };
    static const Pass adam7[7] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
    static const Pass whole = {0, 0, 1, 1};
    const int nPasses = interlace ? 7 : 1;
    auto pass_dim = [](uint32_t full, uint32_t o, uint32_t step) { return full > o ? (full - o + step - 1) / step : 0u; };
    size_t rawSize = 0;
    for (int p = 0; p < nPasses; ++p) {
        const Pass& ps = interlace ? adam7[p] : whole;
        const uint32_t pw = pass_dim(W, ps.x0, ps.dx), ph = pass_dim(H, ps.y0, ps.dy);
        if (pw && ph) rawSize += (((size_t)pw * channels * depth + 7) / 8 + 1) * ph;
    }
    std::vector<unsigned char> raw(rawSize);
    uLongf rawLen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rawLen, idat.data(), (uLong)idat.size()) != Z_OK || rawLen != raw.size()) return false;
    rgba.resize((size_t)W * H * 4);
    size_t passBase = 0;
    for (int p = 0; p < nPasses; ++p) {
        const Pass& ps = interlace ? adam7[p] : whole;
        const uint32_t pw = pass_dim(W, ps.x0, ps.dx), ph = pass_dim(H, ps.y0, ps.dy);
        if (!pw || !ph) continue;
        const size_t stride = ((size_t)pw * channels * depth + 7) / 8;
        unsigned char* base = raw.data() + passBase;
        passBase += (stride + 1) * ph;
        // undo the scanline filters in place
        std::vector<unsigned char> zero(stride, 0);
        for (uint32_t y = 0; y < ph; ++y) {
            unsigned char* row = base + (size_t)y * (stride + 1);
            const int ft = row[0];
            unsigned char* cur = row + 1;
            const unsigned char* up = y ? base + (size_t)(y - 1) * (stride + 1) + 1 : zero.data();
            for (size_t i = 0; i < stride; ++i) {
                const int a = i >= bpp ? cur[i - bpp] : 0, b = up[i], c = i >= bpp ? up[i - bpp] : 0;
                int pred;
                switch (ft) {
                    case 0: pred = 0; break;
                    case 1: pred = a; break;
                    case 2: pred = b; break;
                    case 3: pred = (a + b) >> 1; break;
                    case 4: { const int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c);
                              pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
                    default: return false;
                }
                cur[i] = (unsigned char)(cur[i] + pred);
            }
        }
        for (uint32_t y = 0; y < ph; ++y) {
            const unsigned char* row = base + (size_t)y * (stride + 1) + 1;
            for (uint32_t x = 0; x < pw; ++x) {
                unsigned char* o = &rgba[((size_t)(ps.y0 + y * ps.dy) * W + (ps.x0 + x * ps.dx)) * 4];
                auto sample = [&](int ch) -> unsigned {          // channel value reduced to 8 bits the way stb_image does (high byte)
                    if (depth == 16) return row[((size_t)x * channels + ch) * 2];
                    if (depth == 8) return row[(size_t)x * channels + ch];
                    const size_t bit = ((size_t)x * channels + ch) * depth;
                    const unsigned v = (row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1);
                    return ctype == 3 ? v : v * 255u / ((1u << depth) - 1);
                };
                switch (ctype) {
                    case 0: {
                        const unsigned g = sample(0);
                        o[0] = o[1] = o[2] = (unsigned char)g; o[3] = 255;
                        if (trns.size() >= 2 && depth <= 8 && g == (unsigned)(trns[1] * (depth < 8 ? 255u / ((1u << depth) - 1) : 1u))) o[3] = 0;
                        break;
                    }
                    case 2: o[0] = (unsigned char)sample(0); o[1] = (unsigned char)sample(1); o[2] = (unsigned char)sample(2); o[3] = 255;
                            if (trns.size() >= 6 && depth == 8 && o[0] == trns[1] && o[1] == trns[3] && o[2] == trns[5]) o[3] = 0;
                            break;
                    case 3: { unsigned idx = sample(0);
                              if (idx * 3 + 2 >= plte.size()) { o[0] = o[1] = o[2] = 0; o[3] = 255; break; }
                              o[0] = plte[idx * 3]; o[1] = plte[idx * 3 + 1]; o[2] = plte[idx * 3 + 2]; o[3] = idx < trns.size() ? trns[idx] : 255; break; }
                    case 4: { unsigned g = sample(0); o[0] = o[1] = o[2] = (unsigned char)g; o[3] = (unsigned char)sample(1); break; }
                    default: o[0] = (unsigned char)sample(0); o[1] = (unsigned char)sample(1); o[2] = (unsigned char)sample(2); o[3] = (unsigned char)sample(3);
                }
            }
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------------- glTF
namespace {

struct Document {
    Json root;
    std::string baseDir;
    std::vector<std::vector<unsigned char>> buffers;
    std::vector<unsigned char> glbBin;
    std::deque<std::vector<unsigned char>> dense;      // sparse accessors made dense (a deque: AccessorView keeps pointers into them)
    std::string err;
};

bool fail(Document& d, const std::string& why) { if (d.err.empty()) d.err = why; return false; }

// Accessor element i, component c as double-free arithmetic: the source component is read in its own type and converted
// per fastgltf's convertComponent (tools.hpp:126-149).
struct AccessorView {
    const unsigned char* base = nullptr;
    size_t stride = 0, count = 0;
    int componentType = 0, components = 0;
    bool normalized = false;

    template <class Src> static float toFloat(Src v, bool normalized) {
        if (normalized && std::numeric_limits<Src>::is_integer)
            return std::fmax((float)v / (float)std::numeric_limits<Src>::max(), -1.0f);
        return (float)v;
    }
    float f(size_t i, int c) const {
        const unsigned char* p = base + i * stride;
        switch (componentType) {
            case 5120: { int8_t v; std::memcpy(&v, p + c, 1); return toFloat(v, normalized); }
            case 5121: { uint8_t v; std::memcpy(&v, p + c, 1); return toFloat(v, normalized); }
            case 5122: { int16_t v; std::memcpy(&v, p + 2 * c, 2); return toFloat(v, normalized); }
            case 5123: { uint16_t v; std::memcpy(&v, p + 2 * c, 2); return toFloat(v, normalized); }
            case 5125: { uint32_t v; std::memcpy(&v, p + 4 * c, 4); return toFloat(v, normalized); }
            default: { float v; std::memcpy(&v, p + 4 * c, 4); return v; }
        }
    }
    uint32_t u(size_t i) const {
        const unsigned char* p = base + i * stride;
        switch (componentType) {
            case 5121: return *p;
            case 5123: { uint16_t v; std::memcpy(&v, p, 2); return v; }
            case 5125: { uint32_t v; std::memcpy(&v, p, 4); return v; }
            case 5120: { int8_t v; std::memcpy(&v, p, 1); return (uint32_t)v; }
            case 5122: { int16_t v; std::memcpy(&v, p, 2); return (uint32_t)v; }
            default: { float v; std::memcpy(&v, p, 4); return (uint32_t)v; }
        }
    }
};

int component_size(int t) { return (t == 5120 || t == 5121) ? 1 : (t == 5122 || t == 5123) ? 2 : (t == 5125 || t == 5126) ? 4 : 0; }
int type_components(const std::string& t) {
    static const std::pair<const char*, int> table[] = {{"SCALAR", 1}, {"VEC2", 2}, {"VEC3", 3}, {"VEC4", 4}, {"MAT2", 4}, {"MAT3", 9}, {"MAT4", 16}};
    for (const auto& e : table)
        if (t == e.first) return e.second;
    return 0;
}

bool view_bytes(Document& d, long viewIndex, const unsigned char** ptr, size_t* len, size_t* byteStride) {
    const auto& views = d.root.array("bufferViews");
    if (viewIndex < 0 || (size_t)viewIndex >= views.size()) return fail(d, "bufferView index out of range");
    const Json& v = views[(size_t)viewIndex];
    const long buf = v.integer("buffer", -1);
    if (buf < 0 || (size_t)buf >= d.buffers.size()) return fail(d, "buffer index out of range");
    const size_t off = (size_t)v.integer("byteOffset", 0), n = (size_t)v.integer("byteLength", 0);
    if (off + n > d.buffers[(size_t)buf].size()) return fail(d, "bufferView exceeds its buffer");
    *ptr = d.buffers[(size_t)buf].data() + off;
    *len = n;
    if (byteStride) *byteStride = (size_t)v.integer("byteStride", 0);
    return true;
}

bool accessor_view(Document& d, long index, AccessorView& out) {
    const auto& accs = d.root.array("accessors");
    if (index < 0 || (size_t)index >= accs.size()) return fail(d, "accessor index out of range");
    const Json& a = accs[(size_t)index];
    out.componentType = (int)a.integer("componentType", 0);
    out.components = type_components(a.string("type"));
    out.count = (size_t)a.integer("count", 0);
    const Json* nrm = a.find("normalized");
    out.normalized = nrm && nrm->type == Json::Bool && nrm->b;
    const int cs = component_size(out.componentType);
    if (!cs || !out.components) return fail(d, "accessor has an unknown component type or element type");
    const size_t elem = (size_t)cs * out.components;
    const Json* sparse = a.find("sparse");
    if (!a.has("bufferView") && !sparse) return fail(d, "accessor without a bufferView");
    const unsigned char* p = nullptr; size_t len = 0, bs = 0, off = 0;
    if (a.has("bufferView")) {
        if (!view_bytes(d, a.integer("bufferView", -1), &p, &len, &bs)) return false;
        off = (size_t)a.integer("byteOffset", 0);
        out.stride = bs ? bs : elem;
        if (out.count && off + (out.count - 1) * out.stride + elem > len) return fail(d, "accessor exceeds its bufferView");
        out.base = p + off;
    }
    if (!sparse) return true;
    // A sparse accessor (glTF 2.0, 3.6.2.3; fastgltf's iterateAccessor resolves it the same way): the elements of the buffer view —
    // zeros without one — with `count` of them replaced: element indices[k] takes values[k].  Made dense once, tightly packed.
    if (sparse->type != Json::Object) return fail(d, "accessor.sparse is not an object");
    d.dense.emplace_back(out.count * elem, (unsigned char)0);
    std::vector<unsigned char>& dense = d.dense.back();
    if (out.base)
        for (size_t i = 0; i < out.count; ++i) std::memcpy(dense.data() + i * elem, out.base + i * out.stride, elem);
    const size_t n = (size_t)sparse->integer("count", 0);
    const Json* ji = sparse->find("indices");
    const Json* jv = sparse->find("values");
    if (!ji || !jv || ji->type != Json::Object || jv->type != Json::Object) return fail(d, "sparse accessor without indices or values");
    const int ict = (int)ji->integer("componentType", 0);
    const size_t ics = ict == 5121 ? 1 : ict == 5123 ? 2 : ict == 5125 ? 4 : 0;
    if (!ics) return fail(d, "sparse accessor indices must be unsigned bytes, shorts or ints");
    const unsigned char *ip, *vp; size_t ilen, vlen;
    if (!view_bytes(d, ji->integer("bufferView", -1), &ip, &ilen, nullptr) || !view_bytes(d, jv->integer("bufferView", -1), &vp, &vlen, nullptr)) return false;
    const size_t ioff = (size_t)ji->integer("byteOffset", 0), voff = (size_t)jv->integer("byteOffset", 0);
    if (ioff + n * ics > ilen || voff + n * elem > vlen) return fail(d, "sparse accessor exceeds its bufferViews");
    for (size_t k = 0; k < n; ++k) {
        size_t idx = 0;
        if (ics == 1) idx = ip[ioff + k];
        else if (ics == 2) { uint16_t v; std::memcpy(&v, ip + ioff + 2 * k, 2); idx = v; }
        else { uint32_t v; std::memcpy(&v, ip + ioff + 4 * k, 4); idx = v; }
        if (idx >= out.count) return fail(d, "sparse accessor index out of range");
        std::memcpy(dense.data() + idx * elem, vp + voff + k * elem, elem);
    }
    out.base = dense.data();
    out.stride = elem;
    return true;
}

bool open_document(const std::string& fileName, Document& d) {
    std::vector<unsigned char> file;
    if (!read_file(fileName, file)) return fail(d, "cannot read " + fileName);
    d.baseDir = parent_path(fileName);
    const char* json = nullptr; size_t jsonLen = 0;
    if (file.size() >= 12 && !std::memcmp(file.data(), "glTF", 4)) {           // GLB: header + JSON chunk [+ BIN chunk]
        auto le32 = [&](size_t o) { uint32_t v; std::memcpy(&v, file.data() + o, 4); return v; };
        if (le32(4) != 2) return fail(d, "GLB version is not 2");
        size_t pos = 12;
        while (pos + 8 <= file.size()) {
            const uint32_t len = le32(pos), type = le32(pos + 4);
            if (pos + 8 + (size_t)len > file.size()) return fail(d, "GLB chunk exceeds the file");
            if (type == 0x4E4F534A) { json = (const char*)file.data() + pos + 8; jsonLen = len; }
            else if (type == 0x004E4942 && d.glbBin.empty()) d.glbBin.assign(file.data() + pos + 8, file.data() + pos + 8 + len);
            pos += 8 + (size_t)len;
        }
        if (!json) return fail(d, "GLB without a JSON chunk");
    } else { json = (const char*)file.data(); jsonLen = file.size(); }
    JsonParser jp{json, json + jsonLen};
    d.root = jp.parse();
    if (!jp.ok || d.root.type != Json::Object) return fail(d, "failed to parse glTF");      // "filed to parse glTF" (:246)
    const auto& bufs = d.root.array("buffers");
    d.buffers.resize(bufs.size());
    for (size_t i = 0; i < bufs.size(); ++i) {
        const std::string uri = bufs[i].string("uri");
        if (uri.empty()) {
            if (i == 0 && !d.glbBin.empty()) d.buffers[i] = d.glbBin;      // Options::LoadGLBBuffers
            else return fail(d, "buffer without a uri");
        } else if (!load_uri(uri, d.baseDir, d.buffers[i])) return fail(d, "cannot load buffer " + uri.substr(0, 64));   // LoadExternalBuffers
        if (d.buffers[i].size() < (size_t)bufs[i].integer("byteLength", 0)) return fail(d, "buffer shorter than its byteLength");
    }
    return true;
}

long texture_image(const Document& d, const Json* texInfo) {      // textureInfo -> textures[index].source
    if (!texInfo || texInfo->type != Json::Object) return -1;
    const long t = texInfo->integer("index", -1);
    const auto& textures = d.root.array("textures");
    if (t < 0 || (size_t)t >= textures.size()) return -1;
    return textures[(size_t)t].integer("source", -1);
}

}  // namespace

// Test hook: `count` tightly packed scalar components of glTF component type `componentType` through the accessor conversion
// load_gltf uses for every attribute (fastgltf's convertComponent, external/fastgltf/include/fastgltf/tools.hpp:126-149).
void gltf_convert_components(int componentType, bool normalized, const unsigned char* data, size_t count, float* out) {
    AccessorView v;
    v.base = data; v.count = count; v.componentType = componentType; v.components = 1; v.normalized = normalized;
    v.stride = (componentType == 5120 || componentType == 5121) ? 1 : ((componentType == 5122 || componentType == 5123) ? 2 : 4);
    for (size_t i = 0; i < count; ++i) out[i] = v.f(i, 0);
}

std::shared_ptr<Node> load_gltf(VulkanEngine* engine, const std::string& name, const std::string& fileName, std::string* error) {
    Document d;
    auto bail = [&](const std::string& why) -> std::shared_ptr<Node> { if (error) *error = d.err.empty() ? why : d.err; return nullptr; };
    if (!open_document(fileName, d)) return bail("");
    const Json& gltf = d.root;

    // ---- images -> bindless textures (src/vk_loader.cpp:275-329)
    std::vector<int> textureIndices;
    for (const Json& image : gltf.array("images")) {
        std::vector<unsigned char> bytes, rgba;
        uint32_t w = 0, h = 0;
        bool ok = false;
        if (image.has("uri")) ok = load_uri(image.string("uri"), d.baseDir, bytes);
        else if (image.has("bufferView")) {
            const unsigned char* p; size_t len;
            if (view_bytes(d, image.integer("bufferView", -1), &p, &len, nullptr)) { bytes.assign(p, p + len); ok = true; }
            else d.err.clear();
        }
        ok = ok && (decode_png(bytes.data(), bytes.size(), rgba, w, h) || decode_jpeg(bytes.data(), bytes.size(), rgba, w, h));      // stbi_load (src/vk_loader.cpp:280,293)
        if (ok) textureIndices.push_back(engine->add_texture(rgba.data(), w, h));
        else {   // _errorCheckerboardImage (src/vk_engine_init.cpp:329-341): 16x16 magenta / black, UNORM, no mips, LINEAR sampler here
            uint32_t px[16 * 16];
            for (int x = 0; x < 16; ++x) for (int y = 0; y < 16; ++y) px[y * 16 + x] = ((x % 2) ^ (y % 2)) ? 0xFFFF00FFu : 0xFF000000u;
            VkrImage img = engine->create_image(px, 16, 16, VKR_FORMAT_R8G8B8A8_UNORM, false);
            const int id = engine->_globalTextureIndex++;
            engine->update_bindless_texture(id, img, VKR_SAMPLER_LINEAR);
            textureIndices.push_back(id);
        }
    }
    auto bindless = [&](long imageIndex, int def) { return imageIndex >= 0 && (size_t)imageIndex < textureIndices.size() ? textureIndices[(size_t)imageIndex] : def; };

    // ---- materials (:331-371)
    std::vector<MaterialInstance*> materials;
    for (const Json& mat : gltf.array("materials")) {
        MaterialConstants c{};
        const Json* pbr = mat.find("pbrMetallicRoughness");
        vec4 base{1, 1, 1, 1};
        float metallic = 1.0f, roughness = 1.0f;
        const Json* baseTex = nullptr; const Json* mrTex = nullptr;
        if (pbr && pbr->type == Json::Object) {
            const auto& f = pbr->array("baseColorFactor");
            if (f.size() == 4) base = vec4{(float)f[0].num, (float)f[1].num, (float)f[2].num, (float)f[3].num};
            metallic = (float)pbr->number("metallicFactor", 1.0);
            roughness = (float)pbr->number("roughnessFactor", 1.0);
            baseTex = pbr->find("baseColorTexture");
            mrTex = pbr->find("metallicRoughnessTexture");
        }
        c.colorFactors = base;
        c.metal_rough_factors = vec4{metallic, roughness, 0, 0};
        c.albedoID = bindless(texture_image(d, baseTex), 0);
        c.normalID = bindless(texture_image(d, mrTex), 1);                       // the reference's swap (:353-357)
        c.metalRoughID = bindless(texture_image(d, mat.find("normalTexture")), 0);   // (:359-363)
        const char* tmpl = mat.string("alphaMode") == "BLEND" ? "Transparent" : "Opaque";
        MaterialInstance* inst = engine->_materialSystem.build_instance(engine->_materialSystem.get_template(tmpl), c);
        if (!inst) return bail("build_instance failed");
        materials.push_back(inst);
    }
    if (materials.empty()) {   // (:373-379) value-initialised constants: every factor and id zero
        MaterialInstance* inst = engine->_materialSystem.build_instance(engine->_materialSystem.get_template("Opaque"), MaterialConstants{});
        if (!inst) return bail("build_instance failed");
        materials.push_back(inst);
    }

    // ---- meshes (:381-466)
    std::vector<std::shared_ptr<MeshAsset>> meshes;
    std::vector<uint32_t> indices;
    std::vector<Vertex> vertices;
    const auto& gmeshes = gltf.array("meshes");
    for (size_t i = 0; i < gmeshes.size(); ++i) {
        const Json& mesh = gmeshes[i];
        const std::string meshName = mesh.string("name");
        const std::string uniqueName = name + "_" + (meshName.empty() ? std::to_string(i) : meshName);
        indices.clear(); vertices.clear();
        std::vector<GeoSurface> surfaces;
        for (const Json& p : mesh.array("primitives")) {
            if (!p.has("indices")) continue;                                    // (:399-403)
            const long mode = p.integer("mode", 4);
            if (mode != 4) return bail("only TRIANGLES primitives are supported");
            const Json* attrs = p.find("attributes");
            if (!attrs || !attrs->has("POSITION")) return bail("primitive without POSITION");
            AccessorView ia, pa;
            if (!accessor_view(d, p.integer("indices", -1), ia) || !accessor_view(d, attrs->integer("POSITION", -1), pa)) return bail("");
            if (pa.components != 3 || pa.componentType != 5126) return bail("POSITION must be float VEC3");
            GeoSurface s{};
            s.startIndex = (uint32_t)indices.size();
            s.count = (uint32_t)ia.count;
            const size_t initial = vertices.size();
            indices.reserve(indices.size() + ia.count);
            for (size_t k = 0; k < ia.count; ++k) {
                const uint32_t idx = ia.u(k);
                if (idx >= pa.count) return bail("index exceeds the vertex count");
                indices.push_back(idx + (uint32_t)initial);
            }
            vertices.resize(initial + pa.count);
            for (size_t k = 0; k < pa.count; ++k) {                              // defaults (:417-421)
                Vertex v{};
                v.position = vec3{pa.f(k, 0), pa.f(k, 1), pa.f(k, 2)};
                v.normal = vec3{1, 0, 0};
                v.color = vec4{1, 1, 1, 1};
                v.uv_x = 0; v.uv_y = 0;
                vertices[initial + k] = v;
            }
            AccessorView av;
            if (attrs->has("NORMAL")) {
                if (!accessor_view(d, attrs->integer("NORMAL", -1), av)) return bail("");
                if (av.components != 3 || av.count != pa.count) return bail("NORMAL must be VEC3, one per vertex");
                for (size_t k = 0; k < av.count; ++k) vertices[initial + k].normal = vec3{av.f(k, 0), av.f(k, 1), av.f(k, 2)};
            }
            if (attrs->has("TEXCOORD_0")) {
                if (!accessor_view(d, attrs->integer("TEXCOORD_0", -1), av)) return bail("");
                if (av.components != 2 || av.count != pa.count) return bail("TEXCOORD_0 must be VEC2, one per vertex");
                for (size_t k = 0; k < av.count; ++k) { vertices[initial + k].uv_x = av.f(k, 0); vertices[initial + k].uv_y = av.f(k, 1); }
            }
            if (attrs->has("COLOR_0")) {
                if (!accessor_view(d, attrs->integer("COLOR_0", -1), av)) return bail("");
                if ((av.components != 4 && av.components != 3) || av.count != pa.count) return bail("COLOR_0 must be VEC3 or VEC4, one per vertex");
                for (size_t k = 0; k < av.count; ++k)
                    vertices[initial + k].color = vec4{av.f(k, 0), av.f(k, 1), av.f(k, 2), av.components == 4 ? av.f(k, 3) : 1.0f};
            }
            const long mi = p.integer("material", -1);
            if (mi >= (long)materials.size()) return bail("material index out of range");
            s.material = mi >= 0 ? materials[(size_t)mi] : materials[0];        // (:451-455)
            surfaces.push_back(s);
        }
        meshes.push_back(engine->add_mesh(uniqueName, indices, vertices, surfaces));
    }

    // ---- nodes (:468-515)
    std::vector<std::shared_ptr<Node>> nodes;
    const auto& gnodes = gltf.array("nodes");
    for (const Json& n : gnodes) {
        auto node = std::make_shared<Node>();
        const auto& m = n.array("matrix");
        if (m.size() == 16) {
            for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) node->localTransform.c[c][r] = (float)m[(size_t)c * 4 + r].num;   // column-major
        } else {
            vec3 tl{0, 0, 0}, sc{1, 1, 1};
            quat rot{1, 0, 0, 0};
            const auto& t = n.array("translation"); const auto& r = n.array("rotation"); const auto& s = n.array("scale");
            if (t.size() == 3) tl = vec3{(float)t[0].num, (float)t[1].num, (float)t[2].num};
            if (r.size() == 4) rot = quat{(float)r[3].num, (float)r[0].num, (float)r[1].num, (float)r[2].num};   // glTF xyzw -> glm wxyz
            if (s.size() == 3) sc = vec3{(float)s[0].num, (float)s[1].num, (float)s[2].num};
            node->localTransform = translate(tl) * toMat4(rot) * scale(sc);
        }
        const long mi = n.integer("mesh", -1);
        if (mi >= (long)meshes.size()) return bail("node references an unknown mesh");
        if (mi >= 0) node->mesh = meshes[(size_t)mi].get();
        nodes.push_back(node);
    }
    for (size_t i = 0; i < gnodes.size(); ++i)
        for (const Json& c : gnodes[i].array("children")) {
            const long ci = (long)c.num;
            if (ci < 0 || (size_t)ci >= nodes.size() || (size_t)ci == i) return bail("bad child index");
            nodes[i]->addChild(nodes[(size_t)ci]);
        }
    auto topNode = std::make_shared<Node>();
    for (auto& node : nodes)
        if (node->parent.expired()) topNode->addChild(node);
    return topNode;
}
