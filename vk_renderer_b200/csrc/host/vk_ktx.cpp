// vk_ktx.cpp — see vk_ktx.h.  Layouts follow the Khronos KTX File Format Specification 1.1 and 2.0.
#include <dlfcn.h>
#include <zlib.h>
#include "vk_ktx.h"

#include <algorithm>
#include <cstdio>
#include <cstring>

namespace {

const uint8_t kKtx1Id[12] = {0xAB, 0x4B, 0x54, 0x58, 0x20, 0x31, 0x31, 0xBB, 0x0D, 0x0A, 0x1A, 0x0A};
const uint8_t kKtx2Id[12] = {0xAB, 0x4B, 0x54, 0x58, 0x20, 0x32, 0x30, 0xBB, 0x0D, 0x0A, 0x1A, 0x0A};

// OpenGL enums of the KTX 1 header
const uint32_t GL_UNSIGNED_BYTE = 0x1401, GL_HALF_FLOAT = 0x140B, GL_RGBA = 0x1908;
const uint32_t GL_RGBA8 = 0x8058, GL_SRGB8_ALPHA8 = 0x8C43, GL_RGBA16F = 0x881A;
// VkFormat values of the KTX 2 header
const uint32_t VK_FORMAT_R8G8B8A8_UNORM = 37, VK_FORMAT_R8G8B8A8_SRGB = 43, VK_FORMAT_R16G16B16A16_SFLOAT = 97;

bool fail(std::string* error, const char* why) {
    if (error) *error = why;
    return false;
}
uint32_t bswap32(uint32_t v) { return (v >> 24) | ((v >> 8) & 0xFF00u) | ((v << 8) & 0xFF0000u) | (v << 24); }

struct Reader {
    const uint8_t* p; size_t n; bool swap = false;
    bool u32(size_t off, uint32_t& v) const {
        if (off + 4 > n) return false;
        std::memcpy(&v, p + off, 4);
        if (swap) v = bswap32(v);
        return true;
    }
    bool u64(size_t off, uint64_t& v) const {
        if (off + 8 > n) return false;
        std::memcpy(&v, p + off, 8);     // KTX 2 is little-endian by definition
        return true;
    }
};

size_t level_texels(uint32_t w, uint32_t h, uint32_t l) { return (size_t)std::max(1u, w >> l) * std::max(1u, h >> l); }

bool decode_ktx1(const uint8_t* bytes, size_t size, KtxCube& out, std::string* error) {
    Reader r{bytes, size};
    uint32_t endian = 0;
    if (!r.u32(12, endian)) return fail(error, "KTX1: truncated header");
    if (endian == 0x01020304u) r.swap = true;
    else if (endian != 0x04030201u) return fail(error, "KTX1: bad endianness field");
    uint32_t glType, glTypeSize, glFormat, glInternal, glBase, w, h, d, arrayN, faces, mips, kvBytes;
    if (!r.u32(16, glType) || !r.u32(20, glTypeSize) || !r.u32(24, glFormat) || !r.u32(28, glInternal) || !r.u32(32, glBase) ||
        !r.u32(36, w) || !r.u32(40, h) || !r.u32(44, d) || !r.u32(48, arrayN) || !r.u32(52, faces) || !r.u32(56, mips) || !r.u32(60, kvBytes))
        return fail(error, "KTX1: truncated header");
    if (faces != 6 || arrayN > 0 || d > 1) return fail(error, "KTX1: not a (non-array) cube map");
    if (w == 0 || h == 0 || w != h || w > 16384) return fail(error, "KTX1: bad cube extent");
    size_t bpt;
    if (glInternal == GL_RGBA8 && glType == GL_UNSIGNED_BYTE && glFormat == GL_RGBA) { out.format = VKR_FORMAT_R8G8B8A8_UNORM; bpt = 4; }
    else if (glInternal == GL_SRGB8_ALPHA8 && glType == GL_UNSIGNED_BYTE && glFormat == GL_RGBA) { out.format = VKR_FORMAT_R8G8B8A8_SRGB; bpt = 4; }
    else if (glInternal == GL_RGBA16F && glType == GL_HALF_FLOAT && glFormat == GL_RGBA) { out.format = VKR_FORMAT_R16G16B16A16_SFLOAT; bpt = 8; }
    else return fail(error, "KTX1: texel format is not RGBA8 / SRGB8_ALPHA8 / RGBA16F");
    out.generateMips = mips == 0;
    const uint32_t levels = mips == 0 ? 1 : mips;
    if (levels > 16 || (levels > 1 && (w >> (levels - 1)) == 0)) return fail(error, "KTX1: more mip levels than the extent allows");
    out.width = w; out.height = h; out.levels = levels;
    out.images.assign((size_t)levels * 6, {});
    size_t off = 64 + (size_t)kvBytes;
    for (uint32_t l = 0; l < levels; ++l) {
        uint32_t imageSize = 0;                      // non-array cube map: the size of ONE face
        if (!r.u32(off, imageSize)) return fail(error, "KTX1: truncated level");
        off += 4;
        const size_t want = level_texels(w, h, l) * bpt;
        if (imageSize != want) return fail(error, "KTX1: imageSize does not match the level's extent");
        for (uint32_t f = 0; f < 6; ++f) {
            if (off + want > size) return fail(error, "KTX1: truncated face");
            std::vector<uint8_t>& img = out.images[(size_t)l * 6 + f];
            img.assign(bytes + off, bytes + off + want);
            if (r.swap && glTypeSize == 2)           // a file written on a machine of the other endianness
                for (size_t i = 0; i + 1 < img.size(); i += 2) std::swap(img[i], img[i + 1]);
            off += want;
            off = (off + 3) & ~(size_t)3;             // cubePadding
        }
        off = (off + 3) & ~(size_t)3;                 // mipPadding
    }
    return true;
}

// Zstandard (KTX2 supercompressionScheme 2, what `toktx --zcmp` writes): the system's libzstd, looked up at the first such file —
// the engine library does not link it, so a machine without it still loads every other file and refuses these with a reason.
typedef size_t (*ZstdDecompressFn)(void* dst, size_t dstCapacity, const void* src, size_t compressedSize);
typedef unsigned (*ZstdIsErrorFn)(size_t code);
static bool zstd_functions(ZstdDecompressFn& decompress, ZstdIsErrorFn& isError) {
    static void* lib = dlopen("libzstd.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!lib) return false;
    decompress = (ZstdDecompressFn)dlsym(lib, "ZSTD_decompress");
    isError = (ZstdIsErrorFn)dlsym(lib, "ZSTD_isError");
    return decompress != nullptr && isError != nullptr;
}

bool decode_ktx2(const uint8_t* bytes, size_t size, KtxCube& out, std::string* error) {
    Reader r{bytes, size};
    uint32_t vkFormat, typeSize, w, h, d, layers, faces, levels, scheme;
    if (!r.u32(12, vkFormat) || !r.u32(16, typeSize) || !r.u32(20, w) || !r.u32(24, h) || !r.u32(28, d) || !r.u32(32, layers) ||
        !r.u32(36, faces) || !r.u32(40, levels) || !r.u32(44, scheme))
        return fail(error, "KTX2: truncated header");
    if (faces != 6 || layers > 0 || d > 0) return fail(error, "KTX2: not a (non-array) cube map");
    // supercompressionScheme: 0 none, 2 Zstandard, 3 ZLIB (every level compressed on its own); BasisLZ (1) is refused
    if (scheme != 0 && scheme != 2 && scheme != 3) return fail(error, "KTX2: BasisLZ supercompression is not supported (none, Zstandard and ZLIB are)");
    ZstdDecompressFn zstdDecompress = nullptr; ZstdIsErrorFn zstdIsError = nullptr;
    if (scheme == 2 && !zstd_functions(zstdDecompress, zstdIsError)) return fail(error, "KTX2: Zstandard supercompression needs libzstd.so.1, which this machine does not have");
    if (w == 0 || h == 0 || w != h || w > 16384) return fail(error, "KTX2: bad cube extent");
    size_t bpt;
    if (vkFormat == VK_FORMAT_R8G8B8A8_UNORM) { out.format = VKR_FORMAT_R8G8B8A8_UNORM; bpt = 4; }
    else if (vkFormat == VK_FORMAT_R8G8B8A8_SRGB) { out.format = VKR_FORMAT_R8G8B8A8_SRGB; bpt = 4; }
    else if (vkFormat == VK_FORMAT_R16G16B16A16_SFLOAT) { out.format = VKR_FORMAT_R16G16B16A16_SFLOAT; bpt = 8; }
    else return fail(error, "KTX2: vkFormat is not R8G8B8A8_UNORM / R8G8B8A8_SRGB / R16G16B16A16_SFLOAT");
    out.generateMips = levels == 0;
    if (levels == 0) levels = 1;
    if (levels > 16 || (levels > 1 && (w >> (levels - 1)) == 0)) return fail(error, "KTX2: more mip levels than the extent allows");
    out.width = w; out.height = h; out.levels = levels;
    out.images.assign((size_t)levels * 6, {});
    const size_t indexOff = 12 + 9 * 4 + 4 * 4 + 2 * 8;      // identifier, header, dfd/kvd offsets+lengths, sgd offset+length
    for (uint32_t l = 0; l < levels; ++l) {
        uint64_t byteOffset, byteLength, uncompressed;
        if (!r.u64(indexOff + (size_t)l * 24, byteOffset) || !r.u64(indexOff + (size_t)l * 24 + 8, byteLength) ||
            !r.u64(indexOff + (size_t)l * 24 + 16, uncompressed))
            return fail(error, "KTX2: truncated level index");
        const size_t want = level_texels(w, h, l) * bpt;
        if (byteOffset > size || byteLength > size - byteOffset) return fail(error, "KTX2: truncated level");
        const uint8_t* texels = bytes + byteOffset;
        std::vector<uint8_t> inflated;
        if (scheme == 3 || scheme == 2) {
            if (uncompressed != (uint64_t)want * 6) return fail(error, "KTX2: level length does not match its extent");
            inflated.resize(want * 6);
            if (scheme == 3) {
                uLongf got = (uLongf)inflated.size();
                if (uncompress(inflated.data(), &got, texels, (uLong)byteLength) != Z_OK || got != inflated.size())
                    return fail(error, "KTX2: a ZLIB-compressed level does not inflate to its extent");
            } else {
                const size_t got = zstdDecompress(inflated.data(), inflated.size(), texels, (size_t)byteLength);
                if (zstdIsError(got) || got != inflated.size()) return fail(error, "KTX2: a Zstandard-compressed level does not decompress to its extent");
            }
            texels = inflated.data();
        } else if (byteLength != (uint64_t)want * 6) return fail(error, "KTX2: level length does not match its extent");
        for (uint32_t f = 0; f < 6; ++f)
            out.images[(size_t)l * 6 + f].assign(texels + (size_t)f * want, texels + (size_t)(f + 1) * want);
    }
    return true;
}

}  // namespace

bool decode_ktx_cube(const uint8_t* bytes, size_t size, KtxCube& out, std::string* error) {
    out = KtxCube{};
    if (!bytes || size < 64) return fail(error, "not a KTX file (too short)");
    if (std::memcmp(bytes, kKtx1Id, 12) == 0) return decode_ktx1(bytes, size, out, error);
    if (std::memcmp(bytes, kKtx2Id, 12) == 0) return decode_ktx2(bytes, size, out, error);
    return fail(error, "not a KTX file (bad identifier)");
}

bool load_ktx_cube(const std::string& fileName, KtxCube& out, std::string* error) {
    std::FILE* f = std::fopen(fileName.c_str(), "rb");
    if (!f) return fail(error, "cannot open file");
    std::vector<uint8_t> bytes;
    uint8_t buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof(buf), f)) > 0) bytes.insert(bytes.end(), buf, buf + got);
    std::fclose(f);
    return decode_ktx_cube(bytes.data(), bytes.size(), out, error);
}
