// vk_ktx.h — KTX container decode for VulkanEngine::load_cubemap (src/vk_loader.cpp:521-558).
// The reference hands the file to libktx (ktxTexture_CreateFromNamedFile + ktxTexture_VkUploadEx: every mip level and
// all six faces of the file become one cube image).  libktx is not in this image, so the two container layouts are read
// here: KTX 1.1 (the engine's asset, assets/pisa_cube.ktx) and KTX 2.0 without supercompression, for the three texel
// formats the device path samples (RGBA8 UNORM / sRGB, RGBA16F).  Anything else is refused with a reason, where the
// reference prints "failed to load ktx texture image!" and goes on without a skybox.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../../include/vkr.h"

struct KtxCube {
    uint32_t width = 0, height = 0, levels = 0;
    VkrFormat format = VKR_FORMAT_R8G8B8A8_UNORM;
    bool generateMips = false;                       // KTX1 numberOfMipmapLevels == 0: the loader is asked to build the chain
    std::vector<std::vector<uint8_t>> images;        // [level * 6 + face], face order +X,-X,+Y,-Y,+Z,-Z, rows top to bottom
};

// Decodes a KTX 1.1 / 2.0 cube map held in memory.  False (with *error set) when the bytes are not a cube map this reader handles.
bool decode_ktx_cube(const uint8_t* bytes, size_t size, KtxCube& out, std::string* error = nullptr);
bool load_ktx_cube(const std::string& fileName, KtxCube& out, std::string* error = nullptr);
