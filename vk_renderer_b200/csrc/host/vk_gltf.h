// vk_gltf.h — glTF 2.0 / GLB ingestion for the headless engine: VulkanEngine::load_gltf of the reference
// (src/vk_loader.cpp:227-518) without fastgltf / simdjson / stb_image (none of them is available here).
//
// Same observable behaviour as the reference's loader, quirks included:
//   * every image becomes an R8G8B8A8_SRGB mip-mapped texture in the next bindless slot, LINEAR sampler; an image that
//     cannot be decoded takes the 16x16 magenta/black error checkerboard instead (src/vk_loader.cpp:275-329);
//   * material ids: baseColorTexture -> albedoID, metallicRoughnessTexture -> normalID, normalTexture -> metalRoughID
//     (the swap of src/vk_loader.cpp:347-363), defaults 0 / 1 / 0; alphaMode BLEND -> "Transparent" template;
//   * one MeshAsset per glTF mesh, one GeoSurface per primitive with an index accessor (primitives without one are
//     skipped, :399-403), vertices default normal (1,0,0), colour 1, uv 0 (:417-421), bounds from all positions (:458-460);
//   * accessor conversion follows fastgltf (external/fastgltf/include/fastgltf/tools.hpp:126-149): normalised integers
//     are max(c / MAX, -1), everything else a static_cast;
//   * nodes: `matrix` or T * R * S (:474-490); nodes without a parent hang under one new top node (:509-515).
// Supported containers: .gltf with external or base64 buffers and images, .glb.  Images: PNG (8/16-bit, all colour types,
// non-interlaced), decoded with zlib, and JPEG (baseline and progressive Huffman, grey / YCbCr / RGB; vk_jpeg.cpp) — the two
// formats glTF allows and the reference's own scene (Khronos Sponza) uses.  An image neither decoder takes gets the error
// checkerboard, like a failed stbi_load.  Sparse accessors are not decoded (the load fails).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "vk_engine.h"

// Returns the new top node (not yet attached to the scene) or nullptr, leaving the reason in *error when given.
std::shared_ptr<Node> load_gltf(VulkanEngine* engine, const std::string& name, const std::string& fileName, std::string* error = nullptr);

// PNG -> RGBA8 (test hook and used by load_gltf).  Returns false when the stream is not a PNG this decoder handles.
bool decode_png(const unsigned char* data, size_t size, std::vector<unsigned char>& rgba, uint32_t& width, uint32_t& height);

// JPEG -> RGBA8, alpha 255 (vk_jpeg.cpp).  Returns false when the stream is not a JPEG this decoder handles.
bool decode_jpeg(const unsigned char* data, size_t size, std::vector<unsigned char>& rgba, uint32_t& width, uint32_t& height);

// Test hook (see vk_gltf.cpp): the accessor component conversion on a packed array of scalars.
void gltf_convert_components(int componentType, bool normalized, const unsigned char* data, size_t count, float* out);
