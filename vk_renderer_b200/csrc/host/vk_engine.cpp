// vk_engine.cpp — per-frame host path of the engine (scene update, light matrices, draw) on libvkr.
// Follows src/vk_engine_run.cpp; citations on each function.
#include "vk_engine.h"
#include "vk_ktx.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>

using namespace vkm;

// ---------------------------------------------------------------- MaterialSystem (src/vk_materials.cpp)
void MaterialSystem::init(VulkanEngine* engine) {
    _engine = engine;
    // init_mesh_pipelines registers exactly these two templates (src/vk_engine_init.cpp:543-547)
    _templateCache["Opaque"] = MaterialTemplate{"Opaque", MaterialPass::MainColor};
    _templateCache["Transparent"] = MaterialTemplate{"Transparent", MaterialPass::Transparent};
}
MaterialTemplate* MaterialSystem::get_template(const std::string& name) {
    auto it = _templateCache.find(name);
    return it == _templateCache.end() ? nullptr : &it->second;
}
MaterialInstance* MaterialSystem::build_instance(MaterialTemplate* materialTemplate, MaterialConstants params) {
    MaterialInstance* mat = new MaterialInstance();
    mat->pipeline = materialTemplate;
    mat->params = params;
    mat->passType = materialTemplate->passType;
    VkrMaterialConstants c;
    std::memcpy(&c, &params, sizeof(c));
    if (_engine->_api.build_material(_engine->_vkr, &c, mat->passType == MaterialPass::Transparent ? VKR_PASS_TRANSPARENT : VKR_PASS_MAIN_COLOR,
                           &mat->materialSet) != VKR_SUCCESS) {
        delete mat;
        return nullptr;
    }
    _instances.push_back(mat);
    return mat;
}
MaterialSystem::~MaterialSystem() { for (auto* m : _instances) delete m; }

// ---------------------------------------------------------------- init / resources
int VulkanEngine::init(uint32_t width, uint32_t height, int device, void* stream, uint32_t flags) {
    _windowExtent = {width, height};
    _drawExtent = {width, height};
    _cudaDevice = device; _cudaStream = stream; _vkrFlags = flags;
    VkrConfig cfg{device, width, height, _shadowExtent.width, stream, flags};
    int rc = _api.create(&cfg, &_vkr);
    if (rc != VKR_SUCCESS) return rc;
    _globalTextureIndex = 2;                       // init_default_data (src/vk_engine_init.cpp:351-355)
    _materialSystem.init(this);
    _sceneRoot = std::make_shared<Node>();         // init_scene (:646-647)
    _mainCamera.velocity = vec3{0, 0, 0};          // init_camera (:735-742)
    _mainCamera.position = vec3{0, 0, 0};
    _mainCamera.pitch = 0; _mainCamera.yaw = 0;
    return VKR_SUCCESS;
}
void VulkanEngine::cleanup() {
    if (_vkr) { _api.destroy(_vkr); _vkr = nullptr; }
}
VkrMesh VulkanEngine::upload_mesh(std::span<uint32_t> indices, std::span<Vertex> vertices) {
    VkrMesh m = 0xFFFFFFFFu;
    _api.upload_mesh(_vkr, indices.data(), indices.size(), reinterpret_cast<const VkrVertex*>(vertices.data()), vertices.size(), &m);
    return m;
}
VkrImage VulkanEngine::create_image(const void* data, uint32_t w, uint32_t h, VkrFormat format, bool mipmapped) {
    VkrImage img = 0xFFFFFFFFu;
    _api.create_image(_vkr, data, w, h, format, mipmapped ? 1 : 0, &img);
    return img;
}
void VulkanEngine::update_bindless_texture(int index, VkrImage image, VkrSampler sampler) {
    _api.update_bindless_texture(_vkr, index, image, sampler);
}
int VulkanEngine::add_texture(const void* rgba8, uint32_t w, uint32_t h) {
    // every loaded image is R8G8B8A8_SRGB, mipmapped, bound with the LINEAR sampler (src/vk_loader.cpp:283-320)
    VkrImage img = create_image(rgba8, w, h, VKR_FORMAT_R8G8B8A8_SRGB, true);
    int id = _globalTextureIndex++;
    update_bindless_texture(id, img, VKR_SAMPLER_LINEAR);
    return id;
}
void VulkanEngine::set_cubemap(const void* const faces[6], uint32_t w, uint32_t h) {
    if (_api.set_cubemap(_vkr, faces, w, h, VKR_FORMAT_R8G8B8A8_SRGB, 1) == VKR_SUCCESS) _hasSkybox = true;
}
// load_cubemap (src/vk_loader.cpp:521-558): the KTX file's faces and mip levels become the skybox image as they are.  On
// failure the reference prints a message and carries on without a skybox; so does this (the reason goes to *error).
bool VulkanEngine::load_cubemap(const char* filename, std::string* error) {
    KtxCube cube;
    std::string why;
    if (!load_ktx_cube(filename ? filename : "", cube, &why)) {
        std::fprintf(stderr, "failed to load ktx texture image! (%s)\n", why.c_str());
        if (error) *error = why;
        return false;
    }
    int rc;
    if (cube.generateMips && cube.format != VKR_FORMAT_R16G16B16A16_SFLOAT) {
        const void* faces[6];
        for (int f = 0; f < 6; ++f) faces[f] = cube.images[f].data();
        rc = _api.set_cubemap(_vkr, faces, cube.width, cube.height, cube.format, 1);
    } else {
        if (!_api.set_cubemap_levels) { if (error) *error = "backend has no set_cubemap_levels"; return false; }
        std::vector<const void*> ptrs;
        for (const auto& img : cube.images) ptrs.push_back(img.data());
        rc = _api.set_cubemap_levels(_vkr, ptrs.data(), cube.width, cube.height, cube.levels, cube.format);
    }
    if (rc != VKR_SUCCESS) {
        std::fprintf(stderr, "failed to upload ktx texture!\n");
        if (error) *error = "upload failed";
        return false;
    }
    _hasSkybox = true;
    return true;
}
std::shared_ptr<MeshAsset> VulkanEngine::add_mesh(const std::string& name, std::span<uint32_t> indices, std::span<Vertex> vertices,
                                                   const std::vector<GeoSurface>& surfaces) {
    auto newMesh = std::make_shared<MeshAsset>();
    newMesh->name = name;
    newMesh->surfaces = surfaces;
    vec3 minPos{100000.f, 100000.f, 100000.f}, maxPos{-100000.f, -100000.f, -100000.f};   // src/vk_loader.cpp:391-392
    for (const Vertex& v : vertices) { minPos = min3(minPos, v.position); maxPos = max3(maxPos, v.position); }
    newMesh->bounds.origin = (minPos + maxPos) * 0.5f;                                      // :458-460
    newMesh->bounds.extents = (maxPos - minPos) * 0.5f;
    newMesh->bounds.sphereRadius = length(newMesh->bounds.extents);
    newMesh->meshBuffer = upload_mesh(indices, vertices);
    newMesh->indexOffset = vertices.size() * sizeof(Vertex);
    _meshAssets[name] = newMesh;
    return newMesh;
}
