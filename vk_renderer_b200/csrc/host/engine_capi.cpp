// engine_capi.cpp — extern "C" surface of the C++ host engine (libvkr_engine.so), the call a user of
// this repo makes: create an engine, build a scene, then update_scene() + draw() per frame.
// Python (tests, bench.py) drives it through ctypes; see vk_renderer_b200/engine.py.
#include <cstring>

#include "vk_engine.h"
#include "vk_gltf.h"
#include "vk_ktx.h"
#include "vk_loader.h"

const VkrApi& libvkr_api() {
    static const VkrApi api = {vkr_create, vkr_destroy, vkr_upload_mesh, vkr_create_image, vkr_update_bindless_texture,
                               vkr_set_cubemap, vkr_build_material, vkr_draw, vkr_readback, vkr_readback_async, vkr_sync,
                               vkr_set_cubemap_levels, vkr_scene_create, vkr_scene_set_transforms, vkr_draw_scene, vkr_scene_draw_list,
                               vkr_stats, vkr_scene_destroy};
    return api;
}

extern "C" {

// api == nullptr -> libvkr (the product).  A non-null table is how tests put the oracle behind the engine.
VulkanEngine* vke_create(uint32_t width, uint32_t height, int device, void* stream, uint32_t flags, const VkrApi* api,
                         uint32_t shadowExtent) {
    VulkanEngine* e = new VulkanEngine();
    if (api) e->_api = *api;
    if (shadowExtent) e->_shadowExtent = {shadowExtent, shadowExtent};   // default 2048 (src/vk_engine.h:107)
    if (e->init(width, height, device, stream, flags) != VKR_SUCCESS) { delete e; return nullptr; }
    return e;
}
void vke_destroy(VulkanEngine* e) { delete e; }
VkrDevice* vke_device(VulkanEngine* e) { return e->_vkr; }

uint64_t vke_load_procedural(VulkanEngine* e, int config, uint32_t seed) { e->scene_changed(); return load_procedural(e, config, seed); }
void vke_orbit_camera(VulkanEngine* e, uint32_t seed, int view) { orbit_camera(e, seed, view); }

// init_scene's `_sceneRoot->addChild(load_gltf(name, file))` (src/vk_engine_init.cpp:646-660).  Returns the number of triangles
// added, or -1 with the reason copied to err (may be null).
int64_t vke_load_gltf(VulkanEngine* e, const char* name, const char* fileName, char* err, size_t errBytes) {
    std::string why;
    std::shared_ptr<Node> top = load_gltf(e, name ? name : "gltf", fileName ? fileName : "", &why);
    if (!top) {
        if (err && errBytes) { std::strncpy(err, why.c_str(), errBytes - 1); err[errBytes - 1] = 0; }
        return -1;
    }
    e->_sceneRoot->addChild(top);
    e->scene_changed();
    std::vector<RenderObject> all;
    top->refreshTransform(vkm::mat4(1.0f), all);
    int64_t tris = 0;
    for (const RenderObject& o : all) tris += o.indexCount / 3;
    return tris;
}
// load_cubemap("../../assets/pisa_cube.ktx", _skyboxImage) of init_scene (src/vk_engine_init.cpp:677).  0 on success, -1 with the reason.
int vke_load_cubemap(VulkanEngine* e, const char* fileName, char* err, size_t errBytes) {
    std::string why;
    if (e->load_cubemap(fileName, &why)) return 0;
    if (err && errBytes) { std::strncpy(err, why.c_str(), errBytes - 1); err[errBytes - 1] = 0; }
    return -1;
}
void vke_gltf_convert(int componentType, int normalized, const unsigned char* data, size_t count, float* out) {
    gltf_convert_components(componentType, normalized != 0, data, count, out);
}
// Test hook: decode a KTX cube held in memory.  info = {width, height, levels, format, generateMips}; when `out` is large enough
// (all levels, level-major then face) the texels are copied there.  0 on success, -1 with the reason.
int vke_decode_ktx(const unsigned char* data, size_t size, uint32_t info[5], unsigned char* out, size_t outBytes, char* err, size_t errBytes) {
    KtxCube cube;
    std::string why;
    if (!decode_ktx_cube(data, size, cube, &why)) {
        if (err && errBytes) { std::strncpy(err, why.c_str(), errBytes - 1); err[errBytes - 1] = 0; }
        return -1;
    }
    info[0] = cube.width; info[1] = cube.height; info[2] = cube.levels; info[3] = (uint32_t)cube.format; info[4] = cube.generateMips ? 1u : 0u;
    size_t total = 0;
    for (const auto& img : cube.images) total += img.size();
    if (out && outBytes >= total) {
        size_t off = 0;
        for (const auto& img : cube.images) { std::memcpy(out + off, img.data(), img.size()); off += img.size(); }
    }
    return 0;
}
// The flattened scene after update_scene: object i's mesh/material handles, world matrix, index range and its material constants.
int vke_render_object(VulkanEngine* e, size_t i, VkrRenderObject* out, MaterialConstants* params, int* passType) {
    if (i >= e->_renderObjects.size()) return -1;
    const RenderObject& o = e->_renderObjects[i];
    out->mesh = o.mesh->meshBuffer; out->material = o.material ? o.material->materialSet : 0;
    std::memcpy(out->transform, o.transform.data(), 64);
    out->indexCount = o.indexCount; out->firstIndex = o.firstIndex;
    if (params && o.material) *params = o.material->params;
    if (passType && o.material) *passType = (int)o.material->passType;
    return 0;
}
int vke_decode_png(const unsigned char* data, size_t size, unsigned char* rgbaOut, size_t outBytes, uint32_t* w, uint32_t* h) {
    std::vector<unsigned char> px;
    if (!decode_png(data, size, px, *w, *h)) return -1;
    if (rgbaOut) { if (outBytes < px.size()) return -2; std::memcpy(rgbaOut, px.data(), px.size()); }
    return 0;
}

int vke_decode_jpeg(const unsigned char* data, size_t size, unsigned char* rgbaOut, size_t outBytes, uint32_t* w, uint32_t* h) {
    std::vector<unsigned char> px;
    if (!decode_jpeg(data, size, px, *w, *h)) return -1;
    if (rgbaOut) { if (outBytes < px.size()) return -2; std::memcpy(rgbaOut, px.data(), px.size()); }
    return 0;
}

void vke_set_camera(VulkanEngine* e, float x, float y, float z, float pitch, float yaw) {
    e->_mainCamera.position = vkm::vec3{x, y, z};
    e->_mainCamera.pitch = pitch; e->_mainCamera.yaw = yaw;
}
void vke_set_light(VulkanEngine* e, const float dir[3], const float color[3], const float ambient[3]) {
    if (dir) e->_sunlightDirection = vkm::vec4{dir[0], dir[1], dir[2], 0.0f};
    if (color) e->_sunlightColor = vkm::vec4{color[0], color[1], color[2], 1.0f};
    if (ambient) e->_ambientColor = vkm::vec4{ambient[0], ambient[1], ambient[2], 1.0f};
}
// the ImGui "Settings" toggles (src/vk_engine_run.cpp:200-232); a negative value leaves a field unchanged
void vke_set_toggles(VulkanEngine* e, int enableShadows, int shadowMode, int enableBackground, int enablePostprocess, int fragShader) {
    if (enableShadows >= 0) e->_enableShadows = enableShadows != 0;
    if (shadowMode >= 0) e->_shadowMode = shadowMode;
    if (enableBackground >= 0) e->_enableBackground = enableBackground != 0;
    if (enablePostprocess >= 0) e->_enablePostprocess = enablePostprocess != 0;
    if (fragShader >= 0) e->_fragShader = fragShader;
}
void vke_set_scissor(VulkanEngine* e, int x, int y, int w, int h, int cullToScissor) {
    e->_scissor[0] = x; e->_scissor[1] = y; e->_scissor[2] = w; e->_scissor[3] = h;
    e->_cullToScissor = cullToScissor != 0;
}

void vke_update_scene(VulkanEngine* e, float dt) { e->update_scene(dt); }
int vke_draw(VulkanEngine* e) { return e->draw(); }
int vke_present(VulkanEngine* e, void* pinnedDst, size_t bytes) { return e->present(pinnedDst, bytes); }
int vke_sync(VulkanEngine* e) { return e->_api.sync ? e->_api.sync(e->_vkr) : VKR_SUCCESS; }
int vke_readback(VulkanEngine* e, int what, void* dst, size_t bytes) { return e->_api.readback(e->_vkr, (VkrTarget)what, dst, bytes); }

void vke_stats(VulkanEngine* e, EngineStats* out) {
    if (e->_gpuSceneUpdate && e->_api.stats) {        // the device counted what its is_visible kept (blocks until it has)
        VkrEngineStats st{};
        if (e->_api.stats(e->_vkr, &st) == VKR_SUCCESS) { e->stats.drawcall_count = st.drawcall_count; e->stats.triangle_count = st.triangle_count; }
    }
    *out = e->stats;
}
// update_scene's per-object work on the device (vkr_scene_create / vkr_draw_scene) from the next frame on
void vke_set_gpu_scene_update(VulkanEngine* e, int on) { e->_gpuSceneUpdate = on != 0; }
void vke_scene_changed(VulkanEngine* e) { e->scene_changed(); }
// Node::localTransform of the idx-th node in traversal order (refreshTransform's) := m (column-major); -1 if there is none
int vke_set_node_transform(VulkanEngine* e, size_t idx, const float m[16]) {
    std::vector<Node*> order;
    std::vector<Node*> stack;
    if (e->_sceneRoot) stack.push_back(e->_sceneRoot.get());
    while (!stack.empty()) {
        Node* n = stack.back(); stack.pop_back();
        order.push_back(n);
        for (auto c = n->children.rbegin(); c != n->children.rend(); ++c) stack.push_back(c->get());
    }
    if (idx >= order.size()) return -1;
    std::memcpy(order[idx]->localTransform.data(), m, 64);
    e->transforms_changed();
    return 0;
}
size_t vke_node_count(VulkanEngine* e) {
    size_t n = 0;
    std::vector<Node*> stack;
    if (e->_sceneRoot) stack.push_back(e->_sceneRoot.get());
    while (!stack.empty()) { Node* nd = stack.back(); stack.pop_back(); ++n; for (auto& c : nd->children) stack.push_back(c.get()); }
    return n;
}
// the draw list of the last frame as the device's update_scene left it (vkr_scene_draw_list); returns the count, -1 on failure
int64_t vke_gpu_draw_list(VulkanEngine* e, VkrRenderObject* out, size_t capacity) {
    if (!e->_api.scene_draw_list || !e->_haveScene) return -1;
    size_t count = 0;
    if (e->_api.scene_draw_list(e->_vkr, e->_sceneHandle, out, capacity, &count) != VKR_SUCCESS) return -1;
    return (int64_t)count;
}
size_t vke_visible_objects(VulkanEngine* e) { return e->_renderObjects.size(); }
void vke_scene_data(VulkanEngine* e, GPUSceneData* out) { *out = e->_sceneData; }

// host-math probes (checked against glm golden vectors and the oracle's restatement)
void vke_camera_matrices(VulkanEngine* e, float view[16], float proj[16]) {
    vkm::mat4 v = e->_mainCamera.getViewMatrix();
    vkm::mat4 p = e->_mainCamera.getProjectionMatrix((float)e->_windowExtent.width / (float)e->_windowExtent.height);
    std::memcpy(view, v.data(), 64); std::memcpy(proj, p.data(), 64);
}
void vke_light_matrix(VulkanEngine* e, float out[16]) { vkm::mat4 m = e->compute_light_matrix(); std::memcpy(out, m.data(), 64); }
void vke_csm(VulkanEngine* e, float matrices[64], float distances[4]) {
    CSMData c = e->compute_csmdata();
    for (int i = 0; i < NUM_CASCADES; ++i) { std::memcpy(matrices + 16 * i, c.lightMatrices[i].data(), 64); distances[i] = c.planeDistances[i]; }
}
float vke_compute_split(float n, float f, int i) { return compute_split(n, f, i); }
// extract_frustum_planes(proj * view) of the current camera, 6 planes x 4 floats
void vke_frustum_planes(VulkanEngine* e, float out[24]) {
    vkm::mat4 v = e->_mainCamera.getViewMatrix();
    vkm::mat4 p = e->_mainCamera.getProjectionMatrix((float)e->_windowExtent.width / (float)e->_windowExtent.height);
    Frustum f;
    extract_frustum_planes(p * v, f);
    std::memcpy(out, f.planes.data(), 96);
}
// is_visible for an object with a sphere bound (origin, radius) under `transform` (column-major), current camera
int vke_is_visible(VulkanEngine* e, const float transform[16], const float origin[3], float sphereRadius) {
    vkm::mat4 v = e->_mainCamera.getViewMatrix();
    vkm::mat4 p = e->_mainCamera.getProjectionMatrix((float)e->_windowExtent.width / (float)e->_windowExtent.height);
    Frustum f;
    extract_frustum_planes(p * v, f);
    MeshAsset mesh;
    mesh.bounds.origin = {origin[0], origin[1], origin[2]};
    mesh.bounds.sphereRadius = sphereRadius;
    RenderObject ro{};
    ro.mesh = &mesh;
    std::memcpy(ro.transform.data(), transform, 64);
    return e->is_visible(ro, f) ? 1 : 0;
}

}  // extern "C"
