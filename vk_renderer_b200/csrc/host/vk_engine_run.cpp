// vk_engine_run.cpp — update_scene / light matrices / draw, following src/vk_engine_run.cpp.
#include "vk_engine.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

using namespace vkm;

// Gribb-Hartmann planes from transpose(VP), normalised by |n| (src/vk_engine_run.cpp:420-433)
void extract_frustum_planes(const mat4& VP, Frustum& frustum) {
    mat4 M = transpose(VP);
    frustum.planes[0] = column(M, 3) + column(M, 0);
    frustum.planes[1] = column(M, 3) - column(M, 0);
    frustum.planes[2] = column(M, 3) + column(M, 1);
    frustum.planes[3] = column(M, 3) - column(M, 1);
    frustum.planes[4] = column(M, 2);
    frustum.planes[5] = column(M, 3) - column(M, 2);
    for (auto& plane : frustum.planes) {
        float len = length(xyz(plane));
        plane = plane / len;
    }
}

// src/vk_engine_run.cpp:435-459
void VulkanEngine::update_scene(float dt) {
    _mainCamera.update(dt);
    _renderObjects.clear();
    if (_gpuSceneUpdate) return;        // the traversal, the cull and the sort below run in kernels (draw() -> vkr_draw_scene)

    std::vector<RenderObject> allObjects;
    if (_sceneRoot) _sceneRoot->refreshTransform(mat4(1.0f), allObjects);

    float aspect = (float)_windowExtent.width / (float)_windowExtent.height;
    mat4 viewProj = _mainCamera.getProjectionMatrix(aspect) * _mainCamera.getViewMatrix();
    Frustum camFrustum;
    extract_frustum_planes(viewProj, camFrustum);
    if (_cullToScissor && _scissor[2] > 0 && _scissor[3] > 0) {
        // sort-first sharding: replace the y planes by the row band's (y_ndc in [y0, y1]); the shadow pass
        // still wants every camera-visible caster, so this is only used when shadows are off
        mat4 M = transpose(viewProj);
        const float y0 = 2.0f * (float)_scissor[1] / (float)_drawExtent.height - 1.0f;
        const float y1 = 2.0f * (float)(_scissor[1] + _scissor[3]) / (float)_drawExtent.height - 1.0f;
        vec4 r1 = column(M, 1), r3 = column(M, 3);
        vec4 lo{r1.x - y0 * r3.x, r1.y - y0 * r3.y, r1.z - y0 * r3.z, r1.w - y0 * r3.w};
        vec4 hi{y1 * r3.x - r1.x, y1 * r3.y - r1.y, y1 * r3.z - r1.z, y1 * r3.w - r1.w};
        camFrustum.planes[2] = lo / length(xyz(lo));
        camFrustum.planes[3] = hi / length(xyz(hi));
    }
    for (const auto& obj : allObjects)
        if (is_visible(obj, camFrustum)) _renderObjects.push_back(obj);

    // The reference sorts with std::sort on (MainColor first, then MaterialTemplate address), which leaves
    // the relative order of equal keys unspecified.  There is one template per pass
    // type, so the key reduces to the pass type; we define the order as the STABLE sort of traversal order.
    std::stable_sort(_renderObjects.begin(), _renderObjects.end(), [](const RenderObject& a, const RenderObject& b) {
        const bool am = a.material->passType == MaterialPass::MainColor, bm = b.material->passType == MaterialPass::MainColor;
        return am && !bm;
    });
}

// src/vk_engine_run.cpp:461-480
bool VulkanEngine::is_visible(const RenderObject& obj, const Frustum& frustum) {
    const vec3 o = obj.mesh->bounds.origin;
    vec3 globalCenter = xyz(obj.transform * vec4{o.x, o.y, o.z, 1.f});
    float scaleX = length(xyz(column(obj.transform, 0)));
    float scaleY = length(xyz(column(obj.transform, 1)));
    float scaleZ = length(xyz(column(obj.transform, 2)));
    float maxScale = std::max(std::max(scaleX, scaleY), scaleZ);
    float globalRadius = obj.mesh->bounds.sphereRadius * maxScale;
    for (int i = 0; i < 6; i++) {
        float dist = dot(xyz(frustum.planes[i]), globalCenter) + frustum.planes[i].w;
        if (dist < -globalRadius) return false;
    }
    return true;
}

// src/vk_engine_run.cpp:482-491 — fixed ortho +-100, light at -dir*60, Y flipped
mat4 VulkanEngine::compute_light_matrix() {
    vec3 lightPos = -xyz(_sunlightDirection) * 60.f;
    mat4 lightView = lookAtRH(lightPos, vec3{0.f, 0.f, 0.f}, vec3{0, 1, 0});
    mat4 lightProj = orthoRH_ZO(-100.0f, 100.0f, -100.0f, 100.0f, _mainCamera.zNear, _mainCamera.zFar);
    lightProj.c[1][1] *= -1;
    return lightProj * lightView;
}

// src/vk_engine_run.cpp:493-504 — NDC z in {-1,+1} although depth is 0..1 (Appendix A-6), kept verbatim
static std::vector<vec4> getFrustumCornerWorld(const mat4& proj, const mat4& view) {
    const mat4 inv = inverse(proj * view);
    std::vector<vec4> frustumCorner;
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2; ++j)
            for (int k = 0; k < 2; ++k) {
                const vec4 pt = inv * vec4{2.f * i - 1.f, 2.f * j - 1.f, 2.f * k - 1.f, 1.f};
                frustumCorner.push_back(pt / pt.w);
            }
    return frustumCorner;
}

// src/vk_engine_run.cpp:506-543 — fit an ortho box to the sub-frustum; z range x10; NO Y flip
mat4 VulkanEngine::getLightMatrix(const float zNear, const float zFar) {
    float aspect = (float)_windowExtent.width / (float)_windowExtent.height;
    mat4 proj = perspectiveRH_ZO(radians(_mainCamera.fov), aspect, zNear, zFar);
    const auto corners = getFrustumCornerWorld(proj, _mainCamera.getViewMatrix());
    vec3 center{0.f, 0.f, 0.f};
    for (const auto& v : corners) center = center + xyz(v);
    center = center / (float)corners.size();
    vec3 lightDir = normalize(-xyz(_sunlightDirection));
    const mat4 lightView = lookAtRH(center + lightDir, center, vec3{0, 1, 0});
    float minX = std::numeric_limits<float>::max(), maxX = std::numeric_limits<float>::lowest();
    float minY = minX, maxY = maxX, minZ = minX, maxZ = maxX;
    for (const auto& v : corners) {
        const vec4 trf = lightView * v;
        minX = std::min(minX, trf.x); maxX = std::max(maxX, trf.x);
        minY = std::min(minY, trf.y); maxY = std::max(maxY, trf.y);
        minZ = std::min(minZ, trf.z); maxZ = std::max(maxZ, trf.z);
    }
    float zMult = 10.f;
    if (minZ < 0) minZ *= zMult; else minZ /= zMult;
    if (maxZ < 0) maxZ /= zMult; else maxZ *= zMult;
    const mat4 lightProj = orthoRH_ZO(minX, maxX, minY, maxY, minZ, maxZ);
    return lightProj * lightView;
}

// Practical split scheme, lambda = 0.5 (src/vk_engine_run.cpp:546-552); pow evaluates in double like the reference
float compute_split(float n, float f, int i) {
    float p = (float)i / (float)NUM_CASCADES;
    float C_log = n * pow(f / n, p);
    float C_uni = n + (f - n) * p;
    float lambda = 0.5f;
    return lambda * C_log + (1.f - lambda) * C_uni;
}

// src/vk_engine_run.cpp:554-566
CSMData VulkanEngine::compute_csmdata() {
    CSMData csm;
    float shadowCascadeLevels[NUM_CASCADES];
    for (int i = 0; i < NUM_CASCADES; ++i) {
        shadowCascadeLevels[i] = compute_split(_mainCamera.zNear, _mainCamera.zFar, i + 1);
        csm.planeDistances[i] = shadowCascadeLevels[i];
        float curNear = i == 0 ? _mainCamera.zNear : shadowCascadeLevels[i - 1];
        csm.lightMatrices[i] = getLightMatrix(curNear, shadowCascadeLevels[i]);
    }
    return csm;
}

// The scene graph as arrays, in the order Node::refreshTransform (src/vk_types.h:148-163) visits the nodes and emits their
// surfaces; objects draw() would skip (no material / no pipeline, src/vk_engine_run.cpp:285) are left out.
int VulkanEngine::flatten_scene() {
    if (!_api.scene_create) return VKR_ERROR_INVALID_ARGUMENT;
    std::vector<VkrSceneNode> nodes;
    std::vector<VkrSceneSurface> surfaces;
    _flatNodes.clear();
    struct Item { Node* node; int32_t parent; };
    std::vector<Item> stack;
    if (_sceneRoot) stack.push_back({_sceneRoot.get(), -1});
    while (!stack.empty()) {
        const Item it = stack.back();
        stack.pop_back();
        const int32_t self = (int32_t)nodes.size();
        VkrSceneNode n;
        std::memcpy(n.localTransform, it.node->localTransform.data(), 64);
        n.parent = it.parent;
        nodes.push_back(n);
        _flatNodes.push_back(it.node);
        if (it.node->mesh)
            for (const GeoSurface& surface : it.node->mesh->surfaces) {
                if (!surface.material || !surface.material->pipeline) continue;
                VkrSceneSurface sf;
                sf.node = (uint32_t)self; sf.mesh = it.node->mesh->meshBuffer; sf.material = surface.material->materialSet;
                sf.indexCount = surface.count; sf.firstIndex = surface.startIndex;
                const Bounds& b = it.node->mesh->bounds;
                sf.boundsOrigin[0] = b.origin.x; sf.boundsOrigin[1] = b.origin.y; sf.boundsOrigin[2] = b.origin.z;
                sf.boundsSphereRadius = b.sphereRadius;
                surfaces.push_back(sf);
            }
        for (auto c = it.node->children.rbegin(); c != it.node->children.rend(); ++c) stack.push_back({c->get(), self});   // pre-order, children left to right
    }
    if (nodes.empty()) return VKR_ERROR_INVALID_ARGUMENT;
    if (_haveScene && _api.scene_destroy) { _api.scene_destroy(_vkr, _sceneHandle); _haveScene = false; }     // the graph it was made from is gone
    const int rc = _api.scene_create(_vkr, nodes.data(), nodes.size(), surfaces.data(), surfaces.size(), &_sceneHandle);
    if (rc != VKR_SUCCESS) return rc;
    _haveScene = true; _sceneGraphDirty = false; _sceneTransformsDirty = false;
    return VKR_SUCCESS;
}

// src/vk_engine_run.cpp:68-193 — fill GPUSceneData (:107-125), then one vkr_draw replaces the five passes
int VulkanEngine::draw() {
    stats.triangle_count = 0;
    stats.drawcall_count = 0;
    GPUSceneData& sd = _sceneData;
    sd.view = _mainCamera.getViewMatrix();
    float aspect = (float)_windowExtent.width / (float)_windowExtent.height;
    sd.proj = _mainCamera.getProjectionMatrix(aspect);
    sd.viewproj = sd.proj * sd.view;
    sd.ambientColor = _ambientColor;
    sd.sunlightColor = _sunlightColor;
    sd.sunlightDirection = _sunlightDirection;
    CSMData csm = compute_csmdata();
    for (int i = 0; i < NUM_CASCADES; ++i) {
        sd.lightViewproj[i] = csm.lightMatrices[i];
        sd.cascadeDistances[i] = csm.planeDistances[i];
    }
    if (_shadowMode < 3) sd.lightViewproj[0] = compute_light_matrix();
    sd.sunlightColor.w = _enableShadows ? 1.f : 0.f;
    sd.sunlightDirection.w = (float)_shadowMode;

    _drawList.clear();
    _drawList.reserve(_renderObjects.size());
    for (const auto& object : _renderObjects) {
        if (!object.material || !object.material->pipeline) continue;   // :285
        VkrRenderObject r;
        r.mesh = object.mesh->meshBuffer;
        r.material = object.material->materialSet;
        std::memcpy(r.transform, object.transform.data(), 64);
        r.indexCount = object.indexCount;
        r.firstIndex = object.firstIndex;
        _drawList.push_back(r);
        stats.drawcall_count++;
        stats.triangle_count += object.indexCount / 3;
    }
    VkrFrameParams fp{};
    fp.enableShadows = _enableShadows; fp.shadowMode = _shadowMode;
    fp.enableBackground = _enableBackground; fp.enablePostprocess = _enablePostprocess;
    fp.enableSkybox = _hasSkybox; fp.fragShader = _fragShader;
    std::memcpy(fp.scissor, _scissor, sizeof(fp.scissor));
    std::memcpy(&fp.background, &_backgroundData, sizeof(fp.background));
    if (_gpuSceneUpdate) {
        // update_scene's camera frustum (:443-445): O(1) on the host; everything per object happens in vkr_draw_scene
        if (!_api.draw_scene) return VKR_ERROR_INVALID_ARGUMENT;
        if (_sceneGraphDirty || !_haveScene) { const int rc = flatten_scene(); if (rc != VKR_SUCCESS) return rc; }
        if (_sceneTransformsDirty) {
            std::vector<float> local(_flatNodes.size() * 16);
            for (size_t i = 0; i < _flatNodes.size(); ++i) std::memcpy(&local[i * 16], _flatNodes[i]->localTransform.data(), 64);
            const int rc = _api.scene_set_transforms(_vkr, _sceneHandle, 0, (uint32_t)_flatNodes.size(), local.data());
            if (rc != VKR_SUCCESS) return rc;
            _sceneTransformsDirty = false;
        }
        Frustum camFrustum;
        extract_frustum_planes(sd.viewproj, camFrustum);
        return _api.draw_scene(_vkr, _sceneHandle, reinterpret_cast<const VkrGPUSceneData*>(&sd), &camFrustum.planes[0].x, &fp);
    }
    return _api.draw(_vkr, reinterpret_cast<const VkrGPUSceneData*>(&sd), _drawList.data(), _drawList.size(), &fp);
}

int VulkanEngine::present(void* pinnedDst, size_t bytes) {
    return _api.readback_async ? _api.readback_async(_vkr, VKR_TARGET_RGBA8, pinnedDst, bytes) : _api.readback(_vkr, VKR_TARGET_RGBA8, pinnedDst, bytes);
}
