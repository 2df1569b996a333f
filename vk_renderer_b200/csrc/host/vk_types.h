// vk_types.h — the engine's scene-side types, same names and semantics as the reference's
// src/vk_types.h and src/vk_materials.h with Vulkan handles replaced by libvkr handles.
#pragma once
#include <array>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/vkr.h"
#include "vk_math.h"

#define NUM_CASCADES 4            // src/vk_types.h:16

// struct Vertex — src/vk_types.h:85-91 (48 bytes, layout-identical to VkrVertex)
struct Vertex {
    vkm::vec3 position;
    float uv_x;
    vkm::vec3 normal;
    float uv_y;
    vkm::vec4 color;
};
static_assert(sizeof(Vertex) == 48 && sizeof(Vertex) == sizeof(VkrVertex), "Vertex layout");

// struct GPUSceneData — src/vk_types.h:93-102 (512 bytes)
struct GPUSceneData {
    vkm::mat4 view;
    vkm::mat4 proj;
    vkm::mat4 viewproj;
    vkm::mat4 lightViewproj[NUM_CASCADES];
    float cascadeDistances[NUM_CASCADES];
    vkm::vec4 ambientColor;
    vkm::vec4 sunlightDirection;
    vkm::vec4 sunlightColor;
};
static_assert(sizeof(GPUSceneData) == 512 && sizeof(GPUSceneData) == sizeof(VkrGPUSceneData), "GPUSceneData layout");

struct EngineStats {              // src/vk_types.h:78-83
    float frametime;
    int triangle_count;
    int drawcall_count;
    float mesh_draw_time;
};

struct ComputePushConstants { vkm::vec4 data1, data2, data3, data4; };   // src/vk_types.h:50-55

struct CSMData {                  // src/vk_types.h (lightMatrices + planeDistances)
    vkm::mat4 lightMatrices[NUM_CASCADES];
    float planeDistances[NUM_CASCADES];
};

// ---- materials: src/vk_materials.h:8-36
enum class MaterialPass : uint8_t { MainColor, Transparent, Other };

struct MaterialConstants {
    vkm::vec4 colorFactors;
    vkm::vec4 metal_rough_factors;
    int albedoID;
    int normalID;
    int metalRoughID;
    int padding;
};
static_assert(sizeof(MaterialConstants) == 48, "MaterialConstants layout");

struct MaterialTemplate {         // pipeline + layout collapse to the pass type
    std::string name;
    MaterialPass passType;
};

struct MaterialInstance {
    MaterialTemplate* pipeline = nullptr;
    VkrMaterial materialSet = 0;  // replaces VkDescriptorSet + paramsBuffer
    MaterialPass passType = MaterialPass::MainColor;
    MaterialConstants params{};
};

// ---- geometry: src/vk_types.h:104-137
struct GeoSurface {
    uint32_t startIndex;
    uint32_t count;
    MaterialInstance* material = nullptr;
};

struct Bounds {
    vkm::vec3 origin;
    float sphereRadius;
    vkm::vec3 extents;
};

struct Frustum { std::array<vkm::vec4, 6> planes; };

struct MeshAsset {
    std::string name;
    VkrMesh meshBuffer = 0;       // replaces AllocatedBuffer + device address
    std::vector<GeoSurface> surfaces;
    size_t indexOffset = 0;       // = vertexCount * sizeof(Vertex), kept for parity with the reference
    Bounds bounds{};
};

struct RenderObject {
    MeshAsset* mesh;
    MaterialInstance* material;
    vkm::mat4 transform;
    uint32_t indexCount;
    uint32_t firstIndex;
};

// struct Node — src/vk_types.h:139-169
struct Node : std::enable_shared_from_this<Node> {
    std::weak_ptr<Node> parent;
    std::vector<std::shared_ptr<Node>> children;
    vkm::mat4 localTransform{1.0f};
    vkm::mat4 worldTransform{1.0f};
    MeshAsset* mesh = nullptr;

    void refreshTransform(const vkm::mat4& parentMatrix, std::vector<RenderObject>& outDrawList) {
        worldTransform = parentMatrix * localTransform;
        if (mesh) {
            for (auto& surface : mesh->surfaces) {
                RenderObject obj;
                obj.mesh = mesh;
                obj.transform = worldTransform;
                obj.material = surface.material;
                obj.indexCount = surface.count;
                obj.firstIndex = surface.startIndex;
                outDrawList.push_back(obj);
            }
        }
        for (auto& c : children) c->refreshTransform(worldTransform, outDrawList);
    }
    void addChild(std::shared_ptr<Node> child) {
        child->parent = weak_from_this();
        children.push_back(child);
    }
};
