// vk_engine.h — headless mirror of class VulkanEngine (src/vk_engine.h) on top of libvkr.
// Same member names and per-frame flow (update_scene -> draw); Vulkan objects, the window, the
// swapchain and ImGui are gone, the ImGui toggles stay as public members.
#pragma once
#include <memory>
#include <span>
#include <string>
#include <unordered_map>
#include <vector>

#include "vk_camera.h"
#include "vk_types.h"

struct VkExtent2D { uint32_t width, height; };

// The entry points of include/vkr.h the engine uses, as a table.  The default is libvkr (the CUDA
// path, linked); the test-suite injects the CPU oracle's ork_* functions here to render the same scene
// graph through both.  The engine library itself never links or loads the oracle.
struct VkrApi {
    int (*create)(const VkrConfig*, VkrDevice**);
    void (*destroy)(VkrDevice*);
    int (*upload_mesh)(VkrDevice*, const uint32_t*, size_t, const VkrVertex*, size_t, VkrMesh*);
    int (*create_image)(VkrDevice*, const void*, uint32_t, uint32_t, VkrFormat, int, VkrImage*);
    int (*update_bindless_texture)(VkrDevice*, int, VkrImage, VkrSampler);
    int (*set_cubemap)(VkrDevice*, const void* const*, uint32_t, uint32_t, VkrFormat, int);
    int (*build_material)(VkrDevice*, const VkrMaterialConstants*, VkrMaterialPass, VkrMaterial*);
    int (*draw)(VkrDevice*, const VkrGPUSceneData*, const VkrRenderObject*, size_t, const VkrFrameParams*);
    int (*readback)(VkrDevice*, VkrTarget, void*, size_t);
    int (*readback_async)(VkrDevice*, VkrTarget, void*, size_t);   // may be null
    int (*sync)(VkrDevice*);                                        // may be null
    int (*set_cubemap_levels)(VkrDevice*, const void* const*, uint32_t, uint32_t, uint32_t, VkrFormat);   // may be null
    // update_scene on the device (_gpuSceneUpdate); null in a backend without it (the oracle)
    int (*scene_create)(VkrDevice*, const VkrSceneNode*, size_t, const VkrSceneSurface*, size_t, VkrScene*);
    int (*scene_set_transforms)(VkrDevice*, VkrScene, uint32_t, uint32_t, const float*);
    int (*draw_scene)(VkrDevice*, VkrScene, const VkrGPUSceneData*, const float*, const VkrFrameParams*);
    int (*scene_draw_list)(VkrDevice*, VkrScene, VkrRenderObject*, size_t, size_t*);
    int (*stats)(VkrDevice*, VkrEngineStats*);
    int (*scene_destroy)(VkrDevice*, VkrScene);
};
const VkrApi& libvkr_api();

class MaterialSystem {               // src/vk_materials.h:38-50
public:
    void init(class VulkanEngine* engine);
    MaterialTemplate* get_template(const std::string& name);
    MaterialInstance* build_instance(MaterialTemplate* materialTemplate, MaterialConstants params);
    ~MaterialSystem();
private:
    VulkanEngine* _engine = nullptr;
    std::unordered_map<std::string, MaterialTemplate> _templateCache;
    std::vector<MaterialInstance*> _instances;   // the reference leaks these (src/vk_materials.cpp:62)
};

class VulkanEngine {
public:
    VkExtent2D _windowExtent{1280, 720};            // src/vk_engine.h:38
    VkExtent2D _drawExtent{1280, 720};
    VkExtent2D _shadowExtent{2048, 2048};           // :107

    VkrApi _api = libvkr_api();
    VkrDevice* _vkr = nullptr;
    int _cudaDevice = 0;
    void* _cudaStream = nullptr;
    uint32_t _vkrFlags = 0;

    EngineStats stats{};
    Camera _mainCamera;

    std::shared_ptr<Node> _sceneRoot;
    std::vector<RenderObject> _renderObjects;
    std::vector<VkrRenderObject> _drawList;         // flattened handles, rebuilt every frame

    MaterialSystem _materialSystem;
    std::unordered_map<std::string, std::shared_ptr<MeshAsset>> _meshAssets;
    int _globalTextureIndex = 0;

    // src/vk_engine.h:112-126
    vkm::vec4 _sunlightDirection{0.5f, -1.0f, -0.5f, 0.0f};
    vkm::vec4 _sunlightColor{1.f, 1.f, 1.f, 1.f};
    vkm::vec4 _ambientColor{0.1f, 0.1f, 0.1f, 1.0f};
    bool _enableShadows = false;
    int _shadowMode = 0;
    bool _enablePostprocess = false;
    bool _enableBackground = false;
    bool _hasSkybox = false;                        // _skyboxMesh != nullptr
    int _fragShader = VKR_FRAG_MESH_PBR;
    ComputePushConstants _backgroundData{{1, 0, 0, 1}, {0, 0, 1, 1}, {}, {}};   // src/vk_engine_init.cpp:504-505
    int _scissor[4] = {0, 0, 0, 0};                 // sort-first sharding: rows of this rank
    bool _cullToScissor = false;                    // also cull draws against the scissor's sub-frustum

    GPUSceneData _sceneData{};                      // what draw() last sent

    // update_scene's per-object work (refreshTransform, is_visible, the sort; src/vk_engine_run.cpp:435-480) on the device:
    // the scene graph is flattened and handed over once (vkr_scene_create), update_scene() then only moves the camera and
    // draw() sends the frame's constants and the six frustum planes (vkr_draw_scene).  _renderObjects stays empty.
    bool _gpuSceneUpdate = false;
    bool _sceneGraphDirty = true;                   // nodes were added or removed: flatten again (scene_changed())
    bool _sceneTransformsDirty = false;             // some Node::localTransform changed (transforms_changed())
    bool _haveScene = false;
    VkrScene _sceneHandle = 0;
    std::vector<Node*> _flatNodes;                  // the nodes in the order refreshTransform visits them
    void scene_changed() { _sceneGraphDirty = true; }
    void transforms_changed() { _sceneTransformsDirty = true; }
    int flatten_scene();                            // _sceneRoot -> VkrSceneNode[] + VkrSceneSurface[] -> vkr_scene_create

    int init(uint32_t width, uint32_t height, int device = 0, void* stream = nullptr, uint32_t flags = 0);
    void cleanup();
    ~VulkanEngine() { cleanup(); }

    // src/vk_loader.cpp:109-225
    VkrMesh upload_mesh(std::span<uint32_t> indices, std::span<Vertex> vertices);
    VkrImage create_image(const void* data, uint32_t w, uint32_t h, VkrFormat format, bool mipmapped);
    void update_bindless_texture(int index, VkrImage image, VkrSampler sampler);
    int add_texture(const void* rgba8, uint32_t w, uint32_t h);     // load_gltf's image loop (:275-329)
    void set_cubemap(const void* const faces[6], uint32_t w, uint32_t h);
    bool load_cubemap(const char* filename, std::string* error = nullptr);   // src/vk_loader.cpp:521-558 (KTX 1.1 / 2.0)
    std::shared_ptr<MeshAsset> add_mesh(const std::string& name, std::span<uint32_t> indices, std::span<Vertex> vertices,
                                        const std::vector<GeoSurface>& surfaces);   // load_gltf's mesh loop (:382-466)

    // src/vk_engine_run.cpp
    void update_scene(float dt);                                    // :435-459
    bool is_visible(const RenderObject& obj, const Frustum& frustum);   // :461-480
    vkm::mat4 compute_light_matrix();                               // :482-491
    vkm::mat4 getLightMatrix(const float zNear, const float zFar);  // :506-543
    CSMData compute_csmdata();                                      // :554-566
    int draw();                                                     // :68-193
    int present(void* pinnedDst, size_t bytes);                     // headless vkQueuePresentKHR
};

void extract_frustum_planes(const vkm::mat4& VP, Frustum& frustum);  // src/vk_engine_run.cpp:420-433
float compute_split(float n, float f, int i);                        // :546-552
