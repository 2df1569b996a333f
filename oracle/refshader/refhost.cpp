// refhost.cpp — the REFERENCE's own host math for the path (camera, light/CSM matrices, frustum cull),
// compiled from its source text against its vendored glm (TEST INFRASTRUCTURE ONLY; part of
// oracle/_ref/librefshader.so).  host_math.inc is cut out of /root/reference/src/vk_camera.cpp and
// src/vk_engine_run.cpp by gen.py — whole function definitions, verbatim.  This file supplies only the
// declarations those definitions need (the real headers pull in Vulkan, VMA and SDL).
#define GLM_FORCE_DEPTH_ZERO_TO_ONE   // src/CMakeLists.txt:24
#define GLM_ENABLE_EXPERIMENTAL
#include <glm/glm.hpp>
#include <glm/gtx/transform.hpp>
#include <glm/gtx/quaternion.hpp>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#define NUM_CASCADES 4                                   // src/vk_types.h:16

struct Bounds { glm::vec3 origin; float sphereRadius; glm::vec3 extents; };   // src/vk_types.h:112-116
struct Frustum { std::array<glm::vec4, 6> planes; };                           // :118-120
struct MeshAsset { Bounds bounds; };                                           // :122-128 (bounds only)
struct MaterialInstance;
struct RenderObject { MeshAsset* mesh; MaterialInstance* material; glm::mat4 transform; uint32_t indexCount; uint32_t firstIndex; };
struct CSMData { glm::mat4 lightMatrices[NUM_CASCADES]; float planeDistances[NUM_CASCADES]; };
struct Extent2D { uint32_t width, height; };

class Camera {   // src/vk_camera.h:6-25 without the SDL members
public:
    glm::vec3 position;
    glm::vec3 velocity;
    float pitch = 0;
    float yaw = 0;
    float fov = 60.f;
    float zNear = 0.1f;
    float zFar = 100.f;
    glm::mat4 getViewMatrix();
    glm::mat4 getRotationMatrix();
    glm::mat4 getProjectionMatrix(float aspectRatio);
};

class VulkanEngine {   // the members the extracted functions touch (src/vk_engine.h)
public:
    Extent2D _windowExtent{1280, 720};
    Camera _mainCamera;
    glm::vec3 _sunlightDirection{0.5f, -1.f, -0.5f};
    bool is_visible(const RenderObject& obj, const Frustum& frustum);
    glm::mat4 compute_light_matrix();
    glm::mat4 getLightMatrix(const float zNear, const float zFar);
    CSMData compute_csmdata();
};

using std::pow;
#include "host_math.inc"

extern "C" {

struct RefHostOut {
    float view[16], proj[16], light[16];
    float csm_matrices[64], csm_distances[4];
    float planes[24];
};

void ref_host_math(uint32_t width, uint32_t height, const float position[3], float pitch, float yaw, const float sun[3],
                   RefHostOut* out) {
    VulkanEngine e;
    e._windowExtent = {width, height};
    e._mainCamera.position = glm::vec3(position[0], position[1], position[2]);
    e._mainCamera.velocity = glm::vec3(0.f);
    e._mainCamera.pitch = pitch;
    e._mainCamera.yaw = yaw;
    e._sunlightDirection = glm::vec3(sun[0], sun[1], sun[2]);
    glm::mat4 view = e._mainCamera.getViewMatrix();
    glm::mat4 proj = e._mainCamera.getProjectionMatrix((float)width / (float)height);   // src/vk_engine_run.cpp:108-109
    glm::mat4 light = e.compute_light_matrix();
    CSMData csm = e.compute_csmdata();
    std::memcpy(out->view, &view[0][0], 64);
    std::memcpy(out->proj, &proj[0][0], 64);
    std::memcpy(out->light, &light[0][0], 64);
    std::memcpy(out->csm_matrices, &csm.lightMatrices[0][0][0], 256);
    std::memcpy(out->csm_distances, csm.planeDistances, 16);
    Frustum f;
    extract_frustum_planes(proj * view, f);                                             // src/vk_engine_run.cpp:446-448
    std::memcpy(out->planes, &f.planes[0][0], 96);
}

// is_visible for a sphere-bounded object under `transform` (column-major), camera as above.
int ref_is_visible(uint32_t width, uint32_t height, const float position[3], float pitch, float yaw,
                   const float transform[16], const float origin[3], float sphereRadius) {
    VulkanEngine e;
    e._windowExtent = {width, height};
    e._mainCamera.position = glm::vec3(position[0], position[1], position[2]);
    e._mainCamera.pitch = pitch;
    e._mainCamera.yaw = yaw;
    glm::mat4 viewProj = e._mainCamera.getProjectionMatrix((float)width / (float)height) * e._mainCamera.getViewMatrix();
    Frustum f;
    extract_frustum_planes(viewProj, f);
    MeshAsset mesh;
    mesh.bounds.origin = glm::vec3(origin[0], origin[1], origin[2]);
    mesh.bounds.sphereRadius = sphereRadius;
    RenderObject ro{&mesh, nullptr, glm::mat4(1.f), 0, 0};
    std::memcpy(&ro.transform[0][0], transform, 64);
    return e.is_visible(ro, f) ? 1 : 0;
}

float ref_compute_split(float n, float f, int i) { return compute_split(n, f, i); }

}  // extern "C"
