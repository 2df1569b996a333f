// vk_camera.h — class Camera of src/vk_camera.{h,cpp} without the SDL input half (no window).
#pragma once
#include "vk_math.h"

class Camera {
public:
    vkm::vec3 position;
    vkm::vec3 velocity;
    float pitch = 0;
    float yaw = 0;
    float fov = 60.f;          // src/vk_camera.h:14-16
    float zNear = 0.1f;
    float zFar = 100.f;

    // src/vk_camera.cpp:26-31
    void update(float dt) {
        vkm::mat4 cameraRotation = getRotationMatrix();
        vkm::vec4 t = cameraRotation * vkm::vec4{velocity.x, velocity.y, velocity.z, 0.f};
        position = position + vkm::vec3{t.x, t.y, t.z} * dt * 5.0f;
    }
    // src/vk_camera.cpp:33-38
    vkm::mat4 getViewMatrix() const {
        vkm::mat4 cameraTranslation = vkm::translate(position);
        vkm::mat4 cameraRotation = getRotationMatrix();
        return vkm::inverse(cameraTranslation * cameraRotation);
    }
    // src/vk_camera.cpp:40-46
    vkm::mat4 getRotationMatrix() const {
        vkm::quat pitchRotation = vkm::angleAxis(pitch, vkm::vec3{1.f, 0.f, 0.f});
        vkm::quat yawRotation = vkm::angleAxis(yaw, vkm::vec3{0.f, 1.f, 0.f});
        return vkm::toMat4(yawRotation) * vkm::toMat4(pitchRotation);
    }
    // src/vk_camera.cpp:48-54
    vkm::mat4 getProjectionMatrix(float aspectRatio) const {
        vkm::mat4 proj = vkm::perspectiveRH_ZO(vkm::radians(fov), aspectRatio, zNear, zFar);
        proj.c[1][1] *= -1;    // glm: y up, vulkan: y down
        return proj;
    }
};
