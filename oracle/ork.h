/*
 * ork.h — CPU ORACLE for the per-frame raster + shading path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (libvkr / libvkr_engine /
 * vk_renderer_b200) includes, links or executes anything under oracle/.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may.
 *
 * PARITY UNPINNED for the fixed-function half: the reference's path runs inside
 * a Vulkan driver that cannot be brought up on this box and the reference has no
 * tests or golden images.  What IS pinned: the shader
 * arithmetic (oracle/refshader builds shaders/mesh_pbr.frag, mesh.frag and
 * tonemap.comp from the reference checkout as C++ against the vendored glm and
 * tests/golden/*.json hold its outputs) and the host math (golden vectors from
 * the vendored glm 0.9.9.7).
 *
 * The API mirrors include/vkr.h one to one (ork_ instead of vkr_) so the parity
 * tests drive both with the same calls and the same host buffers.
 */
#ifndef ORK_H
#define ORK_H
#include "../include/vkr.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrkDevice OrkDevice;

int ork_create(const VkrConfig* cfg, OrkDevice** out);
void ork_destroy(OrkDevice* dev);
int ork_set_threads(OrkDevice* dev, int nthreads);   /* 0 -> hardware_concurrency */
int ork_upload_mesh(OrkDevice* dev, const uint32_t* indices, size_t indexCount,
                    const VkrVertex* vertices, size_t vertexCount, VkrMesh* out);
int ork_create_image(OrkDevice* dev, const void* rgba8, uint32_t width, uint32_t height,
                     VkrFormat format, int mipmapped, VkrImage* out);
int ork_update_bindless_texture(OrkDevice* dev, int index, VkrImage image, VkrSampler sampler);
int ork_set_cubemap(OrkDevice* dev, const void* const faces[6], uint32_t width, uint32_t height,
                    VkrFormat format, int mipmapped);
int ork_set_cubemap_levels(OrkDevice* dev, const void* const* levels, uint32_t width, uint32_t height, uint32_t levelCount,
                           VkrFormat format);
int ork_build_material(OrkDevice* dev, const VkrMaterialConstants* params, VkrMaterialPass pass,
                       VkrMaterial* out);
int ork_draw(OrkDevice* dev, const VkrGPUSceneData* scene, const VkrRenderObject* objects,
             size_t objectCount, const VkrFrameParams* params);
int ork_readback(OrkDevice* dev, VkrTarget what, void* dst, size_t dstBytes);
int ork_stats(OrkDevice* dev, VkrEngineStats* out);
/* Read one mip level of an image (tests of generate_mipmaps parity). level<0: returns level count. */
int ork_image_level(OrkDevice* dev, VkrImage image, int level, void* dst, size_t dstBytes,
                    uint32_t* w, uint32_t* h);
const char* ork_last_error(OrkDevice* dev);

/* Shader-level entry points (pinned against oracle/_ref/librefshader.so). */
typedef struct OrkFragIn {
    float normal[3], color[3], uv[2], lightVec[3], posWorld[4], posView[4];
    float duvdx[2], duvdy[2];
} OrkFragIn;
/* Runs the restated mesh_pbr.frag / mesh.frag main() on one fragment against the device's
 * scene state (textures, shadow map as last rendered/loaded). Returns 1 if discarded. */
int ork_shade_fragment(OrkDevice* dev, const VkrGPUSceneData* scene, const VkrMaterialConstants* mat,
                       int fragShader, const OrkFragIn* in, float outColor[4]);
int ork_load_shadow(OrkDevice* dev, const float* depth, size_t count);  /* overwrite shadow layers */
void ork_brdf(const float N[3], const float V[3], const float L[3], float roughness, const float F0[3],
              float* D, float* G, float F[3]);
void ork_tonemap(const float in[3], float out[3]);
/* gradient_color.comp for pixel row y of a height-row image. */
void ork_gradient(const VkrComputePushConstants* push, int y, int height, float out[4]);
/* mesh.vert main(): out = gl_Position[4], normal[3], color[3], uv[2], lightVec[3], posWorld[4], posView[4]. */
void ork_vertex(const VkrGPUSceneData* scene, const float world[16], const VkrVertex* v, float out[23]);
/* shadow.vert main(): gl_Position for cascade `cascade`. */
void ork_shadow_vertex(const VkrGPUSceneData* scene, const float world[16], int cascade, const VkrVertex* v, float out[4]);
/* The per-pixel form of skybox.vert's outUVW at window position (sx, sy). */
void ork_skybox_dir(const VkrGPUSceneData* scene, int width, int height, float sx, float sy, float out[3]);

/* The oracle's texture unit as callbacks, so oracle/_ref/librefshader.so (the reference's shader text)
 * can be run against the same fixed-function sampling when golden vectors are generated. */
typedef struct OrkTexCtx { OrkDevice* dev; float duvdx[2], duvdy[2]; } OrkTexCtx;
void ork_cb_tex2d(void* ctx, int id, float u, float v, float out[4]);
float ork_cb_shadow(void* ctx, float u, float v, int layer);
void ork_cb_cube(void* ctx, const float dir[3], float out[4]);

#ifdef __cplusplus
}
#endif
#endif
