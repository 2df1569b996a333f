// refshader.cpp — the REFERENCE's own shader text, compiled as C++ against the reference's vendored
// glm 0.9.9.7 (TEST INFRASTRUCTURE ONLY; built into oracle/_ref/librefshader.so by `make -C oracle ref`
// where /root/reference is mounted; never shipped, never linked by the product).
//
// The *.inc files are produced from /root/reference/shaders/* by gen.py (syntactic rewrites only).
// This file supplies what the GLSL environment would: sampler/image types, texture() routed to
// caller-supplied callbacks (the texture unit is fixed function, not shader arithmetic), `discard`,
// gl_* built-ins.  Used to (a) validate oracle/ork.cpp's restatement of the shader arithmetic and
// (b) generate tests/golden/shader_*.json (tools/gen_golden.py).
#define GLM_FORCE_SWIZZLE
#define GLM_FORCE_DEPTH_ZERO_TO_ONE   // src/CMakeLists.txt:24
#include <glm/glm.hpp>

#include <cstring>

using namespace glm;

extern "C" {
typedef void (*RefTex2D)(void* user, int id, float u, float v, float out[4]);
typedef float (*RefTexShadow)(void* user, float u, float v, int layer);
typedef void (*RefTexCube)(void* user, const float dir[3], float out[4]);
}

namespace env {
RefTex2D tex2d = nullptr;
RefTexShadow texShadow = nullptr;
RefTexCube texCube = nullptr;
void* user = nullptr;
int shadowSize = 2048;
bool discarded = false;
vec4 imageIn(0.0f), imageOut(0.0f);
ivec2 imageExtent(1, 1);
}  // namespace env

struct sampler2D { int id; };
struct sampler2DTable { sampler2D operator[](int i) const { return sampler2D{i}; } };
struct sampler2DArray {};
struct samplerCube {};
struct image2D {};

inline vec4 texture(sampler2D s, vec2 uv) {
    float o[4];
    env::tex2d(env::user, s.id, uv.x, uv.y, o);
    return vec4(o[0], o[1], o[2], o[3]);
}
inline vec4 texture(sampler2DArray, vec3 c) {
    return vec4(env::texShadow(env::user, c.x, c.y, (int)c.z), 0.0f, 0.0f, 1.0f);
}
inline vec4 texture(samplerCube, vec3 d) {
    float o[4], dir[3] = {d.x, d.y, d.z};
    env::texCube(env::user, dir, o);
    return vec4(o[0], o[1], o[2], o[3]);
}
inline ivec3 textureSize(sampler2DArray, int) { return ivec3(env::shadowSize, env::shadowSize, 4); }
inline ivec2 imageSize(image2D) { return env::imageExtent; }
inline vec4 imageLoad(image2D, ivec2) { return env::imageIn; }
inline void imageStore(image2D, ivec2, vec4 v) { env::imageOut = v; }

#define nonuniformEXT(x) (x)
#define discard do { env::discarded = true; return; } while (0)

static int gl_VertexIndex = 0;
static vec4 gl_Position;
static uvec3 gl_GlobalInvocationID;

namespace ref_pbr {
#include "mesh_pbr_frag.inc"
}
namespace ref_lambert {
#include "mesh_frag.inc"
}
namespace ref_mesh_vert {
#include "mesh_vert.inc"
}
namespace ref_shadow_vert {
#include "shadow_vert.inc"
}
namespace ref_skybox_vert {
#include "skybox_vert.inc"
}
namespace ref_skybox_frag {
#include "skybox_frag.inc"
}
namespace ref_tonemap {
#include "tonemap_comp.inc"
}
namespace ref_gradient {
#include "gradient_color_comp.inc"
}

static_assert(sizeof(ref_pbr::GPUSceneData) == 512, "GPUSceneData scalar layout (SURVEY.md §8a-a4)");
static_assert(sizeof(ref_pbr::MaterialConstants) == 48, "MaterialConstants");
static_assert(sizeof(ref_mesh_vert::Vertex) == 48, "Vertex");

extern "C" {

struct RefFragIn {   // same layout as OrkFragIn (oracle/ork.h); derivatives are consumed by the callbacks only
    float normal[3], color[3], uv[2], lightVec[3], posWorld[4], posView[4];
    float duvdx[2], duvdy[2];
};

void ref_set_env(RefTex2D t2, RefTexShadow ts, RefTexCube tc, void* user, int shadowSize) {
    env::tex2d = t2; env::texShadow = ts; env::texCube = tc; env::user = user; env::shadowSize = shadowSize;
}

// which: 0 = mesh_pbr.frag, 1 = mesh.frag.  Returns 1 when the fragment was discarded.
int ref_frag(int which, const void* scene512, const void* material48, const RefFragIn* in, float out[4]) {
    env::discarded = false;
#define RUN(NS)                                                                        \
    std::memcpy(&NS::sceneData, scene512, 512);                                        \
    std::memcpy(&NS::materialData, material48, 48);                                    \
    NS::inNormal = vec3(in->normal[0], in->normal[1], in->normal[2]);                   \
    NS::inColor = vec3(in->color[0], in->color[1], in->color[2]);                       \
    NS::inUV = vec2(in->uv[0], in->uv[1]);                                             \
    NS::inLightVec = vec3(in->lightVec[0], in->lightVec[1], in->lightVec[2]);           \
    NS::inFragPosWorld = vec4(in->posWorld[0], in->posWorld[1], in->posWorld[2], in->posWorld[3]); \
    NS::inFragPosView = vec4(in->posView[0], in->posView[1], in->posView[2], in->posView[3]);      \
    NS::outFragColor = vec4(0.0f);                                                     \
    NS::shader_main();                                                                 \
    out[0] = NS::outFragColor.x; out[1] = NS::outFragColor.y; out[2] = NS::outFragColor.z; out[3] = NS::outFragColor.w;
    if (which == 0) { RUN(ref_pbr) } else { RUN(ref_lambert) }
#undef RUN
    return env::discarded ? 1 : 0;
}

// mesh.vert: out = gl_Position[4], normal[3], color[3], uv[2], lightVec[3], posWorld[4], posView[4] (23 floats)
void ref_mesh_vert_main(const void* scene512, const float world[16], const void* vertex48, float out[23]) {
    namespace S = ref_mesh_vert;
    std::memcpy(&S::sceneData, scene512, 512);
    std::memcpy(&S::pushConstants.worldMatrix, world, 64);
    S::pushConstants.vertexBuffer.vertices = (const S::Vertex*)vertex48;
    gl_VertexIndex = 0;
    S::shader_main();
    float* o = out;
    for (int i = 0; i < 4; ++i) *o++ = gl_Position[i];
    for (int i = 0; i < 3; ++i) *o++ = S::outNormal[i];
    for (int i = 0; i < 3; ++i) *o++ = S::outColor[i];
    for (int i = 0; i < 2; ++i) *o++ = S::outUV[i];
    for (int i = 0; i < 3; ++i) *o++ = S::outLightVec[i];
    for (int i = 0; i < 4; ++i) *o++ = S::outFragPosWorld[i];
    for (int i = 0; i < 4; ++i) *o++ = S::outFragPosView[i];
}

void ref_shadow_vert_main(const void* scene512, const float world[16], int cascade, const void* vertex48, float out[4]) {
    namespace S = ref_shadow_vert;
    std::memcpy(&S::sceneData, scene512, 512);
    std::memcpy(&S::pushConstants.worldMatrix, world, 64);
    S::pushConstants.vertexBuffer.vertices = (const S::Vertex*)vertex48;
    S::pushConstants.cascadeIndex = cascade;
    gl_VertexIndex = 0;
    S::shader_main();
    for (int i = 0; i < 4; ++i) out[i] = gl_Position[i];
}

// skybox.vert: out = gl_Position[4], outUVW[3]
void ref_skybox_vert_main(const void* scene512, const void* vertex48, float out[7]) {
    namespace S = ref_skybox_vert;
    std::memcpy(&S::sceneData, scene512, 512);
    S::pushConstants.vertexBuffer.vertices = (const S::Vertex*)vertex48;
    gl_VertexIndex = 0;
    S::shader_main();
    for (int i = 0; i < 4; ++i) out[i] = gl_Position[i];
    for (int i = 0; i < 3; ++i) out[4 + i] = S::outUVW[i];
}

void ref_skybox_frag_main(const float uvw[3], float out[4]) {
    namespace S = ref_skybox_frag;
    S::inUVW = vec3(uvw[0], uvw[1], uvw[2]);
    S::shader_main();
    for (int i = 0; i < 4; ++i) out[i] = S::outFragColor[i];
}

void ref_tonemap_main(const float in[4], float out[4]) {
    env::imageExtent = ivec2(1, 1);
    env::imageIn = vec4(in[0], in[1], in[2], in[3]);
    gl_GlobalInvocationID = uvec3(0, 0, 0);
    ref_tonemap::shader_main();
    for (int i = 0; i < 4; ++i) out[i] = env::imageOut[i];
}

void ref_gradient_main(const float push64[16], int x, int y, int width, int height, float out[4]) {
    std::memcpy(&ref_gradient::PushConstants, push64, 64);
    env::imageExtent = ivec2(width, height);
    gl_GlobalInvocationID = uvec3((unsigned)x, (unsigned)y, 0);
    ref_gradient::shader_main();
    for (int i = 0; i < 4; ++i) out[i] = env::imageOut[i];
}

void ref_brdf(const float N[3], const float V[3], const float L[3], float roughness, const float F0[3],
              float* D, float* G, float F[3]) {
    vec3 n(N[0], N[1], N[2]), v(V[0], V[1], V[2]), l(L[0], L[1], L[2]);
    vec3 h = normalize(v + l);
    *D = ref_pbr::DistributionGGX(n, h, roughness);
    *G = ref_pbr::GeometrySmith(n, v, l, roughness);
    vec3 f = ref_pbr::fresnelSchlick(max(dot(h, v), 0.0f), vec3(F0[0], F0[1], F0[2]));
    F[0] = f.x; F[1] = f.y; F[2] = f.z;
}

}  // extern "C"
