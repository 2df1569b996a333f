// vk_loader.h — scene construction for the headless engine.
// The reference builds its scene with load_gltf (src/vk_loader.cpp:227-518) from assets that are not in
// the tree; here deterministic procedural scenes of the BASELINE.json configurations are built through
// the same engine calls load_gltf makes (create_image + update_bindless_texture, build_instance,
// upload_mesh, Node tree).  PRNG = PCG32, seed = 0xB2000000 + config.
#pragma once
#include <cstdint>
#include <memory>
#include <vector>

#include "vk_engine.h"

struct Pcg32 {
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    explicit Pcg32(uint64_t seed, uint64_t seq = 54u) {
        state = 0; inc = (seq << 1u) | 1u; next(); state += seed; next();
    }
    uint32_t next() {
        uint64_t old = state;
        state = old * 6364136223846793005ULL + inc;
        uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((32 - rot) & 31));
    }
    float uniform() { return (float)(next() >> 8) * (1.0f / 16777216.0f); }          // [0,1)
    float uniform(float a, float b) { return a + (b - a) * uniform(); }
};

struct MeshData { std::vector<Vertex> vertices; std::vector<uint32_t> indices; };

MeshData make_cube(float half);                                   // 24 v / 12 t, per-face normals
MeshData make_uv_sphere(int rings, int segments, float radius);   // (r+1)(s+1) v, 2rs t, CCW outside
MeshData make_plane(int nx, int nz, float sizeX, float sizeZ, float uvScale);   // XZ grid facing +Y

std::vector<uint8_t> make_noise_texture(int size, int cells, uint32_t seed, bool cutout);
std::vector<std::vector<uint8_t>> make_sky_faces(int size, uint32_t seed);

// Builds BASELINE config `config` (1..5) into the engine: scene graph under _sceneRoot, camera, light
// and pass toggles.  Returns the number of triangles in the scene, or 0 on failure.
uint64_t load_procedural(VulkanEngine* engine, int config, uint32_t seed);

// Camera `view` of config 5's seeded orbit (radius 30, height U[4,12], looking at the origin).
void orbit_camera(VulkanEngine* engine, uint32_t seed, int view);
