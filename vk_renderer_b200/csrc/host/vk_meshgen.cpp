// vk_meshgen.cpp — procedural meshes and textures (deterministic; see vk_loader.h).
#include <algorithm>
#include <cmath>

#include "vk_loader.h"

using namespace vkm;

static Vertex vtx(vec3 p, vec3 n, float u, float v) {
    Vertex o;
    o.position = p; o.normal = n; o.uv_x = u; o.uv_y = v; o.color = vec4{1, 1, 1, 1};
    return o;
}

MeshData make_cube(float h) {
    MeshData m;
    const vec3 N[6] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    const vec3 A[6] = {{0, 0, -1}, {0, 0, 1}, {1, 0, 0}, {1, 0, 0}, {1, 0, 0}, {-1, 0, 0}};
    const vec3 B[6] = {{0, 1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}, {0, 1, 0}, {0, 1, 0}};
    const float S[4] = {-1, 1, 1, -1}, T[4] = {-1, -1, 1, 1};
    for (int f = 0; f < 6; ++f) {
        uint32_t base = (uint32_t)m.vertices.size();
        for (int k = 0; k < 4; ++k) {
            vec3 p = (N[f] + A[f] * S[k] + B[f] * T[k]) * h;
            m.vertices.push_back(vtx(p, N[f], (S[k] + 1) * 0.5f, (1 - T[k]) * 0.5f));
        }
        const uint32_t q[6] = {0, 1, 2, 0, 2, 3};
        for (uint32_t i : q) m.indices.push_back(base + i);
    }
    return m;
}

MeshData make_uv_sphere(int rings, int segments, float radius) {
    MeshData m;
    const float PI = 3.14159265358979323846f;
    for (int i = 0; i <= rings; ++i)
        for (int j = 0; j <= segments; ++j) {
            float theta = (float)i / (float)rings * PI, phi = (float)j / (float)segments * 2.0f * PI;
            vec3 n{std::sin(theta) * std::cos(phi), std::cos(theta), std::sin(theta) * std::sin(phi)};
            m.vertices.push_back(vtx(n * radius, n, (float)j / (float)segments, (float)i / (float)rings));
        }
    for (int i = 0; i < rings; ++i)
        for (int j = 0; j < segments; ++j) {
            uint32_t a = (uint32_t)(i * (segments + 1) + j), b = a + (uint32_t)segments + 1;
            const uint32_t t[6] = {a, a + 1, b, a + 1, b + 1, b};
            for (uint32_t x : t) m.indices.push_back(x);
        }
    return m;
}

MeshData make_plane(int nx, int nz, float sizeX, float sizeZ, float uvScale) {
    MeshData m;
    for (int i = 0; i <= nz; ++i)
        for (int j = 0; j <= nx; ++j) {
            float fx = (float)j / (float)nx, fz = (float)i / (float)nz;
            m.vertices.push_back(vtx(vec3{(fx - 0.5f) * sizeX, 0.0f, (fz - 0.5f) * sizeZ}, vec3{0, 1, 0}, fx * uvScale, fz * uvScale));
        }
    for (int i = 0; i < nz; ++i)
        for (int j = 0; j < nx; ++j) {
            uint32_t a = (uint32_t)(i * (nx + 1) + j), b = a + (uint32_t)nx + 1;
            const uint32_t t[6] = {a, b, a + 1, a + 1, b, b + 1};
            for (uint32_t x : t) m.indices.push_back(x);
        }
    return m;
}

// value noise + checker, RGBA8; cutout: circular alpha holes (exercises `discard`)
std::vector<uint8_t> make_noise_texture(int size, int cells, uint32_t seed, bool cutout) {
    Pcg32 rng(seed);
    const int G = 17;
    float lattice[G][G];
    for (auto& row : lattice) for (float& v : row) v = rng.uniform();
    const float c0[3] = {rng.uniform(120, 255), rng.uniform(120, 255), rng.uniform(120, 255)};
    const float c1[3] = {rng.uniform(20, 140), rng.uniform(20, 140), rng.uniform(20, 140)};
    std::vector<uint8_t> img((size_t)size * size * 4);
    for (int y = 0; y < size; ++y)
        for (int x = 0; x < size; ++x) {
            float gx = (float)x / (float)size * (G - 1), gy = (float)y / (float)size * (G - 1);
            int ix = std::min((int)gx, G - 2), iy = std::min((int)gy, G - 2);
            float fx = gx - ix, fy = gy - iy;
            float n = (lattice[iy][ix] * (1 - fx) + lattice[iy][ix + 1] * fx) * (1 - fy) +
                      (lattice[iy + 1][ix] * (1 - fx) + lattice[iy + 1][ix + 1] * fx) * fy;
            bool chk = (((x * cells) / size + (y * cells) / size) & 1) != 0;
            uint8_t* p = &img[((size_t)y * size + x) * 4];
            for (int k = 0; k < 3; ++k) {
                float v = (chk ? c1[k] : c0[k]) * (0.7f + 0.3f * n);
                p[k] = (uint8_t)std::min(255.0f, std::max(0.0f, v));
            }
            p[3] = 255;
            if (cutout) {
                float cell = (float)size / 8.0f;
                float cx = std::fmod((float)x, cell) - cell / 2, cy = std::fmod((float)y, cell) - cell / 2;
                if (cx * cx + cy * cy < (cell * 0.3f) * (cell * 0.3f)) p[3] = 0;
            }
        }
    return img;
}

std::vector<std::vector<uint8_t>> make_sky_faces(int size, uint32_t seed) {
    Pcg32 rng(seed);
    const float tint[6][3] = {{255, 180, 150}, {150, 200, 255}, {120, 160, 255}, {90, 80, 70}, {200, 255, 200}, {255, 240, 160}};
    std::vector<std::vector<uint8_t>> faces(6, std::vector<uint8_t>((size_t)size * size * 4));
    for (int f = 0; f < 6; ++f)
        for (int y = 0; y < size; ++y)
            for (int x = 0; x < size; ++x) {
                float g = 0.55f + 0.45f * (1.0f - (float)y / (float)size);
                float stripe = (float)(((x / std::max(1, size / 8)) & 1) * 18);
                float noise = rng.uniform(-10, 10);
                uint8_t* p = &faces[f][((size_t)y * size + x) * 4];
                for (int k = 0; k < 3; ++k) p[k] = (uint8_t)std::min(255.0f, std::max(0.0f, tint[f][k] * g + noise - stripe));
                p[3] = 255;
            }
    return faces;
}
