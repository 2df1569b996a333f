// vk_loader.cpp — procedural BASELINE scenes built through the engine's own loader calls.
#include "vk_loader.h"

#include <cmath>
#include <string>

using namespace vkm;

namespace {

struct Builder {
    VulkanEngine* e;
    Pcg32 rng;
    uint64_t tris = 0;
    int nameCounter = 0;
    Builder(VulkanEngine* eng, uint32_t seed) : e(eng), rng(seed) {}

    MaterialInstance* material(vec4 color, float metallic, float roughness, int albedoID = 0, int metalRoughID = 0,
                               const char* tmpl = "Opaque") {
        MaterialConstants c{};
        c.colorFactors = color;
        c.metal_rough_factors = vec4{metallic, roughness, 0, 0};
        c.albedoID = albedoID; c.normalID = 1; c.metalRoughID = metalRoughID;   // defaults: src/vk_loader.cpp:343-345
        return e->_materialSystem.build_instance(e->_materialSystem.get_template(tmpl), c);
    }
    // one uploaded mesh
    std::shared_ptr<MeshAsset> mesh(const std::string& name, MeshData& md, MaterialInstance* mat) {
        std::vector<GeoSurface> s{GeoSurface{0, (uint32_t)md.indices.size(), mat}};
        return e->add_mesh(name, md.indices, md.vertices, s);
    }
    // same GPU buffers, another material (the reference ties materials to surfaces, so an instance with a
    // different material is a second MeshAsset sharing meshBuffer)
    std::shared_ptr<MeshAsset> variant(const std::shared_ptr<MeshAsset>& base, MaterialInstance* mat) {
        auto m = std::make_shared<MeshAsset>(*base);
        m->name = base->name + "#" + std::to_string(nameCounter++);
        for (auto& s : m->surfaces) s.material = mat;
        e->_meshAssets[m->name] = m;
        return m;
    }
    void place(const std::shared_ptr<MeshAsset>& m, const mat4& local) {
        auto n = std::make_shared<Node>();
        n->mesh = m.get();
        n->localTransform = local;
        e->_sceneRoot->addChild(n);
        for (auto& s : m->surfaces) tris += s.count / 3;
    }
    vec4 randomColor(float lo) { return vec4{rng.uniform(lo, 1.0f), rng.uniform(lo, 1.0f), rng.uniform(lo, 1.0f), 1.0f}; }
};

// config 1: 8x8 grid of cubes / spheres(28,28), untextured, mesh.frag, no shadows — ~50.6k triangles
void config1(Builder& b) {
    VulkanEngine* e = b.e;
    MeshData cube = make_cube(0.9f), sph = make_uv_sphere(28, 28, 1.0f);
    auto mc = b.mesh("cube", cube, nullptr), ms = b.mesh("sphere", sph, nullptr);
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 8; ++j) {
            MaterialInstance* mat = b.material(b.randomColor(0.2f), 0.0f, 1.0f);
            b.place(b.variant(((i + j) & 1) ? ms : mc, mat), translate(vec3{(i - 3.5f) * 3.0f, 0.0f, (j - 3.5f) * 3.0f}));
        }
    e->_mainCamera.position = vec3{0, 6, 22}; e->_mainCamera.pitch = -0.25f; e->_mainCamera.yaw = 0;
    e->_fragShader = VKR_FRAG_MESH_LAMBERT;
    e->_enableShadows = false; e->_enableBackground = false; e->_enablePostprocess = false;
}

// config 2 (and 5): 16x16 spheres(29,29) + 200x200-quad ground, PBR, PCF shadows, tonemap — ~511k triangles
void config2(Builder& b) {
    VulkanEngine* e = b.e;
    MeshData sph = make_uv_sphere(29, 29, 0.8f), ground = make_plane(200, 200, 60.0f, 60.0f, 20.0f);
    auto ms = b.mesh("sphere", sph, nullptr);
    auto mg = b.mesh("ground", ground, b.material(vec4{0.8f, 0.8f, 0.8f, 1}, 0.0f, 0.9f));
    b.place(mg, translate(vec3{0, -0.8f, 0}));
    for (int i = 0; i < 16; ++i)
        for (int j = 0; j < 16; ++j) {
            float metallic = b.rng.uniform() < 0.3f ? 1.0f : 0.0f;
            MaterialInstance* mat = b.material(b.randomColor(0.2f), metallic, b.rng.uniform(0.2f, 0.9f));
            b.place(b.variant(ms, mat), translate(vec3{(i - 7.5f) * 2.5f, 0.0f, (j - 7.5f) * 2.5f}));
        }
    e->_mainCamera.position = vec3{0, 9, 30}; e->_mainCamera.pitch = -0.3f; e->_mainCamera.yaw = 0;
    e->_enableShadows = true; e->_shadowMode = 1; e->_enablePostprocess = true;
}

// config 3: "hall" 60x20x30 — textured walls, two rows of columns, ornaments, alpha-cutout drapes, skybox
void config3(Builder& b, uint32_t seed) {
    VulkanEngine* e = b.e;
    std::vector<int> tex, cut;
    for (int k = 0; k < 32; ++k) { auto img = make_noise_texture(1024, 4 + (k % 5) * 4, seed + 100 + k, false); tex.push_back(e->add_texture(img.data(), 1024, 1024)); }
    for (int k = 0; k < 8; ++k) { auto img = make_noise_texture(1024, 8, seed + 200 + k, true); cut.push_back(e->add_texture(img.data(), 1024, 1024)); }
    auto sky = make_sky_faces(512, seed + 300);
    const void* faces[6];
    for (int f = 0; f < 6; ++f) faces[f] = sky[f].data();
    e->set_cubemap(faces, 512, 512);
    auto texMat = [&](int k) {   // albedo texture k; metal-rough read from another texture's G/B (loader quirk A-2)
        return b.material(vec4{1, 1, 1, 1}, b.rng.uniform(0.0f, 0.6f), b.rng.uniform(0.4f, 1.0f), tex[k % 32], tex[(k * 7 + 3) % 32]);
    };
    MeshData wallL = make_plane(120, 60, 60.0f, 30.0f, 12.0f), wallS = make_plane(40, 60, 20.0f, 30.0f, 6.0f), wallB = make_plane(120, 40, 60.0f, 20.0f, 12.0f);
    auto floorM = b.mesh("floor", wallL, texMat(0));
    b.place(floorM, translate(vec3{0, 0, 0}));
    // side walls and end wall: planes rotated to face inward (the ceiling is left open to the sky)
    mat4 rotX(1.0f); rotX.c[1][1] = 0; rotX.c[1][2] = 1; rotX.c[2][1] = -1; rotX.c[2][2] = 0;      // +Y -> +Z
    mat4 rotXn(1.0f); rotXn.c[1][1] = 0; rotXn.c[1][2] = -1; rotXn.c[2][1] = 1; rotXn.c[2][2] = 0;  // +Y -> -Z
    mat4 rotZ(1.0f); rotZ.c[0][0] = 0; rotZ.c[0][1] = 1; rotZ.c[1][0] = -1; rotZ.c[1][1] = 0;       // +Y -> -X: faces inward at x = +30
    auto back = b.mesh("wall_back", wallB, texMat(1));
    b.place(back, translate(vec3{0, 10, -15}) * rotX);
    auto front = b.variant(back, texMat(2));
    b.place(front, translate(vec3{0, 10, 15}) * rotXn);
    auto side = b.mesh("wall_side", wallS, texMat(3));
    b.place(side, translate(vec3{30, 10, 0}) * rotZ);
    MeshData col = make_uv_sphere(60, 120, 1.0f);                 // 14,400 triangles, stretched into a column
    auto colM = b.mesh("column", col, nullptr);
    for (int r = 0; r < 2; ++r)
        for (int i = 0; i < 20; ++i)
            b.place(b.variant(colM, texMat(4 + i + r * 20)), translate(vec3{-27.0f + i * 2.8f, 7.0f, r ? 8.0f : -8.0f}) * scale(vec3{0.6f, 7.0f, 0.6f}));
    MeshData orn = make_uv_sphere(40, 60, 0.5f);                  // 4,800 triangles
    auto ornM = b.mesh("ornament", orn, nullptr);
    for (int i = 0; i < 268; ++i)
        b.place(b.variant(ornM, texMat(i)), translate(vec3{b.rng.uniform(-28, 28), b.rng.uniform(0.5f, 12.0f), b.rng.uniform(-13, 13)}));
    MeshData drape = make_plane(50, 50, 6.0f, 6.0f, 2.0f);        // 5,000 triangles, alpha cut-out
    auto drapeM = b.mesh("drape", drape, nullptr);
    for (int i = 0; i < 20; ++i) {
        MaterialInstance* m = b.material(vec4{1, 1, 1, 1}, 0.0f, 0.9f, cut[i % 8], 0);
        b.place(b.variant(drapeM, m), translate(vec3{-25.0f + i * 2.6f, 6.0f + (i % 3), (i & 1) ? 3.0f : -3.0f}) * rotX);
    }
    e->_mainCamera.position = vec3{-24, 6, 0}; e->_mainCamera.pitch = -0.1f; e->_mainCamera.yaw = -1.5707963f;
    e->_enableShadows = true; e->_shadowMode = 1; e->_enablePostprocess = true; e->_enableBackground = false;
}

// config 4: one 10k-triangle sphere (rings 50, segments 100) x 1000 instances on a 10x10x10 lattice
void config4(Builder& b, uint32_t seed) {
    VulkanEngine* e = b.e;
    auto sky = make_sky_faces(512, seed + 300);
    const void* faces[6];
    for (int f = 0; f < 6; ++f) faces[f] = sky[f].data();
    e->set_cubemap(faces, 512, 512);
    MeshData sph = make_uv_sphere(50, 100, 1.0f);
    auto ms = b.mesh("sphere", sph, nullptr);
    std::vector<std::shared_ptr<MeshAsset>> variants;
    for (int k = 0; k < 16; ++k) {
        float metallic = b.rng.uniform() < 0.3f ? 1.0f : 0.0f;
        variants.push_back(b.variant(ms, b.material(b.randomColor(0.2f), metallic, b.rng.uniform(0.2f, 0.9f))));
    }
    int k = 0;
    for (int ix = 0; ix < 10; ++ix)
        for (int iy = 0; iy < 10; ++iy)
            for (int iz = 0; iz < 10; ++iz)
                b.place(variants[(k++) % 16], translate(vec3{(ix - 4.5f) * 3.0f, (iy - 4.5f) * 3.0f, (iz - 4.5f) * 3.0f}));
    e->_mainCamera.position = vec3{0, 7, 40}; e->_mainCamera.pitch = -0.17f; e->_mainCamera.yaw = 0;
    e->_enableShadows = true; e->_shadowMode = 1; e->_enablePostprocess = true;
}

// config 6 (not a BASELINE config: the object-count case of SURVEY 8f-2, "the CPU per-frame O(objects) work that dominates the
// reference at high object counts"): 16,000 small spheres (168 triangles each, 2.7 M in all) under 40 group nodes of 400 children
// each, a 40 x 10 x 40 lattice seen from inside, so that is_visible rejects part of it; PBR, PCF shadows, tonemap
void config6(Builder& b) {
    VulkanEngine* e = b.e;
    MeshData sph = make_uv_sphere(8, 12, 0.35f);
    auto ms = b.mesh("sphere", sph, nullptr);
    std::vector<std::shared_ptr<MeshAsset>> variants;
    for (int k = 0; k < 16; ++k) {
        float metallic = b.rng.uniform() < 0.3f ? 1.0f : 0.0f;
        variants.push_back(b.variant(ms, b.material(b.randomColor(0.2f), metallic, b.rng.uniform(0.2f, 0.9f))));
    }
    int k = 0;
    for (int ix = 0; ix < 40; ++ix) {
        auto group = std::make_shared<Node>();
        group->localTransform = translate(vec3{(ix - 19.5f) * 1.5f, 0.0f, 0.0f});
        e->_sceneRoot->addChild(group);
        for (int iy = 0; iy < 10; ++iy)
            for (int iz = 0; iz < 40; ++iz) {
                auto n = std::make_shared<Node>();
                n->mesh = variants[(k++) % 16].get();
                n->localTransform = translate(vec3{0.0f, (iy - 4.5f) * 1.5f, (iz - 19.5f) * 1.5f});
                group->addChild(n);
                for (auto& s : n->mesh->surfaces) b.tris += s.count / 3;
            }
    }
    e->_mainCamera.position = vec3{0, 1.0f, 12.0f}; e->_mainCamera.pitch = -0.05f; e->_mainCamera.yaw = 0.3f;
    e->_enableShadows = true; e->_shadowMode = 1; e->_enablePostprocess = true;
}

}  // namespace

uint64_t load_procedural(VulkanEngine* engine, int config, uint32_t seed) {
    if (!engine || !engine->_vkr) return 0;
    Builder b(engine, seed);
    switch (config) {
        case 1: config1(b); break;
        case 2: case 5: config2(b); break;
        case 3: config3(b, seed); break;
        case 4: config4(b, seed); break;
        case 6: config6(b); break;
        default: return 0;
    }
    return b.tris;
}

void orbit_camera(VulkanEngine* engine, uint32_t seed, int view) {
    Pcg32 rng(seed + 7777u * (uint32_t)view);
    const float ang = rng.uniform(0.0f, 6.2831853f), h = rng.uniform(4.0f, 12.0f), R = 30.0f;
    Camera& c = engine->_mainCamera;
    c.position = vec3{R * std::sin(ang), h, R * std::cos(ang)};
    c.yaw = ang;                                   // looking back at the origin: forward = -Z rotated by yaw
    c.pitch = -std::atan2(h, R);
}
