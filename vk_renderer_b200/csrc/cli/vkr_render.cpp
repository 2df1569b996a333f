// vkr_render — headless stand-in for the reference's window + ImGui panel: builds a scene (a glTF file
// or one of the procedural BASELINE configurations), sets the members the panel exposes (src/vk_engine.h:112-126:
// _enableShadows, _shadowMode, _enableBackground, _enablePostprocess, sun direction/colour, ambient), runs
// update_scene() + draw() for N frames and writes the presented image as a binary PPM; prints EngineStats
// (src/vk_types.h:78-83) the way the panel shows them.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../host/vk_engine.h"
#include "../host/vk_gltf.h"
#include "../host/vk_loader.h"

static void usage() {
    std::puts(
        "usage: vkr_render (--gltf FILE | --config 1..6) [options]\n"
        "  --size WxH            draw extent (default 1280x720, the reference's window)\n"
        "  --shadow-extent N     shadow map size (default 2048)\n"
        "  --shadows 0|1         _enableShadows          --shadow-mode 0..3   hard / PCF / PCSS / CSM\n"
        "  --background 0|1      _enableBackground       --postprocess 0|1    _enablePostprocess\n"
        "  --lambert             mesh.frag instead of mesh_pbr.frag\n"
        "  --cubemap FILE.ktx    skybox cube map (KTX 1.1 / 2.0; RGBA8, sRGB8_A8 or RGBA16F), as load_cubemap\n"
        "  --camera X,Y,Z[,PITCH,YAW]    --sun X,Y,Z    --ambient R,G,B\n"
        "  --frames N            frames to render (default 1; stats are of the last one)\n"
        "  --gpu-scene           update_scene's per-object work (transforms, is_visible, draw list) on the device\n"
        "  --device N            CUDA device ordinal\n"
        "  --out FILE.ppm        write the presented image (P6)\n");
}

static std::vector<float> floats(const char* s) {
    std::vector<float> v;
    while (*s) { char* e; v.push_back(std::strtof(s, &e)); if (e == s) break; s = *e == ',' ? e + 1 : e; }
    return v;
}

int main(int argc, char** argv) {
    std::string gltf, out, cubemap;
    int config = 0, frames = 1, device = 0, shadows = -1, shadowMode = -1, background = -1, post = -1;
    bool lambert = false, gpuScene = false;
    uint32_t W = 1280, H = 720, S = 2048;
    std::vector<float> cam, sun, ambient;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) { usage(); std::exit(2); } return argv[++i]; };
        if (a == "--help" || a == "-h") { usage(); return 0; }
        else if (a == "--gltf") gltf = next();
        else if (a == "--config") config = std::atoi(next());
        else if (a == "--size") { if (std::sscanf(next(), "%ux%u", &W, &H) != 2) { usage(); return 2; } }
        else if (a == "--shadow-extent") S = (uint32_t)std::atoi(next());
        else if (a == "--shadows") shadows = std::atoi(next());
        else if (a == "--shadow-mode") shadowMode = std::atoi(next());
        else if (a == "--background") background = std::atoi(next());
        else if (a == "--postprocess") post = std::atoi(next());
        else if (a == "--lambert") lambert = true;
        else if (a == "--gpu-scene") gpuScene = true;
        else if (a == "--cubemap") cubemap = next();
        else if (a == "--camera") cam = floats(next());
        else if (a == "--sun") sun = floats(next());
        else if (a == "--ambient") ambient = floats(next());
        else if (a == "--frames") frames = std::atoi(next());
        else if (a == "--device") device = std::atoi(next());
        else if (a == "--out") out = next();
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); usage(); return 2; }
    }
    if (gltf.empty() == (config == 0)) { usage(); return 2; }

    VulkanEngine engine;
    engine._shadowExtent = {S, S};
    if (int rc = engine.init(W, H, device, nullptr, VKR_CONFIG_TIMINGS); rc != VKR_SUCCESS) {
        std::fprintf(stderr, "vkr_create failed (%d): a CUDA device of compute capability 10.x is required; there is no CPU path\n", rc);
        return 1;
    }
    uint64_t tris = 0;
    if (!gltf.empty()) {
        std::string why;
        std::shared_ptr<Node> top = load_gltf(&engine, "scene", gltf, &why);
        if (!top) { std::fprintf(stderr, "load_gltf: %s\n", why.c_str()); return 1; }
        engine._sceneRoot->addChild(top);
        engine._enableShadows = true; engine._shadowMode = 1; engine._enablePostprocess = true; engine._enableBackground = true;
        engine._mainCamera.position = vkm::vec3{0, 1, 6};
    } else {
        tris = load_procedural(&engine, config, 0xB2000000u + (uint32_t)config);
        if (!tris) { std::fprintf(stderr, "unknown config %d\n", config); return 1; }
    }
    if (!cubemap.empty()) {      // a failed load leaves the scene without a skybox, as in the reference
        std::string why;
        if (!engine.load_cubemap(cubemap.c_str(), &why)) std::fprintf(stderr, "load_cubemap: %s\n", why.c_str());
    }
    if (shadows >= 0) engine._enableShadows = shadows != 0;
    if (shadowMode >= 0) engine._shadowMode = shadowMode;
    if (background >= 0) engine._enableBackground = background != 0;
    if (post >= 0) engine._enablePostprocess = post != 0;
    if (lambert) engine._fragShader = VKR_FRAG_MESH_LAMBERT;
    engine._gpuSceneUpdate = gpuScene;
    if (cam.size() >= 3) engine._mainCamera.position = vkm::vec3{cam[0], cam[1], cam[2]};
    if (cam.size() >= 5) { engine._mainCamera.pitch = cam[3]; engine._mainCamera.yaw = cam[4]; }
    if (sun.size() >= 3) engine._sunlightDirection = vkm::vec4{sun[0], sun[1], sun[2], 0.0f};
    if (ambient.size() >= 3) engine._ambientColor = vkm::vec4{ambient[0], ambient[1], ambient[2], 1.0f};

    const auto t0 = std::chrono::steady_clock::now();
    for (int f = 0; f < frames; ++f) {
        engine.update_scene(0.0f);
        if (int rc = engine.draw(); rc != VKR_SUCCESS) { std::fprintf(stderr, "draw failed (%d)\n", rc); return 1; }
    }
    std::vector<unsigned char> rgb((size_t)W * H * 3);
    if (engine._api.readback(engine._vkr, VKR_TARGET_RGB8, rgb.data(), rgb.size()) != VKR_SUCCESS) { std::fprintf(stderr, "readback failed\n"); return 1; }
    const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    VkrEngineStats st{};
    vkr_stats(engine._vkr, &st);
    // (with --gpu-scene the host keeps no draw list: the device counted what its is_visible kept)
    const int trisDrawn = gpuScene ? st.triangle_count : engine.stats.triangle_count, draws = gpuScene ? st.drawcall_count : engine.stats.drawcall_count;
    std::printf("frametime %.3f ms\ndraw time %.3f ms\ntriangles %d\ndraws %d\nvisible objects %zu\nwall %.3f ms for %d frame(s)\n",
                st.frametime, st.mesh_draw_time, trisDrawn, draws, gpuScene ? (size_t)draws : engine._renderObjects.size(), wall, frames);
    if (!out.empty()) {
        FILE* f = std::fopen(out.c_str(), "wb");
        if (!f) { std::fprintf(stderr, "cannot write %s\n", out.c_str()); return 1; }
        std::fprintf(f, "P6\n%u %u\n255\n", W, H);
        std::fwrite(rgb.data(), 1, rgb.size(), f);
        std::fclose(f);
    }
    return 0;
}
