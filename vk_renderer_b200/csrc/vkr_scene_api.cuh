// vkr_scene_api.cuh — the C-ABI of update_scene on the device (include/vkr.h: vkr_scene_create, vkr_scene_set_transforms,
// vkr_draw_scene, vkr_scene_draw_list).  Included at the end of vkr_api.cu; kernels in vkr_scene_kernels.cuh.
#pragma once

// The draws' flags follow the bindless slots (alpha-tested / mip-mapped textures): re-derived when a slot changed since.
static int scene_refresh_flags(VkrDevice* dev, SceneRec& s) {
    if (s.flagsGen == dev->bindGen) return VKR_SUCCESS;
    s.anyCutout = false;
    for (size_t i = 0; i < s.tmplHost.size(); ++i) {
        s.tmplHost[i].flags = draw_flags(dev->materials[s.surfaces[i].material], dev->texHost);
        if (s.tmplHost[i].flags & 2u) s.anyCutout = true;
    }
    CK(cudaStreamSynchronize(dev->upStream));          // frames in flight read the templates on the upload stream
    if (!s.tmplHost.empty()) CK(cudaMemcpy(s.dTmpl, s.tmplHost.data(), s.tmplHost.size() * sizeof(DObject), cudaMemcpyHostToDevice));
    s.flagsGen = dev->bindGen;
    return VKR_SUCCESS;
}

extern "C" int vkr_scene_create(VkrDevice* dev, const VkrSceneNode* nodes, size_t nodeCount, const VkrSceneSurface* surfaces, size_t surfaceCount,
                                VkrScene* out) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!out || !nodes || nodeCount == 0 || (surfaceCount && !surfaces)) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: null argument or no nodes");
    if (nodeCount >= 0x7FFFFFFFull || surfaceCount >= 0x7FFFFFFFull) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: too many nodes or surfaces");
    CK(cudaSetDevice(dev->device));
    const uint32_t nn = (uint32_t)nodeCount, ns = (uint32_t)surfaceCount;
    // ---- nodes: parents first (the order refreshTransform visits them), grouped by depth for the level-by-level kernel
    std::vector<float> local((size_t)nn * 16);
    std::vector<int32_t> parent(nn);
    std::vector<uint32_t> depth(nn), levelOrder(nn);
    uint32_t levels = 0;
    for (uint32_t i = 0; i < nn; ++i) {
        const int32_t p = nodes[i].parent;
        if (p < -1 || p >= (int32_t)i) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: a node's parent must precede it (or be -1)");
        std::memcpy(&local[(size_t)i * 16], nodes[i].localTransform, 64);
        parent[i] = p;
        depth[i] = p < 0 ? 0u : depth[p] + 1u;
        levels = std::max(levels, depth[i] + 1u);
    }
    std::unique_ptr<SceneRec> rec(new SceneRec());
    SceneRec& s = *rec;
    s.nNodes = nn; s.nDraws = ns;
    s.levelFirst.assign((size_t)levels + 1, 0u);
    for (uint32_t i = 0; i < nn; ++i) s.levelFirst[depth[i] + 1]++;
    for (uint32_t l = 0; l < levels; ++l) s.levelFirst[l + 1] += s.levelFirst[l];
    {
        std::vector<uint32_t> cursor(s.levelFirst.begin(), s.levelFirst.end() - 1);
        for (uint32_t i = 0; i < nn; ++i) levelOrder[cursor[depth[i]]++] = i;
    }
    // ---- draws: the order of emission, MainColor first (stable): what update_scene's sort leaves (src/vk_engine_run.cpp:454-458)
    for (uint32_t i = 0; i < ns; ++i) {
        const VkrSceneSurface& sf = surfaces[i];
        if (sf.node >= nn) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: surface references an unknown node");
        if (sf.mesh >= dev->meshes.size()) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: surface references an unknown mesh");
        if (sf.material >= dev->materials.size()) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: surface references an unknown material");
        if ((uint64_t)sf.firstIndex + sf.indexCount > dev->meshes[sf.mesh].ni) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: surface index range exceeds the mesh");
    }
    s.surfaces.reserve(ns);
    for (int pass = 0; pass < 2; ++pass)
        for (uint32_t i = 0; i < ns; ++i)
            if ((dev->materials[surfaces[i].material].pass == VKR_PASS_TRANSPARENT) == (pass == 1)) s.surfaces.push_back(surfaces[i]);
    s.opaqueCount = 0;
    while (s.opaqueCount < ns && dev->materials[s.surfaces[s.opaqueCount].material].pass != VKR_PASS_TRANSPARENT) s.opaqueCount++;
    std::vector<SceneDraw> draws(ns);
    std::vector<uint32_t> clusterBase((size_t)ns + 1), seqSlotBase((size_t)ns + 1);
    s.tmplHost.resize(ns);
    DeviceClusterProvider provider(dev);
    uint64_t ncl = 0, nslot = 0, prim = 0;
    for (uint32_t i = 0; i < ns; ++i) {
        const VkrSceneSurface& sf = s.surfaces[i];
        const DMesh& m = dev->meshes[sf.mesh];
        DObject& o = s.tmplHost[i];
        std::memset(&o, 0, sizeof(o));
        o.pos = m.pos; o.attrA = m.attrA; o.attrB = m.attrB;
        o.idx = m.idx + sf.firstIndex;
        o.triCount = sf.indexCount / 3;
        ClusterRange cr{};
        if (o.triCount && !provider.get(sf.mesh, sf.firstIndex, sf.indexCount, cr)) return dev->fail(VKR_ERROR_OUT_OF_MEMORY, "vkr_scene_create: out of memory building clusters");
        o.clVerts = cr.verts; o.clTris = cr.tris; o.clOrig = cr.orig; o.clInfo = cr.info; o.clBounds = cr.bounds; o.clBlob = cr.blob; o.clCone = cr.cone;
        o.material = sf.material;
        SceneDraw& d = draws[i];
        std::memset(&d, 0, sizeof(d));
        d.node = sf.node;
        d.nClusters = o.triCount ? cr.nClusters : 0u;
        std::memcpy(d.origin, sf.boundsOrigin, 12); d.sphereRadius = sf.boundsSphereRadius;
        std::memcpy(d.center, cr.center, 12); d.radius = cr.radius;
        clusterBase[i] = (uint32_t)ncl; seqSlotBase[i] = (uint32_t)nslot;
        ncl += d.nClusters;
        nslot += (o.triCount + SEQ_SLOT - 1) / SEQ_SLOT;
        prim += o.triCount;
        if (i + 1 == s.opaqueCount) s.opaqueClusters = (uint32_t)ncl;
        if (prim >= 0xFFFFFFF0ull || nslot >= 0xFFFFFFull || ncl >= 0x7FFFFFFFull) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_create: too many primitives in one scene");
    }
    clusterBase[ns] = (uint32_t)ncl; seqSlotBase[ns] = (uint32_t)nslot;
    s.nClusters = (uint32_t)ncl; s.nSeqSlots = (uint32_t)nslot; s.totalTris = (uint32_t)prim;
    // ---- device copies (tracked allocations: released by vkr_destroy)
    const size_t nd = std::max<size_t>(ns, 1);
    ALLOC(s.dLocal, (size_t)nn * 64); ALLOC(s.dParent, (size_t)nn * 4); ALLOC(s.dWorld, (size_t)nn * 64); ALLOC(s.dLevelOrder, (size_t)nn * 4);
    ALLOC(s.dDraws, nd * sizeof(SceneDraw)); ALLOC(s.dTmpl, nd * sizeof(DObject));
    ALLOC(s.dClusterBase, (nd + 1) * 4); ALLOC(s.dSeqSlotBase, (nd + 1) * 4);
    ALLOC(s.dKeys, nd * 4 * (1 + VKR_NUM_CASCADES)); ALLOC(s.dOrder, nd * 4 * (1 + VKR_NUM_CASCADES)); ALLOC(s.dVisTris, nd * 4);
    if (ns > SCENE_RANK_MAX) {
        s.nBlocks = (ns + SCENE_BLOCK - 1) / SCENE_BLOCK;
        ALLOC(s.dKeyRange, 2 * 4 * (1 + VKR_NUM_CASCADES)); ALLOC(s.dBucketOf, nd * 2 * (1 + VKR_NUM_CASCADES));
        ALLOC(s.dHist, (size_t)(1 + VKR_NUM_CASCADES) * SCENE_BUCKETS * s.nBlocks * 4);
    }
    CK(cudaMemcpy(s.dLocal, local.data(), (size_t)nn * 64, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.dParent, parent.data(), (size_t)nn * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.dLevelOrder, levelOrder.data(), (size_t)nn * 4, cudaMemcpyHostToDevice));
    if (ns) CK(cudaMemcpy(s.dDraws, draws.data(), (size_t)ns * sizeof(SceneDraw), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.dClusterBase, clusterBase.data(), ((size_t)ns + 1) * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.dSeqSlotBase, seqSlotBase.data(), ((size_t)ns + 1) * 4, cudaMemcpyHostToDevice));
    s.flagsGen = ~0ull;
    {   // (the flags need material_constants-independent state only: the bindless slots as they are now)
        const int rc = scene_refresh_flags(dev, s);
        if (rc != VKR_SUCCESS) return rc;
    }
    s.worldDirty = true;
    dev->scenes.push_back(rec.release());
    *out = (VkrScene)(dev->scenes.size() - 1);
    return VKR_SUCCESS;
}

extern "C" int vkr_scene_destroy(VkrDevice* dev, VkrScene scene) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (scene >= dev->scenes.size() || !dev->scenes[scene]) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_destroy: unknown scene");
    CK(cudaSetDevice(dev->device));
    CK(cudaDeviceSynchronize());                      // frames in flight read the scene's arrays on the upload stream
    SceneRec* s = dev->scenes[scene];
    if (dev->lastPlan && dev->lastPlan->scene == s) { dev->lastPlan = nullptr; dev->lastN = 0; }      // (vkr_sync must not redraw from it)
    for (FramePlan& p : dev->plans) if (p.scene == s) p.scene = nullptr;
    void* arrays[] = {s->dLocal, s->dParent, s->dWorld, s->dLevelOrder, s->dDraws, s->dTmpl, s->dClusterBase, s->dSeqSlotBase,
                      s->dKeys, s->dOrder, s->dVisTris, s->dKeyRange, s->dBucketOf, s->dHist};
    for (void* a : arrays) dev_free(dev, a);
    delete s;
    dev->scenes[scene] = nullptr;
    return VKR_SUCCESS;
}

extern "C" int vkr_scene_set_transforms(VkrDevice* dev, VkrScene scene, uint32_t firstNode, uint32_t count, const float* localTransforms) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);                                       // every frame handed to the submit thread has been issued: it reads the old matrices
    if (scene >= dev->scenes.size() || !dev->scenes[scene]) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_set_transforms: unknown scene");
    SceneRec& s = *dev->scenes[scene];
    if ((uint64_t)firstNode + count > s.nNodes || (count && !localTransforms)) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_set_transforms: node range");
    if (!count) return VKR_SUCCESS;
    CK(cudaSetDevice(dev->device));
    // behind the scene kernels of the frames in flight (upload stream); the source may be pageable: wait for the copy
    CK(cudaMemcpyAsync(s.dLocal + (size_t)firstNode * 16, localTransforms, (size_t)count * 64, cudaMemcpyHostToDevice, dev->upStream));
    CK(cudaStreamSynchronize(dev->upStream));
    s.worldDirty = true;
    return VKR_SUCCESS;
}

extern "C" int vkr_draw_scene(VkrDevice* dev, VkrScene scene, const VkrGPUSceneData* sceneData, const float frustumPlanes[24], const VkrFrameParams* fp) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    if (!sceneData || !frustumPlanes || !fp) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_draw_scene: null argument");
    if (scene >= dev->scenes.size() || !dev->scenes[scene]) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_draw_scene: unknown scene");
    if (dev->shardRanks > 1) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_draw_scene: not available on a sharded device (the band split needs the draw list on the host: use vkr_draw)");
    CK(cudaSetDevice(dev->device));
    const bool async = (dev->flags & VKR_CONFIG_SUBMIT_THREAD) && !dev->shadowExchange;
    if (!async) DRAIN(dev);
    if (async && !dev->worker.joinable()) dev->worker = std::thread(worker_main, dev);
    const auto hostT = std::chrono::steady_clock::now();
    SceneRec& s = *dev->scenes[scene];
    if (s.flagsGen != dev->bindGen) {
        DRAIN(dev);
        const int rc = scene_refresh_flags(dev, s);
        if (rc != VKR_SUCCESS) return rc;
    }
    FramePlan& plan = dev->plans[dev->planIdx];        // the other one may still be in the submit thread's hands
    const size_t n = s.nDraws;
    plan.shadowLayers = fp->enableShadows ? (fp->shadowMode < 3 ? 1 : VKR_NUM_CASCADES) : 0;
    plan.opaqueCount = s.opaqueCount; plan.opaqueClusters = s.opaqueClusters; plan.anyCutout = s.anyCutout; plan.totalTris = s.totalTris;
    plan.passSlots.assign((size_t)(1 + plan.shadowLayers), s.nClusters);      // every draw keeps its work items; hidden ones are culled clusters
    plan.passSlots[0] = s.opaqueClusters;
    plan.bandMain.assign(1, 0); plan.bandShadow.assign(1, 0);
    plan_constants(sceneData, fp, dev->W, dev->H, dev->S, n, plan);
    plan.scene = &s;
    SceneFrame& f = plan.sceneFrame;
    std::memcpy(f.viewproj, sceneData->viewproj, 64);
    std::memcpy(f.lightViewproj, sceneData->lightViewproj, sizeof(f.lightViewproj));
    std::memcpy(f.planes, frustumPlanes, 96);
    f.nDraws = (uint32_t)n; f.opaqueCount = s.opaqueCount; f.shadowLayers = (uint32_t)plan.shadowLayers; f.pad = 0;
    if (dev->hostProf && dev->frameIndex >= 16)
        dev->hostProfUs[0] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - hostT).count();
    if (!async) return submit_frame(dev, plan, n);
    const auto waitT = std::chrono::steady_clock::now();
    DRAIN(dev);
    if (dev->hostProf && dev->frameIndex >= 16)
        dev->hostProfUs[7] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - waitT).count();
    dev->planIdx ^= 1;
    post_job(dev, VkrDevice::Job{0, &plan, n, VKR_TARGET_RGBA8, nullptr, 0});
    return VKR_SUCCESS;
}

extern "C" int vkr_scene_draw_list(VkrDevice* dev, VkrScene scene, VkrRenderObject* out, size_t capacity, size_t* count) {
    if (!dev) return VKR_ERROR_INVALID_ARGUMENT;
    DRAIN(dev);
    if (!count || (capacity && !out)) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_draw_list: null argument");
    if (scene >= dev->scenes.size() || !dev->scenes[scene]) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_draw_list: unknown scene");
    SceneRec& s = *dev->scenes[scene];
    if (!dev->lastPlan || dev->lastPlan->scene != &s || !dev->lastObjects) return dev->fail(VKR_ERROR_INVALID_ARGUMENT, "vkr_scene_draw_list: the last frame was not drawn from this scene");
    CK(cudaSetDevice(dev->device));
    CK(cudaStreamSynchronize(dev->upStream));
    std::vector<DObject> objs(s.nDraws);
    if (s.nDraws) CK(cudaMemcpy(objs.data(), dev->lastObjects, (size_t)s.nDraws * sizeof(DObject), cudaMemcpyDeviceToHost));
    size_t k = 0;
    for (uint32_t i = 0; i < s.nDraws; ++i) {
        if (objs[i].flags & DRAW_HIDDEN) continue;
        if (k < capacity) {
            VkrRenderObject& r = out[k];
            r.mesh = s.surfaces[i].mesh; r.material = s.surfaces[i].material;
            std::memcpy(r.transform, objs[i].world, 64);
            r.indexCount = s.surfaces[i].indexCount; r.firstIndex = s.surfaces[i].firstIndex;
        }
        ++k;
    }
    *count = k;
    return VKR_SUCCESS;
}
