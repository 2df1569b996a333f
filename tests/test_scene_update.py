"""update_scene on the device: vkr_scene_create / vkr_draw_scene against the host's update_scene.

The reference walks the scene graph on the CPU every frame (Node::refreshTransform, src/vk_types.h:148-163), tests each
object with is_visible (src/vk_engine_run.cpp:461-480) and sorts the survivors (:454-458).  With `_gpuSceneUpdate` the
host engine hands the flattened graph over once and the per-object work runs in kernels.  The bar: the draw list the
device keeps equals the host's `_renderObjects` byte for byte (world matrices included), and the frame — colour, depth,
primitive ids, shadow map — equals the frame of vkr_draw with the host's list, bit for bit; against the oracle (which only
ever sees the host's list) it meets the parity bar.

CPU: the engine on the kernels' host emulation (tests/emu runs the same per-thread bodies, vkr_scene.cuh).
GPU: the engine on libvkr, and the C-ABI directly.
"""
import ctypes as C

import numpy as np
import pytest

from vk_renderer_b200 import capi, meshgen as mg
from vk_renderer_b200.engine import Engine
import gltf_util as gu
import parity
import scenes

TARGETS = dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID, shadow=capi.TARGET_SHADOW)

# (position, pitch, yaw) for BASELINE config 1's 8 x 8 grid: the config camera is used when None; inside the grid (part of it
# survives is_visible); grazing; looking away (nothing survives)
CAMERAS = [None, ((0, 1, 4), 0.0, 0.0), ((3.0, 2.0, 9.0), -0.2, 0.5), ((-7.0, 0.5, 2.0), 0.1, -1.2), ((0, 6, 22), -0.25, 3.14159)]


def _frame(e):
    out = {k: e.readback(v) for k, v in TARGETS.items()}
    out["stats"] = e.stats()
    sd = e.scene_data()
    out["shadows"] = sd.sunlightColor[3] > 0.5
    out["layers"] = 4 if int(sd.sunlightDirection[3]) == 3 else 1
    return out


def _assert_same_frame(a, b, what):
    for k in TARGETS:
        assert np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)), f"{what}: {k} differs on {(a[k] != b[k]).sum()} elements"
    assert a["stats"].drawcall_count == b["stats"].drawcall_count, what
    assert a["stats"].triangle_count == b["stats"].triangle_count, what


def _host_list(e):
    return np.array([ro for ro, _, _ in e.render_objects()], capi.RENDER_OBJECT_DTYPE) if e.visible_objects() else np.zeros(0, capi.RENDER_OBJECT_DTYPE)


def _both_engines(lib, prefix, width, height, shadow_extent, load, flags=0):
    """Two engines on the same backend with the same scene: update_scene on the host, and on the device."""
    kw = dict(api_lib=lib, api_prefix=prefix) if lib is not None else {}
    host = Engine(width, height, shadow_extent=shadow_extent, flags=flags, **kw)
    dev = Engine(width, height, shadow_extent=shadow_extent, flags=flags, **kw)
    for e in (host, dev):
        load(e)
    dev.set_gpu_scene_update(True)
    return host, dev


def _check_cameras(host, dev, cameras, what):
    seen = set()
    for cam in cameras:
        for e in (host, dev):
            if cam is not None:
                e.set_camera(cam[0], pitch=cam[1], yaw=cam[2])
            e.update_scene(0.0)
            e.draw()
        a, b = _frame(host), _frame(dev)
        assert dev.visible_objects() == 0, "with the device's update_scene the host keeps no draw list"
        hl, dl = _host_list(host), dev.gpu_draw_list()
        assert hl.size == dl.size, f"{what} {cam}: the device kept {dl.size} draws, update_scene {hl.size}"
        assert hl.tobytes() == dl.tobytes(), f"{what} {cam}: draw lists differ (order, handles or world matrices)"
        assert b["stats"].drawcall_count == hl.size
        _assert_same_frame(a, b, f"{what} {cam}")
        seen.add(int(hl.size))
    return seen


def _config(config, toggles=None):
    def load(e):
        e.load_procedural(config)
        if toggles:
            e.set_toggles(**toggles)
    return load


# ---------------------------------------------------------------- CPU: the kernels' bodies, emulated
@pytest.mark.parametrize("config", [1, 2])
def test_emu_device_update_scene_equals_host_update_scene(emu_lib, config):
    host, dev = _both_engines(emu_lib, "emu", 192, 108, 128, _config(config))
    seen = _check_cameras(host, dev, CAMERAS if config == 1 else CAMERAS[:3], f"config {config}")
    if config == 1:
        assert 64 in seen and 0 in seen and any(0 < s < 64 for s in seen), "the cameras were meant to cull none, some and all of the grid"
    host.close(); dev.close()


def test_emu_many_objects_under_group_nodes(emu_lib):
    """Config 6: 16,000 draws under 40 group nodes (two levels of refreshTransform), seen from inside the lattice."""
    host, dev = _both_engines(emu_lib, "emu", 96, 54, 64, _config(6))
    seen = _check_cameras(host, dev, [None, ((0, 1, 12), -0.05, 2.0)], "config 6")
    assert all(1000 < s < 16000 for s in seen) and len(seen) == 2
    host.close(); dev.close()


def test_emu_device_update_scene_against_the_oracle(ork_lib, emu_lib):
    """The oracle never sees the scene graph: it is given the host's list.  Same frame, parity bar."""
    ref = Engine(192, 108, api_lib=ork_lib, api_prefix="ork", shadow_extent=128)
    dev = Engine(192, 108, api_lib=emu_lib, api_prefix="emu", shadow_extent=128)
    for e in (ref, dev):
        e.load_procedural(2)
        e.set_toggles(shadow_mode=3)              # four shadow layers: five orders on the device
    dev.set_gpu_scene_update(True)
    for e in (ref, dev):
        e.set_camera((2.0, 2.5, 7.0), pitch=-0.2, yaw=0.3)
        e.update_scene(0.0)
        e.draw()
    parity.assert_parity(_frame(ref), _frame(dev))
    ref.close(); dev.close()


def test_emu_nested_nodes_transparent_and_cutout_draws(emu_lib, tmp_path):
    """A glTF scene: nested TRS / matrix nodes (world = parent * local down three levels), a BLEND material (the Transparent
    pass: opaque-first order), textures with cut-outs — and a node moved between frames."""
    asset = gu.build_asset(np.random.default_rng(11))
    path = gu.write_gltf(asset, str(tmp_path), "glb")

    def load(e):
        e.load_gltf(path, "test")
        e.set_toggles(shadows=1, shadow_mode=1, background=1, postprocess=1)
    host, dev = _both_engines(emu_lib, "emu", 200, 120, 128, load)
    cams = [((-1.0, 1.5, 9.0), -0.1, 0.0), ((4.0, 1.0, 3.0), 0.0, 0.9), ((-1.0, 1.5, -14.0), 0.0, 0.0)]
    seen = _check_cameras(host, dev, cams, "glTF")
    assert max(seen) > 3 and min(seen) < max(seen)
    # Node::localTransform changes: a parent with children (its subtree moves) and a leaf
    assert host.node_count() == dev.node_count() > 4
    rng = np.random.default_rng(5)
    for idx in (2, host.node_count() - 1):
        m = np.eye(4, dtype=np.float32)
        m[:3, :3] = np.linalg.qr(rng.normal(size=(3, 3)))[0] * rng.uniform(0.5, 1.5)
        m[:3, 3] = rng.uniform(-1.5, 1.5, 3)
        for e in (host, dev):
            e.set_node_transform(idx, m.T.reshape(-1))
        _check_cameras(host, dev, cams[:2], f"glTF after moving node {idx}")
    host.close(); dev.close()


def _random_tree_asset(seed, count=60):
    """The glTF test asset with its six nodes replaced by a seeded random forest: up to seven levels deep, TRS nodes with
    rotations and non-uniform scales, matrix nodes, empty group nodes, meshes with one or two surfaces, some of them BLEND."""
    rng = np.random.default_rng(seed)
    asset = gu.build_asset(np.random.default_rng(11))
    nodes = []
    for i in range(count):
        q = rng.normal(size=4)
        q = (q / np.linalg.norm(q)).astype(np.float32)
        t = rng.uniform(-2.5, 2.5, 3).astype(np.float32)
        sc = rng.uniform(0.6, 1.4, 3).astype(np.float32)
        kind = rng.integers(0, 4)
        if kind == 0:
            m = (np.diag([*sc, 1.0]) + 0.0).astype(np.float32)
            m[:3, 3] = t
            node = dict(matrix=[float(x) for x in m.T.reshape(-1)])
        elif kind == 1:
            node = dict(trs=dict(t=[float(x) for x in t]))
        else:
            node = dict(trs=dict(t=[float(x) for x in t], r=[float(x) for x in q], s=[float(x) for x in sc]))
        node["mesh"] = [0, 1, 2, None][rng.integers(0, 4)]
        node["children"] = []
        nodes.append(node)
    depth = [0] * count
    for i in range(1, count):                       # a parent among the earlier nodes (or none), at most six levels above
        if rng.random() < 0.15:
            continue
        cands = [j for j in range(i) if depth[j] < 6]
        j = int(cands[rng.integers(0, len(cands))])
        nodes[j]["children"].append(i)
        depth[i] = depth[j] + 1
    asset["nodes"] = nodes
    return asset, max(depth)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_emu_random_scene_graphs(emu_lib, tmp_path, seed):
    asset, deepest = _random_tree_asset(seed)
    assert deepest >= 4
    path = gu.write_gltf(asset, str(tmp_path), "glb")

    def load(e):
        e.load_gltf(path, "test")
        e.set_toggles(shadows=1, shadow_mode=1 if seed != 2 else 3, background=1, postprocess=1)
    host, dev = _both_engines(emu_lib, "emu", 160, 96, 128, load)
    rng = np.random.default_rng(100 + seed)
    cams = [(tuple(rng.uniform(-6, 6, 3) + np.array([0, 0, 6.0])), float(rng.uniform(-0.4, 0.4)), float(rng.uniform(-1.0, 1.0))) for _ in range(3)]
    seen = _check_cameras(host, dev, cams, f"random tree {seed}")
    assert max(seen) > 10
    host.close(); dev.close()


def test_bucket_order_helpers(emu_lib):
    """The order-preserving float -> uint32 map (atomicMax finds the key range with it) and the slice number of a key:
    monotone over negative, zero, denormal and large keys; hidden draws (+inf) in the last slice; every slice in range."""
    emu_lib.emu_scene_ordered.argtypes = [C.c_float]; emu_lib.emu_scene_ordered.restype = C.c_uint32
    emu_lib.emu_scene_unordered.argtypes = [C.c_uint32]; emu_lib.emu_scene_unordered.restype = C.c_float
    emu_lib.emu_scene_buckets.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    rng = np.random.default_rng(9)
    keys = np.concatenate([rng.normal(0, 50, 4000), rng.uniform(-1e-38, 1e-38, 50), [0.0, -0.0, 1e30, -1e30, 3.5, 3.5],
                           np.float32(10.0) + np.arange(64) * np.float32(1e-6)]).astype(np.float32)
    order = np.argsort(keys, kind="stable")
    mapped = np.array([emu_lib.emu_scene_ordered(float(k)) for k in keys], np.uint64)
    ks, ms = keys[order], mapped[order].astype(np.int64)
    strictly = ks[1:] > ks[:-1]                                        # (-0.0 == 0.0 as keys; their images differ by one)
    assert np.all(np.diff(ms)[strictly] > 0) and np.all(np.abs(np.diff(ms)[~strictly]) <= 1), "ordered() must not reorder keys"
    assert all(np.float32(emu_lib.emu_scene_unordered(int(m))).tobytes() == np.float32(k).tobytes() or k == 0.0 for m, k in zip(mapped[:200], keys[:200]))
    for sample in (keys, keys[:4000], np.float32([5.0, 5.0, 5.0]), np.float32([-3.0, 7.0])):
        withinf = np.concatenate([sample, np.float32([np.inf, np.inf])])
        out = np.zeros(withinf.size, np.uint32)
        emu_lib.emu_scene_buckets(withinf.ctypes.data, withinf.size, out.ctypes.data)
        assert out[-1] == out[-2] == 2047 and out[:-2].max() <= 2046
        o = np.argsort(sample, kind="stable")
        assert np.all(np.diff(out[:-2][o].astype(np.int64)) >= 0), "a nearer key never lands in a later slice"
        if sample.min() < sample.max():
            assert out[:-2].min() == 0 and out[:-2].max() == 2046
    out = np.zeros(2, np.uint32)
    emu_lib.emu_scene_buckets(np.float32([np.inf, np.inf]).ctypes.data, 2, out.ctypes.data)
    assert list(out) == [2047, 2047]                                   # no visible draw at all


def _flat_scene(r, objs, bounds_radius=1.0e6):
    """A draw list as a one-level scene graph: node i carries object i's matrix; bounds so large that is_visible keeps all."""
    nodes = np.zeros(len(objs), capi.SCENE_NODE_DTYPE)
    surf = np.zeros(len(objs), capi.SCENE_SURFACE_DTYPE)
    for i, o in enumerate(objs):
        nodes[i]["localTransform"] = o["transform"]
        nodes[i]["parent"] = -1
        surf[i] = (i, o["mesh"], o["material"], o["indexCount"], o["firstIndex"], (0, 0, 0), bounds_radius)
    return r.scene_create(nodes, surf)


def _planes_everything():
    pl = np.zeros((6, 4), np.float32)
    pl[:, 0] = 1.0
    pl[:, 3] = 1.0e9
    return pl


@pytest.mark.parametrize("scene", ["basic_pcf", "transparent", "cutout", "basic_csm"])
def test_emu_draw_scene_equals_draw_through_the_c_abi(emu_lib, scene):
    """vkr_draw_scene with nothing culled against vkr_draw of the same list: every pass (opaque, alpha-tested, Transparent,
    one and four shadow layers) consumes what the scene kernels wrote."""
    if scene not in scenes.SCENES:
        pytest.skip(f"no scene {scene}")
    a = capi.Renderer(emu_lib, "emu", 160, 96, shadow_extent=128)
    b = capi.Renderer(emu_lib, "emu", 160, 96, shadow_extent=128)
    sd, objs, fp = scenes.SCENES[scene](a)
    sd2, objs2, fp2 = scenes.SCENES[scene](b)
    a.draw(sd, objs, fp)
    sc = _flat_scene(b, objs2)
    b.draw_scene(sc, sd2, _planes_everything(), fp2)
    for k, t in TARGETS.items():
        x, y = a.readback(t), b.readback(t)
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{scene}: {k} differs on {(x != y).sum()} elements"
    dl = b.scene_draw_list(sc)
    assert dl.tobytes() == np.ascontiguousarray(objs2, capi.RENDER_OBJECT_DTYPE).tobytes()
    assert a.stats().triangle_count == b.stats().triangle_count and b.stats().drawcall_count == len(objs2)
    a.close(); b.close()


def test_is_visible_on_the_device_follows_the_reference_bit_for_bit(emu_lib):
    """Spheres placed ON the frustum planes (distance = -radius up to a few ulps): the device's verdicts equal the host
    engine's is_visible (golden-pinned against the reference's own text, tests/test_engine.py) for every one of them."""
    e = Engine(320, 180, api_lib=emu_lib, api_prefix="emu", shadow_extent=64)
    e.load_procedural(1)
    e.set_camera((0.3, 1.2, 6.0), pitch=-0.1, yaw=0.2)
    e.update_scene(0.0)
    planes = e.frustum_planes()
    rng = np.random.default_rng(21)
    r = capi.Renderer(emu_lib, "emu", 64, 36, shadow_extent=64)
    sd, objs, fp = scenes.SCENES["basic_pcf"](r)
    base = objs[0]
    n = 400
    nodes = np.zeros(n, capi.SCENE_NODE_DTYPE)
    surf = np.zeros(n, capi.SCENE_SURFACE_DTYPE)
    want = []
    for i in range(n):
        pl = planes[i % 6].astype(np.float64)
        radius = float(rng.uniform(0.2, 2.0))
        scale = rng.uniform(0.5, 2.0, 3)
        origin = rng.uniform(-0.5, 0.5, 3).astype(np.float32)
        # a point in front of the camera, moved along the plane normal until the sphere's centre is at -radius * maxScale (+- a few ulps)
        p = rng.uniform(-3, 3, 3) + np.array([0.0, 1.0, 0.0])
        dist = pl[:3] @ p + pl[3]
        centre = p - pl[:3] * (dist + radius * scale.max() * (1.0 + rng.integers(-3, 4) * 6e-8))
        m = np.diag([*scale, 1.0]).astype(np.float32)
        m[:3, 3] = (centre - (m[:3, :3] @ origin)).astype(np.float32)
        col = m.T.reshape(-1)
        nodes[i]["localTransform"] = col
        nodes[i]["parent"] = -1
        surf[i] = (i, base["mesh"], base["material"], base["indexCount"], base["firstIndex"], tuple(origin), radius)
        want.append(e.is_visible(col, origin, radius))
    sc = r.scene_create(nodes, surf)
    r.draw_scene(sc, sd, planes, fp)
    got = r.scene_draw_list(sc)
    kept = {tuple(np.asarray(g["transform"]).tolist()) for g in got}
    have = [tuple(nodes[i]["localTransform"].tolist()) in kept for i in range(n)]
    assert have == want
    assert 0.2 < np.mean(want) < 0.8, "the spheres were meant to straddle the planes"
    e.close(); r.close()


def test_scene_create_refuses_what_it_cannot_draw(emu_lib):
    r = capi.Renderer(emu_lib, "emu", 64, 36, shadow_extent=64)
    sd, objs, fp = scenes.SCENES["basic_pcf"](r)
    nodes = np.zeros(2, capi.SCENE_NODE_DTYPE)
    nodes["parent"] = (-1, 0)
    surf = np.zeros(1, capi.SCENE_SURFACE_DTYPE)
    surf[0] = (1, objs[0]["mesh"], objs[0]["material"], objs[0]["indexCount"], objs[0]["firstIndex"], (0, 0, 0), 1.0)
    r.scene_create(nodes, surf)
    for bad in ("parent", "node", "mesh", "material", "range"):
        n2, s2 = nodes.copy(), surf.copy()
        if bad == "parent":
            n2["parent"] = (-1, 1)                 # a node must come after its parent
        elif bad == "node":
            s2[0]["node"] = 2
        elif bad == "mesh":
            s2[0]["mesh"] = 999
        elif bad == "material":
            s2[0]["material"] = 999
        else:
            s2[0]["indexCount"] = 1 << 30
        with pytest.raises(capi.VkrError):
            r.scene_create(n2, s2)
    # a graph without a single surface draws the background and counts nothing
    empty = r.scene_create(nodes, surf[:0])
    r.draw_scene(empty, sd, _planes_everything(), fp)
    assert r.scene_draw_list(empty).size == 0 and r.stats().drawcall_count == 0 and r.stats().triangle_count == 0
    assert (r.readback(capi.TARGET_PRIMID) == 0xFFFFFFFF).all()
    # a destroyed scene is dead, the others live on; the engine replaces its scene when the graph changes
    a, b = r.scene_create(nodes, surf), r.scene_create(nodes, surf)
    r.scene_destroy(a)
    with pytest.raises(capi.VkrError):
        r.draw_scene(a, sd, _planes_everything(), fp)
    with pytest.raises(capi.VkrError):
        r.scene_destroy(a)
    r.draw_scene(b, sd, _planes_everything(), fp)
    assert r.scene_draw_list(b).size == 1
    r.close()


# ---------------------------------------------------------------- GPU: the kernels themselves
@pytest.mark.gpu
@pytest.mark.parametrize("config,size,flags", [(1, (640, 360), 0), (2, (960, 540), 0),
                                                (1, (640, 360), capi.CONFIG_SUBMIT_THREAD | capi.CONFIG_TRANSIENT_HDR)])
def test_gpu_device_update_scene_equals_host_update_scene(vkr_lib, config, size, flags):
    host, dev = _both_engines(None, "vkr", *size, 1024, _config(config), flags=flags)
    seen = _check_cameras(host, dev, CAMERAS if config == 1 else CAMERAS[:3], f"config {config}")
    if config == 1:
        assert 64 in seen and 0 in seen and any(0 < s < 64 for s in seen)
    host.close(); dev.close()


@pytest.mark.gpu
def test_gpu_many_objects_under_group_nodes(vkr_lib):
    """Config 6: 16,000 draws under 40 group nodes — the bucket order (more than 2048 draws), 16 blocks per pass."""
    host, dev = _both_engines(None, "vkr", 480, 270, 512, _config(6))
    seen = _check_cameras(host, dev, [None, ((0, 1, 12), -0.05, 2.0), ((30.0, 3.0, 0.0), -0.1, 1.5708)], "config 6")
    assert all(100 < s < 16000 for s in seen) and len(seen) == 3
    host.close(); dev.close()


@pytest.mark.gpu
def test_gpu_device_update_scene_against_the_oracle(ork_lib, vkr_lib):
    ref = Engine(640, 360, api_lib=ork_lib, api_prefix="ork", shadow_extent=512)
    dev = Engine(640, 360, shadow_extent=512)
    for e in (ref, dev):
        e.load_procedural(2)
        e.set_toggles(shadow_mode=3)
    dev.set_gpu_scene_update(True)
    for e in (ref, dev):
        e.set_camera((2.0, 2.5, 7.0), pitch=-0.2, yaw=0.3)
        e.update_scene(0.0)
        e.draw()
    parity.assert_parity(_frame(ref), _frame(dev))
    ref.close(); dev.close()


@pytest.mark.gpu
def test_gpu_nested_nodes_transparent_and_cutout_draws(vkr_lib, tmp_path):
    asset = gu.build_asset(np.random.default_rng(11))
    path = gu.write_gltf(asset, str(tmp_path), "glb")

    def load(e):
        e.load_gltf(path, "test")
        e.set_toggles(shadows=1, shadow_mode=1, background=1, postprocess=1)
    host, dev = _both_engines(None, "vkr", 400, 240, 512, load)
    cams = [((-1.0, 1.5, 9.0), -0.1, 0.0), ((4.0, 1.0, 3.0), 0.0, 0.9), ((-1.0, 1.5, -14.0), 0.0, 0.0)]
    _check_cameras(host, dev, cams, "glTF")
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = (0.5, -0.25, 1.0)
    for e in (host, dev):
        e.set_node_transform(2, m.T.reshape(-1))
    _check_cameras(host, dev, cams[:2], "glTF after moving node 2")
    host.close(); dev.close()


@pytest.mark.gpu
def test_gpu_random_scene_graph(vkr_lib, tmp_path):
    asset, deepest = _random_tree_asset(2)
    path = gu.write_gltf(asset, str(tmp_path), "glb")

    def load(e):
        e.load_gltf(path, "test")
        e.set_toggles(shadows=1, shadow_mode=3, background=1, postprocess=1)
    host, dev = _both_engines(None, "vkr", 480, 270, 512, load)
    rng = np.random.default_rng(102)
    cams = [(tuple(rng.uniform(-6, 6, 3) + np.array([0, 0, 6.0])), float(rng.uniform(-0.4, 0.4)), float(rng.uniform(-1.0, 1.0))) for _ in range(3)]
    _check_cameras(host, dev, cams, "random tree 2")
    host.close(); dev.close()


@pytest.mark.gpu
def test_gpu_scene_handles_live_and_die_independently(vkr_lib):
    """vkr_scene_destroy, a scene without surfaces, and a scene created after another was destroyed, on the device."""
    r = capi.Renderer(vkr_lib, "vkr", 160, 96, shadow_extent=128)
    sd, objs, fp = scenes.SCENES["basic_pcf"](r)
    r.draw(sd, objs, fp)
    want = {k: r.readback(t) for k, t in TARGETS.items()}
    a = _flat_scene(r, objs)
    nodes = np.zeros(1, capi.SCENE_NODE_DTYPE)
    nodes["localTransform"][0] = np.eye(4, dtype=np.float32).reshape(-1)
    nodes["parent"] = -1
    empty = r.scene_create(nodes, np.zeros(0, capi.SCENE_SURFACE_DTYPE))
    r.draw_scene(empty, sd, _planes_everything(), fp)
    assert r.stats().drawcall_count == 0 and (r.readback(capi.TARGET_PRIMID) == 0xFFFFFFFF).all()
    r.draw_scene(a, sd, _planes_everything(), fp)
    r.scene_destroy(empty)
    b = _flat_scene(r, objs)
    r.scene_destroy(a)
    with pytest.raises(capi.VkrError):
        r.draw_scene(a, sd, _planes_everything(), fp)
    r.draw_scene(b, sd, _planes_everything(), fp)
    for k, t in TARGETS.items():
        got = r.readback(t)
        assert np.array_equal(got.view(np.uint8), want[k].view(np.uint8)), k
    assert r.stats().drawcall_count == len(objs)
    r.close()


def _rebinding_case(lib, prefix):
    """The cut-out scene with slots 3 and 4 first bound to an opaque checker: the scene is created while no draw is alpha-tested,
    then the cut-out textures arrive (update_bindless_texture, as a streaming loader would) — the draws' flags, the cut-out
    kernel's launch and the mip-derivative bit must follow the slots."""
    a = capi.Renderer(lib, prefix, 200, 120, shadow_extent=128)
    b = capi.Renderer(lib, prefix, 200, 120, shadow_extent=128)
    frames = []
    for r, use_scene in ((a, False), (b, True)):
        sd, objs, fp = scenes.SCENES["cutout"](r)
        opaque = r.create_image(mg.checker_texture(64, 8, seed=3), capi.FORMAT_SRGB, True)
        cut3 = r.create_image(mg.cutout_texture(128, 4, seed=2), capi.FORMAT_SRGB, True)
        cut4 = r.create_image(mg.cutout_texture(32, 2, seed=5), capi.FORMAT_UNORM, False)
        r.update_bindless_texture(3, opaque, capi.SAMPLER_LINEAR)
        r.update_bindless_texture(4, opaque, capi.SAMPLER_LINEAR)
        sc = _flat_scene(r, objs) if use_scene else None
        out = []
        for phase in range(2):
            if phase == 1:
                r.update_bindless_texture(3, cut3, capi.SAMPLER_LINEAR)
                r.update_bindless_texture(4, cut4, capi.SAMPLER_NEAREST)
            if use_scene:
                r.draw_scene(sc, sd, _planes_everything(), fp)
            else:
                r.draw(sd, objs, fp)
            out.append({k: r.readback(t) for k, t in TARGETS.items()})
        frames.append(out)
        r.close()
    for phase in range(2):
        for k in TARGETS:
            x, y = frames[0][phase][k], frames[1][phase][k]
            assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"phase {phase}: {k} differs on {(x != y).sum()} elements"
    assert not np.array_equal(frames[0][0]["prim"], frames[0][1]["prim"]), "the cut-outs were meant to open holes"


def test_emu_scene_follows_rebound_textures(emu_lib):
    _rebinding_case(emu_lib, "emu")


@pytest.mark.gpu
def test_gpu_scene_follows_rebound_textures(vkr_lib):
    _rebinding_case(vkr_lib, "vkr")


@pytest.mark.gpu
def test_gpu_many_draws_sort_and_scans(vkr_lib):
    """3000 draws (more than one CTA's worth for the rank kernel, three elements per thread in the scans) in a random order
    of depths: vkr_draw_scene against vkr_draw of the same list, and the device's list against the input."""
    from vk_renderer_b200 import meshgen as mg
    a = capi.Renderer(vkr_lib, "vkr", 480, 270, shadow_extent=512)
    b = capi.Renderer(vkr_lib, "vkr", 480, 270, shadow_extent=512)
    rng = np.random.default_rng(3)
    lists = []
    for r in (a, b):
        sd, objs, fp = scenes.SCENES["basic_pcf"](r)
        base = objs[0]
        many = np.zeros(3000, capi.RENDER_OBJECT_DTYPE)
        g = np.random.default_rng(3)
        for i in range(many.size):
            m = np.eye(4, dtype=np.float32) * np.float32(0.15)
            m[3, 3] = 1.0
            m[:3, 3] = g.uniform((-6, -3, -14), (6, 3, 2))
            many[i] = (base["mesh"], base["material"], m.T.reshape(-1), base["indexCount"], base["firstIndex"])
        lists.append((sd, many, fp))
    a.draw(*lists[0])
    sc = _flat_scene(b, lists[1][1])
    b.draw_scene(sc, lists[1][0], _planes_everything(), lists[1][2])
    for k, t in TARGETS.items():
        x, y = a.readback(t), b.readback(t)
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{k} differs on {(x != y).sum()} elements"
    assert b.scene_draw_list(sc).tobytes() == lists[1][1].tobytes()
    assert b.stats().drawcall_count == 3000 and a.stats().triangle_count == b.stats().triangle_count
    a.close(); b.close()
