"""The C++ host engine (csrc/host, libvkr_engine.so): scene graph -> update_scene -> draw.

CPU: host math against golden vectors computed with the reference's vendored glm (SURVEY Appendix C and
tests/golden/host_math.json), culling/sorting behaviour, and the procedural BASELINE scenes rendered
through the engine on the oracle vs. the kernels' host emulation.
GPU: the same scenes, engine on libvkr (CUDA) vs. engine on the oracle.
"""
import json
import os

import numpy as np
import pytest

from vk_renderer_b200 import capi
from vk_renderer_b200.engine import Engine
import parity

HERE = os.path.dirname(os.path.abspath(__file__))


def render_engine(lib, prefix, config, width, height, shadow_extent=512, tweak=None):
    e = Engine(width, height, api_lib=lib, api_prefix=prefix, shadow_extent=shadow_extent)
    tris = e.load_procedural(config)
    if tweak:
        tweak(e)
    e.update_scene(0.0)
    e.draw()
    out = {k: e.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID,
                                             shadow=capi.TARGET_SHADOW).items()}
    out["stats"] = e.stats()
    sd = e.scene_data()
    out["shadows"] = sd.sunlightColor[3] > 0.5
    out["layers"] = 4 if int(sd.sunlightDirection[3]) == 3 else 1
    out["tris"] = tris
    out["visible"] = e.visible_objects()
    e.close()
    return out


# ---------------------------------------------------------------- host math
def test_known_answer_vectors(ork_lib):
    """SURVEY Appendix C (computed from the reference's formulas with its vendored glm 0.9.9.7)."""
    e = Engine(1280, 720, api_lib=ork_lib, api_prefix="ork", shadow_extent=64)
    view, proj = e.camera_matrices()
    assert np.array_equal(view, np.eye(4, dtype=np.float32))          # init_camera: pos 0, pitch 0, yaw 0
    want = np.zeros((4, 4), np.float32)
    want[0, 0], want[1, 1], want[2, 2], want[3, 2], want[2, 3] = 0.97427851, -1.73205066, -1.001001, -1.0, -0.1001001
    np.testing.assert_allclose(proj, want, rtol=2e-7, atol=0)
    splits = [e.lib.vke_compute_split(0.1, 100.0, i) for i in (1, 2, 3, 4)]
    np.testing.assert_allclose(splits, [12.8186712, 26.6061382, 46.4039001, 100.0], rtol=2e-7)
    L = e.light_matrix()          # default sun (0.5,-1,-0.5), zNear 0.1, zFar 100; columns from Appendix C
    cols = np.array([[0.00707106758, -0.0057735024, 0.00408656895, 0], [0, -0.0057735024, -0.0081731379, 0],
                     [0.00707106758, 0.0057735024, -0.00408656895, 0], [0, 0, 0.734581411, 1]], np.float32)
    np.testing.assert_allclose(L.T, cols, rtol=3e-6, atol=1e-9)
    e.close()


def test_golden_host_math(ork_lib):
    """tests/golden/host_math.json: outputs of the reference's own host functions (src/vk_camera.cpp,
    src/vk_engine_run.cpp:420-566) compiled from their source text against the vendored glm
    (oracle/refshader/refhost.cpp; generator tools/gen_golden.py)."""
    path = os.path.join(HERE, "golden", "host_math.json")
    G = json.load(open(path))
    e = Engine(64, 64, api_lib=ork_lib, api_prefix="ork", shadow_extent=64)
    np.testing.assert_allclose([e.lib.vke_compute_split(0.1, 100.0, i) for i in (1, 2, 3, 4)], G["compute_split_0.1_100"], rtol=2e-7)
    e.close()
    n_vis = 0
    for case in G["cases"]:
        e = Engine(case["width"], case["height"], api_lib=ork_lib, api_prefix="ork", shadow_extent=64)
        e.set_camera(case["position"], case["pitch"], case["yaw"])
        e.lib.vke_set_light(e.e, (capi.C.c_float * 3)(*case["sun"]), None, None)
        view, proj = e.camera_matrices()
        np.testing.assert_allclose(view.T.reshape(-1), case["view"], rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(proj.T.reshape(-1), case["proj"], rtol=1e-6, atol=0)
        np.testing.assert_allclose(e.light_matrix().T.reshape(-1), case["light"], rtol=1e-5, atol=2e-6)
        m, d = e.csm()
        np.testing.assert_allclose(d, case["csm_distances"], rtol=1e-6)
        np.testing.assert_allclose(m.transpose(0, 2, 1).reshape(-1), case["csm_matrices"], rtol=2e-4, atol=2e-4)
        np.testing.assert_allclose(e.frustum_planes().reshape(-1), case["planes"], rtol=1e-5, atol=2e-5)
        planes = np.array(case["planes"], np.float64).reshape(6, 4)
        for v in case["is_visible"]:
            # skip spheres within rounding distance of a plane (the engine's own math differs from glm by ulps)
            T = np.array(v["transform"], np.float64).reshape(4, 4).T
            c = (T @ np.append(v["origin"], 1.0))[:3]
            rad = v["radius"] * np.sqrt((T[:3, :3] ** 2).sum(axis=0)).max()
            margin = np.abs(planes[:, :3] @ c + planes[:, 3] + rad).min()
            if margin < 1e-3:
                continue
            assert e.is_visible(v["transform"], v["origin"], v["radius"]) == bool(v["visible"])
            n_vis += 1
        e.close()
    assert n_vis > 200


def test_update_scene_culls_and_sorts(ork_lib):
    e = Engine(320, 180, api_lib=ork_lib, api_prefix="ork", shadow_extent=64)
    e.load_procedural(1)
    e.update_scene(0.0)
    assert e.visible_objects() == 64                       # whole 8x8 grid in view from the config camera
    e.set_camera((0, 6, 22), pitch=-0.25, yaw=3.14159)     # look away: everything is behind the camera
    e.update_scene(0.0)
    assert e.visible_objects() == 0
    e.set_camera((0, 1, 4), pitch=0.0, yaw=0.0)            # inside the grid: only part of it survives
    e.update_scene(0.0)
    assert 0 < e.visible_objects() < 64
    e.draw()
    st = e.stats()
    assert st.drawcall_count == e.visible_objects() and st.triangle_count > 0
    e.close()


# ---------------------------------------------------------------- whole path through the engine
@pytest.mark.parametrize("config", [1, 2])
def test_engine_emu_matches_oracle(ork_lib, emu_lib, config):
    ref = render_engine(ork_lib, "ork", config, 256, 144, shadow_extent=256)
    got = render_engine(emu_lib, "emu", config, 256, 144, shadow_extent=256)
    assert ref["tris"] == got["tris"] and ref["visible"] == got["visible"]
    parity.assert_parity(ref, got)


@pytest.mark.gpu
@pytest.mark.parametrize("config,size", [(1, (1280, 720)), (2, (1920, 1080)), (3, (960, 540)), (3, (3840, 2160))])
def test_engine_gpu_matches_oracle(ork_lib, vkr_lib, config, size):
    ref = render_engine(ork_lib, "ork", config, *size, shadow_extent=2048)
    got = render_engine(None, "vkr", config, *size, shadow_extent=2048)
    assert ref["tris"] == got["tris"] and ref["visible"] == got["visible"]
    parity.assert_parity(ref, got)


@pytest.mark.gpu
def test_engine_gpu_csm_and_pcss(ork_lib, vkr_lib):
    for mode in (2, 3):
        tweak = lambda e: e.set_toggles(shadow_mode=mode)
        ref = render_engine(ork_lib, "ork", 2, 640, 360, shadow_extent=1024, tweak=tweak)
        got = render_engine(None, "vkr", 2, 640, 360, shadow_extent=1024, tweak=tweak)
        parity.assert_parity(ref, got)
