"""Pins the oracle's shader arithmetic to the reference's own shader text.

tests/golden/shader_golden.json holds outputs of the reference's GLSL (shaders/*.frag|vert|comp,
compiled as C++ against its vendored glm by oracle/refshader; generator: tools/gen_golden.py).
Here oracle/ork.cpp's restatement is run on the same inputs.  Where /root/reference is mounted
the vectors are additionally re-derived live (test_golden_is_current).

Tolerance: the restatement uses the same formulas but not glm's exact operation order
(normalize = v * inversesqrt(dot) in glm, v / length here; pow via std::pow in both), so values
agree to a few ulp: |a-b| <= 2e-5 * max(1, |b|).  discard decisions must agree exactly.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "shader_golden.json")
RTOL = 2e-5


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def fixture_dev(ork_lib, golden):
    gu.bind_ork(ork_lib)
    return gu.FixtureDevice(ork_lib, golden["fixture"])


def close(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.all(np.abs(a - b) <= rtol * np.maximum(1.0, np.abs(b)))


def test_fragment_shaders_match_reference(ork_lib, fixture_dev, golden):
    n_shadow_partial = 0
    for i, c in enumerate(golden["frag"]):
        o = (C.c_float * 4)()
        disc = ork_lib.ork_shade_fragment(fixture_dev.dev, C.byref(gu.scene_struct(c["scene"])), C.byref(gu.material_struct(c["material"])),
                                          c["shader"], C.byref(gu.frag_struct(c["in"])), o)
        assert disc == c["discard"], f"case {i}: discard decision differs"
        if not disc:
            assert close(list(o), c["out"]), f"case {i} (shader {c['shader']}): {list(o)} vs {c['out']}"
    assert sum(c["discard"] for c in golden["frag"]) >= 5       # the vectors exercise discard
    assert len({c["shader"] for c in golden["frag"]}) == 2
    del n_shadow_partial


def test_vertex_shaders_match_reference(ork_lib, golden):
    gu.bind_ork(ork_lib)
    for c in golden["vert"]:
        o = (C.c_float * 23)()
        ork_lib.ork_vertex(C.byref(gu.scene_struct(c["scene"])), gu.farr(c["world"]), gu.vertex_bytes(c["vertex"]), o)
        # gl_Position: the oracle (and the kernels) evaluate (viewproj*world)*p as an FMA chain — DESIGN.md §4 — so
        # it differs from glm's unfused evaluation by rounding only
        assert close(list(o), c["out"], 1e-5), (list(o), c["out"])
    for c in golden["shadow_vert"]:
        o = (C.c_float * 4)()
        ork_lib.ork_shadow_vertex(C.byref(gu.scene_struct(c["scene"])), gu.farr(c["world"]), c["cascade"], gu.vertex_bytes(c["vertex"]), o)
        assert close(list(o), c["out"], 1e-5)


def test_skybox_direction_matches_reference(ork_lib, golden):
    """skybox.vert gives (gl_Position, outUVW) per cube vertex; the oracle inverts that per pixel.  For a cube
    vertex in front of the camera the oracle's direction at the vertex's window position must be parallel to outUVW."""
    gu.bind_ork(ork_lib)
    W, H = 1280, 720
    checked = 0
    for c in golden["skybox"]:
        pos, uvw = c["out"][:4], np.array(c["out"][4:7])
        if pos[3] <= 1e-3:
            continue
        assert pos[2] == pos[3]                       # gl_Position = pos.xyww: depth 1.0
        sx = (pos[0] / pos[3] + 1.0) * W / 2
        sy = (pos[1] / pos[3] + 1.0) * H / 2
        o = (C.c_float * 3)()
        ork_lib.ork_skybox_dir(C.byref(gu.scene_struct(c["scene"])), W, H, sx, sy, o)
        d = np.array(list(o))
        cosang = d @ uvw / np.linalg.norm(d) / np.linalg.norm(uvw)
        assert cosang > 1 - 1e-5, (d, uvw)
        checked += 1
    assert checked >= 5


def test_compute_shaders_match_reference(ork_lib, golden):
    gu.bind_ork(ork_lib)
    for c in golden["tonemap"]:
        o = (C.c_float * 3)()
        ork_lib.ork_tonemap(gu.farr(c["in"][:3]), o)
        assert close(list(o), c["out"][:3], 1e-6)
        assert c["out"][3] == 1.0
    for c in golden["gradient"]:
        o = (C.c_float * 4)()
        push = (C.c_float * 16)(*c["push"])
        ork_lib.ork_gradient(push, c["y"], c["height"], o)
        assert close(list(o), c["out"], 1e-6)


def test_brdf_matches_reference(ork_lib, golden):
    gu.bind_ork(ork_lib)
    c0 = golden["brdf"][0]   # SURVEY.md §8c spot values, computed there from the reference's text
    assert abs(c0["D"] - 0.646285474) < 1e-6 and abs(c0["G"] - 0.915602982) < 1e-6 and abs(c0["F"][0] - 0.0400028825) < 1e-7
    for c in golden["brdf"]:
        D, G, F = C.c_float(), C.c_float(), (C.c_float * 3)()
        ork_lib.ork_brdf(gu.farr(c["N"]), gu.farr(c["V"]), gu.farr(c["L"]), c["roughness"], gu.farr(c["F0"]), C.byref(D), C.byref(G), F)
        assert close([D.value, G.value, *F], [c["D"], c["G"], *c["F"]])


@pytest.mark.skipif(not os.path.isdir("/root/reference/shaders"), reason="reference checkout not mounted")
def test_golden_is_current(tmp_path, golden):
    """Where the reference is present: regenerate from its shader text and require identical vectors."""
    import subprocess
    import sys
    import shutil
    HOST = os.path.join(ROOT, "tests", "golden", "host_math.json")
    bak, bak2 = tmp_path / "golden.json", tmp_path / "host.json"
    shutil.copy(GOLDEN, bak)
    shutil.copy(HOST, bak2)
    with open(HOST) as f:
        host_old = json.load(f)
    try:
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_golden.py")], check=True, capture_output=True)
        with open(GOLDEN) as f:
            fresh = json.load(f)
        with open(HOST) as f:
            host_new = json.load(f)
    finally:
        shutil.copy(bak, GOLDEN)
        shutil.copy(bak2, HOST)
    assert host_new == host_old
    for key in ("frag", "vert", "shadow_vert", "skybox", "tonemap", "gradient", "brdf"):
        assert len(fresh[key]) == len(golden[key])
        for a, b in zip(fresh[key], golden[key]):
            assert a["out" if "out" in a else "D"] == b["out" if "out" in b else "D"]
