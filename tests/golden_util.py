"""Shared between tools/gen_golden.py (writes tests/golden/shader_golden.json from the reference's
shader text) and tests/test_golden.py (checks oracle/ork.cpp against it): fixture construction and
ctypes plumbing for the oracle's shader-level entry points (oracle/ork.h)."""
from __future__ import annotations

import base64
import ctypes as C

import numpy as np

from vk_renderer_b200 import capi


class OrkFragIn(C.Structure):
    _fields_ = [("normal", C.c_float * 3), ("color", C.c_float * 3), ("uv", C.c_float * 2), ("lightVec", C.c_float * 3),
                ("posWorld", C.c_float * 4), ("posView", C.c_float * 4), ("duvdx", C.c_float * 2), ("duvdy", C.c_float * 2)]


class OrkTexCtx(C.Structure):
    _fields_ = [("dev", C.c_void_p), ("duvdx", C.c_float * 2), ("duvdy", C.c_float * 2)]


def bind_ork(ork):
    p = C.c_void_p
    ork.ork_shade_fragment.argtypes = [p, p, p, C.c_int, p, p]
    ork.ork_shade_fragment.restype = C.c_int
    ork.ork_load_shadow.argtypes = [p, p, C.c_size_t]
    ork.ork_brdf.argtypes = [p, p, p, C.c_float, p, p, p, p]
    ork.ork_brdf.restype = None
    ork.ork_tonemap.argtypes = [p, p]
    ork.ork_gradient.argtypes = [p, C.c_int, C.c_int, p]
    ork.ork_vertex.argtypes = [p, p, p, p]
    ork.ork_shadow_vertex.argtypes = [p, p, C.c_int, p, p]
    ork.ork_skybox_dir.argtypes = [p, C.c_int, C.c_int, C.c_float, C.c_float, p]


def f32list(a):
    return [float(x) for x in a]


def farr(values):
    return (C.c_float * len(values))(*values)


def b64(a: np.ndarray) -> str:
    return base64.b64encode(np.ascontiguousarray(a).tobytes()).decode()


def unb64(s: str, dtype, shape) -> np.ndarray:
    return np.frombuffer(base64.b64decode(s), dtype=dtype).reshape(shape).copy()


def unit(rng):
    v = rng.normal(size=3)
    return [float(x) for x in np.float32(v / np.linalg.norm(v))]


def normalize32(v):
    v = np.asarray(v, np.float32)
    return [float(x) for x in (v / np.float32(np.sqrt((v * v).sum(dtype=np.float32))))]


# ------------------------------------------------------------------------------- fixture
def make_fixture(rng) -> dict:
    """Device state every fragment case runs against: 4 shadow layers of blocky depth and four textures."""
    S = 32
    coarse = rng.uniform(0.15, 0.85, (4, S // 4, S // 4)).astype(np.float32)
    shadow = np.repeat(np.repeat(coarse, 4, 1), 4, 2)
    shadow += rng.uniform(-0.01, 0.01, shadow.shape).astype(np.float32)
    shadow[:, :2, :] = 1.0  # some untouched (cleared) texels
    tex = {}
    t = rng.integers(0, 256, (8, 8, 4), dtype=np.uint8)
    t[..., 3] = np.where(rng.random((8, 8)) < 0.3, 40, 255)          # alpha cut-outs
    tex["2"] = {"w": 8, "h": 8, "format": capi.FORMAT_SRGB, "mipmapped": 1, "sampler": capi.SAMPLER_LINEAR, "rgba": b64(t)}
    t = rng.integers(0, 256, (4, 4, 4), dtype=np.uint8)
    t[..., 3] = 255
    tex["3"] = {"w": 4, "h": 4, "format": capi.FORMAT_UNORM, "mipmapped": 0, "sampler": capi.SAMPLER_NEAREST, "rgba": b64(t)}
    t = rng.integers(0, 256, (8, 16, 4), dtype=np.uint8)
    t[..., 3] = 255
    tex["5"] = {"w": 16, "h": 8, "format": capi.FORMAT_SRGB, "mipmapped": 1, "sampler": capi.SAMPLER_LINEAR, "rgba": b64(t)}
    return {"shadow_extent": S, "shadow": b64(shadow), "textures": tex}


class FixtureDevice:
    """An OrkDevice holding the fixture's textures and shadow layers."""

    def __init__(self, ork, fixture):
        bind_ork(ork)
        S = fixture["shadow_extent"]
        self.r = capi.Renderer(ork, "ork", 16, 16, shadow_extent=S)
        self.dev = self.r.dev
        for slot, t in fixture["textures"].items():
            img = self.r.create_image(unb64(t["rgba"], np.uint8, (t["h"], t["w"], 4)), t["format"], bool(t["mipmapped"]))
            self.r.update_bindless_texture(int(slot), img, t["sampler"])
        sh = unb64(fixture["shadow"], np.float32, (4, S, S))
        rc = ork.ork_load_shadow(self.dev, sh.ctypes.data, sh.size)
        assert rc == 0


# ------------------------------------------------------------------------------- random inputs
def _look_at(eye, at):
    from vk_renderer_b200 import meshgen as mg
    return mg.look_at_rh(np.asarray(eye, np.float32), at)


def random_scene(rng, mode, shadows) -> dict:
    from vk_renderer_b200 import meshgen as mg
    eye = rng.uniform(-6, 6, 3)
    eye[2] = abs(eye[2]) + 6
    view = _look_at(eye, (0, 0, 0)).astype(np.float32)
    proj = mg.camera_projection(16 / 9)
    lights = []
    for i in range(4):
        ldir = np.array([0.5, -1.0, -0.5]) + rng.uniform(-0.2, 0.2, 3)
        lv = _look_at(-ldir / np.linalg.norm(ldir) * 20.0, (0, 0, 0))
        half = 5.0 * (i + 1)
        lp = mg.ortho_rh_zo(-half, half, -half, half, 0.1, 40.0)
        lights.append((lp @ lv).astype(np.float32))
    sd = capi.make_scene_data(view, proj, lights, shadow_mode=mode, enable_shadows=shadows,
                              cascade_distances=(8.0, 12.0, 16.0, 100.0),
                              ambient=tuple(np.float32(rng.uniform(0.02, 0.2, 3))) + (1.0,),
                              sun_dir=tuple(np.float32(rng.uniform(-1, 1, 3))) + (0.0,),
                              sun_color=tuple(np.float32(rng.uniform(0.5, 3.0, 3))) + (1.0,))
    return {"b64": b64(np.frombuffer(bytes(sd), np.float32))}


def scene_struct(sd: dict) -> capi.GPUSceneData:
    return capi.GPUSceneData.from_buffer_copy(base64.b64decode(sd["b64"]))


def random_material(rng, i) -> dict:
    return {"color": f32list(np.float32(rng.uniform(0.2, 1.0, 4))),
            "mr": [float(np.float32(rng.random() < 0.4)), float(np.float32(rng.uniform(0.1, 1.0))), 0.0, 0.0],
            "albedoID": [0, 2, 3, 5][(i // 8) % 4], "normalID": 1, "metalRoughID": [0, 2, 3][(i // 32) % 3]}


def material_struct(m: dict) -> capi.MaterialConstants:
    mc = capi.MaterialConstants()
    mc.colorFactors[:] = m["color"]
    mc.metal_rough_factors[:] = m["mr"]
    mc.albedoID, mc.normalID, mc.metalRoughID = m["albedoID"], m["normalID"], m["metalRoughID"]
    return mc


def random_frag(rng, sd: dict) -> dict:
    f = unb64(sd["b64"], np.float32, (128,))
    view = f[0:16].reshape(4, 4).T
    pw = np.append(np.float32(rng.uniform(-4, 4, 3)), np.float32(1.0)).astype(np.float32)
    pv = (view @ pw).astype(np.float32)
    sun = f[120:123]
    scale = float(10.0 ** rng.uniform(-3.5, -0.5))
    return {"normal": f32list(np.float32(rng.normal(size=3))), "color": f32list(np.float32(rng.uniform(0.3, 1, 3))),
            "uv": f32list(np.float32(rng.uniform(-2, 3, 2))), "lightVec": normalize32(-sun),
            "posWorld": f32list(pw), "posView": f32list(pv),
            "duvdx": f32list(np.float32(rng.normal(size=2) * scale)), "duvdy": f32list(np.float32(rng.normal(size=2) * scale))}


def frag_struct(fin: dict) -> OrkFragIn:
    s = OrkFragIn()
    for k, v in fin.items():
        getattr(s, k)[:] = v
    return s


def random_world(rng):
    from vk_renderer_b200 import meshgen as mg
    m = mg.translate(*rng.uniform(-5, 5, 3)) @ mg.scale(float(rng.uniform(0.5, 2.0)))
    rot = np.eye(4, dtype=np.float32)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    rot[:3, :3] = q
    return f32list((m @ rot).astype(np.float32).T.reshape(-1))   # column-major


def random_vertex(rng) -> dict:
    return {"position": f32list(np.float32(rng.uniform(-1, 1, 3))), "uv_x": float(np.float32(rng.random())),
            "normal": unit(rng), "uv_y": float(np.float32(rng.random())), "color": f32list(np.float32(rng.random(4)))}


def vertex_bytes(v: dict):
    a = np.zeros(1, capi.VERTEX_DTYPE)
    a["position"], a["uv_x"], a["normal"], a["uv_y"], a["color"] = v["position"], v["uv_x"], v["normal"], v["uv_y"], v["color"]
    return C.create_string_buffer(a.tobytes(), 48)
