"""The facing cull of whole clusters (vkr_raster.cuh cone_culls) may only drop a cluster whose every triangle the per-triangle
face cull — decided on SNAPPED coordinates — would drop as well.  Random meshes (smooth patches, slivers, sub-pixel triangles),
random perspective and orthographic matrices with reflections, both cull modes: no violation, and the cull must actually fire."""
import ctypes as C

import numpy as np
import pytest

from vk_renderer_b200 import meshgen as mg


def _check(lib, idx, pos, mvp, W, H, cull_front):
    lib.emu_cone_check.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.emu_cone_check.restype = C.c_int
    idx = np.ascontiguousarray(idx, np.uint32)
    p4 = np.ones((len(pos), 4), np.float32)
    p4[:, :3] = pos
    m = np.ascontiguousarray(np.asarray(mvp, np.float32).T.reshape(-1))          # column-major
    viol, ncl = C.c_uint32(), C.c_uint32()
    dropped = lib.emu_cone_check(idx.ctypes.data, len(idx) // 3, len(p4), p4.ctypes.data, m.ctypes.data, W, H, int(cull_front), C.byref(viol), C.byref(ncl))
    return dropped, viol.value, ncl.value


def _random_mvp(rng, perspective):
    ang = rng.uniform(0, 2 * np.pi, 3)
    rx = np.array([[1, 0, 0], [0, np.cos(ang[0]), -np.sin(ang[0])], [0, np.sin(ang[0]), np.cos(ang[0])]])
    ry = np.array([[np.cos(ang[1]), 0, np.sin(ang[1])], [0, 1, 0], [-np.sin(ang[1]), 0, np.cos(ang[1])]])
    rz = np.array([[np.cos(ang[2]), -np.sin(ang[2]), 0], [np.sin(ang[2]), np.cos(ang[2]), 0], [0, 0, 1]])
    world = np.eye(4)
    world[:3, :3] = (rx @ ry @ rz) * rng.uniform(0.05, 3.0, 3)[None, :]           # non-uniform scale
    if rng.random() < 0.3:
        world[:3, 0] *= -1.0                                                      # a reflection: winding flips
    world[:3, 3] = rng.uniform(-2, 2, 3)
    if perspective:
        view = mg.camera_view(tuple(rng.uniform(-1, 1, 3) + np.array([0, 0, rng.uniform(3, 60)])), pitch=float(rng.uniform(-0.3, 0.3)), yaw=float(rng.uniform(-0.3, 0.3)))
        proj = mg.camera_projection(16 / 9)
    else:
        view = mg.look_at_rh(rng.uniform(-1, 1, 3) * 30 + np.array([0, 40, 0]), (0, 0, 0), up=(0, 0, 1))
        half = rng.uniform(3, 60)
        proj = mg.ortho_rh_zo(-half, half, -half, half, 0.1, 200.0)
        if rng.random() < 0.5:
            proj[1, :] *= -1.0                                                    # the reference's single-light matrix is Y-flipped
    return (proj.astype(np.float64) @ view.astype(np.float64) @ world).astype(np.float32)


@pytest.mark.parametrize("seed", range(12))
def test_cone_cull_never_drops_a_triangle_the_face_cull_keeps(emu_lib, seed):
    rng = np.random.default_rng(100 + seed)
    fired = 0
    for trial in range(6):
        kind = (seed + trial) % 3
        if kind == 0:       # a tessellated sphere: smooth patches, half of them facing away
            v, i = mg.uv_sphere(int(rng.integers(12, 60)), int(rng.integers(12, 90)), 1.0)
            pos = np.asarray(v["position"], np.float64)
        elif kind == 1:     # a wavy sheet with very anisotropic triangles (slivers)
            n, m = int(rng.integers(8, 40)), int(rng.integers(8, 40))
            v, i = mg.plane(n, m, 2.0, 0.0)
            pos = np.asarray(v["position"], np.float64)
            pos[:, 0] *= rng.uniform(0.002, 1.0)
            pos[:, 1] = 0.2 * np.sin(3 * pos[:, 2] + rng.uniform(0, 6)) + rng.normal(0, 1e-4, len(pos))
        else:               # the sphere with jittered vertices (normals scatter, some triangles nearly degenerate)
            v, i = mg.uv_sphere(int(rng.integers(10, 40)), int(rng.integers(10, 40)), 1.0)
            pos = np.asarray(v["position"], np.float64) * (1.0 + rng.normal(0, 0.01, (len(v), 1)))
        for perspective in (True, False):
            mvp = _random_mvp(rng, perspective)
            W, H = (3840, 2160) if rng.random() < 0.5 else (int(rng.integers(64, 512)), int(rng.integers(64, 512)))
            for cull_front in (0, 1):
                dropped, viol, ncl = _check(emu_lib, i, pos, mvp, W, H, cull_front)
                assert viol == 0, (seed, trial, kind, perspective, cull_front, dropped, viol)
                fired += dropped
    assert fired > 0, "the facing cull never fired"


def test_cone_cull_fires_on_the_bench_sphere(emu_lib):
    """BASELINE config 4's sphere (10 k triangles, ~47 px radius at 4K; 30 of its 79 clusters face away entirely, 33 in the light's
    view): the cull takes those whose cone clears the silhouette and whose smallest triangle is not too small to be argued about
    (in the 2048^2 shadow map the sphere is 14 texels across: most of its triangles could flip under snapping)."""
    v, i = mg.uv_sphere(50, 100, 1.0)
    view = mg.camera_view((0.0, 7.0, 40.0), pitch=-0.17)
    mvp = (mg.camera_projection(3840 / 2160).astype(np.float64) @ view.astype(np.float64)).astype(np.float32)
    dropped, viol, ncl = _check(emu_lib, i, np.asarray(v["position"], np.float64), mvp, 3840, 2160, 0)
    assert viol == 0 and ncl == 79
    assert dropped >= 15, f"only {dropped} of {ncl} clusters dropped"
    L = mg.light_matrix()
    dropped_s, viol_s, _ = _check(emu_lib, i, np.asarray(v["position"], np.float64), L, 2048, 2048, 1)
    assert viol_s == 0 and dropped_s >= 10, f"shadow pass: only {dropped_s} of {ncl} clusters dropped"
