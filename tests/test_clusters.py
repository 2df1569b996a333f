"""build_clusters (csrc/vkr_frameprep.h): the draw's triangles cut into clusters of <= 128 triangles over <= 128 distinct
vertices.  The geometry kernels trust these invariants blindly (7-bit local ids, 2 KB of shared memory per warp, sequence
numbers from `orig`), so they are checked here on ragged inputs: random index soups with repeated and degenerate triangles,
meshes with far more vertices than triangles, counts around the 128 boundaries, the empty draw."""
import ctypes as C

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

CL = 128


def build(emu_lib, idx, pos):
    idx = np.ascontiguousarray(idx, np.uint32)
    pos4 = np.ones((len(pos), 4), np.float32)
    pos4[:, :3] = pos
    ntri = len(idx) // 3
    cap_c = ntri + 2
    verts = np.zeros(max(1, 3 * ntri), np.uint32)
    tris = np.zeros(cap_c * CL, np.uint32)
    orig = np.zeros(cap_c * CL, np.uint32)
    info = np.zeros((cap_c, 2), np.uint32)
    bounds = np.zeros((cap_c, 8), np.float32)
    used = C.c_size_t()
    f = emu_lib.emu_build_clusters
    f.restype = C.c_int
    f.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t,
                  C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    nc = f(idx.ctypes.data, ntri, len(pos), pos4.ctypes.data, verts.ctypes.data, len(verts), tris.ctypes.data, orig.ctypes.data,
           len(tris), info.ctypes.data, bounds.ctypes.data, cap_c, C.byref(used))
    return nc, verts[:used.value], tris, orig, info[:nc], bounds[:nc]


def check(emu_lib, idx, pos):
    idx = np.asarray(idx, np.uint32)
    ntri = len(idx) // 3
    nc, verts, tris, orig, info, bounds = build(emu_lib, idx, pos)
    assert (nc == 0) == (ntri == 0)
    seen = np.zeros(ntri, np.int32)
    off_expected = 0
    for c in range(nc):
        off, counts = int(info[c, 0]), int(info[c, 1])
        nv, nt = counts & 0xFFFF, counts >> 16
        assert 1 <= nt <= CL and 1 <= nv <= CL
        assert off == off_expected, "vertex lists are packed back to back"
        off_expected += nv
        lv = verts[off:off + nv]
        assert len(np.unique(lv)) == nv, "a cluster lists each vertex once"
        p = np.asarray(pos, np.float32)[lv]
        assert np.array_equal(bounds[c, 0:3], p.min(0)) and np.array_equal(bounds[c, 4:7], p.max(0)), "bounds are the exact AABB of the cluster's vertices"
        for t in range(nt):
            packed, o = int(tris[c * CL + t]), int(orig[c * CL + t])
            l = (packed & 127, (packed >> 7) & 127, (packed >> 14) & 127)
            assert max(l) < nv
            assert o < ntri
            assert tuple(int(lv[k]) for k in l) == tuple(int(v) for v in idx[3 * o:3 * o + 3]), "local ids name the triangle's own vertices, in order"
            seen[o] += 1
    assert (seen == 1).all(), "every triangle of the draw is in exactly one cluster"
    return nc


@pytest.mark.parametrize("ntri", [0, 1, 127, 128, 129, 255, 256, 257, 1000])
def test_counts_around_the_cluster_size(emu_lib, ntri):
    rng = np.random.default_rng(ntri)
    nv = max(3, ntri // 2 + 3)
    pos = rng.uniform(-1, 1, (nv, 3))
    idx = rng.integers(0, nv, 3 * ntri)
    nc = check(emu_lib, idx, pos)
    assert nc >= (ntri + CL - 1) // CL


def test_triangles_that_share_nothing_are_cut_by_the_vertex_limit(emu_lib):
    ntri = 300                                  # 900 distinct vertices: at most 42 triangles fit 128 vertices
    pos = np.random.default_rng(1).uniform(-1, 1, (3 * ntri, 3))
    nc = check(emu_lib, np.arange(3 * ntri), pos)
    assert nc >= (3 * ntri + CL - 1) // CL


def test_degenerate_and_repeated_triangles(emu_lib):
    pos = np.random.default_rng(2).uniform(-1, 1, (10, 3))
    idx = np.array([0, 0, 0, 1, 1, 2, 3, 4, 5, 3, 4, 5, 3, 4, 5, 9, 9, 8] * 40)
    check(emu_lib, idx, pos)


def test_a_tessellated_sphere_gets_compact_clusters(emu_lib):
    from vk_renderer_b200 import meshgen as mg
    sv, si = mg.uv_sphere(50, 100, 1.0)          # BASELINE config 4's mesh
    nc = check(emu_lib, si, sv["position"])
    assert nc <= 100, f"{nc} clusters for 10,000 triangles: the Morton order should keep them near the 79 of a perfect cut"
    _, _, _, _, _, bounds = build(emu_lib, si, sv["position"])
    diag = np.linalg.norm(bounds[:, 4:7] - bounds[:, 0:3], axis=1)
    # a disc of a 79th of the sphere's area has a box diagonal of 0.64 to 0.78 depending on its orientation; the Morton order
    # used before gave 0.93 (max 1.42)
    assert np.median(diag) < 0.8 and diag.max() < 1.35, "clusters are surface patches, not slices through the whole sphere"


@settings(max_examples=40, deadline=None)
@given(st.integers(0, 700), st.integers(3, 400), st.integers(0, 2**31 - 1))
def test_random_soups(emu_lib, ntri, nv, seed):
    rng = np.random.default_rng(seed)
    pos = rng.uniform(-5, 5, (nv, 3)) * rng.choice([1.0, 1e-3, 1e3])
    if rng.random() < 0.3:                      # many coincident centroids: the Morton sort has nothing to separate
        pos[:] = pos[0]
    idx = rng.integers(0, nv, 3 * ntri)
    check(emu_lib, idx, pos)
