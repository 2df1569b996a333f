"""Known answers for the ORACLE's fixed-function half (oracle/ork.cpp setup_tri / make_edges / covered / depth_at,
clipping, mip blits, texture unit, cube faces), rendered on `ork` ALONE.

The product is held bit-exact to the oracle; this file is what holds the oracle itself to the Vulkan rules (SURVEY.md
Appendix B) for the state the reference sets — cull BACK / front CCW / depth LESS_OR_EQUAL + write
(/root/reference/src/vk_engine_init.cpp:532-547), LINEAR mip blits (/root/reference/src/vk_images.cpp:64-157), REPEAT / LINEAR
samplers (/root/reference/src/vk_initializers.cpp:228-248).  Expected values are either written out by hand from the rules or
computed by `rule_raster` below: an independent restatement in exact rational arithmetic where the top-left rule is decided
geometrically (a sample ON an edge counts iff that edge is horizontal with the interior below it, or non-horizontal with the
interior to its right) instead of the oracle's integer bias trick.
"""
from fractions import Fraction as Fr
import ctypes as C

import numpy as np
import pytest

from vk_renderer_b200 import capi, meshgen as mg
import scenes

W, H = 64, 64
NONE = 0xFFFFFFFF
BACKEND_PREFIX = "ork"     # tests/test_rules_on_kernels.py re-runs the raster / depth / blend cases below on the kernels themselves


# ------------------------------------------------------------------ the rules, restated independently (exact)
def snap(v):
    """24.8 fixed point, round half to even (subPixelPrecisionBits = 8)."""
    return Fr(int(np.rint(np.float64(np.float32(v)) * 256.0)), 256)


def rule_covers(tri, px, py):
    """tri: three (x, y) Fractions in window coordinates (y down), any winding, non-degenerate."""
    s = (Fr(px) + Fr(1, 2), Fr(py) + Fr(1, 2))
    for i in range(3):
        a, b, c = tri[i], tri[(i + 1) % 3], tri[(i + 2) % 3]
        side = lambda p: (b[0] - a[0]) * (p[1] - a[1]) - (b[1] - a[1]) * (p[0] - a[0])
        inside_sign = side(c)
        v = side(s)
        if v * inside_sign > 0:
            continue
        if v * inside_sign < 0:
            return False
        # exactly on the edge a-b: top edge (horizontal, interior at larger y) or left edge (interior at larger x)
        if a[1] == b[1]:
            if not c[1] > a[1]:
                return False
        else:
            # x of the edge at the height of the opposite vertex vs that vertex
            t = (c[1] - a[1]) / (b[1] - a[1])
            xe = a[0] + t * (b[0] - a[0])
            if not c[0] > xe:
                return False
    return True


def rule_raster(tris, front_only=True):
    """-> dict pixel -> list of triangle numbers producing a fragment there (submission order)."""
    out = {}
    for k, t in enumerate(tris):
        p = [(snap(x), snap(y)) for x, y, *_ in t]
        area2 = (p[1][0] - p[0][0]) * (p[2][1] - p[0][1]) - (p[2][0] - p[0][0]) * (p[1][1] - p[0][1])
        if area2 == 0:
            continue
        if front_only and not area2 < 0:      # front = counter-clockwise as seen on the (y-down) screen
            continue
        x0, x1 = int(min(q[0] for q in p)) - 1, int(max(q[0] for q in p)) + 1
        y0, y1 = int(min(q[1] for q in p)) - 1, int(max(q[1] for q in p)) + 1
        for py in range(max(0, y0), min(H, y1 + 1)):
            for px in range(max(0, x0), min(W, x1 + 1)):
                if rule_covers(p, px, py):
                    out.setdefault((px, py), []).append(k)
    return out


# ------------------------------------------------------------------ driving the oracle
class Ork:
    def __init__(self, lib, tris, colors=None):
        self.r = capi.Renderer(lib, BACKEND_PREFIX, W, H, shadow_extent=64)
        pts = np.array([[tuple(v) + ((0.5,) if len(v) == 2 else ()) for v in t] for t in tris], np.float32)
        v, i = mg.triangles(pts, colors)
        self.mesh = self.r.upload_mesh(i, v)
        self.n = len(tris)
        self.opaque = self.r.build_material((1, 1, 1, 1), 0.0, 0.8)
        self.additive = self.r.build_material((1, 1, 1, 1), 0.0, 0.8, pass_type=capi.PASS_TRANSPARENT)
        view, proj = scenes.pixel_space(self.r)
        # no sun: every fragment's colour is ambient * base = 0.25 * 0.50196 (bindless slot 0) on r, g
        self.sd = capi.make_scene_data(view, proj, None, sun_dir=(0.0, 0.0, 1.0, 0.0), sun_color=(0, 0, 0, 1), ambient=(0.25, 0.25, 0.25, 1.0))
        self.fp = capi.default_frame_params(fragShader=capi.FRAG_LAMBERT)

    def draw(self, which=None, material=None):
        which = range(self.n) if which is None else which
        mat = self.opaque if material is None else material
        objs = np.array([mg.render_object(self.mesh, mat, np.eye(4), 3, first_index=3 * k) for k in which])
        self.r.draw(self.sd, objs, self.fp)
        return self.r.readback(capi.TARGET_PRIMID), self.r.readback(capi.TARGET_DEPTH)

    def fragment_counts(self):
        """Fragments per pixel: the triangles drawn as "Transparent" (additive, no depth write) sum a constant colour."""
        objs = np.array([mg.render_object(self.mesh, self.additive, np.eye(4), 3 * self.n)])
        self.r.draw(self.sd, objs, self.fp)
        hdr = self.r.readback(capi.TARGET_HDR16F).astype(np.float64)[..., 0]
        unit = hdr[hdr > 0].min() if (hdr > 0).any() else 1.0
        return np.rint(hdr / unit).astype(int)

    def close(self):
        self.r.close()


def pixels(prim, k=None):
    ys, xs = np.nonzero(prim != NONE if k is None else prim == k)
    return set(zip(xs.tolist(), ys.tolist()))


def front(x0, y0, x1, y1, z=0.5):
    """The two front-facing triangles of an axis-aligned rectangle."""
    return [[(x0, y0, z), (x0, y1, z), (x1, y0, z)], [(x1, y0, z), (x0, y1, z), (x1, y1, z)]]


# ------------------------------------------------------------------ coverage
def test_rectangle_on_pixel_corners_covers_exactly_its_pixels(ork_lib):
    o = Ork(ork_lib, front(30, 2, 37, 9))
    prim, _ = o.draw()
    want = {(x, y) for x in range(30, 37) for y in range(2, 9)}           # centres 30.5..36.5 x 2.5..8.5 lie strictly inside
    assert pixels(prim) == want
    assert (o.fragment_counts() == 1).sum() == len(want) and o.fragment_counts().max() == 1
    o.close()


def test_rectangle_on_pixel_centres_keeps_top_and_left_drops_bottom_and_right(ork_lib):
    o = Ork(ork_lib, front(40.5, 2.5, 47.5, 9.5))
    prim, _ = o.draw()
    # the samples of column 40 lie ON the left edge (kept), those of column 47 ON the right edge (dropped); row 2 ON the top
    # edge (kept), row 9 ON the bottom edge (dropped)
    assert pixels(prim) == {(x, y) for x in range(40, 47) for y in range(2, 9)}
    o.close()


def test_shared_diagonal_through_pixel_centres_is_covered_exactly_once(ork_lib):
    tris = front(2.5, 2.5, 10.5, 10.5)
    o = Ork(ork_lib, tris)
    prim, _ = o.draw()
    # the diagonal x + y = 12 (pixel indices) runs through sample points; it is a LEFT edge of the lower-right triangle
    a = {(x, y) for x in range(2, 10) for y in range(2, 10) if x + y <= 11}
    b = {(x, y) for x in range(2, 10) for y in range(2, 10) if x + y >= 12}
    assert pixels(prim, 0) == a and pixels(prim, 1) == b
    only0, _ = o.draw([0])
    only1, _ = o.draw([1])
    assert pixels(only0) == a and pixels(only1) == b                  # each alone: disjoint, union = the 8x8 square
    cnt = o.fragment_counts()
    assert cnt.max() == 1 and cnt.sum() == 64
    o.close()


def test_fan_vertex_on_a_pixel_centre_is_covered_exactly_once(ork_lib):
    c = (20.5, 8.5)
    ring = [(14.5, 2.5), (14.5, 14.5), (26.5, 14.5), (26.5, 2.5)]
    tris = [[c, ring[i], ring[(i + 1) % 4]] for i in range(4)]        # left, bottom, right, top
    o = Ork(ork_lib, tris)
    prim, _ = o.draw()
    cnt = o.fragment_counts()
    assert pixels(prim) == {(x, y) for x in range(14, 26) for y in range(2, 14)}
    assert cnt.max() == 1 and cnt.sum() == 144
    # the centre sample lies on two edges of every triangle; only for the RIGHT one are both left edges
    assert prim[8, 20] == 2
    # the diagonals: x - 14 == y - 2 belongs to ... derive every pixel from the rules
    want = rule_raster(tris)
    for (x, y), ks in want.items():
        assert len(ks) == 1 and prim[y, x] == ks[0], (x, y, ks, prim[y, x])
    o.close()


def test_slivers_and_degenerates(ork_lib):
    tris = [
        [(2.0, 20.0), (2.0, 20.003), (30.0, 20.0)],               # 1 sub-pixel high, between sample rows: nothing
        [(2.0, 22.5), (30.0, 22.5), (2.0, 22.5)],                 # zero area, through samples: nothing
        [(5.5, 24.5), (5.5, 24.5), (5.5, 24.5)],                  # a point on a sample: nothing
        [(2.0, 26.5), (2.0, 26.5 + 1 / 256), (30.0, 26.5)],       # 1 sub-pixel high with its TOP edge through the samples of row 26
        [(2.0, 28.5 - 1 / 256), (2.0, 28.5), (30.0, 28.5 - 1 / 256)],   # the same sliver with its BOTTOM vertex on the sample row: nothing
        [(2.0, 30.2), (2.0, 30.2 + 1 / 1024), (30.0, 30.2)],      # thinner than the snapping grid: collapses to zero area
    ]
    o = Ork(ork_lib, tris)
    prim, _ = o.draw()
    assert pixels(prim) == {(x, 26) for x in range(2, 30)}        # samples 2.5 .. 29.5 lie on the top edge, left of the hypotenuse's end
    assert pixels(prim, 3) == pixels(prim)
    assert o.fragment_counts().sum() == 28
    o.close()


def test_back_faces_are_culled_and_front_is_counter_clockwise(ork_lib):
    ccw = [(50.5, 2.5), (50.5, 10.5), (58.5, 2.5)]                # down, then right: counter-clockwise on a y-down screen
    cw = [(50.5, 22.5), (58.5, 22.5), (50.5, 30.5)]
    o = Ork(ork_lib, [ccw, cw])
    prim, _ = o.draw()
    assert len(pixels(prim, 0)) == 36 and len(pixels(prim, 1)) == 0
    o.close()


def test_depth_test_is_less_or_equal_in_submission_order(ork_lib):
    t = [(2.5, 30.5), (2.5, 40.5), (12.5, 30.5)]
    z = lambda zz: [p + (zz,) for p in t]
    o = Ork(ork_lib, [z(0.25), z(0.25), z(0.75), z(0.125), z(0.125)])
    prim, depth = o.draw([0, 1])
    assert pixels(prim) == pixels(prim, 1)                        # equal depth: the later primitive passes LESS_OR_EQUAL and overwrites
    prim, depth = o.draw([0, 2])
    assert pixels(prim) == pixels(prim, 0)                        # farther one later: fails
    prim, depth = o.draw([2, 0, 3, 1, 4])
    assert pixels(prim) == pixels(prim, 4) and len(pixels(prim)) == 55     # nearest wins; among the two nearest, the later (ids are draw-list order)
    assert np.all(depth[prim != NONE].view(np.uint32) == np.float32(0.125).view(np.uint32))
    assert np.all(depth[prim == NONE].view(np.uint32) == 0x3F800000)       # cleared to 1.0
    o.close()


def test_depth_zero_and_one_are_kept(ork_lib):
    o = Ork(ork_lib, [[(30.5, 30.5, 0.0), (30.5, 40.5, 0.0), (40.5, 30.5, 0.0)], [(42.5, 30.5, 1.0), (42.5, 40.5, 1.0), (52.5, 30.5, 1.0)]])
    prim, depth = o.draw()
    assert len(pixels(prim, 0)) == 55 and len(pixels(prim, 1)) == 55       # depth 1.0 passes LEQUAL against the cleared 1.0
    assert np.all(depth[prim == 0] == 0.0) and np.all(depth[prim == 1] == 1.0)
    o.close()


def test_depth_is_the_plane_through_the_vertices_at_the_sample(ork_lib):
    # power-of-two slopes: every intermediate of the f32 evaluation is exact, so the result must equal the exact plane to the bit
    exact = [(16.5, 8.5, 0.25), (16.5, 40.5, 0.75), (48.5, 8.5, 0.5)]        # dz/dy = 1/64, dz/dx = 1/128
    general = [(3.3, 41.7, 0.1), (9.1, 62.2, 0.93), (55.6, 44.4, 0.37)]
    o = Ork(ork_lib, [exact, general])
    prim, depth = o.draw()
    ys, xs = np.nonzero(prim == 0)
    assert len(xs) > 400
    want = 0.25 + (ys + 0.5 - 8.5) / 64.0 + (xs + 0.5 - 16.5) / 128.0
    assert np.array_equal(depth[ys, xs].view(np.uint32), want.astype(np.float32).view(np.uint32))
    # a general plane: float64 evaluation through the SNAPPED vertices, agreement to a few ulp of f32
    p = np.array([[float(snap(x)), float(snap(y)), np.float64(np.float32(zz))] for x, y, zz in general])
    n = np.cross(p[1] - p[0], p[2] - p[0])
    ys, xs = np.nonzero(prim == 1)
    assert len(xs) > 300
    want = p[0, 2] - (n[0] * (xs + 0.5 - p[0, 0]) + n[1] * (ys + 0.5 - p[0, 1])) / n[2]
    assert np.abs(depth[ys, xs].astype(np.float64) - want).max() < 4e-7
    # and its coverage from the rules
    assert pixels(prim, 1) == {k for k, v in rule_raster([exact, general]).items() if v == [1]}      # (the two do not overlap)
    o.close()


def test_random_triangles_match_the_rules(ork_lib):
    rng = np.random.default_rng(42)
    tris = []
    for k in range(40):
        c = rng.uniform(4, 60, 2)
        p = c + rng.normal(0, 5.0, (3, 2))
        if k % 3 == 0:
            p = np.round(p * 2) / 2                                   # vertices on sample points and pixel corners
        if (p[1, 0] - p[0, 0]) * (p[2, 1] - p[0, 1]) - (p[2, 0] - p[0, 0]) * (p[1, 1] - p[0, 1]) > 0:
            p = p[[0, 2, 1]]
        tris.append([(float(a), float(b), 0.5) for a, b in p])
    o = Ork(ork_lib, tris)
    want = rule_raster(tris)
    cnt = o.fragment_counts()
    for y in range(H):
        for x in range(W):
            assert cnt[y, x] == len(want.get((x, y), [])), (x, y)
    prim, _ = o.draw()
    for (x, y), ks in want.items():
        assert prim[y, x] == ks[-1]                                   # equal depths: the last one submitted
    o.close()


def test_near_plane_clip_of_one_triangle(ork_lib):
    # pixel space has w = 1 and z_clip = z: the third vertex is behind the near plane (z < 0).  Clipping at z = 0 cuts both
    # edges at their midpoints: the polygon is (10.5,10.5) (10.5,30.5) (30.5,20.5) (30.5,10.5), depth 0.5 -> 0 along x.
    o = Ork(ork_lib, [[(10.5, 10.5, 0.5), (10.5, 30.5, 0.5), (50.5, 10.5, -0.5)]])
    prim, depth = o.draw()
    # left edge x = 10.5 and top edge y = 10.5 own their samples; the new edge x = 30.5 is a right edge; the slanted lower edge
    # y = 30.5 - (x - 10.5) / 2 is a bottom edge: sample (x+.5, y+.5) is covered iff y + .5 < 30.5 - (x + .5 - 10.5) / 2
    want = {(x, y) for x in range(10, 30) for y in range(10, 31) if 2 * y + 1 < 61 - (x - 10)}
    assert pixels(prim) == want
    ys, xs = np.nonzero(prim == 0)
    zs = 0.5 - (xs + 0.5 - 10.5) / 40.0                               # the unclipped triangle's plane: z falls 1.0 over 40 px
    assert np.abs(depth[ys, xs] - zs).max() < 1e-6
    assert depth[prim == 0].min() >= 0.0
    assert o.fragment_counts().max() == 1                             # the fan's internal edge: no double hits
    o.close()


def test_far_plane_clip(ork_lib):
    o = Ork(ork_lib, [[(10.5, 10.5, 0.5), (10.5, 30.5, 0.5), (50.5, 10.5, 1.5)]])     # z > w beyond x = 30.5
    prim, depth = o.draw()
    want = {(x, y) for x in range(10, 30) for y in range(10, 31) if 2 * y + 1 < 61 - (x - 10)}
    assert pixels(prim) == want and depth[prim == 0].max() <= 1.0
    o.close()


# ------------------------------------------------------------------ texture unit
def srgb_decode(c):
    c = c / 255.0
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)


def srgb_encode(l):
    return np.where(l <= 0.0031308, l * 12.92, 1.055 * np.power(np.maximum(l, 1e-30), 1 / 2.4) - 0.055)


def image_level(lib, r, img, level, w, h):
    fn = getattr(lib, f"{BACKEND_PREFIX}_image_level")
    fn.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    out = np.zeros((h, w, 4), np.uint8)
    ww, hh = C.c_uint32(), C.c_uint32()
    assert fn(r.dev, img, level, out.ctypes.data, out.nbytes, C.byref(ww), C.byref(hh)) == 0
    assert (ww.value, hh.value) == (w, h)
    return out


def test_mip_blit_is_a_linear_light_box_filter(ork_lib):
    r = capi.Renderer(ork_lib, BACKEND_PREFIX, 16, 16, shadow_extent=16)
    tex = np.array([[[255, 0, 10, 255], [0, 0, 200, 0]], [[0, 255, 90, 255], [0, 0, 130, 0]]], np.uint8)
    img = r.create_image(tex, capi.FORMAT_SRGB, True)                 # generate_mipmaps: 2x2 -> 1x1 by one LINEAR blit
    lvl = image_level(ork_lib, r, img, 1, 1, 1)[0, 0]
    lin = srgb_decode(tex[..., :3].astype(np.float64)).reshape(4, 3).mean(0)
    want = np.floor(srgb_encode(lin) * 255.0 + 0.5)
    assert np.array_equal(lvl[:3], want.astype(np.uint8))             # (188, 188, 129): averaging happens in linear light
    assert lvl[3] == 128                                              # alpha is linear: (255 + 0 + 255 + 0) / 4 = 127.5 -> 128
    img2 = r.create_image(tex, capi.FORMAT_UNORM, True)
    lvl2 = image_level(ork_lib, r, img2, 1, 1, 1)[0, 0]
    assert np.array_equal(lvl2, np.floor(tex.reshape(4, 4).astype(np.float64).mean(0) + 0.5).astype(np.uint8))
    # an 8x4 image halves to 4x2, 2x1, 1x1: floor(log2(max)) + 1 = 4 levels
    img3 = r.create_image(np.full((4, 8, 4), 77, np.uint8), capi.FORMAT_UNORM, True)
    assert np.all(image_level(ork_lib, r, img3, 3, 1, 1) == 77) and np.all(image_level(ork_lib, r, img3, 2, 2, 1) == 77)
    r.close()


class TexCtx(C.Structure):
    _fields_ = [("dev", C.c_void_p), ("duvdx", C.c_float * 2), ("duvdy", C.c_float * 2)]


def sample2d(lib, r, slot, u, v, ddx=(0, 0), ddy=(0, 0)):
    fn = getattr(lib, f"{BACKEND_PREFIX}_cb_tex2d")
    fn.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.POINTER(C.c_float)]
    fn.restype = None
    ctx = TexCtx(r.dev, (C.c_float * 2)(*ddx), (C.c_float * 2)(*ddy))
    out = (C.c_float * 4)()
    fn(C.byref(ctx), slot, u, v, out)
    return np.array(out[:], np.float64)


def test_bilinear_repeat_sample_across_the_wrap(ork_lib):
    r = capi.Renderer(ork_lib, BACKEND_PREFIX, 16, 16, shadow_extent=16)
    tex = np.zeros((4, 4, 4), np.uint8)
    tex[..., 0] = np.array([10, 50, 90, 250])[None, :]                # red varies along x
    tex[..., 1] = np.array([0, 100, 200, 40])[:, None]                # green varies along y
    tex[..., 3] = 255
    img = r.create_image(tex, capi.FORMAT_UNORM, False)
    r.update_bindless_texture(9, img, capi.SAMPLER_LINEAR)
    # u = 0 is the border between the last and the first texel column: texel coordinate -0.5 -> half of column 3, half of column 0
    got = sample2d(ork_lib, r, 9, 0.0, 0.375)                         # v: centre of row 1
    assert abs(got[0] - (250 + 10) / 2 / 255) < 1e-6 and abs(got[1] - 100 / 255) < 1e-6
    got = sample2d(ork_lib, r, 9, 0.625, 1.0)                         # centre of column 2; v = 1 wraps: rows 3 and 0
    assert abs(got[0] - 90 / 255) < 1e-6 and abs(got[1] - (40 + 0) / 2 / 255) < 1e-6
    got = sample2d(ork_lib, r, 9, -0.8125, 2.375)                     # REPEAT: -0.8125 = 0.1875 (mod 1): 1/4 of the way from column 0 to 1
    assert abs(got[0] - (0.75 * 10 + 0.25 * 50) / 255) < 1e-6 and abs(got[1] - 100 / 255) < 1e-6
    r.update_bindless_texture(10, img, capi.SAMPLER_NEAREST)
    got = sample2d(ork_lib, r, 10, 0.74, 0.26)                        # NEAREST: floor(0.74 * 4) = 2, floor(0.26 * 4) = 1
    assert abs(got[0] - 90 / 255) < 1e-6 and abs(got[1] - 100 / 255) < 1e-6
    r.close()


def test_trilinear_level_selection(ork_lib):
    r = capi.Renderer(ork_lib, BACKEND_PREFIX, 16, 16, shadow_extent=16)
    tex = np.zeros((8, 8, 4), np.uint8)
    tex[::2, ::2, 0] = 200                                            # level 1 = 50 everywhere, level 2 = 50, level 3 = 50
    tex[..., 3] = 255
    img = r.create_image(tex, capi.FORMAT_UNORM, True)
    r.update_bindless_texture(11, img, capi.SAMPLER_LINEAR)
    # one texel per pixel: lambda = log2(8 * 1/8) = 0 -> level 0 only (magnification / exact): centre of texel (0,0)
    assert abs(sample2d(ork_lib, r, 11, 1 / 16, 1 / 16, (1 / 8, 0), (0, 1 / 8))[0] - 200 / 255) < 1e-6
    # two texels per pixel: lambda = 1 -> level 1 only
    assert abs(sample2d(ork_lib, r, 11, 1 / 16, 1 / 16, (2 / 8, 0), (0, 1 / 8))[0] - 50 / 255) < 1e-6
    # sqrt(2) texels per pixel: lambda = 0.5 -> halfway between level 0 and level 1
    s = 2 ** 0.5
    assert abs(sample2d(ork_lib, r, 11, 1 / 16, 1 / 16, (s / 8, 0), (0, 1 / 8))[0] - (200 + 50) / 2 / 255) < 1e-5
    # the longer footprint axis decides, and lambda clamps to the last level
    assert abs(sample2d(ork_lib, r, 11, 1 / 16, 1 / 16, (0.01, 0), (0, 100.0))[0] - 50 / 255) < 1e-6
    r.close()


def test_cube_face_selection_and_orientation(ork_lib):
    r = capi.Renderer(ork_lib, BACKEND_PREFIX, 16, 16, shadow_extent=16)
    n = 32
    faces = []
    for f in range(6):
        img = np.zeros((n, n, 4), np.uint8)
        img[..., 0] = (np.arange(n) * 8)[None, :]                     # red = 8 * column, green = 8 * row, blue = 40 * face
        img[..., 1] = (np.arange(n) * 8)[:, None]
        img[..., 2] = 40 * f
        img[..., 3] = 255
        faces.append(img)
    r.set_cubemap(faces, capi.FORMAT_UNORM, False)
    cb_cube = getattr(ork_lib, f"{BACKEND_PREFIX}_cb_cube")
    cb_cube.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    cb_cube.restype = None
    ctx = TexCtx(r.dev, (C.c_float * 2)(0, 0), (C.c_float * 2)(0, 0))
    # Vulkan's table: face, (s_c, t_c) for the major axis; s = (s_c / |m| + 1) / 2
    table = [((1, 0, 0), lambda x, y, z: (-z, -y)), ((-1, 0, 0), lambda x, y, z: (z, -y)), ((0, 1, 0), lambda x, y, z: (x, z)),
             ((0, -1, 0), lambda x, y, z: (x, -z)), ((0, 0, 1), lambda x, y, z: (x, -y)), ((0, 0, -1), lambda x, y, z: (-x, -y))]
    for f, (axis, st) in enumerate(table):
        d = np.array(axis, np.float64) * 2.0
        d[(np.argmax(np.abs(d)) + 1) % 3] = 0.5
        d[(np.argmax(np.abs(d)) + 2) % 3] = -0.7
        out = (C.c_float * 4)()
        cb_cube(C.byref(ctx), (C.c_float * 3)(*d), out)
        sc, tc = st(*d)
        s, t = 0.5 * (sc / 2.0 + 1.0), 0.5 * (tc / 2.0 + 1.0)
        assert abs(out[2] * 255 - 40 * f) < 1e-3, f"direction {d} must select face {f}"
        assert abs(out[0] * 255 - 8 * (s * n - 0.5)) < 1e-2 and abs(out[1] * 255 - 8 * (t * n - 0.5)) < 1e-2, (f, out[:])
    r.close()


def test_cube_filtering_is_seamless_across_all_twelve_edges(ork_lib):
    """Vulkan filters cube maps seamlessly ("Cube Map Edge Handling"): a bilinear tap beyond a face edge comes from the face that
    shares the edge.  Stated without any face table: the filtered colour is CONTINUOUS across an edge — two directions a hair
    apart on either side of it (which select different faces) must give the same colour — whereas clamping the taps to the
    selected face jumps by about half the difference of the two faces' edge texels.  Random texels, so that a neighbour
    addressed with a mirrored or transposed coordinate breaks the continuity too.  (Cube corners, where a tap has no texel,
    are kept a texel and a half away.)"""
    r = capi.Renderer(ork_lib, BACKEND_PREFIX, 16, 16, shadow_extent=16)
    n = 8
    rng = np.random.default_rng(7)
    faces = [np.concatenate([rng.integers(0, 256, (n, n, 3), dtype=np.uint8), np.full((n, n, 1), 255, np.uint8)], -1) for _ in range(6)]
    r.set_cubemap(faces, capi.FORMAT_UNORM, False)
    cb_cube = getattr(ork_lib, f"{BACKEND_PREFIX}_cb_cube")
    cb_cube.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    cb_cube.restype = None
    ctx = TexCtx(r.dev, (C.c_float * 2)(0, 0), (C.c_float * 2)(0, 0))

    def sample(d):
        out = (C.c_float * 4)()
        cb_cube(C.byref(ctx), (C.c_float * 3)(*[float(v) for v in d]), out)
        return np.array(out[:3], np.float64)

    axes = np.eye(3)
    edges, jumps = 0, []
    for i in range(3):
        for j in range(i + 1, 3):
            k = 3 - i - j
            for si in (1.0, -1.0):
                for sj in (1.0, -1.0):                       # the edge shared by the faces (si * axis i) and (sj * axis j)
                    edges += 1
                    for along in (-0.55, -0.2, 0.13, 0.4, 0.6):
                        d = si * axes[i] + sj * axes[j] + along * axes[k]
                        eps = 1e-4
                        a, b = sample(d + eps * si * axes[i]), sample(d + eps * sj * axes[j])
                        jumps.append(np.abs(a - b).max())
    assert edges == 12
    assert max(jumps) < 2e-3, f"colour jumps by {max(jumps):.4f} across a cube edge"
    r.close()


# ------------------------------------------------------------------ scissor, attachments, blending, the shadow pass's state
def test_scissor_keeps_exactly_its_pixels(ork_lib):
    """VkRect2D scissor: a fragment at pixel (x, y) survives iff offset <= (x, y) < offset + extent."""
    o = Ork(ork_lib, front(0, 0, 64, 64))
    o.fp.scissor[:] = (5, 7, 20, 11)
    prim, depth = o.draw()
    assert pixels(prim) == {(x, y) for x in range(5, 25) for y in range(7, 18)}
    assert np.all(depth[prim == NONE] == 1.0)
    o.close()


def _f16(x):
    return np.float16(np.float32(x))


def test_colour_attachment_conversions(ork_lib):
    """The draw image is R16G16B16A16_SFLOAT (round to nearest even on store), the presented image 8-bit UNORM
    (round(255 c) on the blit, src/vk_engine_run.cpp:159-161).  Fragment colour = ambient * vertex colour * base texel:
    ambient 0.25, the default texel of bindless slot 0 is (128, 128, 255) / 255."""
    cols = np.array([[0.3, 0.7, 0.9, 1.0]] * 6, np.float32)
    o = Ork(ork_lib, front(8, 8, 24, 24), cols)
    o.draw()
    hdr = o.r.readback(capi.TARGET_HDR16F)
    rgba = o.r.readback(capi.TARGET_RGBA8)
    base = np.array([128, 128, 255], np.float64) / 255.0
    want = (np.float32(0.25) * cols[0, :3].astype(np.float64) * base)
    px = hdr[12, 12]
    for c in range(3):                      # the f32 product rounded once to f16 (the products' own f32 roundings are below half an f16 ulp apart)
        assert abs(float(px[c]) - want[c]) <= abs(float(np.spacing(np.float16(want[c])))) * 0.51
        assert int(rgba[12, 12, c]) == int(np.floor(float(px[c]) * 255.0 + 0.5))
    assert px[3] == 1.0 and rgba[12, 12, 3] == 255
    assert np.all(rgba[40, 40] == (0, 0, 0, 255)) and np.all(hdr[40, 40] == (0, 0, 0, 1))     # clear colour (0, 0, 0, 1)
    o.close()


def test_additive_blend_sums_in_the_f16_attachment(ork_lib):
    """"Transparent" pipeline: colour blend SRC_ALPHA / ONE / ADD, alpha ONE / ZERO (src/vk_pipelines.cpp:108-118): every
    fragment adds src.rgb * src.a to what the RGBA16F attachment holds; with n layers of one constant colour the stored
    value stays within n half-ulps of n times that colour (one f16 rounding per fragment at most, one in all at least)."""
    tri = [(10.5, 10.5), (10.5, 30.5), (30.5, 10.5)]
    for n in (1, 2, 3, 5):
        o = Ork(ork_lib, [[p + (0.5,) for p in tri]] * n)
        o.fragment_counts()                   # draws the n triangles with the additive material
        hdr = o.r.readback(capi.TARGET_HDR16F).astype(np.float64)
        assert hdr[40, 40, 0] == 0.0
        unit = 0.25 * 128.0 / 255.0
        got = hdr[15, 15, 0]
        assert abs(got - n * unit) <= n * 0.5 * float(np.spacing(np.float16(n * unit))) + 1e-12
        assert hdr[15, 15, 3] == 1.0
        o.close()


def test_transparent_pass_tests_depth_but_does_not_write_it(ork_lib):
    """Transparent: depthTest on (LESS_OR_EQUAL), depthWrite off (src/vk_engine_init.cpp:545-547)."""
    tri = [(10.5, 10.5), (10.5, 30.5), (30.5, 10.5)]
    z = lambda zz: [p + (zz,) for p in tri]
    o = Ork(ork_lib, [z(0.5), z(0.75), z(0.25), z(0.25), z(0.5)])
    mk = lambda k, mat: mg.render_object(o.mesh, mat, np.eye(4), 3, first_index=3 * k)
    # opaque at 0.5; transparent ones behind (0.75: fails), in front (0.25, twice: both pass — no depth write between them), equal (0.5: passes)
    objs = np.array([mk(0, o.opaque), mk(1, o.additive), mk(2, o.additive), mk(3, o.additive), mk(4, o.additive)])
    o.r.draw(o.sd, objs, o.fp)
    depth = o.r.readback(capi.TARGET_DEPTH)
    hdr = o.r.readback(capi.TARGET_HDR16F).astype(np.float64)
    unit = 0.25 * 128.0 / 255.0
    assert depth[15, 15] == 0.5                                    # still the opaque surface's depth
    assert abs(hdr[15, 15, 0] - 4 * unit) < 4 * 0.5 * float(np.spacing(np.float16(4 * unit)))      # opaque + three transparent fragments
    o.close()


def test_shadow_pass_culls_front_faces_and_stores_the_plane_depth(ork_lib):
    """draw_shadow's pipeline: cull FRONT, front face counter-clockwise, depth-only, D32 cleared to 1.0
    (src/vk_engine_init.cpp:439-447, src/vk_engine_run.cpp:348-360).  Light matrix = the same pixel-space map onto the 64^2 shadow
    map: a triangle that is counter-clockwise on the light's screen writes nothing, a clockwise one its plane."""
    ccw = [(10.5, 10.5, 0.25), (10.5, 30.5, 0.25), (30.5, 10.5, 0.25)]
    cw = [(40.5, 10.5, 0.25), (60.5, 10.5, 0.5), (40.5, 30.5, 0.75)]
    o = Ork(ork_lib, [ccw, cw])
    view, proj = scenes.pixel_space(o.r)
    o.sd = capi.make_scene_data(view, proj, [proj] * 4, sun_dir=(0.0, 0.0, 1.0, 0.0), sun_color=(0, 0, 0, 1), ambient=(0.25, 0.25, 0.25, 1.0),
                                shadow_mode=0, enable_shadows=True)
    o.fp.enableShadows = 1
    o.fp.shadowMode = 0
    prim, _ = o.draw()
    assert len(pixels(prim, 0)) > 100 and len(pixels(prim, 1)) == 0          # the main pass culls the other one
    sh = o.r.readback(capi.TARGET_SHADOW)[0]
    assert np.all(sh[8:32, 8:32] == 1.0), "a front face (as the light sees it) must not be in the shadow map"
    ys, xs = np.nonzero(sh != 1.0)
    want_px = {p for p, ks in rule_raster([cw], front_only=False).items()}
    assert set(zip(xs.tolist(), ys.tolist())) == want_px and len(want_px) > 150
    want = 0.25 + (xs + 0.5 - 40.5) * (0.25 / 20.0) + (ys + 0.5 - 10.5) * (0.5 / 20.0)
    assert np.max(np.abs(sh[ys, xs].astype(np.float64) - want)) < 4e-7
    o.close()
