"""Seeded test scenes, built identically into any backend exposing the vkr C-ABI
(product `vkr_`, oracle `ork_`, host emulation `emu_`).

Each builder takes a Renderer and returns (scene_data, objects, frame_params).
"""
from __future__ import annotations

import numpy as np

from vk_renderer_b200 import capi, meshgen as mg


def _lights(n=1):
    base = mg.light_matrix()
    if n == 1:
        return [base] * 4
    out = []
    for i in range(4):  # four different ortho fits, stand-ins for the cascade matrices
        view = mg.look_at_rh(np.array([-0.5, 1.0, 0.5], np.float32) * 40, (0, 0, 0))
        half = 8.0 * (i + 1)
        proj = mg.ortho_rh_zo(-half, half, -half, half, 0.1, 100.0)
        out.append((proj @ view).astype(np.float32))
    return out


def basic(r, shadow_mode=1, shadows=True, post=True, skybox=True, frag=capi.FRAG_PBR, background=False):
    W, H = r.width, r.height
    sv, si = mg.uv_sphere(24, 24, 1.0)
    cv, ci = mg.cube(0.8)
    pv, pi = mg.plane(16, 16, 30.0, -1.0, uv_scale=8)
    ms, mc, mp = r.upload_mesh(si, sv), r.upload_mesh(ci, cv), r.upload_mesh(pi, pv)
    tex = r.create_image(mg.checker_texture(128, 8), capi.FORMAT_SRGB, True)
    r.update_bindless_texture(2, tex, capi.SAMPLER_LINEAR)
    tex2 = r.create_image(mg.checker_texture(96, 6, seed=3), capi.FORMAT_UNORM, True)   # npot, UNORM, NEAREST
    r.update_bindless_texture(5, tex2, capi.SAMPLER_NEAREST)
    if skybox == "hdr":     # the engine's real skybox asset is an RGBA16F KTX cube with its own mip chain (load_cubemap)
        r.set_cubemap_levels(mg.sky_levels_f16(32, 4), capi.FORMAT_RGBA16F)
    elif skybox == "levels8":
        r.set_cubemap_levels([[f.astype(np.float32) * 255.0 / max(1.0, float(f.max())) for f in lv] for lv in mg.sky_levels_f16(32, 6)],
                             capi.FORMAT_SRGB)
    elif skybox:
        r.set_cubemap(mg.sky_faces(64))
    m_red = r.build_material((1, 0.3, 0.3, 1), 0.0, 0.6)
    m_metal = r.build_material((0.9, 0.8, 0.4, 1), 1.0, 0.5)
    m_tex = r.build_material((1, 1, 1, 1), 0.0, 1.0, albedo_id=2)
    m_tex2 = r.build_material((1, 1, 1, 1), 0.5, 0.7, albedo_id=5, metal_rough_id=2)
    objs = [mg.render_object(mp, m_tex, np.eye(4), len(pi))]
    for i in range(5):
        objs.append(mg.render_object(ms, m_red if i % 2 else m_metal, mg.translate(-6 + 3 * i, 0, 0), len(si)))
    objs.append(mg.render_object(mc, m_tex2, mg.translate(0, 0.5, 4) @ mg.scale(1.5), len(ci)))
    view = mg.camera_view((0, 3, 12), pitch=-0.2)
    proj = mg.camera_projection(W / H)
    lights = _lights(4 if shadow_mode == 3 else 1)
    sd = capi.make_scene_data(view, proj, lights, shadow_mode=shadow_mode, enable_shadows=shadows,
                              cascade_distances=(6.0, 12.0, 20.0, 100.0))
    fp = capi.default_frame_params(enableShadows=int(shadows), shadowMode=shadow_mode, enablePostprocess=int(post),
                                   enableSkybox=int(bool(skybox)), fragShader=frag, enableBackground=int(background))
    return sd, np.array(objs), fp


def orbit_views(r, n_views, shadow_mode=1, radius=12.0, height=3.0):
    """n camera views on an orbit around the `basic` scene (BASELINE config 5 in miniature): one GPUSceneData per view,
    same light for all."""
    W, H = r.width, r.height
    lights = _lights(4 if shadow_mode == 3 else 1)
    proj = mg.camera_projection(W / H)
    out = []
    for k in range(n_views):
        yaw = 2.0 * np.pi * k / n_views
        pos = (radius * np.sin(yaw), height, radius * np.cos(yaw))
        view = mg.camera_view(pos, pitch=-0.2, yaw=yaw)        # forward = (-sin yaw, ., -cos yaw): towards the centre
        out.append(capi.make_scene_data(view, proj, lights, shadow_mode=shadow_mode, enable_shadows=True,
                                        cascade_distances=(6.0, 12.0, 20.0, 100.0)))
    return out


def pixel_space(r):
    """view = I and a projection that maps object (x, y) in pixels straight to window coordinates,
    z in [0,1] straight to depth — hand-placed raster edge cases."""
    W, H = r.width, r.height
    proj = np.eye(4, dtype=np.float32)
    proj[0, 0], proj[0, 3] = 2.0 / W, -1.0
    proj[1, 1], proj[1, 3] = 2.0 / H, -1.0
    return np.eye(4, dtype=np.float32), proj


def edge_cases(r):
    """Shared edges through pixel centres, top-left ties, sub-pixel slivers, degenerates, coplanar
    duplicates (equal depth: the later primitive must win), both windings."""
    view, proj = pixel_space(r)
    T = []
    z = 0.5

    def tri(a, b, c, zz=z):
        T.append([(a[0], a[1], zz), (b[0], b[1], zz), (c[0], c[1], zz)])

    # CCW-on-screen (front-facing) = clockwise in y-down maths coordinates: (x0,y0),(x0,y1),(x1,y0)
    # two triangles sharing a diagonal that passes exactly through pixel centres
    tri((2.5, 2.5), (2.5, 10.5), (10.5, 2.5))
    tri((10.5, 2.5), (2.5, 10.5), (10.5, 10.5))
    # quad fan sharing a vertex on a pixel centre
    c = (20.5, 8.5)
    ring = [(14.5, 2.5), (14.5, 14.5), (26.5, 14.5), (26.5, 2.5)]
    for i in range(4):
        tri(c, ring[i], ring[(i + 1) % 4])
    # axis-aligned rectangle with integer (pixel-corner) edges and one with centre-aligned edges
    tri((30, 2), (30, 9), (37, 2)); tri((37, 2), (30, 9), (37, 9))
    tri((40.5, 2.5), (40.5, 9.5), (47.5, 2.5)); tri((47.5, 2.5), (40.5, 9.5), (47.5, 9.5))
    # slivers thinner than 1/256 px, and exactly degenerate triangles
    tri((2.0, 20.0), (2.0, 20.003), (30.0, 20.0))
    tri((2.0, 22.5), (30.0, 22.5), (2.0, 22.5))
    tri((5.5, 24.5), (5.5, 24.5), (5.5, 24.5))
    # tiny triangles around one pixel centre
    tri((10.4, 26.4), (10.4, 26.7), (10.7, 26.4))
    tri((12.51, 26.51), (12.51, 26.9), (12.9, 26.51))
    # back-facing (must be culled)
    tri((50.5, 2.5), (58.5, 2.5), (50.5, 10.5))
    # coplanar duplicates at the same depth: later wins; and a nearer one drawn first
    tri((2.5, 30.5), (2.5, 40.5), (12.5, 30.5), 0.25)
    tri((2.5, 30.5), (2.5, 40.5), (12.5, 30.5), 0.25)
    tri((4.5, 30.5), (4.5, 44.5), (18.5, 30.5), 0.75)
    tri((4.5, 30.5), (4.5, 44.5), (18.5, 30.5), 0.125)
    # depth exactly 0 and exactly 1
    tri((30.5, 30.5), (30.5, 40.5), (40.5, 30.5), 0.0)
    tri((42.5, 30.5), (42.5, 40.5), (52.5, 30.5), 1.0)
    # a sloped plane
    T.append([(20.5, 44.5, 0.1), (20.5, 60.5, 0.9), (60.5, 44.5, 0.4)])
    pts = np.array(T, np.float32)
    cols = np.repeat(np.random.default_rng(1).uniform(0.3, 1.0, (len(T), 1, 4)), 3, axis=1).reshape(-1, 4).astype(np.float32)
    v, i = mg.triangles(pts, cols)
    mesh = r.upload_mesh(i, v)
    mat = r.build_material((1, 1, 1, 1), 0.0, 0.8)
    objs = np.array([mg.render_object(mesh, mat, np.eye(4), len(i))])
    sd = capi.make_scene_data(view, proj, None, sun_dir=(0.3, 0.2, 1.0, 0.0), ambient=(0.2, 0.2, 0.2, 1.0))
    fp = capi.default_frame_params(fragShader=capi.FRAG_LAMBERT)
    return sd, objs, fp


def clipping(r, shadow_mode=0):
    """Geometry crossing the near plane, the far plane and the guard band; camera inside a big box."""
    W, H = r.width, r.height
    pv, pi = mg.plane(6, 6, 400.0, -1.5, uv_scale=40)      # huge ground through the camera
    cv, ci = mg.cube(0.5)
    sv, si = mg.uv_sphere(12, 16, 1.0)
    mp, mc, ms = r.upload_mesh(pi, pv), r.upload_mesh(ci, cv), r.upload_mesh(si, sv)
    tex = r.create_image(mg.checker_texture(64, 8, seed=7), capi.FORMAT_SRGB, True)
    r.update_bindless_texture(2, tex, capi.SAMPLER_LINEAR)
    m_tex = r.build_material((1, 1, 1, 1), 0.1, 0.9, albedo_id=2)
    m_a = r.build_material((0.8, 0.9, 0.3, 1), 0.3, 0.4)
    objs = [mg.render_object(mp, m_tex, np.eye(4), len(pi))]
    objs.append(mg.render_object(mc, m_a, mg.translate(0.2, 0, 0.3) @ mg.scale(1.2), len(ci)))     # camera inside
    objs.append(mg.render_object(ms, m_a, mg.translate(0.0, 0.0, -0.9), len(si)))                   # crosses near plane
    objs.append(mg.render_object(ms, m_a, mg.translate(3.0, 0.0, -99.6), len(si)))                  # crosses far plane
    objs.append(mg.render_object(mc, m_a, mg.translate(-2.0, 0.5, -4.0) @ mg.scale(2.0), len(ci)))
    # a needle that reaches far outside the guard band
    needle = np.array([[(-0.05, -0.2, -0.15), (0.05, -0.2, -0.15), (3000.0, 900.0, -5.0)]], np.float32)
    nv, ni = mg.triangles(needle)
    mn = r.upload_mesh(ni, nv)
    objs.append(mg.render_object(mn, m_a, np.eye(4), len(ni)))
    needle2 = np.array([[(0.05, 0.2, -0.15), (-0.05, 0.2, -0.15), (-3000.0, 900.0, -5.0)]], np.float32)
    nv2, ni2 = mg.triangles(needle2)
    objs.append(mg.render_object(r.upload_mesh(ni2, nv2), m_a, np.eye(4), len(ni2)))
    view = mg.camera_view((0, 0, 0), pitch=-0.1, yaw=0.3)
    proj = mg.camera_projection(W / H)
    sd = capi.make_scene_data(view, proj, _lights(1), shadow_mode=shadow_mode, enable_shadows=True)
    fp = capi.default_frame_params(enableShadows=1, shadowMode=shadow_mode, enablePostprocess=1)
    return sd, np.array(objs), fp


def soup(r, n=20000, seed=5):
    """Random small triangles with random depths in pixel space: stresses the atomic visibility path
    (many fragments per pixel, frequent exact-depth ties broken by submission order)."""
    view, proj = pixel_space(r)
    rng = np.random.default_rng(seed)
    W, H = r.width, r.height
    c = rng.uniform((-4, -4), (W + 4, H + 4), (n, 1, 2))
    off = rng.normal(0, 2.5, (n, 3, 2))
    # snap a third of them to quarter pixels so ties / centre hits are common
    xy = c + off
    xy[: n // 3] = np.round(xy[: n // 3] * 4) / 4
    zq = rng.integers(0, 8, (n, 1, 1)) / 8.0            # few distinct depths -> many ties
    zz = np.broadcast_to(zq, (n, 3, 1)).copy()
    zz[n // 2:] += rng.uniform(0, 0.1, (n - n // 2, 3, 1))
    pts = np.concatenate([xy, zz], -1).astype(np.float32)
    # make them all front-facing
    e1 = pts[:, 1, :2] - pts[:, 0, :2]
    e2 = pts[:, 2, :2] - pts[:, 0, :2]
    flip = (e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]) > 0
    pts[flip] = pts[flip][:, [0, 2, 1]]
    cols = np.repeat(rng.uniform(0.2, 1.0, (n, 1, 4)), 3, axis=1).reshape(-1, 4).astype(np.float32)
    v, i = mg.triangles(pts, cols)
    mesh = r.upload_mesh(i, v)
    mats = [r.build_material(tuple(rng.uniform(0.3, 1, 3)) + (1,), 0.0, 0.7) for _ in range(3)]
    third = (n // 3) * 3
    objs = [mg.render_object(mesh, mats[k], np.eye(4), third, first_index=k * third) for k in range(3)]
    sd = capi.make_scene_data(view, proj, None, sun_dir=(0.3, 0.2, 1.0, 0.0), ambient=(0.2, 0.2, 0.2, 1.0))
    fp = capi.default_frame_params(fragShader=capi.FRAG_LAMBERT)
    return sd, np.array(objs), fp


def big_triangles(r):
    """A few screen-filling triangles at various depths (the large-triangle path) over small ones."""
    view, proj = pixel_space(r)
    W, H = r.width, r.height
    T = [
        [(-50, -50, 0.9), (-50, 3 * H, 0.9), (3 * W, -50, 0.9)],
        [(0.5, 0.5, 0.7), (0.5, H - 0.5, 0.2), (W - 0.5, 0.5, 0.8)],
        [(W - 0.5, 0.5, 0.3), (0.5, H - 0.5, 0.6), (W - 0.5, H - 0.5, 0.1)],
        [(W * 0.2, H * 0.1, 0.5), (W * 0.25, H * 0.9, 0.5), (W * 0.8, H * 0.5, 0.05)],
    ]
    pts = np.array(T, np.float32)
    cols = np.repeat(np.array([[1, .5, .5, 1], [.5, 1, .5, 1], [.5, .5, 1, 1], [1, 1, .4, 1]], np.float32), 3, axis=0)
    v, i = mg.triangles(pts, cols)
    mesh = r.upload_mesh(i, v)
    mat = r.build_material((1, 1, 1, 1), 0.0, 0.8)
    sd0, objs0, _ = soup(r, 3000, seed=9)
    objs = np.concatenate([np.array([mg.render_object(mesh, mat, np.eye(4), len(i))]), objs0])
    fp = capi.default_frame_params(fragShader=capi.FRAG_LAMBERT, enableBackground=1)
    return sd0, objs, fp


def cutout(r):
    """Alpha-tested geometry (`discard` when albedo alpha < 0.5): small triangles (tessellated drape), a
    screen-sized quad (large-triangle path), a drape crossing the near plane, all in front of opaque
    geometry that must show through the holes and receive the correct depth."""
    W, H = r.width, r.height
    tex = r.create_image(mg.cutout_texture(128, 4, seed=2), capi.FORMAT_SRGB, True)
    r.update_bindless_texture(3, tex, capi.SAMPLER_LINEAR)
    tex2 = r.create_image(mg.cutout_texture(32, 2, seed=5), capi.FORMAT_UNORM, False)     # single level, NEAREST
    r.update_bindless_texture(4, tex2, capi.SAMPLER_NEAREST)
    back = r.create_image(mg.checker_texture(64, 8, seed=1), capi.FORMAT_SRGB, True)
    r.update_bindless_texture(2, back, capi.SAMPLER_LINEAR)
    m_cut = r.build_material((1, 1, 1, 1), 0.0, 0.9, albedo_id=3)
    m_cut2 = r.build_material((1, 0.8, 0.8, 1), 0.2, 0.7, albedo_id=4)
    m_back = r.build_material((1, 1, 1, 1), 0.0, 1.0, albedo_id=2)
    m_sph = r.build_material((0.9, 0.5, 0.2, 1), 0.0, 0.5)
    pv, pi = mg.plane(24, 24, 6.0, 0.0, uv_scale=3)          # tessellated: small triangles
    qv, qi = mg.plane(1, 1, 14.0, 0.0, uv_scale=4)           # two big triangles
    gv, gi = mg.plane(4, 4, 40.0, -2.0, uv_scale=10)
    sv, si = mg.uv_sphere(16, 16, 1.0)
    mp, mq, mgr, ms = r.upload_mesh(pi, pv), r.upload_mesh(qi, qv), r.upload_mesh(gi, gv), r.upload_mesh(si, sv)
    rot = np.array([[1, 0, 0, 0], [0, 0, -1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], np.float32)     # +Y normal -> +Z (faces the camera)
    objs = [
        mg.render_object(mgr, m_back, np.eye(4), len(gi)),
        mg.render_object(ms, m_sph, mg.translate(-1.5, 0.0, -1.0), len(si)),
        mg.render_object(ms, m_sph, mg.translate(2.0, 0.5, -3.0), len(si)),
        mg.render_object(mq, m_cut, mg.translate(0, 1.0, -6.0) @ rot, len(qi)),            # big, far
        mg.render_object(mp, m_cut, mg.translate(-1.0, 0.5, 1.5) @ rot, len(pi)),          # tessellated, near
        mg.render_object(mp, m_cut2, mg.translate(2.5, 0.0, 2.0) @ rot, len(pi)),
        mg.render_object(mq, m_cut2, mg.translate(0.0, 0.0, 7.95) @ rot, len(qi)),         # crosses the near plane
    ]
    view = mg.camera_view((0, 1.0, 8.0), pitch=-0.05)
    proj = mg.camera_projection(W / H)
    sd = capi.make_scene_data(view, proj, _lights(1), shadow_mode=1, enable_shadows=True)
    fp = capi.default_frame_params(enableShadows=1, shadowMode=1, enablePostprocess=1)
    return sd, np.array(objs), fp


def lattice(r, n=4, rings=24, shadow_mode=1):
    """n^3 instanced spheres seen head-on, submitted back to front (the worst order for the depth test):
    most of them hidden, partly off-screen — exercises instancing, the front-to-back processing order
    and the cluster frustum / occlusion culling, which must not change a single pixel."""
    W, H = r.width, r.height
    sv, si = mg.uv_sphere(rings, 2 * rings, 1.0)
    mesh = r.upload_mesh(si, sv)
    rng = np.random.default_rng(11)
    mats = [r.build_material(tuple(rng.uniform(0.2, 1.0, 3)) + (1.0,), float(rng.random() < 0.3), float(rng.uniform(0.2, 0.9))) for _ in range(5)]
    objs = []
    k = 0
    for iz in range(n):                      # far layers first
        for iy in range(n):
            for ix in range(n):
                t = mg.translate((ix - (n - 1) / 2) * 2.6, (iy - (n - 1) / 2) * 2.6, (iz - (n - 1) / 2) * 2.6)
                objs.append(mg.render_object(mesh, mats[k % 5], t, len(si)))
                k += 1
    view = mg.camera_view((0.4, 0.8, 3.0 + 1.3 * n), pitch=-0.08)      # close: the outer columns leave the frustum
    proj = mg.camera_projection(W / H)
    sd = capi.make_scene_data(view, proj, _lights(4 if shadow_mode == 3 else 1), shadow_mode=shadow_mode, enable_shadows=True,
                              cascade_distances=(6.0, 12.0, 20.0, 100.0))
    fp = capi.default_frame_params(enableShadows=1, shadowMode=shadow_mode, enablePostprocess=1)
    return sd, np.array(objs), fp


def transparent(r, skybox=True):
    """"Transparent" materials (additive blend, depth test, no depth write — src/vk_engine_init.cpp:545-547):
    overlapping tessellated spheres, a screen-sized quad (large-triangle path), one behind opaque geometry
    (must be hidden), one textured with cut-outs (discard inside the transparent pass), over opaque geometry
    and over the background / skybox (where the skybox, drawn last at depth 1.0, wins)."""
    sd, objs, fp = basic(r, 1, skybox=skybox, background=not skybox)
    W, H = r.width, r.height
    sv, si = mg.uv_sphere(20, 20, 1.0)
    qv, qi = mg.plane(1, 1, 9.0, 0.0, uv_scale=3)
    ms, mq = r.upload_mesh(si, sv), r.upload_mesh(qi, qv)
    cut = r.create_image(mg.cutout_texture(64, 3, seed=7), capi.FORMAT_SRGB, True)
    r.update_bindless_texture(7, cut, capi.SAMPLER_LINEAR)
    t_blue = r.build_material((0.1, 0.2, 0.6, 1), 0.0, 0.4, pass_type=capi.PASS_TRANSPARENT)
    t_warm = r.build_material((0.5, 0.3, 0.1, 1), 0.8, 0.3, pass_type=capi.PASS_TRANSPARENT)
    t_cut = r.build_material((0.3, 0.6, 0.3, 1), 0.0, 0.7, albedo_id=7, pass_type=capi.PASS_TRANSPARENT)
    rot = np.array([[1, 0, 0, 0], [0, 0, -1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], np.float32)
    extra = [
        mg.render_object(ms, t_blue, mg.translate(-1.5, 1.0, 3.0) @ mg.scale(1.4), len(si)),
        mg.render_object(ms, t_warm, mg.translate(-0.5, 1.2, 4.0) @ mg.scale(1.2), len(si)),      # overlaps the first
        mg.render_object(ms, t_blue, mg.translate(3.0, 0.0, -4.0), len(si)),                      # behind an opaque sphere row
        mg.render_object(mq, t_warm, mg.translate(0.0, 2.0, 1.0) @ rot, len(qi)),                 # big quad: large-triangle path
        mg.render_object(mq, t_cut, mg.translate(2.0, 1.0, 6.0) @ rot @ mg.scale(0.4), len(qi)),
        mg.render_object(ms, t_cut, mg.translate(5.0, 3.5, 2.0), len(si)),                        # against the sky / background
    ]
    return sd, np.concatenate([objs, np.array(extra)]), fp


SCENES = {
    "cutout": cutout,
    "transparent": transparent,
    "transparent_background": lambda r: transparent(r, skybox=False),
    "lattice": lattice,
    "basic_hard": lambda r: basic(r, 0),
    "basic_pcf": lambda r: basic(r, 1),
    "basic_pcss": lambda r: basic(r, 2),
    "basic_csm": lambda r: basic(r, 3),
    "basic_lambert_noshadow": lambda r: basic(r, 0, shadows=False, post=False, skybox=False, frag=capi.FRAG_LAMBERT),
    # mesh.frag evaluates the same four shadow modes before its Lambert term (shaders/mesh.frag:124-182)
    "basic_lambert_pcf": lambda r: basic(r, 1, frag=capi.FRAG_LAMBERT),
    "basic_lambert_csm": lambda r: basic(r, 3, post=False, skybox=False, background=True, frag=capi.FRAG_LAMBERT),
    "basic_hdr_sky": lambda r: basic(r, 1, skybox="hdr"),
    "basic_sky_levels8": lambda r: basic(r, 1, skybox="levels8", post=False),
    "basic_background": lambda r: basic(r, 1, skybox=False, background=True, post=False),
    "edge_cases": edge_cases,
    "clipping": clipping,
    "soup": soup,
    "big_triangles": big_triangles,
}
