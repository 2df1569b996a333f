"""numpy procedural meshes, textures and camera/light matrices.

Used by the parity tests and by hand-constructed edge cases that talk to the C-ABI
directly.  (The product's scene layer is the C++ engine in csrc/host; bench.py
goes through that.)  Matrices here are math-convention (row, col) numpy arrays.
"""
from __future__ import annotations

import numpy as np

from .capi import VERTEX_DTYPE, RENDER_OBJECT_DTYPE


def _verts(pos, nrm, uv, col=None):
    v = np.zeros(len(pos), VERTEX_DTYPE)
    v["position"] = pos
    v["normal"] = nrm
    v["uv_x"] = uv[:, 0]
    v["uv_y"] = uv[:, 1]
    v["color"] = 1.0 if col is None else col
    return v


def cube(half=0.5):
    """24 vertices / 12 triangles, per-face normals, CCW front faces (seen from outside)."""
    faces = [  # normal, u axis, v axis
        ((1, 0, 0), (0, 0, -1), (0, 1, 0)), ((-1, 0, 0), (0, 0, 1), (0, 1, 0)),
        ((0, 1, 0), (1, 0, 0), (0, 0, -1)), ((0, -1, 0), (1, 0, 0), (0, 0, 1)),
        ((0, 0, 1), (1, 0, 0), (0, 1, 0)), ((0, 0, -1), (-1, 0, 0), (0, 1, 0)),
    ]
    pos, nrm, uv, idx = [], [], [], []
    for n, a, b in faces:
        n, a, b = np.array(n, np.float32), np.array(a, np.float32), np.array(b, np.float32)
        base = len(pos)
        for (s, t) in [(-1, -1), (1, -1), (1, 1), (-1, 1)]:
            pos.append((n + s * a + t * b) * half)
            nrm.append(n)
            uv.append(((s + 1) / 2, (1 - t) / 2))
        idx += [base, base + 1, base + 2, base, base + 2, base + 3]
    return _verts(np.array(pos, np.float32), np.array(nrm, np.float32), np.array(uv, np.float32)), np.array(idx, np.uint32)


def uv_sphere(rings=16, segments=16, radius=0.5):
    """(rings+1)(segments+1) vertices, 2*rings*segments triangles, smooth normals, equirect uv."""
    i, j = np.meshgrid(np.arange(rings + 1), np.arange(segments + 1), indexing="ij")
    theta = (i / rings * np.pi).astype(np.float32)
    phi = (j / segments * 2 * np.pi).astype(np.float32)
    n = np.stack([np.sin(theta) * np.cos(phi), np.cos(theta), np.sin(theta) * np.sin(phi)], -1).astype(np.float32)
    uv = np.stack([j / segments, i / rings], -1).astype(np.float32)
    pos = n * np.float32(radius)
    a = (i[:-1, :-1] * (segments + 1) + j[:-1, :-1]).reshape(-1)
    b = a + segments + 1
    # CCW seen from outside
    idx = np.stack([a, a + 1, b, a + 1, b + 1, b], -1).reshape(-1).astype(np.uint32)
    return _verts(pos.reshape(-1, 3), n.reshape(-1, 3), uv.reshape(-1, 2)), idx


def plane(nx=8, nz=8, size=10.0, y=0.0, uv_scale=1.0):
    """Grid of nx*nz quads on XZ facing +Y."""
    i, j = np.meshgrid(np.arange(nz + 1), np.arange(nx + 1), indexing="ij")
    x = (j / nx - 0.5) * size
    z = (i / nz - 0.5) * size
    pos = np.stack([x, np.full_like(x, y), z], -1).astype(np.float32)
    nrm = np.zeros_like(pos)
    nrm[..., 1] = 1
    uv = np.stack([j / nx * uv_scale, i / nz * uv_scale], -1).astype(np.float32)
    a = (i[:-1, :-1] * (nx + 1) + j[:-1, :-1]).reshape(-1)
    b = a + nx + 1
    idx = np.stack([a, b, a + 1, a + 1, b, b + 1], -1).reshape(-1).astype(np.uint32)
    return _verts(pos.reshape(-1, 3), nrm.reshape(-1, 3), uv.reshape(-1, 2)), idx


def triangles(points, colors=None):
    """Raw triangle soup from an (n,3,3) array of object-space corners (edge-case scenes)."""
    p = np.asarray(points, np.float32).reshape(-1, 3)
    e1 = p[1::3] - p[0::3]
    e2 = p[2::3] - p[0::3]
    n = np.cross(e1, e2)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    n = np.where(ln > 0, n / np.maximum(ln, 1e-30), np.array([[0, 0, 1]], np.float32)).astype(np.float32)
    nrm = np.repeat(n, 3, axis=0)
    uv = np.tile(np.array([[0, 0], [1, 0], [0, 1]], np.float32), (len(p) // 3, 1))
    return _verts(p, nrm, uv, colors), np.arange(len(p), dtype=np.uint32)


# ------------------------------------------------------------------ matrices (float32, glm semantics)
def perspective_rh_zo(fovy_rad, aspect, near, far):
    """glm::perspectiveRH_ZO (external/glm/glm/ext/matrix_clip_space.inl:233-248)."""
    f32 = np.float32
    t = f32(np.tan(f32(fovy_rad) / f32(2)))
    m = np.zeros((4, 4), np.float32)
    m[0, 0] = f32(1) / (f32(aspect) * t)
    m[1, 1] = f32(1) / t
    m[2, 2] = f32(far) / (f32(near) - f32(far))
    m[3, 2] = -f32(1)
    m[2, 3] = -(f32(far) * f32(near)) / (f32(far) - f32(near))
    return m


def camera_projection(aspect, fov_deg=60.0, near=0.1, far=100.0):
    """Camera::getProjectionMatrix (src/vk_camera.cpp:48-54): perspectiveRH_ZO with proj[1][1] *= -1."""
    m = perspective_rh_zo(np.radians(np.float32(fov_deg)), aspect, near, far)
    m[1, 1] *= -1
    return m


def look_at_rh(eye, center, up=(0, 1, 0)):
    """glm::lookAtRH (external/glm/glm/ext/matrix_transform.inl:98-120)."""
    eye, center, up = (np.asarray(a, np.float32) for a in (eye, center, up))
    f = center - eye
    f = f / np.linalg.norm(f)
    s = np.cross(f, up)
    s = s / np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4, dtype=np.float32)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[0, 3], m[1, 3], m[2, 3] = -np.dot(s, eye), -np.dot(u, eye), np.dot(f, eye)
    return m.astype(np.float32)


def ortho_rh_zo(l, r, b, t, n, f):
    """glm::orthoRH_ZO (external/glm/glm/ext/matrix_clip_space.inl:38-48)."""
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = 2 / (r - l)
    m[1, 1] = 2 / (t - b)
    m[2, 2] = -1 / (f - n)
    m[0, 3] = -(r + l) / (r - l)
    m[1, 3] = -(t + b) / (t - b)
    m[2, 3] = -n / (f - n)
    return m.astype(np.float32)


def camera_view(position, pitch=0.0, yaw=0.0):
    """Camera::getViewMatrix (src/vk_camera.cpp:33-46) = inverse(T * Ryaw * Rpitch)."""
    cp, sp, cy, sy = np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    rp = np.array([[1, 0, 0, 0], [0, cp, -sp, 0], [0, sp, cp, 0], [0, 0, 0, 1]], np.float64)
    ry = np.array([[cy, 0, sy, 0], [0, 1, 0, 0], [-sy, 0, cy, 0], [0, 0, 0, 1]], np.float64)
    t = np.eye(4)
    t[:3, 3] = position
    return np.linalg.inv(t @ ry @ rp).astype(np.float32)


def light_matrix(sun_dir=(0.5, -1.0, -0.5), near=0.1, far=100.0):
    """VulkanEngine::compute_light_matrix (src/vk_engine_run.cpp:482-491)."""
    d = np.asarray(sun_dir, np.float32)
    view = look_at_rh(-d * np.float32(60.0), (0, 0, 0))
    proj = ortho_rh_zo(-100.0, 100.0, -100.0, 100.0, near, far)
    proj[1, 1] *= -1
    return (proj @ view).astype(np.float32)


def translate(x, y, z):
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = (x, y, z)
    return m


def scale(s):
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = m[1, 1] = m[2, 2] = s
    return m


def render_object(mesh, material, transform, index_count, first_index=0):
    o = np.zeros((), RENDER_OBJECT_DTYPE)
    o["mesh"], o["material"] = mesh, material
    o["transform"] = np.asarray(transform, np.float32).T.reshape(-1)  # column-major
    o["indexCount"], o["firstIndex"] = index_count, first_index
    return o


# ------------------------------------------------------------------ textures
def checker_texture(size=64, cells=8, c0=(230, 230, 230, 255), c1=(40, 40, 200, 255), seed=0):
    """Checker + value noise, RGBA8."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:size, 0:size]
    m = (((x * cells) // size + (y * cells) // size) & 1).astype(bool)
    img = np.where(m[..., None], np.array(c1, np.int32), np.array(c0, np.int32))
    noise = rng.integers(-24, 25, (size, size, 1))
    img[..., :3] = np.clip(img[..., :3] + noise, 0, 255)
    return img.astype(np.uint8)


def cutout_texture(size=64, holes=4, seed=0):
    """RGBA8 with circular alpha holes (exercises `discard`)."""
    img = checker_texture(size, 4, (200, 180, 60, 255), (60, 160, 90, 255), seed)
    y, x = np.mgrid[0:size, 0:size]
    cell = size / holes
    cx = (x % cell) - cell / 2
    cy = (y % cell) - cell / 2
    img[..., 3] = np.where(cx * cx + cy * cy < (cell * 0.3) ** 2, 0, 255)
    return img


def sky_faces(size=64, seed=0):
    """Six RGBA8 faces (+X,-X,+Y,-Y,+Z,-Z) of a procedural sky: gradient + per-face tint + noise."""
    rng = np.random.default_rng(seed)
    tints = [(255, 180, 150), (150, 200, 255), (120, 160, 255), (90, 80, 70), (200, 255, 200), (255, 240, 160)]
    faces = []
    y, x = np.mgrid[0:size, 0:size]
    for t in tints:
        g = (0.55 + 0.45 * (1 - y / size))[..., None]
        img = np.clip(np.array(t)[None, None, :] * g + rng.integers(-10, 11, (size, size, 1)), 0, 255)
        stripe = ((x // max(1, size // 8)) & 1)[..., None] * 18
        img = np.clip(img - stripe, 0, 255)
        faces.append(np.concatenate([img, np.full((size, size, 1), 255)], -1).astype(np.uint8))
    return faces


def sky_levels_f16(size=32, levels=4, seed=0):
    """An HDR sky as a KTX file would carry it: levels[l][f], RGBA float16, level l = 2x2 box average of level l-1
    (the file's own mip chain; values reach ~6 around a "sun" on the +Z face so that tone mapping matters)."""
    base = [f.astype(np.float32)[..., :3] / 255.0 for f in sky_faces(size, seed)]
    y, x = np.mgrid[0:size, 0:size]
    sun = 6.0 * np.exp(-(((x - 0.6 * size) ** 2 + (y - 0.3 * size) ** 2) / (0.02 * size * size)))
    base[4] = base[4] + sun[..., None].astype(np.float32)
    out = []
    cur = [np.concatenate([b ** 2.2, np.ones((size, size, 1), np.float32)], -1) for b in base]
    for _ in range(levels):
        out.append([c.astype(np.float16) for c in cur])
        if cur[0].shape[0] == 1:
            break
        cur = [0.25 * (c[0::2, 0::2] + c[1::2, 0::2] + c[0::2, 1::2] + c[1::2, 1::2]) for c in cur]
    return out
