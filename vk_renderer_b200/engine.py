"""ctypes binding of libvkr_engine.so — the C++ host engine (csrc/host) that mirrors VulkanEngine.

    eng = Engine(3840, 2160)            # CUDA path (libvkr); raises if it cannot be loaded
    eng.load_procedural(4)              # BASELINE config 4
    eng.update_scene(); eng.draw(); img = eng.readback(capi.TARGET_RGBA8)

`Engine(..., api_lib=<CDLL>, api_prefix="ork")` puts another implementation of the vkr C-ABI behind
the same engine; only the test-suite does that (with the CPU oracle).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi

PKG = os.path.dirname(os.path.abspath(__file__))
_engine_lib = None

_API_FIELDS = ["create", "destroy", "upload_mesh", "create_image", "update_bindless_texture", "set_cubemap",
               "build_material", "draw", "readback", "readback_async", "sync", "set_cubemap_levels",
               "scene_create", "scene_set_transforms", "draw_scene", "scene_draw_list", "stats", "scene_destroy"]


class VkrApi(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _API_FIELDS]


def load_engine() -> C.CDLL:
    global _engine_lib
    if _engine_lib is None:
        from . import load, LIBENGINE_PATH
        load()  # libvkr.so first (RTLD_GLOBAL): no fallback if missing
        if not os.path.exists(LIBENGINE_PATH):
            raise ImportError(f"{LIBENGINE_PATH} is missing: run __graft_entry__.build()")
        lib = C.CDLL(LIBENGINE_PATH)
        p = C.c_void_p
        lib.vke_create.argtypes = [C.c_uint32, C.c_uint32, C.c_int, p, C.c_uint32, C.POINTER(VkrApi), C.c_uint32]
        lib.vke_create.restype = p
        lib.vke_destroy.argtypes = [p]
        lib.vke_device.argtypes = [p]; lib.vke_device.restype = p
        lib.vke_load_procedural.argtypes = [p, C.c_int, C.c_uint32]; lib.vke_load_procedural.restype = C.c_uint64
        lib.vke_orbit_camera.argtypes = [p, C.c_uint32, C.c_int]
        lib.vke_set_camera.argtypes = [p] + [C.c_float] * 5
        lib.vke_set_light.argtypes = [p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.vke_set_toggles.argtypes = [p] + [C.c_int] * 5
        lib.vke_set_scissor.argtypes = [p] + [C.c_int] * 5
        lib.vke_update_scene.argtypes = [p, C.c_float]
        lib.vke_draw.argtypes = [p]; lib.vke_draw.restype = C.c_int
        lib.vke_present.argtypes = [p, p, C.c_size_t]; lib.vke_present.restype = C.c_int
        lib.vke_sync.argtypes = [p]; lib.vke_sync.restype = C.c_int
        lib.vke_readback.argtypes = [p, C.c_int, p, C.c_size_t]; lib.vke_readback.restype = C.c_int
        lib.vke_stats.argtypes = [p, C.POINTER(capi.EngineStats)]
        lib.vke_visible_objects.argtypes = [p]; lib.vke_visible_objects.restype = C.c_size_t
        lib.vke_scene_data.argtypes = [p, C.POINTER(capi.GPUSceneData)]
        lib.vke_camera_matrices.argtypes = [p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.vke_light_matrix.argtypes = [p, C.POINTER(C.c_float)]
        lib.vke_csm.argtypes = [p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.vke_compute_split.argtypes = [C.c_float, C.c_float, C.c_int]; lib.vke_compute_split.restype = C.c_float
        lib.vke_frustum_planes.argtypes = [p, C.POINTER(C.c_float)]
        lib.vke_is_visible.argtypes = [p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float]; lib.vke_is_visible.restype = C.c_int
        lib.vke_load_gltf.argtypes = [p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]; lib.vke_load_gltf.restype = C.c_int64
        lib.vke_load_cubemap.argtypes = [p, C.c_char_p, C.c_char_p, C.c_size_t]; lib.vke_load_cubemap.restype = C.c_int
        lib.vke_render_object.argtypes = [p, C.c_size_t, p, C.POINTER(capi.MaterialConstants), C.POINTER(C.c_int)]
        lib.vke_render_object.restype = C.c_int
        lib.vke_decode_png.argtypes = [C.c_char_p, C.c_size_t, p, C.c_size_t, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        lib.vke_decode_png.restype = C.c_int
        lib.vke_set_gpu_scene_update.argtypes = [p, C.c_int]
        lib.vke_scene_changed.argtypes = [p]
        lib.vke_set_node_transform.argtypes = [p, C.c_size_t, C.POINTER(C.c_float)]; lib.vke_set_node_transform.restype = C.c_int
        lib.vke_node_count.argtypes = [p]; lib.vke_node_count.restype = C.c_size_t
        lib.vke_gpu_draw_list.argtypes = [p, p, C.c_size_t]; lib.vke_gpu_draw_list.restype = C.c_int64
        lib.vke_decode_jpeg.argtypes = lib.vke_decode_png.argtypes
        lib.vke_decode_jpeg.restype = C.c_int
        _engine_lib = lib
    return _engine_lib


def decode_png(data: bytes):
    """The engine's PNG decoder (csrc/host/vk_gltf.cpp) -> (H, W, 4) uint8, or None when it refuses the stream."""
    lib = load_engine()
    w, h = C.c_uint32(), C.c_uint32()
    if lib.vke_decode_png(data, len(data), None, 0, C.byref(w), C.byref(h)) != 0:
        return None
    out = np.empty((h.value, w.value, 4), np.uint8)
    assert lib.vke_decode_png(data, len(data), out.ctypes.data, out.nbytes, C.byref(w), C.byref(h)) == 0
    return out


def decode_jpeg(data: bytes):
    """The engine's JPEG decoder (csrc/host/vk_jpeg.cpp) -> (H, W, 4) uint8, or None when it refuses the stream."""
    lib = load_engine()
    w, h = C.c_uint32(), C.c_uint32()
    if lib.vke_decode_jpeg(data, len(data), None, 0, C.byref(w), C.byref(h)) != 0:
        return None
    out = np.empty((h.value, w.value, 4), np.uint8)
    assert lib.vke_decode_jpeg(data, len(data), out.ctypes.data, out.nbytes, C.byref(w), C.byref(h)) == 0
    return out


class Engine:
    def __init__(self, width, height, device=0, stream=None, flags=0, api_lib=None, api_prefix="ork", shadow_extent=2048):
        self.lib = load_engine()
        self.width, self.height, self.shadow_extent = width, height, shadow_extent
        api = None
        if api_lib is not None:
            api = VkrApi()
            for n in _API_FIELDS:
                fn = getattr(api_lib, f"{api_prefix}_{n}", None)
                setattr(api, n, C.cast(fn, C.c_void_p).value if fn is not None else None)
        self._api = api
        self.e = self.lib.vke_create(width, height, device, stream, flags, C.byref(api) if api is not None else None, shadow_extent)
        if not self.e:
            raise RuntimeError("vke_create failed (no CUDA device?)")
        self.prefix = "vkr" if api_lib is None else api_prefix
        self.backend_lib = api_lib

    def close(self):
        if self.e:
            self.lib.vke_destroy(self.e)
            self.e = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_procedural(self, config, seed=None) -> int:
        seed = (0xB2000000 + config) if seed is None else seed
        n = self.lib.vke_load_procedural(self.e, config, seed)
        if n == 0:
            raise RuntimeError("load_procedural failed")
        return n

    def load_cubemap(self, path):
        """load_cubemap (src/vk_loader.cpp:521-558): a KTX 1.1 / 2.0 cube map becomes the skybox, mip chain included."""
        err = C.create_string_buffer(256)
        if self.lib.vke_load_cubemap(self.e, os.fsencode(path), err, len(err)) != 0:
            raise RuntimeError(f"load_cubemap({path}): {err.value.decode()}")

    def load_gltf(self, path, name="gltf") -> int:
        """load_gltf + _sceneRoot->addChild (src/vk_engine_init.cpp:646-660).  Returns the triangles added."""
        err = C.create_string_buffer(256)
        n = self.lib.vke_load_gltf(self.e, name.encode(), os.fsencode(path), err, len(err))
        if n < 0:
            raise RuntimeError(f"load_gltf({path}): {err.value.decode()}")
        return n

    def render_objects(self):
        """The draw list update_scene produced: [(VkrRenderObject record, MaterialConstants, pass type)]."""
        out = []
        for i in range(self.lib.vke_visible_objects(self.e)):
            ro = np.zeros((), capi.RENDER_OBJECT_DTYPE)
            mc, pt = capi.MaterialConstants(), C.c_int(0)
            assert self.lib.vke_render_object(self.e, i, ro.ctypes.data, C.byref(mc), C.byref(pt)) == 0
            out.append((ro, mc, pt.value))
        return out

    def set_gpu_scene_update(self, on=True):
        """update_scene's per-object work (refreshTransform, is_visible, the sort) on the device from the next frame on
        (vkr_scene_create / vkr_draw_scene); update_scene() then only moves the camera."""
        self.lib.vke_set_gpu_scene_update(self.e, int(on))

    def node_count(self) -> int:
        return self.lib.vke_node_count(self.e)

    def set_node_transform(self, index, local_colmajor):
        """Node::localTransform of the index-th node in traversal order."""
        m = (C.c_float * 16)(*[float(x) for x in local_colmajor])
        if self.lib.vke_set_node_transform(self.e, index, m) != 0:
            raise IndexError(index)

    def gpu_draw_list(self) -> np.ndarray:
        """The visible draws of the last frame drawn with set_gpu_scene_update(True), as the device left them."""
        n = self.lib.vke_gpu_draw_list(self.e, None, 0)
        if n < 0:
            raise RuntimeError("vke_gpu_draw_list failed")
        out = np.zeros(n, capi.RENDER_OBJECT_DTYPE)
        assert self.lib.vke_gpu_draw_list(self.e, out.ctypes.data, out.size) == n
        return out

    def set_camera(self, pos, pitch=0.0, yaw=0.0):
        self.lib.vke_set_camera(self.e, *map(float, pos), float(pitch), float(yaw))

    def orbit_camera(self, seed, view):
        self.lib.vke_orbit_camera(self.e, seed, view)

    def set_toggles(self, shadows=-1, shadow_mode=-1, background=-1, postprocess=-1, frag=-1):
        self.lib.vke_set_toggles(self.e, int(shadows), int(shadow_mode), int(background), int(postprocess), int(frag))

    def set_scissor(self, x, y, w, h, cull=False):
        self.lib.vke_set_scissor(self.e, x, y, w, h, int(cull))

    def update_scene(self, dt=0.0):
        self.lib.vke_update_scene(self.e, dt)

    def draw(self):
        rc = self.lib.vke_draw(self.e)
        if rc != 0:
            raise capi.VkrError(rc, "vke_draw")

    def present(self, pinned_ptr, nbytes):
        rc = self.lib.vke_present(self.e, pinned_ptr, nbytes)
        if rc != 0:
            raise capi.VkrError(rc, "vke_present")

    def sync(self):
        rc = self.lib.vke_sync(self.e)
        if rc != 0:
            raise capi.VkrError(rc, "vke_sync")

    def readback(self, what):
        W, H, S = self.width, self.height, self.shadow_extent
        shapes = {capi.TARGET_RGBA8: ((H, W, 4), np.uint8), capi.TARGET_HDR16F: ((H, W, 4), np.float16),
                  capi.TARGET_DEPTH: ((H, W), np.float32), capi.TARGET_SHADOW: ((4, S, S), np.float32),
                  capi.TARGET_PRIMID: ((H, W), np.uint32)}
        shape, dt = shapes[what]
        out = np.empty(shape, dt)
        rc = self.lib.vke_readback(self.e, what, out.ctypes.data, out.nbytes)
        if rc != 0:
            raise capi.VkrError(rc, "vke_readback")
        return out

    def stats(self):
        st = capi.EngineStats()
        self.lib.vke_stats(self.e, C.byref(st))
        return st

    def visible_objects(self):
        return self.lib.vke_visible_objects(self.e)

    def scene_data(self):
        sd = capi.GPUSceneData()
        self.lib.vke_scene_data(self.e, C.byref(sd))
        return sd

    def device(self):
        return self.lib.vke_device(self.e)

    # host-math probes
    def camera_matrices(self):
        v, p = (C.c_float * 16)(), (C.c_float * 16)()
        self.lib.vke_camera_matrices(self.e, v, p)
        return np.array(v, np.float32).reshape(4, 4).T, np.array(p, np.float32).reshape(4, 4).T

    def light_matrix(self):
        m = (C.c_float * 16)()
        self.lib.vke_light_matrix(self.e, m)
        return np.array(m, np.float32).reshape(4, 4).T

    def frustum_planes(self):
        out = (C.c_float * 24)()
        self.lib.vke_frustum_planes(self.e, out)
        return np.array(out, np.float32).reshape(6, 4)

    def is_visible(self, transform_colmajor, origin, radius) -> bool:
        t = (C.c_float * 16)(*transform_colmajor)
        o = (C.c_float * 3)(*origin)
        return bool(self.lib.vke_is_visible(self.e, t, o, float(radius)))

    def csm(self):
        m, d = (C.c_float * 64)(), (C.c_float * 4)()
        self.lib.vke_csm(self.e, m, d)
        return np.array(m, np.float32).reshape(4, 4, 4).transpose(0, 2, 1), np.array(d, np.float32)
