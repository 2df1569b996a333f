"""ctypes view of include/vkr.h.

`Renderer` binds one shared library exporting the vkr C-ABI.  The product binds
`libvkr.so` (prefix ``vkr_``); the test-suite binds the CPU oracle, which exports
the same entry points with prefix ``ork_`` (oracle/ork.h), through the same class so
that parity tests issue literally the same calls to both.
"""
from __future__ import annotations

import ctypes as C
import numpy as np

NUM_CASCADES = 4
MAX_BINDLESS = 4096

SUCCESS = 0
ERROR_INVALID_ARGUMENT, ERROR_OUT_OF_MEMORY, ERROR_CUDA, ERROR_NO_DEVICE, ERROR_CAPACITY = -1, -2, -3, -4, -5
FORMAT_UNORM, FORMAT_SRGB, FORMAT_RGBA16F = 0, 1, 2
SAMPLER_NEAREST, SAMPLER_LINEAR = 0, 1
PASS_MAIN_COLOR, PASS_TRANSPARENT, PASS_OTHER = 0, 1, 2
FRAG_PBR, FRAG_LAMBERT = 0, 1
TARGET_RGBA8, TARGET_HDR16F, TARGET_DEPTH, TARGET_SHADOW, TARGET_PRIMID, TARGET_RGBA8_GATHERED, TARGET_RGB8 = range(7)
COMM_ID_BYTES = 128
CONFIG_TIMINGS = 1
CONFIG_TIME_GEOM_KERNEL = 2
CONFIG_SUBMIT_THREAD = 4
CONFIG_TRANSIENT_DEPTH = 8
CONFIG_TRANSIENT_HDR = 16
CONFIG_LOCAL_PRESENT = 32
T_NAMES = ["upload", "shadow", "geometry", "shade", "transparent", "post", "total", "geom_kernel", "large_kernel", "clear",
           "shadow_exchange", "frame_gather", "readback"]

# struct Vertex (src/vk_types.h:85-91), 48 bytes
VERTEX_DTYPE = np.dtype(
    [("position", "<f4", 3), ("uv_x", "<f4"), ("normal", "<f4", 3), ("uv_y", "<f4"), ("color", "<f4", 4)]
)
assert VERTEX_DTYPE.itemsize == 48

# struct RenderObject with handles (include/vkr.h), 80 bytes
RENDER_OBJECT_DTYPE = np.dtype(
    [("mesh", "<u4"), ("material", "<u4"), ("transform", "<f4", 16), ("indexCount", "<u4"), ("firstIndex", "<u4")]
)
assert RENDER_OBJECT_DTYPE.itemsize == 80


class GPUSceneData(C.Structure):
    """struct GPUSceneData (src/vk_types.h:93-102) — 512 bytes, column-major matrices."""

    _fields_ = [
        ("view", C.c_float * 16),
        ("proj", C.c_float * 16),
        ("viewproj", C.c_float * 16),
        ("lightViewproj", (C.c_float * 16) * NUM_CASCADES),
        ("cascadeDistances", C.c_float * NUM_CASCADES),
        ("ambientColor", C.c_float * 4),
        ("sunlightDirection", C.c_float * 4),
        ("sunlightColor", C.c_float * 4),
    ]


assert C.sizeof(GPUSceneData) == 512


class MaterialConstants(C.Structure):
    """struct MaterialConstants (src/vk_materials.h:14-21) — 48 bytes."""

    _fields_ = [
        ("colorFactors", C.c_float * 4),
        ("metal_rough_factors", C.c_float * 4),
        ("albedoID", C.c_int32),
        ("normalID", C.c_int32),
        ("metalRoughID", C.c_int32),
        ("padding", C.c_int32),
    ]


assert C.sizeof(MaterialConstants) == 48


class ComputePushConstants(C.Structure):
    _fields_ = [("data1", C.c_float * 4), ("data2", C.c_float * 4), ("data3", C.c_float * 4), ("data4", C.c_float * 4)]


class FrameParams(C.Structure):
    _fields_ = [
        ("enableShadows", C.c_int32),
        ("shadowMode", C.c_int32),
        ("enableBackground", C.c_int32),
        ("enablePostprocess", C.c_int32),
        ("enableSkybox", C.c_int32),
        ("fragShader", C.c_int32),
        ("scissor", C.c_int32 * 4),
        ("background", ComputePushConstants),
        ("shadowScissor", C.c_int32 * 4),
    ]


class Config(C.Structure):
    _fields_ = [
        ("device", C.c_int32),
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("shadowExtent", C.c_uint32),
        ("stream", C.c_void_p),
        ("flags", C.c_uint32),
    ]


class EngineStats(C.Structure):
    _fields_ = [
        ("frametime", C.c_float),
        ("triangle_count", C.c_int32),
        ("drawcall_count", C.c_int32),
        ("mesh_draw_time", C.c_float),
    ]


PACK_STATIC_MAX_BYTES = 6 << 20     # VKR_PACK_STATIC_MAX_BYTES

# struct VkrSceneNode / VkrSceneSurface (include/vkr.h): the flattened scene graph vkr_scene_create takes
SCENE_NODE_DTYPE = np.dtype([("localTransform", "<f4", 16), ("parent", "<i4")])
SCENE_SURFACE_DTYPE = np.dtype([("node", "<u4"), ("mesh", "<u4"), ("material", "<u4"), ("indexCount", "<u4"), ("firstIndex", "<u4"),
                                ("boundsOrigin", "<f4", 3), ("boundsSphereRadius", "<f4")])
assert SCENE_NODE_DTYPE.itemsize == 68 and SCENE_SURFACE_DTYPE.itemsize == 36

# every symbol include/vkr.h declares (tests check the library exports them all)
VKR_SYMBOLS = [
    "create", "destroy", "upload_mesh", "create_image", "update_bindless_texture", "set_cubemap", "set_cubemap_levels",
    "build_material", "draw", "readback", "readback_async", "sync", "stats", "get_timings",
    "set_flags", "device_ptr", "launch_count", "upload_bytes", "cull_stats", "set_shadow_exchange", "comm_unique_id", "shard_init", "last_error", "version",
    "readback_band_async", "shard_info", "pack_rgb8_static", "host_times",
    "scene_create", "scene_set_transforms", "draw_scene", "scene_draw_list", "scene_destroy",
]


# VkrShadowExchangeFn(user, shadow device pointer, bytes per layer, layers, cudaStream_t) -> 0 on success
SHADOW_EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p)


class VkrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vkr error {code}: {msg}")
        self.code = code


def default_frame_params(**kw) -> FrameParams:
    fp = FrameParams()
    fp.background.data1[:] = (1.0, 0.0, 0.0, 1.0)  # src/vk_engine_init.cpp:504-505
    fp.background.data2[:] = (0.0, 0.0, 1.0, 1.0)
    for k, v in kw.items():
        if k == "scissor":
            fp.scissor[:] = v
        elif k == "shadowScissor":
            fp.shadowScissor[:] = v
        else:
            setattr(fp, k, v)
    return fp


def make_scene_data(view, proj, light_viewproj=None, cascade_distances=None, ambient=(0.1, 0.1, 0.1, 1.0),
                    sun_dir=(0.5, -1.0, -0.5, 0.0), sun_color=(1.0, 1.0, 1.0, 1.0), shadow_mode=0,
                    enable_shadows=False, viewproj=None) -> GPUSceneData:
    """Fill GPUSceneData the way VulkanEngine::draw does (src/vk_engine_run.cpp:107-125).
    Matrices are numpy (4,4) arrays in math (row, col) convention; stored column-major."""
    sd = GPUSceneData()
    view = np.asarray(view, np.float32)
    proj = np.asarray(proj, np.float32)
    vp = (proj @ view).astype(np.float32) if viewproj is None else np.asarray(viewproj, np.float32)
    sd.view[:] = view.T.reshape(-1)
    sd.proj[:] = proj.T.reshape(-1)
    sd.viewproj[:] = vp.T.reshape(-1)
    for i in range(NUM_CASCADES):
        m = np.eye(4, dtype=np.float32) if light_viewproj is None else np.asarray(light_viewproj[i], np.float32)
        sd.lightViewproj[i][:] = m.T.reshape(-1)
    cd = cascade_distances if cascade_distances is not None else (12.8186712, 26.6061382, 46.4039001, 100.0)
    sd.cascadeDistances[:] = cd
    sd.ambientColor[:] = ambient
    sd.sunlightDirection[:] = (*sun_dir[:3], float(shadow_mode))
    sd.sunlightColor[:] = (*sun_color[:3], 1.0 if enable_shadows else 0.0)
    return sd


class Renderer:
    """One VkrDevice behind the C-ABI of `lib` (prefix 'vkr' = product, 'ork' = oracle)."""

    def __init__(self, lib: C.CDLL, prefix: str, width: int, height: int, shadow_extent: int = 2048,
                 device: int = 0, stream: int | None = None, flags: int = 0):
        self.lib, self.prefix = lib, prefix
        self.width, self.height, self.shadow_extent = width, height, shadow_extent
        self.stream = stream             # the cudaStream_t frames are submitted on (None: the library's own)
        self._bind()
        cfg = Config(device, width, height, shadow_extent, stream, flags)
        self.dev = C.c_void_p()
        rc = self._f("create")(C.byref(cfg), C.byref(self.dev))
        if rc != SUCCESS:
            raise VkrError(rc, f"{prefix}_create failed (no CUDA device / bad config?)")
        self._keep = []

    # ---- plumbing
    def _f(self, name):
        return getattr(self.lib, f"{self.prefix}_{name}")

    def _bind(self):
        p = C.c_void_p
        sigs = {
            "create": ([C.POINTER(Config), C.POINTER(p)], C.c_int),
            "destroy": ([p], None),
            "upload_mesh": ([p, p, C.c_size_t, p, C.c_size_t, C.POINTER(C.c_uint32)], C.c_int),
            "create_image": ([p, p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(C.c_uint32)], C.c_int),
            "update_bindless_texture": ([p, C.c_int, C.c_uint32, C.c_int], C.c_int),
            "set_cubemap": ([p, C.POINTER(p), C.c_uint32, C.c_uint32, C.c_int, C.c_int], C.c_int),
            "set_cubemap_levels": ([p, C.POINTER(p), C.c_uint32, C.c_uint32, C.c_uint32, C.c_int], C.c_int),
            "build_material": ([p, C.POINTER(MaterialConstants), C.c_int, C.POINTER(C.c_uint32)], C.c_int),
            "draw": ([p, C.POINTER(GPUSceneData), p, C.c_size_t, C.POINTER(FrameParams)], C.c_int),
            "readback": ([p, C.c_int, p, C.c_size_t], C.c_int),
            "stats": ([p, C.POINTER(EngineStats)], C.c_int),
            "last_error": ([p], C.c_char_p),
        }
        if self.prefix == "vkr":
            sigs.update({
                "readback_async": ([p, C.c_int, p, C.c_size_t], C.c_int),
                "readback_band_async": ([p, C.c_int, p, C.c_size_t], C.c_int),
                "shard_info": ([p, C.POINTER(C.c_int32)], C.c_int),
                "sync": ([p], C.c_int),
                "get_timings": ([p, C.POINTER(C.c_float), C.c_int], C.c_int),
                "device_ptr": ([p, C.c_int, C.POINTER(p), C.POINTER(C.c_size_t)], C.c_int),
                "launch_count": ([p], C.c_int),
                "set_flags": ([p, C.c_uint32], C.c_int),
                "upload_bytes": ([p], C.c_size_t),
                "cull_stats": ([p, C.POINTER(C.c_uint32)], C.c_int),
                "set_shadow_exchange": ([p, SHADOW_EXCHANGE_FN, p], C.c_int),
                "comm_unique_id": ([p, C.c_size_t], C.c_int),
                "shard_init": ([p, C.c_int, C.c_int, p, C.c_size_t], C.c_int),
                "version": ([], C.c_char_p),
                "host_times": ([p, C.POINTER(C.c_float), C.c_int, C.c_int], C.c_int),
                "pack_rgb8_static": ([p, p, C.c_uint32, C.c_uint32, p, C.c_size_t, p, C.c_int], C.c_int),
            })
        if self.prefix == "emu":     # the test-only host emulation (tests/emu) exposes the sharding hooks too
            sigs.update({"set_shadow_exchange": ([p, SHADOW_EXCHANGE_FN, p], C.c_int), "host_ptr": ([p, C.c_int], C.c_void_p)})
        if self.prefix in ("vkr", "emu"):     # update_scene on the device
            sigs.update({
                "scene_create": ([p, p, C.c_size_t, p, C.c_size_t, C.POINTER(C.c_uint32)], C.c_int),
                "scene_set_transforms": ([p, C.c_uint32, C.c_uint32, C.c_uint32, p], C.c_int),
                "draw_scene": ([p, C.c_uint32, C.POINTER(GPUSceneData), C.POINTER(C.c_float), C.POINTER(FrameParams)], C.c_int),
                "scene_draw_list": ([p, C.c_uint32, p, C.c_size_t, C.POINTER(C.c_size_t)], C.c_int),
                "scene_destroy": ([p, C.c_uint32], C.c_int),
            })
        for name, (args, res) in sigs.items():
            try:
                fn = self._f(name)
            except AttributeError:
                if name in ("host_times", "pack_rgb8_static", "scene_create", "scene_set_transforms", "draw_scene", "scene_draw_list", "scene_destroy"):      # an older library variant in an A/B run (tools/run_ab.sh)
                    continue
                raise
            fn.argtypes, fn.restype = args, res

    def _check(self, rc):
        if rc != SUCCESS:
            msg = self._f("last_error")(self.dev)
            raise VkrError(rc, msg.decode() if msg else "")

    def close(self):
        if self.dev:
            self._f("destroy")(self.dev)
            self.dev = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- resources
    def upload_mesh(self, indices: np.ndarray, vertices: np.ndarray) -> int:
        idx = np.ascontiguousarray(indices, dtype=np.uint32)
        v = np.ascontiguousarray(vertices, dtype=VERTEX_DTYPE)
        out = C.c_uint32()
        self._check(self._f("upload_mesh")(self.dev, idx.ctypes.data, idx.size, v.ctypes.data, v.size, C.byref(out)))
        return out.value

    def create_image(self, rgba8: np.ndarray, fmt: int = FORMAT_SRGB, mipmapped: bool = True) -> int:
        px = np.ascontiguousarray(rgba8, dtype=np.uint8)
        h, w = px.shape[:2]
        out = C.c_uint32()
        self._check(self._f("create_image")(self.dev, px.ctypes.data, w, h, fmt, int(mipmapped), C.byref(out)))
        return out.value

    def update_bindless_texture(self, index: int, image: int, sampler: int = SAMPLER_LINEAR):
        self._check(self._f("update_bindless_texture")(self.dev, index, image, sampler))

    def set_cubemap(self, faces, fmt: int = FORMAT_SRGB, mipmapped: bool = True):
        fs = [np.ascontiguousarray(f, dtype=np.uint8) for f in faces]
        h, w = fs[0].shape[:2]
        arr = (C.c_void_p * 6)(*[f.ctypes.data for f in fs])
        self._check(self._f("set_cubemap")(self.dev, arr, w, h, fmt, int(mipmapped)))

    def set_cubemap_levels(self, levels, fmt: int):
        """levels[l][f]: face f (+X,-X,+Y,-Y,+Z,-Z) of mip level l, (h, w, 4) uint8 (UNORM / SRGB) or float16
        (FORMAT_RGBA16F): the cube map with the file's own mip chain (load_cubemap, src/vk_loader.cpp:521-558)."""
        dt = np.float16 if fmt == FORMAT_RGBA16F else np.uint8
        keep = [[np.ascontiguousarray(f, dtype=dt) for f in lv] for lv in levels]
        h, w = keep[0][0].shape[:2]
        arr = (C.c_void_p * (6 * len(keep)))(*[f.ctypes.data for lv in keep for f in lv])
        self._check(self._f("set_cubemap_levels")(self.dev, arr, w, h, len(keep), fmt))

    def build_material(self, color=(1, 1, 1, 1), metallic=1.0, roughness=1.0, albedo_id=0, normal_id=1,
                       metal_rough_id=0, pass_type=PASS_MAIN_COLOR) -> int:
        mc = MaterialConstants()
        mc.colorFactors[:] = color
        mc.metal_rough_factors[:] = (metallic, roughness, 0.0, 0.0)
        mc.albedoID, mc.normalID, mc.metalRoughID = albedo_id, normal_id, metal_rough_id
        out = C.c_uint32()
        self._check(self._f("build_material")(self.dev, C.byref(mc), pass_type, C.byref(out)))
        return out.value

    # ---- per frame
    def draw(self, scene: GPUSceneData, objects: np.ndarray, params: FrameParams):
        objs = np.ascontiguousarray(objects, dtype=RENDER_OBJECT_DTYPE)
        self._check(self._f("draw")(self.dev, C.byref(scene), objs.ctypes.data, objs.size, C.byref(params)))

    # ---- update_scene on the device (include/vkr.h: vkr_scene_create ...)
    def scene_create(self, nodes: np.ndarray, surfaces: np.ndarray) -> int:
        nd = np.ascontiguousarray(nodes, dtype=SCENE_NODE_DTYPE)
        sf = np.ascontiguousarray(surfaces, dtype=SCENE_SURFACE_DTYPE)
        out = C.c_uint32()
        self._check(self._f("scene_create")(self.dev, nd.ctypes.data, nd.size, sf.ctypes.data, sf.size, C.byref(out)))
        return out.value

    def scene_destroy(self, scene: int):
        self._check(self._f("scene_destroy")(self.dev, scene))

    def scene_set_transforms(self, scene: int, first_node: int, local_colmajor: np.ndarray):
        m = np.ascontiguousarray(local_colmajor, dtype=np.float32).reshape(-1, 16)
        self._check(self._f("scene_set_transforms")(self.dev, scene, first_node, m.shape[0], m.ctypes.data))

    def draw_scene(self, scene: int, scene_data: GPUSceneData, planes: np.ndarray, params: FrameParams):
        pl = np.ascontiguousarray(planes, dtype=np.float32).reshape(24)
        self._check(self._f("draw_scene")(self.dev, scene, C.byref(scene_data), pl.ctypes.data_as(C.POINTER(C.c_float)), C.byref(params)))

    def scene_draw_list(self, scene: int) -> np.ndarray:
        """The visible draws of the last draw_scene frame, as update_scene would have left them in _renderObjects."""
        n = C.c_size_t()
        self._check(self._f("scene_draw_list")(self.dev, scene, None, 0, C.byref(n)))
        out = np.zeros(n.value, RENDER_OBJECT_DTYPE)
        self._check(self._f("scene_draw_list")(self.dev, scene, out.ctypes.data, out.size, C.byref(n)))
        return out

    def readback(self, what: int) -> np.ndarray:
        W, H, S = self.width, self.height, self.shadow_extent
        if what in (TARGET_RGBA8, TARGET_RGBA8_GATHERED):
            out = np.empty((H, W, 4), np.uint8)
        elif what == TARGET_RGB8:
            out = np.empty((H, W, 3), np.uint8)
        elif what == TARGET_HDR16F:
            out = np.empty((H, W, 4), np.float16)
        elif what == TARGET_DEPTH:
            out = np.empty((H, W), np.float32)
        elif what == TARGET_SHADOW:
            out = np.empty((NUM_CASCADES, S, S), np.float32)
        elif what == TARGET_PRIMID:
            out = np.empty((H, W), np.uint32)
        else:
            raise ValueError(what)
        self._check(self._f("readback")(self.dev, what, out.ctypes.data, out.nbytes))
        return out

    def stats(self) -> EngineStats:
        st = EngineStats()
        self._check(self._f("stats")(self.dev, C.byref(st)))
        return st

    # ---- product-only
    def sync(self):
        self._check(self._f("sync")(self.dev))

    def readback_async(self, what: int, pinned_ptr: int, nbytes: int):
        self._check(self._f("readback_async")(self.dev, what, pinned_ptr, nbytes))

    def readback_band_async(self, what: int, frame_ptr: int, nbytes: int):
        """This rank's rows of the last frame into their place in the whole-frame pinned buffer at frame_ptr."""
        self._check(self._f("readback_band_async")(self.dev, what, frame_ptr, nbytes))

    def shard_info(self) -> dict:
        out = (C.c_int32 * 4)()
        self._check(self._f("shard_info")(self.dev, out))
        return {"rows": (out[0], out[1]), "peer_memory": bool(out[2]), "ranks": out[3]}

    def timings(self) -> dict:
        arr = (C.c_float * len(T_NAMES))()
        self._check(self._f("get_timings")(self.dev, arr, len(T_NAMES)))
        return dict(zip(T_NAMES, list(arr)))

    def set_flags(self, flags: int):
        self._check(self._f("set_flags")(self.dev, flags))

    def device_ptr(self, what: int):
        p, n = C.c_void_p(), C.c_size_t()
        self._check(self._f("device_ptr")(self.dev, what, C.byref(p), C.byref(n)))
        return p.value, n.value

    def pack_rgb8_static(self, rgba8_ptr: int, width: int, height: int, out: np.ndarray, stream: int = 0, wait: bool = True) -> np.ndarray:
        """vkr_pack_rgb8_static: the RGBA8 image at DEVICE address rgba8_ptr -> packed RGB8 in `out` (host array), staged in
        the library's static device array: no allocation (include/vkr.h; PACK_STATIC_MAX_BYTES is the limit)."""
        self._check(self._f("pack_rgb8_static")(self.dev, rgba8_ptr, width, height, out.ctypes.data, out.nbytes, stream, int(wait)))
        return out

    def host_times(self, reset: bool = False) -> dict:
        """Mean host microseconds per frame of vkr_draw's sections since the last reset (vkr_host_times)."""
        out = (C.c_float * 9)()
        self._check(self._f("host_times")(self.dev, out, 9, int(reset)))
        names = ("plan", "upload_expand", "shadow", "main_raster", "transparent_shade", "gather", "async_readbacks",
                 "caller_waiting_for_submit_thread", "submit_thread_waiting_for_device")
        return {k: float(v) for k, v in zip(names, out)}

    def launch_count(self) -> int:
        return self._f("launch_count")(self.dev)

    def set_shadow_exchange(self, fn):
        """fn(shadow_ptr, layer_bytes, layers, stream) -> None, called inside draw() after the shadow pass is enqueued."""
        if fn is None:
            self._shadow_cb = None
            self._check(self._f("set_shadow_exchange")(self.dev, C.cast(None, SHADOW_EXCHANGE_FN), None))
            return

        def tramp(user, shadow, layer_bytes, layers, stream):
            try:
                fn(shadow, layer_bytes, layers, stream)
                return 0
            except Exception as e:      # never let an exception cross the C boundary
                import sys
                print(f"shadow exchange hook failed: {e!r}", file=sys.stderr)
                return 1
        self._shadow_cb = SHADOW_EXCHANGE_FN(tramp)     # keep the trampoline alive
        self._check(self._f("set_shadow_exchange")(self.dev, self._shadow_cb, None))

    def comm_unique_ids(self) -> bytes:
        """Two communicator ids (rank 0 calls this and broadcasts the result)."""
        buf = C.create_string_buffer(2 * COMM_ID_BYTES)
        for i in range(2):
            rc = self._f("comm_unique_id")(C.byref(buf, i * COMM_ID_BYTES), COMM_ID_BYTES)
            if rc != SUCCESS:
                raise VkrError(rc, "vkr_comm_unique_id")
        return buf.raw

    def shard_init(self, rank: int, nranks: int, ids: bytes):
        self._check(self._f("shard_init")(self.dev, rank, nranks, ids, len(ids)))

    def host_ptr(self, what: int):
        """(emulation backend only) host address and size of a target."""
        return self._f("host_ptr")(self.dev, what), self.width * self.height * 4

    def cull_stats(self) -> dict:
        out = (C.c_uint32 * 4)()
        self._check(self._f("cull_stats")(self.dev, out))
        return {"main_clusters": out[0], "main_culled": out[1], "shadow_clusters": out[2], "shadow_culled": out[3]}

    def upload_bytes(self) -> int:
        return self._f("upload_bytes")(self.dev)
