"""KTX cube maps for the skybox (VulkanEngine::load_cubemap, src/vk_loader.cpp:521-558).

The reference gives the file to libktx and uploads every face and mip level it holds; here the containers are decoded by
csrc/host/vk_ktx.cpp and handed to vkr_set_cubemap_levels.  CPU: the decoder against independently written files (both
container versions, three texel formats, other-endian KTX1, malformed input), and a file loaded through the engine against
the same texels uploaded directly, on the oracle and on the kernels' host emulation."""
import ctypes as C
import os

import numpy as np
import pytest

from vk_renderer_b200 import capi, meshgen as mg
from vk_renderer_b200.engine import Engine, load_engine
import ktx_util
import parity


def decode(data: bytes):
    lib = load_engine()
    lib.vke_decode_ktx.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_uint32), C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t]
    lib.vke_decode_ktx.restype = C.c_int
    info = (C.c_uint32 * 5)()
    err = C.create_string_buffer(256)
    out = np.zeros(max(len(data), 1 << 16), np.uint8)      # (a supercompressed file is smaller than its texels)
    rc = lib.vke_decode_ktx(data, len(data), info, out.ctypes.data, out.nbytes, err, len(err))
    if rc != 0:
        raise ValueError(err.value.decode())
    return list(info), out


def levels_of(kind, size=16, n=3):
    lv = mg.sky_levels_f16(size, n, seed=5)
    if kind == "f16":
        return lv
    return [[np.clip(f.astype(np.float32) * 200.0, 0, 255).astype(np.uint8) for f in l] for l in lv]


def flat(levels, kind):
    return np.frombuffer(b"".join(ktx_util._bytes(f, kind) for l in levels for f in l), np.uint8)


@pytest.mark.parametrize("kind", ["unorm", "srgb", "f16"])
@pytest.mark.parametrize("container", ["ktx1", "ktx1_big_endian", "ktx1_key_values", "ktx2", "ktx2_zlib", "ktx2_zstd"])
def test_decoder_returns_the_files_texels(kind, container):
    lv = levels_of(kind)
    if container == "ktx2_zstd" and ktx_util.zstd_compress is None:
        pytest.skip("no libzstd.so.1 on this machine (the loader then refuses such files with that reason)")
    data = {"ktx1": lambda: ktx_util.ktx1_cube(lv, kind),
            "ktx1_big_endian": lambda: ktx_util.ktx1_cube(lv, kind, big_endian=True),
            "ktx1_key_values": lambda: ktx_util.ktx1_cube(lv, kind, key_values=b"KTXorientation\0S=r,T=d\0"),
            "ktx2": lambda: ktx_util.ktx2_cube(lv, kind),
            "ktx2_zlib": lambda: ktx_util.ktx2_cube(lv, kind, scheme=3),
            "ktx2_zstd": lambda: ktx_util.ktx2_cube(lv, kind, scheme=2)}[container]()
    info, out = decode(data)
    want = flat(lv, kind)
    assert info[:3] == [16, 16, 3] and info[3] == {"unorm": 0, "srgb": 1, "f16": 2}[kind] and info[4] == 0
    assert np.array_equal(out[:len(want)], want)


def test_decoder_refuses_what_it_cannot_read():
    lv = levels_of("f16")
    good = ktx_util.ktx1_cube(lv, "f16")
    for bad, why in [(good[:40], "short"), (b"\x89PNG" + good[4:], "identifier"), (good[:-20], "truncated"),
                     (good[:28] + (0x8051).to_bytes(4, "little") + good[32:], "format"),          # GL_RGB8
                     (good[:52] + (1).to_bytes(4, "little") + good[56:], "cube"),                 # one face: a 2-D texture
                     (good[:56] + (9).to_bytes(4, "little") + good[60:], "levels")]:              # 16x16 has five levels
        with pytest.raises(ValueError):
            decode(bad)
    z = ktx_util.ktx2_cube(lv, "f16", scheme=3)
    zs = ktx_util.ktx2_cube(lv, "f16", scheme=2)
    for bad in (ktx_util.ktx2_cube(lv, "f16", scheme=1),                                             # BasisLZ
                z[:-9] + bytes(9), zs[:-9] + bytes(9)):                                               # compressed streams cut short
        with pytest.raises(ValueError):
            decode(bad)
    info, _ = decode(ktx_util.ktx1_cube(lv[:1], "srgb" if False else "f16", stored_levels=0))
    assert info[2] == 1 and info[4] == 1                                                         # "generate the mip chain"


def _engine_frame(lib, prefix, path=None, direct=None):
    e = Engine(192, 128, api_lib=lib, api_prefix=prefix, shadow_extent=128)
    e.load_procedural(1)
    if path is not None:
        e.load_cubemap(path)
    e.update_scene(0.0)
    e.draw()
    img = e.readback(capi.TARGET_RGBA8)
    depth = e.readback(capi.TARGET_DEPTH)
    e.close()
    return img, depth


@pytest.mark.parametrize("kind,container", [("f16", "ktx1"), ("srgb", "ktx2"), ("unorm", "ktx1")])
def test_engine_skybox_from_a_ktx_file(ork_lib, emu_lib, tmp_path, kind, container):
    lv = levels_of(kind, 32, 4)
    path = os.path.join(tmp_path, "sky.ktx")
    with open(path, "wb") as f:
        f.write(ktx_util.ktx1_cube(lv, kind) if container == "ktx1" else ktx_util.ktx2_cube(lv, kind))
    plain, depth = _engine_frame(ork_lib, "ork")
    ref, _ = _engine_frame(ork_lib, "ork", path)
    got, _ = _engine_frame(emu_lib, "emu", path)
    sky = depth.view(np.uint32) == 0x3F800000
    assert sky.mean() > 0.05, "the test camera must see some sky"
    assert not np.array_equal(ref[sky], plain[sky]), "the loaded cube map must reach the sky pixels"
    assert np.array_equal(ref[~sky], plain[~sky]), "... and only those"
    assert (np.abs(ref.astype(np.int32) - got.astype(np.int32)).max(-1) <= 1).mean() >= 0.999
    # the same texels handed to the C-ABI directly give the same sky (the file added nothing and lost nothing)
    r = capi.Renderer(ork_lib, "ork", 64, 64, shadow_extent=64)
    r2 = capi.Renderer(ork_lib, "ork", 64, 64, shadow_extent=64)
    r.set_cubemap_levels(lv, {"unorm": capi.FORMAT_UNORM, "srgb": capi.FORMAT_SRGB, "f16": capi.FORMAT_RGBA16F}[kind])
    info, out = decode(open(path, "rb").read())
    bpt = 8 if kind == "f16" else 4
    dec, off = [], 0
    for l in range(info[2]):
        s = max(1, 32 >> l)
        faces = []
        for _ in range(6):
            raw = out[off:off + s * s * bpt]
            faces.append(raw.view(np.float16).reshape(s, s, 4) if kind == "f16" else raw.reshape(s, s, 4))
            off += s * s * bpt
        dec.append(faces)
    r2.set_cubemap_levels(dec, info[3])
    view = mg.camera_view((0, 0, 0), pitch=0.4, yaw=0.7)
    sd = capi.make_scene_data(view, mg.camera_projection(1.0), [mg.light_matrix()] * 4, shadow_mode=0, enable_shadows=False)
    fp = capi.default_frame_params(enableSkybox=1, enablePostprocess=1)
    empty = np.zeros(0, capi.RENDER_OBJECT_DTYPE)
    r.draw(sd, empty, fp); r2.draw(sd, empty, fp)
    a, b = r.readback(capi.TARGET_RGBA8), r2.readback(capi.TARGET_RGBA8)
    assert np.array_equal(a, b) and a[..., :3].std() > 1.0
    r.close(); r2.close()
