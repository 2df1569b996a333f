"""The drop-in boundary itself: include/vkr.h is plain C, libvkr.so exports every function it declares, the ctypes
mirror in vk_renderer_b200/capi.py has the header's struct layouts, and the error behaviour of the entry points
(no exceptions, no aborts: a negative VkrResult plus vkr_last_error).

The first half needs no GPU (nothing is computed); the second half is marked gpu."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from vk_renderer_b200 import capi, meshgen as mg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vkr.h")


def declared_functions():
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"^(?:int|void|size_t|const char\*)\s+(vkr_[a-z0-9_]+)\s*\(", text, flags=re.M)))


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "vkr.h"\nint main(void) { return (int)sizeof(VkrFrameParams) == 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)], check=True)


def test_library_exports_every_declared_symbol(vkr_lib):
    names = declared_functions()
    assert len(names) >= 20 and "vkr_draw" in names and "vkr_shard_init" in names
    missing = [n for n in names if not hasattr(vkr_lib, n)]
    assert not missing, f"libvkr.so does not export {missing}"
    # the ctypes binding covers the same set
    assert sorted("vkr_" + s for s in capi.VKR_SYMBOLS) == names
    vkr_lib.vkr_version.restype = C.c_char_p
    assert b"sm_100a" in vkr_lib.vkr_version()


def test_struct_layouts_match_the_header(tmp_path):
    """sizeof / offsetof from the C compiler against the ctypes structures the tests and bench.py pass across the ABI,
    and against the reference's layouts (Vertex 48 B, GPUSceneData 512 B, MaterialConstants 48 B)."""
    probes = [("VkrVertex", None), ("VkrGPUSceneData", capi.GPUSceneData), ("VkrMaterialConstants", capi.MaterialConstants),
              ("VkrRenderObject", None), ("VkrComputePushConstants", capi.ComputePushConstants), ("VkrFrameParams", capi.FrameParams),
              ("VkrConfig", capi.Config), ("VkrEngineStats", capi.EngineStats)]
    fields = {"VkrGPUSceneData": ["view", "proj", "viewproj", "lightViewproj", "cascadeDistances", "ambientColor", "sunlightDirection", "sunlightColor"],
              "VkrFrameParams": ["enableShadows", "fragShader", "scissor", "background", "shadowScissor"],
              "VkrRenderObject": ["mesh", "material", "transform", "indexCount", "firstIndex"],
              "VkrVertex": ["position", "uv_x", "normal", "uv_y", "color"]}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "vkr.h"', 'int main(void) {']
    for name, _ in probes:
        lines.append(f'printf("{name} %zu\\n", sizeof({name}));')
        for f in fields.get(name, []):
            lines.append(f'printf("{name}.{f} %zu\\n", offsetof({name}, {f}));')
    lines += ["return 0; }"]
    src, exe = tmp_path / "layout.c", tmp_path / "layout"
    src.write_text("\n".join(lines))
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    got = {k: int(v) for k, v in got.items()}
    assert got["VkrVertex"] == 48 and got["VkrGPUSceneData"] == 512 and got["VkrMaterialConstants"] == 48
    assert [got[f"VkrVertex.{f}"] for f in fields["VkrVertex"]] == [0, 12, 16, 28, 32]                 # src/vk_types.h:85-91
    assert [got[f"VkrGPUSceneData.{f}"] for f in fields["VkrGPUSceneData"]] == [0, 64, 128, 192, 448, 464, 480, 496]
    for name, cls in probes:
        if cls is not None:
            assert C.sizeof(cls) == got[name], name
            for f in fields.get(name, []):
                assert getattr(cls, f).offset == got[f"{name}.{f}"], f"{name}.{f}"
    assert capi.RENDER_OBJECT_DTYPE.itemsize == got["VkrRenderObject"]
    for f in fields["VkrRenderObject"]:
        assert capi.RENDER_OBJECT_DTYPE.fields[f][1] == got[f"VkrRenderObject.{f}"]
    assert mg.VERTEX_DTYPE.itemsize == 48


def test_create_without_a_device_fails_loudly(vkr_lib):
    """No CPU fallback: without a CUDA device vkr_create returns VKR_ERROR_NO_DEVICE and no handle."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    cfg = capi.Config(0, 64, 64, 64, None, 0)
    dev = C.c_void_p()
    vkr_lib.vkr_create.argtypes = [C.c_void_p, C.c_void_p]
    rc = vkr_lib.vkr_create(C.byref(cfg), C.byref(dev))
    assert rc == capi.ERROR_NO_DEVICE and not dev.value
    with pytest.raises(capi.VkrError):
        capi.Renderer(vkr_lib, "vkr", 64, 64)
    # NULL arguments never crash
    assert vkr_lib.vkr_create(None, C.byref(dev)) == capi.ERROR_INVALID_ARGUMENT
    vkr_lib.vkr_destroy.argtypes = [C.c_void_p]
    vkr_lib.vkr_destroy(None)


# ------------------------------------------------------------------------------------------------- on the device
def _small(lib, prefix):
    r = capi.Renderer(lib, prefix, 160, 96, shadow_extent=128)
    sv, si = mg.uv_sphere(8, 8, 1.0)
    mesh = r.upload_mesh(si, sv)
    mat = r.build_material((1, 0.5, 0.25, 1), 0.0, 0.5)
    glass = r.build_material((0.2, 0.4, 0.8, 1), 0.0, 0.5, pass_type=capi.PASS_TRANSPARENT)
    r.set_cubemap(mg.sky_faces(16))
    view, proj = mg.camera_view((0, 0, 5)), mg.camera_projection(160 / 96)
    sd = capi.make_scene_data(view, proj, [mg.light_matrix()] * 4, shadow_mode=1, enable_shadows=True)
    fp = capi.default_frame_params(enableShadows=1, shadowMode=1, enablePostprocess=1, enableSkybox=1)
    return r, mesh, mat, glass, len(si), sd, fp


@pytest.mark.gpu
def test_invalid_arguments_are_reported_not_fatal(vkr_lib):
    r, mesh, mat, glass, ni, sd, fp = _small(vkr_lib, "vkr")
    ok = mg.render_object(mesh, mat, np.eye(4), ni)

    def code(objs):
        with pytest.raises(capi.VkrError) as e:
            r.draw(sd, np.array(objs), fp)
        assert str(e.value).split(":", 1)[1].strip(), "vkr_last_error is empty"
        return e.value.code

    assert code([mg.render_object(mesh + 7, mat, np.eye(4), ni)]) == capi.ERROR_INVALID_ARGUMENT          # unknown mesh
    assert code([mg.render_object(mesh, mat + 9, np.eye(4), ni)]) == capi.ERROR_INVALID_ARGUMENT          # unknown material
    assert code([mg.render_object(mesh, mat, np.eye(4), ni + 3)]) == capi.ERROR_INVALID_ARGUMENT          # index range past the mesh
    bad = mg.render_object(mesh, mat, np.eye(4), 6)
    bad["firstIndex"] = ni - 3
    assert code([bad]) == capi.ERROR_INVALID_ARGUMENT
    # update_scene sorts opaque-first (src/vk_engine_run.cpp:454-458); a list that is not is refused, not mis-rendered
    assert code([mg.render_object(mesh, glass, np.eye(4), ni), ok]) == capi.ERROR_INVALID_ARGUMENT
    f = r._f
    assert f("draw")(r.dev, None, None, 0, C.byref(fp)) == capi.ERROR_INVALID_ARGUMENT
    assert f("draw")(r.dev, C.byref(sd), None, 3, C.byref(fp)) == capi.ERROR_INVALID_ARGUMENT
    assert f("draw")(None, C.byref(sd), None, 0, C.byref(fp)) == capi.ERROR_INVALID_ARGUMENT
    out = np.empty(16, np.uint8)
    assert f("readback")(r.dev, capi.TARGET_RGBA8, out.ctypes.data, out.nbytes) == capi.ERROR_INVALID_ARGUMENT   # too small
    assert f("readback")(r.dev, 99, out.ctypes.data, out.nbytes) == capi.ERROR_INVALID_ARGUMENT
    assert f("update_bindless_texture")(r.dev, capi.MAX_BINDLESS, 0, 0) == capi.ERROR_INVALID_ARGUMENT
    assert f("update_bindless_texture")(r.dev, 3, 12345, 0) == capi.ERROR_INVALID_ARGUMENT
    assert f("shard_init")(r.dev, 2, 2, None, 0) == capi.ERROR_INVALID_ARGUMENT
    # cube maps: RGBA16F only through set_cubemap_levels; level count within the extent's chain; no null level
    face = np.zeros((4, 4, 4), np.uint8)
    six = (C.c_void_p * 6)(*[face.ctypes.data] * 6)
    assert f("set_cubemap")(r.dev, six, 4, 4, capi.FORMAT_RGBA16F, 1) == capi.ERROR_INVALID_ARGUMENT
    assert f("set_cubemap_levels")(r.dev, six, 4, 4, 0, capi.FORMAT_SRGB) == capi.ERROR_INVALID_ARGUMENT
    assert f("set_cubemap_levels")(r.dev, six, 4, 4, 17, capi.FORMAT_SRGB) == capi.ERROR_INVALID_ARGUMENT
    assert f("set_cubemap_levels")(r.dev, six, 4, 4, 1, 7) == capi.ERROR_INVALID_ARGUMENT
    holes = (C.c_void_p * 6)(face.ctypes.data, None, face.ctypes.data, face.ctypes.data, face.ctypes.data, face.ctypes.data)
    assert f("set_cubemap_levels")(r.dev, holes, 4, 4, 1, capi.FORMAT_SRGB) == capi.ERROR_INVALID_ARGUMENT
    assert f("set_cubemap_levels")(r.dev, None, 4, 4, 1, capi.FORMAT_SRGB) == capi.ERROR_INVALID_ARGUMENT
    img = C.c_uint32()
    assert f("create_image")(r.dev, face.ctypes.data, 4, 4, capi.FORMAT_RGBA16F, 0, C.byref(img)) == capi.ERROR_INVALID_ARGUMENT
    # the device is still usable after every refusal
    r.draw(sd, np.array([ok]), fp)
    assert (r.readback(capi.TARGET_PRIMID) != 0xFFFFFFFF).any()
    r.close()


CASES = ["empty_list", "zero_index_draws", "all_offscreen"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_degenerate_draw_lists_match_the_oracle(ork_lib, vkr_lib, case):
    _degenerate(ork_lib, vkr_lib, "vkr", case)


@pytest.mark.parametrize("case", CASES)
def test_degenerate_draw_lists_host_emulation(ork_lib, emu_lib, case):
    _degenerate(ork_lib, emu_lib, "emu", case)


def _degenerate(ork_lib, lib_b, prefix_b, case):
    """Empty and ragged inputs: no draws at all (sky only), draws with indexCount 0 between real ones, and a list whose
    every draw is outside the view volume."""
    out = {}
    for lib, prefix in ((ork_lib, "ork"), (lib_b, prefix_b)):
        r, mesh, mat, glass, ni, sd, fp = _small(lib, prefix)
        if case == "empty_list":
            objs = np.zeros(0, capi.RENDER_OBJECT_DTYPE)
        elif case == "zero_index_draws":
            objs = np.array([mg.render_object(mesh, mat, np.eye(4), 0), mg.render_object(mesh, mat, mg.translate(-1, 0, 0), ni),
                             mg.render_object(mesh, mat, np.eye(4), 0), mg.render_object(mesh, mat, mg.translate(1, 0, 0), 3 * 5),
                             mg.render_object(mesh, glass, mg.translate(0, 0.5, 1), 0)])
        else:
            objs = np.array([mg.render_object(mesh, mat, mg.translate(0, 0, 50), ni), mg.render_object(mesh, mat, mg.translate(500, 0, 0), ni)])
        r.draw(sd, objs, fp)
        out[prefix] = {k: r.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, prim=capi.TARGET_PRIMID, depth=capi.TARGET_DEPTH,
                                                         shadow=capi.TARGET_SHADOW).items()}
        out[prefix]["stats"] = r.stats()
        r.close()
    a, b = out["ork"], out[prefix_b]
    assert a["stats"].triangle_count == b["stats"].triangle_count and a["stats"].drawcall_count == b["stats"].drawcall_count
    assert np.array_equal(a["prim"], b["prim"])
    assert np.array_equal(a["depth"].view(np.uint32), b["depth"].view(np.uint32))
    assert np.array_equal(a["shadow"][:1].view(np.uint32), b["shadow"][:1].view(np.uint32))
    assert (np.abs(a["rgba"].astype(int) - b["rgba"].astype(int)).max(-1) <= 1).mean() >= 0.999
    if case != "zero_index_draws":
        assert (b["prim"] == 0xFFFFFFFF).all()                      # nothing but sky
        assert len(np.unique(b["rgba"].reshape(-1, 4), axis=0)) > 4  # ... and the sky is there
