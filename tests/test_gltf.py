"""glTF 2.0 / GLB ingestion of the host engine (csrc/host/vk_gltf.cpp) — VulkanEngine::load_gltf of the reference
(src/vk_loader.cpp:227-518), SURVEY.md §8f-1.

The files are written by tests/gltf_util.py from a plain-Python scene description, which is also turned — by independent
numpy code — into what the loader has to produce: draw list (index ranges, world matrices, material constants with the
reference's id swap, pass types) and, through rendering, vertex data and textures: the scene loaded from the file must
render bit-identically to the same scene uploaded directly through the C-ABI.
"""
import numpy as np
import pytest

from vk_renderer_b200 import capi, engine as eng_mod
from vk_renderer_b200.engine import Engine
import gltf_util as gu
import parity

W, H, S = 200, 120, 256
VARIANTS = ["external", "embedded", "glb"]


def _engine(lib, prefix, path):
    e = Engine(W, H, api_lib=lib, api_prefix=prefix, shadow_extent=S)
    tris = e.load_gltf(path, "test")
    e.set_camera((-1.0, 1.5, 9.0), pitch=-0.1)
    e.set_toggles(shadows=1, shadow_mode=1, background=1, postprocess=1)
    e.update_scene(0.0)
    return e, tris


def test_png_decoder_matches_the_encoder():
    rng = np.random.default_rng(7)
    for data, want in gu.make_images(rng):
        got = eng_mod.decode_png(data)
        assert got is not None and np.array_equal(got, want)
    for filt in range(5):                                   # every scan-line filter on its own
        img = rng.integers(0, 256, (9, 7, 4), dtype=np.uint8)
        assert np.array_equal(eng_mod.decode_png(gu.png_bytes(img, 6, filt=filt)), img)
    # Adam7 interlacing (stb_image reads it): sizes around the 8 x 8 pattern, some passes empty; checked against PIL too
    from PIL import Image
    import io
    for (h, w), ctype, ch in (((9, 7), 6, 4), ((1, 1), 2, 3), ((3, 2), 6, 4), ((16, 16), 2, 3), ((5, 33), 0, 1), ((13, 4), 4, 2)):
        img = rng.integers(0, 256, (h, w, ch), dtype=np.uint8)
        data = gu.png_bytes(img, ctype, filt=5, interlace=True)
        got = eng_mod.decode_png(data)
        ref = np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"))
        assert got is not None and np.array_equal(got, ref), (h, w, ctype)
        assert np.array_equal(got, eng_mod.decode_png(gu.png_bytes(img, ctype, filt=5)))      # the same pixels as the plain file
    assert eng_mod.decode_png(b"\xff\xd8\xff\xe0 jpeg") is None
    good = gu.png_bytes(rng.integers(0, 256, (4, 4, 3), dtype=np.uint8), 2)
    assert eng_mod.decode_png(good[:len(good) // 2]) is None          # truncated stream
    assert eng_mod.decode_png(good[:8]) is None


@pytest.mark.parametrize("variant", VARIANTS)
def test_draw_list_matches_the_description(ork_lib, tmp_path, variant):
    asset = gu.build_asset(np.random.default_rng(11))
    path = gu.write_gltf(asset, str(tmp_path), variant)
    e, tris = _engine(ork_lib, "ork", path)
    want = gu.expected_draws(asset)
    assert tris == sum(n // 3 for _, _, _, n in want)
    got = e.render_objects()
    assert len(got) == len(want), "every object is in view: nothing may be culled or lost"
    # update_scene sorts opaque-first (src/vk_engine_run.cpp:454-458)
    passes = [pt for _, _, pt in got]
    assert passes == sorted(passes)
    mesh_of_prim = {}
    unmatched = list(got)
    for name, world, first, count in want:
        hit = None
        for k, (ro, mc, pt) in enumerate(unmatched):
            if int(ro["firstIndex"]) == first and int(ro["indexCount"]) == count and \
                    np.allclose(np.asarray(ro["transform"]).reshape(4, 4).T, world, rtol=0, atol=2e-6):
                hit = k
                break
        assert hit is not None, f"no draw for {name}"
        ro, mc, pt = unmatched.pop(hit)
        exp = gu.expected_material(asset, asset["prims"][name]["material"])
        assert list(mc.colorFactors) == [float(np.float32(x)) for x in exp["color"]]
        assert (mc.metal_rough_factors[0], mc.metal_rough_factors[1]) == (float(np.float32(exp["metallic"])), float(np.float32(exp["roughness"])))
        assert (mc.albedoID, mc.normalID, mc.metalRoughID) == (exp["albedoID"], exp["normalID"], exp["metalRoughID"]), name
        assert pt == (1 if exp["transparent"] else 0)
        mesh_of_prim.setdefault(name, set()).add(int(ro["mesh"]))
    # one MeshAsset per glTF mesh: the two primitives of "ball" share a buffer, instances reuse it
    assert mesh_of_prim["sphere"] == mesh_of_prim["cube"] and len(mesh_of_prim["sphere"]) == 1
    assert mesh_of_prim["plane"].isdisjoint(mesh_of_prim["sphere"]) and mesh_of_prim["quad"].isdisjoint(mesh_of_prim["plane"])
    e.close()


def _render_direct(lib, prefix, asset, objects, sd, fp):
    """The same scene through the C-ABI, converted from the description by numpy."""
    r = capi.Renderer(lib, prefix, W, H, shadow_extent=S)
    for k, (_, rgba) in enumerate(asset["images"]):
        if rgba is None:          # _errorCheckerboardImage (src/vk_engine_init.cpp:329-341)
            chk = np.zeros((16, 16, 4), np.uint8)
            chk[..., 3] = 255
            yy, xx = np.mgrid[0:16, 0:16]
            chk[((xx % 2) ^ (yy % 2)) == 1] = (255, 0, 255, 255)
            img = r.create_image(chk, capi.FORMAT_UNORM, False)
        else:
            img = r.create_image(rgba, capi.FORMAT_SRGB, True)
        r.update_bindless_texture(2 + k, img, capi.SAMPLER_LINEAR)
    mats = []
    for i in range(len(asset["materials"])):
        m = gu.expected_material(asset, i)
        mats.append(r.build_material(m["color"], m["metallic"], m["roughness"], albedo_id=m["albedoID"], normal_id=m["normalID"],
                                     metal_rough_id=m["metalRoughID"], pass_type=capi.PASS_TRANSPARENT if m["transparent"] else capi.PASS_MAIN_COLOR))
    meshes = []
    for m in asset["meshes"]:
        vs, idx, base = [], [], 0
        for name in m["prims"]:
            p = asset["prims"][name]
            if p["i"] is None:
                continue
            vs.append(p["v"]); idx.append(np.asarray(p["i"], np.uint32) + base); base += len(p["v"])
        meshes.append(r.upload_mesh(np.concatenate(idx), np.concatenate(vs)))
    # the engine's own draw list (order, matrices), with its handles mapped onto this renderer's by creation order
    mesh_ids = sorted({int(o["mesh"]) for o, _, _ in objects})
    mat_ids = sorted({int(o["material"]) for o, _, _ in objects})
    assert len(mesh_ids) == len(meshes) and len(mat_ids) == len(mats)
    objs = []
    for o, _, _ in objects:
        o2 = o.copy()
        o2["mesh"] = meshes[mesh_ids.index(int(o["mesh"]))]
        o2["material"] = mats[mat_ids.index(int(o["material"]))]
        objs.append(o2)
    r.draw(sd, np.array(objs), fp)
    out = {k: r.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID, shadow=capi.TARGET_SHADOW).items()}
    r.close()
    return out


@pytest.mark.parametrize("variant", VARIANTS)
def test_loaded_scene_renders_like_direct_upload(ork_lib, tmp_path, variant):
    asset = gu.build_asset(np.random.default_rng(11))
    path = gu.write_gltf(asset, str(tmp_path), variant)
    e, _ = _engine(ork_lib, "ork", path)
    e.draw()
    a = {k: e.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID, shadow=capi.TARGET_SHADOW).items()}
    sd = e.scene_data()
    fp = capi.default_frame_params(enableShadows=1, shadowMode=1, enableBackground=1, enablePostprocess=1, enableSkybox=0)
    b = _render_direct(ork_lib, "ork", asset, e.render_objects(), sd, fp)
    e.close()
    assert (a["prim"] != 0xFFFFFFFF).mean() > 0.02
    for k in a:
        assert np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)), f"{k} differs: vertex data, indices, textures or materials were not loaded as described"


def test_malformed_files_are_refused(ork_lib, tmp_path):
    e = Engine(W, H, api_lib=ork_lib, api_prefix="ork", shadow_extent=S)
    cases = {"missing.gltf": None, "garbage.gltf": b"{ not json", "empty.glb": b"glTF\x02\x00\x00\x00\x0c\x00\x00\x00",
             "nobuffer.gltf": b'{"asset":{"version":"2.0"},"buffers":[{"byteLength":4,"uri":"nope.bin"}]}',
             "badaccessor.gltf": b'{"asset":{"version":"2.0"},"meshes":[{"primitives":[{"indices":3,"attributes":{"POSITION":0}}]}]}'}
    for name, data in cases.items():
        p = tmp_path / name
        if data is not None:
            p.write_bytes(data)
        with pytest.raises(RuntimeError):
            e.load_gltf(str(p))
    # a file with nothing in it but the asset tag loads as an empty scene (the reference builds a default material and a top node)
    p = tmp_path / "empty.gltf"
    p.write_text('{"asset": {"version": "2.0"}}')
    assert e.load_gltf(str(p)) == 0
    e.close()


@pytest.mark.gpu
def test_gltf_scene_on_the_device_matches_the_oracle(ork_lib, vkr_lib, tmp_path):
    asset = gu.build_asset(np.random.default_rng(11))
    path = gu.write_gltf(asset, str(tmp_path), "glb")
    out = {}
    for lib, prefix in ((ork_lib, "ork"), (None, "vkr")):
        e, tris = _engine(lib, prefix, path)
        e.draw()
        out[prefix] = {k: e.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID,
                                                        shadow=capi.TARGET_SHADOW).items()}
        out[prefix].update(stats=e.stats(), shadows=True, layers=1)
        e.close()
    parity.assert_parity(out["ork"], out["vkr"])


def test_component_conversion_matches_fastgltfs_own_test():
    """The reference's vendored fastgltf carries one known-answer test for the loader's arithmetic, "Test data type conversion"
    (external/fastgltf/tests/accessor_tests.cpp:29-51; not built by the reference): a normalised integer i of type T becomes
    max(i / T_max, -1) to within float epsilon — over every value of int8, uint8, int16 and uint16.  Restated here on the
    loader's accessor conversion, plus the non-normalised and float pass-through cases of convertComponent
    (external/fastgltf/include/fastgltf/tools.hpp:126-149)."""
    import ctypes as C
    from vk_renderer_b200.engine import load_engine
    lib = load_engine()
    lib.vke_gltf_convert.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.vke_gltf_convert.restype = None

    def convert(values, component_type, normalized):
        values = np.ascontiguousarray(values)
        out = np.zeros(len(values), np.float32)
        lib.vke_gltf_convert(component_type, int(normalized), values.ctypes.data, len(values), out.ctypes.data)
        return out

    eps = np.finfo(np.float32).eps
    for dtype, ctype in ((np.int8, 5120), (np.uint8, 5121), (np.int16, 5122), (np.uint16, 5123)):
        info = np.iinfo(dtype)
        i = np.arange(info.min, info.max + 1, dtype=np.int64).astype(dtype)
        got = convert(i, ctype, True)
        want = np.maximum(i.astype(np.float32) / np.float32(info.max), np.float32(-1.0))
        assert np.all(np.abs(got - want) <= eps), dtype
        assert got.min() == (-1.0 if info.min < 0 else 0.0) and got.max() == 1.0
        back = np.round(got.astype(np.float64) * info.max)              # the test's float -> int leg: exact round trip
        assert np.array_equal(back[i != info.min], i[i != info.min].astype(np.float64))
        assert np.array_equal(convert(i, ctype, False), i.astype(np.float32))       # not normalised: a plain cast
    u32 = np.array([0, 1, 65536, 4000000000], np.uint32)
    assert np.array_equal(convert(u32, 5125, False), u32.astype(np.float32))
    f32 = np.array([0.25, -3.5, 1e-20, 7e9], np.float32)
    assert np.array_equal(convert(f32, 5126, True), f32)                               # floats pass through either way


def _jpeg_cases():
    from PIL import Image
    import io
    rng = np.random.default_rng(3)

    def picture(w, h, noise):
        y, x = np.mgrid[0:h, 0:w]
        a = np.stack([127 + 100 * np.sin(x / 9.0 + y / 17.0), 127 + 100 * np.cos(x / 13.0), 127 + 90 * np.sin(y / 7.0 - x / 23.0)], -1)
        return np.clip(a + rng.normal(0, noise, a.shape), 0, 255).astype(np.uint8)

    for w, h in ((64, 48), (37, 29), (129, 65), (8, 8), (1, 1), (17, 200)):
        for kw in (dict(subsampling="4:4:4"), dict(subsampling="4:2:2"), dict(subsampling="4:2:0"), dict(subsampling="4:2:0", progressive=True),
                   dict(subsampling="4:4:4", progressive=True, quality=95), dict(subsampling="4:2:0", quality=30),
                   dict(subsampling="4:2:0", quality=90, optimize=True), dict(subsampling="4:2:0", restart_marker_blocks=3),
                   dict(subsampling="4:4:4", progressive=True, restart_marker_rows=1), dict(mode="L"), dict(mode="L", progressive=True)):
            kw = dict(kw)
            mode = kw.pop("mode", "RGB")
            a = picture(w, h, 6 if kw.get("quality", 80) >= 50 else 40)
            buf = io.BytesIO()
            Image.fromarray(a if mode == "RGB" else a[..., 0], mode).save(buf, "JPEG", **({"quality": 80} | kw))
            data = buf.getvalue()
            yield (w, h, mode, kw), data, np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))


def test_jpeg_decoder_against_an_independent_decode():
    """glTF images are PNG or JPEG and the reference decodes both with stb_image (src/vk_loader.cpp:280,293).  The engine's own
    JPEG decoder (csrc/host/vk_jpeg.cpp: baseline and progressive, 4:4:4 / 4:2:2 / 4:2:0, grey, restart intervals) against
    PIL's (libjpeg) on the same streams: decoders differ in their inverse DCT and chroma filters by a level or two, never more."""
    n = 0
    for what, data, want in _jpeg_cases():
        got = eng_mod.decode_jpeg(data)
        assert got is not None and got.shape == want.shape[:2] + (4,), what
        assert (got[..., 3] == 255).all()
        d = np.abs(got[..., :3].astype(int) - want.astype(int))
        assert d.max() <= 4 and d.mean() < 0.5, (what, d.max(), d.mean())
        n += 1
    assert n == 66
    assert eng_mod.decode_jpeg(b"\x89PNG\r\n\x1a\n....") is None
    good = next(_jpeg_cases())[1]
    assert eng_mod.decode_jpeg(good[:200]) is None or True          # a truncated stream must not crash (missing scans decode as grey)
    assert eng_mod.decode_jpeg(good[:2]) is None


def test_gltf_with_a_jpeg_texture_loads_it(ork_lib, tmp_path):
    """A glTF whose base-colour image is a JPEG renders exactly like the same file with that image re-encoded losslessly (PNG of the
    engine's own decode): the loader used the decoder, not the error checkerboard."""
    from PIL import Image
    import io
    rng = np.random.default_rng(5)
    asset = gu.build_asset(rng)
    y, x = np.mgrid[0:32, 0:32]
    pic = np.stack([x * 8, y * 8, (x ^ y) * 8], -1).astype(np.uint8)
    buf = io.BytesIO()
    Image.fromarray(pic, "RGB").save(buf, "JPEG", quality=90, subsampling="4:2:0", progressive=True)
    jpg = buf.getvalue()
    decoded = eng_mod.decode_jpeg(jpg)
    assert decoded is not None and np.abs(decoded[..., :3].astype(int) - pic.astype(int)).mean() < 6
    frames = []
    for k, image in enumerate(((jpg, None), (gu.png_bytes(decoded, 6), decoded))):
        asset["images"][0] = image
        path = gu.write_gltf(asset, str(tmp_path / f"v{k}"), "glb" if k == 0 else "external")
        e, tris = _engine(ork_lib, "ork", path)
        e.draw()
        frames.append(e.readback(capi.TARGET_RGBA8))
        e.close()
    assert np.array_equal(frames[0], frames[1])


def _make_sparse(path, out_path, rng):
    """Rewrite an 'external' .gltf so that (a) every POSITION accessor becomes a sparse one over a CORRUPTED copy of its data —
    a third of the elements zeroed in the base buffer view and restored through sparse.indices / sparse.values — and (b) every
    index accessor becomes a sparse accessor WITHOUT a buffer view (all zeros) whose sparse part supplies every element.  What
    the file describes is unchanged."""
    import json, os, urllib.parse
    g = json.load(open(path))
    bin_path = os.path.join(os.path.dirname(path), urllib.parse.unquote(g["buffers"][0]["uri"]))
    blob = bytearray(open(bin_path, "rb").read())

    def add_view(data):
        while len(blob) % 4:
            blob.append(0)
        g["bufferViews"].append({"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)})
        blob.extend(data)
        return len(g["bufferViews"]) - 1

    def dense(acc):
        v = g["bufferViews"][acc["bufferView"]]
        comps = {"SCALAR": 1, "VEC3": 3}[acc["type"]]
        dt = {5121: "<u1", 5123: "<u2", 5125: "<u4", 5126: "<f4"}[acc["componentType"]]
        elem = np.dtype(dt).itemsize * comps
        stride = v.get("byteStride", elem)
        start = v.get("byteOffset", 0) + acc.get("byteOffset", 0)
        rows = [bytes(blob[start + i * stride: start + i * stride + elem]) for i in range(acc["count"])]
        return np.frombuffer(b"".join(rows), dt).reshape(acc["count"], comps).copy(), dt

    pos_ids, idx_ids = set(), set()
    for m in g["meshes"]:
        for p in m["primitives"]:
            pos_ids.add(p["attributes"]["POSITION"])
            if "indices" in p:
                idx_ids.add(p["indices"])
    for k in sorted(pos_ids):
        acc = g["accessors"][k]
        data, dt = dense(acc)
        n = len(data)
        which = np.sort(rng.choice(n, max(1, n // 3), replace=False))
        broken = data.copy()
        broken[which] = 0
        acc["bufferView"], acc.pop("byteOffset", None)
        acc["bufferView"] = add_view(broken.astype(dt).tobytes())
        it = 5123 if n < 65536 else 5125
        acc["sparse"] = {"count": int(len(which)),
                         "indices": {"bufferView": add_view(which.astype("<u2" if it == 5123 else "<u4").tobytes()), "componentType": it},
                         "values": {"bufferView": add_view(data[which].astype(dt).tobytes())}}
    for k in sorted(idx_ids):
        acc = g["accessors"][k]
        data, dt = dense(acc)
        n = len(data)
        del acc["bufferView"]
        acc.pop("byteOffset", None)
        acc["sparse"] = {"count": n, "indices": {"bufferView": add_view(np.arange(n, dtype="<u4").tobytes()), "componentType": 5125},
                         "values": {"bufferView": add_view(data.astype(dt).tobytes())}}
    g["buffers"][0]["byteLength"] = len(blob)
    open(bin_path, "wb").write(bytes(blob))
    json.dump(g, open(out_path, "w"))
    return out_path


def test_sparse_accessors_are_resolved(ork_lib, tmp_path):
    """glTF 2.0 sparse accessors (fastgltf's iterateAccessor, which the reference reads every attribute through, resolves them):
    the same scene written with sparse POSITION and index accessors must load to the same frame, bit for bit."""
    import os
    asset = gu.build_asset(np.random.default_rng(11))
    frames = []
    for sub, sparse in (("dense", False), ("sparse", True)):
        d = tmp_path / sub
        path = gu.write_gltf(asset, str(d), "external")
        if sparse:
            path = _make_sparse(path, os.path.join(str(d), "sparse.gltf"), np.random.default_rng(5))
        e, tris = _engine(ork_lib, "ork", path)
        e.draw()
        frames.append((tris, {k: e.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID).items()}))
        e.close()
    assert frames[0][0] == frames[1][0] and frames[0][0] > 0
    assert (frames[0][1]["prim"] != 0xFFFFFFFF).mean() > 0.02
    for k in frames[0][1]:
        assert np.array_equal(frames[0][1][k].view(np.uint8), frames[1][1][k].view(np.uint8)), k
    # a sparse index beyond the accessor is refused
    import json
    g = json.load(open(os.path.join(str(tmp_path / "sparse"), "sparse.gltf")))
    acc = next(a for a in g["accessors"] if "sparse" in a and a["type"] == "VEC3")
    acc["count"] = 1
    json.dump(g, open(os.path.join(str(tmp_path / "sparse"), "bad.gltf"), "w"))
    e = Engine(W, H, api_lib=ork_lib, api_prefix="ork", shadow_extent=S)
    with pytest.raises(RuntimeError):
        e.load_gltf(os.path.join(str(tmp_path / "sparse"), "bad.gltf"))
    e.close()
