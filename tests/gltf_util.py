"""Test-side glTF 2.0 writer and an independent (numpy) statement of what the engine's loader must make of the file.

`build_asset(rng)` describes a small scene in plain Python (the source of truth); `write_gltf` serialises it three ways
(.gltf + external .bin/.png, .gltf with data: URIs, .glb with images in buffer views), choosing on purpose the encodings
the loader has to convert: 8/16/32-bit indices, normalised ubyte/ushort uv and colours, interleaved (byteStride) and
planar attribute layouts, TRS and matrix nodes, a primitive without indices (skipped by the reference), a primitive
without a material, BLEND materials, PNGs of several colour types and scan-line filters.
"""
import base64
import json
import os
import struct
import zlib

import numpy as np

FLOAT, UBYTE, USHORT, UINT = 5126, 5121, 5123, 5125


# ------------------------------------------------------------------------------------------------------------ PNG
ADAM7 = [(0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)]


def png_bytes(img: np.ndarray, color_type: int, filt: int = 0, palette=None, trns=None, depth: int = 8, interlace: bool = False) -> bytes:
    """img: (H, W, C) uint8 samples already in the PNG's own channel layout (C = 1, 2, 3 or 4; indices for palette).
    interlace: Adam7 — the seven reduced images one after the other, each with its own scan lines (8-bit samples only here)."""
    h, w = img.shape[:2]
    bpp = max(1, img.shape[2] * depth // 8)
    out = bytearray()
    passes = [img[y0::dy, x0::dx] for x0, y0, dx, dy in ADAM7] if interlace else [img]
    for sub in passes:
        if sub.shape[0] == 0 or sub.shape[1] == 0:
            continue
        _filter_rows(sub, bpp, filt, out)

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body))
    data = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, color_type, 0, 0, 1 if interlace else 0))
    if palette is not None:
        data += chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None:
        data += chunk(b"tRNS", np.asarray(trns, np.uint8).tobytes())
    comp = zlib.compress(bytes(out), 6)
    half = len(comp) // 2
    data += chunk(b"IDAT", comp[:half]) + chunk(b"IDAT", comp[half:]) + chunk(b"IEND", b"")     # two IDAT chunks on purpose
    return data


def _filter_rows(img, bpp, filt, out):
    h = img.shape[0]
    rows = np.ascontiguousarray(img).reshape(h, -1).astype(np.uint8)
    prev = np.zeros(rows.shape[1], np.int32)
    for y in range(h):
        cur = rows[y].astype(np.int32)
        a = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]])
        c = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]])
        ft = filt if filt < 5 else y % 5
        if ft == 0:
            pred = 0
        elif ft == 1:
            pred = a
        elif ft == 2:
            pred = prev
        elif ft == 3:
            pred = (a + prev) >> 1
        else:
            p = a + prev - c
            pa, pb, pc = np.abs(p - a), np.abs(p - prev), np.abs(p - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, prev, c))
        out.append(ft)
        out += ((cur - pred) & 255).astype(np.uint8).tobytes()
        prev = cur


def make_images(rng):
    """[(png bytes, expected RGBA8 array)]: RGBA, RGB, palette with tRNS, grey + alpha."""
    out = []
    rgba = rng.integers(0, 256, (16, 8, 4), dtype=np.uint8)
    rgba[..., 3] = np.where(rng.random((16, 8)) < 0.3, 40, 255)             # some texels below the alpha cut-off
    out.append((png_bytes(rgba, 6, filt=5), rgba))
    rgb = rng.integers(0, 256, (8, 8, 3), dtype=np.uint8)
    out.append((png_bytes(rgb, 2, filt=4), np.concatenate([rgb, np.full((8, 8, 1), 255, np.uint8)], -1)))
    pal = rng.integers(0, 256, (5, 3), dtype=np.uint8)
    tr = np.array([255, 0, 128], np.uint8)
    idx = rng.integers(0, 5, (4, 4, 1), dtype=np.uint8)
    alpha = np.concatenate([tr, np.full(2, 255, np.uint8)])
    out.append((png_bytes(idx, 3, filt=1, palette=pal, trns=tr), np.concatenate([pal[idx[..., 0]], alpha[idx[..., 0]][..., None]], -1)))
    ga = rng.integers(0, 256, (4, 8, 2), dtype=np.uint8)
    out.append((png_bytes(ga, 4, filt=3), np.stack([ga[..., 0]] * 3 + [ga[..., 1]], -1)))
    return out


# ---------------------------------------------------------------------------------------------------------- asset
def quat_to_mat(q):           # glTF order x, y, z, w; glm::toMat4
    x, y, z, w = [np.float32(v) for v in q]
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = 1 - 2 * (y * y + z * z); m[1, 0] = 2 * (x * y + w * z); m[2, 0] = 2 * (x * z - w * y)
    m[0, 1] = 2 * (x * y - w * z); m[1, 1] = 1 - 2 * (x * x + z * z); m[2, 1] = 2 * (y * z + w * x)
    m[0, 2] = 2 * (x * z + w * y); m[1, 2] = 2 * (y * z - w * x); m[2, 2] = 1 - 2 * (x * x + y * y)
    return m


def build_asset(rng):
    """Plain-Python scene description.  Vertex attributes are stored as the values the LOADER must produce (float32);
    `enc` says how the writer encodes them, chosen so that the encoding is exact (values on the integer grid)."""
    from vk_renderer_b200 import meshgen as mg
    images = make_images(rng)
    images.append((b"\xff\xd8\xff\xe0 not a png: the loader must fall back to the error checkerboard", None))

    def grid(v, levels):        # snap to k/levels so that a normalised integer encodes the value exactly
        return (np.round(np.asarray(v, np.float64) * levels) / levels).astype(np.float32)

    sv, si = mg.uv_sphere(10, 10, 1.0)
    cv, ci = mg.cube(0.7)
    pv, pi = mg.plane(3, 3, 4.0, -1.0)
    qv, qi = mg.plane(1, 1, 2.0, 0.0)
    sv = sv.copy(); cv = cv.copy(); pv = pv.copy(); qv = qv.copy()
    cv["uv_x"], cv["uv_y"] = grid(cv["uv_x"], 65535), grid(cv["uv_y"], 65535)
    cv["color"] = grid(rng.random((len(cv), 4)), 255)
    pv["uv_x"], pv["uv_y"] = grid(pv["uv_x"] % 1.0, 255), grid(pv["uv_y"] % 1.0, 255)
    pv["normal"] = (1, 0, 0)                      # the file has no NORMAL for this primitive: loader default
    qv["color"] = grid(rng.random((len(qv), 4)), 255)
    qv["color"][:, 3] = 1.0                       # written as VEC3
    prims = {
        "sphere": dict(v=sv, i=si, index_type=USHORT, layout="planar", attrs=dict(NORMAL=FLOAT, TEXCOORD_0=FLOAT), material=0),
        "cube": dict(v=cv, i=ci, index_type=UBYTE, layout="interleaved", attrs=dict(NORMAL=FLOAT, TEXCOORD_0=USHORT, COLOR_0=UBYTE), material=1),
        "plane": dict(v=pv, i=pi, index_type=UINT, layout="planar", attrs=dict(TEXCOORD_0=UBYTE), material=None),
        "no_indices": dict(v=qv, i=None, index_type=None, layout="planar", attrs=dict(), material=1),
        "quad": dict(v=qv, i=qi, index_type=USHORT, layout="planar", attrs=dict(NORMAL=FLOAT, COLOR_0="VEC3_UBYTE"), material=2),
    }
    meshes = [dict(name="ball", prims=["sphere", "cube"]), dict(name=None, prims=["plane", "no_indices"]), dict(name="glass", prims=["quad"])]
    materials = [
        dict(base=[0.9, 0.8, 0.7, 1.0], metallic=0.25, roughness=0.75, baseTex=0, mrTex=1, normalTex=2, blend=False),
        dict(base=[0.5, 1.0, 0.25, 1.0], metallic=1.0, roughness=0.5, baseTex=3, mrTex=None, normalTex=None, blend=False),
        dict(base=[0.25, 0.5, 1.0, 0.5], metallic=0.0, roughness=1.0, baseTex=None, mrTex=None, normalTex=None, blend=True),
    ]
    textures = [0, 1, 2, 4]       # texture -> image; texture 3 points at the undecodable image
    s = float(np.float32(np.sqrt(0.5)))
    nodes = [
        dict(trs=dict(t=[1.0, 0.0, -2.0], r=[0.0, s, 0.0, s], s=[1.0, 2.0, 1.0]), mesh=0, children=[1, 2]),
        dict(matrix=[0.5, 0, 0, 0, 0, 0.5, 0, 0, 0, 0, 0.5, 0, -3.0, 0.5, 1.0, 1.0], mesh=1, children=[]),
        dict(trs=dict(t=[0.0, 1.5, 2.0]), mesh=2, children=[3]),
        dict(trs=dict(), mesh=None, children=[4]),
        dict(trs=dict(s=[0.5, 0.5, 0.5], t=[2.0, 0.0, 0.0]), mesh=0, children=[]),
        dict(trs=dict(t=[-4.0, 0.0, -6.0]), mesh=1, children=[]),
    ]
    return dict(images=images, prims=prims, meshes=meshes, materials=materials, textures=textures, nodes=nodes)


def node_local(n):
    if "matrix" in n:
        return np.array(n["matrix"], np.float32).reshape(4, 4).T          # column-major list -> row-major array
    trs = n["trs"]
    T = np.eye(4, dtype=np.float32); T[:3, 3] = trs.get("t", [0, 0, 0])
    R = quat_to_mat(trs.get("r", [0, 0, 0, 1]))
    S = np.diag(np.array(list(trs.get("s", [1, 1, 1])) + [1], np.float32))
    return (T @ R @ S).astype(np.float32)


def expected_draws(asset):
    """Depth-first flattening in the reference's order (Node::refreshTransform, src/vk_types.h:148-163; top node's children
    are the parentless nodes in index order): [(prim name, world matrix row-major, firstIndex, indexCount)]."""
    nodes = asset["nodes"]
    has_parent = {c for n in nodes for c in n["children"]}
    out = []

    def visit(i, parent):
        world = (parent @ node_local(nodes[i])).astype(np.float32)
        m = nodes[i]["mesh"]
        if m is not None:
            first = 0
            for name in asset["meshes"][m]["prims"]:
                p = asset["prims"][name]
                if p["i"] is None:
                    continue
                out.append((name, world, first, len(p["i"])))
                first += len(p["i"])
        for c in nodes[i]["children"]:
            visit(c, world)
    for i in range(len(nodes)):
        if i not in has_parent:
            visit(i, np.eye(4, dtype=np.float32))
    return out


def expected_material(asset, index, first_slot=2):
    """MaterialConstants the loader must build: image k sits in bindless slot first_slot + k; ids swapped like the reference."""
    if index is None:
        index = 0
    m = asset["materials"][index]
    slot = lambda tex, default: default if tex is None else first_slot + asset["textures"][tex]
    return dict(color=m["base"], metallic=m["metallic"], roughness=m["roughness"], albedoID=slot(m["baseTex"], 0),
                normalID=slot(m["mrTex"], 1), metalRoughID=slot(m["normalTex"], 0), transparent=m["blend"])


# --------------------------------------------------------------------------------------------------------- writer
def _encode(values, ctype):
    """float32 attribute values -> the stored component type (normalised integers for UBYTE / USHORT)."""
    if ctype == FLOAT:
        return np.ascontiguousarray(values, "<f4")
    levels = 255 if ctype == UBYTE else 65535
    return np.round(np.asarray(values, np.float64) * levels).astype("<u1" if ctype == UBYTE else "<u2")


def write_gltf(asset, directory, variant):
    """variant: 'external' (.gltf + scene.bin + image files), 'embedded' (data: URIs), 'glb'.  Returns the file path."""
    os.makedirs(directory, exist_ok=True)
    blob = bytearray()
    views, accessors = [], []

    def add_view(data: bytes, stride=None):
        while len(blob) % 4:
            blob.append(0)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)}
        if stride:
            v["byteStride"] = stride
        blob.extend(data)
        views.append(v)
        return len(views) - 1

    def add_accessor(view, ctype, count, typ, offset=0, normalized=False, minmax=None):
        a = {"bufferView": view, "componentType": ctype, "count": int(count), "type": typ}
        if offset:
            a["byteOffset"] = offset
        if normalized:
            a["normalized"] = True
        if minmax is not None:
            a["min"], a["max"] = [float(x) for x in minmax[0]], [float(x) for x in minmax[1]]
        accessors.append(a)
        return len(accessors) - 1

    prim_json = {}
    for name, p in asset["prims"].items():
        v = p["v"]
        cols = [("POSITION", np.asarray(v["position"], "<f4"), FLOAT, "VEC3")]
        for attr, ctype in p["attrs"].items():
            if attr == "NORMAL":
                cols.append((attr, _encode(v["normal"], ctype), ctype, "VEC3"))
            elif attr == "TEXCOORD_0":
                cols.append((attr, _encode(np.stack([v["uv_x"], v["uv_y"]], -1), ctype), ctype, "VEC2"))
            elif attr == "COLOR_0" and ctype == "VEC3_UBYTE":
                cols.append((attr, _encode(v["color"][:, :3], UBYTE), UBYTE, "VEC3"))
            elif attr == "COLOR_0":
                cols.append((attr, _encode(v["color"], ctype), ctype, "VEC4"))
        attrs = {}
        if p["layout"] == "interleaved":
            offs, off = [], 0
            for _, data, _, _ in cols:
                offs.append(off)
                off += (data.dtype.itemsize * data.shape[1] + 3) & ~3            # every attribute starts 4-byte aligned
            stride = off
            buf = np.zeros((len(v), stride), np.uint8)
            for (attr, data, ctype, typ), o in zip(cols, offs):
                raw = data.view(np.uint8).reshape(len(v), -1)
                buf[:, o:o + raw.shape[1]] = raw
            view = add_view(buf.tobytes(), stride)
            for (attr, data, ctype, typ), o in zip(cols, offs):
                mm = (data.min(0), data.max(0)) if attr == "POSITION" else None
                attrs[attr] = add_accessor(view, ctype, len(v), typ, offset=o, normalized=ctype != FLOAT, minmax=mm)
        else:
            for attr, data, ctype, typ in cols:
                elem = data.dtype.itemsize * data.shape[1]
                if ctype != FLOAT and elem % 4:          # vertex attribute strides must be multiples of 4: pad and say so
                    stride = (elem + 3) & ~3
                    buf = np.zeros((len(v), stride), np.uint8)
                    buf[:, :elem] = data.view(np.uint8).reshape(len(v), -1)
                    view = add_view(buf.tobytes(), stride)
                else:
                    view = add_view(data.tobytes())
                mm = (data.min(0), data.max(0)) if attr == "POSITION" else None
                attrs[attr] = add_accessor(view, ctype, len(v), typ, normalized=ctype != FLOAT, minmax=mm)
        pj = {"attributes": attrs, "mode": 4}
        if p["i"] is not None:
            dt = {UBYTE: "<u1", USHORT: "<u2", UINT: "<u4"}[p["index_type"]]
            pj["indices"] = add_accessor(add_view(np.asarray(p["i"]).astype(dt).tobytes()), p["index_type"], len(p["i"]), "SCALAR")
        if p["material"] is not None:
            pj["material"] = p["material"]
        prim_json[name] = pj

    images_json = []
    for k, (data, _) in enumerate(asset["images"]):
        if variant == "external":
            fn = f"image{k}.png"
            with open(os.path.join(directory, fn), "wb") as f:
                f.write(data)
            images_json.append({"uri": fn})
        elif variant == "embedded":
            images_json.append({"uri": "data:image/png;base64," + base64.b64encode(data).decode()})
        else:
            images_json.append({"bufferView": add_view(data), "mimeType": "image/png"})

    materials_json = []
    for m in asset["materials"]:
        pbr = {"baseColorFactor": m["base"], "metallicFactor": m["metallic"], "roughnessFactor": m["roughness"]}
        if m["baseTex"] is not None:
            pbr["baseColorTexture"] = {"index": m["baseTex"]}
        if m["mrTex"] is not None:
            pbr["metallicRoughnessTexture"] = {"index": m["mrTex"]}
        mj = {"pbrMetallicRoughness": pbr, "name": "mät \"quoted\""}
        if m["normalTex"] is not None:
            mj["normalTexture"] = {"index": m["normalTex"], "scale": 1.0}
        if m["blend"]:
            mj["alphaMode"] = "BLEND"
        materials_json.append(mj)

    nodes_json = []
    for n in asset["nodes"]:
        nj = {}
        if "matrix" in n:
            nj["matrix"] = n["matrix"]
        else:
            for key, name in (("t", "translation"), ("r", "rotation"), ("s", "scale")):
                if key in n["trs"]:
                    nj[name] = n["trs"][key]
        if n["mesh"] is not None:
            nj["mesh"] = n["mesh"]
        if n["children"]:
            nj["children"] = n["children"]
        nodes_json.append(nj)

    meshes_json = []
    for m in asset["meshes"]:
        mj = {"primitives": [prim_json[p] for p in m["prims"]]}
        if m["name"]:
            mj["name"] = m["name"]
        meshes_json.append(mj)

    doc = {"asset": {"version": "2.0", "generator": "tests/gltf_util.py"}, "scene": 0,
           "scenes": [{"nodes": [i for i in range(len(nodes_json)) if all(i not in n["children"] for n in asset["nodes"])]}],
           "nodes": nodes_json, "meshes": meshes_json, "materials": materials_json,
           "textures": [{"source": s, "sampler": 0} for s in asset["textures"]], "samplers": [{"magFilter": 9729, "minFilter": 9987}],
           "images": images_json, "accessors": accessors, "bufferViews": views, "buffers": [{"byteLength": len(blob)}]}
    if variant == "glb":
        js = json.dumps(doc, separators=(",", ":")).encode()
        js += b" " * (-len(js) % 4)
        binp = bytes(blob) + b"\0" * (-len(blob) % 4)
        path = os.path.join(directory, "scene.glb")
        with open(path, "wb") as f:
            f.write(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(binp)))
            f.write(struct.pack("<II", len(js), 0x4E4F534A) + js)
            f.write(struct.pack("<II", len(binp), 0x004E4942) + binp)
        return path
    if variant == "external":
        with open(os.path.join(directory, "scene data.bin"), "wb") as f:      # a space: the uri is percent-encoded
            f.write(blob)
        doc["buffers"][0]["uri"] = "scene%20data.bin"
    else:
        doc["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(bytes(blob)).decode()
    path = os.path.join(directory, "scene.gltf")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    return path
