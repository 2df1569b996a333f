#!/usr/bin/env python
"""Mechanical GLSL -> C++ rewrite of the reference's shaders (TEST INFRASTRUCTURE ONLY).

  python gen.py /root/reference oracle/_ref

Reads <reference>/shaders/*.{vert,frag,comp} where they lie and writes one `<name>.inc` per shader
into the output directory (git-ignored: no reference source enters the history).  The shader
bodies are NOT restated: every statement of the .inc files is the reference's own text, with only
these syntactic rewrites so that g++ accepts it against the reference's vendored glm:

  * float literals get an `f` suffix (`1.0` -> `1.0f`), otherwise max(float,double) is ambiguous;
  * `#version/#extension` lines dropped, `#include "common.glsl"` inlined;
  * `layout(...) in|out T x;`, `layout(...) uniform T x;` -> `static T x;`
  * `layout(...) uniform Block {...} inst;` -> `struct Block {...}; static Block inst;`
  * `layout(buffer_reference...) readonly buffer VertexBuffer {Vertex vertices[];};` -> pointer struct;
  * `uniform sampler2D globalTextures[];` -> an indexable table object;
  * `textureSize(shadowMap, 0).xy` -> `vec2(textureSize(shadowMap, 0))` (glm has no float/ivec-swizzle);
  * `void main()` -> `void shader_main()`;
  * multi-component swizzles `.xyz` -> `.xyz()` (glm's function swizzles under GLM_FORCE_SWIZZLE).
`discard`, `nonuniformEXT`, `texture`, `imageLoad`... are supplied by refshader.cpp as macros/stubs.
"""
import os
import re
import sys

SHADERS = ["mesh_pbr.frag", "mesh.frag", "mesh.vert", "shadow.vert", "skybox.vert", "skybox.frag",
           "tonemap.comp", "gradient_color.comp"]


def rewrite(src: str, shader_dir: str) -> str:
    def inline(m):
        return rewrite_body(open(os.path.join(shader_dir, m.group(1))).read())
    src = re.sub(r'#include\s+"([^"]+)"', inline, src)
    return rewrite_body(src)


def rewrite_body(s: str) -> str:
    s = re.sub(r"^\s*#(version|extension).*$", "", s, flags=re.M)
    # float literals
    s = re.sub(r"(?<![\w.])(\d+\.\d*|\.\d+)(?![\w.])", r"\1f", s)
    # buffer reference block
    s = re.sub(r"layout\s*\(\s*buffer_reference[^)]*\)\s*readonly\s+buffer\s+(\w+)\s*\{\s*(\w+)\s+(\w+)\[\];\s*\}\s*;",
               r"struct \1 { const \2* \3; };", s)
    # compute workgroup size
    s = re.sub(r"layout\s*\([^)]*local_size[^)]*\)\s*in\s*;", "", s)
    # uniform blocks with instance name
    s = re.sub(r"layout\s*\([^)]*\)\s*uniform\s+(\w+)\s*\{([^}]*)\}\s*(\w+)\s*;", r"struct \1 {\2}; static \1 \3;", s)
    # bindless table
    s = re.sub(r"layout\s*\([^)]*\)\s*uniform\s+sampler2D\s+(\w+)\[\]\s*;", r"static sampler2DTable \1;", s)
    # plain uniforms (samplers, images)
    s = re.sub(r"layout\s*\([^)]*\)\s*uniform\s+(?:readonly\s+|writeonly\s+)?(\w+)\s+(\w+)\s*;", r"static \1 \2;", s)
    # stage inputs / outputs
    s = re.sub(r"layout\s*\([^)]*\)\s*(?:in|out)\s+(\w+)\s+(\w+)\s*;", r"static \1 \2;", s)
    s = re.sub(r"textureSize\(\s*(\w+)\s*,\s*0\s*\)\.xy", r"vec2(textureSize(\1, 0))", s)
    s = re.sub(r"void\s+main\s*\(\s*\)", "void shader_main()", s)
    # multi-component swizzles: glm (C++17, no anonymous-struct extension) exposes them as functions
    s = re.sub(r"\.([xyzw]{2,4}|[rgba]{2,4})\b(?!\s*\()", r".\1()", s)
    return s


HOST_FUNCS = [  # (file, function name as it appears before the parameter list)
    ("src/vk_camera.cpp", "Camera::getViewMatrix"), ("src/vk_camera.cpp", "Camera::getRotationMatrix"),
    ("src/vk_camera.cpp", "Camera::getProjectionMatrix"),
    ("src/vk_engine_run.cpp", "extract_frustum_planes"), ("src/vk_engine_run.cpp", "VulkanEngine::is_visible"),
    ("src/vk_engine_run.cpp", "VulkanEngine::compute_light_matrix"), ("src/vk_engine_run.cpp", "getFrustumCornerWorld"),
    ("src/vk_engine_run.cpp", "VulkanEngine::getLightMatrix"), ("src/vk_engine_run.cpp", "compute_split"),
    ("src/vk_engine_run.cpp", "VulkanEngine::compute_csmdata"),
]


def extract_function(text: str, name: str) -> str:
    """The verbatim definition `<ret> name(...) {...}` (brace matched) from a C++ source text."""
    m = re.search(r"^[^\n;{}()]*\b" + re.escape(name) + r"\s*\([^)]*\)\s*\{", text, flags=re.M)
    if not m:
        raise SystemExit(f"gen.py: {name} not found")
    depth, i = 0, m.end() - 1
    while True:
        depth += {"{": 1, "}": -1}.get(text[i], 0)
        i += 1
        if depth == 0:
            return text[m.start():i]


def host_math(ref: str, out: str):
    """Host functions of the path (camera, light/CSM matrices, frustum cull), verbatim, for refhost.cpp.
    Their translation units include <vulkan/vulkan.h>, SDL and VMA, so they cannot be compiled whole."""
    parts = []
    for rel, name in HOST_FUNCS:
        parts.append(f"// ---- {rel}: {name}\n" + extract_function(open(os.path.join(ref, rel)).read(), name))
    with open(os.path.join(out, "host_math.inc"), "w") as f:
        f.write("// generated from the reference's src/*.cpp by oracle/refshader/gen.py — do not commit\n" + "\n\n".join(parts) + "\n")


def main():
    ref, out = sys.argv[1], sys.argv[2]
    sd = os.path.join(ref, "shaders")
    os.makedirs(out, exist_ok=True)
    for name in SHADERS:
        text = rewrite(open(os.path.join(sd, name)).read(), sd)
        with open(os.path.join(out, name.replace(".", "_") + ".inc"), "w") as f:
            f.write(f"// generated from {sd}/{name} by oracle/refshader/gen.py — do not commit\n" + text + "\n")
    host_math(ref, out)


if __name__ == "__main__":
    main()
