"""Build recipes for the native libraries (in-tree, sm_100a only).

  libvkr.so         CUDA kernels + the C-ABI of include/vkr.h        (nvcc)
  libvkr_engine.so  C++ host engine mirroring VulkanEngine on libvkr  (g++)

`python -m vk_renderer_b200.build` builds both.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    # host code feeds the bit-exact position path: never contract a*b+c on the host side
    "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall,-Wno-unknown-pragmas", "-shared",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _sources(dirs, exts):
    out = []
    for d in dirs:
        for r, _, fs in os.walk(d):
            out += [os.path.join(r, f) for f in fs if f.endswith(exts)]
    return out


def build_libvkr(force=False, verbose=False):
    target = os.path.join(PKG, "libvkr.so")
    deps = _sources([CSRC, os.path.join(ROOT, "include")], (".cu", ".cuh", ".h"))
    if not force and _newer(target, deps):
        return target
    # -lnccl: the native sort-first sharding (vkr_shard_init).  In a process that has imported torch the loader
    # resolves libnccl.so.2 to the copy torch already mapped; a plain C++ host gets the system library.
    # (written beside the target and moved into place: a snapshot of the tree never sees a half-written library)
    cmd = ["nvcc", *NVCC_FLAGS, "-o", target + ".tmp", os.path.join(CSRC, "vkr_api.cu"), "-lnccl"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.run(cmd, check=True)
    os.replace(target + ".tmp", target)
    return target


def build_engine(force=False):
    target = os.path.join(PKG, "libvkr_engine.so")
    host = os.path.join(CSRC, "host")
    if not os.path.isdir(host):
        return None
    deps = _sources([host, os.path.join(ROOT, "include")], (".cpp", ".h"))
    if not force and _newer(target, deps):
        return target
    srcs = [s for s in deps if s.endswith(".cpp")]
    cmd = ["g++", "-O2", "-std=c++20", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-I", os.path.join(ROOT, "include"),
           "-o", target + ".tmp", *srcs, "-L", PKG, "-lvkr", "-Wl,-rpath,$ORIGIN", "-lpthread", "-lz", "-ldl"]   # zlib: PNG images of glTF files, KTX2 levels; dl: libzstd looked up when a Zstandard KTX2 file is met
    subprocess.run(cmd, check=True)
    os.replace(target + ".tmp", target)
    return target


def build_cli(force=False):
    """vkr_render: the headless stand-in for the reference's window + ImGui panel (csrc/cli)."""
    target = os.path.join(PKG, "vkr_render")
    src = os.path.join(CSRC, "cli", "vkr_render.cpp")
    if not os.path.exists(src):
        return None
    deps = _sources([os.path.join(CSRC, "host"), os.path.join(CSRC, "cli"), os.path.join(ROOT, "include")], (".cpp", ".h"))
    if not force and _newer(target, deps):
        return target
    cmd = ["g++", "-O2", "-std=c++20", "-ffp-contract=off", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", target, src,
           "-L", PKG, "-lvkr_engine", "-lvkr", "-Wl,-rpath,$ORIGIN", "-lpthread"]
    subprocess.run(cmd, check=True)
    return target


def build_all(force=False, verbose=False):
    build_libvkr(force, verbose)
    build_engine(force)
    build_cli(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
