"""vk_renderer_b200 — B200-native raster + shading path of mzwang34/vk_renderer.

The product is the C-ABI library `libvkr.so` (include/vkr.h; CUDA sm_100a kernels)
and the C++ host engine `libvkr_engine.so` that mirrors VulkanEngine on top of it.
This Python package is only the loader/binding used by tests and bench.py.

There is no CPU fallback: `load()` raises if the CUDA library is missing.
"""
from __future__ import annotations

import ctypes
import os

from . import capi  # noqa: F401

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
LIBVKR_PATH = os.path.join(PKG_DIR, "libvkr.so")
LIBENGINE_PATH = os.path.join(PKG_DIR, "libvkr_engine.so")

_lib = None


def load() -> ctypes.CDLL:
    """dlopen libvkr.so. Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIBVKR_PATH):
            raise ImportError(
                f"{LIBVKR_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU fallback."
            )
        # libvkr links libnccl.so.2 (native sort-first sharding).  PyTorch ships its own copy of that soname and refuses
        # to start on an older one, so when torch is installed let it map its NCCL first; a C++ host gets the system's.
        try:
            import torch  # noqa: F401
        except ImportError:
            pass
        _lib = ctypes.CDLL(LIBVKR_PATH, mode=ctypes.RTLD_GLOBAL)
    return _lib


def Renderer(width: int, height: int, **kw) -> "capi.Renderer":
    """A VkrDevice on the CUDA library."""
    return capi.Renderer(load(), "vkr", width, height, **kw)
