import ctypes
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def build_oracle():
    """CPU oracle (oracle/ork.cpp) -> oracle/_build/libork.so.  Test infrastructure only."""
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    return os.path.join(ROOT, "oracle", "_build", "libork.so")


def build_emu():
    """Host emulation of the kernels' per-thread bodies (tests/emu) -> tests/_build/libvkr_emu.so."""
    src = os.path.join(ROOT, "tests", "emu", "vkr_emu.cpp")
    out = os.path.join(ROOT, "tests", "_build", "libvkr_emu.so")
    deps = [src] + [os.path.join(ROOT, "vk_renderer_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "vk_renderer_b200", "csrc"))
                    if f.endswith((".h", ".cuh"))]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-o", out, src], check=True)
    return out


@pytest.fixture(scope="session")
def ork_lib():
    return ctypes.CDLL(build_oracle())


@pytest.fixture(scope="session")
def emu_lib():
    return ctypes.CDLL(build_emu())


@pytest.fixture(scope="session")
def vkr_lib():
    import vk_renderer_b200
    return vk_renderer_b200.load()
