"""vkr_render (csrc/cli): the headless replacement of the reference's window + ImGui panel."""
import os
import subprocess

import numpy as np
import pytest

from vk_renderer_b200 import capi
from vk_renderer_b200.engine import Engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "vk_renderer_b200", "vkr_render")


def read_ppm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"P6"
        w, h = map(int, f.readline().split())
        assert f.readline().strip() == b"255"
        return np.frombuffer(f.read(), np.uint8).reshape(h, w, 3)


def test_cli_usage_and_no_device_error():
    assert os.path.exists(CLI), "run __graft_entry__.build()"
    out = subprocess.run([CLI, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "--shadow-mode" in out.stdout and "--gltf" in out.stdout
    assert subprocess.run([CLI], capture_output=True).returncode == 2                       # neither --gltf nor --config
    assert subprocess.run([CLI, "--config", "1", "--bogus"], capture_output=True).returncode == 2
    import torch
    if not torch.cuda.is_available():     # no CPU fallback: a clear message and a non-zero exit
        out = subprocess.run([CLI, "--config", "1", "--size", "64x64"], capture_output=True, text=True)
        assert out.returncode == 1 and "no CPU path" in out.stderr


@pytest.mark.gpu
def test_cli_renders_what_the_engine_renders(vkr_lib, tmp_path):
    ppm = tmp_path / "c2.ppm"
    out = subprocess.run([CLI, "--config", "2", "--size", "480x270", "--shadow-extent", "512", "--frames", "2", "--out", str(ppm)],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    stats = {l.split(" ", 1)[0]: l.split(" ", 1)[1] for l in out.stdout.strip().splitlines()}
    e = Engine(480, 270, shadow_extent=512)
    tris = e.load_procedural(2)
    e.update_scene(0.0)
    e.draw()
    want = e.readback(capi.TARGET_RGBA8)[..., :3]
    st = e.stats()
    e.close()
    assert np.array_equal(read_ppm(ppm), want)
    assert int(stats["triangles"]) == st.triangle_count and int(stats["draws"]) == st.drawcall_count and st.triangle_count <= tris


@pytest.mark.gpu
def test_cli_gpu_scene_update_renders_the_same_image(vkr_lib, tmp_path):
    """--gpu-scene: update_scene's per-object work on the device (vkr_draw_scene); same image, same counts."""
    outs = []
    for k, extra in enumerate(([], ["--gpu-scene"])):
        ppm = tmp_path / f"c1_{k}.ppm"
        out = subprocess.run([CLI, "--config", "1", "--size", "480x270", "--camera", "0,1,4,0,0", "--frames", "2", "--out", str(ppm)] + extra,
                             capture_output=True, text=True)
        assert out.returncode == 0, out.stderr
        stats = {l.split(" ", 1)[0]: l.split(" ", 1)[1] for l in out.stdout.strip().splitlines()}
        outs.append((read_ppm(ppm), int(stats["triangles"]), int(stats["draws"])))
    assert np.array_equal(outs[0][0], outs[1][0])
    assert outs[0][1:] == outs[1][1:] and 0 < outs[0][2] < 64        # inside the grid: is_visible rejects part of it


@pytest.mark.gpu
def test_cli_loads_gltf(vkr_lib, tmp_path):
    import gltf_util as gu
    asset = gu.build_asset(np.random.default_rng(11))
    path = gu.write_gltf(asset, str(tmp_path), "glb")
    ppm = tmp_path / "g.ppm"
    out = subprocess.run([CLI, "--gltf", path, "--size", "320x200", "--shadow-extent", "256", "--camera", "-1,1.5,9,-0.1,0", "--out", str(ppm)],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    e = Engine(320, 200, shadow_extent=256)
    e.load_gltf(path, "scene")
    e.set_camera((-1.0, 1.5, 9.0), pitch=-0.1)
    e.set_toggles(shadows=1, shadow_mode=1, background=1, postprocess=1)
    e.update_scene(0.0)
    e.draw()
    want = e.readback(capi.TARGET_RGBA8)[..., :3]
    e.close()
    assert np.array_equal(read_ppm(ppm), want)
    bad = subprocess.run([CLI, "--gltf", str(tmp_path / "missing.glb")], capture_output=True, text=True)
    assert bad.returncode == 1 and "load_gltf" in bad.stderr
