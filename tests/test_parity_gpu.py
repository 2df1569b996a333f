"""GPU parity tests proper: libvkr.so (CUDA, through the C-ABI) against the oracle."""
import numpy as np
import pytest

import parity
import scenes

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("scene", sorted(scenes.SCENES))
def test_gpu_matches_oracle(ork_lib, vkr_lib, scene):
    ref = parity.render(ork_lib, "ork", scene, 320, 200, shadow_extent=512)
    got = parity.render(vkr_lib, "vkr", scene, 320, 200, shadow_extent=512)
    parity.assert_parity(ref, got)


def test_gpu_scissor(ork_lib, vkr_lib):
    sc = (64, 40, 160, 100)
    ref = parity.render(ork_lib, "ork", "basic_pcf", 320, 200, shadow_extent=512, scissor=sc)
    got = parity.render(vkr_lib, "vkr", "basic_pcf", 320, 200, shadow_extent=512, scissor=sc)
    parity.assert_parity(ref, got, region=sc)
    mask = np.ones((200, 320), bool)
    mask[sc[1]:sc[1] + sc[3], sc[0]:sc[0] + sc[2]] = False
    assert (got["prim"][mask] == 0xFFFFFFFF).all()


def test_gpu_1080p_pcf(ork_lib, vkr_lib):
    ref = parity.render(ork_lib, "ork", "basic_pcf", 1920, 1080, shadow_extent=2048)
    got = parity.render(vkr_lib, "vkr", "basic_pcf", 1920, 1080, shadow_extent=2048)
    parity.assert_parity(ref, got)
