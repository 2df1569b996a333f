"""CPU-only: the kernels' per-thread bodies (host emulation, tests/emu) against the oracle.
Catches logic errors in vkr_raster.cuh / vkr_shade.cuh / vkr_frameprep.h without a GPU."""
import pytest

import parity
import scenes


@pytest.mark.parametrize("scene", sorted(scenes.SCENES))
def test_emu_matches_oracle(ork_lib, emu_lib, scene):
    ref = parity.render(ork_lib, "ork", scene, 192, 128, shadow_extent=256)
    got = parity.render(emu_lib, "emu", scene, 192, 128, shadow_extent=256)
    parity.assert_parity(ref, got)


def test_emu_scissor(ork_lib, emu_lib):
    sc = (40, 24, 100, 70)
    ref = parity.render(ork_lib, "ork", "basic_pcf", 192, 128, shadow_extent=256, scissor=sc)
    got = parity.render(emu_lib, "emu", "basic_pcf", 192, 128, shadow_extent=256, scissor=sc)
    parity.assert_parity(ref, got, region=sc)
    # nothing may be rasterised outside the scissor
    import numpy as np
    mask = np.ones((128, 192), bool)
    mask[sc[1]:sc[1] + sc[3], sc[0]:sc[0] + sc[2]] = False
    assert (got["prim"][mask] == 0xFFFFFFFF).all()


def test_emu_cluster_culling_is_exercised(ork_lib, emu_lib):
    """The lattice scene must actually trigger the cluster culls (frustum / no-sample / occlusion) in both
    passes, and — checked by test_emu_matches_oracle[lattice] — leave every pixel unchanged."""
    import ctypes
    from vk_renderer_b200 import capi
    r = capi.Renderer(emu_lib, "emu", 192, 128, shadow_extent=256)
    sd, objs, fp = scenes.SCENES["lattice"](r)
    r.draw(sd, objs, fp)
    out = (ctypes.c_uint32 * 2)()
    emu_lib.emu_culled(r.dev, out)
    total = sum((int(o["indexCount"]) // 3 + 127) // 128 for o in objs)
    assert out[0] > total // 4, f"main pass culled only {out[0]} of {total} clusters"
    assert out[1] > 0, "shadow pass culled nothing"
    r.close()


@pytest.mark.parametrize("seed", range(10))
def test_emu_culling_is_conservative_from_random_cameras(ork_lib, emu_lib, seed):
    """The cluster culls (frustum, no pixel centre, occlusion against the live depth) argue with margins; here
    they are tried from cameras a scene author would not pick (parity.random_camera_frames).  The oracle has no culling:
    every pixel, depth and shadow texel must still be identical."""
    parity.assert_random_camera_parity(ork_lib, "ork", emu_lib, "emu", seed)
