"""Shared parity check: a candidate backend against the oracle on the same scene."""
import numpy as np

from vk_renderer_b200 import capi
import scenes


def render(lib, prefix, scene, width, height, shadow_extent=512, scissor=None):
    r = capi.Renderer(lib, prefix, width, height, shadow_extent=shadow_extent)
    sd, objs, fp = scenes.SCENES[scene](r)
    if scissor is not None:
        fp.scissor[:] = scissor
    r.draw(sd, objs, fp)
    out = {k: r.readback(v) for k, v in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID,
                                             shadow=capi.TARGET_SHADOW, hdr=capi.TARGET_HDR16F, rgb=capi.TARGET_RGB8).items()}
    assert np.array_equal(out["rgb"], out["rgba"][..., :3]), "VKR_TARGET_RGB8 is the presented image without its alpha byte"
    out["stats"] = r.stats()
    out["shadows"] = bool(fp.enableShadows)
    out["layers"] = 4 if fp.shadowMode == 3 else 1
    r.close()
    return out


def psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def assert_parity(ref, got, region=None, rgb_frac=0.999, min_psnr=50.0):
    """Coverage (primitive id) and depth bit-exact; shadow map bit-exact; RGB8 within 1 LSB on
    >= 99.9 % of pixels and PSNR >= 50 dB (BASELINE.json north_star)."""
    sl = (slice(None), slice(None)) if region is None else (slice(region[1], region[1] + region[3]), slice(region[0], region[0] + region[2]))
    assert ref["stats"].triangle_count == got["stats"].triangle_count
    assert ref["stats"].drawcall_count == got["stats"].drawcall_count
    pa, pb = ref["prim"][sl], got["prim"][sl]
    assert (pa != 0xFFFFFFFF).any(), "scene rendered nothing"
    assert np.array_equal(pa, pb), f"coverage / depth-test winner differs on {(pa != pb).sum()} pixels"
    da, db = ref["depth"][sl].view(np.uint32), got["depth"][sl].view(np.uint32)
    assert np.array_equal(da, db), f"depth differs on {(da != db).sum()} pixels"
    if ref["shadows"]:
        n = ref["layers"]
        sa, sb = ref["shadow"][:n].view(np.uint32), got["shadow"][:n].view(np.uint32)
        assert (sa != 0x3F800000).any(), "shadow map is empty"
        assert np.array_equal(sa, sb), f"shadow depth differs on {(sa != sb).sum()} texels"
    ra, rb = ref["rgba"][sl].astype(np.int32), got["rgba"][sl].astype(np.int32)
    diff = np.abs(ra - rb)[..., :3].max(-1)
    frac = (diff <= 1).mean()
    p = psnr(ra[..., :3], rb[..., :3])
    assert frac >= rgb_frac, f"only {frac:.5f} of pixels within 1/255"
    assert p >= min_psnr, f"PSNR {p:.1f} dB"
    return frac, p


def random_camera_frames(lib, prefix, seed, width=160, height=96):
    """The lattice scene (27 instanced spheres) from a seeded camera: inside the lattice or around it, aimed at a point of it
    give or take 0.4 rad (seeds 0-7: clusters cross the near plane, w = 0 and the frustum sides) or anywhere (seeds 8, 9)."""
    from vk_renderer_b200 import meshgen as mg
    rng = np.random.default_rng(1000 + seed)
    pos = rng.uniform(-4.0, 4.0, 3) if seed % 2 else rng.uniform(-1.0, 1.0, 3) * np.array([9.0, 6.0, 9.0])
    if seed < 8:         # forward = (-sin yaw cos pitch, sin pitch, -cos yaw cos pitch)
        d = rng.uniform(-2.6, 2.6, 3) - pos
        d /= np.linalg.norm(d)
        pitch, yaw = float(np.arcsin(d[1]) + rng.uniform(-0.4, 0.4)), float(np.arctan2(-d[0], -d[2]) + rng.uniform(-0.4, 0.4))
    else:
        pitch, yaw = float(rng.uniform(-1.2, 1.2)), float(rng.uniform(-3.1, 3.1))
    r = capi.Renderer(lib, prefix, width, height, shadow_extent=128)
    sd, objs, fp = scenes.lattice(r, n=3, rings=12)
    view = mg.camera_view(tuple(pos), pitch=pitch, yaw=yaw)
    sd = capi.make_scene_data(view, mg.camera_projection(width / height), scenes._lights(1), shadow_mode=1, enable_shadows=True)
    r.draw(sd, objs, fp)
    out = {k: r.readback(v) for k, v in dict(depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID, shadow=capi.TARGET_SHADOW,
                                             rgba=capi.TARGET_RGBA8).items()}
    r.close()
    return out


def assert_random_camera_parity(ref_lib, ref_prefix, lib, prefix, seed):
    a, b = random_camera_frames(ref_lib, ref_prefix, seed), random_camera_frames(lib, prefix, seed)
    if seed < 8:
        assert (a["prim"] != 0xFFFFFFFF).mean() > 0.05, "the camera was meant to see the lattice"
    assert np.array_equal(a["prim"], b["prim"]), f"coverage differs on {(a['prim'] != b['prim']).sum()} pixels"
    assert np.array_equal(a["depth"].view(np.uint32), b["depth"].view(np.uint32))
    assert np.array_equal(a["shadow"][:1].view(np.uint32), b["shadow"][:1].view(np.uint32))
    assert (np.abs(a["rgba"].astype(np.int32) - b["rgba"].astype(np.int32))[..., :3].max(-1) <= 1).mean() >= 0.999
