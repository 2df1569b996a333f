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


@pytest.mark.parametrize("seed", range(10))
def test_gpu_culling_is_conservative_from_random_cameras(ork_lib, vkr_lib, seed):
    """Cameras inside and around the lattice, clusters crossing the near plane and the frustum sides: the device's cluster
    culls must not change a pixel against the oracle, which has none."""
    parity.assert_random_camera_parity(ork_lib, "ork", vkr_lib, "vkr", seed)


def test_gpu_scissor(ork_lib, vkr_lib):
    sc = (64, 40, 160, 100)
    ref = parity.render(ork_lib, "ork", "basic_pcf", 320, 200, shadow_extent=512, scissor=sc)
    got = parity.render(vkr_lib, "vkr", "basic_pcf", 320, 200, shadow_extent=512, scissor=sc)
    parity.assert_parity(ref, got, region=sc)
    mask = np.ones((200, 320), bool)
    mask[sc[1]:sc[1] + sc[3], sc[0]:sc[0] + sc[2]] = False
    assert (got["prim"][mask] == 0xFFFFFFFF).all()


def test_gpu_odd_extent(ork_lib, vkr_lib):
    """An extent that is a multiple of nothing: tile edges, band edges and the tail of the packed RGB8 read-back."""
    ref = parity.render(ork_lib, "ork", "basic_pcf", 397, 211, shadow_extent=67)
    got = parity.render(vkr_lib, "vkr", "basic_pcf", 397, 211, shadow_extent=67)
    parity.assert_parity(ref, got)


def test_gpu_1080p_pcf(ork_lib, vkr_lib):
    ref = parity.render(ork_lib, "ork", "basic_pcf", 1920, 1080, shadow_extent=2048)
    got = parity.render(vkr_lib, "vkr", "basic_pcf", 1920, 1080, shadow_extent=2048)
    parity.assert_parity(ref, got)


def test_gpu_config4_full_size(ork_lib, vkr_lib):
    """BASELINE config 4 at its full size (10 M triangles in 1000 draws, 3840x2160, 2048^2 PCF shadow, skybox, tonemap —
    the frame bench.py times): parity with the oracle, and the size-independent properties of the path:
    idempotence (atomics and culling notwithstanding, a frame is a function of its inputs), scissor halves stitch to the
    frame, and the depth image does not depend on the submission order of the draws."""
    import sys
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from vk_renderer_b200 import capi
    targets = dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID, shadow=capi.TARGET_SHADOW)

    def grab(r, fp):
        out = {k: r.readback(v) for k, v in targets.items()}
        out["stats"], out["shadows"], out["layers"] = r.stats(), bool(fp.enableShadows), 1
        return out

    ro = capi.Renderer(ork_lib, "ork", bench.W, bench.H, shadow_extent=bench.S)
    sd, objs, fp, info = bench.build_scene(ro)
    assert info["tris"] == 10_000_000
    ro.draw(sd, objs, fp)
    ref = grab(ro, fp)
    ro.close()

    r = capi.Renderer(vkr_lib, "vkr", bench.W, bench.H, shadow_extent=bench.S)
    sd, objs, fp, info = bench.build_scene(r)
    r.draw(sd, objs, fp)
    got = grab(r, fp)
    frac, p = parity.assert_parity(ref, got)
    cull = r.cull_stats()
    assert cull["main_culled"] > 0 and cull["shadow_culled"] > 0        # the frame exercised the cluster culls

    r.draw(sd, objs, fp)                                                 # idempotence
    again = grab(r, fp)
    for k in ("rgba", "prim", "shadow"):
        assert np.array_equal(got[k], again[k]), k
    assert np.array_equal(got["depth"].view(np.uint32), again["depth"].view(np.uint32))

    stitched = {k: np.empty_like(got[k]) for k in ("rgba", "prim", "depth")}      # two scissor bands == the frame
    for y0, h in ((0, 1000), (1000, bench.H - 1000)):
        fp.scissor[:] = (0, y0, bench.W, h)
        r.draw(sd, objs, fp)
        for k in stitched:
            stitched[k][y0:y0 + h] = r.readback(targets[k])[y0:y0 + h]
    fp.scissor[:] = (0, 0, 0, 0)
    assert np.array_equal(stitched["prim"], got["prim"]) and np.array_equal(stitched["rgba"], got["rgba"])
    assert np.array_equal(stitched["depth"].view(np.uint32), got["depth"].view(np.uint32))

    r.draw(sd, objs[::-1].copy(), fp)                                    # submission order: same depth image
    rev = grab(r, fp)
    assert np.array_equal(rev["depth"].view(np.uint32), got["depth"].view(np.uint32))
    assert np.array_equal(rev["shadow"][:1].view(np.uint32), got["shadow"][:1].view(np.uint32))
    assert (np.abs(rev["rgba"].astype(np.int32) - got["rgba"].astype(np.int32)).max(-1) <= 1).mean() >= 0.999
    r.close()


def test_gpu_submit_thread_matches_synchronous_submission(vkr_lib):
    """VKR_CONFIG_SUBMIT_THREAD: frames planned on the caller's thread and issued by the submit thread are the same frames;
    an invalid draw list is still refused by the call that receives it; the device survives switching modes."""
    from vk_renderer_b200 import capi
    import scenes as sc
    frames = {}
    for flags in (0, capi.CONFIG_SUBMIT_THREAD):
        r = capi.Renderer(vkr_lib, "vkr", 320, 200, shadow_extent=256, flags=flags)
        sd, objs, fp = sc.SCENES["lattice"](r)
        views = sc.orbit_views(r, 6)
        out = []
        for k in range(12):                                   # back to back: the submit thread is busy while the next frame is planned
            r.draw(views[k % 6], objs, fp)
            if k >= 6:
                out.append((r.readback(capi.TARGET_RGBA8), r.readback(capi.TARGET_PRIMID)))
        if flags:
            # asynchronous read-backs ride the submit thread's queue behind their frame: the caller never waits for a frame to
            # be issued, and every landing buffer still holds exactly its own frame (both packed formats)
            import torch
            land = [torch.zeros((200, 320, 4), dtype=torch.uint8).pin_memory() for _ in range(6)]
            land3 = [torch.zeros((200, 320, 3), dtype=torch.uint8).pin_memory() for _ in range(6)]
            for k in range(6):
                r.draw(views[k], objs, fp)
                r.readback_async(capi.TARGET_RGBA8, land[k].data_ptr(), 320 * 200 * 4)
                r.readback_async(capi.TARGET_RGB8, land3[k].data_ptr(), 320 * 200 * 3)
            r.sync()
            for k in range(6):
                assert np.array_equal(land[k].numpy(), out[k][0]), f"asynchronously read-back frame {k} differs"
                assert np.array_equal(land3[k].numpy(), out[k][0][..., :3]), f"asynchronously read-back RGB8 frame {k} differs"
        bad = objs.copy()
        bad["mesh"][0] = 12345
        with pytest.raises(capi.VkrError):
            r.draw(views[0], bad, fp)
        r.draw(views[0], objs, fp)                            # still usable
        r.set_flags(0)
        r.draw(views[1], objs, fp)
        out.append((r.readback(capi.TARGET_RGBA8), r.readback(capi.TARGET_PRIMID)))
        frames[flags] = out
        r.close()
    for (a_rgba, a_prim), (b_rgba, b_prim) in zip(frames[0], frames[capi.CONFIG_SUBMIT_THREAD]):
        assert np.array_equal(a_prim, b_prim) and np.array_equal(a_rgba, b_rgba)


def test_gpu_transient_depth_fuses_the_clear(vkr_lib):
    """VKR_CONFIG_TRANSIENT_DEPTH: the shading pass leaves the visibility buffer cleared and the next frame skips its clear pass.
    Frames (changing views, a changing scissor rectangle, flag switched on and off between frames) are identical to the ones
    drawn with the separate clear; depth read-back is refused while the flag is set and correct again once it is cleared."""
    from vk_renderer_b200 import capi
    import scenes as sc
    r0 = capi.Renderer(vkr_lib, "vkr", 320, 200, shadow_extent=256, flags=0)
    r1 = capi.Renderer(vkr_lib, "vkr", 320, 200, shadow_extent=256, flags=capi.CONFIG_TRANSIENT_DEPTH)
    sd0, objs0, fp0 = sc.SCENES["lattice"](r0)
    sd1, objs1, fp1 = sc.SCENES["lattice"](r1)
    v0, v1 = sc.orbit_views(r0, 5), sc.orbit_views(r1, 5)
    rects = [(0, 0, 0, 0), (40, 30, 200, 120), (0, 0, 0, 0), (100, 0, 220, 200), (0, 50, 320, 100)]
    for k in range(10):
        for fp in (fp0, fp1):
            fp.scissor[:] = rects[k % 5]
        if k == 6:
            r1.set_flags(0)                       # this frame starts from the cleared buffer the previous one left
        if k == 8:
            r1.set_flags(capi.CONFIG_TRANSIENT_DEPTH)
        r0.draw(v0[k % 5], objs0, fp0)
        r1.draw(v1[k % 5], objs1, fp1)
        assert np.array_equal(r0.readback(capi.TARGET_RGBA8), r1.readback(capi.TARGET_RGBA8)), f"frame {k}"
        if 6 <= k < 8:
            assert np.array_equal(r0.readback(capi.TARGET_DEPTH).view(np.uint32), r1.readback(capi.TARGET_DEPTH).view(np.uint32))
            assert np.array_equal(r0.readback(capi.TARGET_PRIMID), r1.readback(capi.TARGET_PRIMID))
        else:
            with pytest.raises(capi.VkrError):
                r1.readback(capi.TARGET_DEPTH)
    r0.close(); r1.close()


def test_gpu_transient_hdr_keeps_the_presented_image(vkr_lib):
    """VKR_CONFIG_TRANSIENT_HDR: the RGBA16F draw image is not stored; the presented image (both f16 roundings applied in
    registers) is bit-identical, its read-back is refused while the flag is set and works again once it is cleared."""
    from vk_renderer_b200 import capi
    import scenes as sc
    for scene in ("basic_pcf", "transparent", "basic_sky_levels8"):
        r0 = capi.Renderer(vkr_lib, "vkr", 320, 200, shadow_extent=256, flags=0)
        r1 = capi.Renderer(vkr_lib, "vkr", 320, 200, shadow_extent=256, flags=capi.CONFIG_TRANSIENT_HDR | capi.CONFIG_TRANSIENT_DEPTH)
        a, b = sc.SCENES[scene](r0), sc.SCENES[scene](r1)
        for _ in range(2):
            r0.draw(*a)
            r1.draw(*b)
            assert np.array_equal(r0.readback(capi.TARGET_RGBA8), r1.readback(capi.TARGET_RGBA8)), scene
        with pytest.raises(capi.VkrError):
            r1.readback(capi.TARGET_HDR16F)
        r1.set_flags(0)
        r1.draw(*b)
        assert np.array_equal(r0.readback(capi.TARGET_HDR16F).view(np.uint16), r1.readback(capi.TARGET_HDR16F).view(np.uint16))
        r0.close(); r1.close()


def test_gpu_large_triangle_queue_overflow_is_repaired(ork_lib, tmp_path):
    """Queues of 4 triangles / 8 items (VKR_QUEUE_CAPS, read at vkr_create): the scene's screen-filling triangles overflow both.
    The device must not fault, and the frame that is read back must be the complete one (queues grown, frame redrawn)."""
    import subprocess, sys, os, textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = parity.render(ork_lib, "ork", "big_triangles", 320, 200, shadow_extent=128)
    np.save(tmp_path / "prim.npy", ref["prim"]); np.save(tmp_path / "depth.npy", ref["depth"]); np.save(tmp_path / "rgba.npy", ref["rgba"])
    code = textwrap.dedent(f"""
        import sys, numpy as np
        sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'tests')!r})
        import vk_renderer_b200, parity, scenes
        from vk_renderer_b200 import capi
        lib = vk_renderer_b200.load()
        for scene, shadow in (("big_triangles", 128), ("basic_pcf", 256), ("transparent", 256)):
            r = capi.Renderer(lib, "vkr", 320, 200, shadow_extent=shadow)
            r.draw(*scenes.SCENES[scene](r))
            got = {{k: r.readback(v) for k, v in dict(prim=capi.TARGET_PRIMID, depth=capi.TARGET_DEPTH, rgba=capi.TARGET_RGBA8).items()}}
            r.close()
            r = capi.Renderer(lib, "vkr", 320, 200, shadow_extent=shadow)      # and with the submit thread + an asynchronous read-back first
            import torch
            land = torch.zeros((200, 320, 4), dtype=torch.uint8).pin_memory()
            sd, objs, fp = scenes.SCENES[scene](r)
            r.set_flags(capi.CONFIG_SUBMIT_THREAD)
            r.draw(sd, objs, fp)
            r.readback_async(capi.TARGET_RGBA8, land.data_ptr(), 320 * 200 * 4)
            try:
                r.sync()
                reported = False
            except capi.VkrError as e:
                reported = e.code == capi.ERROR_CAPACITY
            assert reported, "an asynchronous read-back of an overflowed frame must be reported"
            r.draw(sd, objs, fp)
            r.readback_async(capi.TARGET_RGBA8, land.data_ptr(), 320 * 200 * 4)
            r.sync()
            assert np.array_equal(land.numpy(), got["rgba"])
            r.close()
            if scene == "big_triangles":
                for k in ("prim", "depth", "rgba"):
                    want = np.load({str(tmp_path)!r} + "/" + k + ".npy")
                    if k == "rgba":
                        assert (np.abs(want.astype(int) - got[k].astype(int))[..., :3].max(-1) <= 1).mean() >= 0.999
                    else:
                        assert np.array_equal(want.view(np.uint32), got[k].view(np.uint32)), k
        print("ok")
    """)
    env = dict(os.environ, VKR_QUEUE_CAPS="4,8")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "ok" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.parametrize("cta_clusters", ["0", "1000000"])
def test_gpu_both_geometry_kernel_forms_give_the_same_frames(vkr_lib, tmp_path, cta_clusters):
    """A pass with fewer than VKR_CTA_CLUSTERS clusters runs one CTA of four warps per cluster (geometry_cta_kernel_*), a larger
    one runs one warp per cluster (geometry_kernel<>): the small parity scenes above therefore only see the first form and the
    4K frame of config 4 only the second.  Here the threshold is forced to 0 (warp form everywhere) and to a million (CTA form
    everywhere) in a child process, and every buffer of every scene must equal, bit for bit, what the default gives."""
    import subprocess, sys, os, textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    names = ["edge_cases", "clipping", "lattice", "big_triangles", "cutout", "basic_csm", "transparent", "soup"]
    for name in names:
        got = parity.render(vkr_lib, "vkr", name, 320, 200, shadow_extent=256)
        np.savez(tmp_path / f"{name}.npz", prim=got["prim"], depth=got["depth"].view(np.uint32), shadow=got["shadow"].view(np.uint32), rgba=got["rgba"])
    code = textwrap.dedent(f"""
        import sys, numpy as np
        sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'tests')!r})
        import vk_renderer_b200, parity
        lib = vk_renderer_b200.load()
        for name in {names!r}:
            got = parity.render(lib, "vkr", name, 320, 200, shadow_extent=256)
            want = np.load({str(tmp_path)!r} + "/" + name + ".npz")
            assert np.array_equal(want["prim"], got["prim"]), name
            assert np.array_equal(want["depth"], got["depth"].view(np.uint32)), name
            assert np.array_equal(want["shadow"], got["shadow"].view(np.uint32)), name
            assert np.array_equal(want["rgba"], got["rgba"]), name
        print("ok")
    """)
    env = dict(os.environ, VKR_CTA_CLUSTERS=cta_clusters)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]


def test_gpu_host_times_are_reported(vkr_lib):
    """vkr_host_times: the host's own time per frame, apart from its waiting (bench.py's host_busy_ms_per_step)."""
    from vk_renderer_b200 import capi
    import scenes as sc
    r = capi.Renderer(vkr_lib, "vkr", 320, 200, shadow_extent=256, flags=capi.CONFIG_SUBMIT_THREAD)
    args = sc.SCENES["basic_pcf"](r)
    for _ in range(20):                      # frames before a device's 16th are never counted
        r.draw(*args)
    r.host_times(reset=True)
    for _ in range(10):
        r.draw(*args)
    r.sync()
    t = r.host_times()
    assert set(t) >= {"plan", "upload_expand", "shadow", "main_raster", "transparent_shade", "submit_thread_waiting_for_device"}
    assert all(v >= 0.0 for v in t.values()) and t["plan"] > 0.0 and t["main_raster"] > 0.0
    assert sum(t.values()) < 50000.0          # microseconds per frame: a 320x200 frame does not take 50 ms of host time
    r.close()
