"""Sort-first sharding (vk_renderer_b200/sharding.py) over a world_size-2 and -3 gloo group on CPU:
every rank renders its row band of the frame and of the shadow map with the host emulation of the kernels,
the bands are all-gathered, and the assembled frame must equal the single-rank frame bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _worker(rank, world, port, scene, W, H, S, out_dir, shard_shadow):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import ctypes
    import conftest
    import scenes
    from vk_renderer_b200 import capi
    from vk_renderer_b200.sharding import TileSharder
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    emu = ctypes.CDLL(os.path.join(ROOT, "tests", "_build", "libvkr_emu.so"))
    r = capi.Renderer(emu, "emu", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES[scene](r)
    sh = TileSharder(r, rank, world, cuda=False, shard_shadow=shard_shadow)
    fp = sh.frame_params(fp)
    for _ in range(2):                       # two frames: exercises the double-buffered gather
        r.draw(sd, objs, fp)
        buf = sh.gather(async_op=True)
    sh.wait()
    img = sh.image(buf).numpy().copy()
    shadow = r.readback(capi.TARGET_SHADOW)
    np.save(os.path.join(out_dir, f"img{rank}.npy"), img)
    np.save(os.path.join(out_dir, f"shadow{rank}.npy"), shadow)
    r.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,scene,shard_shadow", [(2, "basic_pcf", True), (3, "lattice", True), (2, "basic_csm", True), (2, "transparent", False)])
def test_sharded_frame_equals_single_rank(emu_lib, tmp_path, world, scene, shard_shadow):
    import scenes
    from vk_renderer_b200 import capi
    W, H, S = 160, 96 if world == 2 else 100, 192       # 100 rows over 3 ranks: uneven bands; 192 % 3 == 0
    r = capi.Renderer(emu_lib, "emu", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES[scene](r)
    r.draw(sd, objs, fp)
    want = r.readback(capi.TARGET_RGBA8)
    want_shadow = r.readback(capi.TARGET_SHADOW)
    layers = 4 if fp.shadowMode == 3 else 1
    r.close()
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, scene, W, H, S, str(tmp_path), shard_shadow), nprocs=world, join=True)
    for rank in range(world):
        got = np.load(tmp_path / f"img{rank}.npy")
        assert np.array_equal(got, want), f"rank {rank}: assembled frame differs on {(got != want).any(-1).sum()} pixels"
        got_shadow = np.load(tmp_path / f"shadow{rank}.npy")
        if shard_shadow:     # every rank ends up with the complete shadow map
            assert np.array_equal(got_shadow[:layers].view(np.uint32), want_shadow[:layers].view(np.uint32))


def test_row_bands_cover_the_frame():
    from vk_renderer_b200.sharding import row_band
    for rows in (1, 7, 96, 2160, 2048):
        for world in (1, 2, 3, 4, 8, 16):
            bands = [row_band(r, world, rows) for r in range(world)]
            assert sum(h for _, h in bands) == rows
            y = 0
            for y0, h in bands:
                assert y0 == y or h == 0
                y += h


@pytest.mark.parametrize("scene,nranks", [("lattice", 4), ("basic_pcf", 3), ("basic_csm", 2), ("transparent", 5)])
def test_cost_balanced_bands_reassemble_the_frame(emu_lib, scene, nranks):
    """vkr_shard_init's split (plan_frame: one contiguous, cost-balanced band of frame rows and of shadow rows per
    rank, recomputed from the draw list): rendering every rank's band with the host emulation and stacking them must
    reproduce the unsharded frame and shadow map exactly; the bands must tile the rows; the split must be the same on
    every rank."""
    import ctypes
    import scenes
    from vk_renderer_b200 import capi
    W, H, S = 160, 100, 192
    r = capi.Renderer(emu_lib, "emu", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES[scene](r)
    r.draw(sd, objs, fp)
    want = r.readback(capi.TARGET_RGBA8)
    want_shadow = r.readback(capi.TARGET_SHADOW)
    layers = 4 if fp.shadowMode == 3 else 1
    r.close()
    full_shadow = np.ascontiguousarray(want_shadow)
    got = np.zeros_like(want)
    got_shadow = np.ones_like(want_shadow)
    splits = None
    for rank in range(nranks):
        r = capi.Renderer(emu_lib, "emu", W, H, shadow_extent=S)
        sd, objs, fp = scenes.SCENES[scene](r)
        emu_lib.emu_set_shard(r.dev, rank, nranks)
        own = {}

        def exchange(shadow_ptr, layer_bytes, nlayers, stream):
            buf = np.frombuffer((ctypes.c_ubyte * (layer_bytes * nlayers)).from_address(shadow_ptr), dtype=np.float32).reshape(nlayers, S, S)
            own["shadow"] = buf.copy()                # what this rank rasterised (its band; the rest is still cleared)
            buf[:] = full_shadow[:nlayers]            # the exchange: everybody's bands
        r.set_shadow_exchange(exchange)
        r.draw(sd, objs, fp)
        bands = (ctypes.c_int32 * (2 * nranks + 2))()
        assert emu_lib.emu_bands(r.dev, bands) == nranks + 1
        b = list(bands)
        if splits is None:
            splits = b
        assert b == splits
        fb, sb = b[:nranks + 1], b[nranks + 1:]
        assert fb[0] == 0 and fb[-1] == H and all(y1 > y0 for y0, y1 in zip(fb, fb[1:]))
        assert sb[0] == 0 and sb[-1] == S and all(y1 > y0 for y0, y1 in zip(sb, sb[1:]))
        img = r.readback(capi.TARGET_RGBA8)
        got[fb[rank]:fb[rank + 1]] = img[fb[rank]:fb[rank + 1]]
        got_shadow[:layers, sb[rank]:sb[rank + 1]] = own["shadow"][:layers, sb[rank]:sb[rank + 1]]
        # nothing may be rasterised outside the rank's shadow band
        mask = np.ones(S, bool); mask[sb[rank]:sb[rank + 1]] = False
        assert (own["shadow"][:layers, mask] == 1.0).all()
        r.close()
    assert np.array_equal(got, want), f"{(got != want).any(-1).sum()} pixels differ"
    assert np.array_equal(got_shadow[:layers].view(np.uint32), want_shadow[:layers].view(np.uint32))


# ---------------------------------------------------------------------------------------------------------------
# frame-parallel views (BASELINE config 5)
def _view_worker(rank, world, port, n_views, W, H, S, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import ctypes
    import scenes
    from vk_renderer_b200 import capi
    from vk_renderer_b200.sharding import ViewSharder
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    emu = ctypes.CDLL(os.path.join(ROOT, "tests", "_build", "libvkr_emu.so"))
    r = capi.Renderer(emu, "emu", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES["basic_pcf"](r)
    views = scenes.orbit_views(r, n_views)
    done = []

    def sink(v, img):
        np.save(os.path.join(out_dir, f"view{v}.npy"), img)
        done.append(v)
    n = ViewSharder(r, rank, world, cuda=False).render(views, objs, fp, sink)
    assert n == len(done)
    # who rendered what: every view exactly once over the group
    owners = [None] * world
    dist.all_gather_object(owners, done)
    assert sorted(v for o in owners for v in o) == list(range(n_views))
    r.close()
    dist.destroy_process_group()


def test_frame_parallel_views_equal_single_rank(emu_lib, tmp_path):
    import scenes
    from vk_renderer_b200 import capi
    from vk_renderer_b200.sharding import view_slice
    W, H, S, n_views, world = 128, 72, 128, 5, 2
    assert [list(view_slice(k, 3, 7)) for k in range(3)] == [[0, 3, 6], [1, 4], [2, 5]]
    assert list(view_slice(3, 4, 2)) == []                  # more ranks than views: nothing to do, no error
    port = 29500 + (os.getpid() % 2000) + 17
    mp.spawn(_view_worker, args=(world, port, n_views, W, H, S, str(tmp_path)), nprocs=world, join=True)
    r = capi.Renderer(emu_lib, "emu", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES["basic_pcf"](r)
    views = scenes.orbit_views(r, n_views)
    imgs = []
    for v in range(n_views):
        r.draw(views[v], objs, fp)
        want = r.readback(capi.TARGET_RGBA8)
        got = np.load(tmp_path / f"view{v}.npy")
        assert np.array_equal(got, want), f"view {v} differs"
        imgs.append(want)
    r.close()
    assert any(not np.array_equal(imgs[0], im) for im in imgs[1:])     # the views really are different images
