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
