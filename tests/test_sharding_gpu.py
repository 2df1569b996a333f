"""Sort-first sharding on real GPUs over NCCL (needs >= 2 devices; skipped on a single-GPU box):
the frame assembled from the ranks' bands — including the sharded, all-gathered shadow map — must equal
the single-GPU frame bit for bit."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _worker(rank, world, port, scene, W, H, S, out_dir, native):
    if native == "nccl":
        os.environ["VKR_EXCHANGE"] = "nccl"
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist
    import scenes
    import vk_renderer_b200
    from vk_renderer_b200 import capi
    from vk_renderer_b200.sharding import TileSharder
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    torch.cuda.set_stream(torch.cuda.Stream())
    lib = vk_renderer_b200.load()
    r = capi.Renderer(lib, "vkr", W, H, shadow_extent=S, device=rank, stream=torch.cuda.current_stream().cuda_stream)
    sd, objs, fp = scenes.SCENES[scene](r)
    if native:      # libvkr's own NCCL sharding (vkr_shard_init); torch.distributed only carries the communicator ids
        ids = torch.zeros(2 * capi.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            ids.copy_(torch.frombuffer(bytearray(r.comm_unique_ids()), dtype=torch.uint8))
        dist.broadcast(ids, 0)
        r.shard_init(rank, world, ids.cpu().numpy().tobytes())
        for _ in range(3):
            r.draw(sd, objs, fp)
        info = r.shard_info()
        assert info["ranks"] == world and info["peer_memory"] == (native != "nccl"), info
        if rank == 0 or not info["peer_memory"]:
            img = r.readback(capi.TARGET_RGBA8_GATHERED)      # over peer memory only rank 0 presents the assembled frame
            if rank == 0:
                assert np.array_equal(r.readback(capi.TARGET_RGB8), img[..., :3]), "VKR_TARGET_RGB8 of a sharded device is the assembled frame"
        else:
            try:
                r.readback(capi.TARGET_RGBA8_GATHERED)
                raise AssertionError("only rank 0 can read the assembled frame back")
            except capi.VkrError:
                pass
            img = np.zeros((H, W, 4), np.uint8)
        # every rank's own rows, through the band read-back (into one whole-frame buffer, as a shared host frame would be used)
        y0, y1 = info["rows"]
        land = torch.zeros((H, W, 3), dtype=torch.uint8).pin_memory()
        r.readback_band_async(capi.TARGET_RGB8, land.data_ptr(), H * W * 3)
        r.sync()
        own = r.readback(capi.TARGET_RGBA8)
        assert np.array_equal(land.numpy()[y0:y1], own[y0:y1, :, :3]) and not land.numpy()[:y0].any() and not land.numpy()[y1:].any()
        np.save(os.path.join(out_dir, f"band{rank}.npy"), np.concatenate([np.array([[y0, y1]]).repeat(1, 0).reshape(-1), own[y0:y1].reshape(-1)]))
        dist.barrier()
    else:           # the same sharding driven from Python over torch.distributed (vk_renderer_b200/sharding.py)
        sh = TileSharder(r, rank, world, cuda=True, shadow_group=dist.new_group())
        fp = sh.frame_params(fp)
        for _ in range(3):
            r.draw(sd, objs, fp)
            buf = sh.gather(async_op=True)
        sh.wait()
        torch.cuda.synchronize()
        img = sh.image(buf).cpu().numpy()
    np.save(os.path.join(out_dir, f"img{rank}.npy"), img)
    np.save(os.path.join(out_dir, f"shadow{rank}.npy"), r.readback(capi.TARGET_SHADOW))
    r.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("scene,native", [("lattice", True), ("basic_csm", True), ("lattice", "nccl"), ("lattice", False)])
def test_nccl_sharded_frame_equals_single_gpu(vkr_lib, tmp_path, scene, native):
    import torch
    import torch.multiprocessing as mp
    import scenes
    from vk_renderer_b200 import capi
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    W, H, S = 640, 360, 1024 if world != 3 else 1023
    r = capi.Renderer(vkr_lib, "vkr", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES[scene](r)
    r.draw(sd, objs, fp)
    want = r.readback(capi.TARGET_RGBA8)
    want_shadow = r.readback(capi.TARGET_SHADOW)
    layers = 4 if fp.shadowMode == 3 else 1
    r.close()
    port = 29700 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, scene, W, H, S, str(tmp_path), native), nprocs=world, join=True)
    for rank in range(world):
        got = np.load(tmp_path / f"img{rank}.npy")
        if native is True and rank > 0:      # peer memory: rank 0 presents; the other ranks hold (and can read back) their own rows
            band = np.load(tmp_path / f"band{rank}.npy")
            y0, y1 = int(band[0]), int(band[1])
            assert np.array_equal(band[2:].reshape(y1 - y0, W, 4).astype(np.uint8), want[y0:y1]), f"rank {rank}: rows {y0}..{y1} differ"
        else:
            assert np.array_equal(got, want), f"rank {rank}: {(got != want).any(-1).sum()} pixels differ"
        got_shadow = np.load(tmp_path / f"shadow{rank}.npy")
        assert np.array_equal(got_shadow[:layers].view(np.uint32), want_shadow[:layers].view(np.uint32))


def test_frame_parallel_views_on_the_device(vkr_lib):
    """ViewSharder's CUDA path (draw + asynchronous read-back into two pinned buffers) for world = 1 and for rank 1 of 2:
    every view it hands to the sink equals that view drawn and read back on its own."""
    import scenes
    from vk_renderer_b200 import capi
    from vk_renderer_b200.sharding import ViewSharder, view_slice
    W, H, S, n_views = 256, 144, 256, 5
    r = capi.Renderer(vkr_lib, "vkr", W, H, shadow_extent=S)
    sd, objs, fp = scenes.SCENES["basic_pcf"](r)
    views = scenes.orbit_views(r, n_views)
    want = []
    for v in range(n_views):
        r.draw(views[v], objs, fp)
        want.append(r.readback(capi.TARGET_RGBA8))
    for rank, world in ((0, 1), (1, 2)):
        got = {}
        n = ViewSharder(r, rank, world, cuda=True).render(views, objs, fp, lambda v, img: got.__setitem__(v, img.copy()))
        assert n == len(got) and sorted(got) == list(view_slice(rank, world, n_views))
        for v, img in got.items():
            assert np.array_equal(img, want[v]), f"view {v}"
        assert ViewSharder(r, rank, world, cuda=True).render(views, objs, fp) == n      # throughput form: no sink
    r.close()


def test_config5_view_at_1080p_matches_oracle(ork_lib, vkr_lib):
    """BASELINE config 5 at its own size: views of the 511 k-triangle scene at 1920x1080 (2048^2 PCF shadow, tonemap), rendered
    and read back by ViewSharder (rank 1 of 4: views 1, 5, 9, ...), against the oracle's frames of the same views."""
    sys.path.insert(0, ROOT)
    import bench
    import parity
    from vk_renderer_b200 import capi
    from vk_renderer_b200.sharding import ViewSharder
    ro = capi.Renderer(ork_lib, "ork", bench.VW, bench.VH, shadow_extent=bench.S)
    views, objs, fp, tris = bench.build_view_scene(ro)
    picks = [1, 69]                               # two of rank 1's views, a quarter orbit apart
    want = {}
    for v in picks:
        ro.draw(views[v], objs, fp)
        want[v] = {k: ro.readback(t) for k, t in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID,
                                                       shadow=capi.TARGET_SHADOW).items()}
        want[v].update(stats=ro.stats(), shadows=True, layers=1)
    ro.close()
    r = capi.Renderer(vkr_lib, "vkr", bench.VW, bench.VH, shadow_extent=bench.S)
    views, objs, fp, tris2 = bench.build_view_scene(r)
    assert tris == tris2 and 500_000 < tris < 520_000
    sub = [views[v] for v in range(0, 72)]
    got = {}
    ViewSharder(r, 1, 4, cuda=True).render(sub, objs, fp, lambda v, img: got.__setitem__(v, img.copy()) if v in picks else None)
    assert sorted(got) == picks
    for v in picks:                               # the frames the sharder delivered, then the full parity bar on the same view
        d = np.abs(got[v].astype(np.int32) - want[v]["rgba"].astype(np.int32))[..., :3].max(-1)
        assert (d <= 1).mean() >= 0.999, f"view {v}: only {(d <= 1).mean():.5f} of pixels within 1/255"
        r.draw(views[v], objs, fp)
        mine = {k: r.readback(t) for k, t in dict(rgba=capi.TARGET_RGBA8, depth=capi.TARGET_DEPTH, prim=capi.TARGET_PRIMID,
                                                   shadow=capi.TARGET_SHADOW).items()}
        mine.update(stats=r.stats(), shadows=True, layers=1)
        assert np.array_equal(mine["rgba"], got[v])
        parity.assert_parity(want[v], mine)
    r.close()
