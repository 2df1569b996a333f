"""Where does the end-to-end frame time go?  Variants of bench.py's e2e loop on config 4 (one GPU):
draw only / + RGB8 read-back / + RGBA8 read-back / device->host copies alone while the renderer is busy."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vk_renderer_b200
from vk_renderer_b200 import capi, meshgen as mg

W, H, S = bench.W, bench.H, bench.S
lib = vk_renderer_b200.load()
torch.cuda.set_device(0)
torch.cuda.set_stream(torch.cuda.Stream())
r = capi.Renderer(lib, "vkr", W, H, shadow_extent=S, device=0, stream=torch.cuda.current_stream().cuda_stream, flags=bench.BENCH_FLAGS)
sd, objs, fp, info = bench.build_scene(r)
view_b = mg.camera_view((0.6, 7.2, 40.0), pitch=-0.17)
sd_b = capi.make_scene_data(view_b, mg.camera_projection(W / H), [mg.light_matrix()] * 4, shadow_mode=1, enable_shadows=True)
cams = [sd, sd_b]
N = 200
pin3 = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
pin4 = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
for _ in range(10):
    r.draw(sd, objs, fp)
r.sync()

def loop(name, target=None, pins=None, nbytes=0):
    torch.cuda.synchronize(); r.sync()
    t0 = time.perf_counter()
    for i in range(N):
        r.draw(cams[i % 2], objs, fp)
        if target is not None:
            r.readback_async(target, pins[i % 2].data_ptr(), nbytes)
    r.sync()
    dt = time.perf_counter() - t0
    print(f"{name}: {N / dt:.1f} frames/s ({dt / N * 1e3:.3f} ms/frame)", flush=True)

loop("draw only")
r.set_flags(capi.CONFIG_TIMINGS | capi.CONFIG_SUBMIT_THREAD | capi.CONFIG_TRANSIENT_DEPTH)
loop("draw + RGB8 read-back, per-pass events on", capi.TARGET_RGB8, pin3, W * H * 3)
print({k: round(v, 3) for k, v in r.timings().items()}, flush=True)
loop("draw + RGBA8 read-back, per-pass events on", capi.TARGET_RGBA8, pin4, W * H * 4)
print({k: round(v, 3) for k, v in r.timings().items()}, flush=True)
r.set_flags(bench.BENCH_FLAGS)
loop("draw + RGB8 read-back", capi.TARGET_RGB8, pin3, W * H * 3)
loop("draw + RGBA8 read-back", capi.TARGET_RGBA8, pin4, W * H * 4)
# copies alone, idle device and busy device
src = torch.empty(W * H * 3, dtype=torch.uint8, device="cuda")
cs = torch.cuda.Stream()
def copies(k):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(cs):
        e0.record()
        for i in range(k):
            pin3[i % 2].view(-1).copy_(src, non_blocking=True)
        e1.record()
    return e0, e1
e0, e1 = copies(20); torch.cuda.synchronize()
print(f"D2H alone: {20 * W * H * 3 / (e0.elapsed_time(e1) / 1e3) / 1e9:.1f} GB/s", flush=True)
for i in range(60):
    r.draw(cams[i % 2], objs, fp)
e0, e1 = copies(20)
for i in range(60):
    r.draw(cams[i % 2], objs, fp)
r.sync(); torch.cuda.synchronize()
print(f"D2H while rendering: {20 * W * H * 3 / (e0.elapsed_time(e1) / 1e3) / 1e9:.1f} GB/s", flush=True)
r.close()
