"""What one rank of an N-way sort-first split costs, measured on ONE GPU: config 4 with the frame's scissor set to each of N equal
row bands (and the shadow map's to the same fraction), frames drawn one at a time (per-pass event times, nothing pipelined).
The exchange itself is not part of it: this is the raster + shading work that does not shrink with N.
usage: python tools/band_probe.py [N=8] [frames=20]"""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vk_renderer_b200
from vk_renderer_b200 import capi
import bench

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 20
lib = vk_renderer_b200.load()
torch.cuda.set_stream(torch.cuda.Stream())
r = capi.Renderer(lib, "vkr", bench.W, bench.H, shadow_extent=bench.S, device=0, stream=torch.cuda.current_stream().cuda_stream,
                  flags=capi.CONFIG_TIMINGS | capi.CONFIG_TRANSIENT_DEPTH | capi.CONFIG_TRANSIENT_HDR)
sd, objs, fp, info = bench.build_scene(r)
rows = [(0, bench.H)] + [(bench.H * k // N, bench.H * (k + 1) // N) for k in range(N)]
for y0, y1 in rows:
    fp.scissor[:] = (0, y0, bench.W, y1 - y0)
    s0, s1 = bench.S * y0 // bench.H, bench.S * y1 // bench.H
    fp.shadowScissor[:] = (0, s0, bench.S, s1 - s0)
    acc = []
    for i in range(frames + 3):
        r.draw(sd, objs, fp)
        r.sync()
        if i >= 3:
            acc.append(r.timings())
    avg = {k: round(float(np.mean([a[k] for a in acc])), 4) for k in ("shadow", "geometry", "geom_kernel", "large_kernel", "shade", "total")}
    print(json.dumps({"rows": [y0, y1], **avg, "cull": r.cull_stats()}))
r.close()
