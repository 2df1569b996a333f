"""Render the bench scene (or a test scene) on the GPU and save a downscaled PNG + per-pass timings."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from PIL import Image
import vk_renderer_b200
from vk_renderer_b200 import capi
import bench

out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "frame.png")
lib = vk_renderer_b200.load()
r = capi.Renderer(lib, "vkr", bench.W, bench.H, shadow_extent=bench.S, flags=capi.CONFIG_TIMINGS)
sd, objs, fp, info = bench.build_scene(r)
for _ in range(3):
    r.draw(sd, objs, fp)
r.sync()
print(json.dumps(r.timings()))
img = r.readback(capi.TARGET_RGBA8)
prim = r.readback(capi.TARGET_PRIMID)
print("covered", float((prim != 0xFFFFFFFF).mean()), "unique prims", int(len(np.unique(prim))))
Image.fromarray(img[..., :3]).resize((960, 540), Image.BILINEAR).save(out)
