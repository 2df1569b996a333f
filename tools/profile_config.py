#!/usr/bin/env python
"""Per-pass device times of one BASELINE config through the host engine.  usage: python tools/profile_config.py 3 [frames]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vk_renderer_b200
from vk_renderer_b200 import capi
from vk_renderer_b200.engine import Engine

cfg = int(sys.argv[1]); frames = int(sys.argv[2]) if len(sys.argv) > 2 else 5
w, h = {1: (1280, 720), 2: (1920, 1080), 3: (3840, 2160), 4: (3840, 2160)}[cfg]
e = Engine(w, h, flags=capi.CONFIG_TIMINGS)
e.load_procedural(cfg)
lib = vk_renderer_b200.load()
dev = e.lib.vke_device(e.e)
lib.vkr_get_timings.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_int]
lib.vkr_cull_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
for f in range(frames):
    e.update_scene(0.0); e.draw(); e.sync()
    arr = (C.c_float * len(capi.T_NAMES))()
    lib.vkr_get_timings(dev, arr, len(capi.T_NAMES))
    print({k: round(v, 3) for k, v in zip(capi.T_NAMES, arr)})
cs = (C.c_uint32 * 4)(); lib.vkr_cull_stats(dev, cs); print("clusters main/culled shadow/culled", list(cs))
st = e.stats(); print("tris", st.triangle_count, "draws", st.drawcall_count)
