#!/usr/bin/env python
"""Frames/s of BASELINE configs 1-3 (and 4 for reference) through the C++ host engine (update_scene + draw per frame),
device-side pipelined: K frames issued back to back, one sync at the end.  Writes one JSON line per config.
    python tools/bench_configs.py [--frames 300]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vk_renderer_b200.engine import Engine  # noqa: E402

SIZES = {1: (1280, 720), 2: (1920, 1080), 3: (3840, 2160), 4: (3840, 2160), 6: (1920, 1080)}    # 6: the object-count case (16,000 draws)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--configs", default="1,2,3,4")
    ap.add_argument("--loop-flags", action="store_true",
                    help="the flags INTEGRATION.md recommends for a frame loop and bench.py uses: submit thread, transient depth and draw image")
    ap.add_argument("--gpu-scene", action="store_true",
                    help="update_scene's per-object work on the device (vkr_scene_create / vkr_draw_scene) instead of on the host")
    args = ap.parse_args()
    from vk_renderer_b200 import capi
    flags = (capi.CONFIG_SUBMIT_THREAD | capi.CONFIG_TRANSIENT_DEPTH | capi.CONFIG_TRANSIENT_HDR) if args.loop_flags else 0
    for c in [int(x) for x in args.configs.split(",")]:
        w, h = SIZES[c]
        e = Engine(w, h, flags=flags)
        tris = e.load_procedural(c)
        if args.gpu_scene:
            e.set_gpu_scene_update(True)
        for _ in range(10):
            e.update_scene(0.0); e.draw()
        e.sync()
        t0 = time.perf_counter()
        for _ in range(args.frames):
            e.update_scene(0.0); e.draw()
        e.sync()
        dt = time.perf_counter() - t0
        st = e.stats()
        visible = st.drawcall_count if args.gpu_scene else e.visible_objects()
        t1 = time.perf_counter()
        for _ in range(50):
            e.update_scene(0.0)
        host_update_ms = (time.perf_counter() - t1) / 50 * 1e3
        fps = args.frames / dt
        print(json.dumps({"config": c, "resolution": [w, h], "scene_triangles": int(tris), "visible_objects": visible,
                          "update_scene": "device" if args.gpu_scene else "host", "host_update_scene_ms": host_update_ms,
                          "submitted_triangles": st.triangle_count, "frames_per_s": fps, "mtri_per_s": fps * st.triangle_count / 1e6,
                          "gpix_per_s": fps * w * h / 1e9, "frames": args.frames,
                          "engine_flags": "SUBMIT_THREAD | TRANSIENT_DEPTH | TRANSIENT_HDR" if args.loop_flags else "0"}), flush=True)
        e.close()


if __name__ == "__main__":
    main()
