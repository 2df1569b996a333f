#!/usr/bin/env python
"""bench.py — frames/s of the per-frame raster + shading path on BASELINE config 4:
a 4K (3840x2160) PBR + PCF-shadow + skybox + tonemap frame of a 10 M-triangle instanced scene
(one 10 k-triangle sphere x 1000 instances on a 10x10x10 lattice).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N=1: the whole frame on one B200.  N>1 (torchrun, one rank per GPU): sort-first row bands with
replicated geometry and shadow map, framebuffer bands gathered with an NCCL all-gather
("scaling": "strong": the frame is fixed, the pixels are split).
--impl reference: the CPU oracle (oracle/, the restatement of the reference's algorithm — the
reference itself needs a Vulkan driver, unavailable on this box) on all host cores.

One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from vk_renderer_b200 import capi, meshgen as mg  # noqa: E402

W, H, S = 3840, 2160, 2048
# submit thread: planning frame i+1 on the caller's thread overlaps issuing frame i's CUDA / NCCL calls
# TRANSIENT_DEPTH: nothing reads the depth attachment after the frame (as in the reference), so its clear is fused into the
# shading pass instead of being a pass of its own; the work is still done every frame
# TRANSIENT_HDR: nor the RGBA16F draw image once it has been tone-mapped and blitted: its texel stays in registers (the f16
# roundings are still applied); the presented 8-bit image — what the metric's tolerance is stated on — is written as before
BENCH_FLAGS = capi.CONFIG_TIME_GEOM_KERNEL | capi.CONFIG_SUBMIT_THREAD | capi.CONFIG_TRANSIENT_DEPTH | capi.CONFIG_TRANSIENT_HDR
SEED = 0xB2000000 + 4


# ------------------------------------------------------------------------------- workload
def build_scene(r, width=W, height=H):
    """BASELINE config 4: sphere rings=50, segments=100 -> 10,000 triangles, 1000 instances."""
    rng = np.random.default_rng(SEED)
    sv, si = mg.uv_sphere(50, 100, 1.0)
    mesh = r.upload_mesh(si, sv)
    r.set_cubemap(mg.sky_faces(512, seed=SEED & 0xFFFF), capi.FORMAT_SRGB, True)
    mats = []
    for _ in range(16):
        col = tuple(rng.uniform(0.2, 1.0, 3)) + (1.0,)
        mats.append(r.build_material(col, float(rng.random() < 0.3), float(rng.uniform(0.2, 0.9))))
    objs = []
    k = 0
    for ix in range(10):
        for iy in range(10):
            for iz in range(10):
                t = mg.translate((ix - 4.5) * 3.0, (iy - 4.5) * 3.0, (iz - 4.5) * 3.0)
                objs.append(mg.render_object(mesh, mats[k % 16], t, len(si)))
                k += 1
    objs = np.array(objs)
    view = mg.camera_view((0.0, 7.0, 40.0), pitch=-0.17)
    proj = mg.camera_projection(width / height)
    L = mg.light_matrix()
    sd = capi.make_scene_data(view, proj, [L] * 4, shadow_mode=1, enable_shadows=True)
    fp = capi.default_frame_params(enableShadows=1, shadowMode=1, enablePostprocess=1, enableSkybox=1)
    info = dict(tris=int(len(si) // 3 * len(objs)), verts_per_mesh=int(len(sv)), idx_per_mesh=int(len(si)), nobj=int(len(objs)))
    return sd, objs, fp, info


def algorithmic_bytes(info, width=W, height=H, shadow=S, layers=1):
    """B_alg = G_main + G_shadow + Z_shadow + R_shadow + Z_main + C_hdr + T_tex + O_rgb8."""
    n, ni, nv = info["nobj"], info["idx_per_mesh"], info["verts_per_mesh"]
    px = width * height
    g_main = n * (ni * 4 + nv * 48)
    g_shadow = layers * n * (ni * 4 + nv * 12)
    z_shadow = r_shadow = layers * shadow * shadow * 4
    frame = g_main + g_shadow + z_shadow + r_shadow + px * 4 + px * 8 + px * 8 + px * 4
    # dominant kernel (main-pass geometry_kernel): indices + SoA positions read once per instance,
    # visibility buffer (8 B/px) written once
    geom_kernel = n * (ni * 4 + nv * 16) + px * 8
    return frame, geom_kernel


# ------------------------------------------------------------------------------- helpers
class ClockSampler:
    def __init__(self, device_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={device_index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.p.wait()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for nme, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s in sm if s > 0.5 * max(sm)] or sm
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the newest committed ncu capture (profiles/*_traffic.json, written by
    tools/ncu_summary.py), or None."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")))
    if not files:
        return None, None
    d = json.load(open(files[-1]))
    return d["dram_bytes_per_launch"].get(kernel), f"{os.path.basename(files[-1])} (kernels of commit {d.get('kernel_source_commit', 'unrecorded')})"


def load_oracle():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    return ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "libork.so"))


def time_oracle(steps, warmup, budget_s=150.0):
    """The CPU restatement (all host cores) on the bench workload; one step = one full frame."""
    lib = load_oracle()
    r = capi.Renderer(lib, "ork", W, H, shadow_extent=S)
    sd, objs, fp, info = build_scene(r)
    cores = os.cpu_count() or 1
    times = []
    t_begin = time.time()
    for i in range(warmup + steps):
        t0 = time.time()
        r.draw(sd, objs, fp)
        dt = time.time() - t0
        if i >= warmup:
            times.append(dt)
        if time.time() - t_begin > budget_s and times:
            break
    r.close()
    return times, cores, info


# ------------------------------------------------------------------------------- arms
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times, cores, info = time_oracle(max(1, min(args.steps, 3)), min(args.warmup, 1))
    fps = len(times) / sum(times)
    line = {
        "impl": "reference", "metric": "frames_per_sec_4k_pbr_shadow", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": len(times), "warmup": min(args.warmup, 1), "ms_per_step": 1000.0 / fps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(info, args.gpus),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": f"{len(times)} full 4K frame(s) of the same scene, oracle/ork.cpp on {cores} threads "
                                   "(the reference's own renderer needs a Vulkan driver: unavailable on this box)"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mtri_per_s": fps * info["tris"] / 1e6, "gpix_per_s": fps * W * H / 1e9,
    }
    print(json.dumps(line), flush=True)


def workload_config(info, n_gpus):
    return {"workload": "BASELINE config 4: 10M-triangle instanced scene (10k-tri sphere x 1000), 3840x2160, PBR mesh_pbr.frag, "
                        "2048^2 PCF shadow, skybox, tonemap",
            "triangles": info["tris"], "objects": info["nobj"], "resolution": [W, H], "shadow_extent": S,
            "parallelism": "single GPU" if n_gpus == 1 else f"sort-first: {n_gpus} cost-balanced row bands of the frame and of the shadow map, geometry replicated; "
                                                                 "shadow rows and presented rows exchanged over peer memory (CUDA IPC stores over NVLink, flags) "
                                                                 "or, with VKR_EXCHANGE=nccl / without peer access, grouped ncclSend/ncclRecv (see `exchange`)",
            "l2": "inputs larger than L2, no explicit flush: the presented image and the shadow map are double-buffered across frames, so two consecutive "
                  "frames touch visibility 66 MB + 2 x RGBA8 33 MB + 2 x shadow 17 MB = 166 MB against a 126 MB L2 (the RGBA16F draw image is transient: "
                  "VKR_CONFIG_TRANSIENT_HDR); the 0.25 MB of instanced cluster blocks stay in L2, as they do in any frame loop"}


def run_b200(args):
    import torch
    import vk_renderer_b200

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = vk_renderer_b200.load()
    # one explicit stream for everything (kernels, NCCL, events): the legacy default stream has
    # handle 0, which VkrConfig.stream reads as "create your own"
    torch.cuda.set_stream(torch.cuda.Stream())
    stream = torch.cuda.current_stream().cuda_stream
    assert stream != 0
    # timed region: only the two events around the dominant kernel (the full per-pass set costs ~5 % of the frame rate);
    # the per-pass breakdown is taken from a few extra frames after the timed region
    r = capi.Renderer(lib, "vkr", W, H, shadow_extent=S, device=local, stream=stream, flags=BENCH_FLAGS)
    sd, objs, fp, info = build_scene(r)

    # ---- sort-first (native, include/vkr.h vkr_shard_init): this rank renders its rows of the frame and of the shadow
    # map; libvkr all-gathers the shadow bands inside the frame and the RGBA8 bands after it over NCCL, the latter
    # overlapped with the next frame.  torch.distributed is the plumbing: it carries the communicator ids.
    # Every rank keeps the whole draw list: the geometry pass's cluster test drops what falls outside the band.
    if world > 1:
        ids = torch.zeros(2 * capi.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            ids.copy_(torch.frombuffer(bytearray(r.comm_unique_ids()), dtype=torch.uint8))
        dist.broadcast(ids, 0)
        r.shard_init(rank, world, ids.cpu().numpy().tobytes())
    my_objs = objs

    def step():
        r.draw(sd, my_objs, fp)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warmup_run = max(3, args.warmup)                          # the timing rules ask for at least three
    for _ in range(warmup_run):
        step()
    r.sync()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    kern_ms, launches = [], 0
    t_issue = time.perf_counter()
    for _ in range(args.steps):
        step()
    launches = r.launch_count() * args.steps    # (also waits until the submit thread has issued the last frame: the end event
                                                #  below must be recorded after it)
    host_issue_ms = (time.perf_counter() - t_issue) * 1000.0 / args.steps     # host time to ISSUE a step (no device wait)
    if world > 1:
        r.sync()                                 # the last gather (on libvkr's gather stream) is part of the timed region
    e1.record()
    barrier()
    r.sync()
    ms = e0.elapsed_time(e1)
    gk_timed_region_ms = r.timings()["geom_kernel"]      # mean over the frames of the timed region (event pair per frame)
    r.set_flags(capi.CONFIG_TIMINGS | capi.CONFIG_SUBMIT_THREAD | capi.CONFIG_TRANSIENT_DEPTH | capi.CONFIG_TRANSIENT_HDR)
    for _ in range(3):
        step()
    r.sync()
    tm = r.timings()   # per-pass device times of a frame issued back to back with its neighbours
    per_rank = None
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        # every rank's own pass times of that frame: how even the cost-balanced bands came out
        keys = ["shadow", "geometry", "shade", "total", "shadow_exchange", "frame_gather"]
        mine = torch.tensor([tm[k] for k in keys], device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [{k: round(float(v), 4) for k, v in zip(keys, a.tolist())} for a in allr]
    clocks = sampler.stop() if sampler else None

    # ---- dominant-kernel duration, averaged over a few extra frames (events around the kernel itself)
    gk, host_call = [], []
    for _ in range(5):
        t0 = time.perf_counter()
        r.draw(sd, my_objs, fp)
        host_call.append((time.perf_counter() - t0) * 1000.0)      # vkr_draw on an idle device: pure host cost of one call
        r.sync()
        gk.append(r.timings())
    avg = {k: float(np.mean([g[k] for g in gk])) for k in gk[0]}
    cull = r.cull_stats()
    r.set_flags(BENCH_FLAGS)
    # ---- where the host's time goes (vkr_host_times): the caller's planning and the submit thread's issue time per frame,
    # apart from the time either spends waiting (for the other thread, or for the device to free a frame slot).
    # `host_issue_ms_per_step` above is the wall time of the issuing loop, which a device-bound run holds at the frame time.
    try:
        r.host_times(reset=True)
        for _ in range(max(20, min(args.steps, 100))):
            step()
        r.sync()
        host_us = r.host_times()
    except AttributeError:                       # an older library variant in an A/B run
        host_us = dict.fromkeys(("plan", "upload_expand", "shadow", "main_raster", "transparent_shade", "gather", "async_readbacks",
                                 "caller_waiting_for_submit_thread", "submit_thread_waiting_for_device"), float("nan"))
    host_busy = {"caller_plan_ms": host_us["plan"] / 1000.0,
                 "submit_thread_issue_ms": sum(host_us[k] for k in ("upload_expand", "shadow", "main_raster", "transparent_shade", "gather", "async_readbacks")) / 1000.0,
                 "caller_waiting_ms": host_us["caller_waiting_for_submit_thread"] / 1000.0,
                 "submit_thread_waiting_for_device_ms": host_us["submit_thread_waiting_for_device"] / 1000.0}

    # ---- e2e: host draw list in, RGB8 frame out to pinned host memory, every step (FRAME_OVERLAP=2 pipelining).
    # N = 1: the presented image of every frame.  N > 1: ONE host frame shared by the ranks (a /dev/shm segment every process maps
    # and registers with CUDA); every rank copies its own rows into it over its own PCIe link (vkr_readback_band_async,
    # VKR_CONFIG_LOCAL_PRESENT: nothing is assembled on rank 0 on this path), so the frame leaves the box N links wide.
    frame_bytes = W * H * 3
    shm_path, shm_map = None, None
    if world == 1:
        pinned = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
        host_frames = [p_.numpy() for p_ in pinned]
        host_ptrs = [p_.data_ptr() for p_ in pinned]
    else:
        shm_path = f"/dev/shm/vkr_bench_{os.environ.get('MASTER_PORT', '0')}"
        if rank == 0:
            shm_map = np.memmap(shm_path, dtype=np.uint8, mode="w+", shape=(2, H, W, 3))
            shm_map[:] = 0
        barrier()
        if rank != 0:
            shm_map = np.memmap(shm_path, dtype=np.uint8, mode="r+", shape=(2, H, W, 3))
        rc_reg = torch.cuda.cudart().cudaHostRegister(shm_map.ctypes.data, 2 * frame_bytes, 0)
        if int(rc_reg) != 0:
            raise SystemExit(f"bench.py: cudaHostRegister of the shared host frame failed ({rc_reg})")
        host_frames = [shm_map[0], shm_map[1]]
        host_ptrs = [shm_map.ctypes.data, shm_map.ctypes.data + frame_bytes]
        r.set_flags(BENCH_FLAGS | capi.CONFIG_LOCAL_PRESENT)
    n = len(my_objs)
    h2d = r.upload_bytes()                                  # what vkr_draw copied host->device for the last frame (draw records,
                                                            # cluster/slot prefixes, light MVPs, shading constants)
    d2h = W * H * 3                                         # the whole frame, over one link (N = 1) or split into the ranks' rows
    # two cameras alternate so that consecutive frames differ: a read-back that raced with a later frame's rendering
    # (or a gather that mixed frames) is caught by the check after the loop
    view_b = mg.camera_view((0.6, 7.2, 40.0), pitch=-0.17)
    sd_b = capi.make_scene_data(view_b, mg.camera_projection(W / H), [mg.light_matrix()] * 4, shadow_mode=1, enable_shadows=True)
    cams = [sd, sd_b]
    target = capi.TARGET_RGB8          # the presented frame, packed R,G,B (the assembled frame when sharded)
    def present(i):
        if world == 1:
            r.readback_async(target, host_ptrs[i % 2], frame_bytes)
        else:
            r.readback_band_async(target, host_ptrs[i % 2], frame_bytes)

    for i in range(max(3, args.warmup)):                     # untimed: the read-back path allocates its packed buffers on first use
        r.draw(cams[i % 2], my_objs, fp)
        present(i)
    r.sync()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        objs_frame = np.array(my_objs, copy=True)            # the host rebuilds its draw list every frame (update_scene)
        r.draw(cams[i % 2], objs_frame, fp)
        present(i)
    r.sync()
    barrier()
    e2e_s = time.perf_counter() - t0
    my_rows = r.shard_info()["rows"] if world > 1 else (0, H)
    if world > 1:
        r.set_flags(BENCH_FLAGS)
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    checksum = int(host_frames[(args.steps - 1) % 2].reshape(-1)[::4097].astype(np.int64).sum()) if rank == 0 else 0
    # what the bus allows: the same number of bytes, device -> pinned host, timed alone (the e2e frame rate cannot exceed
    # d2h_GBps / d2h_bytes_per_step)
    d2h_gbps = None
    if rank == 0:
        probe_src = torch.empty(W * H * 3, dtype=torch.uint8, device="cuda")
        probe_dst = torch.empty(W * H * 3, dtype=torch.uint8).pin_memory()
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        probe_dst.copy_(probe_src, non_blocking=True)
        torch.cuda.synchronize()
        pe0.record()
        for _ in range(8):
            probe_dst.copy_(probe_src, non_blocking=True)
        pe1.record()
        torch.cuda.synchronize()
        d2h_gbps = 8 * W * H * 3 / (pe0.elapsed_time(pe1) / 1000.0) / 1e9
        del probe_src, probe_dst
    e2e_ok = True
    for i in (args.steps - 2, args.steps - 1):               # the last two frames, re-rendered and read back synchronously
        if i < 0:
            continue
        r.draw(cams[i % 2], my_objs, fp)
        if rank == 0:                                        # (the assembled frame is rank 0's)
            want = r.readback(target)
            if not np.array_equal(want, host_frames[i % 2]):
                e2e_ok = False
        else:
            r.sync()
    if rank == 0 and not e2e_ok:
        raise SystemExit("bench.py: an asynchronously read-back frame differs from the same frame rendered synchronously")

    # ---- N > 1: the assembled frame and the exchanged shadow map against the SAME frame drawn by one unsharded device
    # (a second VkrDevice on rank 0's GPU), byte for byte
    import zlib
    sharded_equals_single, frame_crc = None, None
    r.draw(cams[0], my_objs, fp)
    frame_n = r.readback(target) if rank == 0 or world == 1 else None
    shadow_n = r.readback(capi.TARGET_SHADOW)[:1] if rank == 0 else None
    if rank == 0:
        frame_crc = zlib.crc32(frame_n.tobytes()) & 0xFFFFFFFF
    if world > 1:
        ok = torch.ones(1, device="cuda", dtype=torch.int32)
        if rank == 0:
            r1 = capi.Renderer(lib, "vkr", W, H, shadow_extent=S, device=local, stream=stream, flags=0)
            sd1, objs1, fp1, _ = build_scene(r1)
            r1.draw(cams[0], objs1, fp1)
            frame_1, shadow_1 = r1.readback(target), r1.readback(capi.TARGET_SHADOW)[:1]
            r1.close()
            sharded_equals_single = bool(np.array_equal(frame_1, frame_n) and np.array_equal(shadow_1.view(np.uint32), shadow_n.view(np.uint32)))
            if not sharded_equals_single:
                print(f"bench.py: sharded frame differs from the single-GPU frame on {(frame_1 != frame_n).any(-1).sum()} pixels, "
                      f"shadow map on {(shadow_1.view(np.uint32) != shadow_n.view(np.uint32)).sum()} texels", file=sys.stderr)
                ok.zero_()
        dist.broadcast(ok, 0)
        if int(ok.item()) == 0:
            r.close()
            dist.destroy_process_group()
            raise SystemExit(3)

    line = None
    if rank == 0:
        fps = args.steps / (ms / 1000.0)
        peak, peak_src = measured_peak()
        my_rows_px = (my_rows[1] - my_rows[0]) * W
        frame_bytes, gk_bytes = algorithmic_bytes(info)
        gk_basis = "SURVEY 8d: every instance's indices + positions once, visibility buffer once"
        if world > 1 and cull.get("main_clusters"):
            # one rank's launch: its band of the visibility buffer and the clusters that reach its band
            live = 1.0 - cull["main_culled"] / cull["main_clusters"]
            gk_bytes = int((gk_bytes - W * H * 8) * live + W * H * 8 / world)
            gk_basis = f"rank 0's launch: {live:.3f} of the clusters (the rest fall outside its band or are occluded) + 1/{world} of the visibility buffer"
        gk_ms = gk_timed_region_ms
        traffic, traffic_src = measured_traffic("geometry_kernel<0>")
        achieved = gk_bytes / (gk_ms / 1000.0) / 1e9 if gk_ms > 0 else 0.0
        line = {
            "metric": "frames_per_sec_4k_pbr_shadow", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": warmup_run, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(info, world),
            "mtri_per_s": fps * info["tris"] / 1e6, "gpix_per_s": fps * W * H / 1e9,
            "frame_roofline": {"bound": "hbm", "achieved": frame_bytes * fps / 1e9, "peak": peak, "unit": "GB/s",
                               "frac": frame_bytes * fps / 1e9 / peak, "algorithmic_bytes": frame_bytes,
                               "frac_of_nominal_8TBs": frame_bytes * fps / 1e9 / 8000.0},
            "roofline": {"bound": "hbm", "kernel": "geometry_kernel<main>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes": gk_bytes, "algorithmic_basis": gk_basis, "kernel_ms": gk_ms,
                         "peak_source": peak_src},
            # the other long kernel of the frame, for the record (same byte model as DESIGN.md 5: visibility key 8 + presented
            # texel 4 + material texels 8 per pixel + the shadow map once; its pass is that one kernel)
            "roofline_shade_kernel": (lambda b, t: {"bound": "hbm", "kernel": "shade_kernel", "algorithmic_bytes": b, "kernel_ms": t,
                                                    "achieved": b / (t / 1000.0) / 1e9 if t > 0 else 0.0, "peak": peak, "unit": "GB/s",
                                                    "frac": (b / (t / 1000.0) / 1e9 / peak) if t > 0 else 0.0,
                                                    "traffic": measured_traffic("shade_kernel")[0]})(
                (my_rows_px if world > 1 else W * H) * 20 + S * S * 4, avg["shade"]),
            "pass_ms": avg, "pass_ms_last_pipelined_frame": {k: float(v) for k, v in tm.items()},
            "host_issue_ms_per_step": host_issue_ms, "host_vkr_draw_call_ms": float(np.median(host_call)), "host_busy_ms_per_step": host_busy,
            "cluster_culling": cull, "per_rank_pass_ms_last_pipelined_frame": per_rank,
            "e2e": {"value": args.steps / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "d2h_bytes_per_step_rank0": int((my_rows[1] - my_rows[0]) * W * 3),
                    "readback": ("VKR_TARGET_RGB8: the presented frame as packed R,G,B bytes" if world == 1 else
                                 f"every rank copies its own rows (packed R,G,B) into one host frame shared by the {world} processes"),
                    "checksum": checksum, "frames_verified": 2,
                    "frame_crc32": frame_crc,
                    "d2h_probe_GBps": d2h_gbps, "d2h_bound_frames_per_s": (d2h_gbps * 1e9 / d2h) if d2h_gbps else None},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        if world > 1:
            line["sharded_equals_single"] = sharded_equals_single
            line["exchange"] = "peer memory (CUDA IPC over NVLink)" if r.shard_info()["peer_memory"] else "NCCL send/recv"
        if world == 1 and not args.no_cpu_baseline:
            times, cores, _ = time_oracle(1, 0, budget_s=60.0)
            cfps = len(times) / sum(times)
            line["cpu_baseline"] = {"value": cfps, "unit": "frames/s", "cores": cores, "kind": "port",
                                    "sample": f"{len(times)} full 4K frame of the same scene, oracle/ork.cpp, {cores} threads"}
    r.sync()
    barrier()                            # every rank has finished with its peers' memory and with the shared host frame
    if shm_map is not None:
        torch.cuda.cudart().cudaHostUnregister(shm_map.ctypes.data)
        del host_frames, shm_map
        barrier()
        if rank == 0:
            os.unlink(shm_path)
    r.close()
    # ---- BASELINE config 5 (frame-parallel multi-view batch) on the same ranks, reported inside the same line
    if not args.no_views:
        barrier()
        v = measure_views(lib, dist, rank, world, local, steps=2, warmup_req=3)
        if rank == 0:
            line["views"] = {"metric": v["metric"], "value": v["value"], "unit": v["unit"], "e2e": v["e2e"], "n_gpus": world,
                             "steps": v["steps"], "ms_per_step": v["ms_per_step"], "scaling": "strong", "config": v["config"],
                             "gpu_launches": v["gpu_launches"], "expected": "linear in n_gpus: no exchange step"}
    # ---- SURVEY 8f-2 (update_scene on the device), one GPU only: the object-count case through the host engine, both ways
    if rank == 0 and world == 1 and not args.no_scene_update:
        try:
            line["scene_update"] = measure_scene_update()
        except Exception as e:       # a side measurement: never costs the headline line
            line["scene_update"] = {"error": repr(e)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure_scene_update(frames=200):
    """16,000 draws (procedural config 6: spheres of 192 triangles under 40 group nodes, 1080p, PBR + PCF shadow + tonemap)
    through the C++ host engine: update_scene + draw per frame with the per-object work on the host (the reference's
    update_scene: refreshTransform, is_visible, sort — then the draw list is uploaded) and on the device (vkr_draw_scene).
    Frames issued back to back, one sync at the end, wall clock around the loop."""
    from vk_renderer_b200.engine import Engine
    out = {"workload": "procedural config 6: 16,000 draws (3.07 M triangles) under 40 group nodes, 1920x1080, PBR, 2048^2 PCF shadow, tonemap",
           "frames": frames, "unit": "frames/s", "timing": "wall clock around the loop, one sync at its end"}
    flags = capi.CONFIG_SUBMIT_THREAD | capi.CONFIG_TRANSIENT_DEPTH | capi.CONFIG_TRANSIENT_HDR
    for mode in ("host", "device"):
        e = Engine(1920, 1080, flags=flags)
        try:
            e.load_procedural(6)
            if mode == "device":
                e.set_gpu_scene_update(True)
            for _ in range(10):
                e.update_scene(0.0); e.draw()
            e.sync()
            t0 = time.perf_counter()
            for _ in range(frames):
                e.update_scene(0.0); e.draw()
            e.sync()
            dt = time.perf_counter() - t0
            st = e.stats()
            t1 = time.perf_counter()
            for _ in range(50):
                e.update_scene(0.0)
            out[f"update_scene_on_{mode}"] = {"value": frames / dt, "visible_draws": int(st.drawcall_count), "submitted_triangles": int(st.triangle_count),
                                              "host_update_scene_ms": (time.perf_counter() - t1) / 50 * 1e3}
        finally:
            e.close()
    return out


# ------------------------------------------------------------------------------- frame-parallel views (config 5)
VW, VH, NVIEWS = 1920, 1080, 256


def build_view_scene(r):
    """BASELINE config 5 = config 2's scene: 16x16 spheres (rings = segments = 29 -> 1682 triangles) on a
    200x200-quad ground plane, ~511 k triangles, PCF shadow, tonemap; 256 cameras on an orbit."""
    rng = np.random.default_rng(0xB2000000 + 5)
    sv, si = mg.uv_sphere(29, 29, 1.0)
    pv, pi = mg.plane(200, 200, 60.0, -1.0, uv_scale=1)
    ms, mp_ = r.upload_mesh(si, sv), r.upload_mesh(pi, pv)
    objs = [mg.render_object(mp_, r.build_material((0.8, 0.8, 0.8, 1.0), 0.0, 0.9), np.eye(4), len(pi))]
    for ix in range(16):
        for iz in range(16):
            mat = r.build_material(tuple(rng.uniform(0.2, 1.0, 3)) + (1.0,), float(rng.random() < 0.3), float(rng.uniform(0.2, 0.9)))
            objs.append(mg.render_object(ms, mat, mg.translate((ix - 7.5) * 3.0, 0.0, (iz - 7.5) * 3.0), len(si)))
    L = mg.light_matrix()
    proj = mg.camera_projection(VW / VH)
    views = []
    for k in range(NVIEWS):
        yaw = 2.0 * np.pi * k / NVIEWS
        view = mg.camera_view((34.0 * np.sin(yaw), 9.0 + 3.0 * np.sin(3 * yaw), 34.0 * np.cos(yaw)), pitch=-0.28, yaw=yaw)      # forward = (-sin yaw, ., -cos yaw): towards the orbit's centre
        views.append(capi.make_scene_data(view, proj, [L] * 4, shadow_mode=1, enable_shadows=True))
    fp = capi.default_frame_params(enableShadows=1, shadowMode=1, enablePostprocess=1, enableSkybox=0, enableBackground=1)
    tris = (len(pi) + 256 * len(si)) // 3
    return views, np.array(objs), fp, tris


def run_views(args):
    """--workload views: one step = the batch of 256 views, view i on rank i mod N, no exchange step."""
    import torch
    import vk_renderer_b200

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = vk_renderer_b200.load()
    torch.cuda.set_stream(torch.cuda.Stream())
    line = measure_views(lib, dist, rank, world, local, args.steps, args.warmup)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure_views(lib, dist, rank, world, local, steps, warmup_req):
    """Config 5 on the ranks of an initialised process group (or one GPU): returns the JSON line (rank 0; None elsewhere)."""
    import torch
    from vk_renderer_b200.sharding import ViewSharder, view_slice

    class _A:
        pass
    args = _A()
    args.steps, args.warmup = steps, warmup_req
    r = capi.Renderer(lib, "vkr", VW, VH, shadow_extent=S, device=local, stream=torch.cuda.current_stream().cuda_stream,
                      flags=capi.CONFIG_SUBMIT_THREAD)
    views, objs, fp, tris = build_view_scene(r)
    mine = view_slice(rank, world, NVIEWS)
    sharder = ViewSharder(r, rank, world, cuda=True, rgb8=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        for v in mine:
            r.draw(views[v], objs, fp)

    def maxed(x):
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(3, args.warmup)):
        step()
    r.sync()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = 0
    for _ in range(args.steps):
        step()
        launches += r.launch_count() * len(mine)
    e1.record()
    barrier()
    r.sync()
    ms = maxed(e0.elapsed_time(e1))
    clocks = sampler.stop() if sampler else None
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sharder.render(views, objs, fp)            # draw + read every view back to pinned host memory
    barrier()
    e2e_s = maxed(time.perf_counter() - t0)
    line = None
    if rank == 0:
        vps = args.steps * NVIEWS / (ms / 1000.0)
        line = ({
            "metric": "views_per_sec_1080p_multiview", "value": vps, "unit": "views/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE config 5: batch of 256 camera views at 1920x1080 of the config-2 scene (16x16 spheres + "
                                   "ground plane, PBR, 2048^2 PCF shadow, tonemap), frame-parallel",
                       "triangles": int(tris), "views": NVIEWS, "resolution": [VW, VH], "shadow_extent": S,
                       "parallelism": f"frame-parallel: view i on rank i mod {world}, scene replicated, no collective"},
            "mtri_per_s": vps * tris / 1e6, "gpix_per_s": vps * VW * VH / 1e9,
            "e2e": {"value": args.steps * NVIEWS / e2e_s, "unit": "views/s", "h2d_bytes_per_step": int(r.upload_bytes() * len(mine)),
                    "d2h_bytes_per_step": int(VW * VH * sharder.channels * len(mine)),
                    "readback": "VKR_TARGET_RGB8: every view as packed R,G,B bytes"},
            "gpu_launches": int(launches), "clocks": clocks})
    r.close()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-views", action="store_true", help="skip the config-5 pass that follows the frame workload")
    ap.add_argument("--no-scene-update", action="store_true", help="skip the update_scene (host vs device) side measurement")
    ap.add_argument("--workload", default="frame", choices=["frame", "views"],
                    help="frame: BASELINE config 4 (the headline metric, default); views: config 5, frame-parallel multi-view batch")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "views":
        run_views(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
