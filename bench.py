#!/usr/bin/env python
"""bench.py — Sponza 1080p 8-bounce Mrays/s (device-timed) at 1/2/4/8 B200 vs CPU ref.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Headline workload (BASELINE.json configs[3]): the Sponza-class stand-in (sponza.obj is not shipped:
/root/reference/.MISSING_LARGE_BLOBS) — ~262k triangles, 354 mesh entities, 27 texture layers, procedural sky cube —
1920x1080, maxBounces 8, oct-tree (2,50), BVH (30,5), camera (60,0,0) looking -x.  A step is FRAMES_PER_STEP progressive
frames (1 spp each) of that image.  With N GPUs the image is tile-sharded in interleaved 8-row bands (no data-path
collective); value = ray segments traced by all ranks / max-over-ranks device time.  A ray segment is one findClosestHit
call (RaytracerCS.hlsl:446).

The line also carries, all measured in this run:
  parity      frame 1 of the headline workload rendered by the GPU path and by the oracle (the same frame the cpu_baseline
              leg times), RGBA32F accumulation compared word for word; the bench FAILS (rc 1) if they differ
  e2e         the same K steps through the public calls with host buffers: constants in, result image out, every step
              (okrt_read_result_async: step k's copy overlaps step k+1's kernels)
  roofline    the traversal kernel against unit rates measured on this device (okrt_measure_unit_peaks), from its own
              executed step counts (libokrt_stats.so: the same kernels + counters, one step)
  configs     BASELINE.json configs[0..2] (C1 TestScene, C2 dragon-class, C3 glass) at their full sizes and configs[4]
              (C5: 3840x2160 sample-sharded + one all-reduce of the accumulation image when N>1)
"""
import argparse
import ctypes
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WIDTH, HEIGHT, MAX_BOUNCES = 1920, 1080, 8
FRAMES_PER_STEP = 16         # one wavefront pass at 1080p on one GPU: 2^25 path slots = 16 frames of 2 073 600 pixels
METRIC = "Sponza 1080p 8-bounce Mrays/s (device-timed) at 1/2/4/8 B200 vs CPU ref"
UNIT = "Mrays/s"
C5_FRAMES_PER_RANK = 128     # configs[4]: 1024 spp over 8 ranks


def workload_config(n_gpus):
    return {
        "workload": "sponza-standin 1920x1080 maxBounces=8 (BASELINE configs[3]); %d progressive frames (1 spp) per step" % FRAMES_PER_STEP,
        "scene": "procedural Sponza-class stand-in, seed 11: sponza.obj is listed in .MISSING_LARGE_BLOBS",
        "resolution": [WIDTH, HEIGHT], "max_bounces": MAX_BOUNCES, "frames_per_step": FRAMES_PER_STEP,
        "bvh": [30, 5], "octree": [2, 50],
        "sharding": "tile: interleaved 8-row bands, rank = band %% %d, scene replicated, no collective" % n_gpus,
        "batches_in_flight": "3 (library default, the same for every N): consecutive 16-frame batches overlap on the device, frames are accumulated in frame order",
        "l2": "no flush: the working set (geometry + 27x903x903 RGBA8 textures + 1080p RGBA32F/RGBA8 images, ~230 MB) is larger than the 126 MB L2",
    }


def build_scene(okrt):
    rm, sc, info = okrt.scenes.sponza_standin()
    host = okrt.harness.build_host_side(okrt, rm, sc, info, WIDTH, HEIGHT)
    env = okrt.scenes.procedural_sky_faces(512)
    return host, env, info


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(0.02)

    def finish(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            j = json.load(f)
            return {"hbm_gbs": float(j["hbm_gbs"]), "source": "MEASURED_PEAKS.json"}
    except Exception:
        return {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def oracle_frame(okrt, host, env, threads=0):
    """Frame 1 of the headline workload by the oracle (kind "port": scalar restatement of the HLSL), all host threads:
    the CPU baseline sample AND the image the GPU path is compared with."""
    from oracle import binding as oracle_binding
    oracle = oracle_binding.load()
    scene = oracle_binding.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()
    rd[0]["numAccumulationFrames"] = 1
    t0 = time.perf_counter()
    out = oracle.render(scene, rd, MAX_BOUNCES, threads=threads)
    dt = time.perf_counter() - t0
    cores = threads if threads > 0 else (os.cpu_count() or 1)
    cpu = {"value": out["segments"] / dt / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": "one full 1920x1080 frame, 1 spp, %d ray segments in %.1f s" % (out["segments"], dt)}
    # where the reference's own shader text compiled as C++ travelled with the snapshot (oracle/_ref/libokrt_ref_shader.so), IT is
    # the CPU baseline (kind "reference"); the port's rate stays beside it.  Same frame, same paths: the port's segment count.
    ref_shader = load_reference_shader()
    if ref_shader is not None:
        t0 = time.perf_counter()
        ref_shader.render(scene, rd, MAX_BOUNCES, threads=threads)
        dtr = time.perf_counter() - t0
        cpu = {"value": out["segments"] / dtr / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
               "sample": "one full 1920x1080 frame, 1 spp, %d ray segments in %.1f s by the reference's RaytracerCS.hlsl compiled as C++" % (out["segments"], dtr),
               "port_value": out["segments"] / dt / 1e6}
    return cpu, out


def load_reference_shader():
    try:
        from oracle import binding as oracle_binding
        return oracle_binding.load_ref_shader()
    except Exception:
        return None


def run_reference_arm(args):
    """--impl reference: the reference's own implementation of the path on the host cores — its shader text compiled as C++
    (oracle/_ref/libokrt_ref_shader.so, kind "reference"; running the HLSL itself needs D3D11: unavailable on this box), or the
    oracle port where that library did not travel.  All host threads, each step a bounded sample of the workload.  The compiled
    shader has no segment counter: the port traces the same rows of the same frame outside the timed region (identical paths)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    okrt = importlib.import_module("gpu-raytracer_b200")
    from oracle import binding as oracle_binding
    oracle = oracle_binding.load()
    host, env, info = build_scene(okrt)
    scene = oracle_binding.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()
    rows = np.array([y for y in range(HEIGHT) if (y // 8) % 8 == 0], np.uint32)     # every 8th 8-row band: 1/8 of the frame
    cores = os.cpu_count() or 1

    ref_shader = load_reference_shader()
    kind = "reference" if ref_shader is not None else "port"

    def count(frame):
        rd[0]["numAccumulationFrames"] = frame
        return oracle.render(scene, rd, MAX_BOUNCES, row_list=rows)["segments"]

    def step(frame):
        rd[0]["numAccumulationFrames"] = frame
        if ref_shader is not None:
            ref_shader.render(scene, rd, MAX_BOUNCES, row_list=rows)
            return 0
        return oracle.render(scene, rd, MAX_BOUNCES, row_list=rows)["segments"]
    for w in range(args.warmup):
        step(w + 1)
    t0 = time.perf_counter()
    segments = sum(step(args.warmup + k + 1) for k in range(args.steps))
    dt = time.perf_counter() - t0
    if ref_shader is not None:
        segments = sum(count(args.warmup + k + 1) for k in range(args.steps))      # untimed: the same rows of the same frames
    value = segments / dt / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / max(1, args.steps) * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": "per step: every 8th 8-row band of the 1080p frame (1/8 of the pixels), 1 spp, %d host threads; %s" % (
                                 cores, "the reference's RaytracerCS.hlsl compiled as C++" if kind == "reference" else "oracle port")},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------- legs of the GPU arm
def run_config(okrt, name, rm, sc, info, width, height, bounces, frames, env, device, steps=3, note=None, with_cpu=True):
    """One BASELINE config on one GPU: `steps` timed calls of `frames` progressive frames after one warm-up call; with_cpu: frame 1
    by the oracle on all host threads as well (its rate, and the GPU's frame 1 compared with it word for word)."""
    abi = okrt.abi
    t0 = time.perf_counter()
    host = okrt.harness.build_host_side(okrt, rm, sc, info, width, height, spheres=(name == "C1"))
    build_s = time.perf_counter() - t0
    r = okrt.harness.Renderer(okrt, host, width, height, envFaces=env, device=device)
    r.options(maxBounces=bounces, traversal=abi.TRAVERSAL_FAST, coverAllPixels=0)
    rd = host["rd"].copy()
    t0 = time.perf_counter()
    r.render(rd, frames=frames)                      # warm-up: includes the traversal-layout build (on the device)
    first_ms = (time.perf_counter() - t0) * 1e3
    lb = r.wait_layout()                             # ... and the host SAH tree that replaces the clustered one in the background
    abi.check(r.L.okrt_reset_stats(r.h))
    ms = 0.0
    for _ in range(steps):
        r.render(rd, frames=frames)
        ms += float(r.stats()["lastRenderMs"])
    st = r.stats()
    seg, fb = int(st["raySegments"]), int(st["fallbackSegments"])
    cpu = parity = None
    if with_cpu:
        from oracle import binding as oracle_binding
        oracle = oracle_binding.load()
        scene = oracle_binding.oracle_scene(oracle, host, envFaces=env)
        rd1 = host["rd"].copy()
        rd1[0]["numAccumulationFrames"] = 1
        t0 = time.perf_counter()
        o = oracle.render(scene, rd1, bounces)
        dt = time.perf_counter() - t0
        cpu = {"value": o["segments"] / dt / 1e6, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
               "sample": "frame 1, %d ray segments in %.1f s" % (o["segments"], dt)}
        abi.check(r.L.okrt_reset_accumulation(r.h))
        abi.check(r.L.okrt_reset_stats(r.h))
        r.render(host["rd"].copy(), frames=1)
        acc = r.accumulation()
        bad = np.any(acc.view(np.uint32) != o["accum"].view(np.uint32), axis=2)
        gseg = int(r.stats()["raySegments"])
        parity = {"bit_exact": bool(not bad.any()) and gseg == int(o["segments"]), "differing_pixels": int(bad.sum()),
                  "segments": [gseg, int(o["segments"])]}
        ref_shader = load_reference_shader()
        if ref_shader is not None:                       # the reference's own shader text compiled as C++, same frame
            want = ref_shader.render(scene, rd1, bounces)
            bad_ref = np.any(acc.view(np.uint32) != want["accum"].view(np.uint32), axis=2)
            parity["reference_shader_bit_exact"] = bool(not bad_ref.any())
            parity["bit_exact"] = parity["bit_exact"] and parity["reference_shader_bit_exact"]
    r.close()
    out = {"workload": "%s %dx%d maxBounces=%d, %d frames per call" % (info.get("name", name), width, height, bounces, frames),
           "value": seg / (ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_frame": ms / (steps * frames), "ray_segments": seg,
           "fallback_segments": fb, "triangles": int(len(host["geometry"].trianglesPos)), "stand_in": bool(info.get("stand_in", False)),
           "host_build_s": round(build_s, 2), "first_call_ms": round(first_ms, 1),
           "layout": {"builder": int(lb["builder"]), "device_ms": round(float(lb["deviceMs"]), 2), "refine_ms": round(float(lb["refineMs"]), 1)}}
    if cpu:
        out["cpu_baseline"], out["parity"] = cpu, parity
    if note:
        out["note"] = note
    return out


def run_single_gpu_configs(okrt, device, with_cpu=True):
    sc_mod = okrt.scenes
    configs = {}
    rm, sc, info = sc_mod.test_scene()
    configs["C1"] = run_config(okrt, "C1", rm, sc, info, 512, 512, 0, 16, None, device, with_cpu=with_cpu,
                               note="configs[0]: primary rays only; 512x504 pixels are covered (Dispatch(W/16, H/9), Q9)")
    rm, sc, info = sc_mod.dragon_scene()
    configs["C2"] = run_config(okrt, "C2", rm, sc, info, 1920, 1080, 4, 16, sc_mod.procedural_sky_faces(256), device, with_cpu=with_cpu,
                               note="configs[1]: 871k-triangle displaced knot (stand-in: dragon.obj is not shipped), metallic, sky cube")
    path = os.path.join(ROOT, "tests", "golden", "ikea_glass_mesh.npz")
    rm, sc, info = sc_mod.glass_scene_from_fixture(path) if os.path.exists(path) else sc_mod.glass_scene()
    configs["C3"] = run_config(okrt, "C3", rm, sc, info, 1920, 1080, 16, 16, sc_mod.procedural_sky_faces(256), device, with_cpu=with_cpu,
                               note="configs[2]: the reference's ikea_glass.obj (committed fixture), scale 10, transparency 1, IOR 1.5")
    return configs


def run_c5(okrt, scene_parts, env, device, rank, world, dist):
    """configs[4]: 3840x2160 progressive, sample-sharded: rank r renders global frames r+1, r+1+N, ...; ONE all-reduce of the
    RGBA32F accumulation image through okrt_reduce_accumulation, then the resolve.  Also renders a short sequential run
    for the RMSE of the reduced mean against the frame-ordered sum."""
    import torch
    abi = okrt.abi
    W, H = 3840, 2160
    rm, sc, info = scene_parts
    host = okrt.harness.build_host_side(okrt, rm, sc, info, W, H)
    r = okrt.harness.Renderer(okrt, host, W, H, envFaces=env, device=device)
    opts = dict(maxBounces=MAX_BOUNCES, traversal=abi.TRAVERSAL_FAST, coverAllPixels=0)
    if world > 1:
        opts.update(okrt.sharding.sample_options(rank, world))
    r.options(**opts)
    rd = host["rd"].copy()
    tensor = okrt.sharding.accumulation_tensor(r.L.okrt_accumulation_device_ptr(r.h), W, H, torch.device("cuda", device)) if world > 1 else None
    stream = torch.cuda.ExternalStream(int(r.L.okrt_cuda_stream(r.h)), device=torch.device("cuda", device)) if world > 1 else None
    reduce_events = []

    def allreduce(ptr, count, cuda_stream, user):
        try:
            with torch.cuda.stream(stream):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
                e1.record()
                reduce_events.append((e0, e1))
            return 0
        except Exception as exc:            # noqa: BLE001 — reported through the C ABI's status
            print("all-reduce failed: %r" % (exc,), file=sys.stderr)
            return 1
    cb = abi.ALLREDUCE_FN(allreduce)

    def one_run(frames):
        abi.check(r.L.okrt_reset_accumulation(r.h))
        abi.check(r.L.okrt_reset_stats(r.h))
        abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        abi.check(r.L.okrt_render_frames(r.h, frames))
        if world > 1:
            abi.check(r.L.okrt_reduce_accumulation(r.h, cb, None, frames * world))
        abi.check(r.L.okrt_sync(r.h))
        wall = time.perf_counter() - t0
        st = r.stats()
        return wall, float(st["lastRenderMs"]), int(st["raySegments"])
    one_run(4)                                               # warm-up: layout build, buffers, NCCL channels for this message size
    r.wait_layout()
    reduce_events.clear()
    wall, render_ms, seg = one_run(C5_FRAMES_PER_RANK)
    ar_ms = float(reduce_events[-1][0].elapsed_time(reduce_events[-1][1])) if reduce_events else 0.0
    total_frames = C5_FRAMES_PER_RANK * world
    out = None
    ar_min_ms = ar_ms
    if world > 1:
        t = torch.tensor([wall, render_ms, ar_ms], device="cuda", dtype=torch.float64)
        c = torch.tensor([seg], device="cuda", dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        wall, render_ms, ar_ms, seg = float(t[0]), float(t[1]), float(t[2]), int(c[0])
        tmin = torch.tensor([reduce_events and float(reduce_events[-1][0].elapsed_time(reduce_events[-1][1])) or 0.0], device="cuda", dtype=torch.float64)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        ar_min_ms = float(tmin[0])
    if rank == 0:
        out = {"workload": "sponza-standin 3840x2160 maxBounces=8, %d spp = %d frames on each of %d ranks, sample-sharded" % (total_frames, C5_FRAMES_PER_RANK, world),
               "value": seg / wall / 1e6, "unit": UNIT, "wall_ms": wall * 1e3, "render_ms": render_ms, "all_reduce_ms": ar_ms,
               "all_reduce_ms_last_rank_in": ar_min_ms,
               "all_reduce_bytes": W * H * 16, "all_reduce_gbs": (W * H * 16 / 1e9) / (ar_min_ms * 1e-3) if ar_min_ms > 0 else None,
               "ray_segments": seg, "total_frames": total_frames,
               "note": "configs[4] asks for 1024 spp on 8 ranks; every rank renders 128 spp.  all_reduce_ms: CUDA events around the "
                       "collective on the renderer's stream, after one warm-up reduction of the same size, max over ranks: it contains the wait "
                       "for the slowest rank's frames; all_reduce_ms_last_rank_in (min over ranks) is the collective itself, all_reduce_gbs is built on it"}
    # RMSE of the reduced mean against the sequential frame-ordered sum, on a short run (8 frames per rank)
    if world > 1:
        short = 8
        one_run(short)
        reduced = r.accumulation() if rank == 0 else None
        if rank == 0:
            seq = okrt.harness.Renderer(okrt, host, W, H, envFaces=env, device=device)
            seq.options(maxBounces=MAX_BOUNCES, traversal=abi.TRAVERSAL_FAST, coverAllPixels=0)
            seq.render(rd, frames=short * world)
            want = seq.accumulation().astype(np.float64)
            seq.close()
            n = short * world
            out["rmse_vs_sequential"] = float(np.sqrt(np.mean((reduced.astype(np.float64) / n - want / n) ** 2)))
            out["sample_counts_equal"] = bool(np.array_equal(reduced[..., 3].astype(np.float64), want[..., 3]))
        dist.barrier()
    r.close()
    return out


def trace_roofline(okrt, host, env, device, trace_ms, trace_launches, dev_ms, segments, peaks):
    """The traversal kernel against unit rates measured on this device.  One step of the same workload through
    libokrt_stats.so (same kernels + counters) gives the kernel's own executed work per ray; the timed arms gave its time."""
    abi = okrt.abi
    stats_path = os.path.join(ROOT, "gpu-raytracer_b200", "libokrt_stats.so")
    if not os.path.exists(stats_path):
        raise SystemExit("bench.py: gpu-raytracer_b200/libokrt_stats.so is missing (python -c 'import __graft_entry__ as g; g.build()')")
    slib = abi.load_library(stats_path)
    r = okrt.harness.Renderer(okrt, host, WIDTH, HEIGHT, envFaces=env, device=device, lib=slib)
    r.options(maxBounces=MAX_BOUNCES, traversal=abi.TRAVERSAL_FAST, coverAllPixels=0, batchesInFlight=1)
    rd = host["rd"].copy()
    r.render(rd, frames=FRAMES_PER_STEP)
    r.wait_layout()
    abi.check(slib.okrt_reset_stats(r.h))
    r.render(rd, frames=FRAMES_PER_STEP)
    st = r.trace_stats(MAX_BOUNCES + 1).astype(np.float64)
    r.close()
    tot = st.sum(axis=0)
    rays = max(1.0, tot[0])
    node_steps, tri_steps = tot[1] / rays, tot[3] / rays
    lanes_node, lanes_tri = tot[6] / max(1.0, tot[4]), tot[7] / max(1.0, tot[5])
    # LSU sector requests a ray needs at least, lane by lane: two 32-byte sectors per child-pair node (two 256-bit loads),
    # three for its path state (origin + direction read, hit record written).  Triangle records (48 B: two sectors each) go
    # through the texture pipe and are listed beside it.
    lsu_per_ray = node_steps * 2.0 + 3.0
    tex_per_ray = tri_steps * 2.0
    seconds = max(trace_ms, 1e-9) * 1e-3
    achieved = lsu_per_ray * segments * 32.0 / seconds / 1e9
    peak_l1 = peaks["sectorsPerSecondL1"] * 32.0 / 1e9
    peak_l2 = peaks["sectorsPerSecond"] * 32.0 / 1e9
    wasted = traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")) as f:
            tj = json.load(f)
        traffic = tj["traceFastKernel"] * (trace_ms / max(1, trace_launches)) / tj["traceFastKernel_ms"]
        wasted = tj.get("traceFastKernel_local_sectors_per_global_sector")
    except Exception:
        pass
    return {"bound": "l1-sector-rate", "kernel": "traceFastKernel",
            "achieved": achieved, "peak": peak_l1, "unit": "GB/s of 32-byte LSU sectors (L1 data stage)", "frac": achieved / peak_l1,
            "frac_of_l2_sector_rate": achieved / peak_l2, "peak_l2_resident": peak_l2,
            "peak_source": "okrt_measure_unit_peaks on this device: divergent 256-bit loads, 16 KB window per block (L1 hits) / 16 MB window (L2 hits)",
            "traffic": traffic, "traffic_source": "DRAM bytes of the committed ncu capture, scaled by launch duration (static)",
            "kernel_launches": trace_launches, "kernel_ms_per_launch": trace_ms / max(1, trace_launches), "kernel_share_of_step": trace_ms / dev_ms,
            "per_ray": {"node_steps": node_steps, "triangle_tests": tri_steps, "lsu_sectors": lsu_per_ray, "tex_sectors": tex_per_ray},
            "simt_efficiency": {"node_phase": lanes_node / 32.0, "triangle_phase": lanes_tri / 32.0,
                                "source": "lanes busy per phase iteration, counted by libokrt_stats.so on one step of this workload"},
            "wasted_local_sectors_per_global_sector": wasted,
            "issue": {"peak_warp_instructions_per_s": peaks["warpInstructionsPerSecond"], "sm_clock_mhz": peaks["smClockMHz"]},
            "hbm_peak_gbs": measured_peaks(),
            "note": "achieved = (2 sectors x node steps + 3) per ray x rays / event-timed kernel time: the kernel's own minimum work, not the "
                    "reference order's; the scene is L2-resident, DRAM sees only path state (traffic)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--traversal", default="fast", choices=["fast", "exact", "wide"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU legs (cpu_baseline, parity) — tuning runs only")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1/C2/C3/C5 legs — tuning runs only")
    ap.add_argument("--balance-bands", action="store_true", help="N>1: cost-weighted band table (okrt_set_shard_bands) instead of the interleave")
    ap.add_argument("--cpu-threads", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    okrt = importlib.import_module("gpu-raytracer_b200")
    abi = okrt.abi
    helpers = okrt.harness
    if abi.lib().okrt_device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device visible; the product path has no CPU fallback")

    scene_parts = okrt.scenes.sponza_standin()
    host = helpers.build_host_side(okrt, *scene_parts, WIDTH, HEIGHT)
    env = okrt.scenes.procedural_sky_faces(512)
    r = helpers.Renderer(okrt, host, WIDTH, HEIGHT, envFaces=env, device=local_rank)
    traversal = {"fast": abi.TRAVERSAL_FAST, "exact": abi.TRAVERSAL_EXACT, "wide": abi.TRAVERSAL_WIDE}[args.traversal]
    opts = dict(maxBounces=MAX_BOUNCES, traversal=traversal, coverAllPixels=0)
    band_balance = None
    if world > 1:
        opts.update(shardMode=abi.SHARD_TILE, shardRank=rank, shardCount=world, shardBandRows=8)
    r.options(**opts)
    if world > 1 and args.balance_bands:
        # opt-in: cost-weighted band table instead of the interleave.  Every rank renders ONE unsharded calibration frame (untimed
        # set-up), sums the ray segments of each 8-row band and runs the same deterministic assignment (okrt_shard_balance):
        # no communication, equal tables.  Measured on 8 GPUs: 20 640 vs 21 067 Mrays/s for the plain interleave (segments are
        # a poor proxy for time), hence not the default.
        r.options(writeAov=1, shardMode=abi.SHARD_NONE)
        r.render(host["rd"].copy(), frames=1)
        cost = okrt.sharding.band_costs_from_segments(r.aov()["numSegments"], 8)
        owner = okrt.sharding.balanced_bands(cost, world)
        mine = np.ascontiguousarray(np.nonzero(owner == rank)[0], np.uint32)
        abi.check(r.L.okrt_reset_accumulation(r.h))
        r.options(writeAov=0, shardMode=abi.SHARD_TILE)
        abi.check(r.L.okrt_set_shard_bands(r.h, abi.ptr(mine), len(mine)))
        loads = [float(cost[owner == k].sum()) for k in range(world)]
        band_balance = {"bands_per_rank": [int((owner == k).sum()) for k in range(world)], "predicted_spread": max(loads) / min(loads) - 1.0}
    rd = host["rd"].copy()

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()
        abi.check(r.L.okrt_sync(r.h))

    peaks = None
    if rank == 0:
        pk = np.zeros(1, abi.UNIT_PEAKS)
        abi.check(abi.lib().okrt_measure_unit_peaks(local_rank, abi.ptr(pk)))
        peaks = {k: float(pk[0][k]) for k in ("sectorsPerSecond", "sectorsPerSecondL1", "warpInstructionsPerSecond", "smClockMHz")}

    # ---- device-timed arm: inputs resident, K steps back to back on the renderer's stream
    abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))
    t_first = time.perf_counter()
    abi.check(r.L.okrt_render_frames(r.h, FRAMES_PER_STEP))                   # first call: builds the traversal layout
    abi.check(r.L.okrt_sync(r.h))
    first_call_ms = (time.perf_counter() - t_first) * 1e3
    lb = np.zeros(1, abi.LAYOUT_BUILD)
    abi.check(r.L.okrt_get_layout_build(r.h, abi.ptr(lb)))
    layout_first = {"builder": int(lb[0]["builder"]), "device_ms": float(lb[0]["deviceMs"]), "wall_ms": float(lb[0]["wallMs"]),
                    "pair_nodes": int(lb[0]["numPairNodes"]), "why": bytes(lb[0]["why"]).split(b"\0")[0].decode()}
    t_wait = time.perf_counter()
    abi.check(r.L.okrt_wait_layout(r.h))                                       # the background SAH build replaces the clustered tree
    abi.check(r.L.okrt_get_layout_build(r.h, abi.ptr(lb)))
    layout_now = {"builder": int(lb[0]["builder"]), "refine_ms": float(lb[0]["refineMs"]), "waited_ms": (time.perf_counter() - t_wait) * 1e3,
                  "pair_nodes": int(lb[0]["numPairNodes"]), "stack": int(lb[0]["fastStackDepth"])}
    if args.warmup > 1:
        abi.check(r.L.okrt_render_frames(r.h, (args.warmup - 1) * FRAMES_PER_STEP))
    barrier()
    abi.check(r.L.okrt_reset_stats(r.h))
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    abi.check(r.L.okrt_render_frames(r.h, args.steps * FRAMES_PER_STEP))      # CUDA events bracket exactly these K steps
    barrier()
    clocks = sampler.finish()
    st = r.stats()
    dev_ms = float(st["lastRenderMs"])
    trace_ms, trace_launches = float(st["traceKernelMs"]), int(st["traceKernelLaunches"])
    segments = int(st["raySegments"])
    launches = int(st["kernelLaunches"])
    fallback = int(st["fallbackSegments"])

    # ---- the traversal kernel timed alone: three steps with ONE batch in flight, so that the CUDA-event pair around each
    # traversal launch spans that kernel only (with several batches in flight other batches' kernels run inside the pair)
    alone = None
    if world == 1:
        r.options(batchesInFlight=1)
        barrier()
        abi.check(r.L.okrt_reset_stats(r.h))
        abi.check(r.L.okrt_render_frames(r.h, 3 * FRAMES_PER_STEP))
        barrier()
        sa = r.stats()
        alone = {"dev_ms": float(sa["lastRenderMs"]), "trace_ms": float(sa["traceKernelMs"]), "trace_launches": int(sa["traceKernelLaunches"]),
                 "segments": int(sa["raySegments"]), "steps": 3}
        r.options(batchesInFlight=0)

    # ---- end-to-end arm: host constants in, result image out, every step, wall clock around the public calls.  Step k's
    # device-to-host copy runs on the library's copy stream while the next steps' kernels run; the host waits for copy k-2
    # after it has submitted step k (three page-locked host images in turn), so three steps are on the device at a time.
    pinned = [np.zeros((HEIGHT, WIDTH, 4), np.uint8) for _ in range(3)]
    for buf in pinned:
        abi.check(r.L.okrt_host_register(abi.ptr(buf), buf.nbytes))           # page-locked: the D2H copy is a plain DMA
    moved = ctypes.c_uint64(0)

    def e2e_steps(count):
        for k in range(count):
            rd[0]["numAccumulationFrames"] = r.L.okrt_get_num_accumulation_frames(r.h)
            abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))              # H2D: 240 B of constants travel with the launches
            abi.check(r.L.okrt_render_frames(r.h, FRAMES_PER_STEP))
            # D2H: the RGBA8 result image; a tile-sharded rank moves only the row bands it owns into the shared image layout
            abi.check(r.L.okrt_read_result_async(r.h, abi.ptr(pinned[k % 3]), ctypes.byref(moved), k % 3))
            if k >= 2:
                abi.check(r.L.okrt_wait_read(r.h, (k - 2) % 3))
        for k in range(max(0, count - 2), count):
            abi.check(r.L.okrt_wait_read(r.h, k % 3))
    e2e_steps(args.warmup)                                                     # the same call sequence, untimed
    abi.check(r.L.okrt_reset_stats(r.h))
    barrier()
    t0 = time.perf_counter()
    e2e_steps(args.steps)
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_segments = int(r.stats()["raySegments"])
    for buf in pinned:
        abi.check(r.L.okrt_host_unregister(abi.ptr(buf)))
    d2h_bytes = int(moved.value)                   # per step, this rank; summed over ranks below = one whole image

    rank_ms = rank_segments = None
    if dist is not None:
        import torch
        t = torch.tensor([dev_ms, e2e_s], device="cuda", dtype=torch.float64)
        per_rank = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(per_rank, t)
        rank_ms = [round(float(x[0]), 3) for x in per_rank]
        rank_segments = [torch.zeros(1, device="cuda", dtype=torch.int64) for _ in range(world)]
        dist.all_gather(rank_segments, torch.tensor([segments], device="cuda", dtype=torch.int64))
        rank_segments = [int(x) for x in rank_segments]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        c = torch.tensor([segments, e2e_segments, launches, fallback, d2h_bytes], device="cuda", dtype=torch.int64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        dev_ms, e2e_s = float(t[0]), float(t[1])
        segments, e2e_segments, launches, fallback, d2h_bytes = (int(x) for x in c)

    line = None
    failed = False
    if rank == 0:
        value = segments / (dev_ms * 1e-3) / 1e6
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": workload_config(world), "clocks": clocks,
                "e2e": {"value": e2e_segments / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": 240 + 64,
                        "d2h_bytes_per_step": d2h_bytes,
                        "how": "okrt_set_render_data + okrt_render_frames + okrt_read_result_async per step; the host waits for step k-2's image "
                               "after submitting step k; %d untimed warm-up steps through the same calls" % args.warmup},
                "gpu_launches": launches,
                "ray_segments": segments, "fallback_segments": fallback, "traversal": args.traversal,
                "trace_kernel_ms": trace_ms, "trace_kernel_launches": trace_launches,
                "first_call_ms": first_call_ms, "unit_peaks": peaks,
                "layout_build": {"first_call": layout_first, "timed_region": layout_now,
                                 "note": "builder 1: on the device (clustered tree, ready for the first frame after a scene edit); 2: the host SAH "
                                         "builder's tree, built in the background and adopted before the timed region; 0: host builder, synchronous"}}
        if world > 1:
            line["rank_ms"] = rank_ms                 # device time of each rank's K steps: shows the load balance of the tile split
            line["rank_segments"] = rank_segments
            if band_balance:
                line["band_balance"] = band_balance
    if rank == 0 and world == 1:
        if not args.no_cpu_baseline:
            # ---- parity + CPU baseline: frame 1 of this workload, GPU vs oracle, every accumulation word
            cpu, o = oracle_frame(okrt, host, env, args.cpu_threads)
            line["cpu_baseline"] = cpu
            abi.check(r.L.okrt_reset_accumulation(r.h))
            abi.check(r.L.okrt_reset_stats(r.h))
            r.render(host["rd"].copy(), frames=1)
            acc = r.accumulation()
            gpu_segments = int(r.stats()["raySegments"])
            bad = np.any(acc.view(np.uint32) != o["accum"].view(np.uint32), axis=2)
            ok = bool(not bad.any()) and gpu_segments == int(o["segments"])
            line["parity"] = {"config": "C4 %dx%d mb%d, frame 1" % (WIDTH, HEIGHT, MAX_BOUNCES), "bit_exact": ok,
                              "pixels": int((o["accum"][..., 3] > 0).sum()), "differing_pixels": int(bad.sum()),
                              "segments": [gpu_segments, int(o["segments"])],
                              "against": "oracle (CPU restatement of RaytracerCS.hlsl): every RGBA32F accumulation word and the segment count"}
            failed = not ok
            # the same frame by the reference's own shader text compiled as C++ (oracle/_ref/libokrt_ref_shader.so, when it travelled)
            try:
                from oracle import binding as oracle_binding
                ref_shader = oracle_binding.load_ref_shader()
            except Exception:
                ref_shader = None
            if ref_shader is not None:
                scene = oracle_binding.oracle_scene(oracle_binding.load(), host, envFaces=env)
                rd1 = host["rd"].copy()
                rd1[0]["numAccumulationFrames"] = 1
                t0 = time.perf_counter()
                want = ref_shader.render(scene, rd1, MAX_BOUNCES)
                bad_ref = np.any(acc.view(np.uint32) != want["accum"].view(np.uint32), axis=2)
                line["parity"]["reference_shader"] = {"bit_exact": bool(not bad_ref.any()), "differing_pixels": int(bad_ref.sum()),
                                                      "seconds": round(time.perf_counter() - t0, 1),
                                                      "what": "RaytracerCS.hlsl + GPU-Utilities.hlsli + GPU-Structs.hlsli compiled as C++ (oracle/ref_shader_driver.cpp)"}
                failed = failed or bool(bad_ref.any())
        if args.traversal == "fast":
            line["roofline"] = trace_roofline(okrt, host, env, local_rank, alone["trace_ms"], alone["trace_launches"], alone["dev_ms"], alone["segments"], peaks)
            line["one_batch_in_flight"] = {"ms_per_step": alone["dev_ms"] / alone["steps"], "value": alone["segments"] / (alone["dev_ms"] * 1e-3) / 1e6,
                                           "note": "okrt_options.batchesInFlight = 1: the batches of the step run one after the other; the roofline's kernel times come from this pass"}
    r.close()
    if not args.no_configs:
        if world == 1 and rank == 0:
            line["configs"] = run_single_gpu_configs(okrt, local_rank, with_cpu=not args.no_cpu_baseline)
            failed = failed or any(c.get("parity") and not c["parity"]["bit_exact"] for c in line["configs"].values())
            c5 = run_c5(okrt, scene_parts, env, local_rank, 0, 1, None)
            c5["note"] += "; one GPU: no all-reduce"
            line["configs"]["C5"] = c5
        elif world > 1:
            c5 = run_c5(okrt, scene_parts, env, local_rank, rank, world, dist)
            if rank == 0:
                line["configs"] = {"C5": c5}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    if failed:
        print("bench.py: the GPU image differs from the oracle's (parity.bit_exact = false)", file=sys.stderr)
        sys.exit(1)


if __name__ == "__main__":
    main()
