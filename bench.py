#!/usr/bin/env python
"""bench.py — Sponza 1080p 8-bounce Mrays/s (device-timed) at 1/2/4/8 B200 vs CPU ref.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[3]): the Sponza-class stand-in (sponza.obj is not
shipped: /root/reference/.MISSING_LARGE_BLOBS) — ~262k triangles, 354 mesh entities, 27 texture layers,
procedural sky cube — 1920x1080, maxBounces 8, oct-tree (2,50), BVH (30,5), camera (60,0,0) looking -x.
A step is FRAMES_PER_STEP progressive frames (1 spp each) of that image.  With N GPUs the image is
tile-sharded in interleaved 8-row bands (no data-path collective); value = ray segments traced by all
ranks / max-over-ranks device time.  A ray segment is one findClosestHit call (RaytracerCS.hlsl:446).
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


def workload_config(n_gpus):
    return {
        "workload": "sponza-standin 1920x1080 maxBounces=8 (BASELINE configs[3]); %d progressive frames (1 spp) per step" % FRAMES_PER_STEP,
        "scene": "procedural Sponza-class stand-in, seed 11: sponza.obj is listed in .MISSING_LARGE_BLOBS",
        "resolution": [WIDTH, HEIGHT], "max_bounces": MAX_BOUNCES, "frames_per_step": FRAMES_PER_STEP,
        "bvh": [30, 5], "octree": [2, 50],
        "sharding": "tile: interleaved 8-row bands, rank = band %% %d, scene replicated, no collective" % n_gpus,
        "l2": "no flush: the working set (geometry + 27x903x903 RGBA8 textures + 1080p RGBA32F/RGBA8 images, ~230 MB) is larger than the 126 MB L2",
    }


def build_scene(okrt):
    helpers = okrt.harness
    rm, sc, info = okrt.scenes.sponza_standin()
    host = helpers.build_host_side(okrt, rm, sc, info, WIDTH, HEIGHT)
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


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes_per_segment(visits, pixels_frames):
    """B_seg = 36 N_box + 36 N_tri + 232 N_inst + 48 N_oct + [hit](132 + 68) + 16 per texture tap,
    with the ORACLE's visit counts in reference order, plus 36 B per pixel per frame (accumulate + store).
    Returns (whole path, traversal only): the traversal part (rows a4) is what the dominant kernel is charged with."""
    seg = max(1, visits["segments"])
    trav = 36 * visits["bvhNodes"] + 36 * visits["triangles"] + 232 * visits["instances"] + 48 * visits["octNodes"]
    rest = 200 * visits["hits"] + 16 * visits["textureTaps"] + 36 * pixels_frames
    return (trav + rest) / seg, trav / seg


def cpu_baseline_sample(okrt, host, env, threads=0):
    """The oracle (kind "port": scalar restatement of the HLSL) on one full 1080p frame, 1 spp, all host threads."""
    from oracle import binding as oracle_binding
    oracle = oracle_binding.load()
    scene = oracle_binding.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()
    rd[0]["numAccumulationFrames"] = 1
    t0 = time.perf_counter()
    out = oracle.render(scene, rd, MAX_BOUNCES, threads=threads)
    dt = time.perf_counter() - t0
    cores = threads if threads > 0 else (os.cpu_count() or 1)
    return {"value": out["segments"] / dt / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "one full 1920x1080 frame, 1 spp, %d ray segments in %.1f s" % (out["segments"], dt)}, out["visits"]


def run_reference_arm(args):
    """--impl reference: the reference's CPU-runnable implementation of the path = the oracle port (the HLSL
    itself needs D3D11: unavailable on this box), all host threads, each step a bounded sample of the workload."""
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

    def step(frame):
        rd[0]["numAccumulationFrames"] = frame
        return oracle.render(scene, rd, MAX_BOUNCES, row_list=rows)["segments"]
    for w in range(args.warmup):
        step(w + 1)
    t0 = time.perf_counter()
    segments = sum(step(args.warmup + k + 1) for k in range(args.steps))
    dt = time.perf_counter() - t0
    value = segments / dt / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / max(1, args.steps) * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "per step: every 8th 8-row band of the 1080p frame (1/8 of the pixels), 1 spp, %d host threads" % cores},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--traversal", default="fast", choices=["fast", "exact"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
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

    host, env, info = build_scene(okrt)
    r = helpers.Renderer(okrt, host, WIDTH, HEIGHT, envFaces=env, device=local_rank)
    opts = dict(maxBounces=MAX_BOUNCES, traversal=abi.TRAVERSAL_FAST if args.traversal == "fast" else abi.TRAVERSAL_EXACT, coverAllPixels=0)
    if world > 1:
        opts.update(shardMode=abi.SHARD_TILE, shardRank=rank, shardCount=world, shardBandRows=8)
    r.options(**opts)
    rd = host["rd"].copy()

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()
        abi.check(r.L.okrt_sync(r.h))

    # ---- device-timed arm: inputs resident, K steps back to back on the renderer's stream
    abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))
    abi.check(r.L.okrt_render_frames(r.h, args.warmup * FRAMES_PER_STEP))
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

    # ---- end-to-end arm: host constants in, result image out, every step, wall clock around the public call
    pinned_out = np.zeros((HEIGHT, WIDTH, 4), np.uint8)
    abi.check(r.L.okrt_host_register(abi.ptr(pinned_out), pinned_out.nbytes))     # page-locked: the D2H copy is a plain DMA
    abi.check(r.L.okrt_reset_stats(r.h))
    moved = ctypes.c_uint64(0)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        rd[0]["numAccumulationFrames"] = r.L.okrt_get_num_accumulation_frames(r.h)
        abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))                  # H2D: 240 B of constants travel with the launch
        abi.check(r.L.okrt_render_frames(r.h, FRAMES_PER_STEP))
        # D2H: the RGBA8 result image; a tile-sharded rank moves only the row bands it owns into the shared image layout
        abi.check(r.L.okrt_read_result_owned(r.h, abi.ptr(pinned_out), ctypes.byref(moved)))
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_segments = int(r.stats()["raySegments"])
    abi.check(r.L.okrt_host_unregister(abi.ptr(pinned_out)))
    d2h_bytes = int(moved.value)                   # per step, this rank; summed over ranks below = one whole image

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

    if rank == 0:
        value = segments / (dev_ms * 1e-3) / 1e6
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": workload_config(world), "clocks": clocks,
                "e2e": {"value": e2e_segments / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": 240 + 64,
                        "d2h_bytes_per_step": d2h_bytes},
                "gpu_launches": launches,
                "ray_segments": segments, "fallback_segments": fallback, "traversal": args.traversal,
                "trace_kernel_ms": trace_ms, "trace_kernel_launches": trace_launches}
        if world > 1:
            line["rank_ms"] = rank_ms                 # device time of each rank's K steps: shows the load balance of the tile split
            line["rank_segments"] = rank_segments
        if world == 1 and not args.no_cpu_baseline:
            cpu, visits = cpu_baseline_sample(okrt, host, env, args.cpu_threads)
            line["cpu_baseline"] = cpu
            bseg_path, bseg = algorithmic_bytes_per_segment(visits, WIDTH * HEIGHT)
            peak, which = measured_peak_gbs()
            achieved = bseg * segments / (max(trace_ms, 1e-6) * 1e-3) / 1e9
            traffic, units = None, None
            tpath = os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")
            if os.path.exists(tpath):
                try:
                    with open(tpath) as f:
                        tj = json.load(f)
                        units = tj.get("traceFastKernel_units")
                        # DRAM bytes of the captured launch, scaled to this run's average launch by duration
                        traffic = tj["traceFastKernel"] * (trace_ms / max(1, trace_launches)) / tj["traceFastKernel_ms"]
                except Exception:
                    traffic = None
            line["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                "traffic": traffic, "peak_source": which, "kernel": "traceFastKernel",
                                "kernel_launches": trace_launches, "kernel_ms_per_launch": trace_ms / max(1, trace_launches),
                                "kernel_share_of_step": trace_ms / dev_ms,
                                "algorithmic_bytes_per_launch": bseg * segments / max(1, trace_launches),
                                "unit_utilisation_from_ncu": units,      # the ray-throughput view: how busy the kernel's busiest units are (committed capture, not live)
                                "algorithmic_bytes_per_segment": bseg, "algorithmic_bytes_per_segment_whole_path": bseg_path,
                                "note": "algorithmic bytes = what the reference's traversal order reads per segment (SURVEY.md 8d); the "
                                        "kernel visits far fewer nodes, and the scene is L2-resident, so DRAM traffic is much lower"}
            # north_star: "a fraction of the ray-throughput roofline".  The traversal kernel's busiest unit is the L1 data stage
            # (one 32-byte sector per cycle per SM for divergent lanes); its utilisation in the committed ncu capture of this
            # same command IS achieved / peak of that bound: peak = the traversal rate at which that unit would saturate
            # with this kernel's sectors per ray.
            if units and units.get("l1tex_data_stage_lsu_pct"):
                rate = segments / (max(trace_ms, 1e-6) * 1e-3) / 1e6
                f_l1 = units["l1tex_data_stage_lsu_pct"] / 100.0
                line["roofline_rays"] = {"bound": "l1-data-stage", "achieved": rate, "peak": rate / f_l1, "unit": "Mrays/s (traversal kernel only)",
                                         "frac": f_l1, "issue_slots_frac": units.get("issue_slots_pct", 0) / 100.0,
                                         "lanes_per_instruction": units.get("lanes_per_instruction"), "source": units.get("source"),
                                         "note": "frac = unit utilisation from the committed ncu --set full capture (not live); achieved = live, event-timed"}
            # FP32-issue roofline.  Algorithmic flops per segment in REFERENCE order:
            # 25 per box test (oct-tree, instance root and BVH node boxes), 51 per triangle test, 20 per accepted candidate
            # (3x3 transform + length, counted once per hit segment); peak = SMs x 128 lanes x 2 x f_clk.
            seg_n = max(1, visits["segments"])
            flops_seg = (25.0 * (visits["bvhNodes"] + visits["instances"] + visits["octNodes"]) + 51.0 * visits["triangles"]
                         + 20.0 * visits["hits"]) / seg_n
            sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
            peak_tf = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
            ach_tf = flops_seg * segments / (max(trace_ms, 1e-6) * 1e-3) / 1e12
            line["roofline_fp32"] = {"bound": "fp32-issue", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf,
                                     "algorithmic_flops_per_segment": flops_seg, "kernel": "traceFastKernel",
                                     "visits_per_segment": {k: visits[k] / seg_n for k in ("bvhNodes", "triangles", "instances", "octNodes", "hits", "textureTaps")},
                                     "note": "reference-order flops (SURVEY.md 8d) over the traversal kernel's event-timed launches; the kernel itself runs ~3x "
                                             "fewer box tests than the reference order, so its executed FP32 rate is lower than this figure"}
        print(json.dumps(line))
    r.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
