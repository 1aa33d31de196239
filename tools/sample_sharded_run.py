#!/usr/bin/env python
"""BASELINE.json configs[4] (SURVEY.md §8d C5, §8e "sample mode"): the Sponza-class stand-in at 3840x2160, maxBounces 8,
SPP progressive frames, SAMPLE-sharded over the ranks of one box: rank r renders global frames r+1, r+1+N, ... into its
own RGBA32F accumulation image, ONE NCCL all-reduce (SUM) over NVLink adds the images, okrt_resolve writes
saturate(u1 / SPP) (RaytracerCS.hlsl:529).  Device-timed; one JSON line from rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/sample_sharded_run.py --spp 1024 --check-spp 16

--check-spp K also renders K frames both ways (sample-sharded + all-reduce vs. sequentially on rank 0) and reports the
RMSE of the mean image (tolerance 1e-4: float summation order differs) and whether the sample counts are equal.
"""
import argparse
import ctypes
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--check-spp", type=int, default=16)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--bounces", type=int, default=8)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    rank, local_rank, world = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1")))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    okrt = importlib.import_module("gpu-raytracer_b200")
    abi = okrt.abi
    import importlib; helpers = importlib.import_module("gpu-raytracer_b200.harness")
    W, H = args.width, args.height
    assert args.spp % world == 0 and args.check_spp % world == 0, "frames must split evenly over the ranks"

    rm, sc, info = okrt.scenes.sponza_standin()
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(512)
    r = helpers.Renderer(okrt, host, W, H, envFaces=env, device=local_rank)
    L = r.L
    r.options(maxBounces=args.bounces, coverAllPixels=0, **okrt.sharding.sample_options(rank, world))
    rd = host["rd"].copy()
    acc = okrt.sharding.accumulation_tensor(L.okrt_accumulation_device_ptr(r.h), W, H, torch.device("cuda", local_rank))

    def barrier():
        abi.check(L.okrt_sync(r.h))
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def sharded(total):
        """total frames over all ranks -> (render ms of this rank, all-reduce ms, resolve ms, segments of this rank)"""
        abi.check(L.okrt_reset_accumulation(r.h))
        abi.check(L.okrt_set_render_data(r.h, abi.ptr(rd)))
        abi.check(L.okrt_reset_stats(r.h))
        barrier()
        abi.check(L.okrt_render_frames(r.h, total // world))
        abi.check(L.okrt_sync(r.h))
        st = r.stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        okrt.sharding.all_reduce_accumulation(acc)           # the path's single exchange step
        e1.record()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        abi.check(L.okrt_resolve(r.h, total))
        abi.check(L.okrt_sync(r.h))
        return float(st["lastRenderMs"]), e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3, int(st["raySegments"]), int(st["fallbackSegments"])

    # warm-up: kernels, wavefront buffers, NCCL communicator and its NVLink buffers
    sharded(4 * world)
    sharded(4 * world)

    check = None
    if args.check_spp > 0:
        sharded(args.check_spp)
        summed = r.accumulation()
        if rank == 0:
            r.options(frameOffset=0, frameStride=1)
            abi.check(L.okrt_reset_accumulation(r.h))
            r.render(rd, frames=args.check_spp)
            want = r.accumulation()
            r.options(**okrt.sharding.sample_options(rank, world))
            k = float(args.check_spp)
            rmse = float(np.sqrt(np.mean((summed[..., :3].astype(np.float64) / k - want[..., :3].astype(np.float64) / k) ** 2)))
            check = {"frames": args.check_spp, "rmse_of_mean_vs_sequential": rmse, "tolerance": 1e-4,
                     "sample_counts_equal": bool(np.array_equal(summed[..., 3], want[..., 3])),
                     "bit_identical": bool(np.array_equal(summed.view(np.uint32), want.view(np.uint32)))}
        barrier()

    render_ms, reduce_ms, resolve_ms, segs, fb = sharded(args.spp)
    t = torch.tensor([render_ms, reduce_ms, resolve_ms], device="cuda", dtype=torch.float64)
    c = torch.tensor([segs, fb], device="cuda", dtype=torch.int64)
    per_rank = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, t)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
    else:
        per_rank = [t]
    if rank == 0:
        total_ms = float(t[0] + t[1] + t[2])
        line = {"metric": "Sponza 4K %d spp sample-sharded Mrays/s (device-timed render + NCCL all-reduce + resolve)" % args.spp,
                "value": int(c[0]) / (total_ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
                "config": {"workload": "sponza-standin %dx%d maxBounces=%d, %d progressive frames, %d per rank (BASELINE configs[4])" % (W, H, args.bounces, args.spp, args.spp // world),
                           "sharding": "sample: rank r renders frames r+1, r+1+N, ...; one all_reduce(SUM) of the %d MB RGBA32F accumulation image" % (W * H * 16 // 1000000)},
                "render_ms_max": float(t[0]), "all_reduce_ms_max": float(t[1]), "resolve_ms_max": float(t[2]),
                "render_only_mrays": int(c[0]) / (float(t[0]) * 1e-3) / 1e6,
                "rank_render_ms": [round(float(x[0]), 2) for x in per_rank], "ray_segments": int(c[0]), "fallback_segments": int(c[1]),
                "check": check}
        print(json.dumps(line))
    r.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
