"""One GPU acting as rank 0 of 8 (tile sharding): what a 16-frame step costs against a 128-frame batch, with and without the
result read-back.  usage: python tools/small_batch.py [shardCount]"""
import ctypes, importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
okrt = importlib.import_module("gpu-raytracer_b200")
abi = okrt.abi
W, H = 1920, 1080
count = int(sys.argv[1]) if len(sys.argv) > 1 else 8
parts = okrt.scenes.sponza_standin()
host = okrt.harness.build_host_side(okrt, *parts, W, H)
env = okrt.scenes.procedural_sky_faces(512)
r = okrt.harness.Renderer(okrt, host, W, H, envFaces=env, device=0)
opts = dict(maxBounces=8, traversal=abi.TRAVERSAL_FAST, coverAllPixels=0)
if count > 1:
    opts.update(shardMode=abi.SHARD_TILE, shardRank=0, shardCount=count, shardBandRows=8)
r.options(**opts)
rd = host["rd"].copy()
L = r.L
abi.check(L.okrt_set_render_data(r.h, abi.ptr(rd)))
abi.check(L.okrt_render_frames(r.h, 16)); abi.check(L.okrt_wait_layout(r.h)); abi.check(L.okrt_render_frames(r.h, 64)); abi.check(L.okrt_sync(r.h))
pinned = [np.zeros((H, W, 4), np.uint8) for _ in range(3)]
for b in pinned: abi.check(L.okrt_host_register(abi.ptr(b), b.nbytes))
moved = ctypes.c_uint64(0)

def run(per_call, calls, read):
    abi.check(L.okrt_sync(r.h)); abi.check(L.okrt_reset_stats(r.h))
    t0 = time.perf_counter()
    for k in range(calls):
        abi.check(L.okrt_render_frames(r.h, per_call))
        if read:
            abi.check(L.okrt_read_result_async(r.h, abi.ptr(pinned[k % 3]), ctypes.byref(moved), k % 3))
            if k >= 2: abi.check(L.okrt_wait_read(r.h, (k - 2) % 3))
    if read:
        for k in range(max(0, calls - 2), calls): abi.check(L.okrt_wait_read(r.h, k % 3))
    t_submit = time.perf_counter() - t0
    abi.check(L.okrt_sync(r.h))
    dt = time.perf_counter() - t0
    seg = int(r.stats()["raySegments"])
    print("%4d frames x %3d calls read=%d: %8.3f ms per 16 frames  (%7.1f Mrays/s this shard, submit %.1f ms of %.1f ms)" %
          (per_call, calls, read, dt * 1e3 / (per_call * calls / 16), seg / dt / 1e6, t_submit * 1e3, dt * 1e3), flush=True)

for rep in range(2):
    run(320, 1, 0)
    run(128, 2, 0)
    run(64, 4, 0)
    run(32, 8, 0)
    run(16, 20, 0)
    run(16, 20, 1)
    run(8, 40, 0)
r.close()
