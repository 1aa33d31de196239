"""Per-frame device time and segment counts (diagnostic)."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
okrt = importlib.import_module("gpu-raytracer_b200")
abi = okrt.abi
import importlib; helpers = importlib.import_module("gpu-raytracer_b200.harness")
import bench
host, env, info = bench.build_scene(okrt)
r = helpers.Renderer(okrt, host, bench.WIDTH, bench.HEIGHT, envFaces=env)
r.options(maxBounces=8, traversal=int(sys.argv[1]) if len(sys.argv) > 1 else 0)
rd = host["rd"].copy()
abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))
for f in range(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    abi.check(r.L.okrt_reset_stats(r.h))
    t0 = time.perf_counter()
    abi.check(r.L.okrt_render_frames(r.h, 1))
    abi.check(r.L.okrt_sync(r.h))
    wall = (time.perf_counter() - t0) * 1e3
    st = r.stats()
    print("frame %2d  dev %.3f ms  wall %.3f ms  segments %d  fallback %d  pixels %d" % (f + 1, st["lastRenderMs"], wall, st["raySegments"], st["fallbackSegments"], st["pixelsTraced"]))
