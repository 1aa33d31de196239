"""Row f2: device-timed okrt_geometry_build_device against okrt_geometry_build (the reference's SAH tree, CPU threads)
on the two large meshes of BASELINE.json's configs, plus the render throughput each tree gives.
Usage: python tools/builder_bench.py  -> one JSON line per scene."""
import importlib
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
okrt = importlib.import_module("gpu-raytracer_b200")
import importlib; helpers = importlib.import_module("gpu-raytracer_b200.harness")  # noqa: E402


def main():
    W, H = 1920, 1080
    for name, factory, bounces in (("dragon_871k", okrt.scenes.dragon_scene, 4), ("sponza_standin_262k", okrt.scenes.sponza_standin, 8)):
        rm, sc, info = factory()
        env = okrt.scenes.procedural_sky_faces(256)
        row = {"scene": name, "triangles": int(sum(len(m.trianglesPos) for m in rm.meshes))}
        for label, dev in (("cpu_sah", None), ("device_lbvh", 0)):
            t0 = time.perf_counter()
            host = helpers.build_host_side(okrt, rm, sc, info, W, H, bvhDevice=dev)
            wall = (time.perf_counter() - t0) * 1e3
            geo = host["geometry"]
            if dev is not None:      # second call: context and allocator warm
                geo2 = okrt.Geometry(rm.meshes, *info.get("bvh", (30, 5)), device=dev)
                row[label + "_build_ms_device_timed"] = round(geo2.buildMs, 2)
            row[label + "_host_side_wall_ms"] = round(wall, 1)
            row[label + "_nodes"] = int(len(geo.bvhNodes))
            r = helpers.Renderer(okrt, host, W, H, envFaces=env)
            r.options(maxBounces=bounces)
            r.render(host["rd"].copy(), frames=8)
            r.render(host["rd"].copy(), frames=32)
            st = r.stats()
            row[label + "_mrays_s"] = round(float(st["raySegments"]) / float(st["lastRenderMs"]) / 1e3, 1)
            row[label + "_fallback_share"] = float(st["fallbackSegments"]) / max(float(st["raySegments"]), 1.0)
            r.close()
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
