"""Host-only: traversal-layout statistics (okrt_layout_report) of the bench scene and the dragon scene.
Usage: python tools/layout_report.py [sponza|dragon|glass] ; environment knobs of okrt_accel.cpp apply."""
import importlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
okrt = importlib.import_module("gpu-raytracer_b200")
import importlib; helpers = importlib.import_module("gpu-raytracer_b200.harness")  # noqa: E402


def report(host):
    abi = okrt.abi
    g, t = host["geometry"], host["octree"]
    info = np.zeros(1, abi.LAYOUT_INFO)
    abi.check(abi.lib().okrt_layout_report(abi.ptr(g.trianglesPos), len(g.trianglesPos), abi.ptr(g.bvhNodes), len(g.bvhNodes),
                                           abi.ptr(t.instances), len(t.instances), abi.ptr(t.nodes), len(t.nodes), abi.ptr(info)))
    return info[0]


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "sponza"
    factory = {"sponza": okrt.scenes.sponza_standin, "dragon": okrt.scenes.dragon_scene, "glass": okrt.scenes.glass_scene}[which]
    rm, sc, info = factory()
    host = helpers.build_host_side(okrt, rm, sc, info, 1920, 1080)
    r = report(host)
    print({k: (float(r[k]) if r.dtype[k].kind == "f" else int(r[k])) for k in r.dtype.names})


if __name__ == "__main__":
    main()
