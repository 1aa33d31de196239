"""Headless equivalent of running the reference's application on a scene and pressing "Save screenshot":
OBJ import (row f1) or a built-in scene -> BVH (CPU SAH = the reference's tree, or --device-bvh = row f2) -> oct-tree ->
progressive frames on the GPU -> render.png (row f3), optionally with the BVH / oct-tree box overlay (row f4).

    python tools/render.py --scene sponza --frames 64 --out render.png
    python tools/render.py --obj path/to/mesh.obj --scale 10 --frames 256 --bounces 16 --overlay bvh
"""
import argparse
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
okrt = importlib.import_module("gpu-raytracer_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="test", choices=["test", "sponza", "dragon", "glass"])
    ap.add_argument("--obj", default=None, help="OBJ file (with its .mtl) instead of a built-in scene: glass material, scale --scale")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--bounces", type=int, default=8)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--device-bvh", action="store_true", help="build the BVH on the GPU (LBVH) instead of the reference's SAH tree")
    ap.add_argument("--overlay", default="none", choices=["none", "bvh", "octree"])
    ap.add_argument("--out", default="render.png")
    ap.add_argument("--state", default=None, help="accumulation file: resumed from if it exists, written at the end")
    a = ap.parse_args()

    if a.obj:
        rm, sc, info = okrt.scenes.glass_scene_from_obj(a.obj, a.scale)
    else:
        rm, sc, info = {"test": okrt.scenes.test_scene, "sponza": okrt.scenes.sponza_standin, "dragon": okrt.scenes.dragon_scene,
                        "glass": okrt.scenes.glass_scene}[a.scene]()
    rt = okrt.RayTracer(a.device)
    t0 = time.perf_counter()
    rt.initiate(a.width, a.height, rm, environmentMapFaces=okrt.scenes.procedural_sky_faces(256), maxBounces=a.bounces)
    if a.device_bvh:
        rt.geometry = okrt.Geometry(rm.meshes, 30, 5, device=a.device)
        g, L = rt.geometry, okrt.abi.lib()
        okrt.abi.check(L.okrt_set_triangles(rt._handle, okrt.abi.ptr(g.trianglesPos), okrt.abi.ptr(g.trianglesInfo), len(g.trianglesPos)))
        okrt.abi.check(L.okrt_set_bvh_nodes(rt._handle, okrt.abi.ptr(g.bvhNodes), len(g.bvhNodes)))
    rt.setScene(sc)
    rt.createOctTree(sc, *info.get("octree", (2, 50)))
    t1 = time.perf_counter()
    if a.state and os.path.exists(a.state):
        rt.loadAccumulation(a.state)
    rt.render(frames=a.frames)
    rt.sync()
    t2 = time.perf_counter()
    if a.state:
        rt.saveAccumulation(a.state)
    if a.overlay == "bvh":
        for e in range(len(sc.meshEntities)):
            rt.renderBvhNodeBBs(e, 0, rt.DrawMode_DrawWithChildren)
    elif a.overlay == "octree":
        rt.renderOctTreeNodeBBs(0, rt.DrawMode_DrawWithDecendants)
    rt.saveScreenshot(a.out)
    print("scene setup %.2f s, %d frames in %.2f s (%d accumulated), wrote %s" % (t1 - t0, a.frames, t2 - t1, rt.getNumAccumulationFrames(), a.out))
    rt.shutdown()


if __name__ == "__main__":
    main()
