"""Generates the committed golden fixtures.  Run HERE (where /root/reference exists):
    python tests/golden/make_golden.py
* bvh_*.npz      — node arrays + leaf order produced by the REFERENCE's own BvhBuilder.cpp (compiled unmodified into
                   oracle/_ref by oracle/Makefile) for seeded meshes: the pin for row a10.
* glm_*.npz      — Transform::calculateMatrix / calculateProjectionData values from the reference's vendored glm.
* kat_*.npz      — known-answer vectors of the oracle (pcg_hash sequence, ray/triangle, ray/box, ray/sphere edge
                   cases, canonical math) — regression pins for the oracle itself (the reference ships none).
* render_*.npz   — oracle AOV + radiance of small scenes: lets the GPU box check itself without /root/reference.
"""
import ctypes
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))


def golden_meshes(okrt):
    sc = okrt.scenes
    rng = np.random.default_rng(0)
    c = rng.uniform(-10, 10, (3000, 1, 3))
    soup = c + rng.normal(0, 0.3, (3000, 3, 3))
    flat = sc.make_mesh(*sc.quad((-5, 0, -5), (5, 0, -5), (5, 0, 5), (-5, 0, 5)))
    dup = np.repeat(rng.uniform(-1, 1, (1, 3, 3)), 40, axis=0)           # 40 identical triangles: SAH refuses to split
    return {"icosphere3": (sc.make_mesh(*sc.icosphere(3, 4.0, (0, 5, 4))), 30, 5),
            "soup3000": (sc.make_mesh(soup), 30, 5),
            "soup3000_d10_l2": (sc.make_mesh(soup), 10, 2),
            "knot8k": (sc.knot_mesh(8000), 30, 5),
            "box": (sc.make_mesh(*sc.box((-1, -2, -3), (1, 2, 3))), 30, 5),
            "quad": (flat, 30, 5),
            "duplicates": (sc.make_mesh(dup), 30, 5),
            "depth2": (sc.make_mesh(soup[:500]), 2, 1)}


def main():
    okrt = importlib.import_module("gpu-raytracer_b200")
    abi = okrt.abi
    from tests import oracle_binding as ob, helpers
    ref, orc = ob.load_ref(), ob.load()
    assert ref is not None, "needs /root/reference (oracle/_ref)"

    for name, (mesh, depth, leaf) in golden_meshes(okrt).items():
        tp = mesh.trianglesPos
        n = len(tp)
        lo, hi = mesh.boundingBox
        nodes = np.zeros(2 * n + 1, abi.BVH_NODE)
        order = np.zeros(n, np.uint32)
        cnt = ref.ref_build_bvh(abi.ptr(tp), n, abi.ptr(lo.copy()), abi.ptr(hi.copy()), depth, leaf, abi.ptr(nodes), len(nodes), abi.ptr(order))
        np.savez_compressed(os.path.join(HERE, "bvh_%s.npz" % name), nodes=nodes[:cnt], order=order, maxDepth=depth, maxLeaf=leaf)
        print("bvh", name, n, "tris", cnt, "nodes")

    rng = np.random.default_rng(1)
    rows = []
    for it in range(64):
        pos = rng.uniform(-50, 50, 3).astype(np.float32)
        rot = rng.uniform(-180, 180, 3).astype(np.float32)
        scl = rng.uniform(0.2, 3, 3).astype(np.float32)
        if it < 8:
            rot[:] = 0; scl[:] = 1
        fov = np.float32(rng.uniform(30, 120)); nz = np.float32(rng.uniform(0.01, 1))
        W, H = [(1920, 1080), (512, 512), (1600, 900)][it % 3]
        T = np.zeros(16, np.float32); Ti = np.zeros(16, np.float32)
        ref.ref_transform_matrices(abi.ptr(pos), abi.ptr(rot), abi.ptr(scl), abi.ptr(T), abi.ptr(Ti))
        vp = np.zeros(2, np.float32); ip = np.zeros(16, np.float32); iv = np.zeros(16, np.float32); r3 = np.zeros(3, np.float32); u3 = np.zeros(3, np.float32)
        ref.ref_projection_data(abi.ptr(pos), abi.ptr(rot), float(fov), float(nz), 1000.0, W, H, abi.ptr(vp), abi.ptr(ip), abi.ptr(iv), abi.ptr(r3), abi.ptr(u3))
        rows.append(dict(pos=pos, rot=rot, scl=scl, fov=fov, nz=nz, W=W, H=H, T=T, Ti=Ti, vp=vp, ip=ip, iv=iv, right=r3, up=u3))
    np.savez_compressed(os.path.join(HERE, "glm_transforms.npz"), **{k: np.array([r[k] for r in rows]) for k in rows[0]})
    sizes = np.zeros(10, np.uint32); ref.ref_struct_sizes(abi.ptr(sizes))
    mat = np.zeros(1, abi.MATERIAL); ref.ref_material_default(abi.ptr(mat))
    np.savez_compressed(os.path.join(HERE, "ref_structs.npz"), sizes=sizes, material=mat)

    # oracle KATs
    seed = ctypes.c_uint32(1)
    hashes = np.array([orc.lib.oracle_pcg_hash(ctypes.byref(seed)) for _ in range(16)], np.uint32)
    seed = ctypes.c_uint32(12345)
    floats = np.array([orc.lib.oracle_random_float(ctypes.byref(seed)) for _ in range(16)], np.float32)
    x = np.concatenate([np.linspace(1e-9, 1, 200), np.linspace(1, 50, 200), [0.0, 1e-40, 2.0 ** -32]]).astype(np.float32)
    ang = np.linspace(-7, 7, 501).astype(np.float32)
    out = {}
    for which, name, arg in ((0, "log", x), (1, "cos", ang), (2, "sin", ang), (3, "exp", np.linspace(-90, 80, 301).astype(np.float32))):
        o = np.zeros_like(arg); orc.lib.oracle_math_eval(which, abi.ptr(arg), abi.ptr(arg), abi.ptr(o), len(arg)); out[name + "_x"] = arg; out[name + "_y"] = o
    px = np.linspace(0, 1, 101).astype(np.float32); five = np.full_like(px, 5.0); o = np.zeros_like(px)
    orc.lib.oracle_math_eval(5, abi.ptr(px), abi.ptr(five), abi.ptr(o), len(px)); out["pow5_x"] = px; out["pow5_y"] = o
    np.savez_compressed(os.path.join(HERE, "kat_rng_math.npz"), hashes=hashes, floats=floats, **out)

    # small oracle renders (TestScene C1 at reduced size + sponza cut), canonical math
    for name, build, W, H, bounces, spheres in (("testscene", lambda: okrt.scenes.test_scene(with_textures=True), 96, 54, 4, True),
                                                ("sponza20k", lambda: okrt.scenes.sponza_standin(20000, texture_size=32), 96, 54, 8, False)):
        rm, sc, info = build()
        host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=spheres)
        env = okrt.scenes.procedural_sky_faces(16)
        rd = host["rd"].copy(); rd[0]["numAccumulationFrames"] = 1
        o = orc.render(helpers.oracle_scene(orc, host, envFaces=env), rd, bounces, aov_dtype=abi.AOV)
        np.savez_compressed(os.path.join(HERE, "render_%s.npz" % name), aov=o["aov"], accum=o["accum"], result=o["result"],
                            scene_digest=helpers.scene_digest(host, env),
                            segments=o["segments"], W=W, H=H, bounces=bounces)
        print("render", name, o["segments"], "segments")

    glass_fixture(okrt)


def glass_fixture(okrt):
    """resources/meshes/ikea_glass.obj as the product's OBJ reader delivers it (scale 10, Application.cpp:42):
    positions, normals and uvs of the three cups, so that BASELINE configs[2] runs on the real asset where
    /root/reference does not exist.  Tangents are not stored (no normal map is bound; the loader fills them)."""
    path = "/root/reference/resources/meshes/ikea_glass.obj"
    objs = okrt.loadObjects(path, 10.0)
    np.savez_compressed(os.path.join(HERE, "ikea_glass_mesh.npz"),
                        positions=np.concatenate([o.mesh.trianglesPos["position"] for o in objs]),
                        normals=np.concatenate([o.mesh.trianglesInfo["vertexInfo"]["normal"] for o in objs]),
                        uvs=np.concatenate([o.mesh.trianglesInfo["vertexInfo"]["uv"] for o in objs]),
                        counts=np.array([len(o.mesh.trianglesPos) for o in objs], np.uint32),
                        names=np.array([o.name for o in objs]))
    print("ikea_glass", [len(o.mesh.trianglesPos) for o in objs], "triangles")


if __name__ == "__main__":
    import sys
    if "glass" in sys.argv[1:]:
        glass_fixture(importlib.import_module("gpu-raytracer_b200"))
    else:
        main()
