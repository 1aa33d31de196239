"""Row f2 of SURVEY.md §8: okrt_geometry_build_device, the LBVH builder that runs on the GPU and emits the
reference's own node / triangle formats (RayTracer.h:15-29, BvhBuilder.cpp:12-36 for the conventions).

The tree is not the reference's SAH tree, so the checks are (1) structural: the arrays are a valid reference-format
BVH over exactly the input triangles, with exact float boxes; (2) parity on the same arrays: the CPU oracle walking
THIS node array in the reference's order equals the kernels bit for bit; (3) against the reference's own tree: same
primary hits except at exact distance ties, image close (PSNR)."""
import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu

INVALID = 0xFFFFFFFF


def _check_tree(geo, meshes, maxLeaf):
    nodes, pos = geo.bvhNodes, geo.trianglesPos
    assert len(geo.meshDescs) == len(meshes)
    for desc, mesh in zip(geo.meshDescs, meshes):
        t0, t1, n0, nn = (int(desc[k]) for k in ("startIdx", "endIdx", "bvhTreeStartIdx", "numBvhNodes"))
        assert t1 - t0 == len(mesh.trianglesPos)
        # the leaf-ordered triangles are a permutation of the mesh's triangles (positions and vertex info move together)
        src = np.ascontiguousarray(mesh.trianglesPos).view(np.uint32).reshape(-1, 9)
        dst = np.ascontiguousarray(pos[t0:t1]).view(np.uint32).reshape(-1, 9)
        assert np.array_equal(src[np.lexsort(src.T)], dst[np.lexsort(dst.T)])
        pts = np.ascontiguousarray(pos[t0:t1]).view(np.float32).reshape(-1, 3, 3)
        tlo, thi = pts.min(axis=1), pts.max(axis=1)
        seen = np.zeros(t1 - t0, np.int32)
        reached = np.zeros(nn, bool)
        stack = [(n0, None, None)]
        while stack:
            idx, plo, phi = stack.pop()
            assert n0 <= idx < n0 + nn
            reached[idx - n0] = True
            nd = nodes[idx]
            lo, hi = np.asarray(nd["boundingBox"]["min"]), np.asarray(nd["boundingBox"]["max"])
            if plo is not None:
                assert np.all(lo >= plo) and np.all(hi <= phi), "a parent box contains its child box exactly"
            fc = int(nd["firstChildIdx"])
            if fc == INVALID:
                a, b = int(nd["triStart"]) - t0, int(nd["triEnd"]) - t0
                assert 0 <= a < b <= t1 - t0 and (b - a <= maxLeaf or nn == 1)
                seen[a:b] += 1
                if idx != n0:   # the root keeps the mesh's own bounding box, like the reference's
                    assert np.array_equal(lo, tlo[a:b].min(axis=0)) and np.array_equal(hi, thi[a:b].max(axis=0)), "leaf box = exact min/max"
            else:
                stack.append((fc, lo, hi)); stack.append((fc + 1, lo, hi))
        assert np.all(seen == 1), "every triangle sits in exactly one leaf"
        assert reached.all(), "no unreachable node slots"


def test_device_builder_structure(okrt):
    rm, sc, info = okrt.scenes.sponza_standin(40_000, texture_size=64)
    geo = okrt.Geometry(rm.meshes, 30, 5, device=0)
    assert geo.buildMs is not None and geo.buildMs > 0
    _check_tree(geo, rm.meshes, 5)
    for leaf in (1, 2, 16):
        rm2, _, _ = okrt.scenes.dragon_scene(20_000)
        _check_tree(okrt.Geometry(rm2.meshes, 30, leaf, device=0), rm2.meshes, leaf)


def test_device_builder_degenerate_inputs(okrt):
    """One triangle, all-identical triangles (every Morton code equal: position tie-break), and a flat mesh."""
    scenes = okrt.scenes
    one = scenes.make_mesh(*scenes.quad((0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 0, 1)), name="q")
    same_pos = np.repeat(one.trianglesPos[:1], 100, axis=0).copy()
    same_info = np.repeat(one.trianglesInfo[:1], 100, axis=0).copy()
    same = okrt.Mesh(same_pos, same_info, one.boundingBox, "same")
    for meshes in ([one], [same], [one, same, one]):
        _check_tree(okrt.Geometry(meshes, 30, 1, device=0), meshes, 1)
    geo = okrt.Geometry([], 30, 5, device=0)
    assert len(geo.bvhNodes) == 0 and len(geo.trianglesPos) == 0


def test_device_built_tree_renders_bit_exact_against_oracle(okrt, oracle):
    """The oracle (reference-order traversal of the arrays it is given) and both CUDA traversals agree bit for bit
    on the device-built arrays: the fast path's order-independence argument does not depend on who built the tree."""
    W, H, bounces = 256, 144, 3
    rm, sc, info = okrt.scenes.sponza_standin(30_000, texture_size=64)
    env = okrt.scenes.procedural_sky_faces(32)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, bvhDevice=0)
    rd1 = host["rd"].copy(); rd1[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=env), rd1, bounces, aov_dtype=okrt.abi.AOV)
    for traversal in (0, 1):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=bounces, traversal=traversal, writeAov=1)
        r.render(host["rd"].copy())
        g = r.aov()
        assert np.array_equal(g["triIdx"], o["aov"]["triIdx"])
        assert np.array_equal(g["distance"].view(np.uint32), o["aov"]["distance"].view(np.uint32))
        assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32)), traversal
        assert int(r.stats()["raySegments"]) == o["segments"]
        r.close()


def test_device_built_tree_against_reference_tree(okrt):
    """Same scene through the reference's SAH tree and through the device-built tree: primary-hit distances are
    identical wherever the two trees pick the same triangle, they differ only at exact ties, and the images agree."""
    W, H = 480, 270
    rm, sc, info = okrt.scenes.dragon_scene(120_000)
    env = okrt.scenes.procedural_sky_faces(64)
    imgs, dist, build = [], [], []
    for dev in (None, 0):
        host = helpers.build_host_side(okrt, rm, sc, info, W, H, bvhDevice=dev)
        build.append(host["geometry"].buildMs)
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=4, writeAov=1)
        r.render(host["rd"].copy(), frames=4)
        imgs.append(r.accumulation()[..., :3] / 4.0)
        dist.append(r.aov()["distance"].copy())
        r.close()
    same = dist[0].view(np.uint32) == dist[1].view(np.uint32)
    assert same.mean() > 0.9999, same.mean()
    mse = float(np.mean((np.clip(imgs[0], 0, 1) - np.clip(imgs[1], 0, 1)) ** 2))
    psnr = 10 * np.log10(1.0 / max(mse, 1e-20))
    assert psnr > 45.0, psnr
    assert build[1] is not None and build[1] < 2000.0
