"""Edge cases of the path on the GPU, each bit for bit against the oracle: empty and sphere-only scenes, image sizes the
reference's Dispatch(W/16, H/9) does not cover [Q9], BVHs with huge leaves (depth-limited) and one-triangle leaves,
coincident / degenerate triangles (exact distance ties: the reference's visiting order decides), the camera inside a
mesh's boxes, one bounce and zero bounces."""
import numpy as np
import pytest

from tests import helpers
from tests.test_gpu_configs import _bit_exact_against_oracle

pytestmark = pytest.mark.gpu


def _sky(okrt):
    return okrt.scenes.procedural_sky_faces(32)


def test_empty_scene_is_sky_only(okrt, oracle):
    rm, sc = okrt.ResourceManager(), okrt.Scene()
    sc.setCamera((0.0, 1.0, -5.0), (5.0, 20.0, 0.0), 90.0, 0.1)
    info = dict(name="empty", octree=(2, 50), bvh=(30, 5))
    o = _bit_exact_against_oracle(okrt, oracle, rm, sc, info, 160, 90, 2, _sky(okrt))
    assert o["segments"] == 160 * 90 and int(o["aov"]["numSegments"].max()) == 1


def test_spheres_only(okrt, oracle):
    rm, sc, info = okrt.scenes.test_scene()
    rm2, sc2 = okrt.ResourceManager(), okrt.Scene()
    sc2.sphereEntities = sc.sphereEntities[:24]
    sc2.camera = sc.camera
    W, H, bounces = 160, 90, 3
    host = helpers.build_host_side(okrt, rm2, sc2, info, W, H, spheres=True)
    rd1 = host["rd"].copy(); rd1[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=_sky(okrt)), rd1, bounces, aov_dtype=okrt.abi.AOV)
    for pipeline in (0, 1):
        r = helpers.Renderer(okrt, host, W, H, envFaces=_sky(okrt))
        r.options(maxBounces=bounces, pipeline=pipeline)
        r.render(host["rd"].copy())
        assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32))
        assert int(r.stats()["raySegments"]) == o["segments"]
        r.close()
    assert (o["aov"]["hitKind"] == 1).any()


@pytest.mark.parametrize("size", [(333, 187), (16, 9), (31, 17), (15, 8)])
def test_sizes_the_dispatch_does_not_cover(okrt, oracle, size):
    """[Q9] Dispatch(W/16, H/9): the right and bottom remainders keep the clear colour unless coverAllPixels is set."""
    W, H = size
    rm, sc, info = okrt.scenes.test_scene(num_spheres=0)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    rd1 = host["rd"].copy(); rd1[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=_sky(okrt)), rd1, 2, aov_dtype=okrt.abi.AOV)
    r = helpers.Renderer(okrt, host, W, H, envFaces=_sky(okrt))
    r.options(maxBounces=2)
    r.render(host["rd"].copy())
    acc = r.accumulation()
    assert np.array_equal(acc.view(np.uint32), o["accum"].view(np.uint32))
    cw, ch = 16 * (W // 16), 9 * (H // 9)
    assert not acc[ch:].any() and not acc[:, cw:].any()
    assert int(r.stats()["raySegments"]) == o["segments"]
    r.close()
    r = helpers.Renderer(okrt, host, W, H, envFaces=_sky(okrt))
    r.options(maxBounces=2, coverAllPixels=1)
    r.render(host["rd"].copy())
    full = r.accumulation()
    assert np.array_equal(full[:ch, :cw].view(np.uint32), acc[:ch, :cw].view(np.uint32))
    assert (full[..., 3] == 1.0).all()                      # every pixel received one sample
    r.close()


@pytest.mark.parametrize("bvh", [(3, 5), (30, 1), (30, 64), (1, 1)])
def test_bvh_shapes(okrt, oracle, bvh):
    """maxDepth 3: leaves of hundreds of triangles (chained 8-triangle leaf records in the fast layout);
    maxLeaf 1: one triangle per leaf; maxLeaf 64: leaves far beyond the record size; depth 1: the root is the only node."""
    rm, sc, info = okrt.scenes.test_scene(num_spheres=0)
    info = dict(info, bvh=bvh)
    _bit_exact_against_oracle(okrt, oracle, rm, sc, info, 160, 90, 2, _sky(okrt))


def test_coincident_and_degenerate_triangles(okrt, oracle):
    """Two copies of the same quad in one mesh and in two entities (equal hit distances: the reference's visiting order
    picks the winner), zero-area triangles, and a needle: the fast path must hand every tie to the reference order."""
    scenes = okrt.scenes
    rm, sc = okrt.ResourceManager(), okrt.Scene()
    q = np.asarray(scenes.quad((-4, 0, -4), (4, 0, -4), (4, 0, 4), (-4, 0, 4))[0], np.float64).reshape(-1, 3, 3)
    wall = np.asarray(scenes.quad((-4, 0, 4), (4, 0, 4), (4, 6, 4), (-4, 6, 4))[0], np.float64).reshape(-1, 3, 3)
    zero = np.array([[[0, 1, 0], [0, 1, 0], [0, 1, 0]], [[1, 1, 1], [2, 2, 2], [3, 3, 3]]], np.float64)     # point and line
    needle = np.array([[[-3, 0.5, 0], [3, 0.5, 0], [3, 0.5 + 1e-6, 0]]], np.float64)
    doubled = rm.addMesh(scenes.make_mesh(np.concatenate([q, q, wall, wall[::-1], zero, needle]), name="doubled"))
    single = rm.addMesh(scenes.make_mesh(np.concatenate([q, wall]), name="single"))
    sc.addMeshEntity(doubled, scenes.material(albedo=(0.8, 0.3, 0.3)))
    sc.addMeshEntity(single, scenes.material(albedo=(0.3, 0.8, 0.3), metallic=1.0, roughness=0.1))     # same place, other material
    sc.addMeshEntity(single, scenes.material(albedo=(0.3, 0.3, 0.8)), position=(0.0, 0.0, 0.0), rotation=(0.0, 90.0, 0.0))
    sc.setCamera((0.0, 3.0, -7.0), (15.0, 0.0, 0.0), 90.0, 0.1)
    info = dict(name="ties", octree=(2, 50), bvh=(30, 5))
    o = _bit_exact_against_oracle(okrt, oracle, rm, sc, info, 160, 90, 3, _sky(okrt))
    assert (o["aov"]["hitKind"] == 2).mean() > 0.3


def test_camera_inside_the_geometry(okrt, oracle):
    rm, sc, info = okrt.scenes.sponza_standin(20_000, texture_size=64)
    sc.setCamera((0.0, 2.0, 0.0), (-35.0, 140.0, 0.0), 100.0, 0.1)
    _bit_exact_against_oracle(okrt, oracle, rm, sc, info, 160, 90, 4, _sky(okrt), pipelines=(0,))


@pytest.mark.parametrize("bounces", [0, 1])
def test_few_bounces(okrt, oracle, bounces):
    rm, sc, info = okrt.scenes.test_scene(num_spheres=8, with_textures=True)
    _bit_exact_against_oracle(okrt, oracle, rm, sc, info, 160, 90, bounces, _sky(okrt))
