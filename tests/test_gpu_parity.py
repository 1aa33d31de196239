"""GPU parity tests proper: the CUDA path, driven through the C ABI, against the CPU oracle on the same
seeded inputs.  Bar (BASELINE.json north_star): primary-hit triangle IDs bit-exact, barycentrics and depth
within 1e-5 relative (here: bit-exact), images within a stated RMSE — here the RGBA32F accumulation is
bit-exact because both sides evaluate the same float32 expression trees (canonical math, no FMA)."""
import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=[0, 1], ids=["wavefront", "megakernel"])
def pipeline(request):
    """Every parity test runs through both kernel organisations; they must agree with the oracle bit for bit."""
    old = helpers.harness.DEFAULT_PIPELINE
    helpers.harness.DEFAULT_PIPELINE = request.param
    yield request.param
    helpers.harness.DEFAULT_PIPELINE = old


def _require_gpu(okrt):
    if okrt.abi.lib().okrt_device_count() <= 0:
        pytest.fail("no CUDA device visible: the product path has no CPU fallback")


def _compare_aov(g, o, exact_counters):
    assert np.array_equal(g["hitKind"], o["hitKind"])
    assert np.array_equal(g["hitIdx"], o["hitIdx"])
    assert np.array_equal(g["triIdx"], o["triIdx"]), "primary-hit triangle ids must be bit-exact"
    for f in ("baryU", "baryV", "distance"):
        assert np.array_equal(g[f].view(np.uint32), o[f].view(np.uint32)), f
    assert np.array_equal(g["numSegments"], o["numSegments"])
    assert np.array_equal(g["radiance"].view(np.uint32), o["radiance"].view(np.uint32))
    if exact_counters:
        assert np.array_equal(g["bbCheckCount"], o["bbCheckCount"])
        assert np.array_equal(g["triCheckCount"], o["triCheckCount"])


@pytest.mark.parametrize("traversal", [0, 1])
def test_config1_testscene_primary(okrt, oracle, traversal):
    """BASELINE config 1: TestScene 512x512, 1 spp, primary rays only (KAT-4, KAT-5)."""
    _require_gpu(okrt)
    abi = okrt.abi
    rm, sc, info = okrt.scenes.test_scene()
    W = H = 512
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    env = okrt.scenes.procedural_sky_faces(64)
    r = helpers.Renderer(okrt, host, W, H, envFaces=env)
    r.options(maxBounces=0, traversal=traversal, writeAov=1)
    rd = host["rd"].copy()
    r.render(rd)
    g = r.aov()
    rd[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=env), rd, 0, aov_dtype=abi.AOV)
    assert o["overflow"] == 0
    _compare_aov(g, o["aov"], exact_counters=(traversal == 1))
    res = r.result()
    assert np.array_equal(res, o["result"])
    # [Q9] Dispatch(512/16, 512/9): rows 504..511 keep the clear colour
    assert np.all(res[504:] == np.array([51, 102, 153, 255], np.uint8))
    assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32))
    st = r.stats()
    assert int(st["raySegments"]) == o["segments"] == 512 * 504
    r.close()


@pytest.mark.parametrize("traversal", [0, 1])
def test_testscene_full_paths_bit_exact(okrt, oracle, traversal):
    """Spheres + textured meshes + env map + DOF, 6 bounces, 3 accumulated frames: u1 must match bit for bit."""
    _require_gpu(okrt)
    abi = okrt.abi
    rm, sc, info = okrt.scenes.test_scene(with_textures=True)
    W, H = 256, 144
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    env = okrt.scenes.procedural_sky_faces(32)
    r = helpers.Renderer(okrt, host, W, H, envFaces=env)
    r.options(maxBounces=6, traversal=traversal, writeAov=1)
    rd = host["rd"].copy()
    rd[0]["dofStrength"] = 0.05
    rd[0]["dofDistance"] = 20.0
    r.render(rd, frames=3)
    os_ = helpers.oracle_scene(oracle, host, envFaces=env)
    acc = np.zeros((H, W, 4), np.float32)
    res = np.zeros((H, W, 4), np.uint8)
    for f in (1, 2, 3):
        rd[0]["numAccumulationFrames"] = f
        o = oracle.render(os_, rd, 6, accum=acc, result=res, aov_dtype=abi.AOV)
    _compare_aov(r.aov(), o["aov"], exact_counters=(traversal == 1))
    assert np.array_equal(r.accumulation().view(np.uint32), acc.view(np.uint32))
    assert np.array_equal(r.result(), res)
    r.close()


def test_sponza_standin_small_bit_exact(okrt, oracle):
    """A 60k-triangle cut of the Sponza stand-in (354 instances, 27 texture layers), 8 bounces."""
    _require_gpu(okrt)
    abi = okrt.abi
    rm, sc, info = okrt.scenes.sponza_standin(60000, texture_size=64)
    W, H = 320, 180
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(32)
    os_ = helpers.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()
    rd[0]["numAccumulationFrames"] = 1
    o = oracle.render(os_, rd, 8, aov_dtype=abi.AOV)
    for traversal in (0, 1, 2):
        if traversal == 2 and helpers.harness.DEFAULT_PIPELINE != 0:
            continue                          # OKRT_TRAVERSAL_WIDE: wavefront pipeline only
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=8, traversal=traversal, writeAov=1)
        r.render(host["rd"].copy())
        g = r.aov()
        _compare_aov(g, o["aov"], exact_counters=(traversal == 1))
        assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32))
        st = r.stats()
        assert int(st["raySegments"]) == o["segments"]
        if traversal != 1:
            # the reference-order fallback must stay a rare path
            assert int(st["fallbackSegments"]) < 0.02 * int(st["raySegments"])
        r.close()


def test_lights_and_transforms(okrt, oracle):
    """Point/spot/directional lights [Q11] and rotated+translated instances (rigid: stays on the fast path)."""
    _require_gpu(okrt)
    abi = okrt.abi
    rm, sc = okrt.ResourceManager(), okrt.Scene()
    cube = rm.addMesh(okrt.scenes.make_mesh(*okrt.scenes.box((-1, -1, -1), (1, 1, 1))))
    ico = rm.addMesh(okrt.scenes.make_mesh(*okrt.scenes.icosphere(2, 1.5)))
    rng = np.random.default_rng(5)
    for k in range(40):
        m = okrt.scenes.material(albedo=tuple(rng.uniform(0.2, 0.9, 3)), roughness=float(rng.uniform(0, 1)), metallic=float(k % 2),
                                 transparency=float(k % 5 == 0), indexOfRefraction=1.4)
        sc.addMeshEntity(cube if k % 2 else ico, m, position=tuple(rng.uniform(-12, 12, 3)), rotation=tuple(rng.uniform(-180, 180, 3)))
    sc.setCamera((0, 0, -25), (0, 0, 0), 70.0, 0.1)
    info = dict(bvh=(30, 5), octree=(2, 5))
    W, H = 192, 108
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    dl = np.zeros(1, abi.DIRECTIONAL_LIGHT)
    dl[0]["colour"] = (1, 0.9, 0.8); dl[0]["intensity"] = 2.0; dl[0]["effectiveAngle"] = np.cos(np.radians(20.0)); dl[0]["direction"] = (0, -1, 0)
    pl = np.zeros(2, abi.POINT_LIGHT)
    pl["colour"] = (0.3, 0.6, 1.0); pl["intensity"] = 3.0; pl["radius"] = 2.0; pl[0]["position"] = (5, 6, -5); pl[1]["position"] = (-7, 2, 3)
    sl = np.zeros(1, abi.SPOT_LIGHT)
    sl[0]["colour"] = (1, 1, 1); sl[0]["intensity"] = 4.0; sl[0]["radius"] = 1.5; sl[0]["maxAngle"] = np.cos(np.radians(60.0))
    sl[0]["position"] = (0, 10, -10); sl[0]["direction"] = (0, -0.7071, 0.7071)
    rd = host["rd"].copy()
    rd[0]["numDirLights"], rd[0]["numPointLights"], rd[0]["numSpotLights"] = 1, 2, 1
    os_ = helpers.oracle_scene(oracle, host, dirLights=dl, pointLights=pl, spotLights=sl)
    rdo = rd.copy(); rdo[0]["numAccumulationFrames"] = 1
    o = oracle.render(os_, rdo, 5, aov_dtype=abi.AOV)
    for traversal in (0, 1):
        r = helpers.Renderer(okrt, host, W, H)
        r.set_lights(dl, pl, sl)
        r.options(maxBounces=5, traversal=traversal, writeAov=1)
        r.render(rd.copy())
        _compare_aov(r.aov(), o["aov"], exact_counters=(traversal == 1))
        assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32))
        r.close()


def test_moved_props_under_child_oct_nodes(okrt, oracle):
    """oct-tree nodes that do not contain the (translated) instances they list — the reference culls such a node,
    and everything below it, against a FARTHER hit (RaytracerCS.hlsl:262-264).  The near-first path has to notice (the node's
    distNear joins the winner's gate) and let the reference order decide those rays."""
    _require_gpu(okrt)
    abi = okrt.abi
    rm, sc, info = okrt.scenes.moved_props_scene()
    W, H = 256, 144
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    assert len(host["octree"].nodes) > 1
    env = okrt.scenes.procedural_sky_faces(16)
    rdo = host["rd"].copy(); rdo[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=env), rdo, 4, aov_dtype=abi.AOV)
    for traversal in (0, 1, 2):
        if traversal == 2 and helpers.harness.DEFAULT_PIPELINE != 0:
            continue                          # OKRT_TRAVERSAL_WIDE: wavefront pipeline only
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=4, traversal=traversal, writeAov=1)
        r.render(host["rd"].copy())
        _compare_aov(r.aov(), o["aov"], exact_counters=(traversal == 1))
        assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32))
        r.close()


def test_scaled_instances_use_reference_order(okrt, oracle):
    """Non-unit scale makes the reference's local-vs-world culling order dependent [Q4]: the library must
    switch to the reference visiting order by itself and still match."""
    _require_gpu(okrt)
    abi = okrt.abi
    rm, sc = okrt.ResourceManager(), okrt.Scene()
    ico = rm.addMesh(okrt.scenes.make_mesh(*okrt.scenes.icosphere(3, 1.0)))
    rng = np.random.default_rng(9)
    for k in range(12):
        sc.addMeshEntity(ico, okrt.scenes.material(albedo=(0.8, 0.5, 0.3)), position=tuple(rng.uniform(-8, 8, 3)),
                         rotation=tuple(rng.uniform(-90, 90, 3)), scale=tuple(rng.uniform(0.5, 3.0, 3)))
    sc.setCamera((0, 0, -20), (0, 0, 0), 80.0, 0.1)
    W, H = 160, 90
    host = helpers.build_host_side(okrt, rm, sc, dict(bvh=(30, 5), octree=(2, 4)), W, H)
    rdo = host["rd"].copy(); rdo[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host), rdo, 3, aov_dtype=abi.AOV)
    r = helpers.Renderer(okrt, host, W, H)
    r.options(maxBounces=3, traversal=0, writeAov=1)
    r.render(host["rd"].copy())
    _compare_aov(r.aov(), o["aov"], exact_counters=True)
    r.close()


def test_debug_heatmaps_and_no_accumulation(okrt, oracle):
    """debugMode 1/2 (RaytracerCS.hlsl:516-520) display the reference's own counters; accumulation off [Q14]."""
    _require_gpu(okrt)
    rm, sc, info = okrt.scenes.test_scene()
    W, H = 160, 90
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    os_ = helpers.oracle_scene(oracle, host)
    r = helpers.Renderer(okrt, host, W, H)
    for mode in (1, 2):
        rd = host["rd"].copy()
        rd[0]["debugMode"] = mode
        rd[0]["debugMaxCount"] = 60
        rd[0]["accumulationEnabled"] = 0
        r.options(maxBounces=2, traversal=0)
        r.render(rd)
        o = oracle.render(os_, rd, 2, seedFrame=0)
        assert np.array_equal(r.result(), o["result"])
    r.close()
