"""The oracle against the reference's OWN shader text: resources/shaders/RaytracerCS.hlsl, GPU-Utilities.hlsli and
GPU-Structs.hlsli compiled as C++ where they lie (oracle/ref_shader_driver.cpp, oracle/ref_stubs/hlsl_shim.h ->
oracle/_ref/libokrt_ref_shader.so).  Every RGBA32F accumulation word and every RGBA8 result byte of oracle_render must equal
what the compiled shader writes: spheres, meshes behind child oct-tree nodes, rotated and translated instances, textures and
normal maps, the environment cube, the three light types, depth of field, refraction (16 bounces), the debug heat maps,
accumulation off, progressive frames.  This pins the restatement's control flow, operation order, RNG use and constants to the
reference; the conventions that remain this repository's (intrinsic math, texture filtering, mul's summation order) are listed
in hlsl_shim.h.  Also checks the committed fixtures generated from the compiled shader (tests/golden/ref_shader_*.npz), which
travel to machines without /root/reference."""
import os

import numpy as np
import pytest

from tests import helpers

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def ref_shader():
    from oracle import binding
    lib = binding.load_ref_shader()
    if lib is None:
        pytest.skip("oracle/_ref/libokrt_ref_shader.so is not built and /root/reference is not present")
    return lib


def _equal(o, r):
    bad = np.any(o["accum"].view(np.uint32) != r["accum"].view(np.uint32), axis=2)
    assert not bad.any(), "%d pixels differ from the compiled reference shader (first at %s)" % (int(bad.sum()), tuple(np.argwhere(bad)[0][::-1]))
    assert np.array_equal(o["result"], r["result"])


def _lights(abi):
    dl = np.zeros(1, abi.DIRECTIONAL_LIGHT)
    dl[0]["colour"] = (1, 0.9, 0.8); dl[0]["intensity"] = 2.0; dl[0]["effectiveAngle"] = np.cos(np.radians(20.0)); dl[0]["direction"] = (0, -1, 0)
    pl = np.zeros(2, abi.POINT_LIGHT)
    pl["colour"] = (0.3, 0.6, 1.0); pl["intensity"] = 3.0; pl["radius"] = 2.0; pl[0]["position"] = (5, 6, -5); pl[1]["position"] = (-7, 2, 3)
    sl = np.zeros(1, abi.SPOT_LIGHT)
    sl[0]["colour"] = (1, 1, 1); sl[0]["intensity"] = 4.0; sl[0]["radius"] = 1.5; sl[0]["maxAngle"] = np.cos(np.radians(60.0))
    sl[0]["position"] = (0, 10, -10); sl[0]["direction"] = (0, -0.7071, 0.7071)
    return dl, pl, sl


def scene_cases(okrt):
    """name -> (host arrays, environment faces, lights or None, bounces, W, H)"""
    abi = okrt.abi
    cases = {}
    rm, sc, info = okrt.scenes.test_scene(with_textures=True)
    cases["testscene"] = (helpers.build_host_side(okrt, rm, sc, info, 128, 72, spheres=True), okrt.scenes.procedural_sky_faces(32), None, 8, 128, 72)
    rm, sc, info = okrt.scenes.sponza_standin(6000, texture_size=32)
    cases["sponza"] = (helpers.build_host_side(okrt, rm, sc, info, 112, 63), okrt.scenes.procedural_sky_faces(16), None, 8, 112, 63)
    rm, sc, info = okrt.scenes.glass_scene(segments=24, rings=16)
    cases["glass"] = (helpers.build_host_side(okrt, rm, sc, info, 96, 54), okrt.scenes.procedural_sky_faces(32), None, 16, 96, 54)
    rm, sc, info = okrt.scenes.moved_props_scene()
    cases["moved_props"] = (helpers.build_host_side(okrt, rm, sc, info, 96, 54), okrt.scenes.procedural_sky_faces(16), None, 4, 96, 54)
    rm, sc = okrt.ResourceManager(), okrt.Scene()
    cube = rm.addMesh(okrt.scenes.make_mesh(*okrt.scenes.box((-1, -1, -1), (1, 1, 1))))
    ico = rm.addMesh(okrt.scenes.make_mesh(*okrt.scenes.icosphere(2, 1.5)))
    rng = np.random.default_rng(5)
    for k in range(40):
        m = okrt.scenes.material(albedo=tuple(rng.uniform(0.2, 0.9, 3)), roughness=float(rng.uniform(0, 1)), metallic=float(k % 2),
                                 transparency=float(k % 5 == 0), indexOfRefraction=1.4)
        sc.addMeshEntity(cube if k % 2 else ico, m, position=tuple(rng.uniform(-12, 12, 3)), rotation=tuple(rng.uniform(-180, 180, 3)))
    sc.setCamera((0, 0, -25), (0, 0, 0), 70.0, 0.1)
    cases["lights"] = (helpers.build_host_side(okrt, rm, sc, dict(bvh=(30, 5), octree=(2, 5)), 96, 54), None, _lights(abi), 5, 96, 54)
    return cases


def _scene_and_rd(oracle, case):
    host, env, lights, bounces, W, H = case
    rd = host["rd"].copy()
    kw = {}
    if lights is not None:
        dl, pl, sl = lights
        kw = dict(dirLights=dl, pointLights=pl, spotLights=sl)
        rd[0]["numDirLights"], rd[0]["numPointLights"], rd[0]["numSpotLights"] = len(dl), len(pl), len(sl)
    return helpers.oracle_scene(oracle, host, envFaces=env, **kw), rd, bounces


@pytest.mark.parametrize("name", ["testscene", "sponza", "glass", "moved_props", "lights"])
def test_oracle_equals_compiled_reference_shader(okrt, oracle, ref_shader, name):
    scene, rd, bounces = _scene_and_rd(oracle, scene_cases(okrt)[name])
    for frame in (1, 2, 7):
        rd[0]["numAccumulationFrames"] = frame
        o = oracle.render(scene, rd, bounces)
        assert o["overflow"] == 0, "the scene must fit the reference's fixed traversal stacks"
        _equal(o, ref_shader.render(scene, rd, bounces))
    rd[0]["numAccumulationFrames"] = 1
    _equal(oracle.render(scene, rd, 0), ref_shader.render(scene, rd, 0))        # primary rays only
    _equal(oracle.render(scene, rd, 1), ref_shader.render(scene, rd, 1))        # NUM_BOUNCES as the reference ships it


def test_progressive_frames_depth_of_field_debug_modes(okrt, oracle, ref_shader):
    scene, rd, bounces = _scene_and_rd(oracle, scene_cases(okrt)["testscene"])
    rd[0]["dofStrength"] = 0.35
    rd[0]["dofDistance"] = 9.0
    acc_o = acc_r = None
    for frame in (1, 2, 3):                                                     # u1 carried from frame to frame
        rd[0]["numAccumulationFrames"] = frame
        o = oracle.render(scene, rd, 4, accum=acc_o)
        r = ref_shader.render(scene, rd, 4, accum=acc_r)
        acc_o, acc_r = o["accum"], r["accum"]
        _equal(o, r)
    assert acc_o[..., 3].max() == 3.0
    for mode, limit in ((1, 200), (2, 40)):                                     # RaytracerCS.hlsl:516-520
        rd[0]["debugMode"], rd[0]["debugMaxCount"], rd[0]["numAccumulationFrames"] = mode, limit, 1
        _equal(oracle.render(scene, rd, 2), ref_shader.render(scene, rd, 2))
    rd[0]["debugMode"] = 0
    rd[0]["accumulationEnabled"] = 0                                            # result = saturate(light), u1 untouched
    o, r = oracle.render(scene, rd, 3), ref_shader.render(scene, rd, 3)
    _equal(o, r)
    assert not o["accum"].any()


@pytest.mark.parametrize("name", ["testscene", "sponza", "glass"])
def test_oracle_equals_committed_reference_shader_fixtures(okrt, oracle, name):
    """tests/golden/ref_shader_<name>.npz was written by tests/golden/make_golden_ref_shader.py from the compiled shader."""
    path = os.path.join(GOLDEN, "ref_shader_%s.npz" % name)
    assert os.path.exists(path), "run tests/golden/make_golden_ref_shader.py where /root/reference exists"
    z = np.load(path)
    scene, rd, bounces = _scene_and_rd(oracle, scene_cases(okrt)[name])
    assert int(z["bounces"]) == bounces
    for k, frame in enumerate(z["frames"]):
        rd[0]["numAccumulationFrames"] = int(frame)
        o = oracle.render(scene, rd, bounces)
        assert np.array_equal(o["accum"].view(np.uint32), z["accum"][k].view(np.uint32))
        assert np.array_equal(o["result"], z["result"][k])
