"""Differential stress test of the near-first traversal: GPU fast path vs GPU reference order (`traversal=EXACT`, which
is itself compared with the oracle bit for bit in test_gpu_parity.py) on scenes built to expose order dependence —
flat leaf boxes with triangles in their faces, coplanar layers 1e-7..1e-4 apart, needle triangles, tiny triangles
against a 4000-unit quad, the same scenes translated by 1e3 and 1e4 (What's weak 3).  Together the cases
trace more than 1e8 ray segments; every RGBA32F accumulation word must be equal (RaytracerCS.hlsl:256-334: the
reference culls on `>= payload.distance` and updates on strict `<`)."""
import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu

CASES = [("flat_room", 0.0), ("flat_room", 1.0e3), ("flat_room", 1.0e4), ("layers", 0.0), ("layers", 1.0e3),
         ("slivers", 0.0), ("slivers", 1.0e3), ("contact", 0.0), ("contact", 1.0e4)]


@pytest.mark.parametrize("kind,offset", CASES, ids=["%s@%g" % c for c in CASES])
def test_fast_equals_reference_order(okrt, kind, offset):
    W = H = 768
    frames, bounces = 4, 6
    rm, sc, info = okrt.scenes.adversarial_scene(kind, offset=(offset, offset, offset))
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(32)
    out = {}
    for name, traversal in (("fast", okrt.abi.TRAVERSAL_FAST), ("wide", okrt.abi.TRAVERSAL_WIDE), ("exact", okrt.abi.TRAVERSAL_EXACT)):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=bounces, traversal=traversal, coverAllPixels=1)
        r.render(host["rd"].copy(), frames=frames)
        st = r.stats()
        out[name] = (r.accumulation(), int(st["raySegments"]), int(st["fallbackSegments"]))
        r.close()
    fast, exact = out["fast"], out["exact"]
    bad_wide = np.any(out["wide"][0].view(np.uint32) != exact[0].view(np.uint32), axis=2)
    assert not bad_wide.any() and out["wide"][1] == exact[1], "OKRT_TRAVERSAL_WIDE: %d pixels differ from the reference order" % int(bad_wide.sum())
    bad = np.any(fast[0].view(np.uint32) != exact[0].view(np.uint32), axis=2)
    print("%s@%g: %d segments, %d via the reference-order fallback (%.4f %%), %d differing pixels" %
          (kind, offset, fast[1], fast[2], 100.0 * fast[2] / max(1, fast[1]), int(bad.sum())))
    assert not bad.any(), "%d of %d pixels differ between the near-first and the reference-order traversal (first at %s)" % (
        int(bad.sum()), W * H, tuple(np.argwhere(bad)[0]))
    assert fast[1] == exact[1] and fast[1] >= frames * W * H
    # sheets 1e-7..1e-4 apart are what the reference-order fallback exists for (every ray through a stack is ambiguous)
    assert fast[2] < (0.25 if kind == "layers" else 0.05) * fast[1], "the fallback is meant for rare rays"


def test_bench_scene_many_frames_fast_equals_reference_order(okrt):
    """The bench workload itself (262k triangles, 354 instances, textures, 1920x1080, 8 bounces), six progressive frames: more
    than 5e7 ray segments whose origins sit 0.001 off a surface (every bounce), near-first and wide traversals against the
    reference order, every accumulation word.  (Frame 1 is also compared with the oracle, tests/test_gpu_modes.py.)"""
    import bench
    host, env, info = bench.build_scene(okrt)
    W, H = bench.WIDTH, bench.HEIGHT
    out = {}
    for name, traversal in (("fast", okrt.abi.TRAVERSAL_FAST), ("wide", okrt.abi.TRAVERSAL_WIDE), ("exact", okrt.abi.TRAVERSAL_EXACT)):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=bench.MAX_BOUNCES, traversal=traversal)
        r.render(host["rd"].copy(), frames=6)
        st = r.stats()
        out[name] = (r.accumulation(), int(st["raySegments"]), int(st["fallbackSegments"]))
        r.close()
    exact = out["exact"]
    assert exact[1] > 50_000_000
    for name in ("fast", "wide"):
        bad = np.any(out[name][0].view(np.uint32) != exact[0].view(np.uint32), axis=2)
        assert not bad.any() and out[name][1] == exact[1], "%s: %d pixels differ from the reference order" % (name, int(bad.sum()))
    assert out["fast"][2] < 0.001 * out["fast"][1]


@pytest.mark.parametrize("config", ["C2", "C3"])
def test_config_scenes_many_frames_fast_equals_reference_order(okrt, config):
    """BASELINE configs 2 (871k-triangle dragon-class mesh, 4 bounces) and 3 (the reference's ikea_glass.obj, 16 bounces: every
    refraction path) at 1920x1080, four progressive frames, near-first against the reference order."""
    import os
    if config == "C2":
        rm, sc, info = okrt.scenes.dragon_scene()
        bounces = 4
    else:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ikea_glass_mesh.npz")
        rm, sc, info = okrt.scenes.glass_scene_from_fixture(path) if os.path.exists(path) else okrt.scenes.glass_scene()
        bounces = 16
    W, H = 1920, 1080
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(64)
    out = {}
    for name, traversal in (("fast", okrt.abi.TRAVERSAL_FAST), ("exact", okrt.abi.TRAVERSAL_EXACT)):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=bounces, traversal=traversal)
        r.render(host["rd"].copy(), frames=4)
        out[name] = (r.accumulation(), int(r.stats()["raySegments"]))
        r.close()
    bad = np.any(out["fast"][0].view(np.uint32) != out["exact"][0].view(np.uint32), axis=2)
    assert not bad.any() and out["fast"][1] == out["exact"][1], "%d pixels differ from the reference order" % int(bad.sum())
    assert out["fast"][1] > 10_000_000
