"""The CUDA kernels against the reference's OWN shader text compiled as C++ (oracle/_ref/libokrt_ref_shader.so: built where
/root/reference exists, shipped with the repository snapshot to the GPU box).  Same cases as tests/test_cpu_ref_shader.py, both
traversals and both pipelines, every RGBA32F accumulation word; plus one full-size frame of the bench workload."""
import numpy as np
import pytest

from tests import helpers
from tests.test_cpu_ref_shader import _scene_and_rd, scene_cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_shader():
    from oracle import binding
    lib = binding.load_ref_shader()
    if lib is None:
        pytest.skip("oracle/_ref/libokrt_ref_shader.so did not travel with this snapshot")
    return lib


@pytest.mark.parametrize("name", ["testscene", "sponza", "glass", "moved_props", "lights"])
def test_cuda_equals_compiled_reference_shader(okrt, oracle, ref_shader, name):
    case = scene_cases(okrt)[name]
    host, env, lights, bounces, W, H = case
    scene, rd, _ = _scene_and_rd(oracle, case)
    want = None
    for frame in (1, 2, 3):                                      # three progressive frames, u1 carried along
        rd[0]["numAccumulationFrames"] = frame
        want = ref_shader.render(scene, rd, bounces, accum=None if want is None else want["accum"])
    for traversal, pipeline in ((okrt.abi.TRAVERSAL_FAST, 0), (okrt.abi.TRAVERSAL_EXACT, 0), (okrt.abi.TRAVERSAL_FAST, 1)):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        if lights is not None:
            r.set_lights(*lights)
        r.options(maxBounces=bounces, traversal=traversal, pipeline=pipeline)
        rd0 = rd.copy()
        rd0[0]["numAccumulationFrames"] = 0
        r.render(rd0, frames=3)
        acc, res = r.accumulation(), r.result()
        r.close()
        bad = np.any(acc.view(np.uint32) != want["accum"].view(np.uint32), axis=2)
        assert not bad.any(), "traversal %d pipeline %d: %d pixels differ from the compiled reference shader" % (traversal, pipeline, int(bad.sum()))
        assert np.array_equal(res, want["result"])


def test_bench_frame_equals_compiled_reference_shader(okrt, oracle, ref_shader):
    """Frame 1 of the bench workload (1920x1080, 8 bounces, 262k triangles, 27 textures) by the reference's shader text on the
    host cores against the CUDA path."""
    import bench
    host, env, info = bench.build_scene(okrt)
    scene = helpers.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()
    rd[0]["numAccumulationFrames"] = 1
    want = ref_shader.render(scene, rd, bench.MAX_BOUNCES)
    r = helpers.Renderer(okrt, host, bench.WIDTH, bench.HEIGHT, envFaces=env)
    r.options(maxBounces=bench.MAX_BOUNCES)
    r.render(host["rd"].copy())
    acc, res = r.accumulation(), r.result()
    r.close()
    bad = np.any(acc.view(np.uint32) != want["accum"].view(np.uint32), axis=2)
    assert not bad.any(), "%d of %d pixels differ from the compiled reference shader" % (int(bad.sum()), bad.size)
    assert np.array_equal(res, want["result"])
