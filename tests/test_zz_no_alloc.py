"""okrt_draw_overlay_static (include/okrt.h): the overlay rasteriser with its scratch in a file-scope device array instead of
cudaMalloc'd buffers.  No counterpart in the reference; checked against the
existing entry point, okrt_draw_overlay, bit for bit (the tolerance of every overlay test here), on two calls in a row — the
second with a different, smaller line list, so that whatever the first call left in the shared array would show.

Collected last on purpose (the file name)."""
import numpy as np
import pytest

from tests import helpers
from tests.test_cpu_overlay import product_overlay

pytestmark = pytest.mark.gpu


def _drawn(okrt, r, host, draw, handle):
    """Resets the progressive state, renders frame 1 again (deterministic) and draws the overlay with `draw`."""
    okrt.abi.check(r.L.okrt_reset_accumulation(r.h))
    r.render(host["rd"].copy())
    base = r.result()
    okrt.abi.check(draw(r.h, handle))
    return base, r.result()


def test_static_scratch_overlay_equals_allocating_overlay(okrt):
    W, H = 320, 180
    rm, sc, info = okrt.scenes.test_scene(with_textures=False)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    r = helpers.Renderer(okrt, host, W, H, envFaces=okrt.scenes.procedural_sky_faces(32))
    r.options(maxBounces=1)
    # first call: every BVH box (many lines, most of the image touched); second call: the oct-tree's few boxes
    cases = [(0, host["geometry"].bvhNodes, 3), (1, host["octree"].nodes, 3)]
    try:
        for kind, nodes, mode in cases:
            h, segs, keys, cols = product_overlay(okrt, kind, nodes, 0, mode, None, sc.cameraArray(), W, H)
            try:
                base_a, want = _drawn(okrt, r, host, r.L.okrt_draw_overlay, h)
                base_b, got = _drawn(okrt, r, host, r.L.okrt_draw_overlay_static, h)
            finally:
                r.L.okrt_overlay_free(h)
            assert np.array_equal(base_a, base_b)
            assert (want != base_a).any(), "the overlay must be visible from the test camera"
            assert np.array_equal(got, want)
        # an empty overlay (draw mode "none") is the warm-up call: accepted, draws nothing
        before = r.result()
        h, segs, keys, cols = product_overlay(okrt, 0, host["geometry"].bvhNodes, 0, 0, None, sc.cameraArray(), W, H)
        try:
            assert len(segs) == 0
            okrt.abi.check(r.L.okrt_draw_overlay_static(r.h, h))
        finally:
            r.L.okrt_overlay_free(h)
        assert np.array_equal(r.result(), before)
    finally:
        r.close()


def test_static_scratch_overlay_rejects_sizes_beyond_its_limits(okrt):
    """An image of more than OKRT_OVERLAY_STATIC_MAX_PIXELS pixels is refused with a message; nothing is drawn."""
    W, H = 2048, 1152                     # 2 359 296 pixels > 2^21
    rm, sc, info = okrt.scenes.test_scene(with_textures=False)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    r = helpers.Renderer(okrt, host, W, H, envFaces=okrt.scenes.procedural_sky_faces(32))
    r.options(maxBounces=0)
    try:
        r.render(host["rd"].copy())
        before = r.result()
        h, segs, keys, cols = product_overlay(okrt, 1, host["octree"].nodes, 0, 3, None, sc.cameraArray(), W, H)
        try:
            rc = r.L.okrt_draw_overlay_static(r.h, h)
            message = r.L.okrt_last_error()
            message = message.decode() if isinstance(message, bytes) else str(message)
        finally:
            r.L.okrt_overlay_free(h)
        assert rc == 1 and okrt.abi.ERROR_NAMES[rc] == "OKRT_ERR_INVALID_ARGUMENT" and "exceed the static scratch" in message
        assert np.array_equal(r.result(), before)
    finally:
        r.close()
