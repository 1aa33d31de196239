"""Row f4, host part: okrt_overlay_build (the line list of DebugRenderer's box overlays) against the numpy restatement
oracle/overlay_oracle.py: draw order and colours exactly, screen coordinates within 2e-3 pixel (float32 camera basis in
the product, float64 in the oracle)."""
import ctypes
import importlib.util
import os

import numpy as np
import pytest

_spec = importlib.util.spec_from_file_location("overlay_oracle", os.path.join(os.path.dirname(__file__), "..", "oracle", "overlay_oracle.py"))
ov = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(ov)


def product_overlay(okrt, kind, nodes, node_idx, mode, transform, cam, W, H):
    abi = okrt.abi
    L = abi.lib()
    h = ctypes.c_void_p()
    abi.check(L.okrt_overlay_build(kind, abi.ptr(nodes), len(nodes), node_idx, mode, abi.ptr(transform) if transform is not None else None,
                                   abi.ptr(cam), W, H, ctypes.byref(h)))
    n, d = L.okrt_overlay_num_segments(h), L.okrt_overlay_num_draws(h)
    segs = abi.as_array(L.okrt_overlay_segments(h), np.dtype("<f4"), n * 4).reshape(n, 4).copy() if n else np.zeros((0, 4), np.float32)
    keys = abi.as_array(L.okrt_overlay_keys(h), np.dtype("<u4"), n).copy() if n else np.zeros(0, np.uint32)
    cols = abi.as_array(L.okrt_overlay_colours(h), np.dtype("<u4"), d).copy() if d else np.zeros(0, np.uint32)
    return h, segs, keys, cols


def _scene(okrt):
    from tests import helpers
    rm, sc, info = okrt.scenes.test_scene(with_textures=False)
    host = helpers.build_host_side(okrt, rm, sc, info, 320, 180)
    return rm, sc, host


@pytest.mark.parametrize("mode", [1, 2, 3])
@pytest.mark.parametrize("kind", [0, 1])
def test_overlay_segments_match_oracle(okrt, kind, mode):
    W, H = 320, 180
    rm, sc, host = _scene(okrt)
    nodes = host["geometry"].bvhNodes if kind == 0 else host["octree"].nodes
    cam = sc.cameraArray()
    transform, matrix = None, None
    if kind == 0:
        transform = np.zeros(1, okrt.abi.TRANSFORM)
        transform[0]["position"], transform[0]["rotation"], transform[0]["scale"] = (0.5, 0.25, -0.5), (10, 35, 5), (1.5, 1, 0.75)
        matrix = np.zeros(16, np.float32)
        okrt.abi.lib().okrt_transform_matrix(okrt.abi.ptr(transform), okrt.abi.ptr(matrix))
    h, segs, keys, cols = product_overlay(okrt, kind, nodes, 0, mode, transform, cam, W, H)
    try:
        osegs, okeys, ocols = ov.segments(kind, nodes, 0, mode, matrix, cam[0], W, H)
        assert okrt.abi.lib().okrt_overlay_num_draws(h) == len(ov.draw_order(kind, nodes, 0, mode))
        assert np.array_equal(keys, okeys) and np.array_equal(cols, ocols)
        assert segs.shape == osegs.shape and len(segs) > 0
        scale = np.maximum(1.0, np.abs(osegs))
        assert np.max(np.abs(segs - osegs) / scale) < 2e-3
    finally:
        okrt.abi.lib().okrt_overlay_free(h)


def test_overlay_draw_modes_and_clipping(okrt):
    W, H = 64, 64
    abi = okrt.abi
    nodes = np.zeros(3, abi.BVH_NODE)
    nodes[0]["boundingBox"]["min"], nodes[0]["boundingBox"]["max"] = (-1, -1, 4), (1, 1, 6)
    nodes[0]["firstChildIdx"] = 1
    nodes[1]["boundingBox"]["min"], nodes[1]["boundingBox"]["max"] = (-1, -1, 4), (0, 1, 6)
    nodes[2]["boundingBox"]["min"], nodes[2]["boundingBox"]["max"] = (0, -1, -8), (1, 1, 6)          # reaches behind the camera
    nodes[1]["firstChildIdx"] = nodes[2]["firstChildIdx"] = 0xFFFFFFFF
    cam = np.zeros(1, abi.CAMERA)
    cam[0]["fov"], cam[0]["nearZ"], cam[0]["farZ"] = 90.0, 0.1, 100.0
    for mode, draws in ((0, 0), (1, 1), (2, 3), (3, 3)):
        h, segs, keys, cols = product_overlay(okrt, 0, nodes, 0, mode, None, cam, W, H)
        assert abi.lib().okrt_overlay_num_draws(h) == draws
        assert np.isfinite(segs).all()
        if mode == 3:
            assert list(dict.fromkeys(keys.tolist())) == [1, 2, 3]
            # the stack pops node 2 first, so draw 2 is node 2: its 4 edges at z = -8 are dropped, the 4 that cross the
            # near plane are clipped to it, the 4 at z = 6 stay
            assert (keys == 2).sum() == 8 and (keys == 3).sum() == 12
        abi.lib().okrt_overlay_free(h)
    # a box entirely behind the camera draws nothing; a missing node index draws nothing
    nodes[0]["boundingBox"]["min"], nodes[0]["boundingBox"]["max"] = (-1, -1, -6), (1, 1, -4)
    h, segs, keys, cols = product_overlay(okrt, 0, nodes, 0, 1, None, cam, W, H)
    assert len(segs) == 0 and abi.lib().okrt_overlay_num_draws(h) == 1
    abi.lib().okrt_overlay_free(h)
    h, segs, keys, cols = product_overlay(okrt, 0, nodes, 7, 3, None, cam, W, H)
    assert len(segs) == 0
    abi.lib().okrt_overlay_free(h)
    assert abi.lib().okrt_overlay_build(2, abi.ptr(nodes), 3, 0, 1, None, abi.ptr(cam), W, H, ctypes.byref(ctypes.c_void_p())) == 1


def test_overlay_rasteriser_rule_small_cases():
    img = np.zeros((8, 8, 4), np.uint8)
    colours = np.array([0xFF0000FF, 0xFF00FF00], np.uint32)
    segs = np.array([[0.0, 2.5, 8.0, 2.5], [3.5, 0.0, 3.5, 8.0]], np.float32)          # horizontal then vertical
    ov.rasterize(segs, np.array([1, 2], np.uint32), colours, img)
    assert (img[2, :, 0] == 255).sum() == 7 and tuple(img[2, 3]) == (0, 255, 0, 255)  # the later draw wins the crossing
    assert (img[:, 3, 1] == 255).all()
    img2 = np.zeros((8, 8, 4), np.uint8)
    ov.rasterize(np.array([[2.6, 1.0, 2.9, 1.2]], np.float32), np.array([1], np.uint32), colours, img2)
    assert not img2.any()                                                             # no pixel centre inside [start, end)
