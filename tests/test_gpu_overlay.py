"""Row f4 on the GPU: okrt_draw_overlay rasterises the line list into the RGBA8 result; compared bit for bit with the
numpy restatement of the same pixel rule (oracle/overlay_oracle.py) on the product's own float32 segments."""
import numpy as np
import pytest

from tests import helpers
from tests.test_cpu_overlay import ov, product_overlay

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,mode", [(0, 3), (1, 3), (0, 2)])
def test_overlay_pixels_bit_exact(okrt, kind, mode):
    W, H = 320, 180
    rm, sc, info = okrt.scenes.test_scene(with_textures=False)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    r = helpers.Renderer(okrt, host, W, H, envFaces=okrt.scenes.procedural_sky_faces(32))
    r.options(maxBounces=2)
    r.render(host["rd"].copy())
    before = r.result()
    nodes = host["geometry"].bvhNodes if kind == 0 else host["octree"].nodes
    h, segs, keys, cols = product_overlay(okrt, kind, nodes, 0, mode, None, sc.cameraArray(), W, H)
    try:
        okrt.abi.check(r.L.okrt_draw_overlay(r.h, h))
    finally:
        r.L.okrt_overlay_free(h)
    after = r.result()
    want = ov.rasterize(segs, keys, cols, before.copy())
    assert np.array_equal(after, want)
    assert (after != before).any(), "the overlay must be visible from the test camera"
    acc = r.accumulation()
    r.render(host["rd"].copy())                                  # the next frame rewrites u0; u1 never saw the overlay
    assert int(r.stats()["raySegments"]) > 0 and np.isfinite(acc).all()
    r.close()


def test_overlay_through_the_mirror_class(okrt):
    rm, sc, info = okrt.scenes.test_scene(with_textures=False)
    W, H = 192, 108
    rt = okrt.RayTracer(0)
    rt.initiate(W, H, rm, environmentMapFaces=okrt.scenes.procedural_sky_faces(32), maxBounces=1)
    rt.setScene(sc)
    rt.createOctTree(sc, 2, 50)
    rt.render()
    rt.sync()
    base = rt.readResult()
    rt.renderBvhNodeBBs(0, 0, rt.DrawMode_DrawWithChildren)
    with_bvh = rt.readResult()
    rt.renderOctTreeNodeBBs(0, rt.DrawMode_DrawWithDecendants)
    with_both = rt.readResult()
    changed = (with_bvh != base).any(axis=-1)
    assert changed.any()
    # BVH boxes are drawn in BVH_NODE_COLOUR x (1 - nodeIdx / numNodes), oct-tree boxes in OCT_NODE_COLOUR (DebugRenderer.cpp:10-12)
    oct_px = (with_both != with_bvh).any(axis=-1)
    assert oct_px.any() and (with_both[oct_px] == np.array([102, 179, 102, 255], np.uint8)).all()
    rt.shutdown()


def test_overlay_many_boxes_full_hd(okrt):
    """Every node box of the 60k-triangle dragon-class mesh (tens of thousands of boxes, ~half a million lines) at 1080p."""
    W, H = 1920, 1080
    rm, sc, info = okrt.scenes.dragon_scene(60_000)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    r = helpers.Renderer(okrt, host, W, H, envFaces=okrt.scenes.procedural_sky_faces(32))
    r.options(maxBounces=0)
    r.render(host["rd"].copy())
    before = r.result()
    nodes = host["geometry"].bvhNodes
    h, segs, keys, cols = product_overlay(okrt, 0, nodes, 0, 3, None, sc.cameraArray(), W, H)
    try:
        okrt.abi.check(r.L.okrt_draw_overlay(r.h, h))
    finally:
        r.L.okrt_overlay_free(h)
    after = r.result()
    assert len(segs) > 100_000
    sample = np.random.default_rng(0).choice(len(segs), 4000, replace=False)
    sample.sort()
    # a pixel touched by a sampled segment holds the colour of some draw whose index is >= that segment's
    probe = ov.rasterize(segs[sample], keys[sample], cols, np.zeros_like(before))
    touched = probe.any(axis=-1)
    assert touched.sum() > 1000 and (after[touched] != before[touched]).any(axis=-1).mean() > 0.5
    r.close()
