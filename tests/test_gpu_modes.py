"""GPU tests of the modes around the hot path: sharding, progressive state, the RayTracer mirror, full-size runs."""
import ctypes

import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def _scene(okrt, W, H, tris=20000):
    rm, sc, info = okrt.scenes.sponza_standin(tris, texture_size=64)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(32)
    return host, env


@pytest.mark.parametrize("pipeline", [0, 1])
def test_tile_sharding_union_equals_single(okrt, pipeline):
    """Tile mode: ranks render disjoint interleaved bands; the union is the single-GPU image, bit for bit."""
    W, H = 320, 184
    host, env = _scene(okrt, W, H)
    single = helpers.Renderer(okrt, host, W, H, envFaces=env)
    single.options(maxBounces=4, coverAllPixels=1, pipeline=pipeline)
    single.render(host["rd"].copy(), frames=2)
    ref_acc = single.accumulation()
    ref_img = single.result()
    total = int(single.stats()["raySegments"])
    for world, band in ((2, 8), (3, 16), (8, 8)):
        acc = np.zeros_like(ref_acc)
        img = np.zeros_like(ref_img)          # one host image every rank drops its own bands into (okrt_read_result_owned)
        segs = moved_total = 0
        for rank in range(world):
            r = helpers.Renderer(okrt, host, W, H, envFaces=env)
            r.options(maxBounces=4, coverAllPixels=1, pipeline=pipeline, **okrt.sharding.tile_options(rank, world, band))
            r.render(host["rd"].copy(), frames=2)
            a = r.accumulation()
            rows = okrt.sharding.rows_of_rank(rank, world, H, band)
            other = np.setdiff1d(np.arange(H), rows)
            assert not a[other].any(), "a rank wrote outside its bands"
            acc[rows] = a[rows]
            moved = ctypes.c_uint64(0)
            before = img.copy()
            okrt.abi.check(r.L.okrt_read_result_owned(r.h, okrt.abi.ptr(img), ctypes.byref(moved)))
            assert moved.value == len(rows) * W * 4 and np.array_equal(img[other], before[other])
            moved_total += moved.value
            segs += int(r.stats()["raySegments"])
            r.close()
        assert np.array_equal(acc.view(np.uint32), ref_acc.view(np.uint32))
        assert np.array_equal(img, ref_img) and moved_total == W * H * 4
        assert segs == total
    single.close()


def test_cost_balanced_band_table_union_equals_single(okrt):
    """Tile mode with an explicit band table (okrt_set_shard_bands), bands assigned by okrt_shard_balance from the ray
    segments of a calibration frame: the ranks' rows still tile the image exactly, images equal the single-GPU one."""
    W, H, world = 320, 184, 3
    host, env = _scene(okrt, W, H)
    single = helpers.Renderer(okrt, host, W, H, envFaces=env)
    single.options(maxBounces=4, coverAllPixels=1, writeAov=1)
    single.render(host["rd"].copy(), frames=1)
    cost = okrt.sharding.band_costs_from_segments(single.aov()["numSegments"], 8)
    single.close()
    owner = okrt.sharding.balanced_bands(cost, world)
    assert set(owner.tolist()) == set(range(world)) and len(owner) == (H + 7) // 8
    loads = [cost[owner == k].sum() for k in range(world)]
    assert max(loads) / min(loads) < 1.1
    single = helpers.Renderer(okrt, host, W, H, envFaces=env)
    single.options(maxBounces=4, coverAllPixels=1)
    single.render(host["rd"].copy(), frames=2)
    ref_acc, ref_img = single.accumulation(), single.result()
    single.close()
    acc, img, moved_total = np.zeros_like(ref_acc), np.zeros_like(ref_img), 0
    for rank in range(world):
        bands = np.ascontiguousarray(np.nonzero(owner == rank)[0], np.uint32)
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=4, coverAllPixels=1, **okrt.sharding.tile_options(rank, world, 8))
        okrt.abi.check(r.L.okrt_set_shard_bands(r.h, okrt.abi.ptr(bands), len(bands)))
        r.render(host["rd"].copy(), frames=2)
        rows = np.concatenate([np.arange(b * 8, min((b + 1) * 8, H)) for b in bands])
        a = r.accumulation()
        assert not a[np.setdiff1d(np.arange(H), rows)].any()
        acc[rows] = a[rows]
        moved = ctypes.c_uint64(0)
        okrt.abi.check(r.L.okrt_read_result_async(r.h, okrt.abi.ptr(img), ctypes.byref(moved), 0))
        okrt.abi.check(r.L.okrt_wait_read(r.h, 0))
        assert moved.value == len(rows) * W * 4
        moved_total += moved.value
        r.close()
    assert np.array_equal(acc.view(np.uint32), ref_acc.view(np.uint32)) and np.array_equal(img, ref_img) and moved_total == W * H * 4


def test_sample_sharding_matches_sequential(okrt):
    """Sample mode: rank r renders frames r+1, r+1+G, ...; summed u1 == sequential u1 up to float summation order."""
    W, H, total, world = 256, 144, 8, 4
    host, env = _scene(okrt, W, H)
    seq = helpers.Renderer(okrt, host, W, H, envFaces=env)
    seq.options(maxBounces=4)
    seq.render(host["rd"].copy(), frames=total)
    want = seq.accumulation()
    summed = np.zeros_like(want, dtype=np.float64)
    for rank in range(world):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=4, **okrt.sharding.sample_options(rank, world))
        r.render(host["rd"].copy(), frames=total // world)
        summed += r.accumulation()
        r.close()
    rmse = float(np.sqrt(np.mean((summed / total - want.astype(np.float64) / total) ** 2)))
    assert rmse <= 1e-4, rmse
    assert np.array_equal(summed[..., 3], want[..., 3])
    # merge through the ABI: write the reduced buffer back and resolve (what the all-reduce path does)
    seq.abi.check(seq.L.okrt_write_accumulation(seq.h, seq.abi.ptr(summed.astype(np.float32)), total))
    seq.abi.check(seq.L.okrt_resolve(seq.h, total))
    img = seq.result()
    exp = np.clip(summed[..., :3] / total, 0, 1) * 255.0 + 0.5
    covered = img[:9 * (H // 9), :16 * (W // 16), :3].astype(np.int32)
    assert np.abs(covered - exp[:9 * (H // 9), :16 * (W // 16)].astype(np.int32)).max() <= 1
    seq.close()


def test_progressive_state_reset_resize(okrt, oracle):
    """toggle/resetAccumulation (RayTracer.h:197-209), onResize (RayTracer.cpp:494-501), determinism of re-rendering."""
    W, H = 160, 90
    host, env = _scene(okrt, W, H, tris=8000)
    r = helpers.Renderer(okrt, host, W, H, envFaces=env)
    r.options(maxBounces=3)
    rd = host["rd"].copy()
    r.render(rd, frames=3)
    a3 = r.accumulation()
    assert r.L.okrt_get_num_accumulation_frames(r.h) == 3
    r.abi.check(r.L.okrt_reset_accumulation(r.h))
    assert r.L.okrt_get_num_accumulation_frames(r.h) == 0 and not r.accumulation().any()
    r.abi.check(r.L.okrt_render_frames(r.h, 3)); r.abi.check(r.L.okrt_sync(r.h))
    assert np.array_equal(r.accumulation().view(np.uint32), a3.view(np.uint32))          # same seeds -> same bits
    # accumulation off: the counter stays 0 and every frame is the same image [Q3]
    r.abi.check(r.L.okrt_toggle_accumulation(r.h, 0))
    r.abi.check(r.L.okrt_render_frames(r.h, 1)); r.abi.check(r.L.okrt_sync(r.h)); i1 = r.result()
    r.abi.check(r.L.okrt_render_frames(r.h, 1)); r.abi.check(r.L.okrt_sync(r.h)); i2 = r.result()
    assert np.array_equal(i1, i2) and r.L.okrt_get_num_accumulation_frames(r.h) == 0
    # resize: new images, counter reset, clear colour in the uncovered border [Q9]
    r.abi.check(r.L.okrt_toggle_accumulation(r.h, 1))
    r.abi.check(r.L.okrt_resize(r.h, 100, 50))
    r.W, r.H = 100, 50
    host2 = helpers.build_host_side(okrt, *okrt.scenes.sponza_standin(8000, texture_size=64), 100, 50)
    r.render(host2["rd"].copy())
    img = r.result()
    assert np.all(img[45:, :] == np.array([51, 102, 153, 255], np.uint8)) and np.all(img[:, 96:] == np.array([51, 102, 153, 255], np.uint8))
    o = oracle.render(helpers.oracle_scene(oracle, host2, envFaces=env), _frame1(host2["rd"]), 3)
    assert np.array_equal(img, o["result"])
    r.close()


def _frame1(rd):
    rd = rd.copy(); rd[0]["numAccumulationFrames"] = 1
    return rd


def test_screenshot_and_resume_from_accumulation_file(okrt, tmp_path):
    """Row f3: saveScreenshot writes the RGBA8 target as a PNG (Application.cpp:661-674); the accumulation file resumes a
    progressive render in a new renderer bit for bit (2 frames + file + 2 frames == 4 frames)."""
    from tests.test_cpu_image import _decode_png
    rm, sc, info = okrt.scenes.test_scene(with_textures=True)
    W, H = 192, 108
    env = okrt.scenes.procedural_sky_faces(32)

    def fresh():
        rt = okrt.RayTracer(0)
        rt.initiate(W, H, rm, environmentMapFaces=env, maxBounces=3)
        rt.setScene(sc)
        rt.createOctTree(sc, 2, 50)
        return rt

    whole = fresh()
    whole.render(frames=4)
    first = fresh()
    first.render(frames=2)
    first.saveAccumulation(str(tmp_path / "state.okrtacc"))
    first.shutdown()
    second = fresh()
    second.loadAccumulation(str(tmp_path / "state.okrtacc"))
    assert second.getNumAccumulationFrames() == 2
    second.render(frames=2)
    assert np.array_equal(second.readAccumulation().view(np.uint32), whole.readAccumulation().view(np.uint32))
    assert np.array_equal(second.readResult(), whole.readResult())
    second.saveScreenshot(str(tmp_path / "render.png"))
    assert np.array_equal(_decode_png(tmp_path / "render.png"), whole.readResult())
    whole.shutdown(); second.shutdown()


def test_raytracer_mirror_class(okrt, oracle):
    """The reference-facing class (raytracer.py mirrors RayTracer.h:82-116) end to end."""
    rm, sc, info = okrt.scenes.test_scene(with_textures=True)
    W, H = 192, 108
    rt = okrt.RayTracer(0)
    rt.initiate(W, H, rm, environmentMapFaces=okrt.scenes.procedural_sky_faces(32), maxBounces=3)
    rt.setScene(sc)
    rt.bindSpheres(True)
    rt.createOctTree(sc, 2, 50)
    rt.render(frames=2)
    rt.sync()
    assert rt.getNumAccumulationFrames() == 2
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    os_ = helpers.oracle_scene(oracle, host, envFaces=okrt.scenes.procedural_sky_faces(32))
    acc = np.zeros((H, W, 4), np.float32); res = np.zeros((H, W, 4), np.uint8)
    rd = host["rd"].copy()
    for f in (1, 2):
        rd[0]["numAccumulationFrames"] = f
        oracle.render(os_, rd, 3, accum=acc, result=res)
    assert np.array_equal(rt.readAccumulation().view(np.uint32), acc.view(np.uint32))
    assert np.array_equal(rt.readResult(), res)
    rt.resetAccumulation()
    assert rt.getNumAccumulationFrames() == 0
    rt.shutdown()


def test_full_size_primary_hits_1080p(okrt, oracle):
    """BASELINE size: the 262k-triangle stand-in at 1920x1080, primary rays — every triangle id equals the oracle's."""
    import bench
    host, env, info = bench.build_scene(okrt)
    r = helpers.Renderer(okrt, host, bench.WIDTH, bench.HEIGHT, envFaces=env)
    r.options(maxBounces=0, writeAov=1)
    r.render(host["rd"].copy())
    g = r.aov()
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=env), _frame1(host["rd"]), 0, aov_dtype=okrt.abi.AOV)
    assert np.array_equal(g["triIdx"], o["aov"]["triIdx"]) and np.array_equal(g["hitIdx"], o["aov"]["hitIdx"])
    for f in ("baryU", "baryV", "distance"):
        assert np.array_equal(g[f].view(np.uint32), o["aov"][f].view(np.uint32))
    assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32))
    r.close()


def test_full_size_8_bounces_against_oracle(okrt, oracle):
    """BASELINE size (C4: 1920x1080, 8 bounces, 262k-triangle stand-in): the whole RGBA32F accumulation image of frame 1
    equals the oracle's bit for bit, through both pipelines and both traversal orders; segment counts agree with the
    oracle's, alpha counts samples, the fallback stays rare."""
    import bench
    host, env, info = bench.build_scene(okrt)
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=env), _frame1(host["rd"]), 8)
    for pipeline, traversal in ((0, 0), (1, 0), (0, 1)):
        r = helpers.Renderer(okrt, host, bench.WIDTH, bench.HEIGHT, envFaces=env)
        r.options(maxBounces=8, pipeline=pipeline, traversal=traversal)
        r.render(host["rd"].copy(), frames=1)
        st = r.stats()
        acc = r.accumulation()
        r.close()
        bad = np.any(acc.view(np.uint32) != o["accum"].view(np.uint32), axis=2)
        assert not bad.any(), "pipeline %d traversal %d: %d pixels differ from the oracle" % (pipeline, traversal, int(bad.sum()))
        assert int(st["raySegments"]) == o["segments"]
        assert np.all(acc[..., 3] == 1.0)
        assert int(st["fallbackSegments"]) < 0.01 * o["segments"]


def test_cpp_mirror_on_gpu(okrt, tmp_path):
    from tests.test_cpu_host_mirror import test_cpp_mirror_builds_and_runs
    test_cpp_mirror_builds_and_runs(okrt, tmp_path)


def _lanes_run(okrt, host, W, H, env, lanes, calls, read_every_call):
    """`calls` = frames per okrt_render_frames call; returns u1, the RGBA8 results read after every call, the segment count."""
    import ctypes
    r = helpers.Renderer(okrt, host, W, H, envFaces=env)
    r.options(maxBounces=4, batchesInFlight=lanes)
    rd = host["rd"].copy()
    images = []
    pinned = [np.zeros((H, W, 4), np.uint8) for _ in range(2)]
    moved = ctypes.c_uint64(0)
    for k, frames in enumerate(calls):
        rd[0]["numAccumulationFrames"] = r.L.okrt_get_num_accumulation_frames(r.h)
        okrt.abi.check(r.L.okrt_set_render_data(r.h, okrt.abi.ptr(rd)))
        okrt.abi.check(r.L.okrt_render_frames(r.h, frames))
        if read_every_call:
            okrt.abi.check(r.L.okrt_read_result_async(r.h, okrt.abi.ptr(pinned[k & 1]), ctypes.byref(moved), k & 1))
            okrt.abi.check(r.L.okrt_wait_read(r.h, k & 1))
            images.append(pinned[k & 1].copy())
    acc = r.accumulation()
    seg = int(r.stats()["raySegments"])
    r.close()
    return acc, images, seg


def test_batches_in_flight_bit_identical(okrt):
    """okrt_options::batchesInFlight (launchWavefront: consecutive batches on up to four streams, each with its own path state,
    accumulate stages in frame order): many small calls give the same u1 bits, the same intermediate result images and the same
    segment count with 1, 2, 3 and 4 batches in flight, and the same u1 as ONE call for all the frames."""
    rm, sc, info = okrt.scenes.sponza_standin(8000, texture_size=64)
    W, H = 160, 90
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(32)
    calls = [1, 2, 1, 3, 1, 1, 2, 5, 1, 1, 1, 4]
    ref_acc, ref_images, ref_seg = _lanes_run(okrt, host, W, H, env, 1, calls, True)
    assert ref_acc[..., 3].max() == float(sum(calls))
    for lanes in (2, 3, 4):
        acc, images, seg = _lanes_run(okrt, host, W, H, env, lanes, calls, True)
        assert seg == ref_seg
        assert np.array_equal(acc.view(np.uint32), ref_acc.view(np.uint32))
        assert all(np.array_equal(a, b) for a, b in zip(images, ref_images))
    acc, _, seg = _lanes_run(okrt, host, W, H, env, 0, [sum(calls)], False)
    assert seg == ref_seg and np.array_equal(acc.view(np.uint32), ref_acc.view(np.uint32))


def test_batches_in_flight_with_edits_between_calls(okrt):
    """A call that changes what the kernels read or write (a light, the accumulation reset, an overwritten accumulation image,
    an AOV frame) orders the next batch after itself, without a host synchronisation in between: same images with one batch
    in flight and with four."""
    rm, sc, info = okrt.scenes.test_scene(with_textures=False)
    W, H = 160, 90
    host = helpers.build_host_side(okrt, rm, sc, info, W, H)
    env = okrt.scenes.procedural_sky_faces(32)
    abi = okrt.abi
    outs = []
    for lanes in (1, 4):
        r = helpers.Renderer(okrt, host, W, H, envFaces=env)
        r.options(maxBounces=3, batchesInFlight=lanes)
        rd = host["rd"].copy()

        def go(frames):
            rd[0]["numAccumulationFrames"] = r.L.okrt_get_num_accumulation_frames(r.h)
            abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))
            abi.check(r.L.okrt_render_frames(r.h, frames))          # no okrt_sync: the next call is queued behind it

        got = []
        go(3); go(2); go(1)
        got.append(r.accumulation())
        go(2)
        abi.check(r.L.okrt_reset_accumulation(r.h))
        go(2); go(1)
        got.append(r.accumulation()); got.append(r.result())
        light = np.zeros(1, abi.POINT_LIGHT)
        light[0]["colour"] = (0.9, 0.2, 0.1); light[0]["intensity"] = 3.0; light[0]["radius"] = 0.5; light[0]["position"] = (0.0, 4.0, 0.0)
        go(2)
        r.set_lights(pointLights=light)
        rd[0]["numPointLights"] = 1
        go(2); go(2)
        got.append(r.accumulation())
        go(1)
        abi.check(r.L.okrt_write_accumulation(r.h, abi.ptr(got[0]), 6))
        go(2); go(1)
        got.append(r.accumulation())
        r.options(writeAov=1)
        go(1)
        got.append(r.aov()["numSegments"].copy())
        r.options(writeAov=0)
        go(3)
        got.append(r.accumulation())
        r.close()
        outs.append(got)
    assert outs[0][0][..., 3].max() == 6.0 and outs[0][1][..., 3].max() == 3.0 and outs[0][4][..., 3].max() == 9.0
    for a, b in zip(*outs):
        assert np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))
