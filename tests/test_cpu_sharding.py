"""world_size-2 gloo tests of the N>1 path's host logic: band ownership, seed-frame interleave and
the single all-reduce of sample mode.  The renderer on each rank is the ORACLE here (there is no GPU); the sharding
arithmetic is the product's (libokrt.so)."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_band_ownership_partitions_the_image(okrt):
    sh = okrt.sharding
    for world in (1, 2, 4, 8):
        for height, band in ((1080, 8), (2160, 16), (90, 8)):
            rows = [sh.rows_of_rank(r, world, height, band) for r in range(world)]
            allrows = np.sort(np.concatenate(rows))
            assert np.array_equal(allrows, np.arange(height))
            assert max(len(r) for r in rows) - min(len(r) for r in rows) <= band


def test_cost_balanced_band_table(okrt):
    """okrt_shard_balance (host-only): every band gets one owner, the same table comes out of every call (each rank computes it
    for itself), and no rank carries more than the ideal share plus one band's cost (longest-processing-time-first)."""
    rng = np.random.default_rng(5)
    for bands, world in ((135, 8), (135, 4), (270, 8), (17, 2), (5, 8)):
        cost = rng.uniform(1.0, 3.0, bands) * np.linspace(0.5, 2.0, bands)
        owner = okrt.sharding.balanced_bands(cost, world)
        again = okrt.sharding.balanced_bands(cost.copy(), world)
        assert owner.shape == (bands,) and owner.max() < world and np.array_equal(owner, again)
        loads = np.array([cost[owner == k].sum() for k in range(world)])
        assert abs(loads.sum() - cost.sum()) < 1e-9 * cost.sum()
        assert loads.max() <= cost.sum() / world + cost.max() + 1e-9
    seg = np.arange(20 * 4, dtype=np.uint32).reshape(20, 4)
    per_band = okrt.sharding.band_costs_from_segments(seg, 8)
    assert per_band.tolist() == [float(seg[0:8].sum()), float(seg[8:16].sum()), float(seg[16:20].sum())]


def test_seed_frames_interleave(okrt):
    sh = okrt.sharding
    for world in (1, 2, 8):
        frames = sorted(f for r in range(world) for f in sh.seed_frames_of_rank(r, world, 16))
        assert frames == list(range(1, 16 * world + 1))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    okrt = importlib.import_module("gpu-raytracer_b200")
    from tests import helpers, oracle_binding
    oracle = oracle_binding.load()
    rm, sc, info = okrt.scenes.test_scene()
    W, H, bounces, total = 64, 40, 3, 8
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    env = okrt.scenes.procedural_sky_faces(8)
    scene = helpers.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()

    # ---- sample mode: my frames, one all-reduce, resolve
    acc = np.zeros((H, W, 4), np.float32)
    for k, f in enumerate(okrt.sharding.seed_frames_of_rank(rank, world, total // world), 1):
        rd[0]["numAccumulationFrames"] = k
        oracle.render(scene, rd, bounces, seedFrame=f, coverAll=1, accum=acc, threads=1)
    t = torch.from_numpy(acc)
    okrt.sharding.all_reduce_accumulation(t)

    # ---- tile mode: my bands of frame 1, gathered without any reduction
    rows = okrt.sharding.rows_of_rank(rank, world, H, 8)
    tile = np.zeros((H, W, 4), np.float32)
    rd[0]["numAccumulationFrames"] = 1
    for y in rows:
        oracle.render(scene, rd, bounces, coverAll=1, accum=tile, rows=(int(y), int(y) + 1), threads=1)
    gathered = [torch.zeros(H, W, 4) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(tile))
    if rank == 0:
        np.save(os.path.join(out_dir, "sample.npy"), t.numpy())
        np.save(os.path.join(out_dir, "tile.npy"), sum(g.numpy() for g in gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gloo(okrt, oracle, tmp_path):
    import torch.multiprocessing as mp
    from tests import helpers
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    rm, sc, info = okrt.scenes.test_scene()
    W, H, bounces, total = 64, 40, 3, 8
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    env = okrt.scenes.procedural_sky_faces(8)
    scene = helpers.oracle_scene(oracle, host, envFaces=env)
    rd = host["rd"].copy()
    acc = np.zeros((H, W, 4), np.float32)
    for f in range(1, total + 1):
        rd[0]["numAccumulationFrames"] = f
        one = oracle.render(scene, rd, bounces, coverAll=1, accum=acc) if f > 1 else oracle.render(scene, rd, bounces, coverAll=1, accum=acc)
        if f == 1:
            first = acc.copy()
    sample = np.load(tmp_path / "sample.npy")
    tile = np.load(tmp_path / "tile.npy")
    # KAT-6 (second half): sample-sharded vs sequential mean, RMSE <= 1e-4 (summation order differs)
    rmse = float(np.sqrt(np.mean((sample / total - acc / total) ** 2)))
    assert rmse <= 1e-4, rmse
    assert np.array_equal(sample[..., 3], acc[..., 3])                  # sample counts are exact
    assert np.array_equal(tile.view(np.uint32), first.view(np.uint32))  # tile mode is exact: disjoint pixels
