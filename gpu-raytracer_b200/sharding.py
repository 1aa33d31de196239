"""Multi-GPU sharding of the path, one process per GPU over torch.distributed.

Tile mode   — the image is cut into interleaved row bands; band b belongs to rank b % world.  The scene is
              replicated, there is no data-path collective; a rank's u0/u1 are valid for its own rows only.
Sample mode — rank r renders global frames r+1, r+1+world, ... (so the union of seeds equals the single-GPU
              sequence, [Q3]); the RGBA32F accumulation buffers are summed with ONE all-reduce (NCCL over NVLink on
              GPUs, gloo in the CPU tests) and resolved with saturate(u1 / totalFrames) (RaytracerCS.hlsl:529).
              Float addition order differs from the sequential reference: compare with a tolerance.
The band / frame arithmetic lives in libokrt.so (okrt_shard_owner_of_row, okrt_shard_seed_frame) and is what
the kernels use; these helpers only wrap it.
"""
import numpy as np

from . import _abi as abi


def tile_options(rank, world, band_rows=8):
    return dict(shardMode=abi.SHARD_TILE if world > 1 else abi.SHARD_NONE, shardRank=rank, shardCount=max(world, 1), shardBandRows=band_rows)


def sample_options(rank, world):
    return dict(frameOffset=rank, frameStride=max(world, 1))


def rows_of_rank(rank, world, height, band_rows=8):
    L = abi.lib()
    return np.array([y for y in range(height) if L.okrt_shard_owner_of_row(y, band_rows, world) == rank], np.int64)


def seed_frames_of_rank(rank, world, local_frames):
    L = abi.lib()
    o = abi.default_options()
    o[0]["frameOffset"], o[0]["frameStride"] = rank, max(world, 1)
    return [int(L.okrt_shard_seed_frame(abi.ptr(o), k, 1)) for k in range(1, local_frames + 1)]


def balanced_bands(band_cost, world):
    """Cost-weighted tile sharding: owner rank of every band (okrt_shard_balance: longest-processing-time-first, deterministic).
    band_cost: one number per band, e.g. the ray segments of one calibration frame (aov numSegments summed over the band's rows)."""
    cost = np.ascontiguousarray(band_cost, np.float64)
    owner = np.zeros(len(cost), np.uint32)
    abi.check(abi.lib().okrt_shard_balance(abi.ptr(cost), len(cost), world, abi.ptr(owner)))
    return owner


def band_costs_from_segments(num_segments, band_rows=8):
    """Per-band sums of a per-pixel segment count image (okrt_aov.numSegments of one frame)."""
    h = num_segments.shape[0]
    bands = (h + band_rows - 1) // band_rows
    return np.array([float(num_segments[b * band_rows:(b + 1) * band_rows].sum()) for b in range(bands)], np.float64)


def all_reduce_accumulation(tensor):
    """In-place sum of the RGBA32F accumulation image over all ranks (the single exchange step of sample mode)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
    return tensor


def accumulation_tensor(raytracer_handle_ptr, width, height, device):
    """Wraps u1 of a renderer (device memory owned by libokrt) as a torch tensor without copying."""
    import torch

    class _Holder:
        pass
    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (height, width, 4), "typestr": "<f4", "data": (int(raytracer_handle_ptr), False), "version": 3,
                                  "strides": None}
    return torch.as_tensor(h, device=device)
