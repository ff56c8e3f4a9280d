"""Detector sharding across the GPUs of one box: one process per GPU, contiguous detector ranges,
no data-path collective -- every rank regenerates the (small) layers from the same seed / the same host noise, so the
shards of all ranks tile the single-GPU TOD bit for bit.  ``torch.distributed`` is used only for set-up agreement
checks and for gathering small results (timings, checksums); the TOD itself is never gathered on the device.
"""
from __future__ import annotations

import numpy as np

TILE = 32  # detectors per kernel tile (matm_synth.cu: DT); shard boundaries are aligned to it where possible


def detector_range(n_det, rank, world, align=TILE):
    """(first, count) of the contiguous detector shard of ``rank``.  Boundaries are multiples of ``align`` (except the
    last one), shards differ by at most ``align`` detectors, and together they tile [0, n_det) exactly."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} for world size {world}")
    if n_det < 0:
        raise ValueError("n_det must be non-negative")
    n_tiles = -(-n_det // align)
    lo_t = (n_tiles * rank) // world
    hi_t = (n_tiles * (rank + 1)) // world
    lo, hi = min(lo_t * align, n_det), min(hi_t * align, n_det)
    return lo, hi - lo


def all_ranges(n_det, world, align=TILE):
    return [detector_range(n_det, r, world, align) for r in range(world)]


def env_rank_world():
    """(rank, world, local_rank) from the torchrun environment (RANK / WORLD_SIZE / LOCAL_RANK); (0, 1, 0) without it."""
    import os

    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def assert_same_setup(dist, fingerprint):
    """Every rank must have built the same layer geometry (same host RNG stream): compare a float64 fingerprint."""
    import torch

    t = torch.as_tensor(np.asarray(fingerprint, dtype=np.float64))
    if dist.get_backend() == "nccl":
        t = t.cuda()
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if not torch.equal(lo, hi):
        raise RuntimeError("ranks disagree on the layer geometry: seed the host RNG identically on every rank")


def setup_fingerprint(atmosphere):
    """A few numbers that pin the host set-up of an initialised Atmosphere: grid sizes, transforms, propagator sums."""
    out = [float(atmosphere.timestep), float(len(atmosphere.boresight.t))]
    for k in sorted(atmosphere.processes):
        p = atmosphere.processes[k]
        out += [float(p.n_extrusion), float(p.n_cross_section), float(p.n_sample), float(np.sum(p.transform)),
                float(p.extrusion[0]), float(np.sum(p.A)), float(np.sum(p.B))]
    return np.array(out)
