"""CPU tests of the multi-GPU host logic: detector ranges tile the focal plane, and a world-size-2
``gloo`` job in which every rank builds the set-up from the same seed, takes its shard, and the shards tile the
single-process result.  The per-shard arithmetic on CPU is the oracle (the checker); on the GPU box the same
partitioning is exercised by ``test_detector_shards_tile_the_full_tod`` through the CUDA path."""
import os
import socket
import sys

import numpy as np
import pytest

from maria_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_det", [0, 1, 31, 32, 33, 217, 4764, 100_000])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_ranges_tile_exactly(n_det, world):
    ranges = sharding.all_ranges(n_det, world)
    assert ranges[0][0] == 0 and sum(n for _, n in ranges) == n_det
    for (a0, an), (b0, _) in zip(ranges[:-1], ranges[1:]):
        assert a0 + an == b0
    counts = [n for _, n in ranges]
    assert max(counts) - min(counts) <= sharding.TILE
    for lo, _ in ranges:
        assert lo % sharding.TILE == 0 or lo == n_det


def test_bad_rank_raises():
    with pytest.raises(ValueError):
        sharding.detector_range(100, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    import maria_b200 as M
    from common import small_case
    from oracle import atm_oracle as O

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    case = small_case(n_det=80, duration=6.0, sample_rate=10.0, seed=13)
    dets = case["instrument"].dets
    # host set-up on every rank from the same seed (what bench.py / Simulation(det_range=...) do on the GPU box)
    np.random.seed(13)
    obs = M.Observation(case["instrument"], case["plan"], case["site"])
    atm = M.Atmosphere(weather=case["weather"], spectrum=case["spectrum"])
    atm.initialize(obs, bind=False)
    sharding.assert_same_setup(dist, sharding.setup_fingerprint(atm))
    det0, n = sharding.detector_range(dets.n, rank, world)
    # the checker: the oracle over the whole array with the same noise on every rank; each rank keeps its shard
    np.random.seed(13)
    full = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random,
                      dtype=np.float64)["tod_K_RJ"]
    shard = torch.from_numpy(np.ascontiguousarray(full[det0:det0 + n]))
    # small result exchange: per-rank (det0, n, checksum); the TOD itself is never gathered in production
    info = torch.tensor([det0, n, float(shard.sum())], dtype=torch.float64)
    gathered = [torch.zeros(3, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, info)
    np.save(os.path.join(out_dir, f"shard{rank}.npy"), shard.numpy())
    if rank == 0:
        np.save(os.path.join(out_dir, "info.npy"), torch.stack(gathered).numpy())
        np.save(os.path.join(out_dir, "full.npy"), full)
    # a rank with a different seed must be caught
    np.random.seed(13 + rank)
    atm2 = M.Atmosphere(weather=case["weather"], spectrum=case["spectrum"])
    atm2.initialize(M.Observation(case["instrument"], case["plan"], case["site"]), bind=False)
    caught = False
    try:
        sharding.assert_same_setup(dist, sharding.setup_fingerprint(atm2))
    except RuntimeError:
        caught = True
    open(os.path.join(out_dir, f"caught{rank}"), "w").write(str(caught))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_shards_tile_the_tod(tmp_path):
    import torch.multiprocessing as mp

    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    info = np.load(tmp_path / "info.npy")
    full = np.load(tmp_path / "full.npy")
    parts = [np.load(tmp_path / f"shard{r}.npy") for r in range(world)]
    assert [tuple(map(int, r[:2])) for r in info] == sharding.all_ranges(full.shape[0], world)
    assert np.array_equal(np.concatenate(parts, axis=0), full)
    assert abs(info[:, 2].sum() - full.sum()) < 1e-9 * abs(full.sum())
    assert all(open(tmp_path / f"caught{r}").read() == "True" for r in range(world))
