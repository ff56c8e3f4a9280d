"""Multi-rank parity on the CUDA path: two PROCESSES, each building the set-up from the same seed and
synthesising its detector shard through ``Simulation(..., det_range=sharding.detector_range(n, rank, 2))``; the shards,
concatenated, must equal the single-process TOD bit for bit.  Both ranks may share cuda:0 (the GPU-test box has one GPU);
with two GPUs rank r uses cuda:r.  A rank that seeds its host RNG differently must be caught by the set-up fingerprint
check ``Simulation`` runs under torch.distributed."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _case():
    from maria_b200 import configs

    # C4 geometry (two bands per pixel, Morton tiles with shared lines of sight), ragged detector count, 4 min of scan
    return configs.c4_atlast(n_det=1000, duration=240.0, sample_rate=50.0)


def _simulate(det_range, device, seed_offset=0, exchange=False):
    import maria_b200 as M

    cfg = _case()
    np.random.seed(41 + seed_offset)
    sim = M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=np.float32, det_range=det_range,
                       atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], noise="philox", seed=7,
                                              geometry="fast", device=device, layer_exchange=exchange))
    return sim


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    from maria_b200 import sharding

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    device = rank % torch.cuda.device_count()
    n = _case()["instrument"].dets.n
    det0, cnt = sharding.detector_range(n, rank, world)
    sim = _simulate((det0, cnt), device)
    tod = sim.run(units="K_RJ")[0]
    shard = tod.data["atmosphere"]
    assert shard.shape[0] == cnt
    np.save(os.path.join(out_dir, f"shard{rank}.npy"), shard)
    # SURVEY 8e option B: rank r generates the processes k % 2 == r only, the smoothed layers are broadcast
    sim_x = _simulate((det0, cnt), device, exchange=True)
    atm_x = sim_x.obs_list[0].atmosphere
    np.save(os.path.join(out_dir, f"xshard{rank}.npy"), sim_x.run(units="K_RJ")[0].data["atmosphere"])
    open(os.path.join(out_dir, f"xbytes{rank}"), "w").write(str(getattr(atm_x, "exchanged_bytes", 0)))
    # a rank with a different host seed builds a different geometry: Simulation must refuse
    caught = False
    try:
        _simulate((det0, cnt), device, seed_offset=rank)
    except RuntimeError as e:
        caught = "disagree" in str(e)
    open(os.path.join(out_dir, f"caught{rank}"), "w").write(str(caught))
    dist.barrier()
    dist.destroy_process_group()


def test_two_processes_tile_the_single_process_tod_bit_for_bit(cuda, tmp_path):
    import torch.multiprocessing as mp

    from maria_b200 import sharding

    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sim = _simulate(None, 0)
    full = sim.run(units="K_RJ")[0].data["atmosphere"]
    parts = [np.load(tmp_path / f"shard{r}.npy") for r in range(world)]
    assert [p.shape[0] for p in parts] == [c for _, c in sharding.all_ranges(full.shape[0], world)]
    assert np.array_equal(np.concatenate(parts, axis=0), full)
    assert np.isfinite(full).all() and full.std() > 0
    assert all(open(tmp_path / f"caught{r}").read() == "True" for r in range(world))
    # exchanged layers are bit-identical to self-generated ones, hence so are the shards
    xparts = [np.load(tmp_path / f"xshard{r}.npy") for r in range(world)]
    assert np.array_equal(np.concatenate(xparts, axis=0), full)
    layer_bytes = sum(4 * p.n_extrusion * p.n_cross_section for p in sim.obs_list[0].atmosphere.processes.values())
    assert all(int(open(tmp_path / f"xbytes{r}").read()) == layer_bytes for r in range(world))
