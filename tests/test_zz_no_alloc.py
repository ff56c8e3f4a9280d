"""The allocation-free entry point ``matm_synthesize_noalloc`` (include/maria_atm.h) against ``matm_synthesize``.

No counterpart in the reference: the entry point exists for callers that cannot call cudaMalloc (CUDA graph capture,
tight serving loops).  Its knot-value scratch is one file-scope ``static __device__`` array of
``MATM_NOALLOC_SCRATCH_BYTES`` (32 MiB) shared by every call, so the test makes several calls in a row, with different
detector ranges in between (stale contents of the previous call must not leak into the next one).

Tolerance: the two entry points launch the same kernels on the same inputs in the same detector order, so the bound is
the one the shard tests use -- bit-identical output.

Collected last on purpose (``pytest -x``): a failure here hides no other test.  Needs nothing from the tests before it.
"""
import ctypes as C

import numpy as np
import pytest

import maria_b200 as M
from common import small_case
from maria_b200 import _lib, configs

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 32 << 20  # MATM_NOALLOC_SCRATCH_BYTES
TILE_BYTES_PER_KNOT = 32 * 4  # one knot row of a tile of 32 detectors, float32


def _device(cfg, dtype, np_seed, **akw):
    np.random.seed(np_seed)
    return M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=dtype,
                        atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], **akw))


def _noalloc(atm, det0, n_det, out, units=_lib.UNITS_KRJ, apply_gain=1):
    ctx = atm.ctx
    return ctx.lib.matm_synthesize_noalloc(ctx.h, det0, n_det, C.c_void_p(out.data_ptr()), units, apply_gain, ctx.stream_ptr())


def test_noalloc_entry_point_matches_synthesize_on_consecutive_calls(cuda):
    cfg = configs.c4_atlast(n_det=8192, duration=600.0, sample_rate=50.0, n_groups=2)
    sim = _device(cfg, np.float32, 41, noise="philox", seed=5, geometry="fast")
    obs = sim.obs_list[0]
    atm = obs.atmosphere
    ctx = atm.ctx
    atm.generate()
    n_t, n_atm = len(obs.boresight.t), len(atm.boresight.t)
    # (det0, n_det): whole tiles of co-located pairs; a ragged range (odd start, partial last tile); a small one
    ranges = [(0, 4096), (4099, 1001), (7000, 40)]
    for d0, n in ranges:
        assert -(-n // 32) * n_atm * TILE_BYTES_PER_KNOT <= SCRATCH_BYTES
    # the existing entry point: the reference output, and the set-up call that builds the processing order of each range
    refs = [atm.synthesize(d0, n, units="K_RJ").clone() for d0, n in ranges]
    outs = [ctx.empty((n, n_t)) for _, n in ranges]
    again = ctx.empty((ranges[0][1], n_t))
    cuda.cuda.synchronize()
    free_before = cuda.cuda.mem_get_info(ctx.device)[0]
    launches_before = ctx.lib.matm_launch_count(ctx.h)
    # two calls in a row on the same range, then the other ranges, then the first range once more
    ctx.check(_noalloc(atm, *ranges[0], outs[0]))
    ctx.check(_noalloc(atm, *ranges[0], again))
    ctx.check(_noalloc(atm, *ranges[1], outs[1]))
    ctx.check(_noalloc(atm, *ranges[2], outs[2]))
    cuda.cuda.synchronize()
    # nothing was allocated on the device by the four calls (cudaMalloc works in 2 MiB steps at least); read before any torch
    # operation below, whose temporaries come from torch's own allocator
    assert free_before - cuda.cuda.mem_get_info(ctx.device)[0] < (2 << 20)
    assert cuda.equal(outs[0], refs[0]) and cuda.equal(again, refs[0])
    assert cuda.equal(outs[1], refs[1]) and cuda.equal(outs[2], refs[2])
    again.zero_()
    ctx.check(_noalloc(atm, *ranges[0], again))
    cuda.cuda.synchronize()
    assert cuda.equal(again, refs[0])
    assert bool(cuda.isfinite(refs[0]).all()) and float(refs[0].abs().max()) > 0.0
    # three kernels per call (layer expansion, loading, evaluation), counted by the library: the CUDA path ran
    assert ctx.lib.matm_launch_count(ctx.h) - launches_before == 15
    # pW, no gain: the other units / gain switch goes through the same path
    ref_pw = atm.synthesize(*ranges[2], units="pW", apply_gain=False).clone()
    ctx.check(_noalloc(atm, *ranges[2], outs[2], units=_lib.UNITS_PW, apply_gain=0))
    cuda.cuda.synchronize()
    assert cuda.equal(outs[2], ref_pw)
    atm.check_flags()

    # ---- sizes beyond the static maximum, ranges without a cached order, bad arguments: clear errors, no work ----
    n_all = cfg["instrument"].dets.n
    assert -(-n_all // 32) * n_atm * TILE_BYTES_PER_KNOT > SCRATCH_BYTES
    big = ctx.empty((1, 1))  # never written: the call must fail before any launch
    launches = ctx.lib.matm_launch_count(ctx.h)
    with pytest.raises(ValueError, match="exceeds the static maximum"):
        ctx.check(_noalloc(atm, 0, n_all, big))
    with pytest.raises((ValueError, RuntimeError), match="no cached processing order"):
        ctx.check(_noalloc(atm, 1, 64, big))
    with pytest.raises(ValueError, match="out of bounds"):
        ctx.check(_noalloc(atm, n_all - 10, 64, big))
    assert ctx.lib.matm_launch_count(ctx.h) == launches


def test_noalloc_entry_point_rejects_float64_handles(cuda):
    case = small_case(n_det=64, duration=12.0, sample_rate=20.0, seed=3)
    sim = _device(case, np.float64, 3, geometry="exact")
    atm = sim.obs_list[0].atmosphere
    atm.generate()
    ref = atm.synthesize(units="K_RJ")
    out = atm.ctx.torch.zeros_like(ref)
    with pytest.raises(ValueError, match="float32 handles only"):
        atm.ctx.check(_noalloc(atm, 0, ref.shape[0], out))
    assert float(out.abs().max()) == 0.0
