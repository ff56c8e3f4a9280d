"""Consumers of the path's intermediates against their oracle restatements:
map-step transmission (sim/map.py:117-133), loading-dependent noise scaling (sim/noise.py:35-37,60), full-rate pointing
on demand (sim/observation.py:55-58), and the lazily evaluated ``obs.zenith_scaled_pwv`` / ``TOD.weight``."""
import numpy as np
import pytest

import maria_b200 as M
from common import relmax, small_case
from oracle import atm_oracle as O

pytestmark = pytest.mark.gpu


def _both(dtype, **kw):
    case = small_case(n_det=96, duration=30.0, sample_rate=25.0, pattern="back_and_forth", polarized=True, seed=11, **kw)
    np.random.seed(case["seed"])
    res = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random, dtype=np.float64)
    np.random.seed(case["seed"])
    sim = M.Simulation(case["instrument"], case["plan"], case["site"], atmosphere="2d", dtype=dtype,
                       atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], geometry="exact"))
    tod = sim.run(units="pW")[0]
    return case, res, sim, tod


@pytest.mark.parametrize("dtype, tol", [(np.float64, 1e-10), (np.float32, 3e-6)])
def test_map_transmission_matches_oracle(cuda, dtype, tol):
    case, res, sim, _ = _both(dtype)
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    nu = case["spectrum"].side_nu
    for nu_min, nu_max in [(0.0, np.inf), (float(nu[40]), float(nu[120])), (85e9, 95e9)]:
        ref = O.map_pW_per_K_RJ(res["obs"], res["atmosphere"], nu_min, nu_max)
        dev = atm.transmission(nu_min, nu_max).cpu().numpy()
        assert dev.shape == ref.shape and dev.dtype == dtype
        assert relmax(dev, ref) < tol, (nu_min, nu_max)
    # a detector range equals the rows of the full evaluation, bit for bit
    part = atm.transmission(85e9, 95e9, det0=33, n_det=40).cpu().numpy()
    assert np.array_equal(part, dev[33:73])
    assert atm.check_flags() == 0


def test_noise_scaling_matches_oracle(cuda):
    case, res, sim, tod = _both(np.float32)
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    dets = case["instrument"].dets
    rng = np.random.default_rng(3)
    noise = rng.standard_normal(tod.shape).astype(np.float32)
    nep = {b.name: 1e-17 * (1 + i) for i, b in enumerate(dets.bands)}
    nepl = {b.name: 2e-18 * (1 + i) for i, b in enumerate(dets.bands)}
    loading = tod.device_data["atmosphere"]
    dev = atm.noise_scale(loading, cuda.as_tensor(noise).to(loading.device), [nep[b.name] for b in dets.bands],
                          [nepl[b.name] for b in dets.bands]).cpu().numpy()
    ref = O.noise_loading_pW(res["obs"], loading.cpu().numpy(), noise, nep, nepl)
    assert relmax(dev, ref) < 2e-7


def test_pointing_and_lazy_arrays(cuda):
    case, res, sim, tod = _both(np.float64)
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    o_obs = res["obs"]
    n_det, n_t = tod.shape
    # obs.coords.az / .el: device slices on demand, against the oracle's full-rate pointing
    el = obs.coords.el
    assert el.shape == (n_det, n_t)
    assert relmax(el[10:42, ::7], o_obs.coords.el[10:42, ::7]) < 1e-12
    assert relmax(obs.coords.az[[3, 77, 5]], o_obs.coords.az[[3, 77, 5]]) < 1e-12
    assert relmax(np.asarray(obs.coords.el), o_obs.coords.el) < 1e-12
    a, e = atm.pointing(det0=5, n_det=3, j0=100, n_j=50)
    assert relmax(e.cpu().numpy(), o_obs.coords.el[5:8, 100:150]) < 1e-12 and relmax(a.cpu().numpy(), o_obs.coords.az[5:8, 100:150]) < 1e-12
    # obs.zenith_scaled_pwv evaluates only the detectors a slice touches
    z = obs.zenith_scaled_pwv
    assert z.shape == (n_det, n_t)
    assert relmax(z[40:44], o_obs.zenith_scaled_pwv[40:44]) < 1e-10 and relmax(z[7], o_obs.zenith_scaled_pwv[7]) < 1e-10
    # TOD.weight: ones of the TOD's shape without the memory; TOD.sample: a slice without copying the field
    assert tod.weight.shape == tod.shape and tod.weight.strides == (0, 0) and float(tod.weight[5, 9]) == 1.0
    assert np.array_equal(tod.sample("atmosphere", slice(2, 6), slice(0, 100, 3)), tod.data["atmosphere"][2:6, 0:100:3])
