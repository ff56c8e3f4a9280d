"""GPU parity tests: the CUDA path (through the C-ABI) against the reference's own outputs (tests/golden, written by
oracle/pin_against_reference.py from the unmodified reference) and against the oracle on fresh seeded inputs.

Tolerances (stated here, used below):
  fp64 path vs reference / oracle, identical host noise ........ 1e-10 relative (max |diff| / max |ref|)
  fp32 path vs the float64 oracle ............................... 2e-6 relative on the total TOD (4e-6 on the samples
                                                                  extrapolated beyond the last knot), 5e-4 relative on
                                                                  its fluctuating part (tests/common.py: check32)
"""
import numpy as np
import pytest

import maria_b200 as M
from common import check32, load_case, relmax, small_case
from oracle import atm_oracle as O

pytestmark = pytest.mark.gpu

TOL64 = 1e-10


def run_device(case, dtype, units="K_RJ", **akw):
    np.random.seed(case["seed"])
    sim = M.Simulation(case["instrument"], case["plan"], case["site"], atmosphere="2d", dtype=dtype,
                       atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], geometry="exact", **akw))
    tod = sim.run(units=units)[0]
    return sim, tod


def run_oracle(case):
    np.random.seed(case["seed"])
    rec = {}
    res = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random,
                     dtype=np.float64, record=rec)
    res["record"] = rec
    return res


@pytest.mark.parametrize("name", ["m2_daisy", "1deg_ces"])
def test_fp64_matches_reference_golden(cuda, name):
    case = load_case(name)
    fx = case["fx"]
    sim, tod = run_device(case, np.float64)
    atm = sim.obs_list[0].atmosphere
    for k, p in atm.processes.items():  # a5: the generated layers
        assert relmax(p.values, fx[f"ref_values_{k}"]) < TOL64, k
    assert relmax(atm.zenith_scaled_pwv, fx["ref_zenith_scaled_pwv"]) < TOL64  # a6 + a7 + a8 + a9
    sel = fx["det_sel"]
    krj = tod.data["atmosphere"]
    assert krj.dtype == np.float64 and krj.shape == (len(fx["xi"]), len(fx["time"]))
    assert relmax(krj[sel], fx["ref_tod_K_RJ_sel"]) < TOL64  # a10..a14
    pw = tod.to("pW").data["atmosphere"]
    assert relmax(pw[sel], fx["ref_loading_pW_sel"]) < TOL64
    back = tod.to("pW").to("K_RJ").data["atmosphere"]  # reference: tests/tod/test_tod.py:37-49 round trip
    assert relmax(back, krj) < 1e-14


@pytest.mark.parametrize("name", ["m2_daisy", "1deg_ces"])
def test_fp32_within_stated_bound_of_reference(cuda, name):
    case = load_case(name)
    fx = case["fx"]
    sim, tod = run_device(case, np.float32)
    krj = tod.data["atmosphere"]
    assert krj.dtype == np.float32
    sel = fx["det_sel"]
    obs = sim.obs_list[0]
    check32(krj[sel], fx["ref_tod_K_RJ_sel"], obs.boresight.t, obs.atmosphere.boresight.t[-1])


def test_stages_against_oracle_fresh_inputs(cuda):
    """Synthetic 2-band polarised array, back-and-forth scan: every stage against the oracle with the same noise."""
    case = small_case(n_det=96, duration=30.0, sample_rate=25.0, pattern="back_and_forth", polarized=True, seed=11)
    res = run_oracle(case)
    sim, tod = run_device(case, np.float64)
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    o_atm = res["atmosphere"]
    for k, p in atm.processes.items():
        assert relmax(p.values, o_atm.processes[k].values) < TOL64
    for li, sm in res["record"]["smoothed"].items():  # a8
        assert relmax(atm.smoothed_layer(li), sm) < TOL64
    assert relmax(atm.zenith_scaled_pwv, o_atm.zenith_scaled_pwv) < TOL64
    assert relmax(np.asarray(obs.zenith_scaled_pwv), res["obs"].zenith_scaled_pwv) < TOL64  # a10
    loading = sim._compute_atmospheric_loading(obs).cpu().numpy()  # a11 + a12, no gain
    o_loading = res["loading_pW"] / res["record"]["gain_error"][:, None]
    assert relmax(loading, o_loading) < TOL64
    assert relmax(tod.data["atmosphere"], res["tod_K_RJ"]) < TOL64


def test_detector_shards_tile_the_full_tod(cuda):
    """Detector sharding (multi-GPU partitioning) is exact: shards computed separately equal the full run, bit for bit."""
    case = small_case(n_det=80, duration=12.0, sample_rate=20.0, seed=5)
    sim, tod = run_device(case, np.float32)
    full = tod.data["atmosphere"]
    obs = sim.obs_list[0]
    parts = [sim._compute_atmospheric_loading(obs, d0, n, units="K_RJ", apply_gain=True).cpu().numpy()
             for d0, n in [(0, 33), (33, 7), (40, 40)]]
    assert np.array_equal(np.concatenate(parts, axis=0), full)


def test_philox_normals_are_standard_normal(cuda):
    """matm_fill_normal: moments, independence of neighbours, reproducibility, distinct subsequences."""
    ctx = M.DeviceContext(0, "fp32")
    n = 1 << 22
    x = ctx.empty((n,), cuda.float64)
    y = ctx.empty((n,), cuda.float64)
    ctx.check(ctx.lib.matm_fill_normal(ctx.h, x.data_ptr(), n, 42, 0, ctx.stream_ptr()))
    ctx.check(ctx.lib.matm_fill_normal(ctx.h, y.data_ptr(), n, 42, 1, ctx.stream_ptr()))
    a, b = x.cpu().numpy(), y.cpu().numpy()
    se = 1 / np.sqrt(n)
    assert abs(a.mean()) < 5 * se and abs(a.var() - 1) < 5 * np.sqrt(2) * se
    assert abs(np.mean(a**3)) < 5 * np.sqrt(15) * se and abs(np.mean(a**4) - 3) < 5 * np.sqrt(96) * se
    assert abs(np.mean(a[1:] * a[:-1])) < 5 * se and abs(np.mean(a[::2] * a[1::2])) < 5 * np.sqrt(2) * se
    assert abs(np.mean(a * b)) < 5 * se  # subsequences are independent
    ctx.check(ctx.lib.matm_fill_normal(ctx.h, y.data_ptr(), n, 42, 0, ctx.stream_ptr()))
    assert np.array_equal(a, y.cpu().numpy())


def test_philox_layers_have_matern_structure_function(cuda):
    """Production RNG: the second-order structure function D(r) = 2 (1 - C(r)) of the generated layers along the
    extrusion axis follows the Matern covariance the process was built from.  The marginal
    variance is NOT tested: the reference's propagator is close to a unit root, so the second moment of one realisation
    scatters by factors of several in the reference itself."""
    case = small_case(n_det=32, duration=60.0, sample_rate=10.0, seed=3)
    np.random.seed(3)
    sim = M.Simulation(case["instrument"], case["plan"], case["site"], atmosphere="2d",
                       atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], noise="philox", seed=99))
    atm = sim.obs_list[0].atmosphere
    k = max(atm.processes)
    p = atm.processes[k]
    sf = []
    lags = np.array([1, 2, 4, 8])
    for seed in range(32):
        atm.seed = 1000 + seed
        atm.generate()
        v = p.values
        sf.append([np.mean((v[l:] - v[:-l]) ** 2) for l in lags])
    expected_sf = 2 * (1 - p.callback(lags * p.extrusion_res))
    assert np.all(np.abs(np.mean(sf, axis=0) / expected_sf - 1) < 0.2), (np.mean(sf, axis=0), expected_sf)
    # (seed, realisation) reproduces the layers; the next realisation of the same seed and another seed do not: successive
    # simulate_pwv() calls are fresh draws, like the reference's global np.random stream
    atm.seed, atm.realisation = 5, 0
    atm.generate()
    a = p.values.copy()
    atm.generate()
    assert atm.realisation == 2 and not np.array_equal(a, p.values)
    atm.realisation = 0
    atm.generate()
    assert np.array_equal(a, p.values)
    atm.seed, atm.realisation = 6, 0
    atm.generate()
    assert not np.array_equal(a, p.values)


def test_plans_of_one_simulation_draw_independent_atmospheres(cuda):
    """Two identical plans in one Simulation with one Philox seed: different observations, different atmospheres."""
    case = small_case(n_det=16, duration=10.0, sample_rate=50.0, seed=3)
    np.random.seed(3)
    sim = M.Simulation(case["instrument"], [case["plan"], case["plan"]], case["site"], atmosphere="2d",
                       atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], noise="philox", seed=99))
    t0, t1 = sim.run(units="K_RJ")
    assert t0.shape == t1.shape and not np.array_equal(t0.data["atmosphere"], t1.data["atmosphere"])


def test_out_of_grid_raises_like_the_reference(cuda):
    """A line of sight outside a layer must raise (atmosphere.py:368-369), not clamp."""
    case = small_case(n_det=16, duration=10.0, sample_rate=10.0, seed=2)
    sim, _ = run_device(case, np.float32)
    atm = sim.obs_list[0].atmosphere
    for p in atm.processes.values():
        p.vx = p.vx + 500.0  # blow the footprint out of the generated grids
    atm._layers_set = False
    atm.generate()
    with pytest.raises(RuntimeError, match="introduced nans"):
        atm.zenith_scaled_pwv


def test_3d_model_matches_oracle(cuda):
    """The coupled "3d" model (extrusion.py:69-77, atmosphere.py:249): ONE autoregressive process over all layers, Matern
    nu = 1/3, one layer per resolution element of height (20 layers here).  Its process (n_cs = 300, n_sample ~ 590) does
    not fit the register-resident chain kernel, so this test also covers the general chain kernel (ar_chain_kernel: A
    slices in shared memory staged by the TMA, cluster barrier per step), the column-sliced beam smoothing of a shared
    process, and a sampler with more than 16 layers."""
    case = small_case(n_det=16, duration=8.0, sample_rate=20.0, seed=4, fov_deg=0.2, primary=6.0)
    kw = dict(max_height=300.0)
    np.random.seed(case["seed"])
    res = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random, model="3d",
                     dtype=np.float64, atmosphere_kwargs=kw)
    out = {}
    for dtype in (np.float64, np.float32):
        np.random.seed(case["seed"])
        sim = M.Simulation(case["instrument"], case["plan"], case["site"], atmosphere="3d", dtype=dtype,
                           atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], geometry="exact", **kw))
        atm = sim.obs_list[0].atmosphere
        assert len(atm.processes) == 1 and len(atm.layers["h"]) > 16
        p = atm.processes[0]
        assert p.n_sample > 512 and p.n_cross_section > 256  # beyond the register-resident kernel
        tod = sim.run(units="K_RJ")[0]
        if dtype is np.float64:
            assert relmax(p.values, res["atmosphere"].processes[0].values) < TOL64
            assert relmax(atm.zenith_scaled_pwv, res["atmosphere"].zenith_scaled_pwv) < TOL64
            assert relmax(tod.data["atmosphere"], res["tod_K_RJ"]) < TOL64
        else:
            obs = sim.obs_list[0]
            check32(tod.data["atmosphere"], res["tod_K_RJ"], obs.boresight.t, atm.boresight.t[-1])


def test_unit_carrying_objects_give_the_same_tod(cuda):
    """``Simulation`` fed with maria-style objects (quantities with units, Band objects with ``nu.Hz``; the real ones are
    checked on CPU by tests/test_cpu_adapter.py) produces the same bits as with the plain-array views."""
    from test_cpu_adapter import Q, _Obj

    case = small_case(n_det=32, duration=10.0, sample_rate=50.0, seed=9)
    d = case["instrument"].dets
    bands = [_Obj(name=b.name, nu=Q(b.nu / 1e9, "GHz"), tau=b.tau, efficiency=b.efficiency, gain_error=b.gain_error,
                  center=Q(b.center, "Hz")) for b in d.bands]
    dets = _Obj(xi=Q(np.degrees(d.xi), "deg"), eta=Q(np.degrees(d.eta), "deg"), band_name=d.band_name, bands=bands, gamma=d.gamma,
                primary_size=Q(d.primary_size, "m"), band_center=d.band_center)
    plan = _Obj(time=case["plan"].time, phi=Q(case["plan"].phi, "rad"), theta=Q(case["plan"].theta, "rad"), frame="az/el", roll=0.0)
    site = _Obj(region=case["site"].region, altitude=Q(case["site"].altitude / 1e3, "km"))
    out = []
    for inst, pl, st in [(case["instrument"], case["plan"], case["site"]), (_Obj(dets=dets, name="q"), plan, site)]:
        np.random.seed(case["seed"])
        sim = M.Simulation(inst, pl, st, atmosphere="2d", dtype=np.float32,
                           atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], geometry="exact"))
        out.append(sim.run(units="K_RJ")[0].data["atmosphere"])
    assert relmax(out[1], out[0]) < 1e-6  # degrees -> radians round trip moves offsets by an ulp
