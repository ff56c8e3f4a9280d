"""CPU tests of the oracle: ``oracle/atm_oracle.py`` against the golden vectors the UNMODIFIED reference produced
(tests/golden/ref_*.npz, written by oracle/pin_against_reference.py), against the reference's own pinned helper tests,
and against the domain's size-independent properties.  No GPU, no /root/reference at run time."""
import numpy as np
import pytest
import scipy.interpolate
import scipy.ndimage

from common import load_case, relmax, small_case
from oracle import atm_oracle as O

PIN = 1e-12  # oracle vs reference: same float64 algorithm, same LAPACK -> agreement to rounding


@pytest.fixture(scope="module", params=["m2_daisy", "1deg_ces"])
def golden(request):
    case = load_case(request.param)
    np.random.seed(case["seed"])
    rec = {}
    res = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random,
                     dtype=np.float64, record=rec)
    return case, res, rec


def test_layer_table_matches_reference(golden):  # a1: extrusion.py:27-110
    case, res, _ = golden
    fx, atm = case["fx"], res["atmosphere"]
    assert relmax(atm.layers["h"], fx["ref_layers_h"]) < PIN
    assert relmax(atm.layers["res"], fx["ref_layers_res"]) < PIN
    assert relmax(atm.layers["pwv_rms"], fx["ref_layers_pwv_rms"]) < PIN
    assert abs(atm.timestep - float(fx["ref_timestep"])) < PIN * atm.timestep
    assert abs(atm.weather.pwv - float(fx["ref_weather_pwv"])) < PIN * atm.weather.pwv


def test_geometry_matches_reference(golden):  # a2: atmosphere.py:81-281 (hull, SLSQP alignment, grids)
    case, res, _ = golden
    fx, atm = case["fx"], res["atmosphere"]
    for k, p in atm.processes.items():
        assert relmax(p.transform, fx[f"ref_transform_{k}"]) < 1e-9, k
        assert p.extrusion.shape == fx[f"ref_extrusion_{k}"].shape
        assert relmax(p.extrusion, fx[f"ref_extrusion_{k}"]) < 1e-9
        assert relmax(p.cross_section, fx[f"ref_cross_section_{k}"]) < 1e-9


def test_generated_layers_match_reference(golden):  # a3-a5: process.py:19-209
    case, res, _ = golden
    for k, p in res["atmosphere"].processes.items():
        assert relmax(p.values, case["fx"][f"ref_values_{k}"]) < 1e-10, k


def test_zenith_pwv_and_tod_match_reference(golden):  # a6-a14
    case, res, rec = golden
    fx = case["fx"]
    sel = fx["det_sel"]
    assert relmax(res["atmosphere"].zenith_scaled_pwv, fx["ref_zenith_scaled_pwv"]) < PIN
    assert relmax(rec["gain_error"], fx["ref_gain_error"]) < PIN
    assert relmax(res["loading_pW"][sel], fx["ref_loading_pW_sel"]) < PIN
    assert relmax(res["tod_K_RJ"][sel], fx["ref_tod_K_RJ_sel"]) < PIN
    assert np.isfinite(res["tod_K_RJ"]).all()  # the reference's own test for this path: no NaNs (tests/atmosphere)


# ---- pieces of third-party arithmetic the oracle restates: checked against the library routine itself -------------
def test_rgi_linear_equals_scipy_inside_and_nan_outside():
    """jax RegularGridInterpolator semantics (SURVEY.md section 8a notes): scipy agrees inside the grid; outside -> NaN."""
    rng = np.random.default_rng(0)
    g0, g1, g2 = np.sort(rng.uniform(0, 1, 5)), np.linspace(-2, 3, 7), np.array([0.0, 0.5, 4.0])
    v = rng.standard_normal((5, 7, 3))
    x = (rng.uniform(g0[0], g0[-1], 500), rng.uniform(-2, 3, 500), rng.uniform(0, 4, 500))
    want = scipy.interpolate.RegularGridInterpolator((g0, g1, g2), v)(np.stack(x, -1))
    assert relmax(O.rgi_linear((g0, g1, g2), v, x), want) < 1e-13
    # exactly on the first / last grid line is inside; beyond is NaN (never clamped)
    edge = O.rgi_linear((g0, g1, g2), v, (np.array([g0[0], g0[-1]]), np.array([-2.0, 3.0]), np.array([0.0, 4.0])))
    assert np.allclose(edge, [v[0, 0, 0], v[-1, -1, -1]])
    out = O.rgi_linear((g0, g1, g2), v, (np.array([g0[0] - 1e-9, 0.5 * (g0[0] + g0[1])]), np.array([0.0, 3.0 + 1e-9]),
                                        np.array([1.0, 1.0])))
    assert np.isnan(out).all()
    f32 = O.rgi_linear((g0, g1, g2), v, x, f32=True)
    assert f32.dtype == np.float32 and relmax(f32, want) < 1e-5


def test_offsets_round_trip_like_reference_test():
    """maria/tests/coordinates/test_coordinates.py:7-19: offsets -> (phi, theta) is a rotation of the unit sphere."""
    rng = np.random.default_rng(1)
    dX = np.radians(rng.uniform(-2, 2, (100, 2)))
    cphi, ctheta = np.radians(70.0), np.radians(50.0)
    phi, theta = O.offsets_to_phi_theta(dX[:, None, :], np.array([cphi]), np.array([ctheta]))
    phi, theta = phi.ravel(), theta.ravel()
    # angular distance from the boresight equals |offset| (the map is an isometry of the sphere)
    cosd = np.sin(theta) * np.sin(ctheta) + np.cos(theta) * np.cos(ctheta) * np.cos(phi - cphi)
    assert np.max(np.abs(np.arccos(np.clip(cosd, -1, 1)) - np.hypot(dX[:, 0], dX[:, 1]))) < 1e-7
    zphi, ztheta = O.offsets_to_phi_theta(np.full((1, 1, 2), 1e-12), np.array([cphi]), np.array([ctheta]))
    assert abs(zphi.ravel()[0] - cphi) < 1e-9 and abs(ztheta.ravel()[0] - ctheta) < 1e-9


def test_matern_table_is_accurate_and_normalised():  # functions/__init__.py:30-74
    r = np.geomspace(1e-4, 50, 200)
    for nu in (1 / 3, 5 / 6):
        approx = O.approximate_normalized_matern(r, nu=nu, r0=1.0)
        exact = O.normalized_matern(r, nu)
        assert np.max(np.abs(approx - exact)) < 2e-3
    assert abs(O.approximate_normalized_matern(np.array([1e-9]), nu=5 / 6, r0=1.0)[0] - 1) < 1e-6
    assert O.approximate_normalized_matern(np.array([2e3]), nu=5 / 6, r0=1.0)[0] == 0


def test_fast_psd_inverse():  # utils/linalg.py:95-102
    rng = np.random.default_rng(2)
    M = rng.standard_normal((40, 40))
    M = M @ M.T + 40 * np.eye(40)
    assert relmax(O.fast_psd_inverse(M) @ M, np.eye(40)) < 1e-12


# ---- size-independent properties of the process ------------------------------------------------------------------
def _process(n_cs=12, n_ext=60, res=5.0):
    cs = np.c_[np.linspace(-30, 30, n_cs), 1000.0 * np.ones(n_cs)]
    p = O.AutoregressiveProcess(cs, res * np.arange(n_ext), dict(nu=5 / 6, r0=1000.0), jitter=1e-8)
    p.compute_covariance_matrices(np.random.default_rng(0))
    return p


def test_ar_stencil_and_propagator_properties():  # process.py:44-67,167-174
    p = _process()
    esi = p.extrusion_sample_index
    assert esi[0] == 0 and esi[-1] == p.n_extrusion - 1 and np.all(np.diff(esi) >= 0)
    rows = np.unique(esi)
    assert set(rows[1:-1]) <= {2**i for i in range(12)}  # rows {0, 1, 2, 4, ..., n_ext - 1}
    assert p.A.shape == (p.n_cross_section, p.n_sample) and p.B.shape == (p.n_cross_section,) * 2
    assert np.allclose(p.B, np.tril(p.B)) and np.all(np.diag(p.B) > 0)
    assert np.all(p.A.sum(axis=-1) <= 1 + 1e-9)  # the stability check of process.py:169-172


def test_ar_run_is_linear_in_the_noise():
    """values(noise) is a linear map: values(a n1 + b n2) = a values(n1) + b values(n2)."""
    p = _process()
    rng = np.random.default_rng(3)
    n_ext, n_cs = p.n_extrusion, p.n_cross_section

    def draw():
        return [rng.standard_normal((3 * n_ext, n_cs))] + [rng.standard_normal(n_cs) for _ in range(2 * n_ext)]

    n1, n2 = draw(), draw()
    out = []
    for noise in (n1, n2, [0.3 * a - 1.7 * b for a, b in zip(n1, n2)]):
        p.run(O.ReplayRNG(noise))
        out.append(np.array(p.values))
    assert relmax(0.3 * out[0] - 1.7 * out[1], out[2]) < 1e-12


def test_gaussian_smoothing_is_scipy_reflect_truncate4(golden):  # a8: atmosphere.py:328-344
    case, res, rec = golden
    atm = res["atmosphere"]
    li = max(rec["smoothed"])
    sm = rec["smoothed"][li]
    assert sm.shape[0] == atm.processes[int(atm.layers["process_index"][li])].n_extrusion
    assert np.isfinite(sm).all()
    # smoothing conserves the mean of a constant field and never increases the range
    k = int(atm.layers["process_index"][li])
    v = atm.processes[k].values[:, atm.processes[k].labels == li]
    assert sm.min() >= v.min() - 1e-12 and sm.max() <= v.max() + 1e-12


def test_tod_is_affine_in_detector_gain_and_units_round_trip():
    case = small_case(n_det=24, duration=8.0, sample_rate=10.0, seed=4)
    np.random.seed(4)
    rec = {}
    res = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random,
                     dtype=np.float64, record=rec)
    g = rec["gain_error"]
    ungained = res["loading_pW"] / g[:, None]
    # every detector sees (mean loading) + (few % fluctuation); unpolarised Mueller[0, 0] = 1
    assert np.all(ungained > 0) and np.all(ungained.std(axis=1) / ungained.mean(axis=1) < 0.2)
    k = res["tod_K_RJ"] / res["loading_pW"]  # the calibration factor depends on the elevation only
    assert np.all(k > 0) and np.isfinite(k).all()
    assert res["tod_K_RJ"].shape == (case["instrument"].dets.n, len(case["plan"].time))


def test_empty_and_degenerate_inputs_raise():
    case = small_case(n_det=8, duration=0.2, sample_rate=10.0, seed=1)  # 2 samples: fewer than 4 knots for the cubic
    np.random.seed(1)
    with pytest.raises(Exception):
        O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random)


@pytest.mark.parametrize("name", ["m2_daisy", "1deg_ces"])
def test_as_run_float32_variant_stays_within_the_fp32_bound_of_the_float64_oracle(name):
    """The reference hands its interpolators to jax with x64 disabled; the oracle's ``f32=True`` variant casts where the
    reference does.  On the same geometry and the same generated layers it must sit within the bound this repo states for
    its own float32 device path (tests/common.py), i.e. the reference-as-run and the CUDA path are equally far from the
    float64 reference (scripts/fp32_asrun_table.py)."""
    import importlib.util
    import os

    from common import TOL32_FLUCT, TOL32_TAIL, load_case

    spec = importlib.util.spec_from_file_location(
        "fp32_asrun_table", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts", "fp32_asrun_table.py"))
    table = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(table)
    case = load_case(name)
    ref, as_run, stored, stages = table.as_run_pair(case, case["seed"])
    assert as_run.dtype == np.float32 and as_run.shape == ref.shape
    assert table.rel(stored, ref) < 1e-7                  # storage alone: half an ulp of float32
    assert 1e-8 < table.rel(as_run, ref) < TOL32_TAIL     # the casts do something, and stay inside the stated bound
    assert table.fluct(as_run, ref) < TOL32_FLUCT
    assert stages["zenith_scaled_pwv"] < 1e-6 and stages["az"] < 1e-6 and stages["el"] < 1e-6
