"""Pin the oracle against the reference itself and write tests/golden/ref_*.npz.  TEST INFRASTRUCTURE ONLY.

Runs the UNMODIFIED reference (``/root/reference/maria``, imported in place through ``oracle/_ref_stubs.py``)
for real presets -- ``maria.Simulation(instrument, plans, site, atmosphere="2d", dtype=float64, noise=False)`` --
with ``np.random.seed(seed)``, weather / spectrum tables served from memory in the reference's own h5 layout
(spectrum/atmosphere.py:38-57, weather/__init__.py:105-135), then runs ``oracle/atm_oracle.py`` on the same inputs
with the same global-RNG stream and compares every stage.  The fixture holds the inputs (as plain arrays), the seed of
the legacy np.random stream the reference drew from, and the reference's outputs, so the CPU test-suite and the GPU parity tests can check against
the *reference's* numbers without /root/reference being present.

What this pins: a1-a14 in float64 (the reference with jax_enable_x64; jax.numpy is numpy here).  What it cannot pin:
jax's own RegularGridInterpolator code (restated in atm_oracle.rgi_linear from its published algorithm) and the
float32 "as run" rounding of jax with x64 disabled.

Usage:  python oracle/pin_against_reference.py   (only works where /root/reference exists)
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import _ref_stubs  # noqa: E402

_ref_stubs.install()

import maria  # noqa: E402  (the reference, in place)

from maria_b200 import inputs as I  # noqa: E402
from oracle import atm_oracle as O  # noqa: E402


# ---- in-memory h5 provider ------------------------------------------------------------------------------------
class _DS:
    def __init__(self, a):
        self.a = np.asarray(a)

    def __getitem__(self, k):
        return self.a[()] if k == () else self.a[k]


class _Grp(dict):
    pass


class _FakeH5:
    files = {}

    def __init__(self, path, mode="r"):
        self.d = _FakeH5.files[str(path)]

    def __enter__(self):
        return self.d

    def __exit__(self, *a):
        return False


def _spectrum_file(spec, alts=(0.0, 6000.0)):
    e = np.stack([spec._emission, spec._emission])
    o = np.stack([spec._opacity, spec._opacity])
    el_deg = np.degrees(spec.side_elevation)
    el_deg[-1] = 90.0
    return _Grp(side_altitude_m=_DS(alts), side_base_temperature_K=_DS(spec.side_base_temperature),
                side_elevation_deg=_DS(el_deg), side_zenith_pwv_mm=_DS(spec.side_zenith_pwv),
                side_nu_Hz=_DS(spec.side_nu), rayleigh_jeans_temperature_K=_DS(e), opacity_nepers=_DS(o),
                excess_path_m=_DS(0 * o))


def _weather_file(w):
    q = np.array([0.0, 0.5, 1.0])
    data = _Grp()
    for f in w.fields:
        prof = w.data[f] * (100.0 if f == "humidity" else 1.0)
        nq = np.broadcast_to(prof, (2, 2, 3, len(prof))).copy()
        data[f] = _Grp(normalized_quantiles=_DS(nq), scale=_DS(np.float64(1.0)), mean=_DS(np.float64(0.0)))
    return _Grp(quantile_levels=_DS(q), pressure_levels=_DS(np.linspace(1000, 10, len(w.altitude))),
                year_day_side=_DS(np.array([0.0, 366.0])), day_hour_side=_DS(np.array([0.0, 24.0])),
                year_day_edge_index=_DS(np.array([0, 1])), day_hour_edge_index=_DS(np.array([0, 1])), data=data)


class _RGI:
    """Stand-in for jax.scipy.interpolate.RegularGridInterpolator (restated algorithm, float64)."""

    def __init__(self, points, values, method="linear", bounds_error=False, fill_value=np.nan):
        self.points, self.values = points, values

    def __call__(self, xi):
        if not isinstance(xi, tuple):
            xi = np.asarray(xi)
            xi = tuple(xi[..., k] for k in range(xi.shape[-1]))
        return O.rgi_linear(self.points, self.values, xi, f32=False)


def _install_tables(spec, weather):
    import maria.spectrum.atmosphere as msa
    import maria.weather as mw

    _FakeH5.files["SPEC"] = _spectrum_file(spec)
    _FakeH5.files["WEATHER"] = _weather_file(weather)
    msa.h5py.File = _FakeH5
    mw.h5py.File = _FakeH5
    msa.fetch = lambda path, **k: "SPEC"
    mw.fetch = lambda path, **k: "WEATHER"
    sys.modules["jax"].scipy.interpolate.RegularGridInterpolator = _RGI


# ---- one case -------------------------------------------------------------------------------------------------
def run_case(name, instrument_name, site_name, pattern, scan_options, duration, sample_rate, center, seed,
             weather_override, det_stride):
    site = maria.get_site(site_name)
    altitude = float(site.altitude.m)
    spec = I.make_spectrum(region=site.region)
    weather = I.make_weather(site.region, altitude)  # the table content; the override goes through the reference
    _install_tables(spec, weather)

    inst = maria.get_instrument(instrument_name)
    time = np.arange(1.7e9, 1.7e9 + duration, 1 / sample_rate)
    off = np.radians(I.scan_offsets(pattern, time, **scan_options))
    pt = I.offsets_to_phi_theta_f64(off.T, np.radians(center[0]), np.radians(center[1]))
    plan = maria.Plan(time=time, phi=pt[..., 0], theta=pt[..., 1], frame="az/el", site=site)

    # ---------------- the reference ----------------
    np.random.seed(seed)
    sim = maria.Simulation(instrument=inst, plans=plan, site=site, atmosphere="2d",
                           atmosphere_kwargs={"weather": dict(weather_override)}, dtype=np.float64, noise=False,
                           progress_bars=False)
    obs = sim.obs_list[0]
    atm = obs.atmosphere
    sim._simulate_atmosphere(obs)
    ref_pwv = np.array(atm.zenith_scaled_pwv)
    ref_pwv_full = np.array(obs.zenith_scaled_pwv)
    loading = np.array(sim._compute_atmospheric_loading(obs))
    dets = obs.instrument.dets
    gain_error = np.exp(dets.gain_error * np.random.standard_normal(size=dets.n))  # simulation.py:239-241
    loading_g = loading * gain_error[:, None]
    pwv_r = float(np.round(atm.weather.pwv, 3))
    T0_r = float(np.round(atm.weather.temperature[0], 3))
    ref_krj = np.zeros_like(loading_g)
    el_full = np.array(obs.coords.el)
    for band in dets.bands:  # tod.py:106-142 on float64 data (TOD itself would narrow to float32)
        m = dets.band_name == band.name
        cal = band.cal("pW -> K_RJ", elevation=el_full[m], polarized=~np.isnan(dets.gamma[m]).all(),
                       spectrum=atm.spectrum, zenith_pwv=pwv_r, base_temperature=T0_r)
        ref_krj[m] = np.asarray(cal(loading_g[m]))
    ref_layers = {k: np.array(p.values) for k, p in atm.processes.items()}

    # ---------------- the oracle, same inputs, same RNG stream ----------------
    bands = [I.Band(b.name, nu=np.array(b.nu.Hz), tau=np.array(b.tau), efficiency=float(b.efficiency),
                    gain_error=float(b.gain_error)) for b in dets.bands]
    for b, rb in zip(bands, dets.bands):
        b.efficiency = float(rb.efficiency)  # Band(nu=, tau=) rescales by tau.max(); keep the reference's value
        b.tau = np.array(rb.tau)
        b.center = float(rb.center.Hz)
    arr = I.Array(np.array(dets.xi), np.array(dets.eta), np.array(dets.band_name), bands, np.array(dets.gamma),
                  np.array(dets.primary_size))
    arr.band_center = np.array(dets.band_center)
    o_inst = I.Instrument(arr, name=instrument_name)
    o_site = I.Site(site.region, altitude)
    o_plan = I.Plan(time, pt[..., 0], pt[..., 1])
    # the profiles exactly as the reference holds them after its load (wind correction and pwv override applied)
    ref_profiles = {f: np.array(getattr(atm.weather, f)) for f in I.Weather.fields}
    o_weather = I.Weather(site.region, altitude, ref_profiles, wind_corrected=True)
    np.random.seed(seed)
    rec = {}
    res = O.simulate(o_inst, o_plan, o_site, o_weather, spec, np.random, dtype=np.float64, record=rec)
    o_atm = res["atmosphere"]

    def rel(a, b):
        return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))

    report = {
        "weather.pwv": abs(o_weather.pwv - atm.weather.pwv) / atm.weather.pwv,
        "timestep": abs(o_atm.timestep - atm.timestep) / atm.timestep,
        "layers.res": rel(o_atm.layers["res"], atm.layers.res.values),
        "layers.pwv_rms": rel(o_atm.layers["pwv_rms"], atm.layers.pwv_rms.values),
        "A[7]": rel(o_atm.processes[7].A, atm.processes[7].A),
        "B[7]": rel(o_atm.processes[7].B, atm.processes[7].B),
        "transform[3]": rel(o_atm.processes[3].transform, atm.processes[3].transform),
        "layer values": max(rel(o_atm.processes[k].values, ref_layers[k]) for k in ref_layers),
        "zenith_scaled_pwv": rel(o_atm.zenith_scaled_pwv, ref_pwv),
        "zenith_scaled_pwv full rate": rel(res["obs"].zenith_scaled_pwv, ref_pwv_full),
        "loading pW (gain applied)": rel(res["loading_pW"], loading_g),
        "TOD K_RJ": rel(res["tod_K_RJ"], ref_krj),
    }
    print(f"== {name}: n_det={dets.n} n_t={len(time)} n_t_atm={ref_pwv.shape[1]}")
    for k, v in report.items():
        print(f"   {k:32s} max rel diff oracle vs reference = {v:.3e}")
    assert all(v < 1e-9 for v in report.values()), "oracle does not reproduce the reference"

    # ---------------- fixture ----------------
    sel = np.arange(0, dets.n, det_stride)
    fx = dict(
        seed=seed, instrument=instrument_name, site_region=site.region, site_altitude=altitude,
        time=time, az=pt[..., 0], el=pt[..., 1],
        xi=arr.xi, eta=arr.eta, band_name=arr.band_name.astype("U"), gamma=arr.gamma, primary_size=arr.primary_size,
        band_center=arr.band_center, band_names=np.array([b.name for b in bands]),
        band_efficiency=np.array([b.efficiency for b in bands]), band_gain_error=np.array([b.gain_error for b in bands]),
        **{f"weather_{f}": v for f, v in ref_profiles.items()},
        ref_timestep=atm.timestep, ref_weather_pwv=atm.weather.pwv, ref_T0=atm.weather.temperature[0],
        ref_layers_h=atm.layers.h.values, ref_layers_res=atm.layers.res.values, ref_layers_pwv_rms=atm.layers.pwv_rms.values,
        ref_zenith_scaled_pwv=ref_pwv, ref_loading_atm_rate_sel=rec["loading_atm_rate"][sel] * 0 + np.nan,
        ref_gain_error=gain_error, det_sel=sel, ref_loading_pW_sel=loading_g[sel], ref_tod_K_RJ_sel=ref_krj[sel],
    )
    for i, b in enumerate(bands):
        fx[f"band{i}_nu"] = b.nu
        fx[f"band{i}_tau"] = b.tau
    for k, p in atm.processes.items():
        fx[f"ref_values_{k}"] = np.array(ref_layers[k])
        fx[f"ref_transform_{k}"] = np.array(p.transform)
        fx[f"ref_extrusion_{k}"] = np.array(p.extrusion)
        fx[f"ref_cross_section_{k}"] = np.array(p.cross_section)
    fx.pop("ref_loading_atm_rate_sel")
    path = os.path.join(ROOT, "tests", "golden", f"ref_{name}.npz")
    np.savez_compressed(path, **fx)
    print(f"   wrote {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    run_case("m2_daisy", "MUSTANG-2", "green_bank", "daisy", dict(radius_deg=4 / 60, speed_deg=0.5 / 60), 30.0, 50.0,
             (45.0, 45.0), seed=1234, weather_override={"pwv": 10.0}, det_stride=9)
    run_case("1deg_ces", "test/1deg", "cerro_toco", "back_and_forth", dict(radius_deg=0.5, speed_deg=0.5), 12.0, 25.0,
             (120.0, 50.0), seed=4321, weather_override={}, det_stride=7)
