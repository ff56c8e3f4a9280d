"""Drive maria_b200.adapters with the reference's OWN objects.  TEST INFRASTRUCTURE ONLY (needs /root/reference).

Runs the unmodified reference in place (through oracle/_ref_stubs.py and the in-memory h5 tables of
oracle/pin_against_reference.py) and checks, for the MUSTANG-2 preset at green_bank:

  1. ``adapters.from_maria(instrument, plan, site)`` reproduces the plain arrays of the golden fixture bit for bit;
  2. after ``adapters.install()`` the reference's ``Observation(..., atmosphere="2d", atmosphere_kwargs={"backend": "b200"})``
     builds a ``B200Atmosphere``; its host set-up (``initialize(obs, bind=False)``: no device is touched) equals the
     reference's own ``Atmosphere.initialize`` run with the same seed: layer table, timestep, grids, transforms, A, B to
     1e-12;
  3. without the key (or with backend="reference") the reference's class is built; a bad backend raises ValueError.

Prints one line per check and exits non-zero on the first failure.  Called by tests/test_cpu_adapter.py in a subprocess
(the import stubs must not leak into the test process).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import pin_against_reference as pin  # noqa: E402  (installs the stubs, imports the reference in place)

import maria  # noqa: E402

from maria_b200 import adapters  # noqa: E402
from maria_b200 import inputs as I  # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def check(name, ok, detail=""):
    print(f"{'ok  ' if ok else 'FAIL'} {name} {detail}")
    if not ok:
        sys.exit(1)


def main():
    fx = np.load(os.path.join(ROOT, "tests", "golden", "ref_m2_daisy.npz"), allow_pickle=False)
    seed = int(fx["seed"])
    site = maria.get_site("green_bank")
    altitude = float(site.altitude.m)
    spec = I.make_spectrum(region=site.region)
    pin._install_tables(spec, I.make_weather(site.region, altitude))
    inst = maria.get_instrument("MUSTANG-2")
    plan = maria.Plan(time=fx["time"], phi=fx["az"], theta=fx["el"], frame="az/el", site=site)

    # 1. plain views
    p_inst, p_plan, p_site = adapters.from_maria(inst, plan, site)
    d = p_inst.dets
    check("from_maria: offsets", np.array_equal(d.xi, fx["xi"]) and np.array_equal(d.eta, fx["eta"]))
    check("from_maria: band names / gamma / primary", np.array_equal(d.band_name, fx["band_name"].astype(str))
          and np.array_equal(np.isnan(d.gamma), np.isnan(fx["gamma"])) and np.array_equal(d.primary_size, fx["primary_size"]))
    check("from_maria: band centre / gain error / passband", np.array_equal(d.band_center, fx["band_center"])
          and np.array_equal([b.gain_error for b in d.bands], fx["band_gain_error"])
          and np.array_equal(d.bands[0].nu, fx["band0_nu"]) and np.array_equal(d.bands[0].tau, fx["band0_tau"])
          and float(d.bands[0].efficiency) == float(fx["band_efficiency"][0]))
    check("from_maria: plan / site", np.array_equal(p_plan.time, fx["time"]) and np.array_equal(p_plan.phi, fx["az"])
          and np.array_equal(p_plan.theta, fx["el"]) and p_site.altitude == float(fx["site_altitude"])
          and p_site.region == str(fx["site_region"]))
    check("from_maria: mueller / fov", np.array_equal(d.mueller00(), np.asarray(inst.dets.mueller())[..., 0, 0])
          and abs(d.field_of_view - float(inst.dets.field_of_view.rad)) < 1e-15)

    # 2. the reference's own set-up, then the patched one with the same seed
    import maria.sim.observation as mso

    akw = {"weather": {"pwv": 10.0}}
    np.random.seed(seed)
    ref_obs = mso.Observation(inst, plan, site, atmosphere="2d", atmosphere_kwargs=dict(akw))
    ref_atm = ref_obs.atmosphere
    ref_atm.initialize(ref_obs)
    cls = adapters.install()
    check("install: factory default is the reference", type(mso.Observation(inst, plan, site, atmosphere="2d",
                                                                            atmosphere_kwargs=dict(akw)).atmosphere) is type(ref_atm))
    try:
        mso.Observation(inst, plan, site, atmosphere="2d", atmosphere_kwargs=dict(akw, backend="tpu"))
        check("install: bad backend raises ValueError", False)
    except ValueError:
        check("install: bad backend raises ValueError", True)
    np.random.seed(seed)
    obs = mso.Observation(inst, plan, site, atmosphere="2d", atmosphere_kwargs=dict(akw, backend="b200", geometry="exact"))
    atm = obs.atmosphere
    check("install: backend='b200' builds B200Atmosphere", isinstance(atm, cls))
    check("weather / spectrum views", abs(atm.weather.pwv - ref_atm.weather.pwv) < 1e-12 * ref_atm.weather.pwv
          and np.array_equal(atm.weather.temperature, ref_atm.weather.temperature)
          and np.array_equal(atm.spectrum.side_elevation, ref_atm.spectrum.side_elevation)
          and np.array_equal(atm.spectrum._emission, ref_atm.spectrum._emission))
    atm.initialize(obs, bind=False)  # host set-up only: no device here
    L, RL = atm.layers, ref_atm.layers
    check("initialize: timestep", abs(atm.timestep - ref_atm.timestep) <= 1e-12 * ref_atm.timestep, f"{atm.timestep}")
    for col in ("h", "res", "pwv_rms", "z"):
        check(f"initialize: layers.{col}", rel(L[col], RL[col].values) < 1e-12)
    check("initialize: fixture layer table", rel(L["res"], fx["ref_layers_res"]) < 1e-12 and rel(L["pwv_rms"], fx["ref_layers_pwv_rms"]) < 1e-12)
    worst = 0.0
    for k, rp in ref_atm.processes.items():
        p = atm.processes[k]
        worst = max(worst, rel(p.transform, rp.transform), rel(p.extrusion, rp.extrusion), rel(p.cross_section, rp.cross_section),
                    rel(p.A, rp.A), rel(p.B, rp.B), rel(p.transform, fx[f"ref_transform_{k}"]), rel(p.extrusion, fx[f"ref_extrusion_{k}"]))
        if not (p.A.shape == rp.A.shape and np.array_equal(p.extrusion_sample_index, rp.extrusion_sample_index)):
            check(f"initialize: stencil of process {k}", False)
    check("initialize: transforms, grids, A, B of all processes", worst < 1e-12, f"max rel diff {worst:.2e}")
    check("initialize: boresight", rel(atm.boresight.az, ref_atm.boresight.az) < 1e-14 and rel(atm.boresight.el, ref_atm.boresight.el) < 1e-14)

    # 3. maria_b200.Simulation takes the reference's objects directly (host part only here)
    import maria_b200 as M

    o2 = M.Observation(inst, plan, site)
    check("maria_b200.Observation accepts the reference's objects", np.array_equal(o2.boresight.az, fx["az"]) and o2.instrument.dets.n == len(fx["xi"]))
    adapters.uninstall()
    check("uninstall restores the reference", mso.Atmosphere is type(ref_atm))
    print("adapter checks passed")


if __name__ == "__main__":
    main()
