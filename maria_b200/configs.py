"""The five BASELINE.json configurations as synthetic, network-free inputs.

Each builder returns ``dict(instrument, plan, site, weather, spectrum, name)``.  ``n_det`` / ``duration`` can be
scaled down for tests; the defaults are the full sizes.
"""
from __future__ import annotations

import numpy as np

from . import inputs as I


def _pack(name, instrument, plan, site, pwv=1.0, T0=280.0):
    weather = I.make_weather(site.region, site.altitude, pwv=pwv, T0=T0)
    return dict(name=name, instrument=instrument, plan=plan, site=site, weather=weather, spectrum=I.make_spectrum(site.region))


def c1_mustang2_daisy(duration=600.0, sample_rate=50.0):
    """MUSTANG-2 preset (217 detectors, 100 m primary, fov 0.07 deg, band m2/f093) at green_bank, 10-min daisy."""
    band = I.Band("m2/f093", **I.M2_F093)
    inst = I.make_instrument(217, [band], 0.07, 100.0, name="MUSTANG-2")
    plan = I.Plan.generate(duration=duration, sample_rate=sample_rate, scan_center=(45.0, 45.0), scan_pattern="daisy",
                           scan_options=dict(radius_deg=4 / 60, speed_deg=0.5 / 60))
    return _pack("C1 MUSTANG-2 daisy", inst, plan, I.Site("green_bank", 825.0), pwv=10.0, T0=285.0)


def _act_bands():
    return [I.Band("f090", center=90e9, width=30e9, efficiency=0.5, gain_error=5e-2),
            I.Band("f150", center=150e9, width=40e9, efficiency=0.5, gain_error=5e-2),
            I.Band("f220", center=220e9, width=50e9, efficiency=0.5, gain_error=5e-2)]


def c2_act_ces(n_per_array=397, duration=600.0, sample_rate=200.0):
    """ACT-like: three hex arrays x 2 bands-ish x 2 polarisations (~4.8k detectors), 6 m primary, constant-elevation
    back-and-forth scan at cerro_toco, 3 bands."""
    b = _act_bands()
    groups = [[b[0], b[1]], [b[0], b[1]], [b[1], b[2]]]
    inst = I.make_instrument(n_per_array, groups, 1.0, 6.0, polarized=True, name="ACT-like",
                             focal_plane_offsets_deg=[(0.0, 0.8), (-0.7, -0.4), (0.7, -0.4)])
    plan = I.Plan.generate(duration=duration, sample_rate=sample_rate, scan_center=(90.0, 45.0), scan_pattern="back_and_forth",
                           scan_options=dict(radius_deg=5.0, speed_deg=1.5))
    return _pack("C2 ACT-like CES", inst, plan, I.Site("cerro_toco", 5190.0), pwv=1.0, T0=272.0)


def c3_multiband(n_det=3000, duration=600.0, sample_rate=50.0, pwv=1.0):
    """3-band 90/150/220 GHz array, 8 layers, PWV set through the weather override."""
    inst = I.make_instrument(n_det // 3, [_act_bands()], 0.5, 5.0, name="test/1deg-like")
    plan = I.Plan.generate(duration=duration, sample_rate=sample_rate, scan_center=(150.0, 55.0), scan_pattern="daisy",
                           scan_options=dict(radius_deg=0.5, speed_deg=0.5))
    return _pack(f"C3 multi-band pwv={pwv}", inst, plan, I.Site("cerro_toco", 5190.0), pwv=pwv, T0=272.0)


def c4_atlast(n_det=100_000, duration=3600.0, sample_rate=50.0, n_groups=1):
    """AtLAST-scale: ``n_det`` detectors hex-packed in a 2 deg field, 50 m primary, bands 93 / 150 GHz, 1-hour raster.
    ``n_groups`` > 1 packs ``n_groups`` such sub-arrays into the field (weak-scaling benchmark: one sub-array of ``n_det``
    detectors per GPU); detectors are emitted sub-array by sub-array, band-sorted inside each, so that a contiguous
    detector shard holds every band of its own pixels."""
    bands = [I.Band("f093", center=93e9, width=30e9, efficiency=0.5, gain_error=5e-2),
             I.Band("f150", center=150e9, width=40e9, efficiency=0.5, gain_error=5e-2)]
    inst = I.make_instrument(n_det * n_groups // 2, [bands], 2.0, 50.0, name="AtLAST-like")
    if n_groups > 1:
        n_pos, per = n_det * n_groups // 2, n_det // 2
        perm = np.concatenate([np.concatenate([b * n_pos + np.arange(g * per, (g + 1) * per) for b in range(len(bands))])
                               for g in range(n_groups)])
        inst = I.Instrument(inst.dets._subset(perm), name=inst.name)
    plan = I.Plan.generate(duration=duration, sample_rate=sample_rate, scan_center=(120.0, 50.0), scan_pattern="raster",
                           scan_options=dict(radius_deg=1.0, speed_deg=0.5, n_rows=21))
    return _pack("C4 AtLAST 100k raster", inst, plan, I.Site("chajnantor", 5060.0), pwv=1.0, T0=270.0)


def c5_long(n_det=100_000, duration=21600.0, sample_rate=50.0):
    """C4 geometry for six hours (streamed in time-chunks, detector-sharded)."""
    cfg = c4_atlast(n_det=n_det, duration=duration, sample_rate=sample_rate)
    cfg["name"] = "C5 6-hour streamed"
    return cfg


CONFIGS = {"c1": c1_mustang2_daisy, "c2": c2_act_ces, "c3": c3_multiband, "c4": c4_atlast, "c5": c5_long}
