"""The reference "as run" (float32 where jax computes with x64 disabled) against the float64 oracle, per configuration.

SURVEY.md section 8c asks for this number next to the device path's fp32 bound: how far is the reference's OWN float32
arithmetic from its float64 arithmetic on the same inputs?  ``oracle.atm_oracle`` restates the reference with
``f32=True`` casts exactly where the reference hands arrays to jax (coords/transforms.py:10-29 and the three
RegularGridInterpolator call sites); parity of that variant is NOT pinned against jax itself (jax is absent here).

Both variants are evaluated on the SAME set-up geometry and the SAME generated layers: the footprint optimisation of the
set-up (atmosphere.py:184-194) is chaotic in its inputs, so feeding it float32 pointing changes the grid sizes, hence the
number of random draws and the whole realisation -- that is a different sky, not a rounding error (it shows up as ~0.1-0.3
"relative error" on the constant-elevation scans when the two variants are simply run with the same seed).

Runs on the CPU (no GPU, no reference import):  python scripts/fp32_asrun_table.py [--out profiles/r02_fp32_asrun_table.md]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from common import load_case  # noqa: E402
from maria_b200 import configs  # noqa: E402
from oracle import atm_oracle as O  # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


def fluct(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    m = b.mean(axis=1, keepdims=True)
    return float(np.max(np.abs((a - m) - (b - m))) / np.max(np.abs(b - m)))


def as_run_pair(case, seed):
    """(float64 result, as-run float32 result, float64 arithmetic stored as float32) on one geometry and one realisation."""
    np.random.seed(seed)
    rec = {}
    r64 = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random,
                     dtype=np.float64, record=rec)
    atm, dets = r64["atmosphere"], case["instrument"].dets
    geo = {k: dict(cross_section=p.cross_section, extrusion=p.extrusion, A=p.A, B=p.B, transform=p.transform, vx=p.vx, vy=p.vy,
                   layer_index=p.layer_index, labels=p.labels, extrusion_res=p.extrusion_res) for k, p in atm.processes.items()}
    obs32 = O.Observation(case["instrument"], case["plan"], case["site"], f32=True)
    atm32 = O.atmosphere_from_geometry(case["weather"], case["spectrum"], atm.layers, atm.timestep, atm.boresight.t,
                                       atm.boresight.az, atm.boresight.el, dets.offsets, geo, f32=True)
    for k, p in atm32.processes.items():  # the realisation of the float64 run: nothing is drawn again
        values = atm.processes[k].values
        p.values = values
        p.run = (lambda v: (lambda rng, record=None: v))(values)
    O.simulate_atmosphere(obs32, atm32, None)
    loading = O.compute_atmospheric_loading(obs32, atm32, dtype=np.float32)
    loading *= rec["gain_error"][:, None].astype(np.float32)  # simulation.py:243-247
    tod32 = O.tod_to_K_RJ(loading, obs32, atm32, r64["metadata"], out_dtype=np.float32)
    stages = dict(zenith_scaled_pwv=rel(atm32.zenith_scaled_pwv, atm.zenith_scaled_pwv),
                  az=float(np.max(np.abs(np.asarray(atm32.coords.az, dtype=np.float64) - atm.coords.az))),
                  el=float(np.max(np.abs(np.asarray(atm32.coords.el, dtype=np.float64) - atm.coords.el))))
    return r64["tod_K_RJ"], tod32, r64["tod_K_RJ"].astype(np.float32), stages


CASES = [
    ("golden m2_daisy", lambda: load_case("m2_daisy"), None),
    ("golden 1deg_ces", lambda: load_case("1deg_ces"), None),
    ("C1 (MUSTANG-2 daisy, full size)", lambda: configs.CONFIGS["c1"](), 17),
    ("C2 (ACT-like CES, 24 / array, 120 s)", lambda: configs.CONFIGS["c2"](n_per_array=24, duration=120.0), 17),
    ("C3 (3 bands, 180 det, 200 s, pwv 0.5)", lambda: configs.CONFIGS["c3"](n_det=180, duration=200.0, pwv=0.5), 17),
    ("C4 geometry (AtLAST-like, 64 det, 600 s)", lambda: configs.c4_atlast(n_det=64, duration=600.0, sample_rate=50.0), 17),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    lines = ["| configuration | as-run fp32 vs fp64: total | fluctuating part | zenith_scaled_pwv | az / el [rad] | storage only (fp64 -> fp32) |",
             "|---|---|---|---|---|---|"]
    for name, make, seed in CASES:
        case = make()
        ref, tod32, stored, st = as_run_pair(case, case.get("seed", seed) if seed is None else seed)
        line = (f"| {name} | {rel(tod32, ref):.2e} | {fluct(tod32, ref):.2e} | {st['zenith_scaled_pwv']:.2e} | "
                f"{st['az']:.1e} / {st['el']:.1e} | {rel(stored, ref):.1e} |")
        print(line, flush=True)
        lines.append(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
