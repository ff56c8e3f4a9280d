"""Latency of the AR chain (matm_ar_generate = bz_gemm + chain kernel) per process of the C4 geometry, alone on the GPU,
with the planner's choice of lanes per row / cluster size / threads per CTA.  Prints ms and microseconds per step.
python scripts/chain_bench.py [--config c4|c5]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import maria_b200 as M  # noqa: E402
from maria_b200 import configs  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c4")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--load", action="store_true", help="time the chains underneath a copy loop that saturates HBM")
    args = ap.parse_args()
    dur = 3600.0 if args.config == "c4" else 6 * 3600.0
    cfg = configs.c4_atlast(n_det=2048, duration=dur, sample_rate=50.0)
    np.random.seed(1234)
    sim = M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=np.float32,
                       atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], noise="philox", seed=1, geometry="fast",
                                              pipeline_depth=1))
    atm = sim.obs_list[0].atmosphere
    ctx = atm.ctx
    atm.drain() if hasattr(atm, "drain") else None
    torch.cuda.synchronize()
    ref = {}
    side = torch.cuda.Stream()
    if args.load:
        src = torch.empty(1 << 28, dtype=torch.float32, device="cuda").normal_()
        dst = torch.empty_like(src)
    for k, p in atm.processes.items():
        dv = p._dev
        n_ext, n_cs = p.n_extrusion, p.n_cross_section
        buf, zz = dv["buffer"][0], dv["z"][0]
        tail = buf[2 * n_ext:]
        st = ctx.stream_ptr()
        ctx.check(ctx.lib.matm_fill_normal(ctx.h, tail.data_ptr(), tail.numel(), 7, 2 * k, st))
        ctx.check(ctx.lib.matm_fill_normal(ctx.h, zz.data_ptr(), zz.numel(), 7, 2 * k + 1, st))
        for cl in (0,):
            best = 1e30
            for _ in range(args.reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if args.load:
                    side.wait_stream(torch.cuda.current_stream())
                    with torch.cuda.stream(side):
                        for _ in range(int(3 * 2 * n_ext * 1e-3 / 0.35) + 20):  # ~0.35 ms per 1 GiB copy: 3 us per chain step
                            dst.copy_(src)
                e0.record()
                ctx.check(ctx.lib.matm_ar_generate(ctx.h, n_ext, n_cs, p.n_sample, p.n_near, dv["A"].data_ptr(), dv["B"].data_ptr(),
                                                   dv["esi"].data_ptr(), dv["csi"].data_ptr(), buf.data_ptr(), zz.data_ptr(), st))
                e1.record()
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            vals = buf[:n_ext].clone()
            same = True if k not in ref else bool(torch.equal(vals, ref[k]))
            ref.setdefault(k, vals)
            print(f"process {k}: n_ext {n_ext} n_cs {n_cs} n_sample {p.n_sample}: {best:.3f} ms = {1e3 * best / (2 * n_ext):.3f} us/step"
                  f"  finite {bool(torch.isfinite(vals).all())} same_bits_as_first {same}", flush=True)


if __name__ == "__main__":
    main()
