"""Per-phase clocks of the AR chain's step (library built with -DMATM_AR_TRACE, scripts/build of variants/libmaria_atm_artrace.so):
for warps 0 and 3 of every CTA of the longest C4 chain, 128 steps, SM clocks at: before the LATE wait, after it, after the LATE
dot product, after the butterfly, after the sends, after the EARLY part.  Prints the mean length of every phase in clocks."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import maria_b200 as M  # noqa: E402
from maria_b200 import configs  # noqa: E402


def main():
    which = [int(a) for a in sys.argv[1:]] or [7, 4, 0]
    cfg = configs.c4_atlast(n_det=2048, duration=3600.0, sample_rate=50.0)
    np.random.seed(1234)
    sim = M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=np.float32,
                       atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], noise="philox", seed=1, geometry="fast",
                                              pipeline_depth=1))
    atm = sim.obs_list[0].atmosphere
    ctx = atm.ctx
    torch.cuda.synchronize()
    fn = ctx.lib.matm_debug_ar_trace
    fn.restype, fn.argtypes = C.c_int, [C.c_void_p]
    for k in which:
        p = atm.processes[k]
        dv = p._dev
        n_ext, n_cs = p.n_extrusion, p.n_cross_section
        buf, zz = dv["buffer"][0], dv["z"][0]
        tail = buf[2 * n_ext:]
        st = ctx.stream_ptr()
        ctx.check(ctx.lib.matm_fill_normal(ctx.h, tail.data_ptr(), tail.numel(), 7, 2 * k, st))
        ctx.check(ctx.lib.matm_fill_normal(ctx.h, zz.data_ptr(), zz.numel(), 7, 2 * k + 1, st))
        for _ in range(2):
            ctx.check(ctx.lib.matm_ar_generate(ctx.h, n_ext, n_cs, p.n_sample, p.n_near, dv["A"].data_ptr(), dv["B"].data_ptr(),
                                               dv["esi"].data_ptr(), dv["csi"].data_ptr(), buf.data_ptr(), zz.data_ptr(), st))
        torch.cuda.synchronize()
        tr = np.zeros((16, 128, 6), dtype=np.int64)
        assert fn(tr.ctypes.data) == 0
        n_cta = -(-n_cs // 32) if n_cs > 32 else 1
        tr[2 * n_cta:] = 0  # slots of larger clusters traced earlier
        print(f"process {k}: n_cs {n_cs} n_sample {p.n_sample}")
        names = ["wait LATE", "LATE dot", "butterfly", "send", "EARLY+rest"]
        for slot in range(16):
            t = tr[slot]
            if not t[:, 0].any():
                continue
            t = t[::-1]  # stored by (i & 127): increasing time = decreasing i
            order = np.argsort(t[:, 0])
            t = t[order]
            period = np.diff(t[:, 1])
            good = (period > 0) & (period < 20000)
            seg = np.diff(t, axis=1)
            print(f"  cta {slot // 2} warp {0 if slot % 2 == 0 else 3}: period {period[good].mean():7.1f} clk | " +
                  " | ".join(f"{n} {seg[:, j].mean():6.1f}" for j, n in enumerate(names)))


if __name__ == "__main__":
    main()
