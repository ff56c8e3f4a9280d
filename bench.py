#!/usr/bin/env python
"""Benchmark of the atmospheric TOD synthesis path (BASELINE.json metric: detector-samples/s, % of HBM roofline).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --steps K --warmup W    # the reference's CPU algorithm (oracle port) on host cores

A step = one full pass of the hot path over one observation: Philox noise -> AR recurrence of all layers -> beam
smoothing -> line-of-sight sampling / emission lookup (loading kernel) -> cubic upsample / gain / pW->K_RJ (evaluation
kernel), producing the (n_det, n_t) float32 TOD.  Workload at N = 1: configs[3], the 100k-detector AtLAST-scale 1-hour
raster the metric is quoted on (72 GB of TOD, fits one B200).  At N > 1 every rank synthesises its own 100k-detector shard
of an N x 100k-detector focal plane (weak scaling, no data-path collective: the layers are regenerated from the same
seed); the same line carries a "strong" block (the FIXED 100k-detector focal plane split over the N ranks, layers
generated once and exchanged with NCCL broadcasts) and a "parity_probe" (shards of all ranks against rank 0's full run).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c4", choices=["c1", "c2", "c3", "c4", "c5"],
                    help="BASELINE.json configuration; c4 (default) is the one the metric is quoted on, c5 is its six-hour "
                         "version with the 100k detectors split over the ranks")
    ap.add_argument("--n-det", type=int, default=100_000, help="detectors per GPU (c4) / in total (c5)")
    ap.add_argument("--duration", type=float, default=None, help="scan length [s] (default: the configuration's)")
    ap.add_argument("--sample-rate", type=float, default=50.0)
    ap.add_argument("--cpu-dets", type=int, default=256, help="detectors in the bounded CPU-baseline sample")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the sequential / strong / parity-probe blocks")
    ap.add_argument("--depth", type=int, default=None,
                    help="realisations whose AR chains are in flight underneath the synthesis (default 3; c5: as many as the "
                         "~300 ms six-hour chains need to hide behind the per-rank synthesis)")
    ap.add_argument("--chunk-mib", type=int, default=None, help="scratch per detector chunk of the two-kernel synthesis [MiB]")
    ap.add_argument("--chunk-serial", action="store_true", help="run the detector chunks of one synthesis on one stream")
    ap.add_argument("--kernel-times", action="store_true", help="bracket every launch of the two synthesis kernels with events")
    ap.add_argument("--nccl-max-ctas", type=int, default=2, help="CTAs NCCL may use per collective (the layer exchange)")
    ap.add_argument("--no-overlap", action="store_true",
                    help="generate and synthesise strictly one after the other (default: the AR chains of realisation i+1 run "
                         "underneath the synthesis kernels of realisation i)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            self.path = tempfile.NamedTemporaryFile(delete=False, suffix=".csv").name
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def build_config(args, world):
    """The synthetic inputs of the chosen BASELINE.json configuration.  c4: ``world`` sub-arrays of ``n_det`` detectors (weak
    scaling, one per rank); c5: ONE ``n_det``-detector focal plane, six hours, split over the ranks; c1-c3 at full size."""
    from maria_b200 import configs

    if args.config == "c4":
        return configs.c4_atlast(n_det=args.n_det, duration=args.duration or 3600.0, sample_rate=args.sample_rate, n_groups=world)
    if args.config == "c5":
        return configs.c5_long(n_det=args.n_det, duration=args.duration or 21600.0, sample_rate=args.sample_rate)
    kw = {} if args.duration is None else {"duration": args.duration}
    return configs.CONFIGS[args.config](**kw)


def workload_name(args, cfg, n_det_rank):
    n_t = len(cfg["plan"].time)
    rate = 1.0 / float(np.mean(np.diff(cfg["plan"].time[:100])))
    if args.config == "c4":
        return (f"C4 AtLAST-scale: {n_det_rank} detectors/GPU, 2 bands, 8 turbulent layers, raster {n_t / rate:.0f} s at "
                f"{rate:.0f} Hz ({n_t} samples/detector), units K_RJ")
    return (f"{cfg['name']}: {cfg['instrument'].dets.n} detectors ({n_det_rank} on this rank's GPU), {n_t} samples/detector at "
            f"{rate:.0f} Hz, units K_RJ")


# ---------------------------------------------------------------------------------------------------------------
# CPU baseline: the oracle (a port of the reference's numpy / scipy algorithm) on a bounded detector sample
# ---------------------------------------------------------------------------------------------------------------
def host_geometry(cfg, bind):
    """Run this repo's host set-up (same layer table / stencil / A, B / transform the GPU arm uses)."""
    import maria_b200 as M

    np.random.seed(1234)
    obs = M.Observation(cfg["instrument"], cfg["plan"], cfg["site"])
    atm = M.Atmosphere(weather=cfg["weather"], spectrum=cfg["spectrum"], noise="philox", seed=1234, precision="fp32")
    atm.initialize(obs, bind=bind)
    return obs, atm


def cpu_sample_runner(cfg, atm, n_sub):
    """Returns (run, n_samples): ``run(seed)`` executes stages a5-a14 of the oracle for ``n_sub`` detectors spread over the
    focal plane and both bands, over the full scan (all layers at full size), and returns the K_RJ TOD."""
    from maria_b200 import inputs as I
    from oracle import atm_oracle as O

    dets = cfg["instrument"].dets
    sel = np.unique(np.linspace(0, dets.n - 1, min(n_sub, dets.n)).astype(int))
    sub = I.Array(dets.xi[sel], dets.eta[sel], dets.band_name[sel], [b for b in dets.bands if b.name in set(dets.band_name[sel])],
                  dets.gamma[sel], dets.primary_size[sel])
    inst = I.Instrument(sub, name=cfg["instrument"].name)
    geo = {k: dict(cross_section=p.cross_section, extrusion=p.extrusion, A=p.A, B=p.B, transform=p.transform, vx=p.vx, vy=p.vy,
                   layer_index=p.layer_index, labels=p.labels, extrusion_res=p.extrusion_res) for k, p in atm.processes.items()}
    o_obs = O.Observation(inst, cfg["plan"], cfg["site"])
    o_atm = O.atmosphere_from_geometry(cfg["weather"], cfg["spectrum"], atm.layers, atm.timestep, atm.boresight.t,
                                       atm.boresight.az, atm.boresight.el, sub.offsets, geo)

    def run(seed):
        rng = np.random.default_rng(seed)
        loading, meta = O.run_obs(o_obs, o_atm, rng, dtype=np.float32)
        return O.tod_to_K_RJ(loading, o_obs, o_atm, meta, out_dtype=np.float32)

    return run, len(sel) * len(cfg["plan"].time)


def time_cpu(run, steps, warmup):
    """Wall and process-CPU seconds per step: their ratio is the number of cores the port actually keeps busy."""
    for w in range(warmup):
        run(100 + w)
    t0, c0 = time.perf_counter(), time.process_time()
    for s in range(steps):
        run(200 + s)
    return (time.perf_counter() - t0) / steps, (time.process_time() - c0) / steps


def cpu_threads_available():
    try:
        from threadpoolctl import threadpool_info

        return max([p.get("num_threads", 1) for p in threadpool_info()] + [1])
    except Exception:
        return os.cpu_count() or 1


def cpu_baseline(cfg, atm, n_dets, steps=1, warmup=0, small=32):
    """The oracle on ``n_dets`` detectors (the reported value) and on ``small`` detectors: the two timings separate the fixed
    cost per realisation (layer generation: the sequential AR recurrence, independent of the detector count) from the cost
    per detector-sample, so that the rate a full-size CPU run would approach is reported next to the measured one."""
    n_t = len(cfg["plan"].time)
    run_s, ns_small = cpu_sample_runner(cfg, atm, small)
    wall_s, _ = time_cpu(run_s, 1, 0)
    run, n_samples = cpu_sample_runner(cfg, atm, n_dets)
    wall, cpu = time_cpu(run, steps, warmup)
    nd, ns = n_samples // n_t, ns_small // n_t
    per_det = max((wall - wall_s) / max(nd - ns, 1), 1e-12)
    fixed = max(wall - per_det * nd, 0.0)
    return {"value": n_samples / wall, "unit": "detector-samples/s", "cores": max(1, int(round(cpu / wall))), "kind": "port",
            "sample": f"{nd} of {cfg['instrument'].dets.n} detectors (evenly spaced over the focal plane and the bands) x {n_t} "
                      f"samples, all layers at full size, {wall:.1f} s of CPU work per step",
            "threads_available": cpu_threads_available(), "host_cpus": os.cpu_count(), "cpu_seconds_per_wall_second": cpu / wall,
            "fixed_seconds_per_realisation": fixed, "seconds_per_detector": per_det,
            "asymptotic_value": n_t / per_det, "asymptotic_note": "detector-samples/s a full-size CPU run approaches: the per-"
            f"detector slope between the {ns}- and the {nd}-detector samples; the fixed part is the layer generation"}, wall


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = build_config(args, 1)
    _, atm = host_geometry(cfg, bind=False)
    # K + W steps must end within a few minutes: the sample shrinks with the number of steps (~0.075 s per detector + 2 s)
    total = args.steps + min(args.warmup, 1)
    n_dets = args.cpu_dets if total <= 6 else (128 if total <= 12 else 64)
    if args.config in ("c1", "c2", "c3"):  # the reference's own CPU-runnable cases: the whole configuration, not a sample
        n_dets = cfg["instrument"].dets.n
    cpu, wall = cpu_baseline(cfg, atm, n_dets, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": "atmospheric TOD detector-samples/sec", "value": cpu["value"], "unit": "detector-samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": wall * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, cfg, args.n_det if args.config == "c4" else cfg["instrument"].dets.n // max(1, args.gpus)),
                   "sample": cpu["sample"]},
        "cpu_baseline": cpu,
        "e2e": {"value": cpu["value"], "unit": "detector-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# the CUDA arm
# ---------------------------------------------------------------------------------------------------------------
def make_sim(cfg, det_range, device, depth, exchange=False, np_seed=1234):
    import maria_b200 as M

    np.random.seed(np_seed)  # identical host set-up on every rank
    return M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=np.float32, det_range=det_range,
                        atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], noise="philox", seed=1234,
                                               device=device, pipeline_depth=depth, layer_exchange=exchange))


class Stepper:
    """One step = one fresh realisation end to end: Philox noise -> B z GEMM -> AR chains (one stream per layer) -> beam
    smoothing -> loading + evaluation kernels.  With overlap the chains of the NEXT realisations are enqueued before this
    realisation's synthesis and run underneath it; every step still contains exactly one chain set, one smoothing pass and
    one synthesis (the Philox subsequence advances with every realisation: nothing can be cached)."""

    def __init__(self, torch, sim, out, overlap):
        self.torch, self.sim, self.out, self.overlap = torch, sim, out, overlap
        self.atm = sim.obs_list[0].atmosphere
        self.det0, self.n_det = sim.det_range
        self.primed = False

    def prime(self):
        if self.overlap and not self.primed:
            for _ in range(self.atm.pipeline_depth):
                self.atm.generate_async()
        self.primed = True

    def step(self, ev=None):
        atm = self.atm
        if ev:
            ev[0].record()
        if self.overlap:
            atm.finish_generate()  # beam-smooth realisation i (its chains were started `depth` steps ago)
            atm.generate_async()   # chains of realisation i + depth
        else:
            atm.generate()
        if ev:
            ev[1].record()
        atm.synthesize(self.det0, self.n_det, units="K_RJ", apply_gain=True, out=self.out)
        if ev:
            ev[2].record()

    def finish(self):
        self.atm.join_streams()  # the last chains started end inside the timed region

    def drain(self):
        self.atm.drain()
        self.atm.join_streams()
        self.primed = False


def timed_run(torch, dist, world, device, stepper, steps, warmup, clocks=False):
    """W untimed steps, then exactly K steps between barrier + synchronize on both sides; CUDA events, max over ranks."""

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stepper.prime()
    for _ in range(warmup):
        stepper.step()
    barrier()
    status0 = stepper.atm.check_flags()
    launches0 = stepper.atm.ctx.launch_count
    sampler = ClockSampler(device) if clocks else None
    if sampler:
        sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    per = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
    barrier()
    ev[0].record()
    for i in range(steps):
        stepper.step(per[i])
    stepper.finish()
    ev[1].record()
    barrier()
    res = {"clocks": sampler.stop() if sampler else None, "launches": stepper.atm.ctx.launch_count - launches0}
    total_ms = ev[0].elapsed_time(ev[1])
    gen_ms = float(np.mean([a.elapsed_time(b) for a, b, _ in per]))
    synth_ms = float(np.mean([b.elapsed_time(c) for _, b, c in per]))
    res["status"] = max(status0, stepper.atm.check_flags())
    if world > 1:
        t = torch.tensor([total_ms, gen_ms, synth_ms], device=f"cuda:{device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, gen_ms, synth_ms = (float(x) for x in t)
    res.update(ms_per_step=total_ms / steps, gen_ms=gen_ms, synth_ms=synth_ms)
    return res


def kernel_times(ctx):
    a, b, nch = C.c_double(), C.c_double(), C.c_int64()
    ctx.check(ctx.lib.matm_get_timing(ctx.h, C.byref(a), C.byref(b), C.byref(nch)))
    return a.value, b.value, nch.value


def sequential_block(torch, dist, world, device, sim, out, steps=3):
    """The strictly sequential step (one realisation at a time, nothing overlapped), with the detector chunks of the synthesis
    on one stream in one piece, so that the per-kernel event times are clean: what a single Simulation.run() of one plan
    pays, and the like-for-like numbers behind the pipelined headline."""
    atm = sim.obs_list[0].atmosphere
    ctx = atm.ctx
    stepper = Stepper(torch, sim, out, overlap=False)
    ctx.check(ctx.lib.matm_set_tuning(ctx.h, 60000, 0))
    stepper.step()  # allocates the single-piece scratch
    ctx.check(ctx.lib.matm_set_timing(ctx.h, 1))
    r = timed_run(torch, dist, world, device, stepper, steps, 1)
    a, b, nch = kernel_times(ctx)
    ctx.check(ctx.lib.matm_set_timing(ctx.h, 0))
    ctx.check(ctx.lib.matm_set_tuning(ctx.h, 1024, 1))
    n_calls = max(nch, 1)
    return {"ms_per_step": r["ms_per_step"], "chains_ms": r["gen_ms"], "synth_ms": r["synth_ms"],
            "loading_f32_kernel_ms": a / n_calls, "synth_eval_f32_kernel_ms": b / n_calls,
            "note": "generate() then synthesize(), one realisation at a time; chains_ms = noise + B z GEMM + AR chains + smoothing"}


def shard_checksum(torch, x):
    """64-bit, position-sensitive checksum of the bits of a float32 shard (one int64 per detector row)."""
    v = x.contiguous().view(torch.int32).to(torch.int64)
    w = (torch.arange(v.shape[1], device=v.device, dtype=torch.int64) % 8191) + 1
    return (v * w).sum(dim=1)


def parity_probe(torch, dist, rank, world, device, exchange=True):
    """A fixed 4096-detector focal plane (C4 geometry, 4 minutes): every rank synthesises its detector shard through the
    same code path as the strong-scaling block (layers generated once, exchanged), rank 0 also runs the whole focal plane
    on its own; the per-detector 64-bit checksums of all shards must equal those of the full run."""
    from maria_b200 import configs, sharding

    cfg = configs.c4_atlast(n_det=4096, duration=240.0, sample_rate=50.0)
    n = cfg["instrument"].dets.n
    det0, cnt = sharding.detector_range(n, rank, world)
    sim = make_sim(cfg, (det0, cnt), device, 1, exchange=exchange and world > 1, np_seed=77)
    tod = sim.run(units="K_RJ")[0]
    mine = shard_checksum(torch, tod.device_data["atmosphere"])
    sums = torch.zeros(n, dtype=torch.int64, device=f"cuda:{device}")
    sums[det0:det0 + cnt] = mine
    if world > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    ok = None
    if rank == 0:
        full = make_sim(cfg, None, device, 1, np_seed=77).run(units="K_RJ")[0]
        ok = bool(torch.equal(shard_checksum(torch, full.device_data["atmosphere"]), sums))
    del sim, tod
    torch.cuda.empty_cache()
    return {"bit_identical": ok, "n_det": n, "n_t": len(cfg["plan"].time), "shards": world,
            "layer_exchange": bool(exchange and world > 1), "checksum": "per-detector int64 sum of weighted float32 bit patterns"}


def strong_block(args, torch, dist, rank, world, device, steps, n1_value):
    """The FIXED 100k-detector focal plane of N = 1 split over the ranks (sharding.detector_range); rank k % N generates
    process k (noise + B z + AR chain + smoothing) and its smoothed layers are broadcast with NCCL on the process's own
    stream, underneath the previous realisation's synthesis (SURVEY.md section 8e option B)."""
    from maria_b200 import configs, sharding

    # the N = 1 focal plane (n_det detectors = n_det / 2 dichroic pixels in the same 2-degree field), its detectors emitted
    # in `world` blocks that each hold both bands of their pixels (maria sorts detectors by band inside an array; a shard
    # that kept only one band of every pixel would repeat every line of sight on two ranks)
    per = (args.n_det // world) // 64 * 64
    cfg = configs.c4_atlast(n_det=per, duration=args.duration or 3600.0, sample_rate=args.sample_rate, n_groups=world)
    n, n_t = cfg["instrument"].dets.n, len(cfg["plan"].time)
    det0, cnt = rank * per, per
    # a realisation cannot finish faster than its longest AR chain (~22 ms + noise and smoothing): enough realisations in
    # flight to cover it behind the 1 / world share of a 22 ms synthesis
    depth = int(min(16, max(args.depth, np.ceil(26.0 * world / 22.0) + 1)))
    sim = make_sim(cfg, (det0, cnt), device, depth, exchange=True)
    atm = sim.obs_list[0].atmosphere
    atm.set_gain(np.exp(cfg["instrument"].dets.gain_error * np.random.standard_normal(size=n)))
    out = atm.ctx.empty((cnt, n_t))
    stepper = Stepper(torch, sim, out, overlap=True)
    r = timed_run(torch, dist, world, device, stepper, steps, max(3, depth))
    stepper.drain()
    value = n * n_t / (r["ms_per_step"] * 1e-3)
    layer_bytes = sum(4 * p.n_extrusion * p.n_cross_section for p in atm.processes.values())
    res = {"value": value, "unit": "detector-samples/s", "ms_per_step": r["ms_per_step"], "synth_ms": r["synth_ms"],
           "n_det_total": n, "n_det_per_gpu": cnt, "steps": steps, "pipeline_depth": depth, "layer_exchange": "one torch.distributed broadcast (NCCL) per "
           "layer from rank (process index mod N), on the process's stream", "nccl_bytes_per_step_per_rank": layer_bytes,
           "generate_and_exchange_ms": r["gen_ms"], "status": r["status"],
           "efficiency_vs_n1": (value / (world * n1_value)) if n1_value else None,
           "n1_value": n1_value, "n1_source": "profiles/n1_c4.json (this round's N = 1 line of the same build)" if n1_value else None}
    del out, sim
    torch.cuda.empty_cache()
    return res


def d2h_ceiling(torch, dist, world, device, gib=4):
    """What the box gives a plain pinned cudaMemcpyAsync device -> host, all N ranks at once (the e2e figure's ceiling)."""
    n = gib << 28  # float32 elements
    try:
        host = torch.empty(n, dtype=torch.float32, pin_memory=True)
    except Exception:
        return None
    dev = torch.empty(n, dtype=torch.float32, device=f"cuda:{device}")
    host.copy_(dev, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(2):
        host.copy_(dev, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=f"cuda:{device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t[0])
    del host, dev
    return world * 2 * n * 4 / dt / 1e9


def run_e2e(args, torch, dist, world, local, sim, cfg, gain, barrier):
    """The same step through the public API with host buffers: every step uploads the scan (boresight t/az/el), the
    detector table and the gains from host memory, regenerates the layers and delivers the TOD shard into pinned host
    memory (matm_synthesize_host: chunked, copies overlapped with the kernels)."""
    obs = sim.obs_list[0]
    atm = obs.atmosphere
    det0, n_det = sim.det_range
    n_t = len(cfg["plan"].time)
    # the shard is delivered through a bounded ring of pinned host memory (<= 8 GiB), chunk after chunk, as a consumer that
    # cannot hold N x 72 GB of pinned pages would take it; every byte of the TOD still crosses PCIe inside the timed region
    ring_bytes = min(8 << 30, (32 << 30) // world)  # all ranks share one host: keep the pinned total at <= 32 GiB
    rows = max(32, min(n_det, ring_bytes // (4 * n_t) // 32 * 32))
    try:
        host_out = torch.empty((rows, n_t), dtype=torch.float32, pin_memory=True)
        pinned = True
    except Exception:
        host_out = torch.empty((rows, n_t), dtype=torch.float32)
        pinned = False
    dets = cfg["instrument"].dets
    h2d = 8 * (3 * n_t + 3 * len(atm.boresight.t)) + dets.n * (16 + 16 + 4 + 8 + 8)
    d2h = n_det * n_t * 4
    checksum = [0.0]

    def step(i):
        atm._bind_scan_and_detectors(obs, gain)  # H2D: boresight at both rates, offsets, bands, Mueller, gains
        atm.set_layers()
        atm.generate()
        acc = 0.0
        for off in range(0, n_det, rows):
            nd = min(rows, n_det - off)
            atm.synthesize_host(host_out[:nd], det0 + off, nd, units="K_RJ", apply_gain=True)
            acc += float(host_out[0, :: max(1, n_t // 64)].double().mean())  # the consumer touches the chunk
        checksum[0] = acc / len(range(0, n_det, rows))

    atm.drain()  # realisations still in flight from the pipelined loop
    step(0)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.e2e_steps):
        step(1 + i)
    barrier()
    dt = (time.perf_counter() - t0) / args.e2e_steps
    if world > 1:
        t = torch.tensor([dt], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t[0])
    del host_out
    ceiling = d2h_ceiling(torch, dist, world, local)
    gbs = world * d2h / dt / 1e9
    return {"value": world * n_det * n_t / dt, "unit": "detector-samples/s", "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": dt * 1e3, "steps": args.e2e_steps, "pinned_host_output": pinned,
            "host_ring_rows": rows, "mean_K_RJ": checksum[0], "d2h_gbs": gbs, "d2h_ceiling_gbs": ceiling,
            "frac_of_d2h_ceiling": (gbs / ceiling) if ceiling else None,
            "ceiling_note": "plain pinned cudaMemcpyAsync device -> host by all ranks at once, measured right after"}


def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.depth is None:
        # C5's longest layer is a ~140 ms chain (223 k dependent steps); one rank synthesises its shard in ~180 / N ms
        args.depth = 3 if args.config != "c5" else int(min(16, max(3, np.ceil(160.0 * world / 180.0) + 2)))

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the CUDA path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        # the only data-path collective is a 26 MB layer exchange per realisation: two CTAs per collective are plenty, and
        # NCCL kernels waiting for a peer must not hold the SMs the synthesis runs on
        opts = dist.ProcessGroupNCCL.Options()
        try:
            opts.config.min_ctas, opts.config.max_ctas = 1, args.nccl_max_ctas
        except Exception:
            pass
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"), pg_options=opts)
    from maria_b200 import sharding

    cfg = build_config(args, world)
    dets = cfg["instrument"].dets
    n_t = len(cfg["plan"].time)
    weak = args.config == "c4"
    det_range = (rank * args.n_det, args.n_det) if weak else sharding.detector_range(dets.n, rank, world)
    overlap = not args.no_overlap
    t0 = time.perf_counter()
    sim = make_sim(cfg, det_range, local, max(1, args.depth) if overlap else 1, exchange=(not weak) and world > 1)
    setup_s = time.perf_counter() - t0
    atm = sim.obs_list[0].atmosphere
    ctx = atm.ctx
    det0, n_det = sim.det_range
    gain = np.exp(dets.gain_error * np.random.standard_normal(size=dets.n))  # simulation.py:239-241
    atm.set_gain(gain)
    out = ctx.empty((n_det, n_t))  # the TOD shard, resident in HBM (72 GB at full size)
    if args.chunk_mib is not None or args.chunk_serial:
        ctx.check(ctx.lib.matm_set_tuning(ctx.h, args.chunk_mib or 1024, 0 if args.chunk_serial else 1))
    if args.kernel_times:
        ctx.check(ctx.lib.matm_set_timing(ctx.h, 1))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stepper = Stepper(torch, sim, out, overlap)
    n_warm = max(args.warmup, 3)
    stepper.prime()
    for _ in range(n_warm):
        stepper.step()
    barrier()
    if args.kernel_times:
        kernel_times(ctx)  # reset
    r = timed_run(torch, dist, world, local, stepper, args.steps, 0, clocks=True)
    per_kernel = None
    if args.kernel_times:
        a, b, nch = kernel_times(ctx)
        ctx.check(ctx.lib.matm_set_timing(ctx.h, 0))
        per_kernel = {"loading_f32_kernel_ms": a / args.steps, "synth_eval_f32_kernel_ms": b / args.steps,
                      "chunks_per_step": nch // args.steps, "note": "sums over chunks that overlap in wall time"}
    ms_per_step, synth_ms = r["ms_per_step"], r["synth_ms"]
    n_total = dets.n if not weak else world * n_det
    value = n_total * n_t / (ms_per_step * 1e-3)

    # ---- roofline of the synthesis (loading_f32_kernel + synth_eval_f32_kernel of every chunk of one realisation) ----
    layer_bytes = sum(4 * p.n_extrusion * p.n_cross_section for p in atm.processes.values())
    n_atm = len(atm.boresight.t)
    alg_bytes = 4 * n_det * n_t + layer_bytes + 16 * n_atm + 8 * n_t + 16 * n_det
    peak, peak_src = measured_peak_gbs()
    achieved = alg_bytes / (synth_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None  # DRAM bytes of one synthesis from the committed ncu captures of this very workload
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)["synthesis_f32"]
        if t["n_det"] == n_det and t["n_t"] == n_t:
            traffic, traffic_src = t["traffic"], t["source"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "loading_f32_kernel + synth_eval_f32_kernel (one synthesis)", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": synth_ms,
                "step_ms_outside_kernel": ms_per_step - synth_ms, "per_kernel": per_kernel,
                "note": "kernel_ms = CUDA-event time of one whole synthesis (both kernels, all detector chunks) inside the "
                        "pipelined step, i.e. with the AR chains of the next realisations sharing the SMs"}

    extras = {}
    if not args.no_extras:
        stepper.drain()
        seq = sequential_block(torch, dist, world, local, sim, out)
        seq["roofline_frac"] = alg_bytes / (seq["synth_ms"] * 1e-3) / 1e9 / peak
        seq["eval_kernel_frac"] = 4 * n_det * n_t / (seq["synth_eval_f32_kernel_ms"] * 1e-3) / 1e9 / peak
        extras["sequential"] = seq

    # ---- end to end through the public API with HOST buffers ----
    e2e = None
    del out
    stepper.out = None
    torch.cuda.empty_cache()
    if not args.no_e2e:
        e2e = run_e2e(args, torch, dist, world, local, sim, cfg, gain, barrier)

    if not args.no_extras and world > 1:
        atm.drain()
        del stepper, sim
        torch.cuda.empty_cache()
        extras["parity_probe"] = parity_probe(torch, dist, rank, world, local)
        if weak:
            n1 = None
            try:
                with open(os.path.join(ROOT, "profiles", "n1_c4.json")) as f:
                    j = json.load(f)
                if j["n_det"] == args.n_det and j["n_t"] == n_t:
                    n1 = float(j["value"])
            except Exception:
                pass
            extras["strong"] = strong_block(args, torch, dist, rank, world, local, max(args.steps, 10), n1)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        torch.cuda.empty_cache()
        cpu, _ = cpu_baseline(cfg, atm, cfg["instrument"].dets.n if args.config in ("c1", "c2", "c3") else args.cpu_dets)
    if rank == 0:
        depth = atm.pipeline_depth
        line = {
            "metric": "atmospheric TOD detector-samples/sec", "value": value, "unit": "detector-samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": n_warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args, cfg, n_det), "n_det_per_gpu": n_det, "n_t": n_t, "n_t_atm": n_atm,
                       "layers": [[p.n_extrusion, p.n_cross_section] for p in atm.processes.values()],
                       "l2": "output (n_det*n_t*4 B per step) is far larger than L2; layers are regenerated every step",
                       "rng": "philox (device), a new subsequence every realisation", "setup_s": setup_s, "status": r["status"],
                       "pipeline": (f"AR chains of realisations i+1..i+{depth} run underneath the synthesis of realisation i; the "
                                    "timed region holds exactly K chain sets, K smoothing passes and K syntheses") if overlap
                       else "sequential", "pipeline_depth": depth if overlap else 0},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": r["launches"], "clocks": r["clocks"],
        }
        line.update(extras)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
