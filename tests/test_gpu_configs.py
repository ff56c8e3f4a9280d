"""GPU parity tests on the BASELINE.json configurations (maria_b200/configs.py).

C1 runs at its full size end to end against the oracle.  C2 / C3 keep their scan, bands and layer geometry and thin the
focal plane so that the oracle finishes in seconds.  C4 keeps the full 1-hour time axis and the full-size layers
(n_ext up to 18 684, 37 k recurrence steps): the AR recurrence is checked against the oracle with identical host
noise, and the detector x time stages against the oracle fed with the device's own layers, for a subset of detectors
spread over the whole focal plane.  Size-independent properties (gain linearity, unit round trip, shard bit-equality,
determinism) are checked on the device at C4 geometry.

Tolerances: fp64 path 1e-10 relative; fp32 path 2e-6 on the total TOD (4e-6 on the samples extrapolated beyond the last
knot) and 5e-4 on its fluctuating part -- stated in tests/common.py.
"""
import numpy as np
import pytest

import maria_b200 as M
from common import check32, relmax
from maria_b200 import configs
from maria_b200 import inputs as I
from oracle import atm_oracle as O

pytestmark = pytest.mark.gpu

TOL64 = 1e-10


def _device(cfg, dtype, np_seed, **akw):
    np.random.seed(np_seed)
    sim = M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=dtype,
                       atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], **akw))
    return sim


def _oracle(cfg, seed):
    np.random.seed(seed)
    return O.simulate(cfg["instrument"], cfg["plan"], cfg["site"], cfg["weather"], cfg["spectrum"], np.random,
                      dtype=np.float64)


def _check32(dev, ref, sim=None):
    """common.check32 with the knot range of the simulation's atmosphere-rate axis (None: every sample counts as inside)."""
    if sim is None:
        return check32(dev, ref)
    obs = sim.obs_list[0]
    check32(dev, ref, obs.boresight.t, obs.atmosphere.boresight.t[-1])


@pytest.mark.parametrize("name, kwargs", [
    ("c1", {}),                                         # MUSTANG-2 daisy, full size: 217 detectors x 30 000 samples
    ("c2", dict(n_per_array=24, duration=120.0)),       # ACT-like CES, 3 bands x 2 polarisations, 200 Hz
    ("c3", dict(n_det=180, duration=200.0, pwv=0.5)),   # 3-band array, PWV through the weather override
    ("c3", dict(n_det=90, duration=100.0, pwv=2.0)),
])
def test_baseline_config_matches_oracle(cuda, name, kwargs):
    cfg = configs.CONFIGS[name](**kwargs)
    ref = _oracle(cfg, seed=17)["tod_K_RJ"]
    tod64 = _device(cfg, np.float64, 17, geometry="exact").run(units="K_RJ")[0].data["atmosphere"]
    assert tod64.shape == ref.shape and relmax(tod64, ref) < TOL64
    sim32 = _device(cfg, np.float32, 17, geometry="exact")
    tod32 = sim32.run(units="K_RJ")[0].data["atmosphere"]
    assert tod32.dtype == np.float32
    _check32(tod32, ref, sim32)


# ---- C4 geometry: full time axis, full-size layers ---------------------------------------------------------------
@pytest.fixture(scope="module")
def c4():
    return configs.c4_atlast(n_det=2048, duration=3600.0, sample_rate=50.0)


def _oracle_with_layers(cfg, atm, sel, layer_values):
    """Oracle stages a6-a14 for the detectors ``sel``, with the layer values given (the device's own realisation)."""
    dets = cfg["instrument"].dets
    sub = I.Array(dets.xi[sel], dets.eta[sel], dets.band_name[sel], [b for b in dets.bands if b.name in set(dets.band_name[sel])],
                  dets.gamma[sel], dets.primary_size[sel])
    # the beam smoothing uses the mean beam of the WHOLE array (atmosphere.py:328-339): keep it
    geo = {k: dict(cross_section=p.cross_section, extrusion=p.extrusion, A=p.A, B=p.B, transform=p.transform, vx=p.vx, vy=p.vy,
                   layer_index=p.layer_index, labels=p.labels, extrusion_res=p.extrusion_res) for k, p in atm.processes.items()}
    o_obs = O.Observation(I.Instrument(sub, name="subset"), cfg["plan"], cfg["site"])
    o_atm = O.atmosphere_from_geometry(cfg["weather"], cfg["spectrum"], atm.layers, atm.timestep, atm.boresight.t,
                                       atm.boresight.az, atm.boresight.el, sub.offsets, geo)
    for k, p in o_atm.processes.items():
        vals = np.array(layer_values[k])
        p.run = (lambda rng, record=None, _v=vals, _p=p: setattr(_p, "values", _v))

    class Zeros:  # gain_error = exp(g * 0) = 1
        def standard_normal(self, size=None):
            return np.zeros(size)

    full = cfg["instrument"]  # beam smoothing width from the full array
    o_atm_sim = o_atm.simulate_pwv
    o_atm.simulate_pwv = lambda instrument, rng, record=None: o_atm_sim(full, rng, record=record)
    loading, meta = O.run_obs(o_obs, o_atm, Zeros(), dtype=np.float64)
    return O.tod_to_K_RJ(loading, o_obs, o_atm, meta, out_dtype=np.float64)


def test_c4_full_length_recurrence_matches_oracle(cuda, c4):
    """a5 at C4's layer sizes: 2 x 18 684 dependent steps, identical host noise, float64: device chain == oracle chain."""
    sim = _device(c4, np.float64, 23, noise="host", geometry="fast")
    atm = sim.obs_list[0].atmosphere
    k = max(atm.processes, key=lambda q: atm.processes[q].n_extrusion * atm.processes[q].n_cross_section)
    p = atm.processes[k]
    assert p.n_extrusion > 15000 and p.n_cross_section > 100
    rng = np.random.default_rng(5)
    buffer0 = rng.standard_normal((3 * p.n_extrusion, p.n_cross_section))
    z = rng.standard_normal((2 * p.n_extrusion, p.n_cross_section))
    dev = p.run(rng=O.ReplayRNG([buffer0, z])).cpu().numpy()
    q = O.AutoregressiveProcess(p.cross_section, p.extrusion, {}, jitter=1e-8)
    q.A, q.B = p.A, p.B
    ref = q.run(O.ReplayRNG([buffer0] + [z[i] for i in range(len(z))]))
    assert relmax(dev, ref) < TOL64


def test_c4_full_time_axis_matches_oracle(cuda, c4):
    """a6-a14 on the 1-hour raster (7950 knots -> 180 000 samples) with the production Philox layers: the float32 and the
    float64 device paths against the oracle fed with the device's layers, detectors spread over the 2-degree field."""
    n = c4["instrument"].dets.n
    sel = np.unique(np.linspace(0, n - 1, 24).astype(int))
    out = {}
    for dtype in (np.float64, np.float32):
        sim = _device(c4, dtype, 23, noise="philox", seed=77, geometry="fast")
        obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
        atm.simulate_pwv()
        atm.set_gain(np.ones(n))
        tod = sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=True)
        atm.check_flags()
        out[dtype] = tod[sel].cpu().numpy()
        if dtype is np.float64:
            layers = {k: p.values for k, p in atm.processes.items()}
            ref = _oracle_with_layers(c4, atm, sel, layers)
    assert out[np.float64].shape == ref.shape == (len(sel), 180000)
    assert relmax(out[np.float64], ref) < TOL64
    _check32(out[np.float32], ref, sim)


def test_c4_properties_on_device(cuda, c4):
    """Gain linearity, pW <-> K_RJ round trip, shard bit-equality with ragged boundaries, determinism; float32."""
    n = c4["instrument"].dets.n
    sim = _device(c4, np.float32, 23, noise="philox", seed=3, geometry="fast")
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    atm.simulate_pwv()
    rng = np.random.default_rng(1)
    gain = np.exp(0.05 * rng.standard_normal(n))
    atm.set_gain(gain)
    krj = sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=True)
    again = sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=True)
    assert cuda.equal(krj, again)  # deterministic
    nogain = sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=False)
    ratio = (krj[::37, ::997] / nogain[::37, ::997]).cpu().numpy()
    assert np.max(np.abs(ratio / gain[::37, None] - 1)) < 1e-6  # linear in the detector gain
    # ragged shards: 1 detector, an odd count, a tile-straddling range; every shard equals the full run bit for bit
    for d0, m in [(0, 1), (5, 33), (1000, 531), (n - 7, 7)]:
        part = sim._compute_atmospheric_loading(obs, d0, m, units="K_RJ", apply_gain=True)
        assert cuda.equal(part, krj[d0:d0 + m]), (d0, m)
    # units: pW -> K_RJ -> pW on the device equals the pW synthesis (tests/tod/test_tod.py:37-49 round trip)
    pw = sim._compute_atmospheric_loading(obs, 0, 256, units="pW", apply_gain=True)
    tod = M.TOD({"atmosphere": krj[:256].clone()}, dets=obs.instrument.dets, coords=obs.coords, units="K_RJ", dtype=np.float32,
                metadata={"det0": 0}, calibrator=atm)
    back = tod.to("pW").device_data["atmosphere"]
    assert float((back / pw - 1).abs().max()) < 2e-6
    assert bool(cuda.isfinite(krj).all())
    atm.check_flags()


def test_colocated_detectors_share_the_line_of_sight_bit_exactly(cuda):
    """Weak-scaling layout (sub-array by sub-array, both bands of a pixel in the same shard): a shard whose detectors pair
    up runs the shared-line-of-sight phase 1; dropping its first detector shifts every pair apart (one position per
    detector again).  Both give the same bits, and both match the fp64 oracle."""
    cfg = configs.c4_atlast(n_det=512, duration=120.0, sample_rate=50.0, n_groups=2)
    dets = cfg["instrument"].dets
    assert np.array_equal(dets.offsets[:256], dets.offsets[256:512]) and dets.band_name[0] != dets.band_name[256]
    sim = _device(cfg, np.float32, 29, geometry="exact")
    full = sim.run(units="K_RJ")[0].data["atmosphere"]
    _check32(full, _oracle(cfg, seed=29)["tod_K_RJ"], sim)
    obs = sim.obs_list[0]
    shard = sim._compute_atmospheric_loading(obs, 512, 512, units="K_RJ", apply_gain=True).cpu().numpy()    # rank 1 of 2: paired
    unpaired = sim._compute_atmospheric_loading(obs, 513, 511, units="K_RJ", apply_gain=True).cpu().numpy()
    assert np.array_equal(shard, full[512:]) and np.array_equal(unpaired, full[513:])


def test_pipelined_realisations_equal_sequential_ones(cuda):
    """bench.py and Simulation.run keep up to ``pipeline_depth`` realisations' AR chains in flight underneath the synthesis
    of an earlier one (slots of process buffers, two sets of layer buffers, per-slot streams).  Every realisation must
    come out bit-identical to generating and synthesising it on its own."""
    cfg = configs.c4_atlast(n_det=256, duration=240.0, sample_rate=50.0)
    seeds = [11, 12, 13, 14, 15]

    def make(depth):
        sim = _device(cfg, np.float32, 31, noise="philox", seed=seeds[0], geometry="fast", pipeline_depth=depth)
        return sim, sim.obs_list[0].atmosphere

    sim, atm = make(1)
    sequential = []
    for sd in seeds:
        atm.seed = sd
        atm.generate()
        sequential.append(atm.synthesize(units="K_RJ", apply_gain=False).clone())
    assert not cuda.equal(sequential[0], sequential[1])
    for depth in (2, 3):
        sim, atm = make(depth)
        for r in range(depth):
            atm.seed = seeds[r]
            atm.generate_async()
        for i, sd in enumerate(seeds):
            atm.finish_generate()
            if i + depth < len(seeds):
                atm.seed = seeds[i + depth]
                atm.generate_async()
            out = atm.synthesize(units="K_RJ", apply_gain=False)
            assert cuda.equal(out, sequential[i]), (depth, i)
        atm.join_streams()
        atm.check_flags()


def test_c5_six_hour_scan_runs_and_matches_oracle_subset(cuda):
    """C5's time axis (6 h at 50 Hz = 1.08 M samples per detector, ~48 k knots, layers six times longer than C4's) on a
    thinned focal plane: the float32 path against the oracle fed with the device's layers, and no line of sight leaves a
    layer over the whole scan."""
    cfg = configs.c5_long(n_det=256, duration=21600.0, sample_rate=50.0)
    n = cfg["instrument"].dets.n
    sim = _device(cfg, np.float32, 29, noise="philox", seed=11, geometry="fast")
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    assert max(p.n_extrusion for p in atm.processes.values()) > 80000
    atm.simulate_pwv()
    atm.set_gain(np.ones(n))
    tod = sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=True)
    assert atm.check_flags() == 0
    assert tod.shape == (n, 1080000) and bool(cuda.isfinite(tod).all())
    sel = np.array([0, n // 3, n - 1])
    ref = _oracle_with_layers(cfg, atm, sel, {k: p.values for k, p in atm.processes.items()})
    _check32(tod[sel].cpu().numpy(), ref, sim)


def test_c5_six_hour_recurrence_and_fp64_synthesis_match_oracle(cuda):
    """C5 in validation mode: (1) the AR recurrence of the longest six-hour layer (n_ext > 100 000: > 200 000 dependent
    float64 steps, six times C4's) against the oracle's chain with identical host noise, 1e-10; (2) the float64 device
    synthesis on the 1.08 M-sample axis against the oracle fed with the device's layers, 1e-10."""
    cfg = configs.c5_long(n_det=64, duration=21600.0, sample_rate=50.0)
    n = cfg["instrument"].dets.n
    sim = _device(cfg, np.float64, 37, noise="host", geometry="fast")
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    k = max(atm.processes, key=lambda q: atm.processes[q].n_extrusion * atm.processes[q].n_cross_section)
    p = atm.processes[k]
    assert p.n_extrusion > 100000 and p.n_cross_section > 100
    rng = np.random.default_rng(9)
    buffer0 = np.zeros((3 * p.n_extrusion, p.n_cross_section))
    buffer0[2 * p.n_extrusion:] = rng.standard_normal((p.n_extrusion, p.n_cross_section))  # only these rows are read
    z = rng.standard_normal((2 * p.n_extrusion, p.n_cross_section))
    dev = p.run(rng=O.ReplayRNG([buffer0, z])).cpu().numpy()
    q = O.AutoregressiveProcess(p.cross_section, p.extrusion, {}, jitter=1e-8)
    q.A, q.B = p.A, p.B
    ref = q.run(O.ReplayRNG([buffer0] + [z[i] for i in range(len(z))]))
    assert relmax(dev, ref) < TOL64
    # (2) all processes regenerated (host noise), float64 synthesis of a few detectors over the six hours
    atm.simulate_pwv()
    atm.set_gain(np.ones(n))
    tod = sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=True)
    assert atm.check_flags() == 0 and tod.shape == (n, 1080000)
    sel = np.array([0, n // 2, n - 1])
    ref = _oracle_with_layers(cfg, atm, sel, {kk: pp.values for kk, pp in atm.processes.items()})
    assert relmax(tod[sel].cpu().numpy(), ref) < TOL64


def _log_binned_psd(x, rate, n_bins=12):
    """Hann-windowed, linearly detrended periodogram, log-binned (the recipe of maria/plotting/tod.py:100-112)."""
    n = x.shape[-1]
    t = np.arange(n)
    x = x - np.polyval(np.polyfit(t, x.T, 1), t[:, None]).T if x.ndim > 1 else x - np.polyval(np.polyfit(t, x, 1), t)
    w = np.hanning(n)
    ps = np.abs(np.fft.rfft(x * w, axis=-1)) ** 2 / (rate * np.sum(w**2))
    f = np.fft.rfftfreq(n, 1 / rate)
    edges = np.geomspace(f[1], f[-1], n_bins + 1)
    idx = np.clip(np.digitize(f[1:], edges) - 1, 0, n_bins - 1)
    return np.array([ps[..., 1:][..., idx == b].mean() for b in range(n_bins)])


def test_philox_tod_power_spectrum_matches_oracle_realisations(cuda):
    """Production RNG, statistical check: the log-binned power spectrum of the device TOD over many
    Philox seeds agrees with that of oracle realisations drawn from numpy's generator, within sampling error."""
    cfg = configs.c3_multiband(n_det=24, duration=120.0, sample_rate=20.0)
    np.random.seed(31)
    sim = M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=np.float32,
                       atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], noise="philox", seed=1,
                                              geometry="exact"))
    obs, atm = sim.obs_list[0], sim.obs_list[0].atmosphere
    n = cfg["instrument"].dets.n
    atm.set_gain(np.ones(n))
    n_real = 48
    dev = []
    for s in range(n_real):
        atm.seed = 100 + s
        atm.simulate_pwv()
        dev.append(_log_binned_psd(sim._compute_atmospheric_loading(obs, units="K_RJ", apply_gain=True).cpu().numpy().astype(float), 20.0))
    # oracle realisations on the same geometry (same A, B, transforms), independent noise
    geo = {k: dict(cross_section=p.cross_section, extrusion=p.extrusion, A=p.A, B=p.B, transform=p.transform, vx=p.vx, vy=p.vy,
                   layer_index=p.layer_index, labels=p.labels, extrusion_res=p.extrusion_res) for k, p in atm.processes.items()}
    o_obs = O.Observation(cfg["instrument"], cfg["plan"], cfg["site"])
    o_atm = O.atmosphere_from_geometry(cfg["weather"], cfg["spectrum"], atm.layers, atm.timestep, atm.boresight.t,
                                       atm.boresight.az, atm.boresight.el, cfg["instrument"].dets.offsets, geo)
    rng = np.random.default_rng(7)

    class NoGain:
        def standard_normal(self, size=None):
            return rng.standard_normal(size) if np.ndim(size) or (size or 0) != n else np.zeros(n)

    ref = []
    for s in range(n_real):
        loading, meta = O.run_obs(o_obs, o_atm, NoGain(), dtype=np.float64)
        ref.append(_log_binned_psd(O.tod_to_K_RJ(loading, o_obs, o_atm, meta, out_dtype=np.float64), 20.0))
    dev, ref = np.array(dev), np.array(ref)
    ratio = dev.mean(axis=0) / ref.mean(axis=0)
    # the propagator is close to a unit root, so single realisations scatter by factors; 48 realisations each side bring the
    # ratio of the means within ~40 % per bin
    err = np.sqrt((dev.std(axis=0) / dev.mean(axis=0)) ** 2 + (ref.std(axis=0) / ref.mean(axis=0)) ** 2) / np.sqrt(n_real)
    assert np.all(np.abs(np.log(ratio)) < np.maximum(5 * err, 0.4)), (ratio, err)


def test_layers_do_not_depend_on_what_else_loads_the_memory_system(cuda):
    """The AR chain hands values between the CTAs of a cluster every half microsecond.  Whatever a CTA reads back from global
    memory must be ordered after the store that produced it, however slow the memory system is: the same seed must give the
    same layers bit for bit on an idle GPU and underneath a copy that saturates HBM (regression: far samples read one step
    after they were produced came back stale under load)."""
    cfg = configs.c4_atlast(n_det=256, duration=900.0, sample_rate=50.0)
    sim = _device(cfg, np.float32, 37, noise="philox", seed=21, geometry="fast", pipeline_depth=1)
    atm = sim.obs_list[0].atmosphere

    def layers():
        atm.realisation = 0  # the same Philox subsequence every time
        atm.generate()
        cuda.cuda.synchronize()
        return [p.values_device.clone() for p in atm.processes.values()]

    quiet = layers()
    assert all(bool(cuda.isfinite(v).all()) for v in quiet)
    src = cuda.empty(1 << 28, dtype=cuda.float32, device=atm.ctx.device).normal_()  # 1 GiB
    dst = cuda.empty_like(src)
    side = cuda.cuda.Stream(device=atm.ctx.device)
    for trial in range(3):
        side.wait_stream(cuda.cuda.current_stream(atm.ctx.device))
        with cuda.cuda.stream(side):
            for _ in range(60):  # ~20 ms of HBM-bound traffic: longer than the chains
                dst.copy_(src)
        loaded = layers()
        side.synchronize()
        for k, (a, b) in enumerate(zip(quiet, loaded)):
            assert cuda.equal(a, b), (trial, k)
