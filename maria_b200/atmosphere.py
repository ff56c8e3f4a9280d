"""``Atmosphere`` and ``AutoregressiveProcess`` with maria's interface (maria/atmosphere/atmosphere.py:31-380,
maria/atmosphere/process.py:19-209), running on one B200 through libmaria_atm.so.

Set-up geometry and the propagator matrices are float64 numpy / LAPACK on the host, once per plan (``host.py``).
Layer generation (the AR recurrence), beam smoothing, line-of-sight sampling and everything downstream are CUDA.
There is no CPU path: without the library or without a GPU these classes raise.
"""
from __future__ import annotations

import ctypes as C
import logging
import time as ttime

import numpy as np
import scipy as sp
import scipy.interpolate

from . import _lib, host, inputs

logger = logging.getLogger("maria_b200")

SUPPORTED_MODELS_LIST = ["2d", "3d"]
_PRECISIONS = {"fp32": _lib.F32, "float32": _lib.F32, "fp64": _lib.F64, "float64": _lib.F64}


def _torch():
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("maria_b200 needs a CUDA device (there is no CPU fallback)")
    return torch


class DeviceContext:
    """One libmaria_atm handle (= one GPU, one arithmetic type) plus the status -> exception mapping of
    SURVEY.md section 8b."""

    def __init__(self, device=0, precision="fp32"):
        if precision not in _PRECISIONS:
            raise ValueError(f"Invalid precision '{precision}'. Supported: fp32, fp64.")
        self.torch = _torch()
        self.lib = _lib.load()
        self.device = int(device)
        self.precision = "fp64" if _PRECISIONS[precision] == _lib.F64 else "fp32"
        self.dtype_code = _PRECISIONS[precision]
        self.torch_dtype = self.torch.float64 if self.dtype_code == _lib.F64 else self.torch.float32
        self.np_dtype = np.float64 if self.dtype_code == _lib.F64 else np.float32
        h = C.c_void_p()
        rc = self.lib.matm_create(self.device, self.dtype_code, C.byref(h))
        if rc != _lib.OK:
            raise RuntimeError(f"matm_create failed on cuda:{device} (status {rc})")
        self.h = h

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.lib.matm_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def check(self, rc):
        if rc == _lib.OK:
            return
        buf = C.create_string_buffer(1024)
        self.lib.matm_last_error(self.h, buf, 1024)
        msg = buf.value.decode(errors="replace")
        if rc == _lib.ERR_ARG:
            # call-order errors are the reference's RuntimeError ("Atmosphere must be initialized ...", atmosphere.py:294-295),
            # anything else about the arguments is a ValueError (atmosphere.py:48-51, process.py:169-172)
            if "initialized" in msg or " first" in msg or "not all set" in msg:
                raise RuntimeError(msg)
            raise ValueError(msg)
        if rc == _lib.ERR_TABLE_RANGE:
            raise ValueError(f"(T0, pwv, el) left the spectrum table: {msg}")
        raise RuntimeError(f"libmaria_atm status {rc}: {msg}")

    def stream_ptr(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def empty(self, shape, dtype=None):
        return self.torch.empty(shape, dtype=dtype or self.torch_dtype, device=f"cuda:{self.device}")

    def to_device(self, a, dtype):
        return self.torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(f"cuda:{self.device}")

    def raise_on_flags(self):
        st, layer, det, sample = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
        self.check(self.lib.matm_check_flags(self.h, C.byref(st), C.byref(layer), C.byref(det), C.byref(sample)))
        return st.value, layer.value, det.value, sample.value

    @property
    def launch_count(self):
        return int(self.lib.matm_launch_count(self.h))


class AutoregressiveProcess:
    """Sparse-precision autoregressive Gaussian process of arXiv:2111.01319 (maria/atmosphere/process.py:19-209).
    ``cross_section`` is (n_cs, 2) = (x, height), ``extrusion`` is (n_ext,)."""

    def __init__(self, cross_section, extrusion, callback=None, callback_kwargs=None, jitter=1e-6,
                 MIN_SAMPLES_PER_LAYER=4):
        self.cross_section = np.asarray(cross_section, dtype=float)
        self.extrusion = np.asarray(extrusion, dtype=float)
        self.callback_kwargs = dict(callback_kwargs or {})
        self.callback = callback if callback is not None else host.MaternTable(**self.callback_kwargs)
        self.jitter = jitter
        self.n_cross_section, self.n_extrusion = len(self.cross_section), len(self.extrusion)
        self.extrusion_sample_index, self.cross_section_sample_index = host.build_stencil(
            self.n_extrusion, self.n_cross_section, MIN_SAMPLES_PER_LAYER)
        self.sample_points = np.c_[self.extrusion[self.extrusion_sample_index],
                                   self.cross_section[self.cross_section_sample_index]]  # process.py:72-89
        self.n_sample = len(self.sample_points)
        esi = self.extrusion_sample_index
        # leading samples on the three most recent rows (esi is non-decreasing by construction)
        self.n_near = int(np.sum(esi + 1 <= 3)) if np.all(np.diff(esi) >= 0) else -1
        self.live_edge_points = np.c_[
            self.extrusion[0] * np.ones(self.n_cross_section) - np.gradient(self.extrusion).mean(),
            self.cross_section]  # process.py:91-100
        self.n_live_edge = self.n_cross_section
        if self.n_sample > 4000:  # process.py:102-107
            logger.warning(f"n_sample = {self.n_sample}: inverting the sample covariance is very expensive")
        self._dev = None
        self._values_host = None

    def compute_covariance_matrices(self, rng=np.random):
        """process.py:111-189.  Draws (and discards, as the reference effectively does) the initial slice so that the
        host random stream stays aligned with the reference."""
        t0 = ttime.monotonic()
        self.A, self.B, C_EE = host.propagator_matrices(self.sample_points, self.live_edge_points, self.callback, self.jitter)
        rng.standard_normal(self.n_cross_section)  # process.py:182, overwritten by run()
        logger.debug(f"Computed covariance matrices ({self.n_cross_section} x {self.n_sample}) in "
                     f"{ttime.monotonic() - t0:.3f} s")

    # ---- device side ----
    def bind(self, ctx, n_slots=1):
        """Upload A, B and the stencil; allocate BUFFER (3 n_ext, n_cs) and Z (2 n_ext, n_cs) on the device, once per
        pipeline slot (``n_slots`` realisations can be in flight at the same time)."""
        tc = ctx.torch
        self._ctx = ctx
        self._slot = 0
        self._dev = dict(
            A=ctx.to_device(self.A, tc.float64), B=ctx.to_device(self.B, tc.float64),
            esi=ctx.to_device(self.extrusion_sample_index.astype(np.int32), tc.int32),
            csi=ctx.to_device(self.cross_section_sample_index.astype(np.int32), tc.int32),
            buffer=[ctx.empty((3 * self.n_extrusion, self.n_cross_section), tc.float64) for _ in range(n_slots)],
            z=[ctx.empty((2 * self.n_extrusion, self.n_cross_section), tc.float64) for _ in range(n_slots)],
        )

    def run(self, rng=None, seed=None, subsequence=0, slot=0):
        """process.py:191-209.  ``rng`` (host, e.g. ``np.random``): validation mode, the noise is drawn on the host in
        the reference's order and uploaded.  ``seed``: production mode, Philox noise on the device."""
        if self._dev is None:
            raise RuntimeError("AutoregressiveProcess must be bound to a device context before run().")
        ctx, dv = self._ctx, self._dev
        n_ext, n_cs = self.n_extrusion, self.n_cross_section
        st = ctx.stream_ptr()
        buf, zz = dv["buffer"][slot], dv["z"][slot]
        tail = buf[2 * n_ext:]
        if rng is not None:
            buffer0 = rng.standard_normal((3 * n_ext, n_cs))  # process.py:196-198 (only the last n_ext rows are read)
            z = rng.standard_normal((2 * n_ext, n_cs))        # process.py:206, one row per executed step
            tail.copy_(ctx.torch.from_numpy(buffer0[2 * n_ext:]))
            zz.copy_(ctx.torch.from_numpy(z))
        else:
            if seed is None:
                raise ValueError("either a host rng or a Philox seed is required")
            ctx.check(ctx.lib.matm_fill_normal(ctx.h, tail.data_ptr(), tail.numel(), seed, 2 * subsequence, st))
            ctx.check(ctx.lib.matm_fill_normal(ctx.h, zz.data_ptr(), zz.numel(), seed, 2 * subsequence + 1, st))
        ctx.check(ctx.lib.matm_ar_generate(ctx.h, n_ext, n_cs, self.n_sample, self.n_near, dv["A"].data_ptr(), dv["B"].data_ptr(),
                                           dv["esi"].data_ptr(), dv["csi"].data_ptr(), buf.data_ptr(), zz.data_ptr(), st))
        self._values_host = None
        return buf[:n_ext]

    @property
    def values_device(self):
        """(n_ext, n_cs) float64 view of the generated field on the device."""
        return self._dev["buffer"][self._slot][: self.n_extrusion]

    @property
    def values(self):
        if self._values_host is None:
            self._values_host = self.values_device.cpu().numpy()
        return self._values_host


class LazyCoords:
    """The part of ``Coordinates`` the consumers of ``Atmosphere.coords`` read (``.t``, ``.az``, ``.el``), with the
    (n_det, n_t) arrays evaluated only when asked for."""

    def __init__(self, t, az_b, el_b, offsets):
        self.t, self._az_b, self._el_b, self._offsets = t, az_b, el_b, offsets
        self.shape = (len(offsets), len(t))

    def bind_device(self, atmosphere):
        """Evaluate on the GPU from now on: ``az`` / ``el`` become sliceable lazy arrays (``coords.el[100:132, ::50]``
        touches 32 detectors), nothing of size n_det x n_t is held anywhere."""
        self._atm = atmosphere

    def _eval(self):
        if not hasattr(self, "_az"):
            self._az, self._el = host.offsets_to_az_el(self._offsets, self._az_b, self._el_b)
        return self._az, self._el

    def _lazy(self, which):
        from .sim import LazyArray

        atm = self._atm
        return LazyArray(self.shape, np.float64, lambda det0, n: atm.pointing(det0, n, az=which == 0, el=which == 1)[which])

    @property
    def az(self):
        return self._lazy(0) if getattr(self, "_atm", None) is not None else self._eval()[0]

    @property
    def el(self):
        return self._lazy(1) if getattr(self, "_atm", None) is not None else self._eval()[1]


class Boresight:
    def __init__(self, t, az, el):
        self.t, self.az, self.el = np.asarray(t, dtype=float), np.asarray(az, dtype=float), np.asarray(el, dtype=float)
        self.shape = self.az.shape

    def downsample(self, timestep):
        """coordinates.py:286-304: linear interpolation onto arange(t.min(), t.max(), timestep)."""
        ds_t = np.arange(self.t.min(), self.t.max(), timestep)
        kw = dict(axis=-1, bounds_error=False, fill_value="extrapolate")
        return Boresight(ds_t, sp.interpolate.interp1d(self.t, self.az, **kw)(ds_t),
                         sp.interpolate.interp1d(self.t, self.el, **kw)(ds_t))


class Atmosphere:
    """maria/atmosphere/atmosphere.py:31-380.  Same constructor arguments and attributes as the reference; offline
    there are no ERA5 / ``am`` tables, so ``weather`` may also be a ready ``inputs.Weather`` and ``spectrum`` a ready
    ``inputs.AtmosphericSpectrum`` (a dict ``weather`` is the reference's override, applied to the synthetic stand-in).

    Opt-in keys (all have defaults, so the reference's call sites work unchanged):
      device     CUDA device index                     precision  "fp32" (production) | "fp64" (validation)
      noise      "host": the AR noise is drawn from ``rng`` on the host in the reference's order (validation mode)
                 "philox": counter-based RNG on the device, keyed by ``seed``
      rng        object with ``standard_normal(size)``; default the global ``np.random`` like the reference
      geometry   "exact" | "fast" | "auto": footprint alignment on all points (reference) or on their 3-D hull
      obs_index  index of the observation (plan) this atmosphere belongs to: part of the Philox subsequence, so that the
                 plans of one Simulation draw independent atmospheres from one seed
      layer_exchange  under torch.distributed: rank r generates only the processes k with k % world == r (AR chain + beam
                 smoothing) and the smoothed layers are broadcast (NCCL over NVLink; SURVEY.md section 8e option B)
                 instead of every rank regenerating every layer.  Philox noise only (host noise has a global draw order).

    Philox noise is keyed by (seed, obs_index, realisation, process): every ``simulate_pwv()`` / ``generate_async()`` is a
    fresh realisation (``realisation`` counts them, like successive draws from the reference's global ``np.random``);
    assign ``atm.realisation = n`` to reproduce realisation n.
    """

    def __init__(self, model="2d", timestamp=None, region="princeton", altitude=None, weather=None,
                 weather_quantiles=None, weather_source="era5", spectrum_source="am", pwv_rms_frac=0.03,
                 max_height=5e3, timestep=None, interpolation_method="linear", disable_progress_bars=False, *,
                 spectrum=None, device=0, precision="fp32", noise="host", seed=1234, rng=None, geometry="auto",
                 pipeline_depth=1, obs_index=0, layer_exchange=False):
        if model not in SUPPORTED_MODELS_LIST:
            raise ValueError(f"Invalid model '{model}'. Supported models are {SUPPORTED_MODELS_LIST}.")
        if noise not in ("host", "philox"):
            raise ValueError(f"Invalid noise mode '{noise}'. Supported: 'host', 'philox'.")
        if interpolation_method != "linear":
            raise ValueError("only interpolation_method='linear' is implemented on the device")
        if precision not in _PRECISIONS:
            raise ValueError(f"Invalid precision '{precision}'. Supported: fp32, fp64.")
        if isinstance(weather, inputs.Weather):
            self.weather = weather
        else:
            self.weather = inputs.make_weather(region, altitude or 0.0, pwv=(weather or {}).get("pwv"))
        self.spectrum = spectrum if spectrum is not None else inputs.make_spectrum(region)
        self.timestamp, self.region = timestamp, region
        self.pwv_rms_frac = pwv_rms_frac  # stored, not forwarded: the reference does the same (atmosphere.py:67,86-94)
        self.angular = False
        self.model, self.max_height, self.timestep = model, max_height, timestep
        self.interpolation_method = interpolation_method
        self.disable_progress_bars = disable_progress_bars
        self.device, self.precision, self.noise, self.seed = device, precision, noise, seed
        self.rng = rng if rng is not None else np.random
        self.geometry = geometry
        self.pipeline_depth = max(1, int(pipeline_depth))  # realisations whose AR chains may be in flight at once
        self.obs_index, self.realisation = int(obs_index), 0
        self.layer_exchange = bool(layer_exchange)
        if self.layer_exchange and noise != "philox":
            raise ValueError("layer_exchange needs noise='philox' (host noise is drawn in one global order)")
        self._initialized = False
        self._generated = False
        self._layers_set_once = False

    def __repr__(self):
        return f"Atmosphere(model={self.model}, region={self.region}, pwv={self.weather.pwv:.3f} mm, device=cuda:{self.device})"

    # ---------------------------------------------------------------------------------------------------------
    def initialize(self, obs, bind=True):
        """atmosphere.py:81-281: layer table, timestep, downsampled boresight, one AR process per layer.
        ``bind=False`` stops after the host set-up (no device is touched; used by the CPU tests of the host logic)."""
        dets = obs.instrument.dets
        rng = self.rng
        self.layers = host.generate_layers(dets, obs.boresight.el.min(), self.weather, obs.site.altitude, mode=self.model,
                                           angular=self.angular, max_height=self.max_height)
        L = self.layers
        if self.timestep is None:  # :96-99
            min_fwhm = host.gaussian_beam_angular_fwhm(dets.primary_size, self.max_height, dets.band_center).min()
            self.timestep = max(1e-1, min_fwhm / (L["wind_speed"] / L["h"]).max())
        self.boresight = obs.boresight.downsample(timestep=self.timestep)  # :101
        n_t = len(self.boresight.t)
        self.coords = LazyCoords(self.boresight.t, self.boresight.az, self.boresight.el, dets.offsets)  # :102-105
        outer = dets.offsets[host.outer_detector_mask(dets.offsets)]  # :108-111
        outer_az, outer_el = host.offsets_to_az_el(outer, self.boresight.az, self.boresight.el)

        self.processes = {}
        for process_index in sorted(np.unique(L["process_index"])):
            t0 = ttime.monotonic()
            idx = np.where(L["process_index"] == process_index)[0]
            vx = L["wind_east"][idx][:, None] * np.ones(n_t)  # :149-154 (non-angular branch)
            vy = L["wind_north"][idx][:, None] * np.ones(n_t)
            w = L["absolute_humidity"][idx] * L["temperature"][idx] * L["divergence"][idx]
            vx = (w[:, None] * vx).sum(axis=0) / w.sum()
            vy = (w[:, None] * vy).sum(axis=0) / w.sum()
            drift = np.cumsum(self.timestep * np.c_[vx, vy, np.zeros(n_t)][None], axis=-2)
            hull_layers = idx[[0, -1]] if len(idx) > 1 else idx  # :161-162
            pts = np.concatenate([(host.project_to_height(outer_az, outer_el, L["h"][li]) + drift)[None]
                                  for li in hull_layers], axis=0).reshape(-1, 3)
            pts[..., 2] += 1e-6 * rng.standard_normal(pts[..., 2].shape)  # :184-186
            fast = self.geometry == "fast" or (self.geometry == "auto" and len(pts) > 400_000)
            transform = host.aligning_transform(pts, (True, True, False), rng=rng, reduce_to_hull=fast)  # :191-194
            tp = pts @ transform
            min_tx, max_tx = tp[:, 0].min(), tp[:, 0].max()
            lo_y, hi_y = tp[:, 1].min(), tp[:, 1].max()
            sections, labels = [], []
            for li in idx:  # :203-219
                res = L["res"][li]
                n_cs = int(np.maximum(2, (np.ptp(tp[:, 1]) + 2 * res) / res))
                sections.append(np.stack([np.linspace(lo_y - res, hi_y + res, n_cs), L["h"][li] * np.ones(n_cs)]).T)
                labels.extend(n_cs * [li])
            res_min = L["res"].min()
            extrusion = np.arange(min_tx - 2 * res_min, max_tx + 2 * res_min, res_min)  # :241-245
            outer_scale = np.maximum(1e3, 300 + L["h"][hull_layers].mean() / 10)  # :247
            kw = {"nu": 1 / 3, "r0": outer_scale} if self.model == "3d" else {"nu": 5 / 6, "r0": outer_scale}
            process = AutoregressiveProcess(np.concatenate(sections, axis=0), extrusion, callback_kwargs=kw, jitter=1e-8)
            process.layer_index = idx
            process.labels = np.array(labels)
            process.transform = transform
            process.extrusion_res = np.gradient(extrusion).mean()
            process.vx, process.vy = vx, vy
            process.tp_bounds = (min_tx, max_tx, lo_y, hi_y)
            for attempt, jitter in enumerate([1e-8, 1e-6, 1e-4]):  # :267-277 (the reference never applies `jitter`)
                try:
                    process.compute_covariance_matrices(rng)
                    break
                except np.linalg.LinAlgError:
                    logger.debug(f"Singular covariance matrix with jitter={jitter}")
                    if attempt == 2:
                        raise np.linalg.LinAlgError("Covariance matrix was singular even with max jitter.")
            self.processes[int(process_index)] = process
            logger.debug(f"Constructed process {process_index} ({process.n_extrusion} x {process.n_cross_section}, "
                         f"n_sample={process.n_sample}) in {ttime.monotonic() - t0:.2f} s")
        self._obs = obs
        if bind:
            self._bind_device(obs)
            self._initialized = True

    # ---------------------------------------------------------------------------------------------------------
    def _bind_device(self, obs):
        """Create the device context and upload scan, detectors, band tables and the process matrices."""
        self.ctx = DeviceContext(self.device, self.precision)
        self._bind_scan_and_detectors(obs, None)
        self.set_bands(obs.instrument.dets)
        rank, world = self._rank_world()
        for k, p in self.processes.items():
            # under layer_exchange a rank runs only its own processes: the others need no BUFFER / Z on this GPU
            p.bind(self.ctx, self.pipeline_depth if k % world == rank else 0)
        self._pending_slots, self._issued = [], 0
        self._layer_sets = [{}, {}]   # two sets of smoothed-layer buffers: the sampler reads one, the next is written
        self._layer_cur = 0
        self._reader_done = [None, None]  # event after the last kernel that reads each set

    def _bind_scan_and_detectors(self, obs, gain):
        """Host -> device: boresight at both rates, knot times, detector offsets / band index / Mueller / gain."""
        ctx, dets = self.ctx, obs.instrument.dets
        lib = ctx.lib
        bs, full = self.boresight, obs.boresight
        t_a, az_a, el_a = _lib.f64(bs.t), _lib.f64(bs.az), _lib.f64(bs.el)
        t_f, az_f, el_f = _lib.f64(full.t), _lib.f64(full.az), _lib.f64(full.el)
        ctx.check(lib.matm_set_scan(ctx.h, len(bs.t), _lib.dptr(t_a), float(self.timestep), _lib.dptr(az_a), _lib.dptr(el_a),
                                    len(t_f), _lib.dptr(t_f), _lib.dptr(az_f), _lib.dptr(el_f)))
        self._layers_set = False
        if not hasattr(self, "_det_host"):
            band_names = [b.name for b in dets.bands]
            lut = {n: i for i, n in enumerate(band_names)}
            self._det_host = (np.array([lut[b] for b in dets.band_name], dtype=np.int32), _lib.f64(dets.offsets),
                              _lib.f64(getattr(obs, "offsets", dets.offsets)), _lib.f64(dets.mueller00()))
        band_index, off_atm, off_full, mu = self._det_host
        g = _lib.f64(gain) if gain is not None else None
        ctx.check(lib.matm_set_detectors(ctx.h, dets.n, _lib.dptr(off_atm), _lib.dptr(off_full), _lib.iptr(band_index),
                                         _lib.dptr(mu), _lib.dptr(g) if g is not None else None))

    def set_bands(self, dets):
        """Band-integrated lookup tables (band.py:235-255,264-300), evaluated on the device from the spectrum tables."""
        ctx, s = self.ctx, self.spectrum
        T, W, E, NU = (_lib.f64(a) for a in s.points)
        pb = _lib.f64(np.stack([b.passband(s.side_nu) for b in dets.bands]))
        factor = _lib.f64([0.5 if (~np.isnan(dets.gamma[dets.band_name == b.name])).any() else 1.0 for b in dets.bands])
        em, op = _lib.f64(s._emission), _lib.f64(s._opacity)
        self.cal_pwv = float(np.round(self.weather.pwv, 3))            # simulation.py:258-259 metadata, used by TOD.to
        self.cal_T0 = float(np.round(self.weather.temperature[0], 3))
        ctx.check(ctx.lib.matm_set_bands(ctx.h, len(dets.bands), len(T), len(W), len(E), len(NU), _lib.dptr(T), _lib.dptr(W),
                                         _lib.dptr(E), _lib.dptr(NU), _lib.dptr(em), _lib.dptr(op), _lib.dptr(pb),
                                         float(self.weather.temperature[0]), self.cal_T0, self.cal_pwv, _lib.dptr(factor)))

    def set_gain(self, gain):
        """Per-detector gain factors exp(gain_error N(0,1)) (simulation.py:239-241)."""
        ctx, dets = self.ctx, self._obs.instrument.dets
        g = _lib.f64(gain)
        if g.shape != (dets.n,):
            raise ValueError(f"gain must have one entry per detector ({dets.n}), got {g.shape}")
        self._gain_host = g  # keep the buffer alive until the copy has been issued
        ctx.check(ctx.lib.matm_set_gain(ctx.h, _lib.dptr(g), ctx.stream_ptr()))

    # ---------------------------------------------------------------------------------------------------------
    def generate(self):
        """Stage 1: run every AR process and beam-smooth every layer on the device (atmosphere.py:299-307,328-344)."""
        self.generate_async()
        self.finish_generate()

    def generate_async(self):
        """Start the AR recurrences of all processes for one more realisation, on their own high-priority streams (one
        stream per process and pipeline slot: the recurrences are independent sequential chains of 1-6 CTAs each and
        run side by side).  Nothing the sampler reads is touched, so a synthesis of an EARLIER realisation enqueued after
        this call overlaps with the chains: up to ``pipeline_depth`` realisations are generated underneath the synthesis
        kernels.  ``finish_generate`` completes the oldest one."""
        if not self._initialized:
            raise RuntimeError("Atmosphere must be initialized with a simulation.")
        if len(self._pending_slots) >= self.pipeline_depth:
            raise RuntimeError(f"{len(self._pending_slots)} realisations are already in flight (pipeline_depth={self.pipeline_depth})")
        ctx = self.ctx
        tc = ctx.torch
        cur = tc.cuda.current_stream(ctx.device)
        fork = cur.record_event()
        if not hasattr(self, "_streams"):
            # high priority: a chain CTA needs a whole SM (256 threads x ~190 registers, ~150 KB of shared memory), so when it runs underneath a
            # synthesis kernel it must be first in line for the SMs that drain
            self._streams = {(s, k): tc.cuda.Stream(device=ctx.device, priority=-1)
                             for s in range(self.pipeline_depth) for k in self.processes}
        if not hasattr(self, "_layer_plan"):
            self._build_layer_plan()
        slot = self._issued % self.pipeline_depth
        self._issued += 1
        realisation = self.realisation
        self.realisation += 1
        rank, world = self._rank_world()
        for k, process in self.processes.items():
            if self.layer_exchange and k % world != rank:
                continue  # another rank generates this process; its layers arrive in finish_generate
            stream = self._streams[(slot, k)]
            stream.wait_event(fork)
            with tc.cuda.stream(stream):
                if self.noise == "host":
                    process.run(rng=self.rng, slot=slot)
                else:
                    # subsequence = (observation, realisation, process): no two draws of a Simulation share a Philox stream
                    process.run(seed=self.seed, subsequence=(self.obs_index << 44) | (realisation << 12) | k, slot=slot)
        self._pending_slots.append(slot)

    def _rank_world(self):
        if not self.layer_exchange:
            return 0, 1
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()):
            return 0, 1
        return dist.get_rank(), dist.get_world_size()

    @property
    def _pending(self):
        return len(self._pending_slots) > 0

    def finish_generate(self):
        """Beam-smooth the oldest realisation in flight into the buffers the sampler reads (after everything enqueued so
        far on the current stream, i.e. after any synthesis still reading the previous layers), and join its streams."""
        if not self._pending_slots:
            raise RuntimeError("finish_generate() without generate_async()")
        ctx = self.ctx
        tc = ctx.torch
        cur = tc.cuda.current_stream(ctx.device)
        # the spare set of layer buffers: free once the kernels that read it (two realisations ago) are done, so the
        # smoothing runs underneath the synthesis that still reads the current set
        target = 1 - self._layer_cur if self._layers_set_once else self._layer_cur
        readers_done = self._reader_done[target]
        slot = self._pending_slots.pop(0)
        rank, world = self._rank_world()
        for k, process in self.processes.items():
            stream = self._streams[(slot, k)]
            if readers_done is not None:
                stream.wait_event(readers_done)
            process._slot = slot
            process._values_host = None
            with tc.cuda.stream(stream):
                st = ctx.stream_ptr()
                for (li, col0, n_cols, sig_e, sig_x) in self._layer_plan[k]:
                    out, scratch = self._layer_sets[target][li]
                    if k % world == rank:
                        ctx.check(ctx.lib.matm_smooth_layer(ctx.h, process.values_device.data_ptr(), process.n_extrusion,
                                                            process.n_cross_section, col0, n_cols, sig_e, sig_x,
                                                            out.data_ptr(), scratch.data_ptr(), st))
            if world == 1:
                cur.wait_stream(stream)
        if world > 1:
            # The one exchange step of the path (SURVEY.md section 8e option B): once this rank's layers are smoothed, every
            # owner's segment of the flat layer buffer goes to all ranks -- `world` broadcasts on one side stream,
            # underneath the synthesis that still reads the other set of layer buffers.  (One collective per layer on the
            # process streams was measured first: the waiting NCCL kernels held SMs and slowed the synthesis by 40 %.)
            import torch.distributed as dist

            if not hasattr(self, "_xstream"):
                self._xstream = tc.cuda.Stream(device=ctx.device, priority=-1)
            xs = self._xstream
            if readers_done is not None:
                xs.wait_event(readers_done)
            for k in self.processes:
                if k % world == rank:
                    xs.wait_stream(self._streams[(slot, k)])
            flat = self._flat_sets[target]
            with tc.cuda.stream(xs):
                for r, (off, n) in enumerate(self._owner_segments):
                    if n:
                        dist.broadcast(flat[off:off + n], src=r)
                        self.exchanged_bytes = getattr(self, "exchanged_bytes", 0) + n * flat.element_size()
            cur.wait_stream(xs)
        self._layer_cur = target
        if not getattr(self, "_layers_set", False):
            self.set_layers()
        else:
            ptrs = (C.c_void_p * len(self._layer_order))(*[self._layer_sets[target][li][0].data_ptr() for li in self._layer_order])
            ctx.check(ctx.lib.matm_set_layer_data(ctx.h, len(self._layer_order), ptrs))
        self._layers_set_once = True
        self._generated = True

    def drain(self):
        """Complete every realisation still in flight (the last one stays in the sampler's buffers)."""
        while self._pending_slots:
            self.finish_generate()

    def join_streams(self):
        """Make the current stream wait for everything enqueued on the process streams."""
        cur = self.ctx.torch.cuda.current_stream(self.ctx.device)
        for stream in getattr(self, "_streams", {}).values():
            cur.wait_stream(stream)

    def _build_layer_plan(self):
        """Per layer: its columns in the process, the beam-smoothing widths in cells (atmosphere.py:328-339) and the
        device buffers (smoothed layer in the context's dtype + float64 scratch)."""
        ctx, L, dets = self.ctx, self.layers, self._obs.instrument.dets
        rank, world = self._rank_world()
        self._layer_plan = {}
        shapes = {}
        for k, process in self.processes.items():
            plan = []
            for li in process.layer_index:
                cols = np.where(process.labels == li)[0]
                col0, n_cols = int(cols[0]), len(cols)
                beam_sigma = host.gaussian_beam_physical_fwhm(dets.primary_size, L["z"][li], dets.band_center).mean() / 2.355
                plan.append((int(li), col0, n_cols, float(beam_sigma / process.extrusion_res), float(beam_sigma / L["res"][li])))
                shapes[int(li)] = (k, process.n_extrusion, n_cols)
            self._layer_plan[k] = plan
        # Every set of smoothed layers is ONE flat buffer, the layers of the processes of one owner rank (k % world)
        # contiguous: under layer_exchange the exchange is then one broadcast per owner (one rank: plain views).
        order = sorted(shapes, key=lambda li: (shapes[li][0] % world, li))
        total = sum(shapes[li][1] * shapes[li][2] for li in order)
        self._flat_sets, self._owner_segments = [], []
        for ls in self._layer_sets:
            flat = ctx.empty((total,))
            self._flat_sets.append(flat)
            off = 0
            for li in order:
                k, n_ext, n_cols = shapes[li]
                ls[li] = (flat[off:off + n_ext * n_cols].view(n_ext, n_cols), ctx.empty((n_ext, n_cols), ctx.torch.float64))
                off += n_ext * n_cols
        off = 0
        for r in range(world):
            n = sum(shapes[li][1] * shapes[li][2] for li in order if shapes[li][0] % world == r)
            self._owner_segments.append((off, n))
            off += n

    def set_layers(self):
        """Hand the layer geometry (height, transform, grids, wind) and the device layers to the sampler."""
        ctx, L = self.ctx, self.layers
        descs = (_lib.LayerDesc * len(L["h"]))()
        keep, n_layers, order = [], 0, []
        for k, process in self.processes.items():
            vx, vy = _lib.f64(process.vx), _lib.f64(process.vy)
            keep += [vx, vy]
            for (li, col0, n_cols, _, _) in self._layer_plan[k]:
                d = descs[n_layers]
                d.data_dev, d.n_ext, d.n_cs = self._layer_dev[li][0].data_ptr(), process.n_extrusion, n_cols
                order.append(li)
                d.h, d.pwv_rms = float(L["h"][li]), float(L["pwv_rms"][li])
                d.transform = (C.c_double * 9)(*np.asarray(process.transform, dtype=float).ravel())
                d.ext0, d.ext_step = float(process.extrusion[0]), float(process.extrusion_res)
                xs = process.cross_section[col0:col0 + n_cols, 0]
                d.cs0, d.cs_last = float(xs[0]), float(xs[-1])
                d.vx_host, d.vy_host = _lib.dptr(vx), _lib.dptr(vy)
                n_layers += 1
        ctx.check(ctx.lib.matm_set_layers(ctx.h, n_layers, descs, float(self.weather.pwv)))
        self._layer_order = order
        self._layers_set = True

    @property
    def _layer_dev(self):
        """The set of smoothed-layer buffers the sampler currently reads."""
        return self._layer_sets[self._layer_cur]

    def _mark_readers(self):
        """Record that a kernel reading the current layer set has just been enqueued on the current stream."""
        self._reader_done[self._layer_cur] = self.ctx.torch.cuda.current_stream(self.ctx.device).record_event()

    def smoothed_layer(self, li):
        """Beam-smoothed layer ``li`` as a host array (validation helper)."""
        return self._layer_dev[li][0].cpu().numpy()

    def zenith_pwv_device(self, det0=0, n_det=None):
        """(n_det, n_t_atm) float64 zenith_scaled_pwv on the device for a detector range (atmosphere.py:309-373)."""
        ctx = self.ctx
        n_det = self._obs.instrument.dets.n - det0 if n_det is None else n_det
        out = ctx.empty((n_det, len(self.boresight.t)), ctx.torch.float64)
        ctx.check(ctx.lib.matm_zenith_pwv(ctx.h, det0, n_det, out.data_ptr(), ctx.stream_ptr()))
        self._mark_readers()
        return out

    def synthesize(self, det0=0, n_det=None, units="K_RJ", apply_gain=True, out=None):
        """The fused kernel (a6..a14): (n_det, n_t) TOD of the atmosphere field on the device, in the context's dtype."""
        ctx = self.ctx
        n_det = self._obs.instrument.dets.n - det0 if n_det is None else n_det
        if out is None:
            out = ctx.empty((n_det, len(self._obs.boresight.t)))
        ctx.check(ctx.lib.matm_synthesize(ctx.h, det0, n_det, out.data_ptr(), _lib.UNITS_KRJ if units == "K_RJ" else _lib.UNITS_PW,
                                          1 if apply_gain else 0, ctx.stream_ptr()))
        self._mark_readers()
        return out

    def synthesize_host(self, out_host, det0=0, n_det=None, units="K_RJ", apply_gain=True):
        """Same, delivered into a host tensor / ndarray (pinned for full PCIe speed); returns when the data has landed."""
        ctx = self.ctx
        n_det = self._obs.instrument.dets.n - det0 if n_det is None else n_det
        ptr = out_host.data_ptr() if hasattr(out_host, "data_ptr") else out_host.ctypes.data
        ctx.torch.cuda.current_stream(ctx.device).synchronize()  # the staging streams do not follow torch's stream
        ctx.check(ctx.lib.matm_synthesize_host(ctx.h, det0, n_det, ptr, _lib.UNITS_KRJ if units == "K_RJ" else _lib.UNITS_PW,
                                               1 if apply_gain else 0))
        return out_host

    # ---- consumers of the path's intermediates, tile-wise on the device ----
    def transmission(self, nu_min_Hz=0.0, nu_max_Hz=np.inf, det0=0, n_det=None, out=None):
        """pW per K_RJ of the frequency channel [nu_min, nu_max) for the map step (sim/map.py:117-133): (n_det, n_t) in
        the context's dtype, on the device; zenith_scaled_pwv is sampled and upsampled inside the kernel."""
        ctx = self.ctx
        n_det = self._obs.instrument.dets.n - det0 if n_det is None else n_det
        if out is None:
            out = ctx.empty((n_det, len(self._obs.boresight.t)))
        ctx.check(ctx.lib.matm_transmission(ctx.h, det0, n_det, float(nu_min_Hz), float(min(nu_max_Hz, 1e300)), out.data_ptr(),
                                            ctx.stream_ptr()))
        self._mark_readers()
        return out

    def noise_scale(self, loading, noise, nep, nep_per_loading, det0=0, out=None):
        """1e12 (NEP + NEP_per_loading loading) noise per band (sim/noise.py:35-37,60); device tensors in the context's
        dtype, ``nep`` / ``nep_per_loading`` per band in W sqrt(s), in the order of ``dets.bands``."""
        ctx = self.ctx
        out = ctx.torch.empty_like(noise) if out is None else out
        a, b = _lib.f64(nep), _lib.f64(nep_per_loading)
        ctx.check(ctx.lib.matm_noise_scale(ctx.h, det0, loading.shape[0], loading.data_ptr(), noise.data_ptr(), _lib.dptr(a),
                                           _lib.dptr(b), out.data_ptr(), ctx.stream_ptr()))
        return out

    def pointing(self, det0=0, n_det=None, j0=0, n_j=None, az=True, el=True):
        """Full-rate az / el of a detector range over a sample range (obs.coords, observation.py:55-58), float64, on the
        device, evaluated on demand."""
        ctx = self.ctx
        n_det = self._obs.instrument.dets.n - det0 if n_det is None else n_det
        n_j = len(self._obs.boresight.t) - j0 if n_j is None else n_j
        a = ctx.empty((n_det, n_j), ctx.torch.float64) if az else None
        e = ctx.empty((n_det, n_j), ctx.torch.float64) if el else None
        ctx.check(ctx.lib.matm_pointing(ctx.h, det0, n_det, j0, n_j, a.data_ptr() if az else None, e.data_ptr() if el else None,
                                        ctx.stream_ptr()))
        return a, e

    def simulate_pwv(self, instrument=None):
        """atmosphere.py:293-380.  Generates the layers; ``zenith_scaled_pwv`` is evaluated when read."""
        if not self._initialized:
            raise RuntimeError("Atmosphere must be initialized with a simulation.")
        if self._pending:  # the chains were started ahead (Simulation.run pipelines observations)
            self.finish_generate()
        else:
            self.generate()
        self.__dict__.pop("_zpwv_host", None)

    @property
    def zenith_scaled_pwv(self):
        if not self._generated:
            raise RuntimeError("simulate_pwv() has not been called.")
        if "_zpwv_host" not in self.__dict__:
            z = self.zenith_pwv_device().cpu().numpy()
            self.check_flags()
            self._zpwv_host = z
        return self._zpwv_host

    def check_flags(self):
        """Raise what the reference raises when a line of sight leaves a layer (atmosphere.py:368-369)."""
        st, layer, det, sample = self.ctx.raise_on_flags()
        if st == _lib.ERR_OUT_OF_GRID:
            k = int(self.layers["process_index"][layer]) if layer < len(self.layers["h"]) else -1
            raise RuntimeError(f"Layer {layer} in process {k} introduced nans into PWV simulation "
                               f"(detector {det}, sample {sample}).")
        return st
