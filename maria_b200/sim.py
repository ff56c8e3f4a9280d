"""``Simulation(instrument, plans, site, atmosphere="2d").run()`` for the atmosphere field, with maria's interface
(maria/sim/simulation.py:76-272, maria/sim/observation.py:27-102, maria/sim/atmosphere.py:23-84).

``run()`` goes through ONE fused kernel per observation (line-of-sight sampling of all layers, emission lookup, cubic
upsample, gain, pW -> K_RJ); the mixin methods ``_simulate_atmosphere`` / ``_compute_atmospheric_loading`` keep the
reference's staged interface for callers that want the intermediates.
"""
from __future__ import annotations

import logging
import time as ttime

import numpy as np

from . import _lib, host
from .atmosphere import Atmosphere, Boresight, LazyCoords
from .tod import TOD

logger = logging.getLogger("maria_b200")


class LazyArray:
    """A (n_det, n_t) array that lives nowhere: ``compute(det0, n)`` evaluates the rows [det0, det0 + n) on the device.
    Slicing evaluates only the detector range the key touches (``a[100:132]``, ``a[7]``, ``a[idx, ::10]``); converting the
    whole array (``np.asarray``) evaluates it in detector chunks of at most ``max_chunk_bytes`` on the device."""

    def __init__(self, shape, dtype, compute, max_chunk_bytes=2 << 30):
        self.shape, self.dtype, self._compute = tuple(shape), np.dtype(dtype), compute
        self._rows = max(1, int(max_chunk_bytes // max(1, self.shape[1] * self.dtype.itemsize)))

    @property
    def ndim(self):
        return 2

    def compute_device(self, det0=0, n=None):
        return self._compute(det0, self.shape[0] - det0 if n is None else n)

    def _rows_host(self, d0, d1, rest=None):
        parts = []
        for a in range(d0, d1, self._rows):
            blk = self._compute(a, min(self._rows, d1 - a))
            blk = blk if rest is None else blk[(slice(None),) + rest]
            parts.append(blk.cpu().numpy())
        return np.concatenate(parts, axis=0) if parts else np.empty((0,) + self.shape[1:], self.dtype)

    def __array__(self, dtype=None, copy=None):
        a = self._rows_host(0, self.shape[0])
        return a.astype(dtype) if dtype is not None else a

    def __getitem__(self, key):
        key = key if isinstance(key, tuple) else (key,)
        first, rest = key[0], tuple(key[1:])
        n = self.shape[0]
        if isinstance(first, (int, np.integer)):
            d = int(first) % n
            return self._rows_host(d, d + 1, rest)[0]
        if isinstance(first, slice):
            idx = np.arange(n)[first]
        else:
            idx = np.asarray(first)
            idx = np.nonzero(idx)[0] if idx.dtype == bool else idx % n
        if len(idx) == 0:
            return np.empty((0,) + self.shape[1:], self.dtype)[(slice(None),) + rest]
        d0, d1 = int(idx.min()), int(idx.max()) + 1
        if (d1 - d0) <= 4 * len(idx):  # one covering range
            return self._rows_host(d0, d1, rest)[idx - d0]
        return np.stack([self._rows_host(int(d), int(d) + 1, rest)[0] for d in idx])


class Observation:
    """The fields of maria/sim/observation.py:27-102 that the atmosphere path reads."""

    def __init__(self, instrument, plan, site, atmosphere=None, atmosphere_kwargs=None):
        from . import adapters

        instrument, plan, site = adapters.coerce(instrument, plan, site)  # the reference's own objects are accepted
        self.instrument, self.plan, self.site = instrument, plan, site
        self.boresight = Boresight(plan.time, plan.phi, plan.theta)  # :46-52
        dets = instrument.dets
        self.offsets = dets.offsets @ host.rotation_matrix_2d(np.radians(getattr(plan, "roll", 0.0))).T  # :55-58
        self.coords = LazyCoords(self.boresight.t, self.boresight.az, self.boresight.el, self.offsets)
        min_el = np.degrees(self.boresight.el.min())
        if min_el < 5:  # :67-71 (PointingError in the reference)
            raise ValueError(f"Minimum elevation is {min_el:.2f} degrees; the atmosphere model needs at least 5 degrees.")
        self.loading = {}
        self.atmosphere = None
        if atmosphere:  # :88-96
            kw = dict(atmosphere_kwargs or {})
            kw.setdefault("region", site.region)
            kw.setdefault("altitude", site.altitude)
            kw.setdefault("timestamp", float(np.mean(plan.time)))
            if "weather" in kw and not isinstance(kw["weather"], (dict, type(None))) and not hasattr(kw["weather"], "data"):
                kw["weather"] = adapters.weather_from_maria(kw["weather"])  # a maria.weather.Weather
            if kw.get("spectrum") is not None:
                kw["spectrum"] = adapters.spectrum_from_maria(kw["spectrum"])
            self.atmosphere = Atmosphere(model=atmosphere, **kw)


class AtmosphereMixin:
    """maria/sim/atmosphere.py:23-84."""

    def _simulate_atmosphere(self, obs):
        atm = obs.atmosphere
        atm.simulate_pwv(obs.instrument)
        ctx = atm.ctx
        n_det, n_t = obs.instrument.dets.n, len(obs.boresight.t)

        def upsampled(det0, n):  # linear interp1d to the full rate (:30-37), for the detectors [det0, det0 + n)
            zp = atm.zenith_pwv_device(det0, n)
            out = ctx.empty((n, n_t), ctx.torch.float64)
            ctx.check(ctx.lib.matm_upsample_linear(ctx.h, zp.data_ptr(), n, out.data_ptr(), ctx.stream_ptr()))
            return out

        obs.zenith_scaled_pwv = LazyArray((n_det, n_t), np.float64, upsampled)

    def _compute_atmospheric_loading(self, obs, det0=0, n_det=None, units="pW", apply_gain=False, out=None):
        """(n_det, n_t) loading [pW] in ``self.dtype`` on the device (:39-84): emission lookup x Mueller[0, 0] at the
        atmosphere rate, cubic (not-a-knot) upsample to the full rate."""
        return obs.atmosphere.synthesize(det0, n_det, units=units, apply_gain=apply_gain, out=out)


class Simulation(AtmosphereMixin):
    """maria/sim/simulation.py:76-272, restricted to the atmosphere field.

    ``dtype`` is the TOD dtype as in the reference (float32 default; float64 selects the fp64 validation arithmetic).
    ``det_range=(first, count)`` restricts synthesis to a contiguous detector shard (multi-GPU, one process per GPU).
    """

    def __init__(self, instrument, plans, site, atmosphere=None, atmosphere_kwargs=None, dtype=np.float32, noise=False,
                 progress_bars=False, det_range=None, **kwargs):
        self.instrument, self.site = instrument, site
        self.plans = list(plans) if isinstance(plans, (list, tuple)) or type(plans).__name__ == "PlanList" else [plans]
        self.dtype = np.dtype(dtype)
        self.det_range = det_range
        self.disable_progress_bars = not progress_bars
        akw = dict(atmosphere_kwargs or {})
        akw.setdefault("precision", "fp64" if self.dtype == np.float64 else "fp32")
        self.obs_list = []
        for obs_index, plan in enumerate(self.plans):  # :142-157
            akw_i = dict(akw)
            akw_i.setdefault("obs_index", obs_index)  # plans draw independent atmospheres from one Philox seed
            obs = Observation(instrument, plan, site, atmosphere=atmosphere, atmosphere_kwargs=akw_i)
            if obs.atmosphere is not None:
                t0 = ttime.monotonic()
                obs.atmosphere.initialize(obs)
                obs.coords.bind_device(obs.atmosphere)  # obs.coords.az / .el: device slices on demand
                logger.debug(f"Initialized atmosphere in {ttime.monotonic() - t0:.2f} s")
            self.obs_list.append(obs)
        if det_range is not None:
            self._check_ranks_agree()

    def _check_ranks_agree(self):
        """Detector shards of several ranks tile one TOD only if every rank built the same layer geometry (same host RNG
        stream): under torch.distributed the set-up fingerprints are compared (sharding.assert_same_setup)."""
        try:
            import torch.distributed as dist
        except Exception:
            return
        if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            return
        from . import sharding

        for obs in self.obs_list:
            if obs.atmosphere is not None:
                sharding.assert_same_setup(dist, sharding.setup_fingerprint(obs.atmosphere))

    def run(self, units="K_RJ"):
        """:201-211 -- one TOD per plan, in ``units``."""
        tods = []
        for k, obs in enumerate(self.obs_list):
            nxt = self.obs_list[k + 1].atmosphere if k + 1 < len(self.obs_list) else None
            tods.append(self.run_obs(obs, units=units, prefetch=nxt))
        return tods

    def run_obs(self, obs, units="pW", prefetch=None):
        """:213-272, atmosphere branch; the reference builds the pW TOD and converts, here both happen in the kernel."""
        atm = obs.atmosphere
        if atm is None:
            raise RuntimeError("this build synthesises the atmosphere field only: pass atmosphere='2d'")
        dets = obs.instrument.dets
        self._simulate_atmosphere(obs)  # :216-221 (generates the layers)
        gain_error = np.exp(dets.gain_error * atm.rng.standard_normal(size=dets.n))  # :239-241
        atm.set_gain(gain_error)  # one asynchronous copy (matm_set_gain): nothing below waits for the device
        if prefetch is not None and prefetch.noise == "philox" and prefetch is not atm:
            # device noise has no host draw order to respect: start the next observation's AR chains now, so that they
            # run (on a handful of SMs) underneath this observation's synthesis kernel
            prefetch.generate_async()
        det0, n_det = self.det_range if self.det_range is not None else (0, dets.n)
        data = self._compute_atmospheric_loading(obs, det0, n_det, units=units, apply_gain=True)
        obs.loading["atmosphere"] = data
        atm.check_flags()
        metadata = {"atmosphere": True, "sim_time": ttime.time(), "altitude": float(obs.site.altitude),
                    "region": obs.site.region, "pwv": atm.cal_pwv, "base_temperature": atm.cal_T0, "det0": det0}  # :249-263
        return TOD(data={"atmosphere": data}, dets=dets, coords=obs.coords, units=units, dtype=self.dtype,
                   metadata=metadata, calibrator=atm)
