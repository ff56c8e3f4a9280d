"""Adapters between thomaswmorris/maria's own objects and this package (the seam is duck typing).

``from_maria`` turns a ``maria.Instrument`` / ``maria.Plan`` / ``maria.Site`` (quantities with units, pandas-backed
detector tables, ``Band`` objects) into the plain-array views of ``maria_b200.inputs`` that the host set-up and the CUDA
path read: SI floats and ndarrays, nothing else.  ``weather_from_maria`` / ``spectrum_from_maria`` do the same for the
state of ``maria.weather.Weather`` and ``maria.spectrum.AtmosphericSpectrum`` after their h5 load.

``install()`` patches the reference in place so that

    maria.Simulation(instrument, plans, site, atmosphere="2d", atmosphere_kwargs={"backend": "b200"})

builds a ``B200Atmosphere`` (maria/sim/observation.py:88-96) and routes ``AtmosphereMixin._simulate_atmosphere`` /
``_compute_atmospheric_loading`` (maria/sim/atmosphere.py:23-84) through libmaria_atm.so; everything else of
``Simulation.run`` (gain, ``TOD``, ``TOD.to``) stays the reference's own code.  Without the key the reference runs
unchanged.  ``maria_b200.Simulation`` accepts the reference's objects directly as well (device-resident TOD, fused kernel).

Nothing here imports ``maria`` at module import time: the objects are recognised by their attributes.
"""
from __future__ import annotations

import numpy as np

from . import inputs as I


# ---------------------------------------------------------------------------------------------------------------
# units
# ---------------------------------------------------------------------------------------------------------------
def _si(x, unit=None):
    """A maria ``Quantity`` (or a plain number / array) as float / ndarray in SI base units.  ``unit`` names the attribute
    that converts (``m``, ``Hz``, ``rad``, ``s``); quantities also expose ``base_units_value`` (maria/units/quantity.py)."""
    if hasattr(x, "base_units_value"):
        v = getattr(x, unit) if unit is not None else x.base_units_value
        return np.asarray(v, dtype=float) if np.ndim(v) else float(v)
    return np.asarray(x, dtype=float) if np.ndim(x) else float(x)


def is_plain(instrument):
    """True for this package's own stand-ins (``inputs.Instrument``): nothing to convert."""
    return isinstance(instrument, I.Instrument)


# ---------------------------------------------------------------------------------------------------------------
# instrument / plan / site
# ---------------------------------------------------------------------------------------------------------------
def band_from_maria(b):
    """maria/band/band.py:86-135: sampled passband ``nu`` [Hz], ``tau``, ``efficiency``, ``gain_error``, ``center``."""
    out = I.Band(str(b.name), nu=_si(b.nu, "Hz"), tau=np.asarray(b.tau, dtype=float), efficiency=float(b.efficiency),
                 gain_error=float(getattr(b, "gain_error", 0.0)))
    # Band(nu=, tau=) renormalises tau and scales the efficiency; the reference's values are already in their final state
    out.efficiency = float(b.efficiency)
    out.tau = np.asarray(b.tau, dtype=float).copy()
    out.center = _si(b.center, "Hz")
    return out


def array_from_maria(dets):
    """maria/array/array.py:136-233: the detector table as plain arrays (offsets in radians, primary size in metres)."""
    bands = [band_from_maria(b) for b in dets.bands]
    arr = I.Array(_si(dets.xi, "rad"), _si(dets.eta, "rad"), np.asarray(dets.band_name).astype(str), bands,
                  np.asarray(dets.gamma, dtype=float), _si(dets.primary_size, "m"))
    arr.band_center = np.asarray(dets.band_center, dtype=float).copy()  # array.py:236-240 (Hz)
    return arr


def from_maria(instrument, plan, site):
    """(instrument, plan, site) of the reference -> (inputs.Instrument, inputs.Plan, inputs.Site)."""
    inst = I.Instrument(array_from_maria(instrument.dets), name=str(getattr(instrument, "name", "maria")))
    frame = getattr(plan, "frame", "az/el")
    if frame != "az/el":
        raise ValueError(f"plan frame '{frame}': only az/el plans are supported on this path (sky frames need astropy)")
    p = I.Plan(_si(plan.time, "s"), _si(plan.phi, "rad"), _si(plan.theta, "rad"), roll=float(getattr(plan, "roll", 0.0)))
    s = I.Site(str(site.region), _si(site.altitude, "m"))
    return inst, p, s


def coerce(instrument, plan, site):
    """Pass this package's own stand-ins through, convert the reference's objects."""
    if is_plain(instrument) and isinstance(plan, I.Plan) and isinstance(site, I.Site):
        return instrument, plan, site
    if is_plain(instrument):  # mixed: convert what needs converting
        inst = instrument
    else:
        inst = I.Instrument(array_from_maria(instrument.dets), name=str(getattr(instrument, "name", "maria")))
    p = plan if isinstance(plan, I.Plan) else I.Plan(_si(plan.time, "s"), _si(plan.phi, "rad"), _si(plan.theta, "rad"),
                                                     roll=float(getattr(plan, "roll", 0.0)))
    s = site if isinstance(site, I.Site) else I.Site(str(site.region), _si(site.altitude, "m"))
    return inst, p, s


# ---------------------------------------------------------------------------------------------------------------
# weather / spectrum
# ---------------------------------------------------------------------------------------------------------------
def weather_from_maria(w):
    """State of maria.weather.Weather after its load (weather/__init__.py:137-152): the profiles already carry the wind
    correction and the ``override`` (e.g. pwv), so they are taken as they are."""
    if isinstance(w, I.Weather):
        return w
    profiles = {f: np.asarray(getattr(w, f), dtype=float) for f in I.Weather.fields}
    return I.Weather(str(w.region), _si(w.base_altitude, "m"), profiles, wind_corrected=True)


def spectrum_from_maria(s):
    """maria.spectrum.AtmosphericSpectrum after its load (spectrum/atmosphere.py:38-57): tables interpolated to the site
    altitude, elevation axis in radians with the last point at 90.1 degrees."""
    if isinstance(s, I.AtmosphericSpectrum):
        return s
    out = I.AtmosphericSpectrum(str(s.region), s.side_base_temperature, s.side_zenith_pwv,
                                np.degrees(np.asarray(s.side_elevation, dtype=float)), s.side_nu, s._emission, s._opacity)
    out.side_elevation = np.asarray(s.side_elevation, dtype=float).copy()  # exactly the reference's axis
    return out


# ---------------------------------------------------------------------------------------------------------------
# the drop-in: maria.Simulation(..., atmosphere="2d", atmosphere_kwargs={"backend": "b200"})
# ---------------------------------------------------------------------------------------------------------------
def _b200_atmosphere_class():
    from .atmosphere import Atmosphere

    class B200Atmosphere(Atmosphere):
        """``maria.atmosphere.Atmosphere`` with the reference's constructor (atmosphere.py:32-47), built from the reference's
        own ``Weather`` / ``AtmosphericSpectrum`` (ERA5 / am tables), initialised from the reference's ``Observation``."""

        def __init__(self, model="2d", timestamp=None, region="princeton", altitude=None, weather=None,
                     weather_quantiles=None, weather_source="era5", spectrum_source="am", pwv_rms_frac=0.03,
                     max_height=5e3, timestep=None, interpolation_method="linear", disable_progress_bars=False, **opts):
            spectrum = opts.pop("spectrum", None)
            if not isinstance(weather, I.Weather) or spectrum is None:
                import maria  # the reference's loaders (h5 tables from its cache)

                if spectrum is None:
                    spectrum = maria.spectrum.AtmosphericSpectrum(region=region, source=spectrum_source)
                if not isinstance(weather, I.Weather):
                    weather = maria.weather.Weather(time=timestamp, region=region, altitude=altitude, override=weather or {},
                                                    quantiles=weather_quantiles or {}, source=weather_source)
            self.reference_weather, self.reference_spectrum = weather, spectrum
            super().__init__(model=model, timestamp=timestamp, region=region, altitude=altitude,
                             weather=weather_from_maria(weather), weather_quantiles=weather_quantiles,
                             weather_source=weather_source, spectrum_source=spectrum_source, pwv_rms_frac=pwv_rms_frac,
                             max_height=max_height, timestep=timestep, interpolation_method=interpolation_method,
                             disable_progress_bars=disable_progress_bars, spectrum=spectrum_from_maria(spectrum), **opts)

        def initialize(self, obs, bind=True):
            """``obs`` is the reference's Observation (sim/observation.py:27-102) or this package's."""
            from .sim import Observation

            if not is_plain(obs.instrument):
                inst, plan, site = from_maria(obs.instrument, obs.plan, obs.site)
                view = Observation(inst, plan, site)
                self._reference_obs = obs
                obs = view
            super().initialize(obs, bind=bind)

    return B200Atmosphere


_installed = {}


def install():
    """Patch the imported reference so that ``atmosphere_kwargs={"backend": "b200"}`` selects the CUDA path.  Idempotent;
    returns the ``B200Atmosphere`` class.  ``uninstall()`` restores the reference."""
    if _installed:
        return _installed["cls"]
    import maria.sim.atmosphere as msa
    import maria.sim.observation as mso

    cls = _b200_atmosphere_class()
    orig_atm, orig_sim, orig_load = mso.Atmosphere, msa.AtmosphereMixin._simulate_atmosphere, msa.AtmosphereMixin._compute_atmospheric_loading

    def atmosphere_factory(*args, backend=None, **kwargs):  # observation.py:88-96 calls Atmosphere(model=..., **atmosphere_kwargs)
        if backend in (None, "reference"):
            return orig_atm(*args, **kwargs)
        if backend != "b200":
            raise ValueError(f"Invalid atmosphere backend '{backend}'. Supported: 'reference', 'b200'.")
        return cls(*args, **kwargs)

    def _simulate_atmosphere(self, obs):  # sim/atmosphere.py:24-37
        if not isinstance(obs.atmosphere, cls):
            return orig_sim(self, obs)
        from .sim import LazyArray

        atm = obs.atmosphere
        atm.simulate_pwv(obs.instrument)
        ctx, n_det, n_t = atm.ctx, atm._obs.instrument.dets.n, len(atm._obs.boresight.t)

        def upsampled(det0, n):
            zp = atm.zenith_pwv_device(det0, n)
            out = ctx.empty((n, n_t), ctx.torch.float64)
            ctx.check(ctx.lib.matm_upsample_linear(ctx.h, zp.data_ptr(), n, out.data_ptr(), ctx.stream_ptr()))
            return out

        obs.zenith_scaled_pwv = LazyArray((n_det, n_t), np.float64, upsampled)

    def _compute_atmospheric_loading(self, obs):  # sim/atmosphere.py:39-84: (n_det, n_t) [pW] of self.dtype, no gain
        if not isinstance(obs.atmosphere, cls):
            return orig_load(self, obs)
        atm = obs.atmosphere
        if np.dtype(self.dtype) != np.dtype(atm.ctx.np_dtype):
            raise RuntimeError(f"Simulation dtype {np.dtype(self.dtype)} needs atmosphere_kwargs precision="
                               f"'{'fp64' if np.dtype(self.dtype) == np.float64 else 'fp32'}'")
        data = atm.synthesize(units="pW", apply_gain=False).cpu().numpy()
        atm.check_flags()
        try:  # the reference hands a dask array on (sim/atmosphere.py:71-82)
            import dask.array as da

            return da.asarray(data, dtype=self.dtype)
        except Exception:
            return data

    mso.Atmosphere = atmosphere_factory
    msa.AtmosphereMixin._simulate_atmosphere = _simulate_atmosphere
    msa.AtmosphereMixin._compute_atmospheric_loading = _compute_atmospheric_loading
    _installed.update(cls=cls, mso=mso, msa=msa, orig=(orig_atm, orig_sim, orig_load))
    return cls


def uninstall():
    if not _installed:
        return
    mso, msa = _installed["mso"], _installed["msa"]
    mso.Atmosphere, msa.AtmosphereMixin._simulate_atmosphere, msa.AtmosphereMixin._compute_atmospheric_loading = _installed["orig"]
    _installed.clear()
