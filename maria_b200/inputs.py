"""Synthetic, network-free inputs for the atmospheric TOD path.

The reference builds these objects from YAML presets and from ERA5 / ``am`` h5 tables that it
downloads at run time (maria/io/caching.py:117-169).  Neither exists offline, so this module
provides duck-typed stand-ins that expose exactly the attributes the hot path reads, in SI
units as plain floats / ndarrays (metres, radians, hertz, seconds, kelvin, millimetres PWV):

* ``Band``        -- maria/band/band.py:61-135,317-323 (passband generation + ``passband(nu)``)
* ``Array``       -- maria/array/array.py:136-233 (offsets, mask, outer, fwhm, field_of_view, mueller)
* ``Weather``     -- maria/weather/__init__.py:137-228 (state after the h5 load: profiles, pwv, __call__)
* ``AtmosphericSpectrum`` -- maria/spectrum/atmosphere.py:38-57,97-104 (table layout, ``points``)
* ``Plan`` / scan generators -- maria/plan/plan.py:54-154 (boresight az/el(t) only; an input generator)
* ``Site``, ``Instrument``

They are inputs to both the CUDA path and the CPU oracle; they hold no part of the path itself.
"""
from __future__ import annotations

import dataclasses

import numpy as np
import scipy as sp
import scipy.interpolate
import scipy.spatial

# maria/constants.py
c = 2.99792458e8
g = 9.806651
k_B = 1.380649e-23


# ----------------------------------------------------------------------------------------------
# bands
# ----------------------------------------------------------------------------------------------
def generate_passband(center, width, shape="gaussian", samples=1024):
    """maria/band/band.py:61-83."""
    if shape == "flat":
        nu_min, nu_max = center - 0.6 * width, center + 0.6 * width
    elif shape == "top_hat":
        nu_min, nu_max = center - width, center + width
    else:
        nu_min, nu_max = center - 1.5 * width, center + 1.5 * width
    nu = np.linspace(nu_min, nu_max, samples)
    if shape == "flat":
        tau = np.where((nu > center - 0.5 * width) & (nu < center + 0.5 * width), 1.0, 0.0)
    elif shape == "gaussian":
        tau = np.exp(np.log(0.5) * (2 * (nu - center) / width) ** 2)
    elif shape == "top_hat":
        tau = np.exp(np.log(0.5) * (2 * (nu - center) / width) ** 8)
    else:
        raise ValueError(f"Invalid shape '{shape}'")
    return nu, tau


class Band:
    """The part of maria's Band the atmosphere path reads (band.py:86-135,317-323)."""

    def __init__(self, name, center=None, width=None, nu=None, tau=None, shape="gaussian",
                 efficiency=0.5, gain_error=0.0):
        auto = center is not None and width is not None
        manual = nu is not None and tau is not None
        if not auto ^ manual:
            raise ValueError("You must pass either both 'center' and 'width' or both 'nu' and 'tau'.")
        self.name = name
        if auto:
            self.nu, self.tau = generate_passband(center, width, shape, samples=1024)
        else:
            tau = np.asarray(tau, dtype=float)
            tau_max = tau.max()
            efficiency *= tau_max
            self.nu = np.asarray(nu, dtype=float)
            self.tau = tau / tau_max
        self.efficiency = efficiency
        self.gain_error = gain_error
        # band.py: center = sum(nu * tau) / sum(tau) for manual bands
        self.center = float(center) if auto else float(np.sum(self.nu * self.tau) / np.sum(self.tau))

    def passband(self, nu):
        """maria/band/band.py:317-323."""
        return self.efficiency * sp.interpolate.interp1d(self.nu, self.tau, bounds_error=False, fill_value=0)(nu)

    def __repr__(self):
        return f"Band({self.name}, center={self.center / 1e9:.1f} GHz)"


# MUSTANG-2 band (maria/band/configs/m2.yml:1-8)
M2_F093 = dict(
    efficiency=0.1, gain_error=5e-2,
    nu=[7.40e10, 7.50e10, 7.60e10, 7.70e10, 7.80e10, 7.90e10, 8.00e10, 8.10e10, 8.20e10, 8.30e10, 8.40e10,
        8.50e10, 8.60e10, 8.70e10, 8.80e10, 8.90e10, 9.00e10, 9.10e10, 9.20e10, 9.30e10, 9.40e10, 9.50e10,
        9.60e10, 9.70e10, 9.80e10, 9.90e10, 1.00e11, 1.01e11, 1.02e11, 1.03e11, 1.05e11],
    tau=[0.0, 0.970, 1.000, 0.940, 0.890, 0.940, 0.870, 0.900, 0.850, 0.780, 0.820, 0.710, 0.700, 0.780, 0.700,
         0.680, 0.610, 0.530, 0.640, 0.560, 0.570, 0.550, 0.450, 0.440, 0.370, 0.360, 0.300, 0.310, 0.340, 0.310,
         0.0],
)


# ----------------------------------------------------------------------------------------------
# detector array
# ----------------------------------------------------------------------------------------------
def hex_offsets(n, field_of_view_rad):
    """n points hex-packed in a disc whose diameter is the field of view (input generator)."""
    if n == 1:
        return np.zeros((1, 2))
    m = int(np.ceil(np.sqrt(n / 0.9069) / 2)) + 2
    i, j = np.meshgrid(np.arange(-2 * m, 2 * m + 1), np.arange(-2 * m, 2 * m + 1))
    x = (i + 0.5 * (j % 2)).ravel()
    y = (j * np.sqrt(3) / 2).ravel()
    r = np.hypot(x, y)
    order = np.argsort(r, kind="stable")[:n]
    x, y = x[order], y[order]
    scale = 0.5 * field_of_view_rad / np.hypot(x, y).max()
    return np.c_[x * scale, y * scale]


class Array:
    """Detector table, maria/array/array.py.  Everything is a plain ndarray in SI units."""

    def __init__(self, xi, eta, band_name, bands, gamma, primary_size):
        self.xi = np.asarray(xi, dtype=float)
        self.eta = np.asarray(eta, dtype=float)
        self.band_name = np.asarray(band_name)
        self.bands = list(bands)
        self.gamma = np.asarray(gamma, dtype=float)  # NaN = unpolarised
        self.primary_size = np.broadcast_to(np.asarray(primary_size, dtype=float), self.xi.shape).copy()
        by_name = {b.name: b for b in self.bands}
        self.band_center = np.array([by_name[b].center for b in self.band_name])
        self.gain_error = np.array([by_name[b].gain_error for b in self.band_name])

    @property
    def n(self):
        return len(self.xi)

    @property
    def offsets(self):  # array.py:169-170
        return np.c_[self.xi, self.eta]

    def mask(self, band_name):  # array.py:136-140
        return self.band_name == band_name

    def _subset(self, mask):
        names = set(self.band_name[mask])
        return Array(self.xi[mask], self.eta[mask], self.band_name[mask], [b for b in self.bands if b.name in names],
                     self.gamma[mask], self.primary_size[mask])

    def one_detector_from_each_band(self):  # array.py:149-154
        first = np.isin(np.arange(self.n), np.unique(self.band_name, return_index=True)[1])
        return self._subset(first)

    def outer(self):  # array.py:156-162
        try:
            hull = sp.spatial.ConvexHull(self.offsets)
        except sp.spatial.QhullError:
            return self
        return self._subset(np.isin(np.arange(self.n), hull.vertices))

    @property
    def field_of_view(self):  # array.py:176-178 ; utils/__init__.py:56-90
        X = self.offsets
        if len(X) < 16:
            i, j = np.triu_indices(n=len(X), k=1)
            return float(np.sqrt(np.sum(np.square(X[i] - X[j]), axis=-1).max())) if len(i) else 0.0
        hull = sp.spatial.ConvexHull(X)
        v = hull.points[hull.vertices]
        i, j = np.triu_indices(len(v), k=1)
        return float(np.sqrt(np.max(np.square(v[i] - v[j]).sum(axis=-1))))

    def angular_fwhm(self, z=np.inf):  # array.py:223-227 ; beam/__init__.py:9-25
        w_0 = self.primary_size / 2
        z_r = np.pi * w_0**2 / (c / self.band_center)
        with np.errstate(divide="ignore"):
            return 2 * w_0 * np.sqrt(1 / np.asarray(z, dtype=float) ** 2 + 1 / z_r**2)

    def physical_fwhm(self, z):  # array.py:229-233
        return z * self.angular_fwhm(z)

    def mueller00(self):  # array.py:204-218, element [0, 0]
        m0 = np.where(np.isnan(self.gamma), np.sqrt(2), 1.0)
        return 0.5 * m0 * m0


@dataclasses.dataclass
class Instrument:
    dets: Array
    name: str = "synthetic"


@dataclasses.dataclass
class Site:
    region: str
    altitude: float  # metres


def make_instrument(n_per_band, bands, field_of_view_deg, primary_size, polarized=False, name="synthetic",
                    focal_plane_offsets_deg=None):
    """Hex-packed arrays, one per entry of ``bands`` groups.  ``bands`` is a list of lists of Band: each inner
    list is one sub-array whose ``n_per_band`` positions each carry one detector per band (and two orthogonal
    polarisations when ``polarized``), which is how maria's presets lay out dichroic pixels
    (array/configs/act.yml:1-26).  Detectors are emitted sorted by band name (array.py:504)."""
    groups = bands if isinstance(bands[0], (list, tuple)) else [bands]
    fpo = focal_plane_offsets_deg or [(0.0, 0.0)] * len(groups)
    xi, eta, names, gamma, all_bands = [], [], [], [], []
    for grp, (ox, oy) in zip(groups, fpo):
        off = hex_offsets(n_per_band, np.radians(field_of_view_deg))
        off = off + np.radians([ox, oy])
        for b in grp:
            all_bands.append(b)
            pols = [0.0, np.pi / 2] if polarized else [np.nan]
            for pa in pols:
                xi.append(off[:, 0])
                eta.append(off[:, 1])
                names.extend([b.name] * len(off))
                gamma.append(np.full(len(off), pa))
    xi, eta, gamma = np.concatenate(xi), np.concatenate(eta), np.concatenate(gamma)
    names = np.array(names)
    order = np.argsort(names, kind="stable")
    uniq = []
    for b in all_bands:
        if b.name not in [u.name for u in uniq]:
            uniq.append(b)
    uniq = sorted(uniq, key=lambda b: b.name)
    return Instrument(Array(xi[order], eta[order], names[order], uniq, gamma[order], primary_size), name=name)


# ----------------------------------------------------------------------------------------------
# scan plans (input generator only: boresight az/el as a function of time)
# ----------------------------------------------------------------------------------------------
def offsets_to_phi_theta_f64(dX, cphi, ctheta):
    """float64 statement of maria/coords/transforms.py:10-29 (used to place the scan on the sky)."""
    dx, dy = dX[..., 0], dX[..., 1]
    r = np.sqrt(dx**2 + dy**2)
    p = np.arctan2(-dx, -dy)
    q = (np.sin(r) * np.cos(p) + 1j * np.cos(r)) * np.exp(1j * (ctheta - np.pi / 2))
    return np.stack([np.arctan2(np.sin(r) * np.sin(p), q.real) + cphi, np.arcsin(q.imag)], axis=-1)


def _smooth_sawtooth(p, delta=0.01):  # maria/plan/patterns.py:158-160
    norm = 1 / (2 * np.arccos(delta - 1) / np.pi - 1)
    return norm * (1 - 2 * np.arccos((delta - 1) * np.cos(p)) / np.pi)


def scan_offsets(pattern, time, radius_deg=1.0, speed_deg=0.5, y_throw_deg=None, n_rows=11):
    t = time - time[0]
    R = radius_deg
    if pattern == "stare":
        return np.zeros((2, len(t)))
    if pattern == "daisy":  # petal curve in the spirit of patterns.py:108-155
        petals = np.sqrt(np.e)
        phase = t * speed_deg / max(R, 1e-16) / petals
        x = R * np.cos(petals * phase) * np.sin(phase)
        y = R * np.cos(petals * phase) * np.cos(phase)
        return np.stack([x, y])
    if pattern == "back_and_forth":  # constant-elevation scan, patterns.py:163-179
        period = 4 * R / speed_deg
        x = R * _smooth_sawtooth(2 * np.pi * t / period)
        return np.stack([x, np.zeros_like(x)])
    if pattern == "raster":  # patterns.py:182-239 in spirit: fast axis x, slow axis y
        Ry = R if y_throw_deg is None else y_throw_deg
        period = 4 * R / speed_deg
        x = R * _smooth_sawtooth(2 * np.pi * t / period)
        y = Ry * _smooth_sawtooth(2 * np.pi * t / (period * n_rows))
        return np.stack([x, y])
    raise ValueError(f"unknown scan pattern '{pattern}'")


class Plan:
    """maria/plan/plan.py: only ``time``, ``phi`` (az), ``theta`` (el), ``frame``, ``roll``."""

    def __init__(self, time, phi, theta, frame="az/el", roll=0.0):
        self.time = np.asarray(time, dtype=float)
        self.phi = np.asarray(phi, dtype=float)
        self.theta = np.asarray(theta, dtype=float)
        if frame != "az/el":
            raise ValueError("only az/el plans are supported on this path (sky frames need astropy)")
        self.frame = frame
        self.roll = roll

    @classmethod
    def generate(cls, start_time=1.7e9, duration=60.0, sample_rate=50.0, scan_center=(45.0, 45.0),
                 scan_pattern="daisy", scan_options=None, roll=0.0):
        time = np.arange(start_time, start_time + duration, 1 / sample_rate)  # plan.py:83
        off = np.radians(scan_offsets(scan_pattern, time, **(scan_options or {})))
        pt = offsets_to_phi_theta_f64(off.T, np.radians(scan_center[0]), np.radians(scan_center[1]))
        return cls(time, pt[..., 0], pt[..., 1], roll=roll)

    @property
    def sample_rate(self):
        return 1 / np.mean(np.gradient(self.time))


# ----------------------------------------------------------------------------------------------
# weather
# ----------------------------------------------------------------------------------------------
def saturation_pressure(temperature):  # weather/__init__.py:31-34 (Pa; Magnus form)
    T = temperature - 273.15
    a, b, c_ = 611.21, 17.67, 238.88
    return a * np.exp(b * T / (c_ + T))


def relative_to_absolute_humidity(temperature, humidity):  # weather/__init__.py:55-56
    return humidity * saturation_pressure(temperature) / (461.5 * temperature)


def absolute_to_relative_humidity(temperature, abs_hum):  # weather/__init__.py:59-60
    return 461.5 * temperature * abs_hum / saturation_pressure(temperature)


class Weather:
    """State of maria's Weather after the h5 load (weather/__init__.py:137-228): vertical profiles on pressure
    levels, ``pwv`` as the integral of absolute humidity above the site, ``__call__(altitude)`` interpolation."""

    fields = ["temperature", "humidity", "wind_east", "wind_north", "wind_speed", "geopotential", "divergence"]

    def __init__(self, region, base_altitude, profiles, override=None, wind_corrected=False):
        self.region = region
        self.base_altitude = float(base_altitude)
        self.data = {k: np.asarray(v, dtype=float).copy() for k, v in profiles.items()}
        if not wind_corrected:  # :139-143
            factor = self.data["wind_speed"] / np.sqrt(self.data["wind_east"] ** 2 + self.data["wind_north"] ** 2)
            self.data["wind_north"] = self.data["wind_north"] * factor
            self.data["wind_east"] = self.data["wind_east"] * factor
        self.altitude = self.data["geopotential"] / g  # :145
        override = override or {}
        if "pwv" in override:  # :150-152
            force = override["pwv"] / self.pwv * self.absolute_humidity
            self.data["humidity"] = absolute_to_relative_humidity(self.temperature, force)

    def __getattr__(self, attr):
        data = self.__dict__.get("data", {})
        if attr in data:
            return data[attr]
        raise AttributeError(attr)

    @property
    def absolute_humidity(self):  # :160-162
        return relative_to_absolute_humidity(self.temperature, self.humidity)

    @property
    def pwv(self):  # :176-220 (layers().total_water.sum())
        alt = self.altitude
        keep = alt > self.base_altitude
        h_values = alt[keep]
        h_bins = np.array([self.base_altitude, *(h_values[:-1] + h_values[1:]) / 2, h_values[-1] + 100])
        total = 0.0
        f = sp.interpolate.interp1d(self.altitude, self.absolute_humidity, bounds_error=False, fill_value="extrapolate")
        for h1, h2 in zip(h_bins[:-1], h_bins[1:]):
            h = np.linspace(h1, h2, 256)
            total += np.trapezoid(f(h), x=h)
        return float(total)

    def __call__(self, altitude):  # :222-228
        res = {}
        for field in [*self.fields, "absolute_humidity"]:
            res[field] = sp.interpolate.interp1d(self.altitude, getattr(self, field))(altitude)
        return res


def make_weather(region="synthetic", base_altitude=0.0, pwv=None, T0=280.0, wind0=5.0, wind_top=30.0):
    """Physically shaped stand-in for an ERA5 quantile draw: wind rising with height and
    veering 90 degrees over 12 km, 6.5 K/km lapse rate to a 215 K floor, humidity scale height 2 km."""
    alt = base_altitude + np.concatenate([np.arange(-200.0, 3000.0, 250.0), np.arange(3000.0, 24000.0, 1000.0)])
    T = np.maximum(T0 - 6.5e-3 * (alt - base_altitude), 215.0)
    rel_hum = 0.6 * np.exp(-np.maximum(alt - base_altitude, 0) / 2000.0) + 0.01
    h_rel = np.clip((alt - base_altitude) / 12000.0, 0, 1.5)
    speed = wind0 + (wind_top - wind0) * np.clip(h_rel, 0, 1)
    bearing = 0.5 * np.pi * h_rel
    profiles = dict(
        temperature=T, humidity=rel_hum, wind_east=speed * np.cos(bearing), wind_north=speed * np.sin(bearing),
        wind_speed=speed, geopotential=alt * g, divergence=1e-5 * (1 + h_rel),
    )
    return Weather(region, base_altitude, profiles, override=({"pwv": pwv} if pwv is not None else None))


# ----------------------------------------------------------------------------------------------
# spectrum tables
# ----------------------------------------------------------------------------------------------
class AtmosphericSpectrum:
    """Table layout of maria/spectrum/atmosphere.py:38-57 after interpolation to the site altitude:
    ``_emission`` [K_RJ] and ``_opacity`` [nepers] with dims (T, pwv, el, nu)."""

    def __init__(self, region, side_base_temperature, side_zenith_pwv, side_elevation_deg, side_nu, emission, opacity):
        self.region = region
        self.side_base_temperature = np.asarray(side_base_temperature, dtype=float)
        self.side_zenith_pwv = np.asarray(side_zenith_pwv, dtype=float)
        self.side_elevation = np.radians(np.asarray(side_elevation_deg, dtype=float))
        self.side_elevation[-1] = np.radians(90.1)  # spectrum/atmosphere.py:46-48
        self.side_nu = np.asarray(side_nu, dtype=float)
        self._emission = np.asarray(emission, dtype=float)
        self._opacity = np.asarray(opacity, dtype=float)

    @property
    def points(self):  # spectrum/atmosphere.py:97-104
        return (self.side_base_temperature, self.side_zenith_pwv, self.side_elevation, self.side_nu)


def make_spectrum(region="synthetic", nu_min=20e9, nu_max=320e9, n_nu=301):
    """A smooth, physically shaped stand-in for the ``am`` tables: water lines at 22 and 183 GHz plus a nu^2
    continuum scaled by PWV, O2 at 60 and 118 GHz, plane-parallel airmass, isothermal emission."""
    Ts = np.array([240.0, 255.0, 270.0, 285.0, 300.0])
    pwvs = np.array([0.0, 0.3, 0.6, 1.0, 1.5, 2.0, 3.0, 4.5, 7.0, 10.0, 15.0, 25.0])
    els = np.array([5.0, 7.0, 10.0, 14.0, 20.0, 27.0, 35.0, 45.0, 56.0, 68.0, 80.0, 90.0])
    nu = np.linspace(nu_min, nu_max, n_nu)
    ghz = nu / 1e9

    def lor(f0, w, a):
        return a / (1 + ((ghz - f0) / w) ** 2)

    tau_dry = lor(60.0, 4.0, 8.0) + lor(118.75, 1.5, 2.0) + 0.004 * (ghz / 100) ** 2 + 0.005
    tau_wet = lor(22.235, 2.0, 0.012) + lor(183.31, 4.0, 2.5) + lor(325.15, 4.0, 1.2) + 0.03 * (ghz / 100) ** 2
    T, W, E = np.meshgrid(Ts, pwvs, np.radians(els), indexing="ij")
    tau = (tau_dry[None, None, None, :] * (T[..., None] / 270.0) ** -0.5 + W[..., None] * tau_wet) / np.sin(E[..., None])
    emission = 0.95 * T[..., None] * (1 - np.exp(-tau))
    return AtmosphericSpectrum(region, Ts, pwvs, els, nu, emission, tau)
