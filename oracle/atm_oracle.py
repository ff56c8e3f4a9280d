"""CPU oracle for maria's atmospheric TOD synthesis path.  TEST INFRASTRUCTURE ONLY.

This file is a plain numpy/scipy restatement of the reference algorithm
(thomaswmorris/maria, paths cited as ``maria/<file>:<lines>``).  It is the
checker the CUDA path is compared with; nothing under ``maria_b200/`` may
import it.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` use it.

Parity status: PINNED for the float64 variant.  The reference cannot be imported as-is in the build container (jax,
dask, h5py, astropy, arrow, matplotlib are absent, its weather/spectrum tables are downloaded at run time, and its own
tests hold no golden vectors for atmosphere values -- SURVEY.md section 4).  ``oracle/pin_against_reference.py``
therefore runs the UNMODIFIED reference source in place through import stubs (jax.numpy -> numpy float64, dask.array ->
numpy, in-memory tables in the reference's h5 layout) for two real presets and checks this file against it stage by
stage: every stage a1-a14 agrees to <= 2e-15 relative (bit-equal up to the last calibration multiply), and the
reference's outputs are committed as ``tests/golden/ref_*.npz``.  Not pinned: jax's own RegularGridInterpolator code
(``rgi_linear`` restates its published algorithm) and the float32 rounding of jax with x64 disabled (``f32=True``).

Two arithmetic variants, selected with ``f32``:
  * ``f32=False``: everything in float64 (the 1e-10 target of the fp64 device path);
  * ``f32=True`` : "as run" -- float32 casts exactly where the reference goes
    through jax with x64 disabled (pointing, layer interpolation, table lookups).

Every random array is drawn from an explicit ``rng`` object exposing
``standard_normal(size)``, in the reference's draw order,
so seeding ``np.random`` and passing the ``np.random`` module reproduces the
reference's global-RNG stream.
"""
from __future__ import annotations

import numpy as np
import scipy as sp
import scipy.interpolate
import scipy.linalg
import scipy.ndimage
import scipy.optimize
import scipy.spatial
import scipy.special

# maria/constants.py:1-28
c = 2.99792458e8
g = 9.806651
k_B = 1.380649e-23
WATER_VAPOR_SPECIFIC_GAS_CONSTANT = 461.495
DRY_AIR_SPECIFIC_GAS_CONSTANT = 287.05

COV_MAT_JITTER = 1e-6  # maria/atmosphere/process.py:16


# --------------------------------------------------------------------------
# beam / array helpers
# --------------------------------------------------------------------------
def compute_angular_fwhm(fwhm_0, z=np.inf, n=1.0, nu=None):
    """maria/beam/__init__.py:9-25."""
    w_0 = fwhm_0 / 2
    z_r = np.pi * w_0**2 * n / (c / nu)
    with np.errstate(divide="ignore"):
        return 2 * w_0 * np.sqrt(1 / z**2 + 1 / z_r**2)


def angular_fwhm(primary_size, band_center, z=np.inf):
    """maria/array/array.py:223-227."""
    return compute_angular_fwhm(z=z, fwhm_0=primary_size, n=1, nu=band_center)


def physical_fwhm(primary_size, band_center, z):
    """maria/array/array.py:229-233."""
    return z * angular_fwhm(primary_size, band_center, z)


def compute_diameter(points):
    """maria/utils/__init__.py:56-90 (lazy=False, jitter=0)."""
    X = np.array(points, dtype=float).reshape(-1, points.shape[-1])
    dim_mask = np.ptp(X, axis=0) > 0
    if not dim_mask.any():
        return 0.0
    if len(X) < 16:
        i, j = np.triu_indices(n=len(X), k=1)
        return np.sqrt(np.sum(np.square(X[i] - X[j]), axis=-1).max())
    hull = sp.spatial.ConvexHull(X[:, dim_mask])
    vertices = hull.points[hull.vertices]
    i, j = np.triu_indices(len(vertices), k=1)
    return float(np.sqrt(np.max(np.square(vertices[i] - vertices[j]).sum(axis=-1))))


def outer_detector_mask(offsets):
    """maria/array/array.py:156-162 -- detectors on the convex hull of the offsets."""
    try:
        hull = sp.spatial.ConvexHull(offsets)
    except sp.spatial.QhullError:
        return np.ones(len(offsets), dtype=bool)
    return np.isin(np.arange(len(offsets)), hull.vertices)


def mueller00(gamma):
    """maria/array/array.py:204-218, element [0, 0] only."""
    m0 = np.where(np.isnan(gamma), np.sqrt(2), 1)
    return 0.5 * m0 * m0


# --------------------------------------------------------------------------
# Matern covariance and linear algebra
# --------------------------------------------------------------------------
def normalized_matern(r, nu):
    """maria/functions/__init__.py:30-39."""
    return (
        2 ** (1 - nu)
        / sp.special.gamma(nu)
        * sp.special.kv(nu, np.sqrt(2 * nu) * r + 1e-16)
        * (np.sqrt(2 * nu) * r + 1e-16) ** nu
    )


def approximate_normalized_matern(r, nu=1 / 3, r0=1e0, n_test_points=1024):
    """maria/functions/__init__.py:42-74."""
    r_eff = r / r0
    r_eff_min, r_eff_max = 1e-6, 1e3
    r_eff_safe = np.atleast_1d(np.abs(r_eff)).clip(min=r_eff_min)
    r_eff_nonzero = r_eff_safe[r_eff_safe < r_eff_max]
    r_eff_samples = np.geomspace(r_eff_min, r_eff_max, n_test_points)
    cov_samples = normalized_matern(r_eff_samples, nu=nu)
    with np.errstate(divide="ignore"):
        sf = np.exp(np.interp(np.log(r_eff_nonzero), np.log(r_eff_samples), np.log(1 - cov_samples)))
        cov = np.exp(np.interp(np.log(r_eff_nonzero), np.log(r_eff_samples), np.log(cov_samples)))
    t = 1 / (1 + r_eff_nonzero**2)
    res = np.zeros(r.shape)
    res[r_eff_safe < r_eff_max] = t * (1 - sf) + (1 - t) * cov
    return res


def fast_psd_inverse(M):
    """maria/utils/linalg.py:95-102."""
    cholesky, dpotrf_info = sp.linalg.lapack.dpotrf(M)
    invM, dpotri_info = sp.linalg.lapack.dpotri(cholesky)
    return np.where(invM, invM, invM.T)


def get_orthogonal_transform(signature, entries):
    """maria/utils/rotations.py:25-42."""
    axes = np.where(signature)[0]
    n_dim = len(signature)
    n_axes = sum(signature)
    i, j = np.triu_indices(n=n_axes, k=1)
    S = np.zeros((n_dim, n_dim))
    S[axes[i], axes[j]] = entries
    return sp.linalg.expm(S - S.T)


def compute_aligning_transform(points, signature, rng, n_init=16):
    """maria/utils/rotations.py:45-77 (n_dim == 3 branch)."""
    *input_shape, n_dim = points.shape
    args = points.reshape(-1, n_dim)

    def loss(entries, *a):
        tp = a[0] @ get_orthogonal_transform(signature=signature, entries=entries)
        if n_dim > 2:
            ch = sp.spatial.ConvexHull(tp[..., 1:])
            return np.log(ch.volume)
        return np.log(np.ptp(tp[..., 1:]))

    n_axes = sum(signature)
    n_dof = int(n_axes * (n_axes - 1) / 2)
    x0_samples = rng.standard_normal(size=(n_init, n_dof))
    best_index = np.argmin([loss(x0, args) for x0 in x0_samples])
    res = sp.optimize.minimize(loss, x0=x0_samples[best_index], args=args, tol=1e-6, method="SLSQP")
    if not res.success:
        raise RuntimeError("Could not find optimal rotation.")
    return get_orthogonal_transform(signature=signature, entries=res.x)


# --------------------------------------------------------------------------
# pointing
# --------------------------------------------------------------------------
def offsets_to_phi_theta(dX, cphi, ctheta, f32=False):
    """maria/coords/transforms.py:10-29.  With ``f32`` every array is narrowed to float32 first, which is
    what jax does with x64 disabled (the reference never enables it)."""
    ft = np.float32 if f32 else np.float64
    ct = np.complex64 if f32 else np.complex128
    dX = np.asarray(dX, dtype=ft)
    cphi = np.asarray(cphi, dtype=ft)
    ctheta = np.asarray(ctheta, dtype=ft)
    dx, dy = dX[..., 0], dX[..., 1]
    r = np.sqrt(dx**2 + dy**2)
    p = np.arctan2(-dx, -dy)
    proj_from_east = ((np.sin(r) * np.cos(p)).astype(ct) + (ct(1j) * np.cos(r))) * np.exp(
        ct(1j) * (ctheta - ft(np.pi / 2))
    ).astype(ct)
    phi = np.arctan2(np.sin(r) * np.sin(p), np.real(proj_from_east)) + cphi
    theta = np.arcsin(np.imag(proj_from_east))
    return phi.astype(ft), theta.astype(ft)


def get_rotation_matrix_2d(a):
    """maria/utils/rotations.py:5-8."""
    return np.array([[np.cos(a), -np.sin(a)], [np.sin(a), np.cos(a)]])


class Coords:
    """az/el subset of maria/coords/coordinates.py: t (n_t,), az/el (..., n_t), telescope at the origin."""

    def __init__(self, t, az, el, f32=False):
        self.t, self.az, self.el, self.f32 = np.asarray(t, dtype=float), az, el, f32

    def downsample(self, timestep):
        """coordinates.py:286-304."""
        ds_t = np.arange(self.t.min(), self.t.max(), timestep)
        kw = dict(axis=-1, bounds_error=False, fill_value="extrapolate")
        return Coords(ds_t, sp.interpolate.interp1d(self.t, self.az, **kw)(ds_t),
                      sp.interpolate.interp1d(self.t, self.el, **kw)(ds_t), self.f32)

    def broadcast(self, offsets):
        """coordinates.py:378-386."""
        az, el = offsets_to_phi_theta(offsets[..., None, :], self.az, self.el, f32=self.f32)
        return Coords(self.t, az, el, self.f32)

    def project(self, z):
        """coordinates.py:333-349 with self.z = 0 and self.xyz = 0; the result is float64 (the ones column)."""
        tan_theta = np.tan(self.el)[..., None]
        p = np.float64(z) * np.concatenate(
            [np.cos(self.az)[..., None] / tan_theta, np.sin(self.az)[..., None] / tan_theta,
             np.ones((*self.az.shape, 1))], axis=-1)
        return p


# --------------------------------------------------------------------------
# jax RegularGridInterpolator(method="linear", bounds_error=False, fill_value=nan)
# --------------------------------------------------------------------------
def rgi_linear(grid, values, xi, f32=False):
    """jax.scipy.interpolate.RegularGridInterpolator.__call__ (jax is not vendored in the reference; this
    restates its published algorithm: per axis i = clip(searchsorted(g, x) - 1, 0, n - 2),
    w = (x - g[i]) / (g[i+1] - g[i]); sum over the 2^d corners in itertools.product order of the value times
    prod(1-w | w); NaN where x < g[0] or x > g[-1] on any axis).  Call sites: atmosphere/atmosphere.py:359-366,
    band/band.py:255,283-286.  ``xi`` is a tuple of broadcastable arrays, one per axis."""
    import itertools

    ft = np.float32 if f32 else np.float64
    grid = [np.asarray(gr, dtype=ft) for gr in grid]
    values = np.asarray(values, dtype=ft)
    xs = np.broadcast_arrays(*[np.asarray(x, dtype=ft) for x in xi])
    shape = xs[0].shape
    xs = [x.ravel() for x in xs]
    indices, dists = [], []
    oob = np.zeros(xs[0].shape, dtype=bool)
    for x, gr in zip(xs, grid):
        i = np.searchsorted(gr, x) - 1
        i = np.where(i < 0, 0, i)
        i = np.where(i > gr.size - 2, gr.size - 2, i)
        indices.append(i)
        dists.append(((x - gr[i]) / (gr[i + 1] - gr[i])).astype(ft))
        oob |= x < gr[0]
        oob |= x > gr[-1]
    out = np.zeros(xs[0].shape, dtype=ft)
    for edge in itertools.product(*[[0, 1] for _ in indices]):
        weight = np.ones(xs[0].shape, dtype=ft)
        for e, w in zip(edge, dists):
            weight = weight * (w if e else (ft(1) - w))
        out = out + values[tuple(i + e for i, e in zip(indices, edge))] * weight
    out = np.where(oob, ft(np.nan), out)
    return out.reshape(shape)


# --------------------------------------------------------------------------
# a1: layer table
# --------------------------------------------------------------------------
MIN_RES = {"2d": 2, "3d": 15}
MIN_RES_PER_BEAM = {"2d": 0.1, "3d": 0.5}
MIN_RES_PER_FOV = {"2d": 0.02, "3d": 0.1}


def generate_layers(instrument, min_el, weather, site_altitude, mode="2d", angular=False, max_height=2e3,
                    pwv_rms_frac=3e-2):
    """maria/atmosphere/extrusion.py:27-110 ("2d" and "3d" branches).  Returns a dict of equal-length arrays
    (the reference's DataFrame columns)."""
    min_res, min_res_per_beam, min_res_per_fov = MIN_RES[mode], MIN_RES_PER_BEAM[mode], MIN_RES_PER_FOV[mode]
    h_samples = np.arange(0.0, 20000.0, 1e0)
    z_samples = h_samples / np.sin(min_el)
    fov = instrument.dets.field_of_view
    one = instrument.dets.one_detector_from_each_band()
    fwhm = physical_fwhm(one.primary_size, one.band_center, z_samples[:, None] + 1e-16).min(axis=1)
    r1 = min_res * np.ones(len(z_samples))
    r2 = min_res_per_beam * fwhm
    r3 = min_res_per_fov * z_samples * fov
    res_samples = np.minimum(1e3, np.maximum.reduce([r1, r2, r3]))

    def res_func(h):
        return sp.interpolate.interp1d(h_samples, res_samples)(h)

    if mode == "2d":
        h_boundaries = np.array([0.0, 500.0, 1000.0, 1500.0, 2000.0, 3000.0, 5000.0, 8000.0, 12000.0])
        process_index = np.arange(len(h_boundaries) - 1)
    else:
        h_boundaries = [0]
        while True:
            new_h = h_boundaries[-1] + res_func(h_boundaries[-1])
            if new_h > max_height:
                break
            h_boundaries.append(h_boundaries[-1] + res_func(h_boundaries[-1]))
        h_boundaries = np.array(h_boundaries)
        process_index = np.zeros(len(h_boundaries) - 1, dtype=int)
    h_centers = (h_boundaries[1:] + h_boundaries[:-1]) / 2
    layers = {k: np.asarray(v, dtype=float) for k, v in weather(altitude=site_altitude + h_centers).items()}
    layers["process_index"] = np.asarray(process_index)
    layers["h"] = h_centers
    layers["dh"] = np.diff(h_boundaries)
    layers["res"] = np.asarray(res_func(h_centers), dtype=float)
    layers["z"] = h_centers / np.sin(min_el)
    layers["angular"] = np.full(len(h_centers), angular)
    hb = [0, *(h_centers[:-1] + h_centers[1:]) / 2, 1e5]
    total_water = np.zeros(len(h_centers))
    for li, (h1, h2) in enumerate(zip(hb[:-1], hb[1:])):
        dummy_h = site_altitude + np.linspace(h1, h2, 1024)
        total_water[li] = np.trapezoid(np.interp(dummy_h, weather.altitude, weather.absolute_humidity), x=dummy_h)
    layers["total_water"] = total_water
    rel_var = (np.exp(-h_centers / 1e3) * h_centers ** (1 / 7)) ** 2
    pwv_var = (weather.pwv * pwv_rms_frac) ** 2 * rel_var / sum(rel_var)
    layers["pwv_rms"] = np.sqrt(pwv_var)
    if angular:
        layers["res"] = layers["res"] / layers["z"]
    return layers


# --------------------------------------------------------------------------
# a3-a5: autoregressive process
# --------------------------------------------------------------------------
class AutoregressiveProcess:
    """maria/atmosphere/process.py:19-209."""

    def __init__(self, cross_section, extrusion, callback_kwargs, jitter=1e-6, MIN_SAMPLES_PER_LAYER=4):
        self.cross_section, self.extrusion = cross_section, extrusion
        self.callback_kwargs, self.jitter = callback_kwargs, jitter
        self.n_cross_section, self.n_extrusion = len(cross_section), len(extrusion)
        extrusion_indices = [0, *(2 ** np.arange(0, np.log(self.n_extrusion) / np.log(2))).astype(int),
                             self.n_extrusion - 1]  # :44-48
        esi, csi = [], []
        for i, extrusion_index in enumerate(extrusion_indices):  # :52-67
            n_ribbon = np.minimum(np.maximum(int(self.n_cross_section * 2.0 ** -i), MIN_SAMPLES_PER_LAYER),
                                  self.n_cross_section)
            cs_idx = np.unique(np.linspace(0, self.n_cross_section - 1, n_ribbon).astype(int))
            csi.extend(cs_idx)
            esi.extend(np.repeat(extrusion_index, len(cs_idx)))
        self.cross_section_sample_index = np.array(csi)
        self.extrusion_sample_index = np.array(esi)
        # points[j, i] = (extrusion[j], cross_section[i, 0], cross_section[i, 1])  (:38-42)
        self.sample_points = np.c_[self.extrusion[self.extrusion_sample_index],
                                   self.cross_section[self.cross_section_sample_index]]  # :72-89
        self.n_sample = len(self.sample_points)
        self.live_edge_points = np.c_[
            self.extrusion[0] * np.ones(self.n_cross_section) - np.gradient(self.extrusion).mean(),
            self.cross_section]  # :91-100
        self.n_live_edge = self.n_cross_section

    def compute_covariance_matrices(self, rng):
        """process.py:111-185."""
        cb = lambda r: approximate_normalized_matern(r, **self.callback_kwargs)  # noqa: E731
        i, j = np.triu_indices(self.n_live_edge, k=1)
        COV_E_E = np.eye(self.n_live_edge) + self.jitter
        COV_E_E[i, j] = cb(np.sqrt(np.square(self.live_edge_points[j] - self.live_edge_points[i]).sum(axis=1)))
        COV_E_E[j, i] = COV_E_E[i, j]
        COV_E_E += np.diag(COV_MAT_JITTER * np.diag(COV_E_E))
        COV_E_S = cb(np.sqrt(np.square(self.sample_points[None] - self.live_edge_points[:, None]).sum(axis=2)))
        i, j = np.triu_indices(self.n_sample, k=1)
        COV_S_S = np.eye(self.n_sample) + self.jitter
        COV_S_S[i, j] = cb(np.sqrt(np.square(self.sample_points[j] - self.sample_points[i]).sum(axis=1)))
        COV_S_S[j, i] = COV_S_S[i, j]
        COV_S_S += np.diag(COV_MAT_JITTER * np.diag(COV_S_S))
        inv_COV_S_S = fast_psd_inverse(COV_S_S)
        self.A = COV_E_S @ inv_COV_S_S
        if (self.A.sum(axis=-1) > 1.0).any():
            raise ValueError(f"Propagation operator is unstable (A_max = {self.A.sum(axis=-1).max()}).")
        self.B = np.linalg.cholesky(COV_E_E - self.A @ COV_E_S.T)
        initial_slice = self.B @ rng.standard_normal(self.n_cross_section)  # :182 (overwritten by run())
        initial_slice *= np.sqrt(np.diag(COV_E_E) / initial_slice.var())
        self.values = np.zeros((self.n_extrusion, self.n_cross_section))
        self.values[:] = initial_slice

    def run(self, rng, record=None):
        """process.py:191-209.  ``record`` (a dict) receives the drawn noise: BUFFER0 (3 n_ext, n_cs) and
        Z (2 n_ext, n_cs) with Z[k] the vector drawn for the k-th executed step (buffer_index 2 n_ext-1-k)."""
        n_steps = 2 * self.n_extrusion
        BUFFER = rng.standard_normal((self.n_extrusion + n_steps, self.n_cross_section))
        if record is not None:
            record["BUFFER0"] = BUFFER.copy()
            record["Z"] = np.zeros((n_steps, self.n_cross_section))
        for k, buffer_index in enumerate(np.arange(n_steps)[::-1]):
            z = rng.standard_normal(size=self.n_live_edge)
            if record is not None:
                record["Z"][k] = z
            BUFFER[buffer_index] = self.A @ BUFFER[
                buffer_index + self.extrusion_sample_index + 1, self.cross_section_sample_index] + self.B @ z
        self.values = BUFFER[: self.n_extrusion]
        return self.values


class ReplayRNG:
    """An ``rng`` that replays pre-drawn standard-normal arrays in order (validation mode)."""

    def __init__(self, arrays):
        self.arrays, self.k = list(arrays), 0

    def standard_normal(self, size=None):
        a = self.arrays[self.k]
        self.k += 1
        assert size is None or tuple(np.atleast_1d(size)) == a.shape or np.prod(size) == a.size, (size, a.shape)
        return a.reshape(size) if size is not None else a


# --------------------------------------------------------------------------
# a2, a8, a9: Atmosphere.initialize / simulate_pwv
# --------------------------------------------------------------------------
class Observation:
    """The fields of maria/sim/observation.py:27-102 the path reads."""

    def __init__(self, instrument, plan, site, f32=False):
        self.instrument, self.plan, self.site = instrument, plan, site
        self.boresight = Coords(plan.time, plan.phi, plan.theta, f32)  # observation.py:46-52
        self.offsets = instrument.dets.offsets @ get_rotation_matrix_2d(np.radians(plan.roll)).T  # :55-58
        self.coords = self.boresight.broadcast(self.offsets)  # full rate, (n_det, n_t)
        self.loading = {}


class Atmosphere:
    """maria/atmosphere/atmosphere.py:31-380, "2d"/"3d", non-angular.  ``weather`` and ``spectrum`` are passed in
    ready-made (the reference loads them from h5 in __init__, :53-65)."""

    def __init__(self, weather, spectrum, model="2d", max_height=5e3, timestep=None, f32=False):
        if model not in ["2d", "3d"]:
            raise ValueError(f"Invalid model '{model}'. Supported models are ['2d', '3d'].")
        self.weather, self.spectrum, self.model = weather, spectrum, model
        self.max_height, self.timestep, self.f32 = max_height, timestep, f32
        self.angular = False
        self.interpolation_method = "linear"
        self._initialized = False

    def initialize(self, obs, rng):
        dets = obs.instrument.dets
        self.layers = generate_layers(obs.instrument, obs.boresight.el.min(), self.weather, obs.site.altitude,
                                      mode=self.model, angular=self.angular, max_height=self.max_height)  # :86-94
        L = self.layers
        if self.timestep is None:  # :96-99
            min_fwhm = angular_fwhm(dets.primary_size, dets.band_center, z=self.max_height).min()
            max_wind = (L["wind_speed"] / L["h"]).max()
            self.timestep = max(1e-1, min_fwhm / max_wind)
        self.boresight = obs.boresight.downsample(timestep=self.timestep)  # :101
        self.coords = self.boresight.broadcast(dets.offsets)  # :102-105  (note: un-rolled offsets, as the reference)
        outer_coords = self.boresight.broadcast(dets.offsets[outer_detector_mask(dets.offsets)])  # :108-111
        self.processes = {}
        n_layers = len(L["h"])
        for process_index in sorted(np.unique(L["process_index"])):
            idx = np.where(L["process_index"] == process_index)[0]
            n_t = self.boresight.t.shape[-1]
            vx = L["wind_east"][idx][:, None] * np.ones(n_t)  # :149-154
            vy = L["wind_north"][idx][:, None] * np.ones(n_t)
            w = L["absolute_humidity"][idx] * L["temperature"][idx] * L["divergence"][idx]  # :156-159
            vx = (w[:, None] * vx).sum(axis=0) / w.sum()
            vy = (w[:, None] * vy).sum(axis=0) / w.sum()
            vz = np.zeros(vx.shape)
            hull_idx = idx[[0, -1]] if len(idx) > 1 else idx  # :161-162
            pts = []
            for li in hull_idx:  # :169-172
                p = outer_coords.project(z=L["h"][li])
                p = p + np.cumsum(self.timestep * np.c_[vx, vy, vz][None], axis=-2)
                pts.append(p[None])
            pts = np.concatenate(pts, axis=0).reshape(-1, 3)
            pts[..., 2] += 1e-6 * rng.standard_normal(pts[..., 2].shape)  # :184-186
            transform = compute_aligning_transform(pts, signature=(True, True, False), rng=rng)  # :191-194
            tp = pts @ transform
            min_tx, max_tx = tp[:, 0].min(), tp[:, 0].max()
            cs_list, labels = [], []
            for li in idx:  # :203-219
                res = L["res"][li]
                n_cs = int(np.maximum(2, (np.ptp(tp[:, 1]) + 2 * res) / res))
                cs_list.append(np.stack([np.linspace(tp[:, 1].min() - res, tp[:, 1].max() + res, n_cs),
                                         L["h"][li] * np.ones(n_cs)]).T)
                labels.extend(n_cs * [li])
            cross_section_points = np.concatenate(cs_list, axis=0)
            res_min = L["res"].min()
            extrusion_points = np.arange(min_tx - 2 * res_min, max_tx + 2 * res_min, res_min)  # :241-245
            outer_scale = np.maximum(1e3, 300 + L["h"][hull_idx].mean() / 10)  # :247
            matern_kwargs = {"nu": 1 / 3, "r0": outer_scale} if self.model == "3d" else {"nu": 5 / 6, "r0": outer_scale}
            process = AutoregressiveProcess(cross_section_points, extrusion_points, matern_kwargs, jitter=1e-8)
            process.layer_index = idx
            process.labels = np.array(labels)
            process.transform = transform
            process.extrusion_res = np.gradient(extrusion_points).mean()
            process.vx, process.vy, process.tp = vx, vy, tp
            process.compute_covariance_matrices(rng)  # :267-277 (the retry loop never changes the jitter)
            self.processes[int(process_index)] = process
        self._initialized = True

    def simulate_pwv(self, instrument, rng, record=None):
        """atmosphere.py:293-380.  ``record`` (dict) receives per-process noise and the smoothed layers."""
        if not self._initialized:
            raise RuntimeError("Atmosphere must be initialized with a simulation.")
        pp = self.coords.project(z=1)  # :297
        for k, process in self.processes.items():  # :299-307
            rec = {} if record is not None else None
            process.run(rng, record=rec)
            if record is not None:
                record.setdefault("noise", {})[k] = rec
        self.zenith_scaled_pwv = self.weather.pwv * np.ones(shape=pp.shape[:-1])  # :309
        L = self.layers
        for k, process in self.processes.items():
            wind_vector = np.c_[process.vx, process.vy, np.zeros(process.vx.shape)]
            translation = np.cumsum(self.timestep * wind_vector[None], axis=-2)  # :318-319
            for li in process.layer_index:
                layer_mask = process.labels == li
                if instrument is not None:  # :328-339
                    d = instrument.dets
                    beam_fwhm = physical_fwhm(d.primary_size, d.band_center, L["z"][li]).mean()
                else:
                    beam_fwhm = 0
                beam_sigma = beam_fwhm / 2.355
                extrusion_sigma = beam_sigma / process.extrusion_res
                x_sigma = beam_sigma / L["res"][li]
                smoothed = sp.ndimage.gaussian_filter(process.values[:, layer_mask], sigma=(extrusion_sigma, x_sigma))
                if record is not None:
                    record.setdefault("smoothed", {})[int(li)] = smoothed
                p = L["h"][li] * pp + translation  # :346
                tl = p @ process.transform  # :347
                y = rgi_linear((process.extrusion, process.cross_section[layer_mask, 0]), smoothed,
                               (tl[..., 0], tl[..., 1]), f32=self.f32)  # :359-366
                if np.isnan(y).any():
                    raise RuntimeError(f"Layer {li} in process {k} introduced nans into PWV simulation.")
                self.zenith_scaled_pwv += L["pwv_rms"][li] * y  # :373
        return self.zenith_scaled_pwv


# --------------------------------------------------------------------------
# a10-a14: sim mixin, gain, TOD.to("K_RJ")
# --------------------------------------------------------------------------
def band_emission_grid(band, spectrum):
    """maria/band/band.py:272-280: (nT, npwv, nel) grid in pW."""
    return 1e12 * k_B * np.trapezoid(spectrum._emission * band.passband(spectrum.side_nu), spectrum.side_nu, axis=-1)


def band_transmission_grid(band, spectrum):
    """maria/band/band.py:248-253 with nu_min = 0, nu_max = inf: (nT, npwv, nel)."""
    nu = spectrum.side_nu
    return np.trapezoid(y=band.passband(nu) * np.exp(-spectrum._opacity), x=nu, axis=-1)


def simulate_atmosphere(obs, atmosphere, rng, record=None):
    """maria/sim/atmosphere.py:24-37."""
    atmosphere.simulate_pwv(obs.instrument, rng, record=record)
    obs.zenith_scaled_pwv = sp.interpolate.interp1d(
        atmosphere.coords.t, atmosphere.zenith_scaled_pwv, bounds_error=False, fill_value="extrapolate")(obs.coords.t)


def compute_atmospheric_loading(obs, atmosphere, dtype=np.float32, record=None):
    """maria/sim/atmosphere.py:39-84."""
    f32 = atmosphere.f32
    loading = np.zeros(atmosphere.zenith_scaled_pwv.shape, dtype=dtype)
    dets = obs.instrument.dets
    sp_ = atmosphere.spectrum
    for band in dets.bands:
        band_index = dets.mask(band_name=band.name)
        values = band_emission_grid(band, sp_)
        el = atmosphere.coords.el[band_index].clip(max=np.pi / 2)
        p = rgi_linear(sp_.points[:3], values,
                       (atmosphere.weather.temperature[0], atmosphere.zenith_scaled_pwv[band_index], el), f32=f32)
        loading[band_index] = dets.mueller00()[band_index][:, None] * p
    if record is not None:
        record["loading_atm_rate"] = loading.copy()
    out = sp.interpolate.interp1d(atmosphere.coords.t, loading, kind="cubic", bounds_error=False,
                                  fill_value="extrapolate", axis=-1)(obs.coords.t)
    return out.astype(dtype)


def run_obs(obs, atmosphere, rng, dtype=np.float32, record=None):
    """maria/sim/simulation.py:213-272, atmosphere branch: returns (loading_pW (n_det, n_t) of ``dtype``, metadata)."""
    simulate_atmosphere(obs, atmosphere, rng, record=record)
    loading = compute_atmospheric_loading(obs, atmosphere, dtype=dtype, record=record)
    dets = obs.instrument.dets
    gain_error = np.exp(dets.gain_error * rng.standard_normal(size=dets.n))  # :239-241
    loading *= gain_error[:, None].astype(dtype) if dtype == np.float32 else gain_error[:, None]  # :243-247
    metadata = {"atmosphere": True, "altitude": float(obs.site.altitude), "region": obs.site.region,
                "pwv": float(np.round(atmosphere.weather.pwv, 3)),
                "base_temperature": float(np.round(atmosphere.weather.temperature[0], 3))}
    if record is not None:
        record["gain_error"] = gain_error
    return loading, metadata


def tod_to_K_RJ(data_pW, obs, atmosphere, metadata, out_dtype=np.float32):
    """maria/tod/tod.py:90-142 -> calibration/calibration.py:108-118 -> calibration/functions.py:73-90 ->
    band/band.py:235-255, for the "pW -> K_RJ" edge.  ``data_pW`` is TOD.data["atmosphere"] (n_det, n_t)."""
    f32 = atmosphere.f32
    dets = obs.instrument.dets
    out = np.array(data_pW, dtype=out_dtype)  # content(): deep copy in the TOD dtype
    for band in dets.bands:
        band_mask = dets.band_name == band.name
        if band_mask.sum() == 0:
            continue
        polarized = ~np.isnan(dets.gamma[band_mask]).all()
        grid = band_transmission_grid(band, atmosphere.spectrum)
        integral = rgi_linear(atmosphere.spectrum.points[:3], grid,
                              (metadata["base_temperature"], metadata["pwv"], obs.coords.el[band_mask]), f32=f32)
        P = np.asarray(data_pW[band_mask], dtype=np.float64) * 1e-12  # Quantity(x, "pW").base_units_value
        out[band_mask] = P / ((0.5 if polarized else 1.0) * k_B * integral)
    return out


def simulate(instrument, plan, site, weather, spectrum, rng, model="2d", f32=False, dtype=np.float32,
             atmosphere_kwargs=None, record=None):
    """Simulation(instrument, plan, site, atmosphere=model).run(units="K_RJ") for the atmosphere field
    (maria/sim/simulation.py:76-211).  Returns a dict with the K_RJ TOD and the intermediates."""
    obs = Observation(instrument, plan, site, f32=f32)
    atm = Atmosphere(weather, spectrum, model=model, f32=f32, **(atmosphere_kwargs or {}))
    atm.initialize(obs, rng)
    loading, metadata = run_obs(obs, atm, rng, dtype=dtype, record=record)
    tod = tod_to_K_RJ(loading, obs, atm, metadata, out_dtype=dtype)
    return dict(obs=obs, atmosphere=atm, loading_pW=loading, tod_K_RJ=tod, metadata=metadata)


def atmosphere_from_geometry(weather, spectrum, layers, timestep, boresight_t, boresight_az, boresight_el, det_offsets,
                             processes, f32=False):
    """An initialised oracle ``Atmosphere`` from already-computed set-up geometry (layer table, downsampled boresight,
    per-process stencil / A / B / transform), so that the CPU baseline of ``bench.py`` times stages a5-a14 without
    repeating the footprint optimisation at 100k-detector scale.  ``processes``: {k: dict(cross_section, extrusion, A, B,
    transform, vx, vy, layer_index, labels, extrusion_res)}."""
    atm = Atmosphere(weather, spectrum, f32=f32)
    atm.layers, atm.timestep = layers, timestep
    atm.boresight = Coords(boresight_t, boresight_az, boresight_el, f32)
    atm.coords = atm.boresight.broadcast(det_offsets)
    atm.processes = {}
    for k, g in processes.items():
        p = AutoregressiveProcess(g["cross_section"], g["extrusion"], {}, jitter=1e-8)
        p.A, p.B = g["A"], g["B"]
        p.transform, p.vx, p.vy = g["transform"], g["vx"], g["vy"]
        p.layer_index, p.labels, p.extrusion_res = g["layer_index"], g["labels"], g["extrusion_res"]
        atm.processes[k] = p
    atm._initialized = True
    return atm


# ---------------------------------------------------------------------------------------------------------------
# consumers of the path's intermediates
# ---------------------------------------------------------------------------------------------------------------
def map_pW_per_K_RJ(obs, atmosphere, nu_min_Hz, nu_max_Hz):
    """maria/sim/map.py:117-133 with band.py:235-255: per band, 1e12 k_B x the trilinear lookup of
    trapz(passband(nu) exp(-opacity[..., nu]), nu) over spectrum.side_nu in [nu_min, nu_max) at (T0, full-rate
    zenith_scaled_pwv, full-rate elevation).  (n_det, n_t) float64; ``simulate_atmosphere`` must have run."""
    dets, sp_ = obs.instrument.dets, atmosphere.spectrum
    out = np.zeros(obs.zenith_scaled_pwv.shape)
    nu_mask = (sp_.side_nu >= nu_min_Hz) & (sp_.side_nu < nu_max_Hz)
    nu = sp_.side_nu[nu_mask]
    for band in dets.bands:
        m = dets.mask(band_name=band.name)
        grid = np.trapezoid(band.passband(nu) * np.exp(-sp_._opacity[..., nu_mask]), x=nu, axis=-1)
        xi = (atmosphere.weather.temperature[0], obs.zenith_scaled_pwv[m], obs.coords.el[m])
        out[m] = 1e12 * k_B * rgi_linear(sp_.points[:3], grid, xi, f32=atmosphere.f32)
    return out


def noise_loading_pW(obs, loading_pW, unscaled_noise, nep, nep_per_loading):
    """maria/sim/noise.py:35-37,60: per band, 1e12 (NEP + NEP_per_loading * sum of the loadings) * unscaled_noise.
    ``nep`` / ``nep_per_loading``: {band name: W sqrt(s)}."""
    dets = obs.instrument.dets
    out = np.zeros(np.shape(loading_pW))
    for band in dets.bands:
        m = dets.mask(band_name=band.name)
        total = nep[band.name] + nep_per_loading[band.name] * np.asarray(loading_pW, dtype=np.float64)[m]
        out[m] = 1e12 * total * np.asarray(unscaled_noise, dtype=np.float64)[m]
    return out
