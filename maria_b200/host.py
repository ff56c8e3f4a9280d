"""Host-side set-up numerics of the atmosphere path: layer table, stencil,
Matern covariances, the propagator matrices A and B, and the footprint-aligning rotation.  These run once per plan in
float64 numpy / LAPACK, as in the reference; everything per detector-sample runs on the device.

Every function cites the reference lines it mirrors (``maria/<file>:<lines>``).
"""
from __future__ import annotations

import numpy as np
import scipy as sp
import scipy.interpolate
import scipy.linalg
import scipy.optimize
import scipy.spatial
import scipy.special

C_LIGHT = 2.99792458e8
K_B = 1.380649e-23
COV_MAT_JITTER = 1e-6  # maria/atmosphere/process.py:16

MIN_RES = {"2d": 2, "3d": 15}            # maria/atmosphere/extrusion.py:20-22
MIN_RES_PER_BEAM = {"2d": 0.1, "3d": 0.5}
MIN_RES_PER_FOV = {"2d": 0.02, "3d": 0.1}
LAYER_BOUNDARIES_2D = (0.0, 500.0, 1000.0, 1500.0, 2000.0, 3000.0, 5000.0, 8000.0, 12000.0)  # extrusion.py:66-68


# ---------------------------------------------------------------------------------------------------------------
# beams
# ---------------------------------------------------------------------------------------------------------------
def gaussian_beam_angular_fwhm(fwhm_0, z, nu, n=1.0):
    """Gaussian-beam width at distance z (maria/beam/__init__.py:9-25)."""
    waist = fwhm_0 / 2
    rayleigh = np.pi * waist**2 * n / (C_LIGHT / nu)
    with np.errstate(divide="ignore"):
        return 2 * waist * np.sqrt(1 / np.asarray(z, dtype=float) ** 2 + 1 / rayleigh**2)


def gaussian_beam_physical_fwhm(fwhm_0, z, nu):
    """maria/array/array.py:229-233."""
    return z * gaussian_beam_angular_fwhm(fwhm_0, z, nu)


# ---------------------------------------------------------------------------------------------------------------
# a1: layer table (maria/atmosphere/extrusion.py:27-110)
# ---------------------------------------------------------------------------------------------------------------
def generate_layers(dets, min_el, weather, site_altitude, mode="2d", angular=False, max_height=2e3, pwv_rms_frac=3e-2):
    """Returns a dict of equal-length float arrays, one entry per layer (the reference's DataFrame columns)."""
    h_grid = np.arange(0.0, 20000.0, 1.0)
    z_grid = h_grid / np.sin(min_el)
    # one detector from each band (array.py:149-154)
    _, first = np.unique(dets.band_name, return_index=True)
    first = np.sort(first)
    beam = gaussian_beam_physical_fwhm(dets.primary_size[first], z_grid[:, None] + 1e-16, dets.band_center[first]).min(axis=1)
    res_grid = np.minimum(1e3, np.maximum.reduce([
        MIN_RES[mode] * np.ones(len(z_grid)), MIN_RES_PER_BEAM[mode] * beam, MIN_RES_PER_FOV[mode] * z_grid * dets.field_of_view]))
    res_of = sp.interpolate.interp1d(h_grid, res_grid)

    if mode == "2d":
        edges = np.array(LAYER_BOUNDARIES_2D)
        process_index = np.arange(len(edges) - 1)
    else:  # "3d": one coupled process, layer spacing follows the resolution (extrusion.py:69-77)
        edges = [0.0]
        while edges[-1] + res_of(edges[-1]) <= max_height:
            edges.append(edges[-1] + float(res_of(edges[-1])))
        edges = np.array(edges)
        process_index = np.zeros(len(edges) - 1, dtype=int)
    h = 0.5 * (edges[1:] + edges[:-1])

    layers = {k: np.asarray(v, dtype=float) for k, v in weather(altitude=site_altitude + h).items()}
    layers.update(process_index=np.asarray(process_index), h=h, dh=np.diff(edges), res=np.asarray(res_of(h), dtype=float),
                  z=h / np.sin(min_el), angular=np.full(len(h), angular))
    cuts = [0, *(0.5 * (h[:-1] + h[1:])), 1e5]  # extrusion.py:91-98
    water = np.empty(len(h))
    for i in range(len(h)):
        col = site_altitude + np.linspace(cuts[i], cuts[i + 1], 1024)
        water[i] = np.trapezoid(np.interp(col, weather.altitude, weather.absolute_humidity), x=col)
    layers["total_water"] = water
    rel_var = (np.exp(-h / 1e3) * h ** (1 / 7)) ** 2  # extrusion.py:100-105
    layers["pwv_rms"] = np.sqrt((weather.pwv * pwv_rms_frac) ** 2 * rel_var / sum(rel_var))
    if angular:
        layers["res"] = layers["res"] / layers["z"]
    return layers


# ---------------------------------------------------------------------------------------------------------------
# Matern covariance (maria/functions/__init__.py:30-74)
# ---------------------------------------------------------------------------------------------------------------
class MaternTable:
    """Log-log table of the normalised Matern covariance; ``__call__(r)`` interpolates it.  Building the 1024 Bessel
    samples once per (nu) instead of on every call is the only difference to the reference's function."""

    R_MIN, R_MAX, N = 1e-6, 1e3, 1024

    def __init__(self, nu, r0):
        self.nu, self.r0 = nu, r0
        x = np.geomspace(self.R_MIN, self.R_MAX, self.N)
        arg = np.sqrt(2 * nu) * x + 1e-16
        cov = 2 ** (1 - nu) / sp.special.gamma(nu) * sp.special.kv(nu, arg) * arg**nu
        self._logx = np.log(x)
        with np.errstate(divide="ignore"):
            self._log_cov, self._log_sf = np.log(cov), np.log(1 - cov)

    def __call__(self, r):
        r = np.asarray(r, dtype=float)
        x = np.atleast_1d(np.abs(r / self.r0)).clip(min=self.R_MIN)
        inside = x < self.R_MAX
        xi = x[inside]
        lx = np.log(xi)
        sf = np.exp(np.interp(lx, self._logx, self._log_sf))
        cov = np.exp(np.interp(lx, self._logx, self._log_cov))
        t = 1 / (1 + xi**2)
        out = np.zeros(r.shape)
        out[inside] = t * (1 - sf) + (1 - t) * cov
        return out


# ---------------------------------------------------------------------------------------------------------------
# a3 + a4: stencil and propagator (maria/atmosphere/process.py:20-189)
# ---------------------------------------------------------------------------------------------------------------
def build_stencil(n_ext, n_cs, min_samples_per_row=4):
    """Rows {0, 1, 2, 4, ..., < n_ext, n_ext - 1}; row i keeps clamp(int(n_cs 2^-i), 4, n_cs) evenly spaced, distinct
    columns (process.py:44-67).  Returns (extrusion_sample_index, cross_section_sample_index)."""
    rows = [0, *(2 ** np.arange(0, np.log(n_ext) / np.log(2))).astype(int), n_ext - 1]
    esi, csi = [], []
    for i, row in enumerate(rows):
        keep = int(np.minimum(np.maximum(int(n_cs * 2.0**-i), min_samples_per_row), n_cs))
        cols = np.unique(np.linspace(0, n_cs - 1, keep).astype(int))
        csi.append(cols)
        esi.append(np.full(len(cols), row))
    return np.concatenate(esi), np.concatenate(csi)


def psd_inverse(M):
    """Inverse of a positive-definite matrix through dpotrf + dpotri (maria/utils/linalg.py:95-102)."""
    chol, info = sp.linalg.lapack.dpotrf(M)
    if info != 0:
        raise np.linalg.LinAlgError(f"dpotrf failed (info={info})")
    inv, info = sp.linalg.lapack.dpotri(chol)
    if info != 0:
        raise np.linalg.LinAlgError(f"dpotri failed (info={info})")
    return np.where(inv, inv, inv.T)


def _pairwise_cov(points, cov, jitter):
    n = len(points)
    i, j = np.triu_indices(n, k=1)
    M = np.eye(n) + jitter
    M[i, j] = cov(np.sqrt(np.square(points[j] - points[i]).sum(axis=1)))
    M[j, i] = M[i, j]
    M += np.diag(COV_MAT_JITTER * np.diag(M))
    return M


def propagator_matrices(sample_points, live_edge_points, cov, jitter):
    """A = C_ES C_SS^-1 and B = chol(C_EE - A C_ES^T) (process.py:111-174).  Raises ValueError when a row of A sums to
    more than one (unstable propagator, :169-172) and LinAlgError when a factorisation fails."""
    C_EE = _pairwise_cov(live_edge_points, cov, jitter)
    C_ES = cov(np.sqrt(np.square(sample_points[None] - live_edge_points[:, None]).sum(axis=2)))
    C_SS = _pairwise_cov(sample_points, cov, jitter)
    A = C_ES @ psd_inverse(C_SS)
    row_sum = A.sum(axis=-1)
    if (row_sum > 1.0).any():
        raise ValueError(f"Propagation operator is unstable (A_max = {row_sum.max()}).")
    B = np.linalg.cholesky(C_EE - A @ C_ES.T)
    return A, B, C_EE


# ---------------------------------------------------------------------------------------------------------------
# footprint alignment (maria/utils/rotations.py:5-8,25-77)
# ---------------------------------------------------------------------------------------------------------------
def rotation_matrix_2d(a):
    return np.array([[np.cos(a), -np.sin(a)], [np.sin(a), np.cos(a)]])


def orthogonal_transform(signature, entries):
    """expm of the antisymmetric generator over the axes selected by ``signature`` (rotations.py:25-42)."""
    axes = np.where(signature)[0]
    n_dim, n_axes = len(signature), int(np.sum(signature))
    i, j = np.triu_indices(n=n_axes, k=1)
    S = np.zeros((n_dim, n_dim))
    S[axes[i], axes[j]] = entries
    return sp.linalg.expm(S - S.T)


def aligning_transform(points, signature=(True, True, False), rng=np.random, n_init=16, reduce_to_hull=False):
    """Rotation minimising the hull area of the projected footprint (rotations.py:45-77; 16 random starts, SLSQP).

    ``reduce_to_hull`` first replaces the points by the vertices of their 3-D convex hull: the projected hull, hence the
    loss, is mathematically unchanged, and each of the ~100 loss evaluations touches hundreds of points instead of
    millions.  The loss then differs by rounding only, so the optimiser may stop a hair (< its 1e-6 tolerance) away from
    where the reference stops; exact-parity runs keep it off."""
    pts = np.asarray(points, dtype=float).reshape(-1, points.shape[-1])
    n_dim = pts.shape[-1]
    if reduce_to_hull and n_dim == 3 and len(pts) > 64:
        pts = pts[sp.spatial.ConvexHull(pts).vertices]

    def loss(entries):
        tp = pts @ orthogonal_transform(signature, entries)
        if n_dim > 2:
            return np.log(sp.spatial.ConvexHull(tp[..., 1:]).volume)
        return np.log(np.ptp(tp[..., 1:]))

    n_axes = int(np.sum(signature))
    starts = rng.standard_normal(size=(n_init, n_axes * (n_axes - 1) // 2))
    best = starts[int(np.argmin([loss(x0) for x0 in starts]))]
    res = sp.optimize.minimize(lambda x, *a: loss(x), x0=best, args=(pts,), tol=1e-6, method="SLSQP")
    if not res.success:
        raise RuntimeError("Could not find optimal rotation.")
    return orthogonal_transform(signature, res.x)


def outer_detector_mask(offsets):
    """Detectors on the convex hull of the focal plane (maria/array/array.py:156-162)."""
    try:
        hull = sp.spatial.ConvexHull(offsets)
    except sp.spatial.QhullError:
        return np.ones(len(offsets), dtype=bool)
    return np.isin(np.arange(len(offsets)), hull.vertices)


# ---------------------------------------------------------------------------------------------------------------
# pointing on the host, for the footprint hull only (a few hundred detectors; coords/transforms.py:10-29)
# ---------------------------------------------------------------------------------------------------------------
def offsets_to_az_el(offsets, az_b, el_b):
    dx, dy = offsets[..., 0, None], offsets[..., 1, None]
    r = np.sqrt(dx**2 + dy**2)
    p = np.arctan2(-dx, -dy)
    q = (np.sin(r) * np.cos(p) + 1j * np.cos(r)) * np.exp(1j * (el_b - np.pi / 2))
    return np.arctan2(np.sin(r) * np.sin(p), q.real) + az_b, np.arcsin(q.imag)


def project_to_height(az, el, z):
    """coordinates.py:333-349 for a telescope at the origin."""
    tan_el = np.tan(el)
    return z * np.stack([np.cos(az) / tan_el, np.sin(az) / tan_el, np.ones(az.shape)], axis=-1)
