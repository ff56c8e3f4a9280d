"""CPU tests of the host side: the set-up code of ``maria_b200`` (layer table, stencil, propagator matrices, geometry)
against the reference's golden vectors and the oracle; the reference's constructor / error behaviour; the C-ABI library
(loads, exports every symbol ``include/maria_atm.h`` declares); and that there is no CPU fallback."""
import ctypes
import os
import re

import numpy as np
import pytest

import maria_b200 as M
from common import load_case, relmax, small_case
from maria_b200 import _build, _lib, host
from oracle import atm_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- host set-up against the reference's golden vectors -----------------------------------------------------------
@pytest.fixture(scope="module", params=["m2_daisy", "1deg_ces"])
def host_setup(request):
    case = load_case(request.param)
    np.random.seed(case["seed"])
    obs = M.Observation(case["instrument"], case["plan"], case["site"])
    atm = M.Atmosphere(weather=case["weather"], spectrum=case["spectrum"], geometry="exact")
    atm.initialize(obs, bind=False)  # host only: no device is touched
    return case, obs, atm


def test_host_layer_table_matches_reference(host_setup):
    case, _, atm = host_setup
    fx = case["fx"]
    assert relmax(atm.layers["h"], fx["ref_layers_h"]) < 1e-12
    assert relmax(atm.layers["res"], fx["ref_layers_res"]) < 1e-12
    assert relmax(atm.layers["pwv_rms"], fx["ref_layers_pwv_rms"]) < 1e-12
    assert abs(atm.timestep - float(fx["ref_timestep"])) < 1e-12 * atm.timestep


def test_host_geometry_matches_reference(host_setup):
    case, _, atm = host_setup
    fx = case["fx"]
    assert sorted(atm.processes) == list(range(8))
    for k, p in atm.processes.items():
        assert relmax(p.transform, fx[f"ref_transform_{k}"]) < 1e-9
        assert p.extrusion.shape == fx[f"ref_extrusion_{k}"].shape
        assert relmax(p.extrusion, fx[f"ref_extrusion_{k}"]) < 1e-9
        assert relmax(p.cross_section, fx[f"ref_cross_section_{k}"]) < 1e-9


def test_host_propagators_match_oracle(host_setup):
    """A, B and the stencil of every process equal the oracle's (process.py:44-67,111-174)."""
    case, _, atm = host_setup
    np.random.seed(case["seed"])
    o_obs = O.Observation(case["instrument"], case["plan"], case["site"])
    o_atm = O.Atmosphere(case["weather"], case["spectrum"])
    o_atm.initialize(o_obs, np.random)
    for k, p in atm.processes.items():
        q = o_atm.processes[k]
        assert np.array_equal(p.extrusion_sample_index, q.extrusion_sample_index)
        assert np.array_equal(p.cross_section_sample_index, q.cross_section_sample_index)
        assert relmax(p.A, q.A) < 1e-9 and relmax(p.B, q.B) < 1e-9
        assert np.all(p.A.sum(axis=-1) <= 1.0)
        n_near = int(np.sum(p.extrusion_sample_index + 1 <= 3))
        assert p.n_near == n_near and 0 < n_near <= p.n_sample


def test_stencil_shapes_and_edge_sizes():
    for n_ext, n_cs in [(4, 2), (5, 3), (64, 4), (100, 37), (1000, 251)]:
        esi, csi = host.build_stencil(n_ext, n_cs, 4)
        assert esi[0] == 0 and esi[-1] == n_ext - 1 and np.all(np.diff(esi) >= 0)
        assert csi.min() == 0 and csi.max() == n_cs - 1
        first_row = csi[esi == 0]
        assert np.array_equal(first_row, np.arange(n_cs))  # row 0 keeps every column
        # identical to the oracle's restatement of process.py:44-67 (which may repeat row n_ext - 1, as the reference)
        q = O.AutoregressiveProcess(np.zeros((n_cs, 2)), np.arange(n_ext, dtype=float), dict(nu=5 / 6, r0=1e3))
        assert np.array_equal(esi, q.extrusion_sample_index) and np.array_equal(csi, q.cross_section_sample_index)


def test_matern_table_equals_oracle_callback():
    r = np.geomspace(1e-3, 3e4, 500)
    tab = host.MaternTable(nu=5 / 6, r0=1200.0)
    assert relmax(tab(r), O.approximate_normalized_matern(r, nu=5 / 6, r0=1200.0)) < 1e-13


def test_pointing_equals_oracle():
    rng = np.random.default_rng(5)
    off = np.radians(rng.uniform(-1, 1, (50, 2)))
    az_b, el_b = np.radians(rng.uniform(0, 360, 40)), np.radians(rng.uniform(20, 80, 40))
    az, el = host.offsets_to_az_el(off, az_b, el_b)
    phi, theta = O.offsets_to_phi_theta(off[:, None, :], az_b[None], el_b[None])
    assert np.max(np.abs(el - theta)) < 1e-13
    assert np.max(np.abs(np.angle(np.exp(1j * (az - phi))))) < 1e-13


# ---- the reference's constructor and error behaviour (atmosphere.py:32-51,294-295) --------------------------------
def test_atmosphere_constructor_contract():
    case = small_case(n_det=8, duration=5.0, sample_rate=10.0)
    with pytest.raises(ValueError, match="Invalid model"):
        M.Atmosphere(model="4d", weather=case["weather"], spectrum=case["spectrum"])
    atm = M.Atmosphere(model="2d", weather=case["weather"], spectrum=case["spectrum"], pwv_rms_frac=0.5)
    assert atm.pwv_rms_frac == 0.5 and atm.max_height == 5e3 and atm.interpolation_method == "linear"
    with pytest.raises(RuntimeError, match="must be initialized"):
        atm.simulate_pwv()
    assert "Atmosphere(" in repr(atm)
    # pwv_rms_frac is stored but not forwarded (the reference's behaviour): the layer table uses 3e-2
    obs = M.Observation(case["instrument"], case["plan"], case["site"])
    np.random.seed(0)
    atm.initialize(obs, bind=False)
    L = atm.layers
    assert abs(np.sqrt(np.sum(L["pwv_rms"] ** 2)) - 3e-2 * atm.weather.pwv) < 1e-12


def test_low_elevation_is_rejected():
    case = small_case(n_det=8, duration=5.0, sample_rate=10.0, center=(60.0, 3.0))
    with pytest.raises(ValueError, match="elevation"):
        M.Observation(case["instrument"], case["plan"], case["site"])


def test_no_cpu_fallback():
    """Without a CUDA device the product path raises; it never routes to the oracle or any CPU implementation."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    case = small_case(n_det=8, duration=5.0, sample_rate=10.0)
    with pytest.raises(RuntimeError, match="CUDA"):
        M.Simulation(case["instrument"], case["plan"], case["site"], atmosphere="2d",
                     atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"]))
    for fn in os.listdir(os.path.join(ROOT, "maria_b200")):
        if fn.endswith(".py"):
            src = open(os.path.join(ROOT, "maria_b200", fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f"{fn} imports the oracle"


# ---- the C-ABI ----------------------------------------------------------------------------------------------------
def _header_symbols():
    text = open(os.path.join(ROOT, "include", "maria_atm.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(matm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_build.LIB):
        _build.build()
    lib = ctypes.CDLL(_build.LIB)
    names = _header_symbols()
    assert len(names) >= 18
    for name in names:
        assert hasattr(lib, name), f"{name} is declared in include/maria_atm.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes binding and header disagree"
    bound = _lib.load(build_if_missing=False)
    assert bound.matm_abi_version() == 1
    assert bound.matm_destroy(None) == _lib.ERR_ARG  # null handle: status, not a crash


def test_layer_desc_layout_matches_header():
    """sizeof / field order of the ctypes mirror of matm_layer_desc."""
    names = [f[0] for f in _lib.LayerDesc._fields_]
    assert names == ["data_dev", "n_ext", "n_cs", "h", "pwv_rms", "transform", "ext0", "ext_step", "cs0", "cs_last",
                     "vx_host", "vy_host"]
    assert ctypes.sizeof(_lib.LayerDesc) == 8 + 4 + 4 + 8 + 8 + 72 + 4 * 8 + 2 * 8
