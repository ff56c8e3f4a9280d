"""Import stubs that let the UNMODIFIED reference source under /root/reference run in the build container.

TEST INFRASTRUCTURE ONLY (used by oracle/pin_against_reference.py, never by maria_b200/ or at GPU run time).
The container lacks jax, dask, h5py, astropy, arrow and matplotlib.  The stubs are:
  * jax.numpy -> numpy (float64, i.e. the reference as it would run with jax_enable_x64), jax.jit -> identity;
    jax.scipy.interpolate.RegularGridInterpolator is installed by the pin script from the oracle's restatement;
  * dask.array -> eager numpy arrays with a .compute() method;
  * arrow -> a tiny UTC timestamp class;  maria._version -> a constant;
  * h5py / astropy / matplotlib / healpy / reproject -> MagicMock modules (only touched at import time on az/el plans;
    h5py.File is replaced by an in-memory table provider in the pin script).
No reference source is copied: the files are imported from where they lie.
"""
import sys, types, importlib.abc, importlib.machinery
from unittest.mock import MagicMock
import numpy as np

MOCK_ROOTS = ("h5py", "astropy", "matplotlib", "healpy", "reproject", "cmocean", "PIL", "skimage", "torch_unused")

class _MockLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = MagicMock(name=spec.name); m.__path__ = []; m.__name__ = spec.name; m.__spec__ = spec; m.__loader__ = self
        return m
    def exec_module(self, module): pass

class MockFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path, target=None):
        if name.split(".")[0] in MOCK_ROOTS:
            return importlib.machinery.ModuleSpec(name, _MockLoader(), is_package=True)
        return None

class _Arrow:
    def __init__(self, ts): self.ts = float(ts)
    def timestamp(self): return self.ts
    def to(self, tz): return self
    def _dt(self):
        import datetime
        return datetime.datetime.fromtimestamp(self.ts, datetime.timezone.utc)
    hour = property(lambda self: self._dt().hour)
    minute = property(lambda self: self._dt().minute)
    second = property(lambda self: self._dt().second)
    year = property(lambda self: self._dt().year)
    def format(self, fmt="", *a, **k):
        d = self._dt()
        if fmt == "SSS": return f"{d.microsecond // 1000:03d}"
        if fmt == "DDD": return str(d.timetuple().tm_yday)
        return d.isoformat()
    def shift(self, seconds=0, **k): return _Arrow(self.ts + seconds)
    def __repr__(self): return f"<Arrow {self.ts}>"

def _arrow_get(x=None, *a, **k):
    if isinstance(x, _Arrow): return x
    if x is None: return _Arrow(1.7e9)
    return _Arrow(float(x))

def install():
    arrow = types.ModuleType("arrow"); arrow.get = _arrow_get; arrow.now = lambda *a, **k: _Arrow(1.7e9); arrow.Arrow = _Arrow
    sys.modules["arrow"] = arrow
    sys.meta_path.insert(0, MockFinder())
    jax = types.ModuleType("jax")
    jax.numpy = np; jax.jit = lambda f=None, **k: (f if f is not None else (lambda g: g))
    jax.scipy = types.ModuleType("jax.scipy"); jax.scipy.interpolate = types.ModuleType("jax.scipy.interpolate")
    jax.scipy.ndimage = types.ModuleType("jax.scipy.ndimage")
    sys.modules.update({"jax": jax, "jax.numpy": np, "jax.scipy": jax.scipy, "jax.scipy.interpolate": jax.scipy.interpolate, "jax.scipy.ndimage": jax.scipy.ndimage})
    dask = types.ModuleType("dask"); da = types.ModuleType("dask.array")
    class DArr(np.ndarray):
        def compute(self): return np.asarray(self)
    da.asarray = lambda x, dtype=None, **k: np.asarray(x, dtype=dtype).view(DArr); da.from_array = lambda x, **k: np.asarray(x).view(DArr)
    da.Array = DArr; da.ones = lambda *a, **k: np.ones(*a, **{kk: v for kk, v in k.items() if kk != "chunks"}).view(DArr); da.zeros = np.zeros
    dask.array = da; sys.modules.update({"dask": dask, "dask.array": da})
    dd = types.ModuleType("dask.diagnostics"); sys.modules["dask.diagnostics"] = dd
    v = types.ModuleType("maria._version"); v.__version__ = "0.0.0-pin"; v.version = "0.0.0-pin"; v.__version_tuple__ = (0,0,0); v.version_tuple=(0,0,0)
    sys.modules["maria._version"] = v
    sys.path.insert(0, "/root/reference")
