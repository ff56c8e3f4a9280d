"""The adapter between the reference's own objects and this package (maria_b200/adapters.py).

With /root/reference present (the build container) the real ``maria.get_instrument("MUSTANG-2")`` / ``maria.Plan`` /
``maria.get_site`` go through ``from_maria`` and ``install()`` in a subprocess (oracle/check_adapter.py: the import stubs
must not leak into this process).  Everywhere, unit-carrying stand-ins check the conversion rules of ``coerce``."""
import os
import subprocess
import sys

import numpy as np
import pytest

from maria_b200 import adapters
from maria_b200 import inputs as I

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not os.path.isdir("/root/reference/maria"), reason="the reference is only mounted in the build container")
def test_real_maria_objects_through_the_adapter():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "check_adapter.py")], capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "adapter checks passed" in r.stdout and "FAIL" not in r.stdout


class Q:
    """A quantity with units, the way maria/units/quantity.py exposes it: ``.<unit>`` converts, ``base_units_value``."""

    FACTORS = {"m": 1.0, "km": 1e3, "Hz": 1.0, "GHz": 1e9, "rad": 1.0, "deg": np.pi / 180, "s": 1.0}

    def __init__(self, value, unit):
        self.base_units_value = np.asarray(value, dtype=float) * self.FACTORS[unit]

    def __getattr__(self, unit):
        if unit in Q.FACTORS:
            return self.base_units_value / Q.FACTORS[unit]
        raise AttributeError(unit)


class _Obj:
    def __init__(self, **kw):
        self.__dict__.update(kw)


def _maria_like():
    band = _Obj(name="f150", nu=Q([120.0, 150.0, 180.0], "GHz"), tau=np.array([0.0, 1.0, 0.0]), efficiency=0.4, gain_error=0.05,
                center=Q(150.0, "GHz"))
    dets = _Obj(xi=Q([0.0, 0.1], "deg"), eta=Q([0.0, -0.1], "deg"), band_name=np.array(["f150", "f150"]), bands=[band],
                gamma=np.array([np.nan, 0.0]), primary_size=Q([6.0, 6.0], "m"), band_center=np.array([150e9, 150e9]))
    inst = _Obj(dets=dets, name="toy")
    t = np.arange(0.0, 5.0, 0.1)
    plan = _Obj(time=t, phi=Q(np.full(len(t), 30.0), "deg"), theta=Q(np.full(len(t), 45.0), "deg"), frame="az/el", roll=0.0)
    site = _Obj(region="chajnantor", altitude=Q(5.06, "km"))
    return inst, plan, site


def test_quantities_become_si_arrays():
    inst, plan, site = adapters.from_maria(*_maria_like())
    assert isinstance(inst, I.Instrument) and isinstance(plan, I.Plan) and isinstance(site, I.Site)
    assert np.allclose(inst.dets.offsets, np.radians([[0.0, 0.0], [0.1, -0.1]]))
    assert inst.dets.bands[0].center == 150e9 and inst.dets.bands[0].efficiency == 0.4
    assert np.allclose(inst.dets.bands[0].nu, [120e9, 150e9, 180e9]) and inst.dets.gain_error[0] == 0.05
    assert np.allclose(inst.dets.mueller00(), [1.0, 0.5])  # unpolarised / polarised (array.py:204-218)
    assert site.altitude == 5060.0 and np.allclose(plan.phi, np.radians(30.0)) and np.allclose(plan.theta, np.radians(45.0))


def test_coerce_passes_plain_objects_through_and_rejects_sky_frames():
    inst, plan, site = adapters.from_maria(*_maria_like())
    assert adapters.coerce(inst, plan, site) == (inst, plan, site)
    m_inst, m_plan, m_site = _maria_like()
    m_plan.frame = "ra/dec"
    with pytest.raises(ValueError, match="az/el"):
        adapters.from_maria(m_inst, m_plan, m_site)


def test_observation_takes_unit_carrying_objects():
    import maria_b200 as M

    obs = M.Observation(*_maria_like())
    assert obs.instrument.dets.n == 2 and obs.site.altitude == 5060.0 and obs.boresight.el[0] == np.radians(45.0)
