"""maria_b200 -- B200-native atmospheric time-ordered-data synthesis behind maria's Python interface.

Public surface (same names and argument meaning as thomaswmorris/maria for this path):
``Simulation``, ``Observation``, ``Atmosphere``, ``AutoregressiveProcess``, ``TOD``; ``inputs`` holds network-free
stand-ins for the instrument / plan / site / weather / spectrum objects the path reads.
"""
from . import inputs  # noqa: F401
from .atmosphere import Atmosphere, AutoregressiveProcess, DeviceContext  # noqa: F401
from .sim import Observation, Simulation  # noqa: F401
from .tod import TOD  # noqa: F401

__version__ = "0.1.0"
