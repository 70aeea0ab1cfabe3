"""`import scarlet` alias: the reference package name, served by scarlet_b200."""
import sys
import scarlet_b200
from scarlet_b200 import *  # noqa: F401,F403
from scarlet_b200 import (fft, psf, operator, update, measurement, observation, component, source, blend, _native,  # noqa: F401
                          operators_pybind11)

for _n in ("fft", "psf", "operator", "update", "measurement", "observation", "component", "source", "blend",
           "operators_pybind11"):
    sys.modules["scarlet." + _n] = getattr(scarlet_b200, _n)
