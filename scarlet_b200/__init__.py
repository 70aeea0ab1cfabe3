"""scarlet_b200: B200-native implementation of scarlet's `Blend.fit` hot path.

Drop-in surface of the reference package (scarlet/__init__.py:1-9): Frame, Observation,
Blend, the sources, and the `operator` / `update` / `measurement` / `psf` / `fft` modules.
All array work dispatches to hand-written sm_100a CUDA kernels through the C ABI of
libscarlet_b200.so; there is no CPU fallback.
"""
from . import _native  # noqa: F401
from . import fft, psf, operator, update, measurement, bbox, operators_pybind11  # noqa: F401
from .observation import Frame, Observation, LowResObservation  # noqa: F401
from .component import Component, ComponentTree, BlendFlag, Prior, AffinePrior  # noqa: F401
from . import schedule  # noqa: F401
from .source import (PointSource, ExtendedSource, MultiComponentSource, CombinedExtendedSource,  # noqa: F401
                     RandomSource, SourceInitError, extended_sources)
from .blend import Blend, fit_many  # noqa: F401

__version__ = "0.1.0"
