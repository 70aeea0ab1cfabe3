"""scarlet_b200: B200-native implementation of scarlet's `Blend.fit` hot path."""
from . import _native  # noqa: F401
