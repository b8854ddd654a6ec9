"""Drop-in module name for the reference's tracer/cone.py; see geometry.py."""
from .geometry import InfiniteCone, FiniteCone, ConicalFrustum, RectCutConicalFrustum  # noqa: F401
