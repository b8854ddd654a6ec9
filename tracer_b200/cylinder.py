"""Drop-in module name for the reference's tracer/cylinder.py; see geometry.py."""
from .geometry import InfiniteCylinder, FiniteCylinder, RectCutCylinder  # noqa: F401
