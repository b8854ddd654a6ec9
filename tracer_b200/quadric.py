"""Drop-in module name for the reference's tracer/quadric.py; see geometry.py."""
from .geometry import QuadricGM  # noqa: F401
