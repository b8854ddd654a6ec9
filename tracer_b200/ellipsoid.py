"""Drop-in module name for the reference's tracer/ellipsoid.py; see geometry.py."""
from .geometry import Ellipsoid, EllipsoidGM  # noqa: F401
