"""Drop-in module name for the reference's tracer/quadratic_surface.py; see geometry.py."""
from .geometry import FlatQuadricSurfaceGM, RectFlatQuadricSurfaceGM  # noqa: F401
