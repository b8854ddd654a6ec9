"""Drop-in module name for the reference's tracer/geometry_manager.py; see geometry.py."""
from .geometry import GeometryManager  # noqa: F401
