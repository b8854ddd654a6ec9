"""Drop-in module name for the reference's tracer/sphere_surface.py; see geometry.py."""
from .geometry import SphericalGM, HemisphereGM, CutSphereGM, SphericalRectFacet  # noqa: F401
