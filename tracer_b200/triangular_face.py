"""Drop-in module name for the reference's tracer/triangular_face.py; see geometry.py."""
from .geometry import TriangularFace  # noqa: F401
