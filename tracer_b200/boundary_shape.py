"""Drop-in module name for the reference's tracer/boundary_shape.py; see scenegraph.py."""
from .scenegraph import BoundaryShape, BoundaryBox, BoundarySphere, BoundaryCylinder, BoundaryPlane  # noqa: F401
