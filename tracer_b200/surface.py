"""Drop-in module name for the reference's tracer/surface.py; see scenegraph.py."""
from .scenegraph import Surface  # noqa: F401
