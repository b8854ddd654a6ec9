"""Drop-in module name for the reference's tracer/has_frame.py; see scenegraph.py."""
from .scenegraph import HasFrame  # noqa: F401
