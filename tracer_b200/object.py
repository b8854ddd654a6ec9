"""Drop-in module name for the reference's tracer/object.py; see scenegraph.py."""
from .scenegraph import AssembledObject  # noqa: F401
