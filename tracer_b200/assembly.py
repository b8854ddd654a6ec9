"""Drop-in module name for the reference's tracer/assembly.py; see scenegraph.py."""
from .scenegraph import Assembly  # noqa: F401
