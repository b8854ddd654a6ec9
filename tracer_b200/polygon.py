"""Drop-in module name for the reference's tracer/polygon.py; see geometry.py."""
from .geometry import FlatSimplePolygonGM, PerforatedPolygonGM  # noqa: F401
