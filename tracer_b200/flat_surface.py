"""Drop-in module name for the reference's tracer/flat_surface.py; see geometry.py."""
from .geometry import (FlatGeometryManager, FiniteFlatGM, RectPlateGM, ExtrudedRectPlateGM,  # noqa: F401
                       RoundPlateGM, StraightCutRoundPlateGM, PerforatedRectPlateGM)
