"""Drop-in module name for the reference's tracer/paraboloid.py; see geometry.py."""
from .geometry import (Paraboloid, ParabolicDishGM, HexagonalParabolicDishGM,  # noqa: F401
                       RectangularParabolicDishGM, ParabolicCylinder, ParabolicTroughGM)
