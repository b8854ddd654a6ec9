"""One-sided mirror / receiver builders (reference tracer/models/one_sided_mirror.py:33-107).

Each returns an AssembledObject with a single surface whose optics zero the
energy of rays arriving from behind (OneSided*), so no back surface is needed.
"""
from ..scenegraph import AssembledObject, Surface
from ..geometry import RectPlateGM, RectangularParabolicDishGM, RectFlatQuadricSurfaceGM
from .. import optics_callables as opt


def _mirror_optics(absorptivity, sigma, bi_var, option):
    if option == 'fast':        # no per-hit storage on the mirrors
        return opt.OneSidedRealReflective(absorptivity, sigma, bi_var)
    if option == 'receiver':
        return opt.OneSidedRealReflectiveReceiver(absorptivity, sigma, bi_var)
    return opt.OneSidedRealReflectiveDetector(absorptivity, sigma, bi_var)


def rect_one_sided_mirror(width, height, absorptivity=0, sigma=0., bi_var=True, option=None, location=None,
                          rotation=None, bounds=None):
    """Flat rectangular mirror on the local XY plane, reflective from +z."""
    if option == 'receiver':
        option = None           # the reference only distinguishes 'fast' here (:47-50)
    surf = Surface(RectPlateGM(width, height), _mirror_optics(absorptivity, sigma, bi_var, option))
    return AssembledObject(surfs=[surf], location=location, rotation=rotation, bounds=bounds)


def rect_para_one_sided_mirror(width, height, focal_length, absorptivity=0., sigma=0., bi_var=True, option=None,
                               location=None, rotation=None, bounds=None):
    """Rectangular cut of a paraboloid of the given focal length."""
    if option == 'receiver':
        option = None
    surf = Surface(RectangularParabolicDishGM(width, height, focal_length),
                   _mirror_optics(absorptivity, sigma, bi_var, option))
    return AssembledObject(surfs=[surf], location=location, rotation=rotation, bounds=bounds)


def flat_quad_one_sided_mirror(width, height, quad_params, absorptivity=0., sigma=0., bi_var=True, option=None,
                               location=None, rotation=None, bounds=None):
    """Rectangular facet z = a x^2 + b y^2 + c x y + d x + e y + f."""
    a, b, c, d, e, f = quad_params
    surf = Surface(RectFlatQuadricSurfaceGM(width, height, a, b, c, d, e, f),
                   _mirror_optics(absorptivity, sigma, bi_var, option))
    return AssembledObject(surfs=[surf], location=location, rotation=rotation, bounds=bounds)


def one_sided_receiver(width, height, absorptivity=1, location=None, rotation=None):
    """Rectangular receiver whose front (+z) records hits; ``get_all_hits()`` of
    its surface returns [absorbed energy, hit locations]."""
    front = Surface(RectPlateGM(width, height), opt.OneSidedReflectiveReceiver(absorptivity))
    return AssembledObject(surfs=[front], location=location, rotation=rotation)
