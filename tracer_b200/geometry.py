"""Geometry-manager descriptions with the reference's class names, constructor
signatures, validation errors and attribute names.

In the reference every GeometryManager also *computes* (find_intersections,
select_rays, get_normals ...; geometry_manager.py:8-63).  Here the computation
lives in csrc/geom.cuh; these classes only validate and hold parameters, which
the scene compiler (scene.py) turns into TrcSurfaceRec.gp.  Attribute names are
the reference's so that either family of objects compiles.

Reference classes: flat_surface.py (FlatGeometryManager :11, FiniteFlatGM :115,
RectPlateGM :181, ExtrudedRectPlateGM :253, RoundPlateGM :457,
StraightCutRoundPlateGM :548), triangular_face.py:12, paraboloid.py (:11, :71,
:174, :225, :328, :386), sphere_surface.py (:9, :117, :168, :206), cone.py (:7,
:74, :261, :356), cylinder.py (:12, :59, :161), quadratic_surface.py (:4, :63),
flat_surface.py PerforatedRectPlateGM :357, ellipsoid.py (:5, :63), polygon.py (:7, :139).
"""
import numpy as np


class GeometryManager(object):
    """Base of all geometry descriptions (geometry_manager.py:7)."""

    def find_intersections(self, frame, ray_bundle):
        raise TypeError(
            "tracer_b200 geometry managers are descriptions; intersections run on the GPU "
            "through TracerEngine.ray_tracer (no CPU path)")


# ------------------------------------------------------------------ flat
class FlatGeometryManager(GeometryManager):
    """Infinite plane, the local XY plane."""


class FiniteFlatGM(FlatGeometryManager):
    pass


def _cartesian_fluxmap(half_dims, eners, local_coords, resolution):
    xs = np.linspace(-half_dims[0], half_dims[0], resolution + 1)
    ys = np.linspace(-half_dims[1], half_dims[1], resolution + 1)
    binned = np.histogram2d(local_coords[0], local_coords[1], bins=[xs, ys], weights=eners)[0]
    areas = np.outer(np.abs(np.diff(xs)), np.abs(np.diff(ys)))
    return np.hstack(binned / areas)


class RectPlateGM(FiniteFlatGM):
    def __init__(self, width, height):
        if width <= 0:
            raise ValueError("Width must be positive")
        if height <= 0:
            raise ValueError("Height must be positive")
        self.width = width
        self.height = height
        self._half_dims = np.c_[[width, height]] / 2.

    def get_fluxmap(self, eners, local_coords, resolution):
        """Flux (energy / bin area) on a resolution x resolution cartesian grid,
        flattened row-major in x (flat_surface.py:237-251)."""
        return _cartesian_fluxmap(self._half_dims[:, 0], eners, local_coords, resolution)


class ExtrudedRectPlateGM(RectPlateGM):
    def __init__(self, width, height, extr_center, extr_width, extr_height):
        RectPlateGM.__init__(self, width, height)
        self.extr_center = extr_center
        self.extr_half_dims = np.c_[[extr_width, extr_height]] / 2.
        assert ((extr_center + self.extr_half_dims) < self._half_dims).all()


class PerforatedRectPlateGM(RectPlateGM):
    """Rectangle with circular holes (flat_surface.py:357-377).  ``extr_centers`` is an (n, 2)
    array, one hole centre per row (that is how find_intersections :372 reads it, whatever the
    reference docstring says), ``extr_radii`` an (n,) array."""

    def __init__(self, width, height, extr_centers, extr_radii):
        RectPlateGM.__init__(self, width, height)
        self.extr_centers = extr_centers
        self.extr_radii = extr_radii


class FlatSimplePolygonGM(FiniteFlatGM):
    """Simple polygon in the local XY plane, crossing-number test (polygon.py:7-63).
    ``profile`` = [[xs], [ys]], vertices in sequence."""

    def __init__(self, profile):
        self.profile = profile


class PerforatedPolygonGM(FlatSimplePolygonGM):
    """Polygon with circular holes (polygon.py:139-162); centres (n, 2), radii (n,)."""

    def __init__(self, profile, extr_centers, extr_radii):
        self.extr_centers = extr_centers
        self.extr_radii = extr_radii
        FlatSimplePolygonGM.__init__(self, profile)


class RoundPlateGM(FiniteFlatGM):
    def __init__(self, Re, Ri=None):
        if Re <= 0.:
            raise ValueError("Radius must be positive")
        if Ri is not None:
            if Ri >= Re:
                raise ValueError("Inner Radius must be lower than the outer one")
            if Ri <= 0.:
                raise ValueError("Radius must be positive")
        self._Ri = Ri
        self._Re = Re


class StraightCutRoundPlateGM(RoundPlateGM):
    def __init__(self, Re, x_cut=None):
        RoundPlateGM.__init__(self, Re)
        self.x_cut = x_cut


class TriangularFace(FiniteFlatGM):
    """Triangle = local origin + the two columns of ``verts`` (3x2)."""

    def __init__(self, verts):
        self.set_vertices(verts)

    def set_vertices(self, verts):
        self._verts = verts


# --------------------------------------------------------------- quadrics
class QuadricGM(GeometryManager):
    pass


class Paraboloid(QuadricGM):
    """z = a x^2 + b y^2 with a = 1/a_arg^2, b = 1/b_arg^2 (legacy arguments)."""

    def __init__(self, a=1., b=None):
        if b is None:
            b = a
        self.a = 1. / (a ** 2)
        self.b = 1. / (b ** 2)


class ParabolicDishGM(Paraboloid):
    def __init__(self, diameter, focal_length):
        par_param = 2. * np.sqrt(focal_length)
        Paraboloid.__init__(self, par_param, par_param)
        self._R = float(diameter / 2.)
        self._h = float((diameter / 2. / par_param) ** 2)


class HexagonalParabolicDishGM(Paraboloid):
    def __init__(self, diameter, focal_length):
        par_param = 2 * np.sqrt(focal_length)
        Paraboloid.__init__(self, par_param, par_param)
        self._R = diameter / 2.


class RectangularParabolicDishGM(Paraboloid):
    def __init__(self, width, height, focal_length, off_axis_normal=None):
        self._off_axis_normal = off_axis_normal
        if off_axis_normal is not None:
            raise NotImplementedError(
                "RectangularParabolicDishGM(off_axis_normal=...) is not supported by the GPU engine")
        self._w, self._h = width / 2., height / 2.
        self._half_dims = np.c_[[width, height]] / 2
        par_param = 2. * np.sqrt(focal_length)
        Paraboloid.__init__(self, par_param, par_param)


class ParabolicCylinder(QuadricGM):
    def __init__(self, a=1.):
        self.a = 1. / (a ** 2)


class ParabolicTroughGM(ParabolicCylinder):
    def __init__(self, aperture, focal_length, length):
        par_param = 2. * np.sqrt(focal_length)
        ParabolicCylinder.__init__(self, par_param)
        self._l = length
        self._w = float(aperture)
        self._h = float((aperture / 2. / par_param) ** 2)


class Ellipsoid(QuadricGM):
    """(x/a)^2 + (y/b)^2 + (z/c)^2 = 1; holds the inverse squares like ellipsoid.py:16-18."""

    def __init__(self, a, b, c):
        self.a = 1. / (a ** 2)
        self.b = 1. / (b ** 2)
        self.c = 1. / (c ** 2)


class EllipsoidGM(Ellipsoid):
    """Ellipsoid truncated along the local axes (ellipsoid.py:63-118).  As in the reference
    (:70-75) the limits only act when at least one of them is None; with all three given they
    are replaced by the full extent."""

    def __init__(self, a, b, c, xlim=None, ylim=None, zlim=None):
        Ellipsoid.__init__(self, a, b, c)
        if any(lim is None for lim in (xlim, ylim, zlim)):
            self.xlim, self.ylim, self.zlim = xlim, ylim, zlim
            self.truncated = True
        else:
            self.xlim, self.ylim, self.zlim = [-a, a], [-b, b], [-c, c]
            self.truncated = False
        self._semi_axes = (float(a), float(b), float(c))


class SphericalGM(QuadricGM):
    def __init__(self, radius=1.):
        self.set_radius(radius)

    def get_radius(self):
        return self._rad

    def set_radius(self, rad):
        if rad <= 0:
            raise ValueError("Radius must be positive")
        self._rad = rad


class HemisphereGM(SphericalGM):
    pass


class CutSphereGM(SphericalGM):
    def __init__(self, radius=1., bounding_volume=None):
        SphericalGM.__init__(self, radius)
        self._bound = bounding_volume


class SphericalRectFacet(SphericalGM):
    def __init__(self, radius, lx, ly):
        SphericalGM.__init__(self, radius)
        self.lx = lx
        self.ly = ly


class InfiniteCone(QuadricGM):
    """x^2 + y^2 = (c (z - a))^2"""

    def __init__(self, c, a=0):
        self.c = float(c)
        self.a = float(a)


class FiniteCone(InfiniteCone):
    def __init__(self, r, h):
        if h <= 0. or r <= 0.:
            raise AttributeError
        self.h = float(h)
        self.r = float(r)
        InfiniteCone.__init__(self, c=self.r / self.h)


class ConicalFrustum(InfiniteCone):
    def __init__(self, z1, r1, z2, r2):
        r1, r2, z1, z2 = float(r1), float(r2), float(z1), float(z2)
        if r1 <= 0. or r2 <= 0.:
            raise AttributeError
        if r1 == r2 or z1 == z2:
            raise AttributeError
        InfiniteCone.__init__(self, c=float((r2 - r1) / (z2 - z1)), a=float((r2 * z1 - r1 * z2) / (r2 - r1)))
        self.r1, self.r2, self.z1, self.z2 = r1, r2, z1, z2
        self.zmin, self.zmax = np.sort([z1, z2])


class RectCutConicalFrustum(ConicalFrustum):
    def __init__(self, z1, r1, z2, r2, w, h):
        ConicalFrustum.__init__(self, z1, r1, z2, r2)
        self.half_dims = np.array([w / 2., h / 2.])
        if np.sqrt(np.sum(self.half_dims ** 2)) <= min(r1, r2):
            raise ValueError('Bad rectangular cut frustum shape, width and height too small')


class InfiniteCylinder(QuadricGM):
    def __init__(self, diameter):
        self._R = diameter / 2.


class FiniteCylinder(InfiniteCylinder):
    def __init__(self, diameter, height, ang_range=[0., 2. * np.pi]):
        self._half_h = height / 2.
        assert len(ang_range) == 2
        self._ang_range = ang_range
        InfiniteCylinder.__init__(self, diameter)


class RectCutCylinder(FiniteCylinder):
    def __init__(self, diameter, height, w, h):
        FiniteCylinder.__init__(self, diameter, height)
        self.half_dims = np.array([w / 2., h / 2.])
        if np.sqrt(np.sum(self.half_dims ** 2)) <= self._R:
            raise ValueError('Bad rectangular cut cylindershape, width and height too small')


class FlatQuadricSurfaceGM(QuadricGM):
    """z = a x^2 + b y^2 + c x y + d x + e y + f"""

    def __init__(self, a=1., b=1., c=1., d=0., e=0., f=0.):
        self.a, self.b, self.c, self.d, self.e, self.f = a, b, c, d, e, f


class RectFlatQuadricSurfaceGM(FlatQuadricSurfaceGM):
    def __init__(self, width, height, a=1., b=1., c=1., d=1., e=1., f=1.):
        FlatQuadricSurfaceGM.__init__(self, a, b, c, d, e, f)
        self._half_dims = np.c_[[width, height]] / 2.
        self._w, self._h = width / 2., height / 2.
