"""Scene builders of the BASELINE configurations and of the test scenes, shared by
``bench.py``, ``__graft_entry__.smoke()``, the golden-vector generator (real reference classes,
``NS('tracer')``) and the tests (``NS('tracer_b200')``).  Every builder returns
(assembly, source_factory) where source_factory() draws the source bundle from
numpy.random in the reference's order.  Scenes follow SURVEY.md section 8(d).
"""
import importlib
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HSTAT_CSV = os.path.join(HERE, 'data', 'sandia_hstat_coordinates.csv')   # examples/sandia_hstat_coordinates.csv


class NS(object):
    """Lazy namespace: NS('tracer').flat_surface.RectPlateGM etc."""

    def __init__(self, pkg):
        self.pkg = pkg

    def __getattr__(self, mod):
        m = importlib.import_module(self.pkg + '.' + mod)
        if hasattr(m, '__path__'):          # sub-package (tracer.models)
            m = NS(self.pkg + '.' + mod)
        setattr(self, mod, m)
        return m


def c1_flat_mirror(T, n_rays=2000):
    """Config 1: flat mirror + flat receiver, pillbox disk source."""
    mirror = T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(2., 2.),
                                                               T.optics_callables.Reflective(0.05))])
    recv = T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(4., 4.), T.optics_callables.ReflectiveReceiver(1.))],
        transform=np.dot(T.spatial_geometry.translate(0, 0, 5.), T.spatial_geometry.rotx(np.pi)))
    asm = T.assembly.Assembly(objects=[mirror, recv])

    def source():
        return T.sources.disk_bundle(n_rays, np.c_[[0., 0., 4.]], np.r_[0., 0., -1.], 0.8, 4.65e-3, flux=1000.)
    return asm, source


def c2_dish(T, n_rays=2000, bi_var=True, sigma=2e-3):
    """Config 2: parabolic dish with slope error + flat receiver, Buie sunshape."""
    dish = T.object.AssembledObject(surfs=[T.surface.Surface(
        T.paraboloid.ParabolicDishGM(5., 3.), T.optics_callables.RealReflective(0.06, sigma, bi_var=bi_var))])
    recv = T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(0.3, 0.3), T.optics_callables.ReflectiveReceiver(1.))],
        transform=np.dot(T.spatial_geometry.translate(0, 0, 3.), T.spatial_geometry.rotx(np.pi)))
    asm = T.assembly.Assembly(objects=[dish, recv])

    def source():
        return T.sources.buie_sunshape(n_rays, np.c_[[0., 0., 10.]], np.r_[0., 0., -1.], 2.5, 0.05, flux=1000.)
    return asm, source


def nsttf_positions(n_hstats=None):
    pos = np.loadtxt(HSTAT_CSV, delimiter=',')
    pos[:, 1] -= 4.       # examples/Sandia_NSTTF_field example.py:98
    return pos if n_hstats is None else pos[:n_hstats].copy()


def sun_vector(azimuth, zenith):
    az = np.pi / 2. - azimuth
    return np.r_[np.sin(zenith) * np.cos(az), np.sin(zenith) * np.sin(az), np.cos(zenith)]


def c3_nsttf(T, n_rays=5000, n_hstats=None, dense=False, sigma=1e-3):
    """Config 3: Sandia NSTTF heliostat field + tower receiver, Buie CSR 0.01."""
    pos = nsttf_positions(n_hstats)
    hf = T.models.heliostat_field
    field = hf.HeliostatField(pos, 6.1, 6.1, absorptivity=0.04, sigma=sigma, bi_var=True, MCRT_option='fast')
    zenith = np.radians(35.05)
    field.track_sun(0., zenith, aim_points=np.tile(np.r_[0., 0., 60.], (len(pos), 1)))
    recv = T.models.one_sided_mirror.one_sided_receiver(11., 11., 1.)
    recv.set_transform(np.dot(T.spatial_geometry.translate(0., 0., 60.), T.spatial_geometry.rotx(-np.pi / 2.)))
    asm = T.assembly.Assembly(objects=[recv], subassemblies=[field])
    sun = sun_vector(0., zenith)
    lo, hi = pos.min(axis=0), pos.max(axis=0)
    centre = 0.5 * (lo + hi)
    radius = 1.1 * 0.5 * np.sqrt(np.sum((hi[:2] - lo[:2]) ** 2))
    if dense:
        radius = 0.25 * radius

    def source():
        return T.sources.buie_sunshape(n_rays, np.vstack(300. * sun + centre), -sun, radius, 0.01, flux=1000.,
                                       pre_process_CSR=False)
    return asm, source


def cylinder_mesh(n_theta, n_z, radius=1., length=2.):
    """Closed triangulated cavity: side wall + back cap at z = length; open at z=0.
    Returns a list of (p0, p1, p2) global triangles with inward-facing CCW order."""
    tris = []
    th = np.linspace(0., 2. * np.pi, n_theta + 1)
    zs = np.linspace(0., length, n_z + 1)
    for i in range(n_theta):
        for j in range(n_z):
            a = np.r_[radius * np.cos(th[i]), radius * np.sin(th[i]), zs[j]]
            b = np.r_[radius * np.cos(th[i + 1]), radius * np.sin(th[i + 1]), zs[j]]
            c = np.r_[radius * np.cos(th[i + 1]), radius * np.sin(th[i + 1]), zs[j + 1]]
            d = np.r_[radius * np.cos(th[i]), radius * np.sin(th[i]), zs[j + 1]]
            tris += [(a, c, b), (a, d, c)]
    centre = np.r_[0., 0., length]
    for i in range(n_theta):
        b = np.r_[radius * np.cos(th[i]), radius * np.sin(th[i]), length]
        c = np.r_[radius * np.cos(th[i + 1]), radius * np.sin(th[i + 1]), length]
        tris.append((centre, c, b))
    return tris


def triangle_object(T, p0, p1, p2, optics):
    """One AssembledObject per triangle, as ray_trace_utils/stl_utils.py:213-227:
    frame with origin p0, x along p1-p0, z along the normal; BoundaryBox = AABB."""
    e0, e1 = p1 - p0, p2 - p0
    z = np.cross(e0, e1)
    z /= np.sqrt(np.sum(z ** 2))
    x = e0 / np.sqrt(np.sum(e0 ** 2))
    y = np.cross(z, x)
    rot = np.vstack((x, y, z)).T
    frame = np.eye(4)
    frame[:3, :3] = rot
    frame[:3, 3] = p0
    verts = np.dot(rot.T, np.vstack((e0, e1)).T)
    verts[2] = 0.
    surf = T.surface.Surface(T.triangular_face.TriangularFace(verts), optics)
    pts = np.vstack((p0, p1, p2))
    local = np.dot(rot.T, (pts - p0).T)
    box = T.boundary_shape.BoundaryBox(np.array([local.min(axis=1) - 1e-9, local.max(axis=1) + 1e-9]))
    return T.object.AssembledObject(surfs=[surf], bounds=[box], transform=frame)


def c4_cavity(T, n_rays=2000, n_theta=8, n_z=4, absorptivity=0.7):
    """Config 4 (reduced): triangulated cylindrical cavity, Lambertian multi-bounce."""
    objs = [triangle_object(T, p0, p1, p2, T.optics_callables.LambertianReceiver(absorptivity))
            for (p0, p1, p2) in cylinder_mesh(n_theta, n_z)]
    asm = T.assembly.Assembly(objects=objs)

    def source():
        return T.sources.disk_bundle(n_rays, np.c_[[0., 0., -0.01]], np.r_[0., 0., 1.], 0.9, np.pi / 4., flux=1000.)
    return asm, source


def cylinder_mesh_array(n_theta, n_z, radius=1., length=2.):
    """cylinder_mesh as one (n,3,3) array [face, vertex, xyz], built without Python loops
    (1e6 faces in a fraction of a second); same faces, side wall first (theta-major) then cap."""
    th = np.linspace(0., 2. * np.pi, n_theta + 1)
    zs = np.linspace(0., length, n_z + 1)
    x, y = radius * np.cos(th), radius * np.sin(th)
    i, j = np.meshgrid(np.arange(n_theta), np.arange(n_z), indexing='ij')
    i, j = i.ravel(), j.ravel()
    a = np.stack([x[i], y[i], zs[j]], axis=1)
    b = np.stack([x[i + 1], y[i + 1], zs[j]], axis=1)
    c = np.stack([x[i + 1], y[i + 1], zs[j + 1]], axis=1)
    d = np.stack([x[i], y[i], zs[j + 1]], axis=1)
    side = np.stack([np.stack([a, c, b], axis=1), np.stack([a, d, c], axis=1)], axis=1).reshape(-1, 3, 3)
    k = np.arange(n_theta)
    centre = np.tile(np.r_[0., 0., length], (n_theta, 1))
    cap = np.stack([centre, np.stack([x[k + 1], y[k + 1], np.full(n_theta, length)], axis=1),
                    np.stack([x[k], y[k], np.full(n_theta, length)], axis=1)], axis=1)
    return np.concatenate([side, cap])


def c4_cavity_mesh(T, n_rays=2000, n_theta=8, n_z=4, absorptivity=0.7, per_face_objects=False):
    """Config 4 through the STL route (ray_trace_utils/stl_utils.py:213): the cavity of
    c4_cavity as ONE mesh object, or (per_face_objects) as one Surface per face built from
    the same per-face frames -- the two must trace identically."""
    tris = cylinder_mesh_array(n_theta, n_z)
    opt = T.optics_callables.LambertianReceiver
    if per_face_objects:
        geoms, locs, rots = T.stl_utils.stl_to_tracer_geom(tris, option='triangle')
        surfs = [T.surface.Surface(g, opt(absorptivity), location=l, rotation=r) for g, l, r in zip(geoms, locs, rots)]
        obj = T.object.AssembledObject(surfs=surfs)
    else:
        obj = T.stl_utils.make_stl_tracer_object(tris, opt, dict(absorptivity=absorptivity), option='triangle')
    asm = T.assembly.Assembly(objects=[obj])

    def source():
        return T.sources.disk_bundle(n_rays, np.c_[[0., 0., -0.01]], np.r_[0., 0., 1.], 0.9, np.pi / 4., flux=1000.)
    return asm, source


def c5_cavity(T, n_rays=2000, spectral=False):
    """Config 5 (one pass): Holman ex. 8.17-like cylinder cavity made of quadric
    surfaces: aperture disk, three cylinder sections, flat back
    (emissive_losses/emissive_losses_test.py:38).  Rays are emitted from the
    aperture into the cavity.  ``spectral`` puts directional-spectral optics on the walls."""
    oc = T.optics_callables
    R, depth = 0.01, 0.01
    if spectral:
        thetas = np.linspace(0., np.pi / 2., 10)
        wls = np.linspace(0.3e-6, 2.5e-6, 12)
        table = 0.5 + 0.4 * np.outer(np.cos(thetas), np.linspace(1., 0.5, 12))

        def wall():
            return oc.Lambertian_directional_axisymmetric_piecewise_spectralReceiver(thetas, table, wls)
    else:
        def wall():
            return oc.LambertianReceiver(1.)
    objs = [T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RoundPlateGM(R), oc.LambertianReceiver(1.))],
                                     transform=T.spatial_geometry.rotx(np.pi))]
    for k in range(3):
        objs.append(T.object.AssembledObject(
            surfs=[T.surface.Surface(T.cylinder.FiniteCylinder(2. * R, depth), wall())],
            transform=T.spatial_geometry.translate(0., 0., (k + 0.5) * depth)))
    objs.append(T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RoundPlateGM(R), wall())],
                                         transform=T.spatial_geometry.translate(0., 0., 3. * depth)))
    asm = T.assembly.Assembly(objects=objs)

    def source():
        b = T.sources.disk_bundle(n_rays, np.c_[[0., 0., 1e-6]], np.r_[0., 0., 1.], R, np.pi / 2., flux=1.)
        if spectral:
            wl = np.random.uniform(0.3e-6, 2.5e-6, size=n_rays)
            b = T.ray_bundle.RayBundle(vertices=b.get_vertices(), directions=b.get_directions(),
                                       energy=b.get_energy(), wavelengths=wl)
        return b
    return asm, source


def c6_optics_zoo(T, n_rays=2000):
    """Five plates under a diffuse lid, one optics class of the IAM / piecewise /
    attenuating family per plate, each with accountants."""
    oc = T.optics_callables
    thetas = np.linspace(0., np.pi / 2., 7)
    abs_th = np.linspace(0.2, 0.7, 7)
    spec_th = np.linspace(0.9, 0.1, 7)
    optics = [
        oc.Lambertian_IAMReceiver(0.3, 0.16, 1.3),
        oc.RealReflective_IAMReceiver(0.1, 0.2, 2e-3, True),
        oc.LambertianSpecular_IAMReceiver(0.2, 0.5, 0.16),
        oc.Lambertian_piecewise_Specular_directional_axisymmetric_piecewiseReceiver(thetas, abs_th, spec_th),
        oc.LambertianAbsorbantAbsorberAttenuatorLocation(0.25, [0.4], np.pi / 2., 1.5),
    ]
    objs = []
    for k, opt in enumerate(optics):
        objs.append(T.object.AssembledObject(
            surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(1., 3.), opt)],
            transform=T.spatial_geometry.translate(k - 2., 0., 0.)))
    objs.append(T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(7., 5.), oc.LambertianReceiver(0.3))],
        transform=np.dot(T.spatial_geometry.translate(0., 0., 1.5), T.spatial_geometry.rotx(np.pi))))
    asm = T.assembly.Assembly(objects=objs)

    def source():
        direction = np.r_[0.3, 0.1, -1.]
        direction = direction / np.sqrt(np.sum(direction ** 2))
        return T.sources.disk_bundle(n_rays, np.c_[[-0.4, -0.1, 1.4]], direction, 2.4, 0.02, flux=1000.)
    return asm, source


def c7_periodic_cell(T, n_rays=2000):
    """A unit cell: Lambertian floor, PeriodicBoundary side walls (period 1 along x and y),
    oblique pillbox source above.  Rays leaving through a wall re-enter from the opposite one."""
    oc = T.optics_callables
    sg = T.spatial_geometry
    objs = [T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(1., 1.), oc.LambertianReceiver(0.6))])]
    walls = [
        np.dot(sg.translate(0.5, 0., 0.5), sg.roty(np.pi / 2.)),
        np.dot(sg.translate(-0.5, 0., 0.5), sg.roty(-np.pi / 2.)),
        np.dot(sg.translate(0., 0.5, 0.5), sg.rotx(-np.pi / 2.)),
        np.dot(sg.translate(0., -0.5, 0.5), sg.rotx(np.pi / 2.)),
    ]
    for frame in walls:
        objs.append(T.object.AssembledObject(
            surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(1., 1.), oc.PeriodicBoundary(1.))], transform=frame))
    asm = T.assembly.Assembly(objects=objs)

    def source():
        direction = np.r_[0.7, 0.35, -1.]
        direction = direction / np.sqrt(np.sum(direction ** 2))
        return T.sources.disk_bundle(n_rays, np.c_[[0., 0., 0.95]], direction, 0.45, 0.3, flux=1000.)
    return asm, source


def c9_refractive(T, n_rays=2000):
    """Refraction: a glass slab that splits every ray into a reflected and a
    refracted child (single_ray=False), a glass ball that picks one of them per ray with a
    rough surface (single_ray=True, sigma), an absorbing slab (RefractiveTransmissiveHomogenous,
    one attenuation coefficient per medium) and a diffuse receiver underneath.  The source
    bundle carries ref_index = 1 (air)."""
    oc = T.optics_callables
    sg = T.spatial_geometry
    slab = T.object.AssembledObject(surfs=[
        T.surface.Surface(T.flat_surface.RectPlateGM(1.2, 2.), oc.RefractiveHomogenous(1., 1.5, single_ray=False)),
        T.surface.Surface(T.flat_surface.RectPlateGM(1.2, 2.), oc.RefractiveHomogenous(1., 1.5, single_ray=False),
                          location=np.r_[0., 0., -0.3])],
        transform=sg.translate(-0.7, 0., 0.))
    ball = T.object.AssembledObject(surfs=[
        T.surface.Surface(T.sphere_surface.SphericalGM(0.35), oc.RefractiveHomogenous(1., 1.33, single_ray=True, sigma=2e-3))],
        transform=sg.translate(0.45, 0.5, -0.1))
    tinted = T.object.AssembledObject(surfs=[
        T.surface.Surface(T.flat_surface.RectPlateGM(0.9, 0.9),
                          oc.RefractiveTransmissiveHomogenous(1., 1.45, [0.05, 2.5], single_ray=True)),
        T.surface.Surface(T.flat_surface.RectPlateGM(0.9, 0.9),
                          oc.RefractiveTransmissiveHomogenous(1., 1.45, [0.05, 2.5], single_ray=True),
                          location=np.r_[0., 0., -0.25])],
        transform=np.dot(sg.translate(0.5, -0.55, 0.), sg.rotx(0.1)))
    floor = T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(4., 4.), oc.LambertianReceiver(0.7))],
        transform=sg.translate(0., 0., -0.8))
    asm = T.assembly.Assembly(objects=[slab, ball, tinted, floor])

    def source():
        direction = np.r_[0.25, -0.1, -1.]
        direction = direction / np.sqrt(np.sum(direction ** 2))
        b = T.sources.disk_bundle(n_rays, np.c_[[-0.2, 0.05, 0.9]], direction, 1.1, 0.2, flux=1000.)
        return T.ray_bundle.RayBundle(vertices=b.get_vertices(), directions=b.get_directions(),
                                      energy=b.get_energy(), ref_index=np.ones(n_rays))
    return asm, source


def c10_polychromatic(T, n_rays=2000, n_samples=5):
    """Rays carrying spectra (bundle properties ``wavelengths`` and ``spectra``, both (W, n)): a floor with the
    polychromatic directional optics and Absorber + Polychromatic + Location accountants, a wall with the same
    optics and the Polychromatic accountant alone, a specular plate and a diffuse lid that pass the spectra on
    unchanged (optics_callables.py:393-425, 1825-1848)."""
    oc = T.optics_callables
    thetas = np.linspace(0., np.pi / 2., 9)
    wls = np.linspace(0.25e-6, 3.0e-6, 14)
    table = 0.15 + 0.7 * np.outer(np.cos(thetas) ** 0.5, np.linspace(1., 0.35, 14))
    poly = oc.Lambertian_directional_axisymmetric_piecewise_PolychromaticAbsorberPolychromaticLocation
    poly2 = oc.Lambertian_directional_axisymmetric_piecewise_PolychromaticPolychromatic
    objs = [
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(4., 4.), poly(thetas, table, wls))]),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(4., 2.), poly2(thetas, table[::-1], wls))],
                                 transform=np.dot(T.spatial_geometry.translate(2., 0., 1.), T.spatial_geometry.roty(-np.pi / 2.))),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(4., 2.), oc.ReflectiveReceiver(0.2))],
                                 transform=np.dot(T.spatial_geometry.translate(-2., 0., 1.), T.spatial_geometry.roty(np.pi / 2.))),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(4., 4.), oc.LambertianPolychromatic(0.3))],
                                 transform=np.dot(T.spatial_geometry.translate(0., 0., 2.), T.spatial_geometry.rotx(np.pi))),
    ]
    asm = T.assembly.Assembly(objects=objs)

    def source():
        direction = np.r_[0.35, 0.1, -1.]
        direction = direction / np.sqrt(np.sum(direction ** 2))
        b = T.sources.disk_bundle(n_rays, np.c_[[-0.3, 0., 1.9]], direction, 1.2, 0.05, flux=1000.)
        # every ray samples its own wavelengths (sorted, inside the table's range) and a smooth spectrum
        wl = np.sort(np.random.uniform(0.3e-6, 2.8e-6, size=(n_samples, n_rays)), axis=0)
        spectra = (1. + 0.5 * np.sin(wl * 3e6)) * np.random.uniform(0.5, 1.5, size=n_rays) * b.get_energy() / 2.5e-6
        return T.ray_bundle.RayBundle(vertices=b.get_vertices(), directions=b.get_directions(),
                                      energy=b.get_energy(), wavelengths=wl, spectra=spectra)
    return asm, source


def c11_bifacial(T, n_rays=2000):
    """BiFacial optics managers (optics_callables.py:1877-1924) on plates that are hit from both sides: a tilted
    plate diffuse in front / specular with slope error at the back (both sides with accountants), a plate
    specular in front / absorbing at the back, inside a diffuse box (floor + lid)."""
    oc = T.optics_callables
    thetas = np.linspace(0., np.pi / 2., 6)
    abs_th = np.linspace(0.1, 0.6, 6)
    two_a = oc.BiFacial(oc.LambertianReceiver(0.25), oc.RealReflectiveReceiver(0.1, 3e-3, True))
    two_b = oc.BiFacial(oc.ReflectiveDetector(0.15), oc.Lambertian_directional_axisymmetric_piecewiseAbsorber(thetas, abs_th))
    two_c = oc.BiFacial(oc.Reflective(0.05), oc.LambertianSpecular(0.3, 0.4))
    objs = [
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(2., 2.), two_a)],
                                 transform=np.dot(T.spatial_geometry.translate(-0.8, 0., 1.), T.spatial_geometry.roty(0.9))),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(2., 2.), two_b)],
                                 transform=np.dot(T.spatial_geometry.translate(0.9, 0., 1.1), T.spatial_geometry.roty(-1.1))),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RoundPlateGM(0.7), two_c)],
                                 transform=np.dot(T.spatial_geometry.translate(0., 0.2, 1.55), T.spatial_geometry.rotx(0.3))),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(6., 6.), oc.LambertianReceiver(0.2))]),
        T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(6., 6.), oc.LambertianReceiver(0.2))],
                                 transform=np.dot(T.spatial_geometry.translate(0., 0., 2.2), T.spatial_geometry.rotx(np.pi))),
    ]
    asm = T.assembly.Assembly(objects=objs)

    def source():
        return T.sources.disk_bundle(n_rays, np.c_[[0., 0., 0.05]], np.r_[0., 0., 1.], 1.6, np.pi / 3., flux=1000.)
    return asm, source


def c12_refractive_materials(T, n_rays=2000):
    """Refraction between media given by material objects (complex m(lambda) from a table; optics_callables.py
    :726-858, :908-944): a glass slab split in the two modes of ``Refractive`` (both children | one child drawn from
    the Fresnel reflectance, rough surface), under it an absorbing slab (``RefractiveAbsorbant``: attenuation from
    the imaginary index), a diffuse receiver below.  The rays carry one wavelength each and start in air."""
    oc = T.optics_callables
    ocn = _optical_constants(T)
    here = os.path.dirname(os.path.abspath(__file__))
    glass = ocn.OpticalMaterialFromFile(os.path.join(here, 'data', 'glass_nk.csv'))
    air = ocn.Air()
    sg = T.spatial_geometry

    def plate(opt, z, flip=False, w=3., h=3.):
        tr = sg.translate(0., 0., z)
        if flip:
            tr = np.dot(tr, sg.rotx(np.pi))
        return T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(w, h), opt)], transform=tr)
    objs = [
        plate(oc.Refractive(air, glass, single_ray=False), 1.0),
        plate(oc.RefractiveDetector(air, glass, single_ray=True, sigma=3e-3), 0.7, flip=True),     # accountants see the
                                                                         # [reflected][refracted] child bundle as it is
        plate(oc.RefractiveAbsorbant(air, glass, single_ray=False, attenuation_coefficient_1=1., scaling=2.), 0.4),
        # smooth single_ray surface with accountants of the incoming rays: wavelengths and normals in hit order
        plate(oc.RefractiveAbsorbantSpectralNormal(air, glass, single_ray=True, attenuation_coefficient_2=1.), 0.1,
              flip=True),
        plate(oc.LambertianReceiver(0.6), -0.3, w=8., h=8.),
    ]
    asm = T.assembly.Assembly(objects=objs)

    def source():
        direction = np.r_[0.25, 0.1, -1.]
        direction = direction / np.sqrt(np.sum(direction ** 2))
        b = T.sources.disk_bundle(n_rays, np.c_[[-0.2, 0., 1.6]], direction, 0.9, 0.1, flux=1000.)
        wl = np.random.uniform(3.5e-7, 2.4e-6, size=n_rays)
        return T.ray_bundle.RayBundle(vertices=b.get_vertices(), directions=b.get_directions(), energy=b.get_energy(),
                                      wavelengths=wl, ref_index=air.m(wl))
    return asm, source


def _optical_constants(T):
    """The reference keeps its materials outside the tracer package."""
    if T.pkg == 'tracer':
        return importlib.import_module('ray_trace_utils.optical_constants')
    return T.optical_constants


def c8_fresnel_conductor(T, n_rays=2000):
    """A metal plate (Fresnel reflectance from tabulated n, k) under a spectrally selective
    mirror lid; the source bundle carries one wavelength per ray."""
    oc = T.optics_callables
    here = os.path.dirname(os.path.abspath(__file__))
    metal = _optical_constants(T).OpticalMaterialFromFile(os.path.join(here, 'data', 'conductor_nk.csv'))
    wls = np.linspace(3e-7, 2.5e-6, 9)
    lid_abs = np.linspace(0.05, 0.6, 9)
    objs = [
        T.object.AssembledObject(surfs=[T.surface.Surface(
            T.flat_surface.RectPlateGM(4., 4.), oc.FresnelConductorHomogenousReceiver(1., metal))]),
        T.object.AssembledObject(
            surfs=[T.surface.Surface(T.flat_surface.RoundPlateGM(3.), oc.Reflective_spectralReceiver(lid_abs, wls))],
            transform=np.dot(T.spatial_geometry.translate(0., 0., 1.), T.spatial_geometry.rotx(np.pi))),
    ]
    asm = T.assembly.Assembly(objects=objs)

    def source():
        direction = np.r_[0.5, 0.2, -1.]
        direction = direction / np.sqrt(np.sum(direction ** 2))
        b = T.sources.disk_bundle(n_rays, np.c_[[-0.4, -0.15, 0.9]], direction, 0.8, 0.3, flux=1000.)
        wl = np.random.uniform(3.2e-7, 2.4e-6, size=n_rays)
        return T.ray_bundle.RayBundle(vertices=b.get_vertices(), directions=b.get_directions(),
                                      energy=b.get_energy(), wavelengths=wl)
    return asm, source


def gm_zoo(T):
    """name -> (geometry manager factory, frame) covering every supported GM kind."""
    sg = T.spatial_geometry
    tilt = np.dot(sg.translate(0.1, -0.2, 0.3), np.dot(sg.rotx(0.3), sg.roty(-0.2)))
    F, P, S, C, Y, Q = T.flat_surface, T.paraboloid, T.sphere_surface, T.cone, T.cylinder, T.quadratic_surface
    return {
        'flat': (lambda: F.FlatGeometryManager(), tilt),
        'rect': (lambda: F.RectPlateGM(1.5, 1.), tilt),
        'rect_extruded': (lambda: F.ExtrudedRectPlateGM(1.5, 1., np.c_[[0.1, 0.05]], 0.4, 0.3), tilt),
        'round': (lambda: F.RoundPlateGM(0.9, 0.3), tilt),
        'round_cut': (lambda: F.StraightCutRoundPlateGM(0.9, 0.4), tilt),
        'triangle': (lambda: T.triangular_face.TriangularFace(np.array([[1., 0.2], [0., 0.9], [0., 0.]])), tilt),
        'paraboloid': (lambda: P.Paraboloid(1.5, 2.), tilt),
        'dish': (lambda: P.ParabolicDishGM(2., 1.), tilt),
        'hex_dish': (lambda: P.HexagonalParabolicDishGM(2., 1.), tilt),
        'rect_dish': (lambda: P.RectangularParabolicDishGM(1.5, 1., 1.), tilt),
        'parab_cyl': (lambda: P.ParabolicCylinder(1.5), tilt),
        'trough': (lambda: P.ParabolicTroughGM(2., 1., 1.5), tilt),
        'sphere': (lambda: S.SphericalGM(0.8), tilt),
        'hemisphere': (lambda: S.HemisphereGM(0.8), tilt),
        'cut_sphere': (lambda: S.CutSphereGM(0.8, T.boundary_shape.BoundarySphere(radius=1.6, location=np.r_[0.1, -0.2, -1.1])), tilt),
        'cut_sphere_plane': (lambda: S.CutSphereGM(0.8, T.boundary_shape.BoundaryPlane(
            location=np.r_[0.05, 0., -0.2], rotation=sg.rotx(0.4)[:3, :3])), tilt),
        'sphere_rect': (lambda: S.SphericalRectFacet(0.8, 0.9, 0.7), tilt),
        'cone_inf': (lambda: C.InfiniteCone(0.5, 0.1), tilt),
        'cone': (lambda: C.FiniteCone(0.7, 1.2), tilt),
        'frustum': (lambda: C.ConicalFrustum(0.1, 0.3, 0.9, 0.8), tilt),
        'frustum_rect': (lambda: C.RectCutConicalFrustum(0.1, 0.5, 0.9, 0.9, 1.2, 1.5), tilt),
        'cyl_inf': (lambda: Y.InfiniteCylinder(1.2), tilt),
        'cyl': (lambda: Y.FiniteCylinder(1.2, 1., [0.3, 5.5]), tilt),
        'cyl_rect': (lambda: Y.RectCutCylinder(1.2, 1., 1.0, 1.1), tilt),
        'quadric_flat': (lambda: Q.FlatQuadricSurfaceGM(0.2, 0.1, 0.05, 0.02, -0.03, 0.01), tilt),
        'quadric_rect': (lambda: Q.RectFlatQuadricSurfaceGM(1.5, 1., 0.2, 0.1, 0.05, 0.02, -0.03, 0.01), tilt),
        'rect_perforated': (lambda: F.PerforatedRectPlateGM(1.8, 1.4, np.array([[0.3, 0.2], [-0.4, -0.3], [0.5, -0.4]]),
                                                           np.array([0.25, 0.2, 0.12])), tilt),
        'polygon': (lambda: T.polygon.FlatSimplePolygonGM(_STAR), tilt),
        'polygon_perforated': (lambda: T.polygon.PerforatedPolygonGM(_STAR, np.array([[0.1, 0.1], [-0.3, 0.35]]),
                                                                     np.array([0.2, 0.1])), tilt),
        'ellipsoid': (lambda: T.ellipsoid.EllipsoidGM(0.9, 0.6, 0.75), tilt),
        'ellipsoid_cut': (lambda: T.ellipsoid.EllipsoidGM(0.9, 0.6, 0.75, xlim=[-0.5, 0.7], zlim=[-0.6, 0.2]), tilt),
        'ellipsoid_all_limits': (lambda: T.ellipsoid.EllipsoidGM(0.9, 0.6, 0.75, [-0.2, 0.2], [-0.2, 0.2], [-0.2, 0.2]), tilt),
    }


# a concave 8-vertex star, clockwise, with one vertical and one horizontal edge
_STAR = np.array([[0.0, 0.3, 0.9, 0.35, 0.5, 0.0, -0.8, -0.8],
                  [0.9, 0.3, 0.2, -0.2, -0.8, -0.35, -0.6, 0.4]])


def zoo_scene(T, name, optics=None):
    make, frame = gm_zoo(T)[name]
    opt = optics if optics is not None else T.optics_callables.ReflectiveReceiver(0.1)
    # ReflectiveReceiver exists in both packages; add normals through a second accountant class
    obj = T.object.AssembledObject(surfs=[T.surface.Surface(make(), opt)], transform=frame.copy())
    return T.assembly.Assembly(objects=[obj])


def zoo_rays(T, n, seed):
    """Rays from a shell of radius ~3 around the origin aimed at a jittered
    point near the surface, plus rays starting inside the shapes."""
    rng = np.random.RandomState(seed)
    p = rng.normal(size=(3, n))
    p = 3. * p / np.sqrt(np.sum(p ** 2, axis=0))
    inside = rng.rand(n) < 0.25
    p[:, inside] = rng.uniform(-0.5, 0.5, size=(3, int(inside.sum())))
    target = rng.uniform(-0.8, 0.8, size=(3, n)) + np.c_[[0.1, -0.2, 0.3]]
    d = target - p
    d /= np.sqrt(np.sum(d ** 2, axis=0))
    return T.ray_bundle.RayBundle(vertices=p, directions=d, energy=np.ones(n))


def fluxmap_cases(T):
    """name -> (geometry manager, frame, resolution, (3,n) local sample points) for the
    get_fluxmap meshes: rectangle, annulus, dish, cylinder sector, sphere."""
    sg = T.spatial_geometry
    rng = np.random.RandomState(77)
    n = 6000
    frame = np.dot(sg.translate(0.3, -1.2, 2.), np.dot(sg.rotx(0.4), sg.rotz(1.1)))
    cases = {}
    xy = rng.uniform(-1.1, 1.1, size=(2, n)) * np.c_[[1.5, 1.0]]
    cases['rect'] = (T.flat_surface.RectPlateGM(3., 2.), frame, 12, np.vstack([xy, np.zeros(n)]))
    r, a = rng.uniform(0.3, 1.3, n), rng.uniform(0, 2 * np.pi, n)
    cases['annulus'] = (T.flat_surface.RoundPlateGM(1.2, 0.4), frame, 9, np.vstack([r * np.cos(a), r * np.sin(a), np.zeros(n)]))
    cases['disk'] = (T.flat_surface.RoundPlateGM(1.2), frame, 10, np.vstack([r * np.cos(a), r * np.sin(a), np.zeros(n)]))
    dish = T.paraboloid.ParabolicDishGM(2.4, 1.5)
    rd = rng.uniform(0., 1.25, n)
    x, y = rd * np.cos(a), rd * np.sin(a)
    cases['dish'] = (dish, frame, 8, np.vstack([x, y, dish.a * x ** 2 + dish.b * y ** 2]))
    cyl = T.cylinder.FiniteCylinder(2., 1.5, ang_range=[0.5, 4.])
    ac = rng.uniform(0., 2 * np.pi, n)
    cases['cylinder'] = (cyl, frame, 11, np.vstack([np.cos(ac), np.sin(ac), rng.uniform(-0.8, 0.8, n)]))
    d = rng.normal(size=(3, n))
    d /= np.sqrt(np.sum(d ** 2, axis=0))
    cases['sphere'] = (T.sphere_surface.SphericalGM(0.7), frame, 6, 0.7 * d)
    # four meshes around a rectangular hole (flat_surface.py:321-356)
    cases['rect_extruded'] = (T.flat_surface.ExtrudedRectPlateGM(3., 2., np.array([0.3, -0.2]), 1., 0.8), frame, 6,
                              np.vstack([xy, np.zeros(n)]))
    return cases


# ------------------------------------------------------------------ config 5 (view factors)
# Holman, Heat Transfer 8th ed., example 8.17 as the reference holds it (emissive_losses/emissive_losses_test.py:40-44):
# aperture disc, three cylinder bands, back disc of a cylinder with radius = band depth.  Textbook precision
# (2-3 digits); holman_exact() gives the closed forms.
HOLMAN_VF = np.array([[0., 0.63, 0.195, 0.075, 0.1],
                      [0.315, 0.37, 0.2175, 0.06, 0.0375],
                      [0.0975, 0.2175, 0.37, 0.2175, 0.0975],
                      [0.0375, 0.06, 0.2175, 0.37, 0.315],
                      [0.1, 0.075, 0.195, 0.63, 0.]])


def holman_exact(radius=0.01, depth=0.01, sections=3):
    """Exact view factors of the cavity (coaxial-disc formula + enclosure algebra): elements
    0 = aperture disc, 1..sections = bands, sections + 1 = back disc.  Returns (F, areas)."""
    n = sections

    def f(h):                   # disc -> coaxial disc of the same radius at distance h
        if h == 0.:
            return 1.
        X = 2. + (h / radius) ** 2.
        return 0.5 * (X - np.sqrt(X * X - 4.))
    z = depth * np.arange(n + 1)
    a_disc, a_band = np.pi * radius ** 2., 2. * np.pi * radius * depth

    def disc_to_band(a, j):     # virtual disc at z[a] -> band j (between z[j-1] and z[j])
        return abs(f(abs(z[j - 1] - z[a])) - f(abs(z[j] - z[a])))
    F = np.zeros((n + 2, n + 2))
    for j in range(1, n + 1):
        F[0, j] = disc_to_band(0, j)
        F[n + 1, j] = disc_to_band(n, j)
        F[j, 0] = a_disc / a_band * F[0, j]
        F[j, n + 1] = a_disc / a_band * F[n + 1, j]
    F[0, n + 1] = F[n + 1, 0] = f(z[n])
    for i in range(1, n + 1):
        for j in range(1, n + 1):
            if i != j:
                F[i, j] = a_disc / a_band * abs(disc_to_band(i - 1, j) - disc_to_band(i, j))
        F[i, i] = 1. - F[i].sum()
    return F, np.r_[a_disc, [a_band] * n, a_disc]


def c5_holman(T, radius=0.01, depth=0.01, sections=3, device_sources=True):
    """Config 5: the Holman cavity as black Lambertian receivers (view_factors_3D.py:435-475 builds the same
    shapes), one surface per element.  Returns (assembly, element_of_surface, areas, emit) for AssemblyRTVF;
    emit(i, n, seed) is a unit-energy Lambertian bundle leaving element i."""
    oc, sg = T.optics_callables, T.spatial_geometry
    n = sections
    objs = [T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RoundPlateGM(radius), oc.LambertianReceiver(1.))])]
    for k in range(n):
        objs.append(T.object.AssembledObject(
            surfs=[T.surface.Surface(T.cylinder.FiniteCylinder(2. * radius, depth), oc.LambertianReceiver(1.))],
            transform=sg.translate(0., 0., (k + 0.5) * depth)))
    objs.append(T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.RoundPlateGM(radius), oc.LambertianReceiver(1.))],
        transform=np.dot(sg.translate(0., 0., n * depth), sg.rotx(np.pi))))
    asm = T.assembly.Assembly(objects=objs)
    _, areas = holman_exact(radius, depth, n)

    def emit(i, num, seed=0):
        if i == 0 or i == n + 1:
            c = np.c_[[0., 0., 0. if i == 0 else n * depth]]
            d = np.r_[0., 0., 1. if i == 0 else -1.]
            if device_sources:
                return T.sources.disk_bundle_device(num, c, d, radius, np.pi / 2., seed=seed)
            np.random.seed(seed)
            return T.sources.disk_bundle(num, c, d, radius, np.pi / 2.)
        np.random.seed(seed)
        return T.sources.vf_cylinder_bundle(num, radius, depth, np.c_[[0., 0., (i - 0.5) * depth]], np.r_[0., 0., 1.],
                                            rays_in=True)
    return asm, list(range(n + 2)), areas, emit
