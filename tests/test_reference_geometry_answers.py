"""Known answers of the reference's geometry-manager tests -- ray parameters, hit points and
normals written in tests/test_hemisphere_gm.py, test_paraboloid_gm.py, test_cylinder.py,
test_flat_geometry_manager.py, test_rect_plate.py, test_round_plate.py, test_triangular_face.py, test_cut_sphere.py, test_parab_dish.py of the
reference
-- against the oracle's find_intersections (CPU) and the CUDA kernels (GPU: a one-bounce trace whose
Location / Normal accountants and child vertices carry what the kernels computed).  These vectors come on
top of tests/golden/gm_zoo.npz (4000 random rays per geometry kind, recorded from the reference)."""
import numpy as np
import pytest

from tests import scenes

T = scenes.NS('tracer_b200')


def _ring(num=10):
    theta = np.linspace(0, 2 * np.pi, num, endpoint=False)
    pos = np.vstack((np.cos(theta), np.sin(theta), np.ones(num)))
    return pos, np.tile(np.c_[[0, 0, -1.]], (1, num))


def case_hemisphere():
    """tests/test_hemisphere_gm.py:9-39: prm = 1 + 2 sin(pi/3), z = -2 sin(pi/3), normals point to the axis."""
    pos, d = _ring()
    prm = np.full(10, 1 + 2 * np.sin(np.pi / 3))
    pts = np.vstack((pos[:2], np.full(10, -2 * np.sin(np.pi / 3))))
    return (lambda: T.sphere_surface.HemisphereGM(radius=2.)), np.eye(4), pos, d, prm, pts, None


def case_paraboloid():
    """tests/test_paraboloid_gm.py:8-42: Paraboloid(a=5, b=5): prm = 0.96, z = 0.04."""
    pos, d = _ring()
    pts = np.vstack((pos[:2], np.full(10, 0.04)))
    return (lambda: T.paraboloid.Paraboloid(a=5., b=5.)), np.eye(4), pos, d, np.full(10, 0.96), pts, None


def _dish_rays():
    d = np.c_[[0., 0, -1], [0, 1, -1], [0, 11, -2], [0, 1, 0]]
    d /= np.sqrt(np.sum(d ** 2, axis=0))
    pos = np.c_[[0., 0, 1], [0, -1, 1], [0, -11, 2], [0, 1, 1]]
    return pos, d, np.r_[1., np.sqrt(2), np.sqrt(11 ** 2 + 4), np.inf]


def case_dish_circular():
    """tests/test_paraboloid_gm.py:44-61: rays above the cut plane of a ParabolicDishGM miss."""
    pos, d, prm = _dish_rays()
    return (lambda: T.paraboloid.ParabolicDishGM(diameter=10., focal_length=12.)), np.eye(4), pos, d, prm, None, None


def case_dish_hexagonal():
    """tests/test_paraboloid_gm.py:63-69."""
    pos, d, prm = _dish_rays()
    return (lambda: T.paraboloid.HexagonalParabolicDishGM(diameter=10., focal_length=12.)), np.eye(4), pos, d, prm, None, None


def _cylinder_rays():
    pos = np.c_[[0., 0., 0.], [0., 0., 0.], [0., 1., 0.], [0., 1., 0.]]
    d = np.c_[[0., 1., 0.], [0., 1., 1.], [0., -1., 0.], [0., -1., 1.]]
    return pos, d / np.sqrt(np.sum(d ** 2, axis=0))


def case_cylinder_vertical():
    """tests/test_cylinder.py:9-36: InfiniteCylinder(diameter=1), two rays inside, two outside."""
    pos, d = _cylinder_rays()
    prm = np.r_[0.5, np.sqrt(0.5), 0.5, np.sqrt(0.5)]
    nrm = np.c_[[0., -1., 0.], [0., -1., 0.], [0., 1., 0.], [0., 1., 0.]]
    pts = np.c_[[0., 0.5, 0.], [0., 0.5, 0.5], [0., 0.5, 0.], [0., 0.5, 0.5]]
    return (lambda: T.cylinder.InfiniteCylinder(diameter=1.)), np.eye(4), pos, d, prm, pts, nrm


def case_cylinder_rotated():
    """tests/test_cylinder.py:38-62: the same cylinder rotated by 30 degrees about x."""
    pos, d = _cylinder_rays()
    ydist = 0.5 / np.cos(np.pi / 6)
    p1 = np.sin(2 * np.pi / 3) * (1 - ydist) / np.sin(np.pi / 12)
    p2 = np.sin(np.pi / 3) * ydist / np.sin(5 * np.pi / 12)
    prm = np.r_[ydist, p2, 1 - ydist, p1]
    out = np.r_[0., np.cos(np.pi / 6), 0.5]
    nrm = np.c_[-out, -out, out, out]
    pts = np.c_[[0., ydist, 0.], [0., p2 * np.sqrt(0.5), p2 * np.sqrt(0.5)],
                [0., ydist, 0.], [0., 1 - p1 * np.sqrt(0.5), p1 * np.sqrt(0.5)]]
    return (lambda: T.cylinder.InfiniteCylinder(diameter=1.)), T.spatial_geometry.rotx(np.pi / 6), pos, d, prm, pts, nrm


def case_flat():
    """tests/test_flat_geometry_manager.py:10-52: prm = sqrt(3), normals +z, all points at the origin but one."""
    d = np.c_[[1, 1, -1], [-1, 1, -1], [-1, -1, -1], [1, -1, -1]] / np.sqrt(3)
    pos = np.c_[[0, 0, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]].astype(float)
    pts = np.zeros((3, 4))
    pts[:2, 0] = 1
    return (lambda: T.flat_surface.FlatGeometryManager()), np.eye(4), pos, d, np.full(4, np.sqrt(3)), pts, \
        np.tile(np.c_[[0, 0, 1.]], (1, 4))


def case_flat_tilted():
    """tests/test_flat_geometry_manager.py:54-107: the plane rotated by -45 degrees about x."""
    s2 = np.sqrt(2)
    d = np.c_[[1, 0, -s2], [-1, 0, -s2], [-1, -s2, 0], [1, -s2, 0]] / np.sqrt(3)
    pos = np.c_[[0, 1 / s2, 1 / s2], [1, 0, s2], [1, s2, 0], [-1, s2, 0]]
    pts = np.zeros((3, 4))
    pts[:, 0] = np.r_[1, 1 / s2, -1 / s2]
    frame = T.spatial_geometry.generate_transform(np.r_[1., 0, 0], -np.pi / 4., np.zeros((3, 1)))
    return (lambda: T.flat_surface.FlatGeometryManager()), frame, pos, d, np.full(4, np.sqrt(3)), pts, \
        np.tile(np.c_[[0, 1 / s2, 1 / s2]], (1, 4))


def _down_rays(xs):
    pos = np.zeros((3, len(xs)))
    pos[0] = xs
    pos[2] = 1.
    return pos, np.tile(np.c_[[0, 0, -1.]], (1, len(xs)))


def _mask_case(make, frame, pos, d, misses):
    prm = np.where(misses, np.inf, np.nan)          # nan: hit, parameter not stated by the reference test
    return make, frame, pos, d, prm, None, None


def case_rect_plate():
    """tests/test_rect_plate.py:17-27: misses = [F, F, T, T]."""
    pos, d = _down_rays(np.r_[0, 0.5, 2, -2])
    return _mask_case(lambda: T.flat_surface.RectPlateGM(1, 0.25), np.eye(4), pos, d, np.r_[False, False, True, True])


def case_rect_plate_extruded():
    """tests/test_rect_plate.py:43-55: a ray inside the rectangular hole misses."""
    pos, d = _down_rays(np.r_[0, 0.05, 0.2, -0.3])
    make = (lambda: T.flat_surface.ExtrudedRectPlateGM(width=1, height=1, extr_center=np.vstack([0.2, 0.1]),
                                                       extr_width=0.2, extr_height=0.4))
    return _mask_case(make, np.eye(4), pos, d, np.r_[False, False, True, False])


def case_round_plate():
    """tests/test_round_plate.py:17-27."""
    pos, d = _down_rays(np.r_[0, 0.5, 2, -2])
    return _mask_case(lambda: T.flat_surface.RoundPlateGM(1), np.eye(4), pos, d, np.r_[False, False, True, True])


def _triangle(frame):
    x, y = np.mgrid[-1:1:2j, -1:1:3j]
    pos = np.vstack((x.flatten(), y.flatten(), np.ones(6) * 10))
    pos[1] -= 10.
    d = np.tile(np.c_[[0., 10., -10.]], (1, 6))
    verts = np.zeros((3, 2))
    verts[1] = 2.
    verts[0] = np.r_[4., -4.]
    return _mask_case(lambda: T.triangular_face.TriangularFace(verts), frame, pos, d,
                      ~np.r_[False, False, True, False, False, True])


def case_triangle():
    """tests/test_triangular_face.py:9-29: exclusion by the barycentric test, global frame."""
    return _triangle(np.eye(4))


def case_triangle_transformed():
    """tests/test_triangular_face.py:31-37."""
    sg = T.spatial_geometry
    return _triangle(np.dot(sg.translate(0, -1, 0.5), sg.rotx(np.pi / 4.)))


def case_cut_sphere():
    """tests/test_cut_sphere.py:10-43: a bounding sphere placed so that the lower hemisphere remains."""
    pos, d = _ring()
    prm = np.full(10, 1 + 2 * np.sin(np.pi / 3))
    pts = np.vstack((pos[:2], np.full(10, -2 * np.sin(np.pi / 3))))
    make = (lambda: T.sphere_surface.CutSphereGM(
        2., T.boundary_shape.BoundarySphere(radius=4., location=np.r_[0., 0., -4 * np.sqrt(3) / 2.])))
    return make, np.eye(4), pos, d, prm, pts, None


def _dish_selection(frame):
    pos = np.zeros((3, 4))
    pos[0] = np.r_[0, 0.5, 2, -2]
    pos[2] = 2.
    d = np.tile(np.c_[[0, 0, -1.]], (1, 4))
    return _mask_case(lambda: T.paraboloid.ParabolicDishGM(2., 1.), frame, pos, d, np.r_[False, False, True, True])


def case_parab_dish_selection():
    """tests/test_parab_dish.py:11-28."""
    return _dish_selection(np.eye(4))


def case_parab_dish_transformed():
    """tests/test_parab_dish.py:30-36: translated and rotated dish."""
    return _dish_selection(T.spatial_geometry.generate_transform(np.r_[1., 0., 0.], np.pi / 4., np.c_[[0., 0., 1.]]))


CASES = [case_hemisphere, case_paraboloid, case_dish_circular, case_dish_hexagonal, case_cylinder_vertical,
         case_cylinder_rotated, case_flat, case_flat_tilted, case_rect_plate, case_rect_plate_extruded,
         case_round_plate, case_triangle, case_triangle_transformed, case_cut_sphere, case_parab_dish_selection,
         case_parab_dish_transformed]


def _surface(make, frame, optics):
    surf = T.surface.Surface(make(), optics)
    obj = T.object.AssembledObject(surfs=[surf], transform=frame)
    return T.assembly.Assembly(objects=[obj]), surf


@pytest.mark.parametrize('case', CASES, ids=lambda c: c.__name__[5:])
def test_oracle_geometry_known_answers(case):
    from oracle.numpy_engine import find_intersections
    make, frame, pos, d, prm, pts, nrm = case()
    _, surf = _surface(make, frame, T.optics_callables.Reflective(0.))
    got, points, normals = find_intersections(surf, pos, d)
    got = np.where(got == 0., np.inf, got)
    assert np.array_equal(np.isinf(got), np.isinf(prm)), 'hit / miss pattern'
    stated = np.isfinite(prm)
    np.testing.assert_array_almost_equal(got[stated], prm[stated])
    hit = np.nonzero(~np.isinf(prm))[0]
    if pts is not None:
        np.testing.assert_array_almost_equal(points(hit), pts[:, hit])
    if nrm is not None:
        np.testing.assert_array_almost_equal(normals(hit), nrm[:, hit])


@pytest.mark.gpu
@pytest.mark.parametrize('case', CASES, ids=lambda c: c.__name__[5:])
def test_gpu_geometry_known_answers(case):
    from tracer_b200.tracer_engine import TracerEngine
    make, frame, pos, d, prm, pts, nrm = case()
    opt = T.optics_callables.ReflectiveNormalReceiver(0.)      # accountants: Absorber, Location, Normal
    asm, _ = _surface(make, frame, opt)
    bundle = T.ray_bundle.RayBundle(pos.copy(), d.copy(), energy=np.ones(pos.shape[1]))
    eng = TracerEngine(asm, device_accel='brute')
    eng.ray_tracer(bundle, reps=1, min_energy=-1., tree=True)
    hit = np.nonzero(~np.isinf(prm))[0]
    child = eng.tree[1]
    assert np.array_equal(child.get_parents(), hit)
    stated = np.isfinite(prm[hit])
    np.testing.assert_array_almost_equal(child.get_vertices()[:, stated],
                                         (pos[:, hit] + prm[hit] * d[:, hit])[:, stated])
    _, where, normals = opt.get_all_hits()
    if pts is not None:
        np.testing.assert_array_almost_equal(where, pts[:, hit])
    if nrm is not None:
        np.testing.assert_array_almost_equal(normals, nrm[:, hit])
