"""Known answers of the reference's geometry-manager tests -- ray parameters, hit points and
normals written in tests/test_hemisphere_gm.py, test_paraboloid_gm.py, test_cylinder.py of the reference
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


CASES = [case_hemisphere, case_paraboloid, case_dish_circular, case_dish_hexagonal, case_cylinder_vertical,
         case_cylinder_rotated]


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
    np.testing.assert_array_almost_equal(got, prm)
    hit = np.nonzero(np.isfinite(prm))[0]
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
    hit = np.nonzero(np.isfinite(prm))[0]
    child = eng.tree[1]
    assert np.array_equal(child.get_parents(), hit)
    np.testing.assert_array_almost_equal(child.get_vertices(), pos[:, hit] + prm[hit] * d[:, hit])
    _, where, normals = opt.get_all_hits()
    if pts is not None:
        np.testing.assert_array_almost_equal(where, pts[:, hit])
    if nrm is not None:
        np.testing.assert_array_almost_equal(normals, nrm[:, hit])
