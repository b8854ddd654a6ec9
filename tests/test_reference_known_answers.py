"""The reference's own known-answer tests of the bounce loop, restated against (a) the
NumPy oracle on the CPU and (b) the CUDA path on the GPU.  Each case cites the reference test it
comes from; expected values are the ones written in the reference's test files, tolerances theirs
(assert_array_almost_equal, 6 decimals unless stated).

  tests/test_tracer_engine.py  TestTraceProtocol1 / 3 / 4 / 5 / 6   (lines 26-232)
  tests/test_objects.py        TestObjectBuilding1 / 2, TestAssemblyBuilding4   (lines 20-170)
  tests/models/test_one_sided_mirror.py  TestRectOneSided (lines 10-52)
  tests/test_tracer_tree.py    TestTraceTree... final order / absorbed to back (lines 106-153) are in
                               tests/test_gpu_parity.py::test_culling_and_parent_bookkeeping
  tests/test_opt_callable.py, test_tracer_engine.py nested lens: tests/test_gpu_refractive.py

Not restated: tests/test_tracer_tree.py TestTree.test_tree1 / test_tree2 and TestTree2 (:19-104) expect
parent lists of an earlier child order (e.g. [2, 3, 1] inside one bounce) and go through CutSphereGM,
whose find_intersections raises inside the reference's intersect_ray under numpy 2; the
current reference does not produce those lists either.
"""
import math

import numpy as np
import pytest

from tests import scenes

T = scenes.NS('tracer_b200')


def _planes(*rotations):
    sg = T.spatial_geometry
    obj = T.object.AssembledObject()
    for rot in rotations:
        kw = {} if rot is None else dict(rotation=sg.general_axis_rotation([1, 0, 0], rot))
        obj.add_surface(T.surface.Surface(T.flat_surface.FlatGeometryManager(), T.optics_callables.perfect_mirror, **kw))
    asm = T.assembly.Assembly()
    asm.add_object(obj)
    return asm


def case_protocol1():
    """TestTraceProtocol1.test_ray_tracer (:26-62): one flat mirror, one iteration."""
    d = np.array([[1, 1, -1], [-1, 1, -1], [-1, -1, -1], [1, -1, -1]]).T / math.sqrt(3)
    pos = np.c_[[0, 0, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]].astype(float)
    correct = np.zeros((3, 4))
    correct[:2, 0] = 1
    return _planes(None), T.ray_bundle.RayBundle(pos, d, energy=np.ones(4)), 1, correct, 6


def case_protocol3():
    """TestTraceProtocol3.test_ray_tracer1 (:97-135): two planes at +-45 degrees."""
    x = 1 / math.sqrt(2)
    b = T.ray_bundle.RayBundle(np.c_[[0, 0, 1.], [0, 0, 1.]], np.c_[[0, x, -x], [0, 1., 0]], energy=np.ones(2))
    return _planes(np.pi / 4, np.pi / (-4)), b, 1, np.c_[[0, .5, .5], [0, 1, 1]], 6


def _protocol4(reps, correct):
    x = 1 / math.sqrt(2)
    b = T.ray_bundle.RayBundle(np.c_[[0, 2., 1.], [0, 2., 1.]], np.c_[[0, -x, x], [0, 0, -1.]], energy=np.ones(2))
    return _planes(np.pi / 4, None), b, reps, correct, 6


def case_protocol4_one():
    """TestTraceProtocol4.test_ray_tracer1 (:137-164): the rays hit different surfaces."""
    return _protocol4(1, np.c_[[0, 1.5, 1.5], [0, 2, 0]])


def case_protocol4_two():
    """TestTraceProtocol4.test_ray_tracer2 (:166-170)."""
    return _protocol4(2, np.c_[[0, 2, 2], [0, 3, 0]])


def case_protocol4_escape():
    """TestTraceProtocol4.test_too_much_intersections (:172-177): the tracer stops when every ray escapes."""
    return _protocol4(42, np.array([]).reshape(3, 0))


def case_protocol5():
    """TestTraceProtocol5.test_ray_tracer1 (:179-200): a hemisphere seen from inside and outside, 3 decimals."""
    sg = T.spatial_geometry
    surf = T.surface.Surface(T.sphere_surface.HemisphereGM(1.), T.optics_callables.perfect_mirror,
                             rotation=sg.general_axis_rotation(np.r_[1, 0, 0], np.pi))
    b = T.ray_bundle.RayBundle(energy=np.ones(3))
    b.set_directions(np.c_[[0, 1., 0], [0, 1., 0], [0, -1., 0]])
    b.set_vertices(np.c_[[0, -2., 0.001], [0, 0, 0.001], [0, 2, 0.001]])
    obj = T.object.AssembledObject()
    obj.add_surface(surf)
    asm = T.assembly.Assembly()
    asm.add_object(obj)
    return asm, b, 1, np.c_[[0, -1, 0], [0, 1, 0], [0, 1, 0]], 3


def case_protocol6():
    """TestTraceProtocol6.test_ray_tracers1 (:202-232): two facing hemispheres in two objects."""
    sg = T.spatial_geometry
    s1 = T.surface.Surface(T.sphere_surface.HemisphereGM(2.), T.optics_callables.perfect_mirror,
                           rotation=sg.general_axis_rotation(np.r_[1, 0, 0], np.pi / 2.))
    s2 = T.surface.Surface(T.sphere_surface.HemisphereGM(2.), T.optics_callables.perfect_mirror,
                           location=np.array([0, -2., 0]), rotation=sg.general_axis_rotation(np.r_[1, 0, 0], -np.pi / 2.))
    b = T.ray_bundle.RayBundle()
    b.set_directions(np.c_[[0, 1., 0]])
    b.set_vertices(np.c_[[0, -1., 0]])
    b.set_energy(np.r_[[1.]])
    b.set_ref_index(np.r_[[1.]])
    asm = T.assembly.Assembly()
    for s in (s1, s2):
        obj = T.object.AssembledObject()
        obj.add_surface(s)
        asm.add_object(obj)
    return asm, b, 1, np.c_[[0, 2, 0]], 6


def _two_hemispheres():
    sg = T.spatial_geometry
    s1 = T.surface.Surface(T.sphere_surface.HemisphereGM(3.), T.optics_callables.perfect_mirror,
                           location=np.array([0, 0, -1.]), rotation=sg.general_axis_rotation(np.r_[1, 0, 0], np.pi))
    s2 = T.surface.Surface(T.sphere_surface.HemisphereGM(3.), T.optics_callables.perfect_mirror,
                           location=np.array([0, 0, 1.]))
    obj = T.object.AssembledObject()
    obj.add_surface(s1)
    obj.add_surface(s2)
    asm = T.assembly.Assembly()
    asm.add_object(obj)
    return asm


def case_objects_hemispheres():
    """test_objects.py TestObjectBuilding1.test_object (:20-46): two hemispheres in one object."""
    b = T.ray_bundle.RayBundle(np.c_[[0, 0, -3.], [0, 0, -1.]], np.c_[[0, 0, 1.], [0, 0, 1.]], energy=np.ones(2))
    return _two_hemispheres(), b, 1, np.c_[[0, 0, 2], [0, 0, -2]], 6


def case_objects_translated():
    """TestObjectBuilding1.test_translation (:48-58): the assembly moved by transform_children."""
    asm = _two_hemispheres()
    asm.transform_children(np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 1], [0, 0, 0, 1.]]))
    b = T.ray_bundle.RayBundle(np.c_[[0, 0, -3.], [0, 0, -1.]], np.c_[[0, 0, 1.], [0, 0, 1.]], energy=np.ones(2))
    return asm, b, 1, np.c_[[0, 0, 3], [0, 0, -1]], 6


def case_objects_rotated_and_translated():
    """TestObjectBuilding1.test_rotation_and_translation (:60-76)."""
    asm = _two_hemispheres()
    asm.transform_children(T.spatial_geometry.generate_transform(np.r_[[1, 0, 0]], np.pi / 2, np.c_[[0, 0, 1]]))
    b = T.ray_bundle.RayBundle()
    b.set_vertices(np.c_[[0, -5., 1], [0, 5., 1]])
    b.set_directions(np.c_[[0, 1., 0], [0, 1., 0]])
    b.set_energy(np.r_[[1., 1.]])
    b.set_ref_index(np.r_[[1., 1.]])
    return asm, b, 1, np.c_[[0, -2, 1]], 6


def case_objects_paraboloid():
    """TestAssemblyBuilding4.test_paraboloid1 (:152-170): the default Paraboloid."""
    x = 1. / math.sqrt(2)
    obj = T.object.AssembledObject()
    obj.add_surface(T.surface.Surface(T.paraboloid.Paraboloid(), T.optics_callables.perfect_mirror))
    asm = T.assembly.Assembly()
    asm.add_object(obj)
    b = T.ray_bundle.RayBundle(np.c_[[0, 0, 1.], [0, 0, 1.]], np.c_[[0, 0, -1.], [0, x, -x]], energy=np.ones(2),
                               ref_index=np.ones(2))
    return asm, b, 1, np.c_[[0, 0, 0], [0, 0.618033989, 0.381966011]], 6


def _glass_slab():
    """TestObjectBuilding2 (:78-110): two refractive planes; the expected angles are those of both children
    of every ray (the test predates single_ray mode)."""
    oc = T.optics_callables
    s1 = T.surface.Surface(T.flat_surface.FlatGeometryManager(), oc.RefractiveHomogenous(1., 1.5, single_ray=False),
                           location=np.array([0, 0, -1.]))
    s2 = T.surface.Surface(T.flat_surface.FlatGeometryManager(), oc.RefractiveHomogenous(1.5, 1., single_ray=False),
                           location=np.array([0, 0, 1.]))
    asm = T.assembly.Assembly()
    asm.add_object(T.object.AssembledObject(surfs=[s1, s2]))
    x = 1 / math.sqrt(2)
    b = T.ray_bundle.RayBundle(np.c_[[0, 1., -2.]], np.c_[[0, -x, x]], energy=np.r_[1.], ref_index=np.r_[1.])
    return asm, b


REFRACTION_CASES = [(1, np.r_[0.785398163, -.4908826]), (2, np.r_[-0.7853981])]


@pytest.mark.parametrize('reps,angles', REFRACTION_CASES, ids=['refraction1', 'refraction2'])
def test_oracle_refraction_angles(reps, angles):
    from oracle.numpy_trace import OracleEngine
    asm, bundle = _glass_slab()
    d = OracleEngine(asm).ray_tracer(bundle, reps, .05)[1]
    np.testing.assert_array_almost_equal(np.arctan(d[1] / d[2]), angles)


@pytest.mark.gpu
@pytest.mark.parametrize('reps,angles', REFRACTION_CASES, ids=['refraction1', 'refraction2'])
def test_gpu_refraction_angles(reps, angles):
    from tracer_b200.tracer_engine import TracerEngine
    asm, bundle = _glass_slab()
    d = TracerEngine(asm).ray_tracer(bundle, reps, .05)[1]
    np.testing.assert_array_almost_equal(np.arctan(d[1] / d[2]), angles)


CASES = [case_protocol1, case_protocol3, case_protocol4_one, case_protocol4_two, case_protocol4_escape,
         case_protocol5, case_protocol6, case_objects_hemispheres, case_objects_translated,
         case_objects_rotated_and_translated, case_objects_paraboloid]


@pytest.mark.parametrize('case', CASES, ids=lambda c: c.__name__[5:])
def test_oracle_reproduces_the_references_known_answers(case):
    from oracle.numpy_trace import OracleEngine
    asm, bundle, reps, correct, decimal = case()
    params = OracleEngine(asm).ray_tracer(bundle, reps, .05)[0]
    np.testing.assert_array_almost_equal(params, correct, decimal=decimal)


@pytest.mark.gpu
@pytest.mark.parametrize('accel', ['brute', 'bvh'])
@pytest.mark.parametrize('case', CASES, ids=lambda c: c.__name__[5:])
def test_gpu_reproduces_the_references_known_answers(case, accel):
    from tracer_b200.tracer_engine import TracerEngine
    asm, bundle, reps, correct, decimal = case()
    params = TracerEngine(asm, device_accel=accel).ray_tracer(bundle, reps, .05)[0]
    np.testing.assert_array_almost_equal(params, correct, decimal=decimal)


def _one_sided_case(rotated):
    """tests/models/test_one_sided_mirror.py TestRectOneSided (:10-52): rays from above are reflected with
    10 % of their energy (absorptivity 0.9), rays from below lose everything, rays beside the plate miss."""
    mirror = T.models.one_sided_mirror.rect_one_sided_mirror(1.5, 1.5, 0.9)
    if rotated:
        mirror.set_transform(T.spatial_geometry.roty(np.pi / 4.))
    pos = np.zeros((3, 8))
    pos[0] = np.tile(np.r_[0, 0.5, 2, -2], 2)
    pos[2] = np.repeat(np.r_[1, -1], 4)
    d = np.zeros((3, 8))
    d[2] = np.repeat(np.r_[-1, 1], 4)
    b = T.ray_bundle.RayBundle()
    b.set_vertices(pos)
    b.set_directions(d)
    b.set_energy(np.ones(8) * 1000)
    b.set_ref_index(np.ones(8))
    verts = np.array([[0., 0.5], [0., 0.], [0., -0.5]]) if rotated else np.array([[0., 0.5], [0., 0.], [0., 0.]])
    return T.assembly.Assembly(objects=[mirror]), b, verts


def _check_one_sided(outg, verts):
    np.testing.assert_array_almost_equal(outg.get_vertices()[:, outg.get_energy() > 0], verts)
    np.testing.assert_array_almost_equal(outg.get_energy(), np.r_[100., 100., 0, 0])


@pytest.mark.parametrize('rotated', [False, True], ids=['regular', 'rotated'])
def test_oracle_one_sided_mirror(rotated):
    from oracle.numpy_trace import OracleEngine
    asm, b, verts = _one_sided_case(rotated)
    e = OracleEngine(asm)
    e.ray_tracer(b, 1, 0.05)
    _check_one_sided(e.tree[-1], verts)


@pytest.mark.gpu
@pytest.mark.parametrize('rotated', [False, True], ids=['regular', 'rotated'])
def test_gpu_one_sided_mirror(rotated):
    from tracer_b200.tracer_engine import TracerEngine
    asm, b, verts = _one_sided_case(rotated)
    e = TracerEngine(asm)
    e.ray_tracer(b, 1, 0.05)
    _check_one_sided(e.tree[-1], verts)


@pytest.mark.gpu
def test_gpu_surface_trace_protocol():
    """tests/test_flat_surface.py:12-44 of the reference: Surface.register_incoming / select_rays /
    get_outgoing on a flat mirror, plus the miss masks of tests/test_rect_plate.py:17-27 and
    tests/test_parab_dish.py:25-36 obtained through register_incoming."""
    surf = T.surface.Surface(T.flat_surface.FlatGeometryManager(), T.optics_callables.perfect_mirror)
    d = np.array([[1, 1, -1], [-1, 1, -1], [-1, -1, -1], [1, -1, -1]]).T / math.sqrt(3)
    pos = np.c_[[0, 0, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]].astype(float)
    bund = T.ray_bundle.RayBundle(pos, d, energy=np.ones(4) * 100)
    params = surf.register_incoming(bund)
    np.testing.assert_array_almost_equal(params, np.r_[[math.sqrt(3)] * 4])
    surf.select_rays(np.arange(4))
    outg = surf.get_outgoing()
    correct_pts = np.zeros((3, 4))
    correct_pts[:2, 0] = 1
    np.testing.assert_array_almost_equal(outg.get_vertices(), correct_pts, decimal=12)
    np.testing.assert_array_almost_equal(
        outg.get_directions(), np.c_[[1, 1, 1], [-1, 1, 1], [-1, -1, 1], [1, -1, 1]] / np.sqrt(3), decimal=12)
    np.testing.assert_array_equal(outg.get_energy(), np.ones(4) * 100)
    np.testing.assert_array_equal(outg.get_parents(), np.arange(4))
    surf.select_rays(np.r_[3, 1])
    outg = surf.get_outgoing()
    np.testing.assert_array_equal(outg.get_parents(), np.r_[3, 1])
    surf.done()
    assert not hasattr(surf, '_current_bundle')

    pos = np.zeros((3, 4))
    pos[0] = np.r_[0, 0.5, 2, -2]
    pos[2] = 1.
    down = T.ray_bundle.RayBundle(pos, np.tile(np.c_[[0, 0, -1.]], (1, 4)))
    rect = T.surface.Surface(T.flat_surface.RectPlateGM(1, 0.25), T.optics_callables.perfect_mirror)
    np.testing.assert_array_equal(np.isinf(rect.register_incoming(down)), np.r_[False, False, True, True])
    pos[2] = 2.
    dish = T.surface.Surface(T.paraboloid.ParabolicDishGM(2., 1.), T.optics_callables.perfect_mirror)
    high = T.ray_bundle.RayBundle(pos, np.tile(np.c_[[0, 0, -1.]], (1, 4)))
    np.testing.assert_array_equal(np.isinf(dish.register_incoming(high)), np.r_[False, False, True, True])
    dish.transform_frame(T.spatial_geometry.generate_transform(np.r_[1., 0., 0.], np.pi / 4., np.c_[[0., 0., 1.]]))
    np.testing.assert_array_equal(np.isinf(dish.register_incoming(high)), np.r_[False, False, True, True])
