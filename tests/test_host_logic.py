"""CPU: host-side logic -- C ABI surface, scene compiler, BVH builder, the
API mirror (RayBundle, optics name factory, scene graph)."""
import os
import re
import warnings

import numpy as np
import pytest

from tests import scenes
from tests.golden.make_golden import ZOO_RAYS, ZOO_SEED

T = scenes.NS('tracer_b200')
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    """The shared library loads and exports each function of include/tracer_b200.h
    (no compute call: there is no GPU here)."""
    from tracer_b200 import _lib
    header = open(os.path.join(ROOT, 'include', 'tracer_b200.h')).read()
    header = re.sub(r'/\*.*?\*/', '', header, flags=re.S)
    declared = set(re.findall(r'\b(trc_[a-z0-9_]+)\s*\(', header))
    assert declared, 'no declarations parsed'
    lib = _lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f'{name} declared in the header but not exported'
    assert declared == set(_lib.SIGNATURES), 'ctypes SIGNATURES out of sync with the header'
    assert lib.trc_abi_version() == _lib.ABI_VERSION
    assert _lib.SURFACE_DTYPE.itemsize == 384 and _lib.BVH_NODE_DTYPE.itemsize == 64


def test_bad_arguments_return_status_not_crash():
    from tracer_b200 import _lib
    lib = _lib.load()
    assert lib.trc_sort_work_bytes(-1, 4) == -1
    assert lib.trc_sort_work_bytes(1000, 219) > 0
    with pytest.raises(ValueError):
        _lib.check(lib.trc_nearest_hit(None, 10, None, None, 10, 0, None, None))
    assert b'trc_nearest_hit' in lib.trc_last_error()


@pytest.mark.parametrize('name', sorted(scenes.gm_zoo(T)))
def test_compiled_aabb_is_conservative(name):
    """Every oracle hit point of a bounded surface lies inside its compiled AABB,
    so BVH culling can never drop a real hit."""
    from oracle.numpy_engine import find_intersections
    from tracer_b200.scene import CompiledScene
    asm = scenes.zoo_scene(T, name)
    sc = CompiledScene(asm, upload=False)
    rays = scenes.zoo_rays(T, ZOO_RAYS, ZOO_SEED)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        prm, points, _ = find_intersections(asm.get_surfaces()[0], rays.get_vertices(), rays.get_directions())
    hit = np.nonzero(np.isfinite(prm))[0]
    assert len(hit) > 50
    if len(sc.unbounded):
        assert name in ('flat', 'paraboloid', 'parab_cyl', 'cone_inf', 'cyl_inf', 'quadric_flat')
        return
    p = points(hit)
    assert np.all(p >= sc.aabb_lo[0][:, None]) and np.all(p <= sc.aabb_hi[0][:, None])


def _bvh_reach(nodes, leaf_surfs):
    seen = []
    stack = [0]
    while stack:
        i = stack.pop()
        nd = nodes[i]
        if nd['count'] != 0:
            for k in range(max(nd['count'], 0)):
                seen.append(int(leaf_surfs[nd['first'] + k]))
        else:
            for c in (nd['first'], nd['first'] + 1):
                ch = nodes[c]
                if ch['count'] >= 0:
                    assert np.all(ch['lo'] >= nd['lo']) and np.all(ch['hi'] <= nd['hi'])
                stack.append(int(c))
    return seen


@pytest.mark.parametrize('n,leaf', [(1, 2), (2, 2), (3, 1), (5, 2), (219, 2), (1000, 4)])
def test_bvh_builder_invariants(n, leaf):
    from tracer_b200.scene import build_bvh
    rng = np.random.RandomState(n)
    c = rng.uniform(-10, 10, size=(n, 3))
    half = rng.uniform(0.1, 1.0, size=(n, 3))
    nodes, order = build_bvh(c - half, c + half, leaf)
    seen = _bvh_reach(nodes, order)
    assert sorted(seen) == list(range(n)), 'every box in exactly one leaf'
    for nd in nodes:
        if nd['count'] > 0:
            idx = order[nd['first']:nd['first'] + nd['count']]
            assert np.all((c - half)[idx] >= nd['lo']) and np.all((c + half)[idx] <= nd['hi'])
            assert nd['count'] <= leaf


def test_nsttf_scene_compiles_like_the_survey_says():
    from tracer_b200 import _lib as L
    from tracer_b200.scene import CompiledScene
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    assert len(asm.get_objects()) == 655
    sc = CompiledScene(asm, upload=False)
    assert sc.num_surfaces == 219
    r = sc.records
    # receiver first? no: sub-assemblies come first (assembly.py:60-65) -> heliostats 0..217, receiver 218
    assert np.all(r['gkind'] == L.G_RECT_PLATE)
    assert np.all(r['okind'][:218] == L.O_REAL_REFLECTIVE) and r['okind'][218] == L.O_REFLECTIVE
    assert np.all(r['flags'] & L.F_ONE_SIDED)
    assert sc.optics[218]['accountants'] and not sc.optics[0]['accountants']
    inv = r['inv'].reshape(-1, 3, 4)
    fr = r['frame'].reshape(-1, 3, 4)
    for k in (0, 100, 218):
        f4 = np.vstack((fr[k], [0, 0, 0, 1]))
        assert np.allclose(np.vstack((inv[k], [0, 0, 0, 1])) @ f4, np.eye(4), atol=1e-12)


def test_optics_name_factory():
    oc = T.optics_callables
    r = oc.ReflectiveReceiver(0.3)
    assert [type(a).__name__ for a in r.accountants] == ['AbsorptionAccountant', 'LocationAccountant']
    d = oc.OneSidedRealReflectiveDetector(0.1, 1e-3, True)
    assert [a.shorthand for a in d.accountants] == ['Absorber', 'Location', 'Directional']
    assert type(d._opt).__name__ == 'OneSidedRealReflective' and d._opt.bi_var is True
    t = oc.LambertianTransmitter(0.2)
    assert [a.shorthand for a in t.accountants] == ['Scatterer', 'Location']
    x = oc.LambertianSpecularDirectionalReceiver(0.1, 0.5)
    assert type(x._opt).__name__ == 'LambertianSpecular'
    assert [a.shorthand for a in x.accountants] == ['Absorber', 'Location', 'Directional']
    assert r.get_all_hits()[0].shape == (0,) and r.get_all_hits()[1].shape == (3, 0)
    with pytest.raises(AttributeError):
        oc.NoSuchOpticsReceiver
    with pytest.raises(NotImplementedError):
        oc.RefractiveScatteringHomogenous
    rh = oc.RefractiveHomogenous(1., 1.5, single_ray=False, sigma=1e-3)
    assert rh._ref_idxs == (1., 1.5) and rh._single_ray is False and rh._sigma == 1e-3
    rt = oc.RefractiveTransmissiveHomogenous(1., 1.5, [0.1, 2.])
    assert isinstance(rt, oc.RefractiveHomogenous) and list(rt.a_c) == [0.1, 2.] and rt._scaling == 1.
    with pytest.raises(TypeError):
        r(None, None, None)            # descriptions are not callable on the host


def test_unknown_classes_and_hooks_raise_at_compile_time():
    from tracer_b200.scene import CompiledScene

    class Weird(T.geometry_manager.GeometryManager):
        pass

    def asm_with(gm, opt):
        return T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[T.surface.Surface(gm, opt)])])
    with pytest.raises(NotImplementedError, match='Weird'):
        CompiledScene(asm_with(Weird(), T.optics_callables.Reflective(0.)), upload=False)

    class WeirdOptics(object):
        pass
    with pytest.raises(NotImplementedError, match='WeirdOptics'):
        CompiledScene(asm_with(T.flat_surface.RectPlateGM(1, 1), WeirdOptics()), upload=False)

    class TiltedMirror(T.optics_callables.Reflective):          # a subclass that changes the optics ...
        def __call__(self, geometry, rays, selector):
            return None
    with pytest.raises(NotImplementedError, match='TiltedMirror'):  # ... must not be traced as its parent
        CompiledScene(asm_with(T.flat_surface.RectPlateGM(1, 1), TiltedMirror(0.1)), upload=False)

    class RenamedMirror(T.optics_callables.Reflective):         # same __call__: fine
        pass
    sc = CompiledScene(asm_with(T.flat_surface.RectPlateGM(1, 1), RenamedMirror(0.1)), upload=False)
    assert sc.records[0]['okind'] == 1
    with pytest.raises(NotImplementedError):
        T.optics_callables.Reflective_IAM
    with pytest.raises(NotImplementedError, match='one-element'):
        CompiledScene(asm_with(T.flat_surface.RectPlateGM(1, 1),
                               T.optics_callables.LambertianAbsorbant(0.1, [0.1, 0.2])), upload=False)
    sc = CompiledScene(asm_with(T.flat_surface.RectPlateGM(1, 1),
                                T.optics_callables.RealReflective_IAM(0.1, 0.16, 1e-3)), upload=False)
    assert sc.records[0]['flags'] & 4 and sc.records[0]['op'][4] == 0.16 and sc.records[0]['op'][5] == 1.

    class Owner(T.object.AssembledObject):
        def own_rays(self, rays, surface_id):
            return np.ones(rays.get_num_rays(), dtype=bool)
    a = T.assembly.Assembly(objects=[Owner(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(1, 1),
                                                                    T.optics_callables.Reflective(0.))])])
    with pytest.raises(NotImplementedError, match='own_rays'):
        CompiledScene(a, upload=False)
    shared = T.optics_callables.ReflectiveReceiver(0.)
    objs = [T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(1, 1), shared)])
            for _ in range(2)]
    with pytest.raises(NotImplementedError, match='shared'):
        CompiledScene(T.assembly.Assembly(objects=objs), upload=False)


def test_constructor_validation_matches_reference_errors():
    with pytest.raises(ValueError):
        T.flat_surface.RectPlateGM(-1, 1)
    with pytest.raises(ValueError):
        T.flat_surface.RoundPlateGM(1., 2.)
    with pytest.raises(ValueError):
        T.sphere_surface.SphericalGM(0)
    with pytest.raises(AttributeError):
        T.cone.FiniteCone(0, 1)
    with pytest.raises(AttributeError):
        T.cone.ConicalFrustum(0, 1, 0, 2)
    with pytest.raises(ValueError):
        T.surface.Surface(T.flat_surface.RectPlateGM(1, 1), None, location=np.zeros(2))


def test_ray_bundle_api():
    RB = T.ray_bundle.RayBundle
    v = np.arange(12.).reshape(3, 4)
    b = RB(vertices=v, directions=v + 1, energy=np.arange(4.), wavelengths=np.arange(4.) * 1e-6)
    assert b.get_num_rays() == 4 and b.has_property('wavelengths') and b.has_property('parents')
    with pytest.raises(KeyError):
        b.get_parents()
    c = b.inherit(np.array([2, 0]), energy=np.array([9., 8.]), parents=np.array([2, 0]))
    assert np.array_equal(c.get_vertices(), v[:, [2, 0]]) and np.array_equal(c.get_energy(), [9., 8.])
    assert np.array_equal(c.get_wavelengths(), [2e-6, 0.])
    d = b.delete_rays(np.array([1]))
    assert d.get_num_rays() == 3 and np.array_equal(d.get_energy(), [0., 2., 3.])
    s = b + b
    assert s.get_num_rays() == 8
    cat = T.ray_bundle.concatenate_rays([b, d])
    assert cat.get_num_rays() == 7
    assert T.ray_bundle.concatenate_rays([]).get_num_rays() == 0
    b.set_energy(5., selector=np.array([0]))
    assert b.get_energy()[0] == 5.
    assert b.get_energy(selector=np.array([1, 2])).shape == (2,)


def test_ray_tree_history():
    RT, RB = T.trace_tree.RayTree, T.ray_bundle.RayBundle
    t = RT()
    z = np.zeros((3, 4))
    t.append(RB(vertices=z, directions=z, energy=np.ones(4)))
    t.append(RB(vertices=z, directions=z, energy=np.ones(4), parents=np.array([1, 0, 3, 2])))
    t.append(RB(vertices=z[:, :2], directions=z[:, :2], energy=np.ones(2), parents=np.array([2, 1])))
    assert t.num_bunds() == 3
    assert [list(p) for p in t.ordered_parents()] == [[1, 0, 3, 2], [2, 1]]
    assert list(t.ray_history(0)) == [0, 2, 3]


def test_triangle_mesh_object_compiles_like_one_surface_per_face():
    """stl_utils.TriangleMeshObject (array route) gives the scene compiler the records that one
    Surface per face gives it, under a tilted assembly transform and next to an ordinary
    object; its boxes are the exact face boxes (inside the per-surface ones); surfaces made on
    demand carry the frames of the per-face objects."""
    from tracer_b200.scene import CompiledScene
    from tracer_b200.spatial_geometry import rotx, roty
    scn = []
    tilt = np.dot(rotx(0.3), roty(-0.4))
    tilt[:3, 3] = [0.5, -1., 2.]
    for per_face in (False, True):
        asm, _ = scenes.c4_cavity_mesh(T, n_theta=12, n_z=5, per_face_objects=per_face)
        plate = T.object.AssembledObject(surfs=[T.surface.Surface(T.flat_surface.RectPlateGM(1., 1.),
                                                                  T.optics_callables.ReflectiveReceiver(0.2))],
                                         transform=tilt)
        asm.add_object(plate)
        asm.set_transform(tilt)
        scn.append((asm, CompiledScene(asm, upload=False)))
    (asm_m, m), (asm_o, o) = scn
    assert len(m.records) == len(o.records) == 12 * 5 * 2 + 12 + 1
    assert m.mesh_segments[0][:2] == (0, 132) and o.mesh_segments == []
    for f in ('gkind', 'okind', 'flags', 'table_off', 'table_len', 'op'):
        assert np.array_equal(m.records[f], o.records[f]), f
    for f in ('frame', 'inv', 'gp'):
        np.testing.assert_allclose(m.records[f], o.records[f], rtol=0, atol=1e-14)
    assert np.all(m.aabb_lo >= o.aabb_lo - 1e-12) and np.all(m.aabb_hi <= o.aabb_hi + 1e-12)
    assert m.account_list == [132] and o.account_list == list(range(133))
    surfs_m, surfs_o = asm_m.get_surfaces(), asm_o.get_surfaces()
    assert len(surfs_m) == 133
    for i in (0, 57, 131, 132, -1):
        np.testing.assert_allclose(surfs_m[i]._temp_frame, surfs_o[i]._temp_frame, atol=1e-14)
        assert type(surfs_m[i].get_geometry_manager()).__name__ == type(surfs_o[i].get_geometry_manager()).__name__
    with pytest.raises(ValueError):
        T.stl_utils.make_stl_tracer_object(np.zeros((2, 3, 3)), T.optics_callables.LambertianReceiver, {})


def test_load_stl_binary_and_ascii(tmp_path):
    tris = scenes.cylinder_mesh_array(5, 2).astype('<f4')
    rec = np.zeros(len(tris), dtype=np.dtype([('n', '<f4', 3), ('v', '<f4', (3, 3)), ('a', '<u2')]))
    rec['v'] = tris
    p = tmp_path / 'm.stl'
    p.write_bytes(b' ' * 80 + np.uint32(len(tris)).tobytes() + rec.tobytes())
    assert np.array_equal(T.stl_utils.load_stl(str(p)), tris.astype(float))
    q = tmp_path / 'a.stl'
    q.write_text('solid x\n' + ''.join(
        'facet normal 0 0 0\n outer loop\n' + ''.join('  vertex %r %r %r\n' % tuple(float(c) for c in v) for v in t) +
        ' endloop\nendfacet\n' for t in tris) + 'endsolid x\n')
    assert np.array_equal(T.stl_utils.load_stl(str(q)), tris.astype(float))


def test_refractive_records():
    """Scene compiler, row O8: op = n_a, n_b, single_ray, sigma (-1 = None), n_att, a_c0, a_c1, scaling;
    two child slots per hit; accountants and complex indices are refused."""
    from tracer_b200 import _lib as L
    from tracer_b200.scene import CompiledScene
    oc = T.optics_callables
    flat = T.flat_surface.RectPlateGM

    def scene(*optics):
        objs = [T.object.AssembledObject(surfs=[T.surface.Surface(flat(1., 1.), o, location=np.r_[0., 0., -float(k)])])
                for k, o in enumerate(optics)]
        return CompiledScene(T.assembly.Assembly(objects=objs), upload=False)
    sc = scene(oc.RefractiveHomogenous(1., 1.5), oc.RefractiveHomogenous(1.33, 1., single_ray=False, sigma=2e-3),
               oc.RefractiveTransmissiveHomogenous(1., 1.45, [0.05, 2.5], scaling=10.), oc.Reflective(0.1))
    r = sc.records
    assert list(r['okind']) == [L.O_REFRACTIVE] * 3 + [L.O_REFLECTIVE]
    assert list(r['op'][0]) == [1., 1.5, 1., -1., 0., 0., 0., 0.]
    assert list(r['op'][1]) == [1.33, 1., 0., 2e-3, 0., 0., 0., 0.]
    assert list(r['op'][2]) == [1., 1.45, 1., -1., 2., 0.05, 2.5, 10.]
    assert list(sc.children_per_hit) == [2, 2, 2, 1] and sc.has_split and sc.has_refractive
    assert not scene(oc.Reflective(0.1)).has_refractive
    assert scene(oc.RefractiveHomogenousReceiver(1., 1.5)).has_accountants       # one child per hit: count() is defined
    with pytest.raises(NotImplementedError):                                     # two: the reference raises ValueError
        scene(oc.RefractiveHomogenousReceiver(1., 1.5, single_ray=False))
    with pytest.raises(NotImplementedError):
        scene(oc.RefractiveHomogenous(1., 1.5 + 0.1j))
    # accountants of the incoming rays (wavelengths, normals): fine on a smooth surface; on a rough one the reference
    # perturbs the geometry manager's normals in place before NormalAccountant reads them -- refused
    smooth = scene(oc.RefractiveHomogenousSpectralNormal(1., 1.5))
    assert smooth.has_accountants and {'wavelengths', 'normals'} <= set(smooth.accountant_sources)
    scene(oc.RefractiveHomogenousSpectral(1., 1.5, sigma=1e-3))
    with pytest.raises(NotImplementedError):
        scene(oc.RefractiveHomogenousNormal(1., 1.5, sigma=1e-3))
    with pytest.raises(AttributeError):                     # SpectralAccountant.get_data, as the reference's (:1813-1823)
        oc.RefractiveHomogenousSpectralNormal(1., 1.5).get_all_hits()
    # media given by material tables: knots of both materials in the pool, flag TRC_F_TABULATED
    ocn = T.optical_constants
    glass = ocn.OpticalMaterialTable([3e-7, 1e-6, 2.5e-6], [1.55, 1.52, 1.5], [1e-8, 2e-8, 4e-8])
    tab = scene(oc.RefractiveAbsorbant(ocn.Air(), glass, single_ray=False, attenuation_coefficient_1=1., scaling=2.))
    r0 = tab.records[0]
    assert r0['okind'] == L.O_REFRACTIVE and r0['flags'] & L.F_TABULATED and tab.has_tabulated
    assert r0['table_len'] == 2 and int(r0['op'][1]) == 3 and r0['op'][4] == 1. and r0['op'][7] == 2.
    off_b = int(r0['op'][0])
    assert list(tab.tables[off_b:off_b + 9]) == [3e-7, 1e-6, 2.5e-6, 1.55, 1.52, 1.5, 1e-8, 2e-8, 4e-8]
    with pytest.raises(NotImplementedError):                                     # the reference raises AttributeError
        scene(oc.RefractiveAbsorbant(ocn.Air(), glass))
    with pytest.raises(NotImplementedError):
        scene(oc.RefractiveTransmissiveHomogenous(1., 1.5, None))


def test_kdtree_handle():
    """accel_tree.KdTree keeps the reference's constructor and root-box test; the tree is the BVH the
    nearest-hit kernel walks."""
    asm, _ = scenes.c3_nsttf(T, n_rays=10, n_hstats=40)
    kd = T.accel_tree.KdTree(asm, min_leaf=1, t_isec=500.)
    assert kd.n_surfs == 41 and len(kd.nodes) == 2 * 41 - 1 and kd.always_relevant == []
    assert sorted(kd.leaf_surfaces.tolist()) == list(range(41)) and 6 <= kd.depth() <= 12
    root = kd.nodes[0]
    assert np.allclose(root['lo'], kd.minpoint) and np.allclose(root['hi'], kd.maxpoint)
    poss = np.c_[[0., 40., 100.], [0., 40., 100.], [500., 40., 30.], [0., 40., 30.]]
    dirs = np.c_[[0., 0., -1.], [0., 0., 1.], [0., 1., 0.], [1., 0., 0.]]
    hit, t0, t1 = kd.intersect_bounds(poss, dirs)
    assert hit.tolist() == [True, False, False, True]
    assert t0[0] == pytest.approx(100. - kd.maxpoint[2]) and t0[3] == 0. and t1[3] == pytest.approx(kd.maxpoint[0])
    with pytest.raises(NotImplementedError):
        kd.traversal(None)
    flat = T.assembly.Assembly(objects=[T.object.AssembledObject(
        surfs=[T.surface.Surface(T.flat_surface.FlatGeometryManager(), T.optics_callables.Reflective(0.))])])
    with pytest.raises(Exception, match='No boundary defined'):
        T.accel_tree.KdTree(flat)


HAVE_REF = os.path.isdir('/root/reference/tracer')


@pytest.mark.skipif(not HAVE_REF, reason='reference tree only exists in the build container')
@pytest.mark.parametrize('name', ['c1_flat', 'c2_dish_bivar', 'c3_nsttf_dense', 'c4_cavity', 'c5_cavity_spectral',
                                  'c6_optics_zoo', 'c7_periodic_cell', 'c8_fresnel_conductor', 'c9_refractive'])
def test_reference_built_scene_compiles_to_the_same_tables(name):
    """The drop-in claim: an Assembly built with the REFERENCE's own classes (tracer.*) goes through the
    scene compiler unchanged and yields the device tables of the same scene built with this package."""
    from oracle import ref_import
    from tests.golden.make_golden import TRACE_CASES
    from tracer_b200.scene import CompiledScene
    ref_import.enable()
    builder, kw, seed, _, _ = TRACE_CASES[name]
    small = dict(kw)
    small['n_rays'] = 10
    ours, _ = builder(T, **small)
    theirs, _ = builder(scenes.NS('tracer'), **small)
    a, b = CompiledScene(ours, upload=False), CompiledScene(theirs, upload=False)
    assert a.num_surfaces == b.num_surfaces
    for field in a.records.dtype.names:
        assert np.array_equal(a.records[field], b.records[field]), field
    assert np.array_equal(a.tables, b.tables)
    assert np.array_equal(a.children_per_hit, b.children_per_hit)
    assert a.accountant_sources == b.accountant_sources
    assert np.array_equal(a.bvh_leaf_surfaces, b.bvh_leaf_surfaces)


def test_ray_bundle_known_answers_of_the_reference():
    """tests/test_ray_bundle.py:12-99 of the reference: inheritance (empty, full, positional overrides,
    partial), delete_rays, extra properties through inherit and concatenate_rays."""
    RB = T.ray_bundle
    RB.RayBundle().inherit()
    pos, d, energy, ref_ind, prn = np.ones((3, 4)), np.zeros((3, 4)), np.ones(4), np.zeros(4), np.ones(4)
    father = RB.RayBundle(pos, d, energy, prn, ref_ind)
    child = father.inherit()
    for got, want in ((child.get_vertices(), pos), (child.get_directions(), d), (child.get_energy(), energy),
                      (child.get_ref_index(), ref_ind), (child.get_parents(), prn)):
        np.testing.assert_array_equal(got, want)
    child = father.inherit(np.s_[:], d, pos, prn, energy, np.ones(4) * 5)
    np.testing.assert_array_equal(child.get_vertices(), d)
    np.testing.assert_array_equal(child.get_directions(), pos)
    np.testing.assert_array_equal(child.get_energy(), prn)
    np.testing.assert_array_equal(child.get_ref_index(), np.ones(4) * 5)
    np.testing.assert_array_equal(child.get_parents(), energy)
    child = RB.RayBundle(pos, d).inherit(selector=np.r_[0, 1], direction=np.ones((3, 2)))
    np.testing.assert_array_equal(child.get_vertices(), pos[:, [0, 1]])
    np.testing.assert_array_equal(child.get_directions(), np.ones((3, 2)))
    child = RB.RayBundle(pos, d, parents=np.arange(4)).delete_rays(np.r_[0])
    np.testing.assert_array_equal(child.get_vertices(), np.ones((3, 3)))
    np.testing.assert_array_equal(child.get_parents(), np.arange(1, 4))
    child = RB.RayBundle(pos, d, wavelength=np.arange(4)).inherit()
    np.testing.assert_array_equal(child.get_wavelength(), np.arange(4))
    con = RB.concatenate_rays((RB.RayBundle(np.ones((3, 4)), np.ones((3, 4)), wavelen=np.ones((3, 4))),
                               RB.RayBundle(np.zeros((3, 4)), np.zeros((3, 4)), wavelen=np.zeros((3, 4)))))
    correct = np.hstack((np.ones((3, 4)), np.zeros((3, 4))))
    np.testing.assert_array_equal(con.get_vertices(), correct)
    np.testing.assert_array_equal(con.get_wavelen(), correct)


@pytest.mark.skipif(not HAVE_REF, reason='reference tree only exists in the build container')
def test_spatial_geometry_equals_live_reference():
    """Rotations and frames array-equal to tracer/spatial_geometry.py, incl. the 14-decimal
    rounding of general_axis_rotation and the degenerate rows of rotation_to_z."""
    from oracle import ref_import
    ref_import.enable()
    from tracer import spatial_geometry as ref
    sg = T.spatial_geometry
    rs = np.random.RandomState(6)
    for _ in range(20):
        axis = rs.normal(size=3)
        axis /= np.linalg.norm(axis)
        ang = rs.uniform(-4., 4.)
        assert np.array_equal(sg.general_axis_rotation(axis, ang), ref.general_axis_rotation(axis, ang))
        tr = rs.normal(size=(3, 1))
        assert np.array_equal(sg.generate_transform(axis, ang, tr), ref.generate_transform(axis, ang, tr))
        for name in ('rotx', 'roty', 'rotz'):
            assert np.array_equal(getattr(sg, name)(ang), getattr(ref, name)(ang)), name
        assert np.array_equal(sg.translate(*tr[:, 0]), ref.translate(*tr[:, 0]))
    vecs = rs.normal(size=(50, 3))
    vecs /= np.sqrt(np.sum(vecs ** 2, axis=1))[:, None]
    vecs[3] = [0., 0., 1.]
    vecs[7] = [0., 0., -1.]
    assert np.array_equal(sg.rotation_to_z(vecs), ref.rotation_to_z(vecs))
    assert np.array_equal(sg.rotation_to_z(vecs[5]), ref.rotation_to_z(vecs[5]))


def test_boundary_sphere_known_answer():
    """tests/test_boundary_surface.py:10-19 of the reference (the bound CutSphereGM uses)."""
    pts = np.array([[0., 0., 0.], [1., 1., 1.], [2., 2., 2.]])
    np.testing.assert_array_equal(T.boundary_shape.BoundarySphere(radius=2.).in_bounds(pts), [True, True, False])
