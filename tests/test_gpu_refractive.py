"""GPU: refractive optics -- RefractiveHomogenous and
RefractiveTransmissiveHomogenous in the optics kernel.

Known-answer vectors are the reference's own tests (tests/test_opt_callable.py:63-111,
tests/test_tracer_tree.py:65-97, tests/test_tracer_engine.py:236-283); the golden trace
c9_refractive (recorded from the reference, tests/golden/make_golden.py) is covered by
tests/test_gpu_parity.py.  Tolerances: indices and ref_index bit-exact, floats 1e-9.
"""
import math

import numpy as np
import pytest

from tests import scenes
from tests.helpers import close

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')
TIGHT = 1e-9


def _engine(asm, **kw):
    from tracer_b200.tracer_engine import TracerEngine
    return TracerEngine(asm, **kw)


def _flat_interface(**kw):
    oc = T.optics_callables
    surf = T.surface.Surface(T.flat_surface.FlatGeometryManager(), oc.RefractiveHomogenous(1., 1.5, **kw))
    return T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[surf])])


def test_all_refracted_split():
    """tests/test_opt_callable.py:64-89 of the reference (both children of every ray)."""
    d = np.c_[[1, 1, -1], [-1, 1, -1], [-1, -1, -1], [1, -1, -1]] / np.sqrt(3)
    pos = np.c_[[0, 0, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]].astype(float)
    en = np.r_[100., 200., 300., 400.]
    bund = T.ray_bundle.RayBundle(pos, d, energy=en, ref_index=np.ones(4))
    eng = _engine(_flat_interface(single_ray=False))
    eng.ray_tracer(bund, reps=1, min_energy=0., tree=True)
    out = eng.tree[1]
    assert out.get_num_rays() == 8
    correct = np.zeros((3, 4))
    correct[:2, 0] = 1
    close(out.get_vertices(), np.hstack((correct, correct)), TIGHT, 'hit points')
    norm = np.c_[[0., 0., 1.]]
    refl_cos = -(d * norm).sum(axis=0)
    refr_cos = -np.sqrt(1 - (1. / 1.5) ** 2 * (1 - refl_cos ** 2))
    close((out.get_directions() * norm).sum(axis=0), np.r_[refl_cos, refr_cos], TIGHT, 'cosines')
    close(out.get_energy().reshape(2, -1).sum(axis=0), en, TIGHT, 'reflection + refraction = 100 %')
    assert np.array_equal(out.get_parents(), np.tile(np.arange(4), 2))
    assert np.array_equal(out.get_ref_index(), np.r_[np.ones(4), 1.5 * np.ones(4)])


@pytest.mark.parametrize('single_ray', [True, False])
def test_total_internal_reflection(single_ray):
    """tests/test_opt_callable.py:91-111: a grazing ray inside the glass stays inside."""
    d = np.c_[[0, math.cos(math.pi / 180), -math.sin(math.pi / 180)]]
    bund = T.ray_bundle.RayBundle(np.c_[[0., 0., 1.]], d, energy=np.r_[100.], ref_index=np.r_[1.5])
    eng = _engine(_flat_interface(single_ray=single_ray))
    eng.ray_tracer(bund, reps=1, min_energy=0., tree=True)
    out = eng.tree[1]
    assert out.get_vertices().shape == (3, 1)
    close(out.get_directions(), np.c_[[0, math.cos(math.pi / 180), math.sin(math.pi / 180)]], TIGHT, 'direction')
    close(out.get_energy(), np.r_[100.], TIGHT, 'energy')
    assert np.array_equal(out.get_parents(), np.r_[0])
    assert np.array_equal(out.get_ref_index(), np.r_[1.5])


def test_tree_through_a_slab():
    """The scene of tests/test_tracer_tree.py:65-97 -- two refractive planes under a mirror (there a
    CutSphereGM; the reference's intersect_ray raises on that scene under numpy 2, so the known answer
    is the oracle's, itself pinned by the c9 golden trace) -- with every ray split, three generations."""
    from oracle.numpy_trace import OracleEngine
    oc = T.optics_callables
    s1 = T.surface.Surface(T.flat_surface.FlatGeometryManager(), oc.RefractiveHomogenous(1., 1.5, single_ray=False),
                           location=np.array([0, 0, -1.]))
    s2 = T.surface.Surface(T.flat_surface.FlatGeometryManager(), oc.RefractiveHomogenous(1., 1.5, single_ray=False),
                           location=np.array([0, 0, 1.]))
    o1 = T.object.AssembledObject(surfs=[s1, s2])
    s3 = T.surface.Surface(T.flat_surface.RectPlateGM(3., 3.), oc.perfect_mirror)
    o2 = T.object.AssembledObject(surfs=[s3], transform=T.spatial_geometry.translate(0., 0., 4.))
    asm = T.assembly.Assembly(objects=[o1, o2])
    x = 1. / math.sqrt(2)
    d = np.c_[[0, 0.6, -0.8], [0, x, x], [0, 0, 1.]]
    pos = np.c_[[0, 0, 2.], [0, 0, 2.], [0, 0., 2.]]

    def bundle():
        return T.ray_bundle.RayBundle(pos.copy(), d.copy(), ref_index=np.ones(3), energy=np.ones(3))
    eng = _engine(asm)
    eng.ray_tracer(bundle(), 3, .05)
    ora = OracleEngine(asm)
    ora.ray_tracer(bundle(), 3, .05)
    assert eng.tree.num_bunds() == len(ora.tree) == 4
    assert [eng.tree[k].get_num_rays() for k in range(4)] == [t.n for t in ora.tree] == [3, 3, 4, 2]
    assert np.array_equal(ora.tree[2].get_parents(), np.r_[0, 1, 0, 1])     # [reflected][refracted]
    for k in range(1, 4):
        mine, ref = eng.tree[k], ora.tree[k]
        assert np.array_equal(mine.get_parents(), ref.get_parents()), f'level {k} parents'
        assert np.array_equal(mine.get_ref_index(), ref.get_ref_index()), f'level {k} ref_index'
        close(mine.get_vertices(), ref.get_vertices(), TIGHT, f'level {k} vertices')
        close(mine.get_directions(), ref.get_directions(), TIGHT, f'level {k} directions')
        close(mine.get_energy(), ref.get_energy(), TIGHT, f'level {k} energy')


def test_paraxial_ray_through_a_lens():
    """tests/test_tracer_engine.py:236-283: glass screen + bi-convex lens (f = 1) in nested
    assemblies, mirror at the focus: a paraxial ray returns on the other side of the axis."""
    oc = T.optics_callables
    sg = T.spatial_geometry
    R = 1. / 6.

    def refr():
        return oc.RefractiveHomogenous(1., 1.5, single_ray=False)
    back = T.surface.Surface(T.sphere_surface.HemisphereGM(R), refr(), location=np.r_[0., 0., -R / 2.])
    front = T.surface.Surface(T.sphere_surface.HemisphereGM(R), refr(), location=np.r_[0., 0., R / 2.],
                              rotation=sg.rotx(np.pi / 2.)[:3, :3])
    lens = T.object.AssembledObject(surfs=[back, front])
    back = T.surface.Surface(T.flat_surface.RoundPlateGM(R), refr(), location=np.r_[0., 0., -0.01])
    front = T.surface.Surface(T.flat_surface.RoundPlateGM(R), refr(), location=np.r_[0., 0., 0.01])
    screen = T.object.AssembledObject(surfs=[back, front], transform=sg.translate(0., 0., 0.5))
    lens_asm = T.assembly.Assembly(objects=[screen, lens])
    lens_asm.set_transform(sg.translate(0., 0., 1.))
    mirror = T.models.one_sided_mirror.rect_one_sided_mirror(1., 1., 0.)
    asm = T.assembly.Assembly(objects=[mirror], subassemblies=[lens_asm])
    bund = T.ray_bundle.RayBundle()
    bund.set_vertices(np.c_[[0.01, 0., 2.]])
    bund.set_directions(np.c_[[0., 0., -1.]])
    bund.set_energy(np.r_[100.])
    bund.set_ref_index(np.r_[1.])
    eng = _engine(asm)
    eng.ray_tracer(bund, 15, 10.)
    last = eng.tree[-1]
    keep = last.get_energy() > 10.
    assert keep.sum() == 1
    np.testing.assert_array_almost_equal(last.get_vertices()[:, keep], np.c_[[-0.01, 0., 1.5]], 2)
    np.testing.assert_array_almost_equal(last.get_directions()[:, keep], np.c_[[0., 0., 1.]], 2)


def test_philox_refraction_statistics():
    """Production mode (device draws): a beam at 45 degrees onto glass in single_ray mode reflects the
    Fresnel fraction of its rays (binomial 5-sigma bound), energy is conserved exactly ray by ray,
    and the split mode carries the same expected reflected energy."""
    n = 400000
    d = np.tile(np.c_[[math.sin(math.pi / 4), 0., -math.cos(math.pi / 4)]], (1, n))
    pos = np.vstack([np.random.RandomState(3).uniform(-1, 1, size=(2, n)), np.ones(n)])
    en = np.ones(n)
    th1 = math.pi / 4
    th2 = math.asin(math.sin(th1) / 1.5)
    rs = ((math.cos(th1) - 1.5 * math.cos(th2)) / (math.cos(th1) + 1.5 * math.cos(th2))) ** 2
    rp = ((math.cos(th2) - 1.5 * math.cos(th1)) / (math.cos(th2) + 1.5 * math.cos(th1))) ** 2
    R = 0.5 * (rs + rp)
    bund = T.ray_bundle.RayBundle(pos, d, energy=en, ref_index=np.ones(n))
    eng = _engine(_flat_interface(single_ray=True), seed=5)
    eng.ray_tracer(bund, reps=1, min_energy=0., tree=True)
    out = eng.tree[1]
    assert out.get_num_rays() == n
    ri = out.get_ref_index()
    n_refl = int((ri == 1.).sum())
    assert n_refl + int((ri == 1.5).sum()) == n
    assert abs(n_refl - R * n) < 5. * math.sqrt(n * R * (1. - R))
    assert np.array_equal(ri[:n_refl], np.ones(n_refl)), 'reflected block first'
    up = out.get_directions()[2] > 0
    assert np.array_equal(up, ri == 1.)
    close(out.get_energy(), np.ones(n), TIGHT, 'single_ray keeps the whole energy on one child')
    assert np.array_equal(np.sort(out.get_parents()), np.arange(n))
    # refracted directions follow Snell's law
    sin_out = np.sqrt(1. - out.get_directions()[2, ~up] ** 2)
    close(sin_out, np.full(int((~up).sum()), math.sin(th1) / 1.5), TIGHT, 'Snell')

    bund = T.ray_bundle.RayBundle(pos, d, energy=en, ref_index=np.ones(n))
    eng2 = _engine(_flat_interface(single_ray=False), seed=5)
    eng2.ray_tracer(bund, reps=1, min_energy=0., tree=True)
    out2 = eng2.tree[1]
    assert out2.get_num_rays() == 2 * n
    close(out2.get_energy()[:n], np.full(n, R), TIGHT, 'reflected energy')
    close(out2.get_energy()[n:], np.full(n, 1. - R), TIGHT, 'refracted energy')


def test_attenuating_slab_beer_lambert():
    """RefractiveTransmissiveHomogenous: rays crossing a 0.2-thick slab at normal incidence lose
    exp(-a_glass * 0.2) of what enters, on top of the two Fresnel transmittances (split mode)."""
    oc = T.optics_callables
    a_air, a_glass = 0.0, 3.0

    def opt():
        return oc.RefractiveTransmissiveHomogenous(1., 1.5, [a_air, a_glass], single_ray=False)
    top = T.surface.Surface(T.flat_surface.RectPlateGM(2., 2.), opt())
    bottom = T.surface.Surface(T.flat_surface.RectPlateGM(2., 2.), opt(), location=np.r_[0., 0., -0.2])
    catcher = T.surface.Surface(T.flat_surface.RectPlateGM(4., 4.), oc.LambertianReceiver(1.),
                                location=np.r_[0., 0., -1.])
    asm = T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[top, bottom]),
                                       T.object.AssembledObject(surfs=[catcher])])
    n = 1000
    pos = np.vstack([np.random.RandomState(4).uniform(-0.5, 0.5, size=(2, n)), np.ones(n)])
    d = np.tile(np.c_[[0., 0., -1.]], (1, n))
    bund = T.ray_bundle.RayBundle(pos, d, energy=np.ones(n), ref_index=np.ones(n))
    eng = _engine(asm)
    eng.ray_tracer(bund, reps=3, min_energy=1e-3, tree=True)
    absorbed = catcher.get_optics_manager().get_all_hits()[0]
    Rn = ((1. - 1.5) / (1. + 1.5)) ** 2
    # first arrivals: (1 - R) in, attenuated over 0.2 in glass, (1 - R) out, then 0.8 of air at a_air = 0
    expect = (1. - Rn) * math.exp(-a_glass * 0.2) * (1. - Rn)
    first = absorbed[:n]
    close(first, np.full(n, expect), TIGHT, 'transmitted energy')


def test_refractive_needs_ref_index_and_refuses_accountants():
    bund = T.ray_bundle.RayBundle(np.c_[[0., 0., 1.]], np.c_[[0., 0., -1.]], energy=np.r_[1.])
    eng = _engine(_flat_interface())
    with pytest.raises(ValueError):
        eng.ray_tracer(bund, reps=1, min_energy=0.)
    oc = T.optics_callables
    # accountants need one child per hit (single_ray): with two the reference's count() raises ValueError
    surf = T.surface.Surface(T.flat_surface.FlatGeometryManager(), oc.RefractiveHomogenousReceiver(1., 1.5, single_ray=False))
    asm = T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[surf])])
    with pytest.raises(NotImplementedError):
        _engine(asm).compile()


def test_accountants_on_a_single_ray_refractive_surface():
    """The reference's accountants see the child bundle as the callable returns it, [reflected][refracted]
    (optics_callables.py:1638-1643): Location in hit order, Scatterer the children's energies in child order, Absorber
    their difference to the hit-ordered incoming energies -- zero for a source of equal energies, a permutation
    artefact otherwise.  (Bundle-by-bundle parity with the reference: golden c12, surface 1.)"""
    oc = T.optics_callables
    opt = oc.RefractiveHomogenousAbsorberScattererLocation(1., 1.5, single_ray=True)
    surf = T.surface.Surface(T.flat_surface.RectPlateGM(4., 4.), opt)
    asm = T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[surf])])
    n = 5000
    rng = np.random.RandomState(1)
    v = np.vstack([rng.uniform(-1, 1, n), rng.uniform(-1, 1, n), np.ones(n)])
    d = np.tile(np.c_[[0.6, 0., -0.8]], (1, n))
    e = rng.uniform(0.5, 1.5, n)
    bund = T.ray_bundle.RayBundle(vertices=v, directions=d, energy=e, ref_index=np.ones(n))
    eng = _engine(asm, rng='philox', seed=3)
    eng.ray_tracer(bund, reps=1, min_energy=0., tree=True)
    child = eng.tree[1]
    absorbed, scattered, hits = opt.get_all_hits()
    assert child.get_num_rays() == n and hits.shape == (3, n)
    close(hits[:2], v[:2] + 0.75 * np.r_[1., 0.][:, None], 1e-12, 'hit points in hit order')
    close(scattered, child.get_energy(), 0., 'Scatterer = child energies in child order')
    close(absorbed, e - child.get_energy(), 1e-15, 'Absorber = hit-ordered e_in - child-ordered e_out')
    par = child.get_parents()
    assert not np.array_equal(par, np.arange(n))          # reflected rays first: the child order is not the hit order
    close(child.get_energy(), e[par], 0., 'single_ray children keep their parent energy')


def test_tabulated_media_snell_fresnel_and_k_attenuation():
    """Refractive / RefractiveAbsorbant with material tables (optics_callables.py:726-858, 908-944), device draws:
    Snell's law with Re m(lambda), energy split R + T = 1 with the complex-index Fresnel reflectance, the child's
    complex index = the material's at its wavelength, attenuation exp(-4 pi k path / lambda) of the children."""
    import os
    from tracer_b200.tracer_engine import TracerEngine
    oc, ocn, sg = T.optics_callables, T.optical_constants, T.spatial_geometry
    from tracer_b200.models import scenes as ms
    glass = ocn.OpticalMaterialFromFile(os.path.join(os.path.dirname(os.path.abspath(ms.__file__)), 'data', 'glass_nk.csv'))
    air = ocn.Air()
    n = 20000
    rng = np.random.RandomState(4)
    wl = rng.uniform(4e-7, 2.3e-6, n)
    d = np.tile(np.c_[[0.5, 0.2, -1.]] / np.sqrt(1.29), (1, n))
    v = np.vstack([rng.uniform(-0.5, 0.5, n), rng.uniform(-0.5, 0.5, n), np.ones(n)])
    for cls, kw in ((oc.Refractive, {}), (oc.RefractiveAbsorbant, dict(attenuation_coefficient_1=1., scaling=3.))):
        surf = T.surface.Surface(T.flat_surface.RectPlateGM(6., 6.), cls(air, glass, single_ray=False, **kw))
        asm = T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[surf])])
        b = T.ray_bundle.RayBundle(vertices=v, directions=d, energy=np.ones(n), wavelengths=wl, ref_index=air.m(wl))
        eng = TracerEngine(asm, rng='philox', seed=2)
        eng.ray_tracer(b, reps=1, min_energy=1e-12, tree=True)
        child = eng.tree[1]
        assert child.get_num_rays() == 2 * n
        par = child.get_parents()
        assert np.array_equal(par[:n], np.arange(n)) and np.array_equal(par[n:], np.arange(n))
        m2 = glass.m(wl)
        ri = child.get_ref_index()
        close(np.real(ri[n:]), m2.real, 1e-12, 'Re m2')
        assert np.allclose(np.imag(ri[n:]), m2.imag, rtol=1e-9, atol=0.)
        assert np.all(ri[:n] == 1.)                                   # reflected rays stay in air
        close(child.get_wavelengths(), np.r_[wl, wl], 1e-15, 'wavelengths')
        # Snell: sin(theta_2) = sin(theta_1) / Re m2
        dr = child.get_directions()[:, n:]
        sin1 = np.sqrt(1. - d[2] ** 2)
        sin2 = np.sqrt(dr[0] ** 2 + dr[1] ** 2)
        close(sin2, sin1 / m2.real, 1e-12, 'Snell')
        # Fresnel with complex indices, real part kept
        th = np.arccos(np.abs(d[2]))
        bar = np.sqrt(1 - (1. / m2 * np.sin(th)) ** 2)
        rs = ((np.cos(th) - m2 * bar) / (np.cos(th) + m2 * bar)) ** 2
        rp = ((bar - m2 * np.cos(th)) / (bar + m2 * np.cos(th))) ** 2
        R = ((rs + rp) / 2).real
        e = child.get_energy()
        if cls is oc.Refractive:
            close(e[:n], R, 1e-12, 'reflected energy')
            close(e[n:], 1. - R, 1e-12, 'refracted energy')
        else:
            path = 3. * np.sqrt(np.sum((child.get_vertices()[:, n:] - v) ** 2, axis=0))
            close(e[:n], R, 1e-12, 'reflected energy (k = 0 in air)')
            close(e[n:], (1. - R) * np.exp(-4. * np.pi * path * m2.imag / wl), 1e-12, 'attenuated refracted energy')
    # without a coefficient the reference raises AttributeError on the first hit: refused at compile time
    surf = T.surface.Surface(T.flat_surface.RectPlateGM(6., 6.), oc.RefractiveAbsorbant(air, glass))
    asm = T.assembly.Assembly(objects=[T.object.AssembledObject(surfs=[surf])])
    with pytest.raises(NotImplementedError):
        TracerEngine(asm).ray_tracer(b, reps=1, min_energy=1e-12)
