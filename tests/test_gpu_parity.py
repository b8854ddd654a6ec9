"""GPU: the CUDA path (through the C ABI, via tracer_b200.TracerEngine) against
(a) golden vectors recorded from the real reference and (b) the NumPy oracle on
the same seeded inputs with injected draws (rng='numpy').

Bar: hit-surface indices, parent indices and bundle sizes bit-exact; positions,
directions, energies within 1e-5 relative (BASELINE.json north_star) -- the fp64
kernels are checked at 1e-9.
"""
import warnings

import numpy as np
import pytest

from tests import scenes
from tests.golden.make_golden import TRACE_CASES, ZOO_RAYS, ZOO_SEED
from tests.helpers import (close, compare_tree, golden_accountant_lists, golden_accountants, golden_bifacial,
                           load_golden)

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')
TIGHT = 1e-9


def _engine(asm, **kw):
    from tracer_b200.tracer_engine import TracerEngine
    return TracerEngine(asm, **kw)


@pytest.mark.parametrize('accel', ['brute', 'bvh'])
@pytest.mark.parametrize('name', sorted(TRACE_CASES))
def test_trace_matches_reference_golden(name, accel):
    builder, kw, seed, reps, min_energy = TRACE_CASES[name]
    gold = load_golden(name)
    np.random.seed(seed)
    asm, source = builder(T, **kw)
    bundle = source()
    eng = _engine(asm, rng='numpy', device_accel=accel)
    eng.ray_tracer(bundle, reps=reps, min_energy=min_energy, tree=True)
    compare_tree(gold, eng.tree, rtol=TIGHT)
    for s, arrays in golden_accountants(gold).items():
        mine = asm.get_surfaces()[s].get_optics_manager().get_all_hits()
        assert len(mine) == len(arrays)
        for a, b in zip(mine, arrays):
            close(a, b, TIGHT, f'accountant of surface {s}')
    for s, sides in golden_bifacial(gold).items():          # BiFacial managers: (front, back) lists
        front, back = asm.get_surfaces()[s].get_optics_manager().get_all_hits()
        for side, mine in (('front', front), ('back', back)):
            arrays = sides.get(side, [])
            assert len(mine) == len(arrays), f'BiFacial surface {s} {side}'
            for a, b in zip(mine, arrays):
                close(a, b, TIGHT, f'BiFacial surface {s} {side}')
    for s, lists in golden_accountant_lists(gold).items():  # managers with a PolychromaticAccountant: raw lists
        opt = asm.get_surfaces()[s].get_optics_manager()
        with pytest.raises(AttributeError):                 # as in the reference (optics_callables.py:1840-1848)
            opt.get_all_hits()
        for (j, field), ref in lists.items():
            close(np.hstack(getattr(opt.accountants[j], field)), ref, TIGHT, f'accountant {j} list {field} of surface {s}')


@pytest.mark.parametrize('name', sorted(scenes.gm_zoo(T)))
def test_geometry_kinds_against_oracle(name):
    """One surface of every GM kind: hit/miss pattern bit-exact, hit points and
    normals within tolerance of the oracle (intended per-ray root choice)."""
    from oracle.numpy_engine import find_intersections
    opt = T.optics_callables.ReflectiveNormalReceiver(0.1)
    asm = scenes.zoo_scene(T, name, opt)
    rays = scenes.zoo_rays(T, ZOO_RAYS, ZOO_SEED)
    surf = asm.get_surfaces()[0]
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        prm, points, normals = find_intersections(surf, rays.get_vertices(), rays.get_directions())
    prm = prm.copy()
    prm[prm == 0.] = np.inf
    hit = np.nonzero(np.isfinite(prm))[0]
    eng = _engine(asm, device_accel='brute')
    eng.ray_tracer(rays, reps=1, min_energy=-1., tree=True)
    child = eng.tree[1]
    assert np.array_equal(child.get_parents(), hit), 'hit/miss pattern differs from the oracle'
    close(child.get_vertices(), points(hit), TIGHT, 'hit points')
    absorbed, normal_acc, hits_acc = None, None, None
    data = opt.get_all_hits()          # order: Absorber, Location, Normal
    close(data[1], points(hit), TIGHT, 'Location accountant')
    close(data[2], normals(hit), TIGHT, 'Normal accountant')
    close(data[0], 0.1 * np.ones(len(hit)), TIGHT, 'Absorber accountant')
    # same scene through the BVH (or the unbounded list) gives the same answer
    eng2 = _engine(asm, device_accel='bvh')
    eng2.ray_tracer(rays, reps=1, min_energy=-1., tree=True)
    assert np.array_equal(eng2.tree[1].get_parents(), hit)


@pytest.mark.parametrize('name', ['hemisphere', 'cut_sphere', 'cut_sphere_plane', 'sphere_rect'])
@pytest.mark.parametrize('accel', ['brute', 'bvh'])
def test_sphere_family_against_reference_ray_by_ray(name, accel):
    """CUDA path against the REAL reference for the sphere-family classes, on the behaviour the reference has
    where its root selection is well defined: one ray at a time (tests/golden/sphere_family.npz)."""
    from tests.golden.make_sphere_family import N_RAYS, SEED
    gold = load_golden('sphere_family')
    opt = T.optics_callables.ReflectiveNormalReceiver(0.1)
    asm = scenes.zoo_scene(T, name, opt)
    rays = scenes.zoo_rays(T, N_RAYS, SEED)
    ref = gold[name + '_params'].copy()
    ref[ref == 0.] = np.inf
    hit = np.nonzero(np.isfinite(ref))[0]
    eng = _engine(asm, device_accel=accel)
    eng.ray_tracer(rays, reps=1, min_energy=-1., tree=True)
    child = eng.tree[1]
    assert np.array_equal(child.get_parents(), hit), 'hit/miss pattern differs from the reference'
    close(child.get_vertices(), gold[name + '_points'], TIGHT, 'hit points')
    close(opt.get_all_hits()[2], gold[name + '_normals'], TIGHT, 'normals')


def test_culling_and_parent_bookkeeping():
    """Survivors first, culled rays at the back, parents index the previous record
    (tests/test_tracer_tree.py:106-153 of the reference)."""
    from oracle.numpy_trace import OracleEngine
    np.random.seed(5)
    asm, source = scenes.c4_cavity(T, n_rays=3000, n_theta=6, n_z=3, absorptivity=0.9)
    bundle = source()
    e0 = bundle.get_energy()[0]
    state = np.random.get_state()
    orc = OracleEngine(asm)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        orc.ray_tracer(bundle, reps=6, min_energy=0.05 * e0)
    np.random.set_state(state)
    eng = _engine(asm, rng='numpy')
    eng.ray_tracer(bundle, reps=6, min_energy=0.05 * e0)
    assert eng.tree.num_bunds() == len(orc.tree)
    for k in range(1, len(orc.tree)):
        assert np.array_equal(eng.tree[k].get_parents(), orc.tree[k].parents)
        close(eng.tree[k].get_vertices(), orc.tree[k].vertices, TIGHT)
        close(eng.tree[k].get_energy(), orc.tree[k].energy, TIGHT)
    # ray_history walks back to the source bundle
    last = eng.tree.num_bunds() - 1
    if eng.tree[last].get_num_rays():
        hist = eng.tree.ray_history(0)
        assert len(hist) == eng.tree.num_bunds()


def test_tree_false_keeps_last_record_only():
    np.random.seed(3)
    asm, source = scenes.c1_flat_mirror(T, n_rays=500)
    eng = _engine(asm)
    v, d = eng.ray_tracer(source(), reps=3, tree=False)
    assert eng.tree.num_bunds() == 1
    assert v.shape[0] == 3 and d.shape == v.shape


def test_empty_and_missing_bundles():
    asm, _ = scenes.c1_flat_mirror(T)
    eng = _engine(asm)
    # rays that miss everything: tree holds only the source, result is empty
    b = T.ray_bundle.RayBundle(vertices=np.c_[[10., 10., 10.]], directions=np.c_[[0., 0., 1.]],
                               energy=np.ones(1))
    v, d = eng.ray_tracer(b, reps=3)
    assert v.shape == (3, 0) and eng.tree.num_bunds() == 1
    # empty bundle
    e = T.ray_bundle.RayBundle(vertices=np.zeros((3, 0)), directions=np.zeros((3, 0)), energy=np.zeros(0))
    v, d = eng.ray_tracer(e, reps=3)
    assert v.shape == (3, 0)


def test_nearest_hit_tie_goes_to_lowest_index():
    """Two coincident plates: strict '<' keeps the first surface (tracer_engine.py:59)."""
    mk = lambda: T.object.AssembledObject(surfs=[T.surface.Surface(  # noqa: E731
        T.flat_surface.RectPlateGM(2., 2.), T.optics_callables.ReflectiveReceiver(0.5))])
    asm = T.assembly.Assembly(objects=[mk(), mk()])
    b = T.ray_bundle.RayBundle(vertices=np.array([[0., 0.3], [0., -0.2], [1., 2.]]),
                               directions=np.array([[0., 0.], [0., 0.], [-1., -1.]]), energy=np.ones(2))
    for accel in ('brute', 'bvh'):
        eng = _engine(asm, device_accel=accel)
        eng.ray_tracer(b, reps=1)
        s0, s1 = [s.get_optics_manager().get_all_hits() for s in asm.get_surfaces()]
        assert s0[1].shape[1] == 2 and s1[1].shape[1] == 0
        asm.reset_all_optics()


def test_unsupported_classes_fail_loudly():
    class Weird(T.geometry_manager.GeometryManager):
        pass
    asm = T.assembly.Assembly(objects=[T.object.AssembledObject(
        surfs=[T.surface.Surface(Weird(), T.optics_callables.Reflective(0.))])])
    eng = _engine(asm)
    b = T.ray_bundle.RayBundle(vertices=np.zeros((3, 1)), directions=np.c_[[0., 0., 1.]], energy=np.ones(1))
    with pytest.raises(NotImplementedError):
        eng.ray_tracer(b, reps=1)


def test_philox_energy_conservation_and_statistics():
    """Independent RNG: closed-ish cavity, absorbed + escaped + remaining = emitted;
    slope-error spread recovered from reflected rays."""
    np.random.seed(1)
    asm, source = scenes.c4_cavity(T, n_rays=200000, n_theta=12, n_z=6, absorptivity=0.7)
    bundle = source()
    e_tot = bundle.get_energy().sum()
    eng = _engine(asm, rng='philox', seed=7)
    eng.ray_tracer(bundle, reps=40, min_energy=1e-12 * e_tot)
    absorbed = sum(s.get_optics_manager().get_all_hits()[0].sum() for s in asm.get_surfaces())
    # escaped = energy of rays that left through the aperture: emitted - absorbed - remaining
    remaining = eng.tree[eng.tree.num_bunds() - 1].get_energy().sum() if eng.tree.num_bunds() > 1 else 0.
    assert absorbed <= e_tot * (1 + 1e-12)
    # Apparent absorptance of a diffuse cylinder cavity L/D=1, alpha=0.7 is ~0.9 (Sparrow): sanity bracket
    assert 0.80 < (absorbed + remaining) / e_tot < 0.97

    # slope error: flat mirror, normal incidence, sigma = 2 mrad bi-variate
    sig = 2e-3
    mirror = T.object.AssembledObject(surfs=[T.surface.Surface(
        T.flat_surface.RectPlateGM(10., 10.), T.optics_callables.RealReflective(0., sig, bi_var=True))])
    asm2 = T.assembly.Assembly(objects=[mirror])
    n = 400000
    b = T.ray_bundle.RayBundle(vertices=np.vstack((np.zeros((2, n)), np.ones(n))),
                               directions=np.tile(np.c_[[0., 0., -1.]], (1, n)), energy=np.ones(n))
    eng2 = _engine(asm2, rng='philox', seed=11)
    eng2.ray_tracer(b, reps=1)
    dirs = eng2.tree[1].get_directions()
    # reflected ray deviates by twice the normal error: std of each transverse component ~ 2 sigma
    sx, sy = dirs[0].std(), dirs[1].std()
    assert abs(sx - 2 * sig) < 0.02 * 2 * sig and abs(sy - 2 * sig) < 0.02 * 2 * sig


@pytest.mark.parametrize('name', ['c3_nsttf', 'c5_cavity_spectral', 'c6_optics_zoo'])
def test_fused_bounce_equals_step_by_step(name):
    """trc_trace (device-driven: every bounce enqueued at once) and the step-by-step entry points driven
    from Python give identical trees, tallies and accountant data (Philox draws)."""
    builder, kw, seed, reps, min_energy = TRACE_CASES[name]
    out = []
    for fused in (True, False):
        np.random.seed(seed)
        asm, source = builder(T, **kw)
        bundle = source()
        eng = _engine(asm, rng='philox', seed=21)
        eng.fused_bounce = fused
        eng.ray_tracer(bundle, reps=reps, min_energy=min_energy, tree=True)
        tree = [(eng.tree[k].get_vertices(), eng.tree[k].get_directions(), eng.tree[k].get_energy(),
                 eng.tree[k].get_parents()) for k in range(1, eng.tree.num_bunds())]
        accs = [s.get_optics_manager().get_all_hits() for s in asm.get_surfaces()
                if hasattr(s.get_optics_manager(), 'get_all_hits')]
        out.append((eng.stats['per_bounce'], tree, accs, eng.hits_per_surface.cpu().numpy(),
                    eng.absorbed_per_surface.cpu().numpy()))
    assert out[0][0] == out[1][0]
    assert len(out[0][1]) == len(out[1][1]) >= 2
    for a, b in zip(out[0][1], out[1][1]):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
    for a, b in zip(out[0][2], out[1][2]):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
    assert np.array_equal(out[0][3], out[1][3])
    np.testing.assert_allclose(out[0][4], out[1][4], rtol=1e-12)


@pytest.mark.parametrize('fused', [True, False])
def test_hit_list_overflow_is_retried(fused):
    """The hit list is sized from the hit fraction of the previous trace; a bundle that hits
    far more than the last one overflows it and the bounce is re-run with the exact size."""
    n = (1 << 21) + 1234
    asm, _ = scenes.c1_flat_mirror(T, n_rays=10)
    centre, down = np.c_[[0., 0., 4.]], np.r_[0., 0., -1.]
    sparse = T.sources.disk_bundle_device(n, centre, down, 12., 4.65e-3, flux=1000., seed=1)    # ~3 % hit the mirror
    dense = T.sources.disk_bundle_device(n, centre, down, 0.8, 4.65e-3, flux=1000., seed=2)     # all hit
    eng = _engine(asm, rng='philox', seed=3)
    eng.fused_bounce = fused
    eng.ray_tracer(sparse, reps=3, min_energy=1e-10, tree=False)
    assert eng.stats['per_bounce'][0][1] < 0.1 * n
    asm.reset_all_optics()
    eng.ray_tracer(dense, reps=3, min_energy=1e-10, tree=True)
    got = (eng.stats['per_bounce'], eng.hits_per_surface.cpu().numpy(), eng.absorbed_per_surface.cpu().numpy(),
           eng.tree[1].get_parents().copy())
    assert got[0][0][1] == n
    asm.reset_all_optics()
    fresh = _engine(asm, rng='philox', seed=3)
    fresh.fused_bounce = fused
    fresh.ray_tracer(dense, reps=3, min_energy=1e-10, tree=True)
    assert got[0] == fresh.stats['per_bounce']
    assert np.array_equal(got[1], fresh.hits_per_surface.cpu().numpy())
    np.testing.assert_allclose(got[2], fresh.absorbed_per_surface.cpu().numpy(), rtol=1e-12)
    assert np.array_equal(got[3], fresh.tree[1].get_parents())
