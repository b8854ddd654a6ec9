"""CPU: the NumPy oracle (oracle/) against golden vectors recorded from the real
reference (tests/golden/make_golden.py).  This is what pins the oracle."""
import warnings

import numpy as np
import pytest

from oracle.numpy_engine import find_intersections
from oracle.numpy_trace import OracleEngine
from tests import scenes
from tests.golden.make_golden import TRACE_CASES, ZOO_RAYS, ZOO_SEED
from tests.helpers import (close, compare_tree, golden_accountant_lists, golden_accountants, golden_bifacial,
                           load_golden)

T = scenes.NS('tracer_b200')


@pytest.mark.parametrize('name', sorted(TRACE_CASES))
def test_trace_matches_reference(name):
    builder, kw, seed, reps, min_energy = TRACE_CASES[name]
    gold = load_golden(name)
    np.random.seed(seed)
    asm, source = builder(T, **kw)
    bundle = source()
    eng = OracleEngine(asm)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        eng.ray_tracer(bundle, reps=reps, min_energy=min_energy, tree=True)
    compare_tree(gold, eng.tree, rtol=1e-12)
    for s, arrays in golden_accountants(gold).items():
        mine = eng.get_all_hits(s)
        assert len(mine) == len(arrays)
        for a, b in zip(mine, arrays):
            close(a, b, 1e-12, f'accountant of surface {s}')
    for s, sides in golden_bifacial(gold).items():
        front, back = eng.get_all_hits(s)
        for side, mine in (('front', front), ('back', back)):
            arrays = sides.get(side, [])
            assert len(mine) == len(arrays), f'BiFacial surface {s} {side}'
            for a, b in zip(mine, arrays):
                close(a, b, 1e-12, f'BiFacial surface {s} {side}')
    for s, lists in golden_accountant_lists(gold).items():
        with pytest.raises(AttributeError):      # as in the reference (optics_callables.py:1840-1848)
            eng.get_all_hits(s)
        tall = eng.tallies[s]
        for (j, field), ref in lists.items():
            close(np.hstack(tall[field]), ref, 1e-12, f'accountant {j} list {field} of surface {s}')


@pytest.mark.parametrize('name', sorted(scenes.gm_zoo(T)))
def test_geometry_matches_reference(name, monkeypatch):
    import oracle.numpy_engine as ne
    # the three sphere-family classes have a bundle-order-dependent root-selection
    # bug in the reference; reproduce it here so the masks are still pinned exactly
    monkeypatch.setattr(ne, 'REPRODUCE_SPHERE_SELECT_BUG', True)
    gold = load_golden('gm_zoo')
    make, frame = scenes.gm_zoo(T)[name]
    rays = scenes.zoo_rays(T, ZOO_RAYS, ZOO_SEED)
    surf = T.surface.Surface(make(), T.optics_callables.Reflective(0.))
    surf._temp_frame = frame
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        prm, points, normals = find_intersections(surf, rays.get_vertices(), rays.get_directions())
    ref = gold[name + '_params']
    assert np.array_equal(np.isfinite(prm), np.isfinite(ref)), 'hit/miss pattern differs'
    fin = np.isfinite(ref)
    close(prm[fin], ref[fin], 1e-12, 'params')
    hit = np.nonzero(fin & (ref != 0.))[0]
    if len(hit) and name != 'flat':
        close(points(hit), gold[name + '_points'], 1e-12, 'points')
        close(normals(hit), gold[name + '_normals'], 1e-12, 'normals')


@pytest.mark.parametrize('name', ['hemisphere', 'cut_sphere', 'cut_sphere_plane', 'sphere_rect'])
def test_sphere_family_matches_reference_ray_by_ray(name):
    """The branch the oracle (and the CUDA path) follow for the three sphere-family classes -- the single valid
    root chosen per ray -- is pinned by the REAL reference run one ray at a time, where its root selection is
    well defined (tests/golden/make_sphere_family.py).  The whole-bundle call of the reference crosses roots
    between rays (recorded next to it): with the quirk reproduced the oracle matches that too."""
    import oracle.numpy_engine as ne
    from tests.golden.make_sphere_family import N_RAYS, SEED
    gold = load_golden('sphere_family')
    make, frame = scenes.gm_zoo(T)[name]
    rays = scenes.zoo_rays(T, N_RAYS, SEED)
    surf = T.surface.Surface(make(), T.optics_callables.Reflective(0.))
    surf._temp_frame = frame
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        prm, points, normals = find_intersections(surf, rays.get_vertices(), rays.get_directions())
    ref = gold[name + '_params']
    assert np.array_equal(np.isfinite(prm), np.isfinite(ref)), 'hit/miss pattern differs'
    fin = np.isfinite(ref)
    close(prm[fin], ref[fin], 1e-12, 'params')
    hit = np.nonzero(fin & (ref != 0.))[0]
    close(points(hit), gold[name + '_points'], 1e-12, 'points')
    close(normals(hit), gold[name + '_normals'], 1e-12, 'normals')
    whole = gold[name + '_params_whole_bundle']
    assert (~((ref == whole) | (np.isinf(ref) & np.isinf(whole)))).sum() > 100      # the quirk is not a corner case
    old = ne.REPRODUCE_SPHERE_SELECT_BUG
    ne.REPRODUCE_SPHERE_SELECT_BUG = True
    try:
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            prm_q, _, _ = find_intersections(surf, rays.get_vertices(), rays.get_directions())
    finally:
        ne.REPRODUCE_SPHERE_SELECT_BUG = old
    assert np.array_equal(np.isfinite(prm_q), np.isfinite(whole))
    close(prm_q[np.isfinite(whole)], whole[np.isfinite(whole)], 1e-12, 'whole-bundle params with the quirk')
