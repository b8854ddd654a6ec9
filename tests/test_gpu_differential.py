"""GPU: large differentials against the oracle (>= 1e6 random rays per geometry kind, C2 and C3 at
1e6 rays): hit indices, parents and child counts bit-exact, positions / directions / energies within 1e-9
(the contract allows 1e-5).  The golden fixtures pin the oracle to the reference; these runs put ~250x more
rays through every kind than the fixtures hold (edge cases of apertures, root choices and thresholds)."""
import warnings

import numpy as np
import pytest

from tests import scenes
from tests.helpers import close

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')
N = 1000000
TIGHT = 1e-9


@pytest.mark.parametrize('name', sorted(scenes.gm_zoo(T)))
def test_geometry_kind_1e6_rays(name):
    from oracle.numpy_engine import find_intersections
    from tracer_b200.tracer_engine import TracerEngine
    opt = T.optics_callables.ReflectiveNormalReceiver(0.1)
    asm = scenes.zoo_scene(T, name, opt)
    rays = scenes.zoo_rays(T, N, 2024)
    surf = asm.get_surfaces()[0]
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        prm, points, normals = find_intersections(surf, rays.get_vertices(), rays.get_directions())
    prm = prm.copy()
    prm[prm == 0.] = np.inf
    hit = np.nonzero(np.isfinite(prm))[0]
    assert len(hit) > 20000
    for accel in ('brute', 'bvh'):
        asm.reset_all_optics()
        eng = TracerEngine(asm, rng='numpy', device_accel=accel)
        eng.ray_tracer(rays, reps=1, min_energy=-1., tree=True)
        parents = eng.tree[1].get_parents()
        assert np.array_equal(parents, hit), '%s: %d hits differ from the oracle' % (
            accel, len(np.setxor1d(parents, hit)))
    close(eng.tree[1].get_vertices(), points(hit), TIGHT, 'hit points')
    close(opt.get_all_hits()[2], normals(hit), TIGHT, 'normals')


@pytest.mark.parametrize('name', ['c2_dish', 'c3_nsttf'])
def test_config_trace_1e6_rays(name):
    """C2 (dish, slope-error draws injected in the reference's order) and C3 (NSTTF field, tree walk) at 1e6 source
    rays: the whole tree against the oracle."""
    from oracle.numpy_trace import OracleEngine
    from tracer_b200.tracer_engine import TracerEngine
    np.random.seed(31)
    asm, source = (scenes.c2_dish if name == 'c2_dish' else scenes.c3_nsttf)(T, n_rays=N)
    bundle = source()
    state = np.random.get_state()
    orc = OracleEngine(asm)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        orc.ray_tracer(bundle, reps=3, min_energy=1e-10, tree=True)
    np.random.set_state(state)
    eng = TracerEngine(asm, rng='numpy', device_accel='bvh' if name == 'c3_nsttf' else 'auto')
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        eng.ray_tracer(bundle, reps=3, min_energy=1e-10, tree=True)
    assert eng.tree.num_bunds() == len(orc.tree) >= 3
    for k in range(1, len(orc.tree)):
        assert np.array_equal(eng.tree[k].get_parents(), orc.tree[k].parents), 'parents of bundle %d' % k
        close(eng.tree[k].get_vertices(), orc.tree[k].vertices, TIGHT, 'vertices %d' % k)
        close(eng.tree[k].get_directions(), orc.tree[k].directions, TIGHT, 'directions %d' % k)
        close(eng.tree[k].get_energy(), orc.tree[k].energy, TIGHT, 'energy %d' % k)
    recv = len(asm.get_surfaces()) - 1
    mine = asm.get_surfaces()[recv].get_optics_manager().get_all_hits()
    theirs = orc.get_all_hits(recv)
    assert len(mine[0]) == len(theirs[0]) > 1000
    close(mine[0], theirs[0], TIGHT, 'receiver absorbed energies')
    close(mine[1], theirs[1], TIGHT, 'receiver hit points')
