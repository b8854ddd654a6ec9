"""GPU: the shadow grid of a coherent bundle (csrc/shadow_grid.cu, stage A of the tree-walk kernel) only prunes:
hit indices through trc_nearest_hit(use_bvh=2) are those of the walk (use_bvh=1) and of the brute-force kernel
(use_bvh=0) bit for bit -- for rays inside the cone, outside it, far away, parallel to the reference plane, and
for a hint that is wrong."""
import ctypes as C

import numpy as np
import pytest

from tests import scenes
from tests.test_gpu_tree_walk import _hits, _random_rays, _triangle_soup

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')


def _nsttf_source(n, seed, csr=0.01):
    pos = scenes.nsttf_positions()
    sun = scenes.sun_vector(0., np.radians(35.05))
    lo, hi = pos.min(axis=0), pos.max(axis=0)
    centre = 0.5 * (lo + hi)
    radius = 1.1 * 0.5 * np.sqrt(np.sum((hi[:2] - lo[:2]) ** 2))
    return T.sources.buie_sunshape_device(n, np.vstack(300. * sun + centre), -sun, radius, csr, flux=1000.,
                                          pre_process_CSR=False, seed=seed), sun


def test_nsttf_grid_equals_walk_equals_brute_1e6():
    from tracer_b200.scene import CompiledScene
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    src, sun = _nsttf_source(1000000, 5)
    V, D = src.get_vertices(), src.get_directions()
    sc = CompiledScene(asm)
    axis, half = src._dir_cone
    assert np.allclose(axis / np.linalg.norm(axis), -sun) and 0.0436 <= half < 0.0437
    assert sc.set_shadow_grid(axis, half)
    brute, walk, grid = _hits(sc, V, D, 0), _hits(sc, V, D, 1), _hits(sc, V, D, 2)
    assert np.array_equal(walk, brute)
    bad = np.nonzero(grid != brute)[0]
    assert bad.size == 0, 'grid != brute at rays %s: %s vs %s' % (bad[:5], grid[bad[:5]], brute[bad[:5]])
    assert 0.04 < (grid >= 0).mean() < 0.09
    # a hint narrower than the bundle: the rays outside the slope box take the walk, same answer
    assert sc.set_shadow_grid(axis, 0.004)
    assert np.array_equal(_hits(sc, V, D, 2), brute)
    # a hint pointing elsewhere: every ray takes the walk
    assert sc.set_shadow_grid(np.r_[0.3, 0.2, -0.9], 0.01)
    assert np.array_equal(_hits(sc, V, D, 2), brute)
    # no grid set: use_bvh = 2 is an ordinary walk
    assert not sc.set_shadow_grid(None, 0.)
    assert np.array_equal(_hits(sc, V, D, 2), brute)
    sc.close()


@pytest.mark.parametrize('axis', [np.r_[0., 0., -1.], np.r_[1., 0.2, 0.1], np.r_[-0.5, -0.6, 0.62]])
def test_soup_with_mixed_rays(axis):
    """Triangle soup + an unbounded plane; a bundle that mixes rays inside the cone (from near and from 100 km
    away), random rays, rays parallel to the grid's reference plane and rays of the opposite sense."""
    from tracer_b200.scene import CompiledScene
    rng = np.random.RandomState(31)
    asm = _triangle_soup(300, rng, with_plane=True)
    axis = axis / np.linalg.norm(axis)
    n = 200000
    V, D = _random_rays(n, rng, 10.)
    # first 120000: inside a 30 mrad cone around `axis`, origins on a disc 40 units upstream
    k = 120000
    perp = np.linalg.svd(axis[None, :])[2][1:]                      # two unit vectors perpendicular to the axis
    th = 0.03 * np.sqrt(rng.uniform(size=k))
    ph = rng.uniform(0., 2. * np.pi, size=k)
    D[:, :k] = (np.cos(th) * axis[:, None] + np.sin(th) * (np.cos(ph) * perp[0][:, None] + np.sin(ph) * perp[1][:, None]))
    r = 14. * np.sqrt(rng.uniform(size=k))
    a = rng.uniform(0., 2. * np.pi, size=k)
    V[:, :k] = -40. * axis[:, None] + r * (np.cos(a) * perp[0][:, None] + np.sin(a) * perp[1][:, None])
    V[:, :1000] -= 1e5 * D[:, :1000]                               # same lines, origins 100 km upstream: beyond rmax
    D[:, k:k + 1000] = -D[:, :1000]                                # opposite sense
    D[:, k + 1000:k + 2000] = perp[0][:, None]                     # perpendicular to the axis
    sc = CompiledScene(asm, leaf_size=2)
    assert sc.set_shadow_grid(axis, 0.03)
    brute, grid = _hits(sc, V, D, 0), _hits(sc, V, D, 2)
    bad = np.nonzero(grid != brute)[0]
    assert bad.size == 0, 'grid != brute at rays %s: %s vs %s' % (bad[:5], grid[bad[:5]], brute[bad[:5]])
    assert (brute[:k] >= 0).sum() > 1000
    sc.close()


def test_engine_uses_the_grid_and_results_do_not_change():
    """Same Philox trace with and without the grid: identical counts, parents and tallies."""
    from tracer_b200.tracer_engine import TracerEngine
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    out = []
    for use in (True, False):
        asm.reset_all_optics()
        src, _ = _nsttf_source(2000000, 9)
        eng = TracerEngine(asm, rng='philox', seed=3, device_accel='bvh')
        eng.shadow_grid = use
        eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=True)
        assert eng.stats['shadow_grid'] == use
        out.append((eng.stats['per_bounce'], eng.hits_per_surface.cpu().numpy(), eng.absorbed_per_surface.cpu().numpy(),
                    eng.tree[1].get_parents().copy(), eng.tree[2].get_vertices().copy()))
    assert out[0][0] == out[1][0]
    assert np.array_equal(out[0][1], out[1][1])
    assert np.array_equal(out[0][3], out[1][3])
    assert np.array_equal(out[0][4], out[1][4])
    assert np.allclose(out[0][2], out[1][2], rtol=1e-12, atol=0.)
