"""GPU: the headline configuration at FULL size (C3 NSTTF field, 1e8 source rays,
BASELINE.json configs[1]) checked through size-independent properties -- the CPU
oracle cannot finish this size, so parity at scale rests on:

  * energy conservation per bounce: E_in of the hit rays = absorbed + E_out;
  * bookkeeping: hit counts per surface add up, parents ascend inside each surface
    segment (the reference's child order), parents index the parent bundle;
  * determinism: the same seed gives the same tallies;
  * linearity: twice the flux gives twice every absorbed energy, same hits;
  * sharding: two half-size shards (the multi-GPU decomposition: global Philox ray
    ids via first_ray/total_rays) give exactly the hit counts of the whole bundle;
  * tree walk = brute force on the per-surface hit counts (2e7 rays).
"""
import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')
N_FULL = 100000000


def _geometry():
    pos = scenes.nsttf_positions()
    sun = scenes.sun_vector(0., np.radians(35.05))
    lo, hi = pos.min(axis=0), pos.max(axis=0)
    centre = 0.5 * (lo + hi)
    radius = 1.1 * 0.5 * np.sqrt(np.sum((hi[:2] - lo[:2]) ** 2))
    return np.vstack(300. * sun + centre), -sun, radius


def _source(n, flux=1000., first_ray=0, total=None, seed=99):
    centre, direction, radius = _geometry()
    return T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=flux, pre_process_CSR=False,
                                          seed=seed, first_ray=first_ray, total_rays=total)


def _trace(src, accel='bvh', seed=7, tree=True):
    from tracer_b200.tracer_engine import TracerEngine
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    eng = TracerEngine(asm, rng='philox', seed=seed, device_accel=accel)
    eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=tree)
    return eng, asm


def test_full_size_conservation_and_bookkeeping():
    import torch
    src = _source(N_FULL)
    e_ray = float(src._props['energy'].pieces[0][0].item())
    eng, asm = _trace(src)
    per_bounce = eng.stats['per_bounce']
    assert per_bounce[0][0] == N_FULL
    hits = eng.hits_per_surface.cpu().numpy()
    assert hits.sum() == sum(h for _, h, _ in per_bounce)
    assert 0.05 < per_bounce[0][1] / float(N_FULL) < 0.08
    absorbed = eng.absorbed_per_surface.cpu().numpy()
    # bounce 1: every hit ray carried e_ray; what is not absorbed leaves with the children
    b1 = eng.tree[1]
    E1 = b1._props['energy'].pieces[0]
    P1 = b1._props['parents'].pieces[0]
    assert int(P1.min().item()) >= 0 and int(P1.max().item()) < N_FULL
    H1 = per_bounce[0][1]
    assert E1.shape[0] == H1
    e_out1 = float(E1.sum().item())
    b2 = eng.tree[2]
    e_out2 = float(b2._props['energy'].pieces[0].sum().item())
    # survivors of bounce 1 are the parents' energy pool of bounce 2 (culled rays stay in the record only)
    n_surv = per_bounce[0][2]
    e_in2 = float(E1[:n_surv][b2._props['parents'].pieces[0]].sum().item())
    total_in = H1 * e_ray + e_in2
    np.testing.assert_allclose(absorbed.sum() + e_out1 + e_out2, total_in, rtol=1e-9)
    # child order (surface-major, parent-minor): parents ascend strictly inside each surface segment,
    # so the record of a bounce where nothing survives (bounce 2: slot order) has fewer descents than surfaces
    P2 = b2._props['parents'].pieces[0]
    assert per_bounce[1][2] == 0
    descents = int((P2[1:] <= P2[:-1]).sum().item())
    assert descents < len(hits)
    # determinism
    eng_b, _ = _trace(src)
    assert np.array_equal(hits, eng_b.hits_per_surface.cpu().numpy())
    np.testing.assert_allclose(absorbed, eng_b.absorbed_per_surface.cpu().numpy(), rtol=1e-12)
    torch.cuda.empty_cache()


def test_linearity_and_sharding_at_2e7():
    n = 20000000
    whole, _ = _trace(_source(n))
    hits = whole.hits_per_surface.cpu().numpy()
    absorbed = whole.absorbed_per_surface.cpu().numpy()
    double, _ = _trace(_source(n, flux=2000.))
    assert np.array_equal(hits, double.hits_per_surface.cpu().numpy())
    np.testing.assert_allclose(double.absorbed_per_surface.cpu().numpy(), 2. * absorbed, rtol=1e-12)
    # two shards = the two ranks of a 2-GPU run.  The source rays are identical by construction (global ray
    # ids) and the optics draws are keyed by (seed, source-ray id, bounce), so every ray of a shard does
    # exactly what it does in the whole bundle: the tallies agree exactly in the counts and to summation
    # order in the energies.
    half = n // 2
    a, _ = _trace(_source(half, first_ray=0, total=n))
    b, _ = _trace(_source(half, first_ray=half, total=n))
    for k in range(len(whole.stats['per_bounce'])):
        assert [x + y for x, y in zip(a.stats['per_bounce'][k], b.stats['per_bounce'][k])] == \
            list(whole.stats['per_bounce'][k])
    hs = a.hits_per_surface.cpu().numpy() + b.hits_per_surface.cpu().numpy()
    assert np.array_equal(hs, hits)
    np.testing.assert_allclose((a.absorbed_per_surface + b.absorbed_per_surface).cpu().numpy(), absorbed,
                               rtol=1e-9, atol=1e-9 * absorbed.max())


def test_walk_equals_brute_on_tallies_at_2e7():
    n = 20000000
    src = _source(n, seed=5)
    walk, _ = _trace(src, accel='bvh')
    brute, _ = _trace(src, accel='brute')
    assert walk.stats['per_bounce'] == brute.stats['per_bounce']
    assert np.array_equal(walk.hits_per_surface.cpu().numpy(), brute.hits_per_surface.cpu().numpy())
    np.testing.assert_allclose(walk.absorbed_per_surface.cpu().numpy(), brute.absorbed_per_surface.cpu().numpy(),
                               rtol=1e-12)


def test_c2_dish_full_size_properties():
    """Config 2 at BASELINE size: 1e7 Buie rays (CSR 0.05) onto the 5 m paraboloid dish with 2 mrad
    bi-variate slope error and the 0.3 m receiver at the focus.  Size-independent checks: energy
    bookkeeping per bounce, the dish's reflectance exactly 1 - 0.06 ray by ray, parents ascending per
    surface, the receiver's shadow on the dish (its area fraction of the source disk) and the
    intercept of the focal spot; generic-kind kernels (quadric + flat) with brute = tree walk."""
    import torch
    from tracer_b200.tracer_engine import TracerEngine
    n = 10000000
    asm, _ = scenes.c2_dish(T, n_rays=10)
    src = T.sources.buie_sunshape_device(n, np.c_[[0., 0., 10.]], np.r_[0., 0., -1.], 2.5, 0.05, flux=1000.,
                                         seed=21)
    e_ray = float(src._props['energy'].pieces[0][0].item())
    np.testing.assert_allclose(e_ray * n, 1000. * np.pi * 2.5 ** 2, rtol=1e-12)
    eng = TracerEngine(asm, rng='philox', seed=3, device_accel='brute')
    eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=True)
    pb = eng.stats['per_bounce']
    hits = eng.hits_per_surface.cpu().numpy()
    absorbed = eng.absorbed_per_surface.cpu().numpy()
    assert pb[0][0] == n and hits.sum() == sum(h for _, h, _ in pb)
    b1 = eng.tree[1]
    E1, P1 = b1._props['energy'].pieces[0], b1._props['parents'].pieces[0]
    # bounce 1: rays hit the dish (surface 0) or the back of the receiver (surface 1, shadowing the dish)
    shadow = 0.3 * 0.3 / (np.pi * 2.5 ** 2)
    first_recv = pb[0][1] - int((E1 == e_ray * (1. - 0.06)).sum().item())
    assert abs(first_recv / float(n) - shadow) < 5. * np.sqrt(shadow / n)
    assert 0.985 * n < pb[0][1] < n                   # the source disk is the dish aperture; rays of the rim tilted
                                                      # outward by the sunshape (10 m x ~2.5 mrad) miss: ~1 %
    dish_children = E1[:pb[0][1] - first_recv]
    assert bool((dish_children == e_ray * (1. - 0.06)).all()), 'dish children carry (1 - absorptivity) E'
    assert bool((P1[1:pb[0][1] - first_recv] > P1[:pb[0][1] - first_recv - 1]).all()), 'parents ascend per surface'
    # energy: everything that hit is either absorbed or leaves with the children, bounce by bounce
    e_in, e_children = pb[0][1] * e_ray, float(E1.sum().item())
    for k in range(2, eng.tree.num_bunds()):
        bk = eng.tree[k]
        prev = eng.tree[k - 1]._props['energy'].pieces[0]
        e_in += float(prev[bk._props['parents'].pieces[0]].sum().item())
        e_children += float(bk._props['energy'].pieces[0].sum().item())
    np.testing.assert_allclose(absorbed.sum() + e_children, e_in, rtol=1e-9)
    # focal spot: 2 mrad slope error + Buie disk on a 0.3 m target 3 m away -> most reflected rays arrive
    intercept = (hits[1] - first_recv) / float(pb[0][1] - first_recv)
    assert 0.97 < intercept <= 1.0
    # the tree walk gives the same tallies
    eng2 = TracerEngine(asm, rng='philox', seed=3, device_accel='bvh')
    eng2.ray_tracer(src, reps=3, min_energy=1e-10, tree=False)
    assert eng2.stats['per_bounce'] == pb
    assert np.array_equal(eng2.hits_per_surface.cpu().numpy(), hits)
    np.testing.assert_allclose(eng2.absorbed_per_surface.cpu().numpy(), absorbed, rtol=1e-12)
    torch.cuda.empty_cache()
