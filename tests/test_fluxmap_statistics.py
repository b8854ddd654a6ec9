"""Flux maps from INDEPENDENT random numbers (BASELINE.json north_star: "flux maps from
independent RNG must agree statistically within stated Monte Carlo error"; metric "flux-map L2
vs ref").  The reference side is tests/golden/c3_fluxmap_stats.npz: 64 batches x 2e5 rays of the
C3 NSTTF scene traced by the real reference with numpy.random (make_fluxmap_stats.py).

Stated bar: every bin of the 10 x 10 receiver map within 4.5 combined standard errors (100 bins:
a 4.5-sigma excursion has probability 7e-4 per test), total power within 4 standard errors, and
the whole-map relative L2 distance below 1.5 x the distance Monte-Carlo noise alone explains.
"""
import warnings

import numpy as np
import pytest

from tests import scenes
from tests.golden.make_fluxmap_stats import RES, receiver_map
from tests.helpers import batch_stats, fluxmap_agreement, load_golden

T = scenes.NS('tracer_b200')
Z_BIN, Z_TOTAL, L2_FACTOR = 4.5, 4.0, 1.5


def _check(maps, label):
    gold = load_golden('c3_fluxmap_stats')['maps']
    ref_mean, ref_sem = batch_stats(gold)
    mean, sem = batch_stats(maps)
    z, l2, l2_exp = fluxmap_agreement(mean, sem, ref_mean, ref_sem)
    tot, tot_ref = np.sum(maps, axis=(1, 2)), np.sum(gold, axis=(1, 2))
    s_tot = np.sqrt(tot.var(ddof=1) / len(tot) + tot_ref.var(ddof=1) / len(tot_ref))
    z_tot = abs(tot.mean() - tot_ref.mean()) / s_tot
    print('%s: receiver power %.4e W vs reference %.4e W (z = %.2f); max bin z = %.2f; '
          'flux-map L2 vs ref = %.4f (Monte-Carlo noise alone: %.4f)' % (
              label, tot.mean(), tot_ref.mean(), z_tot, z, l2, l2_exp))
    assert z_tot < Z_TOTAL, 'receiver power differs from the reference beyond Monte-Carlo error'
    assert z < Z_BIN, 'a flux-map bin differs from the reference beyond Monte-Carlo error'
    assert l2 < L2_FACTOR * l2_exp, 'flux-map L2 distance larger than Monte-Carlo noise explains'
    return l2, l2_exp


def test_fixture_is_self_consistent():
    """Two halves of the reference batches agree with each other under the same bar."""
    gold = load_golden('c3_fluxmap_stats')['maps']
    assert gold.shape[1:] == (RES, RES) and gold.shape[0] >= 16
    a, b = batch_stats(gold[0::2]), batch_stats(gold[1::2])
    z, l2, l2_exp = fluxmap_agreement(a[0], a[1], b[0], b[1])
    assert z < Z_BIN and l2 < L2_FACTOR * l2_exp


def test_oracle_fluxmap_within_mc_error_of_reference():
    """CPU: the NumPy oracle with its own seeds against the reference's map."""
    from oracle.numpy_trace import OracleEngine
    maps = []
    for b in range(6):
        np.random.seed(7000 + b)
        asm, source = scenes.c3_nsttf(T, n_rays=60000)
        recv_idx = len(asm.get_surfaces()) - 1
        eng = OracleEngine(asm)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            eng.ray_tracer(source(), reps=3, min_energy=1e-10, tree=False)
        absorbed, hits = eng.get_all_hits(recv_idx)
        maps.append(receiver_map(asm.get_surfaces()[recv_idx], absorbed, hits))
    _check(maps, 'oracle 6 x 6e4 rays')


@pytest.mark.gpu
def test_gpu_fluxmap_within_mc_error_of_reference():
    """GPU production mode: Philox Buie source in HBM, Philox slope errors, flux map binned on
    the device -- no random number shared with the reference."""
    import torch
    from bench import nsttf_source_geometry
    from tracer_b200.fluxmap import receiver_energy_bins
    from tracer_b200.tracer_engine import TracerEngine
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    recv = asm.get_surfaces()[-1]
    centre, direction, radius = nsttf_source_geometry()
    dev = torch.device('cuda', 0)
    n = 4000000
    maps = []
    for b in range(16):
        eng = TracerEngine(asm, rng='philox', seed=500 + b)
        src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False,
                                             seed=900 + b, device=dev)
        asm.reset_all_optics()
        eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=False)
        maps.append(receiver_energy_bins(recv, RES, device=dev).cpu().numpy())
    l2, l2_exp = _check(maps, 'GPU 16 x 4e6 rays')
    assert l2 < 0.05
