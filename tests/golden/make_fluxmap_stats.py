"""Generate tests/golden/c3_fluxmap_stats.npz: the receiver flux map of the C3 NSTTF scene
(218 heliostats + tower receiver, tests/scenes.py:c3_nsttf) traced by the REAL reference
(/root/reference through oracle/ref_import.py) with numpy.random, in independent batches.

    python tests/golden/make_fluxmap_stats.py [batches] [rays_per_batch]

Each batch is one ``TracerEngine.ray_tracer(reps=3)`` of the reference on a fresh
``buie_sunshape`` bundle (seed = 1000 + batch); the receiver's absorbed energies are binned
on its local (x, y) into RES x RES bins over the plate (numpy.histogram2d, the rule of
RectPlateGM.get_fluxmap, flat_surface.py:237-251, without the division by the bin area).
The file holds the per-batch maps [W per bin]; the tests derive the mean map and its
Monte-Carlo standard error from them.  The .npz is committed; the GPU box never runs this.
"""
import multiprocessing as mp
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

RES = 10


def receiver_map(surface, absorbed, hits, res=RES):
    """Absorbed energy per bin of a RectPlateGM receiver, local frame of the surface."""
    gm = surface.get_geometry_manager()
    hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
    inv = np.linalg.inv(surface._temp_frame)
    local = np.dot(inv, np.vstack([hits, np.ones(hits.shape[1])]))
    m, _, _ = np.histogram2d(local[0], local[1], bins=res, range=[[-hx, hx], [-hy, hy]], weights=absorbed)
    return m


def one_batch(args):
    batch, n_rays = args
    import logging
    from oracle import ref_import
    ref_import.enable()
    from tests import scenes
    from tracer.tracer_engine import TracerEngine
    T = scenes.NS('tracer')
    np.random.seed(1000 + batch)
    asm, source = scenes.c3_nsttf(T, n_rays=n_rays)
    recv = asm.get_surfaces()[-1]
    out = np.zeros((RES, RES))
    chunk = 50000                      # the reference's (S, N) relevance mask bounds the bundle size
    for a in range(0, n_rays, chunk):
        n = min(chunk, n_rays - a)
        _, src = scenes.c3_nsttf(T, n_rays=n)
        bundle = src()
        bundle.set_energy(bundle.get_energy() * (float(n) / n_rays))
        asm.reset_all_optics()
        eng = TracerEngine(asm, loglevel=logging.DEBUG)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            eng.ray_tracer(bundle, reps=3, min_energy=1e-10, tree=False)
        absorbed, hits = recv.get_optics_manager().get_all_hits()
        out += receiver_map(recv, absorbed, hits)
    return out


def main():
    batches = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    n_rays = int(float(sys.argv[2])) if len(sys.argv) > 2 else 200000
    with mp.get_context('fork').Pool(min(8, os.cpu_count() or 1)) as pool:
        maps = pool.map(one_batch, [(b, n_rays) for b in range(batches)])
    maps = np.array(maps)
    np.savez_compressed(os.path.join(HERE, 'c3_fluxmap_stats.npz'), maps=maps, rays_per_batch=n_rays, res=RES)
    mean = maps.mean(axis=0)
    sem = maps.std(axis=0, ddof=1) / np.sqrt(batches)
    print('power on receiver %.1f W +- %.1f; median relative bin error %.3f' % (
        mean.sum(), maps.sum(axis=(1, 2)).std(ddof=1) / np.sqrt(batches), np.median(sem[mean > 0] / mean[mean > 0])))


if __name__ == '__main__':
    main()
