"""Generate the golden vectors of tests/golden/*.npz by running the REAL reference
(imported from /root/reference through oracle/ref_import.py).  Run here, in the
build container, whenever a scene in tests/scenes.py changes:

    python tests/golden/make_golden.py

The .npz files are committed; the GPU box never runs this script.
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from tests import scenes  # noqa: E402

# name -> (builder, kwargs, seed, reps, min_energy)
TRACE_CASES = {
    'c1_flat': (scenes.c1_flat_mirror, dict(n_rays=2000), 11, 3, 1e-10),
    'c2_dish_bivar': (scenes.c2_dish, dict(n_rays=1500, bi_var=True), 12, 3, 1e-10),
    'c2_dish_conical': (scenes.c2_dish, dict(n_rays=1500, bi_var=False), 13, 3, 1e-10),
    'c3_nsttf': (scenes.c3_nsttf, dict(n_rays=20000), 14, 3, 1e-10),
    'c3_nsttf_dense': (scenes.c3_nsttf, dict(n_rays=4000, n_hstats=40, dense=True), 15, 3, 1e-10),
    'c4_cavity': (scenes.c4_cavity, dict(n_rays=1500, n_theta=8, n_z=4), 16, 5, 1e-3 * 1000. * np.pi * 0.81 / 1500),
    'c5_cavity': (scenes.c5_cavity, dict(n_rays=1500), 17, 6, 1e-10),
    'c5_cavity_spectral': (scenes.c5_cavity, dict(n_rays=1500, spectral=True), 18, 6, 1e-10),
    'c6_optics_zoo': (scenes.c6_optics_zoo, dict(n_rays=3000), 19, 5, 1e-10),
    'c7_periodic_cell': (scenes.c7_periodic_cell, dict(n_rays=2000), 20, 6, 1e-10),
    'c8_fresnel_conductor': (scenes.c8_fresnel_conductor, dict(n_rays=2000), 21, 6, 1e-10),
    'c9_refractive': (scenes.c9_refractive, dict(n_rays=2000), 22, 7, 1e-3 * 1000. * np.pi * 1.21 / 2000),
    'c10_polychromatic': (scenes.c10_polychromatic, dict(n_rays=2000), 23, 5, 1e-3 * 1000. * np.pi * 1.44 / 2000),
    'c11_bifacial': (scenes.c11_bifacial, dict(n_rays=3000), 24, 6, 1e-3 * 1000. * np.pi * 2.56 / 3000),
    'c12_refractive_materials': (scenes.c12_refractive_materials, dict(n_rays=2000), 25, 8, 1e-4 * 1000. * np.pi * 0.81 / 2000),
}
# raw lists of the accountants (get_all_hits() raises in the reference when a PolychromaticAccountant is there)
_ACC_LISTS = ('_absorbed', '_attenuated', '_received', '_scattered', '_wavelengths', '_spectra', '_hits',
              '_directions', '_normals')
ZOO_RAYS = 4000
ZOO_SEED = 99


def trace_reference(name):
    import logging
    T = scenes.NS('tracer')
    builder, kw, seed, reps, min_energy = TRACE_CASES[name]
    np.random.seed(seed)
    asm, source = builder(T, **kw)
    bundle = source()
    from tracer.tracer_engine import TracerEngine
    eng = TracerEngine(asm, loglevel=logging.DEBUG)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        eng.ray_tracer(bundle, reps=reps, min_energy=min_energy, tree=True)
    out = {'n_tree': eng.tree.num_bunds(), 'seed': seed, 'reps': reps, 'min_energy': min_energy}
    for k in range(eng.tree.num_bunds()):
        b = eng.tree[k]
        out['v%d' % k] = b.get_vertices()
        out['d%d' % k] = b.get_directions()
        out['e%d' % k] = b.get_energy()
        if k > 0:
            out['p%d' % k] = b.get_parents()
        if name == 'c9_refractive':
            out['r%d' % k] = b.get_ref_index()
        if name == 'c12_refractive_materials':          # complex indices: real and imaginary parts, compared as floats
            out['rr%d' % k] = np.real(b.get_ref_index())
            out['ri%d' % k] = np.imag(b.get_ref_index())
            out['wl%d' % k] = b.get_wavelengths()
        if name == 'c10_polychromatic':
            out['w%d' % k] = b.get_wavelengths()
            out['s%d' % k] = b.get_spectra()
    for s, surf in enumerate(asm.get_surfaces()):
        opt = surf.get_optics_manager()
        if type(opt).__name__ == 'BiFacial':
            # (front, back) lists; a side without accountants gives []
            for side, arrays in zip(('front', 'back'), opt.get_all_hits()):
                for j, arr in enumerate(arrays):
                    out['accb_%d_%s_%d' % (s, side, j)] = arr
        elif hasattr(opt, 'get_all_hits'):
            try:
                for j, arr in enumerate(opt.get_all_hits()):
                    out['acc_%d_%d' % (s, j)] = arr
            except AttributeError:
                # PolychromaticAccountant.get_data raises (optics_callables.py:1840-1848): record the raw lists
                for j, acc in enumerate(opt.accountants):
                    for field in _ACC_LISTS:
                        chunks = getattr(acc, field, None)
                        if chunks:
                            out['accl_%d_%d%s' % (s, j, field)] = np.hstack(chunks)
    return out


def zoo_reference():
    """find_intersections / normals / hit points of every GM class on fixed rays."""
    T = scenes.NS('tracer')
    out = {}
    for name, (make, frame) in scenes.gm_zoo(T).items():
        rays = scenes.zoo_rays(T, ZOO_RAYS, ZOO_SEED)
        gm = make()
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            prm = np.array(gm.find_intersections(frame, rays), dtype=float)
            hit = np.nonzero(np.isfinite(prm) & (prm != 0.))[0]
            out[name + '_params'] = prm
            if len(hit) and name != 'flat':
                gm.select_rays(hit)
                out[name + '_normals'] = np.array(gm.get_normals())
                out[name + '_points'] = np.array(gm.get_intersection_points_global())
    return out


def fluxmap_reference():
    """get_fluxmap of the reference's geometry managers on fixed hit points."""
    T = scenes.NS('tracer')
    out = {}
    for name, (gm, frame, res, local) in scenes.fluxmap_cases(T).items():
        surf = T.surface.Surface(gm, T.optics_callables.Reflective(0.))
        surf._temp_frame = frame
        hits = np.dot(frame, np.vstack([local, np.ones(local.shape[1])]))[:3]
        eners = np.random.RandomState(5).uniform(0.5, 1.5, size=local.shape[1])
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            flux = gm.get_fluxmap(eners.copy(), surf.global_to_local(hits)[:3], res)
        out[name + '_hits'], out[name + '_eners'], out[name + '_flux'] = hits, eners, np.asarray(flux)
    return out


def main():
    ref_import.enable()
    only = sys.argv[1:]          # optional: names of the trace cases to (re)generate, nothing else
    if only == ['--zoo']:        # only the geometry zoo
        z = zoo_reference()
        np.savez_compressed(os.path.join(HERE, 'gm_zoo.npz'), **z)
        print('gm_zoo', {k[:-7]: int(np.isfinite(v).sum()) for k, v in z.items() if k.endswith('_params')})
        return
    for name in (only or TRACE_CASES):
        data = trace_reference(name)
        np.savez_compressed(os.path.join(HERE, name + '.npz'), **data)
        print(name, 'bundles:', [data['v%d' % k].shape[1] for k in range(data['n_tree'])])
    if only:
        return
    np.savez_compressed(os.path.join(HERE, 'fluxmaps.npz'), **fluxmap_reference())
    z = zoo_reference()
    np.savez_compressed(os.path.join(HERE, 'gm_zoo.npz'), **z)
    print('gm_zoo', {k[:-7]: int(np.isfinite(v).sum()) for k, v in z.items() if k.endswith('_params')})


if __name__ == '__main__':
    main()
