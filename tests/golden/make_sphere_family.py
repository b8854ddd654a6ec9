"""Golden vectors for the three sphere-family geometry managers whose root selection is bundle-order dependent in
the reference (sphere_surface.py:138, :203, :227: ``select[one_hit] = N.nonzero(mask[:, one_hit])[0]`` hands the
row indices out in row-major order, so a bundle that mixes rays whose only valid root is #0 with rays whose only
valid root is #1 gets them crossed).  On a bundle of ONE ray the line is well defined, so the REAL reference is
run ray by ray here: that is the behaviour the reference itself has whenever it is well defined, and it is what
the oracle ("intended per-ray choice") and the CUDA path implement.  Also records what the reference returns for
the whole bundle at once, to show how many rays the quirk crosses.

    python tests/golden/make_sphere_family.py        (needs /root/reference; ~1 min)
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref_import  # noqa: E402
from tracer_b200.models import scenes  # noqa: E402

KINDS = ('hemisphere', 'cut_sphere', 'cut_sphere_plane', 'sphere_rect')
N_RAYS, SEED = 6000, 321


def main():
    ref_import.enable()
    T = scenes.NS('tracer')
    out = {}
    for name in KINDS:
        make, frame = scenes.gm_zoo(T)[name]
        rays = scenes.zoo_rays(T, N_RAYS, SEED)
        V, D = rays.get_vertices(), rays.get_directions()
        prm = np.empty(N_RAYS)
        pts, nrm = [], []
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            whole = np.array(make().find_intersections(frame, rays), dtype=float)
            for i in range(N_RAYS):
                gm = make()
                one = T.ray_bundle.RayBundle(vertices=V[:, i:i + 1].copy(), directions=D[:, i:i + 1].copy(),
                                             energy=np.ones(1))
                p = float(np.asarray(gm.find_intersections(frame, one), dtype=float)[0])
                prm[i] = p
                if np.isfinite(p) and p != 0.:
                    gm.select_rays(np.array([0]))
                    nrm.append(np.array(gm.get_normals())[:, 0])
                    pts.append(np.array(gm.get_intersection_points_global())[:, 0])
        out[name + '_params'] = prm
        out[name + '_points'] = np.array(pts).T
        out[name + '_normals'] = np.array(nrm).T
        out[name + '_params_whole_bundle'] = whole
        differ = ~((prm == whole) | (np.isinf(prm) & np.isinf(whole)))
        print(name, 'hits', int(np.isfinite(prm).sum()), 'rays the whole-bundle call gets differently:', int(differ.sum()))
    np.savez_compressed(os.path.join(HERE, 'sphere_family.npz'), **out)


if __name__ == '__main__':
    main()
