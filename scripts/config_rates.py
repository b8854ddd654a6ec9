"""Ray-bounces/s of the other BASELINE configurations at their named sizes (device-resident Philox
sources, CUDA-event timed, best of 5 after 2 warm-up traces).  Not bench lines: context for DESIGN.md.

    python scripts/config_rates.py
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tests import scenes  # noqa: E402
from tracer_b200.tracer_engine import TracerEngine  # noqa: E402

T = scenes.NS('tracer_b200')


def rate(asm, src, reps, accel):
    eng = TracerEngine(asm, rng='philox', seed=1, device_accel=accel)
    best = None
    for it in range(7):
        asm.reset_all_optics()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.ray_tracer(src, reps=reps, min_energy=1e-10, tree=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if it >= 2:
            best = ms if best is None else min(best, ms)
    return dict(ms=best, ray_bounces=eng.stats['ray_bounces'], rate=eng.stats['ray_bounces'] / (best * 1e-3),
                per_bounce=eng.stats['per_bounce'], surfaces=eng.stats['surfaces'])


def main():
    out = {}
    asm, _ = scenes.c1_flat_mirror(T, n_rays=10)
    src = T.sources.disk_bundle_device(100000, np.c_[[0., 0., 4.]], np.r_[0., 0., -1.], 0.8, 4.65e-3, flux=1000., seed=1) \
        if hasattr(T.sources, 'disk_bundle_device') else None
    if src is not None:
        out['C1 flat mirror, 1e5 rays, reps=3'] = rate(asm, src, 3, 'auto')
    asm, _ = scenes.c2_dish(T, n_rays=10)
    src = T.sources.buie_sunshape_device(10000000, np.c_[[0., 0., 10.]], np.r_[0., 0., -1.], 2.5, 0.05, flux=1000., seed=2)
    out['C2 dish, 1e7 Buie rays, reps=3'] = rate(asm, src, 3, 'auto')
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
