"""C2 (dish, 1e7 Buie rays) once, for an ncu capture of the kernels the headline configuration does not use:
buie_disk_kernel, nearest_hit_brute_kernel, the optics kernel on a quadric, fluxmap_rect_kernel.
    ncu --set full -k regex:"buie_disk|nearest_hit_brute|fluxmap_rect" --launch-skip 6 -c 4 python scripts/c2_profile.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tests import scenes  # noqa: E402
from tracer_b200.fluxmap import receiver_energy_bins  # noqa: E402
from tracer_b200.tracer_engine import TracerEngine  # noqa: E402

T = scenes.NS('tracer_b200')
asm, _ = scenes.c2_dish(T, n_rays=10)
eng = TracerEngine(asm, rng='philox', seed=1)
for it in range(3):
    src = T.sources.buie_sunshape_device(10000000, np.c_[[0., 0., 10.]], np.r_[0., 0., -1.], 2.5, 0.05, flux=1000., seed=2)
    asm.reset_all_optics()
    eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=True)
    bins = receiver_energy_bins(asm.get_surfaces()[1], 50)
torch.cuda.synchronize()
print(eng.stats['per_bounce'], float(bins.sum()))
