"""Configuration 4 at full size on one GPU: 1 001 000-face cylinder cavity (TriangleMeshObject),
Lambertian receiver, pillbox disk source; prints one JSON line with the per-bounce
nearest-hit kernel times.  Usage: python scripts/c4_full_size.py [rays] [reps] [n_theta] [n_z] [leaf_size]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tracer_b200.models import scenes  # noqa: E402


def main():
    import torch
    from tracer_b200.tracer_engine import TracerEngine
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 20000000
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    n_theta = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
    n_z = int(sys.argv[4]) if len(sys.argv) > 4 else 500
    leaf = int(sys.argv[5]) if len(sys.argv) > 5 else 1
    T = scenes.NS('tracer_b200')
    t0 = time.time()
    asm, _ = scenes.c4_cavity_mesh(T, n_theta=n_theta, n_z=n_z)
    t_build = time.time() - t0
    eng = TracerEngine(asm, rng='philox', seed=9, device_accel='bvh', leaf_size=leaf)
    t0 = time.time()
    eng.compile()
    t_compile = time.time() - t0
    src = T.sources.disk_bundle_device(n, np.c_[[0., 0., -0.01]], np.r_[0., 0., 1.], 0.9, np.pi / 4., flux=1000., seed=5)
    e_ray = float(src._props['energy'].pieces[0][0].item())
    out = []
    for it in range(2 if n > 50000000 else 3):
        asm.reset_all_optics()
        eng.kernel_events = []
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        eng.ray_tracer(src, reps=reps, min_energy=1e-3 * e_ray, tree=False)     # the config's min_energy
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        pb = eng.stats['per_bounce']
        out.append(dict(ms=ms, ray_bounces_per_s=sum(p[0] for p in pb) / ms * 1e3,
                        kernel_ms=[round(e[3](), 3) for e in eng.kernel_events], per_bounce=pb,
                        peak_mem_gb=round(torch.cuda.max_memory_allocated() / 1e9, 1)))
    best = min(out, key=lambda o: o['ms'])
    print(json.dumps(dict(workload='C4 cylinder cavity mesh, LambertianReceiver(0.7)', faces=len(asm.get_surfaces()),
                          rays=n, reps=reps, leaf_size=leaf, host_build_s=round(t_build, 2), host_compile_upload_s=round(t_compile, 2),
                          **best)))


if __name__ == '__main__':
    main()
