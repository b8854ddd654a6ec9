"""Host time of the sections of one bench step (perf_counter, no added synchronisation)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests import scenes
from bench import nsttf_source_geometry, REPS, FLUX_RES
from tracer_b200 import tracer_engine as TE, fluxmap
T = scenes.NS("tracer_b200")
asm, _ = scenes.c3_nsttf(T, n_rays=10)
recv = asm.get_surfaces()[-1]
eng = TE.TracerEngine(asm, rng="philox", seed=1, device_accel="bvh")
centre, direction, radius = nsttf_source_geometry()
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000000
src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False, seed=99)
acc = {}
def timed(obj, name, label=None):
    fn = getattr(obj, name)
    def w(*a, **k):
        t0 = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            acc[label or name] = acc.get(label or name, 0.) + time.perf_counter() - t0
    setattr(obj, name, w)
for nm in ('_trace_block', '_account', '_as_bundle', '_to_host', '_to_device', 'compile'):
    timed(eng, nm)
timed(asm, 'reset_all_optics')
timed(fluxmap, 'receiver_energy_bins')
lib = TE.L.load()
def step():
    asm.reset_all_optics()
    t0 = time.perf_counter()
    eng.ray_tracer(src, reps=REPS, min_energy=1e-10, tree=True)
    acc['ray_tracer'] = acc.get('ray_tracer', 0.) + time.perf_counter() - t0
    return fluxmap.receiver_energy_bins(recv, FLUX_RES)
for _ in range(3): step()
torch.cuda.synchronize(); acc.clear()
N = 10
t0 = time.perf_counter()
for _ in range(N): step()
torch.cuda.synchronize()
print('step %.3f ms' % ((time.perf_counter() - t0) / N * 1e3))
for k, v in sorted(acc.items(), key=lambda kv: -kv[1]):
    print('  %-24s %8.1f us/step' % (k, v / N * 1e6))
