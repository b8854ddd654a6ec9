import os, sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from tests import scenes
from bench import nsttf_source_geometry, REPS, FLUX_RES
from tracer_b200 import tracer_engine as TE
from tracer_b200 import _lib as L
from tracer_b200.fluxmap import receiver_energy_bins
T = scenes.NS("tracer_b200")
asm, _ = scenes.c3_nsttf(T, n_rays=10)
recv = asm.get_surfaces()[-1]
eng = TE.TracerEngine(asm, rng="philox", seed=1, device_accel="bvh")
centre, direction, radius = nsttf_source_geometry()
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000000
src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False, seed=99)
lib = L.load()
# wrap C entry points with sync timing
acc = {}
def wrap(name):
    fn = getattr(lib, name)
    def w(*a):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = fn(*a)
        torch.cuda.synchronize(); acc[name] = acc.get(name, 0.) + time.perf_counter() - t0
        return r
    return w
class Proxy(object):
    def __getattr__(self, k):
        return wrap(k) if k.startswith('trc_') and k not in ('trc_last_error','trc_order_work_bytes','trc_sort_work_bytes') else getattr(lib, k)
def step():
    asm.reset_all_optics()
    eng.ray_tracer(src, reps=REPS, min_energy=1e-10, tree=True)
    return receiver_energy_bins(recv, FLUX_RES)
for _ in range(3): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): step()
torch.cuda.synchronize(); base = (time.perf_counter() - t0) / 5
print('step ms (async)', base * 1e3)
orig = L.load
L._lib = Proxy()
acc.clear()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): step()
torch.cuda.synchronize(); tot = (time.perf_counter() - t0) / 5
print('step ms (serialised)', tot * 1e3)
for k, v in sorted(acc.items(), key=lambda kv: -kv[1]):
    print('  %-28s %8.3f ms/step' % (k, v / 5 * 1e3))
print('  other (python, torch ops, syncs) %8.3f ms/step' % ((tot - sum(acc.values()) / 5) * 1e3))
