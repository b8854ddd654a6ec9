"""Host-side cost of one step: cProfile over 30 steps at a ray count where the GPU is not the limit."""
import cProfile, pstats, sys, time
sys.path.insert(0, '/root/repo')
import torch
from tests import scenes
from bench import nsttf_source_geometry, REPS, FLUX_RES
from tracer_b200 import tracer_engine as TE
from tracer_b200.fluxmap import receiver_energy_bins
T = scenes.NS("tracer_b200")
asm, _ = scenes.c3_nsttf(T, n_rays=10)
recv = asm.get_surfaces()[-1]
eng = TE.TracerEngine(asm, rng="philox", seed=1, device_accel="bvh")
centre, direction, radius = nsttf_source_geometry()
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1000000
src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False, seed=99)
def step():
    asm.reset_all_optics()
    eng.ray_tracer(src, reps=REPS, min_energy=1e-10, tree=True)
    return receiver_energy_bins(recv, FLUX_RES)
for _ in range(5): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(30): step()
torch.cuda.synchronize(); print('ms/step', (time.perf_counter() - t0) / 30 * 1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(30): step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(28)
