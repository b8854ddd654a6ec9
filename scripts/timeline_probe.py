"""GPU timeline of bench steps from torch.profiler (CUPTI): warm per-kernel time, idle gaps.
    python scripts/timeline_probe.py [rays] [c3|c4]"""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
from tests import scenes
from bench import nsttf_source_geometry, REPS, FLUX_RES
from tracer_b200 import tracer_engine as TE
from tracer_b200.fluxmap import receiver_energy_bins
T = scenes.NS("tracer_b200")
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000000
if len(sys.argv) > 2 and sys.argv[2] == 'c4':
    asm, _ = scenes.c4_cavity_mesh(T, n_theta=1000, n_z=500)
    eng = TE.TracerEngine(asm, rng="philox", seed=9, device_accel="bvh")
    src = T.sources.disk_bundle_device(n, np.c_[[0., 0., -0.01]], np.r_[0., 0., 1.], 0.9, np.pi / 4., flux=1000., seed=5)
    def step():
        asm.reset_all_optics()
        eng.ray_tracer(src, reps=5, min_energy=1e-12, tree=False)
else:
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    recv = asm.get_surfaces()[-1]
    eng = TE.TracerEngine(asm, rng="philox", seed=1, device_accel="bvh")
    centre, direction, radius = nsttf_source_geometry()
    src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False, seed=99)
    def step():
        asm.reset_all_optics()
        eng.ray_tracer(src, reps=REPS, min_energy=1e-10, tree=True)
        return receiver_energy_bins(recv, FLUX_RES)
for _ in range(3): step()
torch.cuda.synchronize()
STEPS = 4
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(STEPS): step()
    torch.cuda.synchronize()
f = tempfile.mktemp(suffix='.json')
prof.export_chrome_trace(f)
ev = [e for e in json.load(open(f))['traceEvents'] if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
ev.sort(key=lambda e: e['ts'])
t0, t1 = ev[0]['ts'], max(e['ts'] + e['dur'] for e in ev)
busy = sum(e['dur'] for e in ev)
print('span %.3f ms/step, busy %.3f ms/step, idle %.3f ms/step, %d launches/step' % (
    (t1 - t0) / STEPS / 1e3, busy / STEPS / 1e3, (t1 - t0 - busy) / STEPS / 1e3, len(ev) // STEPS))
agg = {}
for e in ev:
    k = e['name'][:70]
    a = agg.setdefault(k, [0, 0.])
    a[0] += 1; a[1] += e['dur']
for k, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print('  %-70s %4d  %8.1f us/step' % (k, c // STEPS, d / STEPS))
if os.environ.get('TRC_PROBE_LIST'):
    per = len(ev) // STEPS
    print('launches of the last step (us):')
    for e in ev[-per:]:
        print('  %8.1f  %s' % (e['dur'], e['name'][:60]))
# largest idle gaps of one step
gaps = []
end = ev[0]['ts'] + ev[0]['dur']
for a, b in zip(ev[:-1], ev[1:]):
    g = b['ts'] - max(end, a['ts'] + a['dur'])
    end = max(end, a['ts'] + a['dur'])
    if g > 0: gaps.append((g, a['name'][:40], b['name'][:40]))
gaps.sort(reverse=True)
print('largest gaps (us): ')
for g, a, b in gaps[:16]:
    print('  %7.1f  after %-40s before %s' % (g, a, b))
