"""Where a small trace (C1: 1e5 rays, 2 live bounces) spends its time: wall per trace, cProfile of the host side,
and the GPU timeline (busy time, launches) from torch.profiler."""
import cProfile, pstats, sys, time, os, json, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
from tracer_b200.models import scenes
from tracer_b200.tracer_engine import TracerEngine
T = scenes.NS("tracer_b200")
asm, _ = scenes.c1_flat_mirror(T, n_rays=10)
src = T.sources.disk_bundle_device(100000, np.c_[[0., 0., 4.]], np.r_[0., 0., -1.], 0.8, 4.65e-3, flux=1000., seed=1)
eng = TracerEngine(asm, rng='philox', seed=1)
def step():
    asm.reset_all_optics()
    eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=True)
for _ in range(5): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(200): step()
torch.cuda.synchronize(); print('wall us/trace %.1f' % ((time.perf_counter() - t0) / 200 * 1e6))
pr = cProfile.Profile(); pr.enable()
for _ in range(100): step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(14)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(4): step()
    torch.cuda.synchronize()
f = tempfile.mktemp(suffix='.json'); prof.export_chrome_trace(f)
ev = [e for e in json.load(open(f))['traceEvents'] if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
ev.sort(key=lambda e: e['ts'])
per = len(ev) // 4
last = ev[-per:]
print('GPU: %d launches/trace, busy %.1f us, span %.1f us' % (per, sum(e['dur'] for e in last), last[-1]['ts'] + last[-1]['dur'] - last[0]['ts']))
for e in last: print('  %7.1f  +%7.1f  %s' % (e['dur'], e['ts'] - last[0]['ts'], e['name'][:60]))
