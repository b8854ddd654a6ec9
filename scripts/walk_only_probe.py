"""Device time of the nearest-hit pass alone (trc_nearest_hit, tree walk) over a C3 source bundle, and the
occupancy the launch gets.  python scripts/walk_only_probe.py [rays]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tracer_b200.models import scenes
from tracer_b200 import _lib as L
from tracer_b200.scene import CompiledScene
from bench import nsttf_source_geometry
T = scenes.NS("tracer_b200")
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000000
asm, _ = scenes.c3_nsttf(T, n_rays=10)
sc = CompiledScene(asm)
centre, direction, radius = nsttf_source_geometry()
src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False, seed=99)
V, D = src._props['vertices'].pieces[0], src._props['directions'].pieces[0]
hit = torch.empty(n, dtype=torch.int32, device=V.device)
lib = L.load()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
best = 1e9
for it in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    L.check(lib.trc_nearest_hit(sc.handle, n, L.dptr(V), L.dptr(D), V.stride(0), 1, L.dptr(hit), st))
    e1.record(); torch.cuda.synchronize()
    if it >= 2: best = min(best, e0.elapsed_time(e1))
print('walk %.3f ms for %d rays; hits %d' % (best, n, int((hit >= 0).sum().item())))
