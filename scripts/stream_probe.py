"""Streaming rate of a ray bundle through persistent warps (trc_stream_probe): groups in flight per warp by register
prefetch vs bulk copies of whole chunk rows.   python scripts/stream_probe.py [rays]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tracer_b200 import _lib as L  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000000
lib = L.load()
V = torch.rand((3, n), dtype=torch.float64, device='cuda')
D = torch.rand((3, n), dtype=torch.float64, device='cuda')
for variant in (1, 2, 4, 11, 12, 14, 108):
    ms = C.c_float(0.)
    L.check(lib.trc_stream_probe(variant, L.dptr(V), L.dptr(D), n, n, C.byref(ms)))
    print('variant %3d: %.3f ms  %.2f TB/s' % (variant, ms.value, 48. * n / (ms.value * 1e-3) / 1e12))
