"""On-device flux-map binning for rectangular receivers.

Replaces ``Surface.global_to_local`` + ``RectPlateGM.get_fluxmap``'s
numpy.histogram2d (reference surface.py:114-126, flat_surface.py:237-251) for
accountant data that is still in HBM: the per-bounce hit/absorbed chunks are
binned by trc_fluxmap_rect and only the (resolution x resolution) map leaves
the GPU.  Bins are energies; divide by ``bin_area`` for W/m^2.
"""
import ctypes as C

import numpy as np

from . import _lib as L


def _chunks(acc_list):
    return [c for c in acc_list if c.shape[-1]]


def receiver_energy_bins(surface, resolution, device=None):
    """(resolution, resolution) float64 CUDA tensor: absorbed energy per bin
    [ix][iy] of a RectPlateGM surface whose optics carries Absorber+Location
    accountants (e.g. ReflectiveReceiver)."""
    import torch
    lib = L.load()
    opt = surface.get_optics_manager()
    accs = {a.shorthand: a for a in opt.accountants}
    if 'Absorber' not in accs or 'Location' not in accs:
        raise ValueError("receiver_energy_bins needs Absorber and Location accountants")
    gm = surface.get_geometry_manager()
    hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
    proj = np.round(np.linalg.inv(surface._temp_frame), decimals=9)[:3]   # surface.py:125
    _, pp = L.host_doubles(proj, 12)
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    bins = torch.zeros((resolution, resolution), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for hits, absorbed in zip(_chunks(accs['Location']._hits), _chunks(accs['Absorber']._absorbed)):
            th = hits.t if getattr(hits, 't', None) is not None else torch.from_numpy(np.asarray(hits)).to(dev)
            te = absorbed.t if getattr(absorbed, 't', None) is not None else torch.from_numpy(np.asarray(absorbed)).to(dev)
            th, te = th.contiguous(), te.contiguous()
            n = int(te.shape[0])
            L.check(lib.trc_fluxmap_rect(n, L.dptr(th), th.stride(0), L.dptr(te), pp, float(hx), float(hy),
                                         resolution, resolution, L.dptr(bins), st))
    return bins


def bin_area(surface, resolution):
    gm = surface.get_geometry_manager()
    hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
    return (2. * hx / resolution) * (2. * hy / resolution)
