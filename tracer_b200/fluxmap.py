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
    frame = surface._temp_frame
    cached = surface.__dict__.get('_b200_proj')          # frames are replaced, never edited in place (has_frame.py:75)
    if cached is None or cached[0] is not frame:
        proj = np.round(np.linalg.inv(frame), decimals=9)[:3]   # surface.py:125
        cached = (frame, L.host_doubles(proj, 12))
        surface.__dict__['_b200_proj'] = cached
    _, pp = cached[1]
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    bins = torch.zeros((resolution, resolution), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for hits, absorbed in zip(_chunks(accs['Location']._hits), _chunks(accs['Absorber']._absorbed)):
            th = hits.t if getattr(hits, 't', None) is not None else torch.from_numpy(np.asarray(hits)).to(dev)
            te = absorbed.t if getattr(absorbed, 't', None) is not None else torch.from_numpy(np.asarray(absorbed)).to(dev)
            th = th if th.stride(-1) == 1 else th.contiguous()      # rows may be strided views of a bounce arena
            te = te.contiguous()
            n = int(te.shape[0])
            L.check(lib.trc_fluxmap_rect(n, L.dptr(th), th.stride(0), L.dptr(te), pp, float(hx), float(hy),
                                         resolution, resolution, L.dptr(bins), st))
    return bins


def bin_area(surface, resolution):
    gm = surface.get_geometry_manager()
    hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
    return (2. * hx / resolution) * (2. * hy / resolution)


# ---------------------------------------------------------------- all mesh kinds
def _mesh(gm, resolution):
    """(coord kind, param, u edges, v edges, bin areas (nu, nv)) of the mesh the
    reference's ``get_fluxmap`` of this geometry manager uses."""
    N = np
    names = [c.__name__ for c in type(gm).__mro__]
    if 'StraightCutRoundPlateGM' in names and getattr(gm, 'x_cut', None) is not None:
        # the reference's own get_fluxmap raises TypeError here (flat_surface.py:620-632 slices and sizes arrays
        # with resolution / 3, a float): there is no map to reproduce
        raise NotImplementedError("flux map of StraightCutRoundPlateGM: the reference's get_fluxmap raises TypeError")
    if 'ExtrudedRectPlateGM' in names:
        raise NotImplementedError("ExtrudedRectPlateGM has four meshes: use _mesh_regions / surface_fluxmap")
    if 'RectPlateGM' in names:                                        # flat_surface.py:237-251
        xs = N.linspace(-gm._half_dims[0, 0], gm._half_dims[0, 0], resolution + 1)
        ys = N.linspace(-gm._half_dims[1, 0], gm._half_dims[1, 0], resolution + 1)
        dxs = N.tile(N.abs(xs[1:] - xs[:-1]), (len(ys) - 1, 1))
        dys = N.tile(N.abs(ys[1:] - ys[:-1]), (len(xs) - 1, 1)).T
        return L.FLUX_XY, 0., xs, ys, dxs * dys
    if 'RoundPlateGM' in names:                                       # flat_surface.py:524-545
        if resolution is None:
            resolution = 40
        angs = N.r_[0.:2. * N.pi + 2. * N.pi / resolution:2. * N.pi / resolution]
        if gm._Ri is not None:
            rs = gm._Ri + (gm._Re - gm._Ri) / resolution * N.arange(0, resolution + 1)
        else:
            rs = gm._Re / resolution * N.arange(0, resolution + 1)
        drs = N.tile(N.abs(rs[1:] - rs[:-1]), (len(angs) - 1, 1)).T
        ravgs = N.tile((rs[1:] + rs[:-1]) / 2., (len(angs) - 1, 1)).T
        dangs = N.tile(N.abs(angs[1:] - angs[:-1]), (len(rs) - 1, 1))
        return L.FLUX_R_ATAN_XY, 0., rs, angs, drs * ravgs * dangs
    if 'ParabolicDishGM' in names:                                    # paraboloid.py:151-172
        r_end = gm._R * (1. + 1. / resolution)
        r = N.r_[0:r_end:gm._R / resolution]
        ang_end = 2. * N.pi * (1 + 1. / resolution)
        ang = N.r_[0:ang_end:2 * N.pi / resolution][:resolution + 1]
        rs = N.tile(r, (len(ang) - 1, 1)).T
        dangs = N.repeat(N.vstack(ang[1:] - ang[:-1]), len(r) - 1, axis=1)
        areas = dangs * ((4. * gm.a ** 2 * rs[1:] ** 2 + 1.) ** (3. / 2.) - (4. * gm.a ** 2 * rs[:-1] ** 2 + 1.) ** (3. / 2.))
        return L.FLUX_R_ATAN_YX, 0., r, ang, areas
    if 'FiniteCylinder' in names:                                     # cylinder.py:139-159
        h_bins = N.linspace(-gm._half_h, gm._half_h, resolution + 1)
        ang_bins = N.linspace(gm._ang_range[0], gm._ang_range[1], resolution + 1)
        dh = N.abs(N.repeat(N.vstack(h_bins[1:] - h_bins[:-1]), len(ang_bins) - 1, axis=1)).T
        dangs = N.repeat(N.vstack(ang_bins[1:] - ang_bins[:-1]), len(h_bins) - 1, axis=1)
        return L.FLUX_Z_PHI, 0., h_bins, ang_bins, dangs * gm._R * dh
    if 'SphericalGM' in names:                                        # sphere_surface.py:100-115
        ths_bin = N.linspace(0., N.pi, resolution + 1)
        phis_bin = N.linspace(0., 2. * N.pi, resolution * 2 + 1)
        tb = N.vstack(ths_bin)
        dphi = phis_bin[1:] - phis_bin[:-1]
        areas = gm._rad ** 2. * dphi * (N.cos(tb[:-1]) - N.cos(tb[1:]))
        return L.FLUX_THETA_PHI, float(gm._rad), ths_bin, phis_bin, areas
    raise NotImplementedError(f"{type(gm).__name__} has no get_fluxmap in the reference")


def _mesh_regions(gm, resolution):
    """The meshes of ``gm.get_fluxmap`` as a list of (coord kind, param, u edges, v edges, bin areas): one entry
    for every class but ExtrudedRectPlateGM, whose plate is cut in four rectangles around the hole -- bottom,
    middle left, middle right, top, each resolution x resolution (flat_surface.py:321-356)."""
    N = np
    names = [c.__name__ for c in type(gm).__mro__]
    if 'ExtrudedRectPlateGM' not in names:
        return [_mesh(gm, resolution)]
    hx, hy = gm._half_dims[0, 0], gm._half_dims[1, 0]
    cx, cy = gm.extr_center[0], gm.extr_center[1]
    ex, ey = gm.extr_half_dims[0, 0], gm.extr_half_dims[1, 0]
    spans = [((-hx, hx), (-hy, cy - ey)), ((-hx, cx - ex), (cy - ey, cy + ey)),
             ((cx + ex, hx), (cy - ey, cy + ey)), ((-hx, hx), (cy + ey, hy))]
    out = []
    for (x0, x1), (y0, y1) in spans:
        xs, ys = N.linspace(x0, x1, resolution + 1), N.linspace(y0, y1, resolution + 1)
        dxs = N.tile(N.abs(xs[1:] - xs[:-1]), (len(ys) - 1, 1))
        dys = N.tile(N.abs(ys[1:] - ys[:-1]), (len(xs) - 1, 1)).T
        out.append((L.FLUX_XY, 0., xs, ys, dxs * dys))
    return out


def binned_energy(surface, hits, energy, resolution, device=None, region=0):
    """Energy per mesh bin, (nu, nv) float64 CUDA tensor, of hit points (3,n, global
    coordinates) weighted by ``energy`` (n,), both CUDA tensors or host arrays.
    Returns (bins, areas).  ``region`` picks one of the meshes of a multi-mesh class (_mesh_regions)."""
    import torch
    lib = L.load()
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    coord, param, eu, ev, areas = _mesh_regions(surface.get_geometry_manager(), resolution)[region]
    proj = np.round(np.linalg.inv(surface._temp_frame), decimals=9)[:3]   # surface.py:125
    _, pp = L.host_doubles(proj, 12)
    nu, nv = len(eu) - 1, len(ev) - 1
    bins = torch.zeros((nu, nv), dtype=torch.float64, device=dev)
    d_eu = torch.from_numpy(np.ascontiguousarray(eu, dtype=np.float64)).to(dev)
    d_ev = torch.from_numpy(np.ascontiguousarray(ev, dtype=np.float64)).to(dev)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        chunks = zip(hits, energy) if isinstance(hits, (list, tuple)) else [(hits, energy)]
        for h, e in chunks:
            th = h if torch.is_tensor(h) else torch.from_numpy(np.ascontiguousarray(h, dtype=np.float64))
            te = e if torch.is_tensor(e) else torch.from_numpy(np.ascontiguousarray(e, dtype=np.float64))
            th, te = th.to(dev).contiguous(), te.to(dev).contiguous()
            n = int(te.shape[0])
            if n == 0:
                continue
            L.check(lib.trc_fluxmap_binned(coord, n, L.dptr(th), th.stride(0), L.dptr(te), pp, float(param),
                                           L.dptr(d_eu), nu, L.dptr(d_ev), nv, L.dptr(bins), st))
    return bins, areas


def surface_fluxmap(surface, resolution, device=None):
    """What ``GM.get_fluxmap(absorbed, Surface.global_to_local(hits), resolution)`` returns in
    the reference (flat_surface.py:237,524; paraboloid.py:151; cylinder.py:139;
    sphere_surface.py:100) -- the flattened flux array -- computed from the accountant
    data of the last trace while it is still in HBM; only the map is copied to the host."""
    opt = surface.get_optics_manager()
    accs = {a.shorthand: a for a in opt.accountants}
    if 'Location' not in accs or not any(k in accs for k in ('Absorber', 'Receptor', 'Scatterer')):
        raise ValueError("surface_fluxmap needs a Location accountant and an energy accountant")
    from .scene import _ACC_FIELDS
    en = accs.get('Absorber') or accs.get('Receptor') or accs.get('Scatterer')
    hits = [c.t if getattr(c, 't', None) is not None else np.asarray(c) for c in _chunks(accs['Location']._hits)]
    ener = [c.t if getattr(c, 't', None) is not None else np.asarray(c)
            for c in _chunks(getattr(en, _ACC_FIELDS[en.shorthand][0]))]
    out = []
    for region in range(len(_mesh_regions(surface.get_geometry_manager(), resolution))):
        bins, areas = binned_energy(surface, hits, ener, resolution, device, region=region)
        out.append(np.hstack(bins.cpu().numpy() / areas))
    return out[0] if len(out) == 1 else tuple(out)      # ExtrudedRectPlateGM returns its four maps (flat_surface.py:356)
