"""Binary ray-source files: the data format on the input side of the bounce loop
(ray_trace_utils/Source_formatter.py:166-233 of the reference).

A source file is ``7 n`` little-endian float64 values, rows one after the other:
``x[n] y[n] z[n] dx[n] dy[n] dz[n] energy[n]`` (``binarize_source``, :166-172).  That is exactly
the SoA layout of a bundle in HBM (vertices (3, n), directions (3, n), energy (n,)), so
``load_source_bundle`` reads the file into page-locked memory and moves it with ONE
host-to-device copy of 56 B/ray into one device allocation that the three bundle columns view;
``binarize_source`` accepts host arrays or a device-backed ``RayBundle``'s columns.

``load_source`` / ``load_sources`` / ``adjust_ener`` / ``push_back`` / ``push_to`` keep the
reference's names, arguments and file effects (the reference computes ``len(X)/7`` with true
division, which no longer indexes under Python 3; integer division is used here and a size that
is not a multiple of 7 raises ValueError).
"""
import numpy as np

from .ray_bundle import RayBundle, _DeviceColumns

_ROWS = 7


def _rows_of(fname_or_array):
    X = np.fromfile(fname_or_array) if isinstance(fname_or_array, (str, bytes)) or hasattr(
        fname_or_array, '__fspath__') else np.asarray(fname_or_array, dtype=np.float64)
    if X.size % _ROWS:
        raise ValueError("source file holds %d float64 values: not 7 rows of equal length" % X.size)
    return X.reshape(_ROWS, X.size // _ROWS)


def _as_host(a):
    if isinstance(a, _DeviceColumns):
        return a.to_host()
    if hasattr(a, 'detach'):             # torch tensor
        return a.detach().cpu().numpy()
    return np.asarray(a, dtype=np.float64)


def binarize_source(vs, ds, es, filename):
    """Write positions (3, n), directions (3, n) and energies (n,) as one source file."""
    vs, ds, es = _as_host(vs), _as_host(ds), _as_host(es)
    if vs.ndim != 2 or vs.shape[0] != 3 or ds.shape != vs.shape or es.shape != (vs.shape[1],):
        raise ValueError("binarize_source needs vs (3,n), ds (3,n), es (n,)")
    np.concatenate((np.ravel(vs), np.ravel(ds), es)).astype(np.float64).tofile(filename)


def save_bundle(bundle, filename):
    """``binarize_source`` of a RayBundle (host- or device-backed)."""
    props = getattr(bundle, '_props', None)
    if props is not None:
        binarize_source(props['vertices'], props['directions'], props['energy'], filename)
    else:
        binarize_source(bundle.get_vertices(), bundle.get_directions(), bundle.get_energy(), filename)


def load_source(fname):
    """-> pos (3, n), dirs (3, n), ener (n,) as NumPy arrays."""
    X = _rows_of(fname)
    return X[0:3].copy(), X[3:6].copy(), X[6].copy()


def load_sources(list_fnames):
    """Several source files as ONE source: arrays concatenated, energies divided by the number
    of files (Source_formatter.py:185-201)."""
    if isinstance(list_fnames, (str, bytes)):
        list_fnames = [list_fnames]
    parts = [_rows_of(f) for f in list_fnames]
    X = np.concatenate(parts, axis=1)
    return X[0:3].copy(), X[3:6].copy(), X[6] / len(parts)


def load_source_bundle(fname, device=None, energy_scale=1.):
    """The file as a device-backed RayBundle: read into page-locked memory, one H2D copy."""
    import torch
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    n_values = int(np.memmap(fname, dtype=np.float64, mode='r').size)
    if n_values % _ROWS:
        raise ValueError("source file holds %d float64 values: not 7 rows of equal length" % n_values)
    n = n_values // _ROWS
    host = torch.empty(n_values, dtype=torch.float64, pin_memory=True)
    with open(fname, 'rb') as fh:
        got = fh.readinto(host.numpy())
    if got != 8 * n_values:
        raise IOError("short read of %s" % fname)
    buf = torch.empty(n_values, dtype=torch.float64, device=dev)
    buf.copy_(host, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()       # the pinned block is recycled after this
    V, D, E = buf[:3 * n].view(3, n), buf[3 * n:6 * n].view(3, n), buf[6 * n:]
    if energy_scale != 1.:
        E = E * energy_scale
    return RayBundle(vertices=_DeviceColumns([V]), directions=_DeviceColumns([D]), energy=_DeviceColumns([E]))


def load_sources_bundle(list_fnames, device=None):
    """``load_sources`` on the device: one bundle, energies divided by the number of files."""
    import torch
    bundles = [load_source_bundle(f, device, 1. / len(list_fnames)) for f in list_fnames]
    cat = (lambda name, dim: torch.cat([b._props[name].pieces[0] for b in bundles], dim=dim))
    return RayBundle(vertices=_DeviceColumns([cat('vertices', 1)]), directions=_DeviceColumns([cat('directions', 1)]),
                     energy=_DeviceColumns([cat('energy', 0)]))


def adjust_ener(fname, new_ener_tot):
    """Rewrite the file with a uniform energy per ray summing to ``new_ener_tot``."""
    pos, dirs, ener = load_source(fname)
    binarize_source(pos, dirs, new_ener_tot / len(ener) * np.ones(len(ener)), fname)


def push_back(fname, dist):
    """Move every ray origin back by ``dist`` along its direction."""
    pos, dirs, ener = load_source(fname)
    binarize_source(pos - dist * dirs, dirs, ener, fname)


def push_to(fname, z):
    """Move every ray origin along its direction onto the plane of height ``z``."""
    pos, dirs, ener = load_source(fname)
    binarize_source(pos + (z - pos[2]) / dirs[2] * dirs, dirs, ener, fname)
