"""Shared comparison helpers for the parity tests."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
# stated tolerance (BASELINE.json north_star): positions, directions, energies
# within 1e-5 relative; indices bit-exact.  The fp64 kernels are far inside it.
RTOL = 1e-5


def load_golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + '.npz'))


def close(a, b, rtol=RTOL, what=''):
    a, b = np.asarray(a, float), np.asarray(b, float)
    assert a.shape == b.shape, f'{what}: shape {a.shape} != {b.shape}'
    if a.size == 0:
        return
    scale = np.maximum(1.0, np.abs(b))
    err = np.abs(a - b) / scale
    assert np.all(err <= rtol), f'{what}: max rel err {err.max():.3e} at {np.argmax(err)}'


def compare_tree(gold, tree, n_levels=None, rtol=RTOL):
    """tree: sequence of bundles with get_vertices/... ; gold: npz of make_golden.py.
    Parents and bundle sizes bit-exact, floats within rtol."""
    n = int(gold['n_tree'])
    have = tree.num_bunds() if hasattr(tree, 'num_bunds') else len(tree)
    assert have == n, f'tree depth {have} != golden {n}'
    for k in range(n if n_levels is None else n_levels):
        b = tree[k]
        assert b.get_num_rays() == gold['v%d' % k].shape[1], f'level {k}: ray count'
        if k > 0:
            assert np.array_equal(np.asarray(b.get_parents()), gold['p%d' % k]), f'level {k}: parents differ'
        close(b.get_vertices(), gold['v%d' % k], rtol, f'level {k} vertices')
        close(b.get_directions(), gold['d%d' % k], rtol, f'level {k} directions')
        close(b.get_energy(), gold['e%d' % k], rtol, f'level {k} energy')
        if 'r%d' % k in gold.files:
            assert np.array_equal(np.asarray(b.get_ref_index()), gold['r%d' % k]), f'level {k}: ref_index differs'
        if 'rr%d' % k in gold.files:                # complex refractive indices from material tables
            ri = np.asarray(b.get_ref_index())
            close(np.real(ri), gold['rr%d' % k], rtol, f'level {k} Re(ref_index)')
            im, ref_im = np.imag(ri), gold['ri%d' % k]          # k ~ 1e-8: relative, not absolute
            assert np.all(np.abs(im - ref_im) <= max(rtol, 1e-12) * np.abs(ref_im)), f'level {k} Im(ref_index)'
            close(b.get_wavelengths(), gold['wl%d' % k], rtol, f'level {k} wavelengths')
        if 'w%d' % k in gold.files:                 # polychromatic bundles: (W, n) wavelengths and spectra
            close(b.get_wavelengths(), gold['w%d' % k], rtol, f'level {k} wavelengths')
            close(b.get_spectra(), gold['s%d' % k], rtol, f'level {k} spectra')


def golden_accountants(gold):
    out = {}
    for key in gold.files:
        if key.startswith('acc_'):
            _, s, j = key.split('_')
            out.setdefault(int(s), {})[int(j)] = gold[key]
    return {s: [v[j] for j in sorted(v)] for s, v in out.items()}


def golden_bifacial(gold):
    """{surface: {'front': [arrays], 'back': [arrays]}} of the BiFacial managers of a golden trace."""
    out = {}
    for key in gold.files:
        if key.startswith('accb_'):
            _, s, side, j = key.split('_')
            out.setdefault(int(s), {}).setdefault(side, {})[int(j)] = gold[key]
    return {s: {side: [v[j] for j in sorted(v)] for side, v in d.items()} for s, d in out.items()}


def golden_accountant_lists(gold):
    """{surface: {(accountant index, list attribute): array}} of managers whose get_all_hits() raises in the
    reference (PolychromaticAccountant): the accountants' raw lists, stacked over the bounces."""
    out = {}
    for key in gold.files:
        if key.startswith('accl_'):
            _, s, rest = key.split('_', 2)
            j, field = rest.split('_', 1)
            out.setdefault(int(s), {})[(int(j), '_' + field)] = gold[key]
    return out


def fluxmap_agreement(mean_a, sem_a, mean_b, sem_b):
    """Two Monte-Carlo estimates of the same flux map from independent random numbers
    (BASELINE metric "flux-map L2 vs ref").  Returns (z_max, l2, l2_expected):
    the largest per-bin deviation in units of the combined standard error, the relative L2
    distance of the maps and the L2 distance pure Monte-Carlo noise would produce."""
    sigma = np.sqrt(np.asarray(sem_a, float) ** 2 + np.asarray(sem_b, float) ** 2)
    diff = np.asarray(mean_a, float) - np.asarray(mean_b, float)
    live = sigma > 0
    z_max = float(np.max(np.abs(diff[live]) / sigma[live])) if live.any() else 0.
    norm = np.sqrt(np.sum(np.asarray(mean_b, float) ** 2))
    return z_max, float(np.sqrt(np.sum(diff ** 2)) / norm), float(np.sqrt(np.sum(sigma ** 2)) / norm)


def batch_stats(maps):
    """Mean map and its standard error from per-batch maps (B, nx, ny)."""
    maps = np.asarray(maps, float)
    return maps.mean(axis=0), maps.std(axis=0, ddof=1) / np.sqrt(maps.shape[0])
