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


def golden_accountants(gold):
    out = {}
    for key in gold.files:
        if key.startswith('acc_'):
            _, s, j = key.split('_')
            out.setdefault(int(s), {})[int(j)] = gold[key]
    return {s: [v[j] for j in sorted(v)] for s, v in out.items()}
