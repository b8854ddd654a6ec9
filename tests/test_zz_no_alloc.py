"""GPU: the allocation-free entry point trc_sort_pairs_static (work area in a static device array of the
library) gives exactly what trc_sort_pairs gives, on two calls in a row that share the array.  Collected last
on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sort(lib, L, torch, keys, n_keys, static):
    n = int(keys.shape[0])
    dev = keys.device
    ko = torch.full((n,), -7, dtype=torch.int32, device=dev)
    vo = torch.full((n,), -7, dtype=torch.int32, device=dev)
    count = torch.zeros(1, dtype=torch.int64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    if static:
        L.check(lib.trc_sort_pairs_static(n, n_keys, L.dptr(keys), None, L.dptr(ko), L.dptr(vo), L.dptr(count), st))
    else:
        work = torch.empty(int(lib.trc_sort_work_bytes(n, n_keys)), dtype=torch.uint8, device=dev)
        L.check(lib.trc_sort_pairs(n, n_keys, L.dptr(keys), None, L.dptr(ko), L.dptr(vo), L.dptr(count), L.dptr(work), st))
    torch.cuda.synchronize()
    c = int(count.item())
    return c, ko[:c].cpu().numpy(), vo[:c].cpu().numpy()


def test_static_scratch_sort_equals_sort_pairs():
    import torch
    from tracer_b200 import _lib as L
    lib = L.load()
    dev = torch.device('cuda', torch.cuda.current_device())
    rng = np.random.RandomState(11)
    # two calls in a row: a large one, then a smaller one with another key range (sees the first one's leftovers)
    for n, n_keys in ((1500007, 219), (40001, 2)):
        keys_h = rng.randint(-1, n_keys, size=n).astype(np.int32)          # -1 = dropped (a ray that hit nothing)
        keys = torch.from_numpy(keys_h).to(dev)
        c0, k0, v0 = _sort(lib, L, torch, keys, n_keys, static=False)
        c1, k1, v1 = _sort(lib, L, torch, keys, n_keys, static=True)
        assert c0 == c1 == int((keys_h >= 0).sum())
        assert np.array_equal(k0, k1) and np.array_equal(v0, v1)          # integer work: bit-exact
        order = np.argsort(keys_h[keys_h >= 0], kind='stable')             # and both are THE stable sort
        assert np.array_equal(v1, np.nonzero(keys_h >= 0)[0][order])
    # beyond the stated maxima: a clear error, nothing launched
    with pytest.raises(MemoryError, match='TRC_SORT_STATIC_MAX_PAIRS'):
        L.check(lib.trc_sort_pairs_static(4194305, 4, None, None, None, None, None, None))
    with pytest.raises(MemoryError, match='TRC_SORT_STATIC_MAX_KEYS'):
        L.check(lib.trc_sort_pairs_static(10, 257, None, None, None, None, None, None))
