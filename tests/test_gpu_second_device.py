"""GPU (needs two devices, skipped otherwise): an engine built for cuda:1 while cuda:0 is the current device compiles its
scene tables, BVH and shadow grid on cuda:1, draws its device source there and traces there -- same counts, parents
and tallies as the same trace on cuda:0 (Philox is keyed by the source ray, not by the device)."""
import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')


def _trace(device, n=300000):
    import torch
    from tracer_b200.tracer_engine import TracerEngine
    from tests.test_gpu_shadow_grid import _nsttf_source
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    with torch.cuda.device(device):
        src, _ = _nsttf_source(n, 11)          # device sources draw on the current device
    eng = TracerEngine(asm, rng='philox', seed=5, device=device, device_accel='bvh')
    eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=True)
    assert eng.absorbed_per_surface.device == torch.device(device)
    return (eng.stats['per_bounce'], eng.hits_per_surface.cpu().numpy(), eng.absorbed_per_surface.cpu().numpy(),
            eng.tree[1].get_parents().copy(), eng.tree[1].get_vertices().copy())


def test_engine_on_a_device_that_is_not_current():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    torch.cuda.set_device(0)
    a = _trace('cuda:0')
    b = _trace('cuda:1')
    assert torch.cuda.current_device() == 0
    assert a[0] == b[0] and a[0][0][1] > 1000
    assert np.array_equal(a[1], b[1])
    assert np.array_equal(a[3], b[3])
    assert np.array_equal(a[4], b[4])
    assert np.allclose(a[2], b[2], rtol=1e-12, atol=0.)
    # and brute force on the second device (small scene, no tree)
    from tracer_b200.tracer_engine import TracerEngine
    asm, source = scenes.c1_flat_mirror(T, n_rays=20000)
    eng = TracerEngine(asm, rng='philox', seed=1, device='cuda:1')
    eng.ray_tracer(source(), reps=3, min_energy=1e-10, tree=True)
    assert eng.stats['per_bounce'][0][1] > 0
