"""GPU: a large HOST bundle is uploaded in slices and traced slice by slice
(trc_trace_hits_slice) while the rest is still on the wire.  The result must be
identical to the same trace with the whole bundle uploaded first."""
import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')


def _trace(asm, host_bundle, streamed):
    import torch
    from tracer_b200 import tracer_engine as TE
    old = TE.STREAM_MIN_RAYS
    TE.STREAM_MIN_RAYS = 1 << 16 if streamed else 1 << 62
    try:
        asm.reset_all_optics()
        eng = TE.TracerEngine(asm, rng='philox', seed=11, device_accel='bvh')
        eng.ray_tracer(host_bundle, reps=3, min_energy=1e-10, tree=True)
        torch.cuda.synchronize()
        out = []
        for k in range(1, len(eng.tree._bunds)):
            b = eng.tree[k]
            out.append((b.get_vertices().copy(), b.get_directions().copy(), b.get_energy().copy(),
                        b.get_parents().copy()))
        return out, eng.absorbed_per_surface.cpu().numpy(), eng.hits_per_surface.cpu().numpy()
    finally:
        TE.STREAM_MIN_RAYS = old


@pytest.mark.parametrize('n', [300000, 1000003])
def test_streamed_upload_equals_whole_upload(n):
    import torch
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    pos = scenes.nsttf_positions()
    sun = scenes.sun_vector(0., np.radians(35.05))
    lo, hi = pos.min(axis=0), pos.max(axis=0)
    centre = 0.5 * (lo + hi)
    radius = 1.1 * 0.5 * np.sqrt(np.sum((hi[:2] - lo[:2]) ** 2))
    src = T.sources.buie_sunshape_device(n, np.vstack(300. * sun + centre), -sun, radius, 0.01, flux=1000.,
                                         pre_process_CSR=False, seed=21)
    hV = torch.empty((3, n), dtype=torch.float64, pin_memory=True)
    hD = torch.empty((3, n), dtype=torch.float64, pin_memory=True)
    hE = torch.empty(n, dtype=torch.float64, pin_memory=True)
    hV.copy_(torch.from_numpy(src.get_vertices()))
    hD.copy_(torch.from_numpy(src.get_directions()))
    hE.copy_(torch.from_numpy(src.get_energy()))
    host = T.ray_bundle.RayBundle(vertices=hV.numpy(), directions=hD.numpy(), energy=hE.numpy())
    whole, a0, h0 = _trace(asm, host, streamed=False)
    sliced, a1, h1 = _trace(asm, host, streamed=True)
    assert len(whole) == len(sliced) >= 2
    for w, s in zip(whole, sliced):
        for x, y in zip(w, s):
            assert np.array_equal(x, y)
    assert np.array_equal(h0, h1)
    np.testing.assert_allclose(a0, a1, rtol=1e-12)      # float64 atomics: summation order differs
    # pageable (not pinned) host arrays take the same path
    host2 = T.ray_bundle.RayBundle(vertices=hV.numpy().copy(), directions=hD.numpy().copy(), energy=hE.numpy().copy())
    sliced2, _, h2 = _trace(asm, host2, streamed=True)
    assert np.array_equal(h0, h2)
    for w, s in zip(whole, sliced2):
        assert np.array_equal(w[3], s[3])
