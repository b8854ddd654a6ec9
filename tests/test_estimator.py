"""Estimator / MCRT_to_CI against the reference's own class, imported
from /root/reference when it is there (build container); golden numbers otherwise."""
import os

import numpy as np
import pytest

from tracer_b200.estimator import Estimator, MCRT_to_CI

# recorded from ray_trace_utils/estimator.py with the batches below (see test body)
GOLD_MEAN = np.array([2.0387096774193547, 4.27741935483871])
GOLD_CI = np.array([0.0889416226791784, 0.14550775495427481])
BATCHES = [(np.array([2., 4.]), 100), (np.array([2.2, 4.1]), 50), (np.array([1.9, 4.9]), 70), (np.array([2.1, 4.2]), 90)]


def _feed(est):
    for v, k in BATCHES:
        est.update(v.copy(), k)
    return est


def test_estimator_matches_recorded_reference_values():
    est = _feed(Estimator(3.))
    np.testing.assert_allclose(est.mean, GOLD_MEAN, rtol=1e-14)
    np.testing.assert_allclose(est.get_CI(), GOLD_CI, rtol=1e-13)
    assert np.all(np.isinf(Estimator().get_CI()))
    one = Estimator()
    one.update(np.array([1., 2.]), 10)
    assert np.all(np.isinf(one.get_CI()))          # a single batch has no spread yet


@pytest.mark.skipif(not os.path.isdir('/root/reference'), reason='reference tree only exists in the build container')
def test_estimator_equals_live_reference():
    from oracle import ref_import
    ref_import.enable()
    from ray_trace_utils.estimator import Estimator as Ref
    a, b = _feed(Estimator(3.)), _feed(Ref(3.))
    np.testing.assert_array_equal(a.mean, b.mean)
    np.testing.assert_array_equal(a.M2, b.M2)
    np.testing.assert_array_equal(a.get_CI(), b.get_CI())
    np.testing.assert_allclose(b.mean, GOLD_MEAN, rtol=1e-14)
    np.testing.assert_allclose(b.get_CI(), GOLD_CI, rtol=1e-13)


def test_mcrt_to_ci_stops_at_target():
    rng = np.random.RandomState(0)
    calls = []

    def fun(num_rays):
        calls.append(num_rays)
        return np.array([rng.normal(5., 2. / np.sqrt(num_rays))])
    est = MCRT_to_CI(fun, 0.02, 400)
    assert len(calls) >= 2 and float(est.get_CI()[0]) <= 0.02
    assert abs(est.mean[0] - 5.) < 0.2


@pytest.mark.gpu
def test_trace_until_ci_on_device_tallies():
    """The receiver power of C1 estimated to 1 % with tallies that stay on the GPU."""
    import torch
    from tests import scenes
    from tracer_b200.tracer_engine import TracerEngine
    T = scenes.NS('tracer_b200')
    asm, _ = scenes.c1_flat_mirror(T, n_rays=10)
    eng = TracerEngine(asm, rng='philox', seed=1)
    state = {'k': 0}

    def fun(num_rays):
        state['k'] += 1
        src = T.sources.disk_bundle_device(num_rays, np.c_[[0., 0., 4.]], np.r_[0., 0., -1.], 0.8, 4.65e-3,
                                           flux=1000., seed=100 + state['k'])
        asm.reset_all_optics()
        eng.ray_tracer(src, reps=3, min_energy=1e-10, tree=False)
        return eng.absorbed_per_surface[1:2].clone()          # CUDA tensor: absorbed power of the receiver
    est = MCRT_to_CI(fun, 0.01, 20000)
    assert torch.is_tensor(est.mean) and est.mean.is_cuda
    assert float(est.get_CI().max()) <= 0.01
    assert abs(float(est.mean[0]) - 1000. * np.pi * 0.8 ** 2 * 0.95) < 0.02 * 1900.
