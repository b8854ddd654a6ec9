"""Sources.

CPU: the host mirrors draw from numpy.random in the reference's order, so with the same seed
they reproduce the reference's bundles bit for bit (checked against the live reference in the
build container).  GPU: the Philox kernels are checked statistically against the closed-form
moments of the same distributions, for sharding reproducibility and for their geometry."""
import os

import numpy as np
import pytest

from tests import scenes

T = scenes.NS('tracer_b200')
HAVE_REF = os.path.isdir('/root/reference/tracer')


def _ref():
    from oracle import ref_import
    ref_import.enable()
    return scenes.NS('tracer')


@pytest.mark.skipif(not HAVE_REF, reason='reference tree only exists in the build container')
@pytest.mark.parametrize('case', ['disk', 'rect', 'rect_down', 'oblique', 'buie', 'rect_buie', 'vf_frustum', 'vf_cylinder'])
def test_host_sources_equal_live_reference(case):
    R = _ref()
    d = np.r_[0.3, -0.2, -1.]
    d /= np.linalg.norm(d)
    rd = np.r_[0.25, -0.2, -1.]
    rd /= np.linalg.norm(rd)
    c = np.c_[[1., 2., 30.]]

    def make(M):
        np.random.seed(5)
        s = M.sources
        if case == 'disk':
            return s.disk_bundle(500, c, d, 2.5, 0.01, flux=900., radius_in=0.5)
        if case == 'rect':
            return s.rect_bundle(500, c, d, 3., 2., 0.01, flux=900.)
        if case == 'rect_down':
            return s.rect_bundle(500, c, np.r_[0., 0., -1.], 3., 2., 0.01, flux=900.)
        if case == 'oblique':
            return s.oblique_solar_rect_bundle(500, c, d, rd, 3., 2., 0.01, flux=900., wavelength=5e-7)
        if case == 'buie':
            return s.buie_sunshape(500, c, d, 2.5, 0.05, flux=900.)
        if case == 'vf_frustum':
            return s.vf_frustum_bundle(300, 0.5, 1.2, 0.8, c, d, rays_in=False, angular_span=[0.3, 4.])
        if case == 'vf_cylinder':
            return s.vf_cylinder_bundle(300, 0.7, 1.5, c, d, flux=3., rays_in=True)
        return s.rect_buie_sunshape(500, c, d, 3., 2., 0.05, flux=900., rays_direction=rd)
    a, b = make(T), make(R)
    assert np.array_equal(a.get_vertices(), b.get_vertices())
    assert np.array_equal(a.get_directions(), b.get_directions())
    assert np.array_equal(a.get_energy(), b.get_energy())
    if case == 'oblique':
        assert np.array_equal(a.get_wavelengths(), b.get_wavelengths())


def _local(frame_normal, centre, V):
    rot = T.spatial_geometry.rotation_to_z(frame_normal)
    return np.dot(rot.T, V - centre)


@pytest.mark.gpu
def test_device_rect_sources_geometry_and_moments():
    n = 400000
    d = np.r_[0.3, -0.2, -1.]
    d /= np.linalg.norm(d)
    rd = np.r_[0.25, -0.2, -1.]
    rd /= np.linalg.norm(rd)
    c = np.c_[[1., 2., 30.]]
    for maker, normal, rays_dir, area in [
        (lambda: T.sources.rect_bundle_device(n, c, d, 3., 2., 0.01, flux=900., seed=3), d, d, 6.),
        (lambda: T.sources.oblique_solar_rect_bundle_device(n, c, d, rd, 3., 2., 0.01, flux=900., seed=4), d, rd, 6.),
        (lambda: T.sources.rect_buie_sunshape_device(n, c, d, 3., 2., 0.05, flux=900., rays_direction=rd, seed=5),
         d, rd, 6.),
    ]:
        b = maker()
        V, D, E = b.get_vertices(), b.get_directions(), b.get_energy()
        loc = _local(normal, c, V)
        assert np.abs(loc[2]).max() < 1e-9                       # on the source plane
        ext = np.sort([np.abs(loc[0]).max(), np.abs(loc[1]).max()])
        assert 0.99 < ext[0] <= 1. + 1e-12 and 1.49 < ext[1] <= 1.5 + 1e-12
        for k in range(2):                                       # uniform: mean 0, variance a^2/3
            a = np.abs(loc[k]).max()
            assert abs(loc[k].mean()) < 5e-3
            assert abs(loc[k].var() - a * a / 3.) < 5e-3
        np.testing.assert_allclose(np.sqrt(np.sum(D ** 2, axis=0)), 1., rtol=1e-12)
        cosz = np.dot(rays_dir, D)
        assert cosz.min() > np.cos(0.05) - 1e-9                  # inside the pillbox / Buie cone
        cosang = np.cos(2. * np.arcsin(0.5 * np.linalg.norm(rays_dir - normal)))
        np.testing.assert_allclose(E.sum(), 900. * area * cosang, rtol=1e-3)


@pytest.mark.gpu
def test_device_sources_are_shard_invariant():
    """The multi-GPU decomposition: shards generated with first_ray/total_rays concatenate to the
    whole bundle bit for bit."""
    n = 100000
    d = np.r_[0., 0., -1.]
    c = np.c_[[0., 0., 10.]]
    whole = T.sources.rect_buie_sunshape_device(n, c, d, 3., 2., 0.05, flux=900., seed=9)
    parts = [T.sources.rect_buie_sunshape_device(n // 4, c, d, 3., 2., 0.05, flux=900., seed=9,
                                                 first_ray=k * (n // 4), total_rays=n) for k in range(4)]
    assert np.array_equal(whole.get_vertices(), np.hstack([p.get_vertices() for p in parts]))
    assert np.array_equal(whole.get_directions(), np.hstack([p.get_directions() for p in parts]))
    assert np.array_equal(whole.get_energy(), np.hstack([p.get_energy() for p in parts]))
    # straight-down rectangle: x extent along local x (the reference's axis swap, sources.py:251-252)
    b = T.sources.rect_bundle_device(n, c, d, 3., 2., 0.01, flux=None, seed=1)
    loc = _local(d, c, b.get_vertices())
    assert 1.49 < np.abs(loc[0]).max() <= 1.5 and 0.99 < np.abs(loc[1]).max() <= 1.
    np.testing.assert_allclose(b.get_energy().sum(), 1., rtol=1e-12)
