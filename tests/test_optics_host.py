"""Host optics laws (tracer_b200/optics.py) against the known answers of the reference's
tests/test_optics.py and, in the build container, against the live reference functions."""
import math
import os

import numpy as np
import pytest

from tracer_b200 import optics

HAVE_REF = os.path.isdir('/root/reference/tracer')


def test_reflections_known_answers():
    # tests/test_optics.py:13-56
    d = np.array([[0, 1, -1]]).T / math.sqrt(2)
    assert (optics.reflections(d, np.array([[0, 0, 1]]).T) == np.array([[0, 1, 1]]).T / math.sqrt(2)).all()
    d4 = np.array([[1, 1, -1], [-1, 1, -1], [-1, -1, -1], [1, -1, -1]]).T / math.sqrt(3)
    want = np.array([[1, 1, 1], [-1, 1, 1], [-1, -1, 1], [1, -1, 1]]).T / math.sqrt(3)
    assert (optics.reflections(d4, np.array([[0, 0, 1]]).T) == want).all()
    tangent = np.array([[1, 1, 0], [-1, 1, 0], [-1, -1, 0], [1, -1, 0]]).T / math.sqrt(2)
    assert np.allclose(optics.reflections(tangent, np.array([[0, 0, 1]]).T), tangent)
    normals = np.array([[0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]]).T / np.sqrt([1, 2, 3, 2])
    assert np.allclose(optics.reflections(d4, normals), np.tile([1, 1, 1], (4, 1)).T / math.sqrt(3))


def test_fresnel_known_answers():
    # tests/test_optics.py:58-83 (Fowles, Introduction to Modern Optics, p. 44-45)
    z = np.c_[[0, 0, 1.]]
    assert optics.fresnel(z, z, np.r_[1.], np.r_[1.5])[0] == pytest.approx(0.04)
    assert optics.fresnel(z, np.c_[[0, 1., 0]], np.r_[1.], np.r_[1.5])[0] == pytest.approx(1.)
    assert optics.fresnel(np.c_[[0, 1, 1.]] / math.sqrt(2), np.c_[[0, 1., 0]], np.r_[1.], np.r_[1.])[0] == pytest.approx(0.)


def test_refraction_known_answers():
    # tests/test_optics.py:85-131
    d = np.c_[[0, -1, 1], [0, -1, -1]] / math.sqrt(2)
    for normal in (np.c_[[0, 1, 0]], np.c_[[0, -1, 0]]):
        refr, dirs = optics.refractions(np.r_[1., 1.], np.r_[1., 1.5], d, normal)
        assert refr.all()
        np.testing.assert_array_almost_equal(dirs, np.c_[d[:, 0], -np.sqrt([0, 7. / 9, 2. / 9])])
    refr, dirs = optics.refractions(1., 1.5, d, np.c_[[0, -1, 0]])
    assert refr.all()
    np.testing.assert_array_almost_equal(
        dirs, np.c_[np.sqrt([0, 7. / 9, 2. / 9]) * np.r_[0, -1, 1], -np.sqrt([0, 7. / 9, 2. / 9])])
    refr, dirs = optics.refractions(1.5, 1., np.c_[[0, -1, 0], np.r_[0, -1, -1] / math.sqrt(2)], np.c_[[0, -1, 0]])
    np.testing.assert_array_equal(refr, np.r_[True, False])
    np.testing.assert_array_equal(dirs, np.c_[[0, -1, 0]])


@pytest.mark.skipif(not HAVE_REF, reason='reference tree only exists in the build container')
def test_equal_to_live_reference():
    from oracle import ref_import
    ref_import.enable()
    from tracer import optics as ref
    rs = np.random.RandomState(2)
    n = 500
    d = rs.normal(size=(3, n))
    d /= np.sqrt(np.sum(d ** 2, axis=0))
    nrm = rs.normal(size=(3, n))
    nrm /= np.sqrt(np.sum(nrm ** 2, axis=0))
    n1, n2 = rs.uniform(1., 1.6, n), rs.uniform(1., 1.6, n)
    assert np.array_equal(optics.reflections(d, nrm), ref.reflections(d, nrm))
    with np.errstate(invalid='ignore'):
        assert np.array_equal(optics.fresnel(d, nrm, n1, n2), ref.fresnel(d, nrm, n1, n2), equal_nan=True)
    a, b = optics.refractions(n1, n2, d, nrm), ref.refractions(n1, n2, d, nrm)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    m2 = rs.uniform(0.2, 3., n) + 1j * rs.uniform(0.5, 6., n)
    th = rs.uniform(0., 1.5, n)
    for x, y in zip(optics.fresnel_to_attenuating(1.2, m2, th), ref.fresnel_to_attenuating(1.2, m2, th)):
        assert np.array_equal(x, y)
    for x, y in zip(optics.fresnel_conductor(d, nrm, None, None, 1.1, m2), ref.fresnel_conductor(d, nrm, None, None, 1.1, m2)):
        assert np.array_equal(x, y)
    wl = rs.uniform(0.3, 2., n)
    assert np.array_equal(optics.refr_idx_hartmann(wl, 1.5, .01, .1, .02, .2), ref.refr_idx_hartmann(wl, 1.5, .01, .1, .02, .2))
    L = rs.uniform(0., 2., n)
    assert np.array_equal(optics.attenuations(L, 1e-6, wl * 1e-6, n1), ref.attenuations(L, 1e-6, wl * 1e-6, n1))
