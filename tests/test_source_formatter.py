"""Binary source files (ray_trace_utils/Source_formatter.py:166-233 of the reference): file layout,
the host helpers, and the one-copy device load feeding a trace."""
import numpy as np
import pytest

from tests import scenes

T = scenes.NS('tracer_b200')


def _rays(n, seed=4):
    rs = np.random.RandomState(seed)
    d = rs.normal(size=(3, n))
    d[2] = -np.abs(d[2]) - 0.5
    d /= np.sqrt(np.sum(d ** 2, axis=0))
    return rs.uniform(-1, 1, size=(3, n)) + np.c_[[0., 0., 3.]], d, rs.uniform(0.5, 1.5, size=n)


def test_file_layout_and_host_helpers(tmp_path):
    sf = T.source_formatter
    v, d, e = _rays(1000)
    f = str(tmp_path / 'src0')
    sf.binarize_source(v, d, e, f)
    raw = np.fromfile(f)
    assert np.array_equal(raw, np.hstack((np.ravel(v), np.ravel(d), e)))       # Source_formatter.py:170
    pos, dirs, ener = sf.load_source(f)
    assert np.array_equal(pos, v) and np.array_equal(dirs, d) and np.array_equal(ener, e)
    g = str(tmp_path / 'src1')
    v2, d2, e2 = _rays(600, seed=5)
    sf.binarize_source(v2, d2, e2, g)
    pos, dirs, ener = sf.load_sources([f, g])
    assert pos.shape == (3, 1600) and np.array_equal(pos[:, 1000:], v2)
    np.testing.assert_array_equal(ener, np.hstack((e, e2)) / 2.)
    sf.adjust_ener(f, 500.)
    assert np.allclose(sf.load_source(f)[2], 0.5) and np.array_equal(sf.load_source(f)[0], v)
    sf.push_back(g, 2.)
    np.testing.assert_allclose(sf.load_source(g)[0], v2 - 2. * d2, rtol=0, atol=1e-15)
    sf.push_to(g, 1.25)
    np.testing.assert_allclose(sf.load_source(g)[0][2], 1.25, rtol=0, atol=1e-12)
    bad = str(tmp_path / 'bad')
    np.arange(10.).tofile(bad)
    with pytest.raises(ValueError):
        sf.load_source(bad)
    b = T.ray_bundle.RayBundle(vertices=v, directions=d, energy=e)
    h = str(tmp_path / 'src2')
    sf.save_bundle(b, h)
    assert np.array_equal(np.fromfile(h), raw)


@pytest.mark.gpu
def test_device_load_feeds_the_trace(tmp_path):
    import torch
    from tracer_b200.tracer_engine import TracerEngine
    sf = T.source_formatter
    np.random.seed(31)
    asm, source = scenes.c1_flat_mirror(T, n_rays=50000)
    host = source()
    f = str(tmp_path / 'c1_source')
    sf.save_bundle(host, f)
    dev_bundle = sf.load_source_bundle(f)
    assert isinstance(dev_bundle._props['vertices'].pieces[0], torch.Tensor)
    assert dev_bundle._props['vertices'].pieces[0].is_cuda and dev_bundle.get_num_rays() == 50000
    a = TracerEngine(asm, rng='philox', seed=9)
    a.ray_tracer(dev_bundle, reps=3, min_energy=1e-10)
    asm.reset_all_optics()
    b = TracerEngine(asm, rng='philox', seed=9)
    b.ray_tracer(host, reps=3, min_energy=1e-10)
    assert a.tree.num_bunds() == b.tree.num_bunds()
    for k in range(1, a.tree.num_bunds()):
        assert np.array_equal(a.tree[k].get_parents(), b.tree[k].get_parents())
        assert np.array_equal(a.tree[k].get_vertices(), b.tree[k].get_vertices())
        assert np.array_equal(a.tree[k].get_energy(), b.tree[k].get_energy())
    # a device bundle goes back to a file unchanged
    g = str(tmp_path / 'again')
    sf.save_bundle(dev_bundle, g)
    assert np.array_equal(np.fromfile(g), np.fromfile(f))
    two = sf.load_sources_bundle([f, g])
    assert two.get_num_rays() == 100000
    np.testing.assert_allclose(two.get_energy().sum(), host.get_energy().sum(), rtol=1e-12)
