"""GPU: the float32-pruned tree walk must return exactly the hit indices of the
brute-force float64 kernel (both through the C ABI) -- on the NSTTF field at
1e6 rays, on random triangle soups with axis-parallel rays, rays starting inside
boxes and far-away origins, and with unbounded surfaces mixed in.
Bar: bit-exact int32 hit indices (the accelerator only prunes)."""
import ctypes as C

import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')


def _hits(scene, V, D, use_bvh):
    import torch
    from tracer_b200 import _lib as L
    lib = L.load()
    dev = torch.device('cuda', 0)
    tv = torch.from_numpy(np.ascontiguousarray(V)).to(dev)
    td = torch.from_numpy(np.ascontiguousarray(D)).to(dev)
    n = tv.shape[1]
    hit = torch.full((n,), -7, dtype=torch.int32, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L.check(lib.trc_nearest_hit(scene.handle, n, L.dptr(tv), L.dptr(td), tv.stride(0), use_bvh, L.dptr(hit), st))
    torch.cuda.synchronize()
    return hit.cpu().numpy()


def _compare(asm, V, D, leaf_size=2):
    from tracer_b200.scene import CompiledScene
    sc = CompiledScene(asm, leaf_size=leaf_size)
    a = _hits(sc, V, D, 0)
    b = _hits(sc, V, D, 1)
    assert a.min() >= -1
    bad = np.nonzero(a != b)[0]
    assert bad.size == 0, 'walk != brute at rays %s: %s vs %s' % (bad[:5], b[bad[:5]], a[bad[:5]])
    sc.close()
    return a


def _triangle_soup(n_tri, rng, extent=10., with_plane=False):
    objs = []
    opt = T.optics_callables.Reflective(0.1)
    for _ in range(n_tri):
        verts = rng.uniform(-0.6, 0.6, size=(3, 2))
        verts[2] = 0.                      # the triangle lies in its local XY plane (triangular_face.py:12-27)
        tr = T.spatial_geometry.translate(*rng.uniform(-extent, extent, size=3))
        axis = rng.normal(size=3)
        rot = T.spatial_geometry.general_axis_rotation(axis / np.linalg.norm(axis), rng.uniform(0, 2 * np.pi))
        frame = tr.copy()
        frame[:3, :3] = rot[:3, :3]
        objs.append(T.object.AssembledObject(
            surfs=[T.surface.Surface(T.triangular_face.TriangularFace(verts), opt)], transform=frame))
    if with_plane:
        objs.append(T.object.AssembledObject(
            surfs=[T.surface.Surface(T.flat_surface.FlatGeometryManager(), opt)],
            transform=T.spatial_geometry.translate(0, 0, -extent - 1.)))
    return T.assembly.Assembly(objects=objs)


def _random_rays(n, rng, extent):
    V = rng.uniform(-1.5 * extent, 1.5 * extent, size=(3, n))
    D = rng.normal(size=(3, n))
    # axis-parallel and plane-parallel directions (1/d = inf in the slab test)
    k = n // 8
    D[:, :k] = 0.
    D[rng.randint(0, 3, size=k), np.arange(k)] = rng.choice([-1., 1.], size=k)
    D[0, k:2 * k] = 0.
    D[1, 2 * k:3 * k] = 1e-300
    D /= np.sqrt(np.sum(D ** 2, axis=0))
    return V, D


@pytest.mark.parametrize('leaf', [1, 2, 4])
def test_triangle_soup_walk_equals_brute(leaf):
    rng = np.random.RandomState(100 + leaf)
    asm = _triangle_soup(400, rng)
    V, D = _random_rays(300000, rng, 10.)
    hits = _compare(asm, V, D, leaf_size=leaf)
    assert (hits >= 0).sum() > 1000


def test_walk_with_unbounded_surface_and_far_origins():
    rng = np.random.RandomState(7)
    asm = _triangle_soup(200, rng, with_plane=True)
    V, D = _random_rays(100000, rng, 10.)
    V[:, :20000] *= 1e4          # origins 100 km away: the per-ray pad must follow |v|
    D[:, :20000] = -V[:, :20000] + rng.uniform(-10, 10, size=(3, 20000))
    D /= np.sqrt(np.sum(D ** 2, axis=0))
    hits = _compare(asm, V, D)
    assert (hits == 200).sum() > 1000          # the infinite plane catches many rays


def test_rays_starting_on_surfaces_and_small_bundles():
    rng = np.random.RandomState(8)
    asm = _triangle_soup(64, rng, extent=3.)
    for n in (1, 31, 32, 33, 255, 256, 257, 1000):
        V, D = _random_rays(max(n, 8), rng, 3.)
        _compare(asm, V[:, :n], D[:, :n])


def test_nsttf_walk_equals_brute_1e6():
    import torch
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    pos = scenes.nsttf_positions()
    sun = scenes.sun_vector(0., np.radians(35.05))
    lo, hi = pos.min(axis=0), pos.max(axis=0)
    centre = 0.5 * (lo + hi)
    radius = 1.1 * 0.5 * np.sqrt(np.sum((hi[:2] - lo[:2]) ** 2))
    src = T.sources.buie_sunshape_device(1000000, np.vstack(300. * sun + centre), -sun, radius, 0.01, flux=1000.,
                                         pre_process_CSR=False, seed=5)
    V = src.get_vertices()
    D = src.get_directions()
    hits = _compare(asm, V, D)
    frac = (hits >= 0).mean()
    assert 0.04 < frac < 0.09, frac
    # second bounce geometry: rays leaving the heliostats towards the tower
    from tracer_b200.tracer_engine import TracerEngine
    eng = TracerEngine(asm, rng='philox', seed=3, device_accel='brute')
    eng.ray_tracer(src, reps=1, tree=True)
    b1 = eng.tree[1]
    _compare(asm, b1.get_vertices(), b1.get_directions())
    torch.cuda.synchronize()


def test_triangle_cavity_20k_faces_walk_equals_brute_multi_bounce():
    """Config 4 at reduced scale (2e4 triangular faces: the tree no longer fits the
    shared-memory cache, leaves come from global memory): identical per-bounce bundle
    sizes and per-surface hit counts with and without the tree over 5 Lambertian bounces."""
    from tracer_b200.tracer_engine import TracerEngine
    np.random.seed(4)
    asm, source = scenes.c4_cavity(T, n_rays=200000, n_theta=200, n_z=50)
    bundle = source()
    out = []
    for accel in ('bvh', 'brute'):
        asm.reset_all_optics()
        eng = TracerEngine(asm, rng='philox', seed=9, device_accel=accel)
        eng.ray_tracer(bundle, reps=5, min_energy=1e-6, tree=True)
        out.append((eng.stats['per_bounce'], eng.hits_per_surface.cpu().numpy(),
                    eng.absorbed_per_surface.cpu().numpy(), eng.tree[3].get_parents().copy()))
    assert out[0][0] == out[1][0]
    assert len(out[0][0]) == 5 and out[0][0][4][1] > 10000          # still bouncing after 5 iterations
    assert np.array_equal(out[0][1], out[1][1])
    np.testing.assert_allclose(out[0][2], out[1][2], rtol=1e-12)
    assert np.array_equal(out[0][3], out[1][3])


def test_prebuilt_kdtree_handle_selects_the_walk():
    """ray_tracer(accel=True, Kd_Tree=KdTree(assembly)) as in the reference (tracer_engine.py:171-185)."""
    from tracer_b200.tracer_engine import TracerEngine
    np.random.seed(8)
    asm, source = scenes.c3_nsttf(T, n_rays=20000, n_hstats=40, dense=True)
    bundle = source()
    kd = T.accel_tree.KdTree(asm)
    a = TracerEngine(asm, rng='philox', seed=4, device_accel='auto')
    a.ray_tracer(bundle, reps=3, min_energy=1e-10, accel=True, Kd_Tree=kd)
    assert a.Kd_Tree is kd and a.stats['use_bvh']
    asm.reset_all_optics()
    b = TracerEngine(asm, rng='philox', seed=4, device_accel='brute')
    b.ray_tracer(bundle, reps=3, min_energy=1e-10)
    assert not b.stats['use_bvh'] and a.stats['per_bounce'] == b.stats['per_bounce']
    for k in range(1, a.tree.num_bunds()):
        assert np.array_equal(a.tree[k].get_parents(), b.tree[k].get_parents())
        assert np.array_equal(a.tree[k].get_vertices(), b.tree[k].get_vertices())


def test_reflected_bundle_walk_equals_brute():
    """The second bounce of a heliostat field: child bundles are surface-major, so consecutive rays leave one mirror
    towards the receiver, every ray starts ON a surface (its own mirror is a candidate that must lose) and nearly all
    reach the receiver.  Same hit indices as brute force -- for the reflected bundle itself, with the origins moved
    back behind the mirrors (the own mirror and its neighbours become real candidates), reversed, with the order of
    half the rays scrambled and with non-finite rays mixed in."""
    from tracer_b200.scene import CompiledScene
    from tracer_b200.tracer_engine import TracerEngine
    from tests.test_gpu_shadow_grid import _nsttf_source
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    src, _ = _nsttf_source(2000000, 7)
    eng = TracerEngine(asm, rng='philox', seed=3, device_accel='bvh')
    eng.ray_tracer(src, reps=1, min_energy=1e-10, tree=True)
    b = eng.tree[1]
    V, D = np.ascontiguousarray(b.get_vertices()), np.ascontiguousarray(b.get_directions())
    n = V.shape[1]
    assert n > 100000
    sc = CompiledScene(asm)

    def same(V, D, frac_lo=None):
        brute, walk = _hits(sc, V, D, 0), _hits(sc, V, D, 1)
        bad = np.nonzero(walk != brute)[0]
        assert bad.size == 0, 'walk != brute at rays %s: %s vs %s' % (bad[:5], walk[bad[:5]], brute[bad[:5]])
        if frac_lo is not None:
            assert (brute >= 0).mean() > frac_lo
        return brute
    same(V, D, 0.9)                                   # nearly all reach the receiver
    back = same(V - 30. * D, D, 0.9)                  # from 30 m behind: the own mirror (or a neighbour) comes first
    assert (back < 218).mean() > 0.5
    same(V, -D)                                       # away from the tower
    rng = np.random.RandomState(4)
    perm = np.arange(n)
    half = rng.permutation(n // 2)
    perm[:n // 2] = half
    same(V[:, perm], D[:, perm], 0.9)
    V2, D2 = V.copy(), D.copy()
    V2[0, 1000::7777] = np.nan
    D2[2, 5000::9999] = np.inf
    same(V2, D2)
    # a short bundle whose last batch is partial and odd
    same(V[:, :64 * 37 + 13], D[:, :64 * 37 + 13])
    sc.close()
