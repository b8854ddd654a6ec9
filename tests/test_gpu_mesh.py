"""GPU: configuration 4 (STL triangle-mesh cavity, Lambertian multi-bounce) through the mesh
object of tracer_b200/stl_utils.py.

  * a small mesh traced as ONE TriangleMeshObject equals the same faces traced as one Surface
    per face (hit counts per face, parents, tallies, per-face accountant data);
  * at FULL size (1 001 000 faces, BASELINE.json configs[3]) the CPU oracle cannot run, so:
    tree walk = brute force on the hit face of 1e5 rays (source rays and Lambertian children),
    energy bookkeeping of a 2e7-ray, 6-bounce trace in closed form (every child of bounce k
    carries e0 * 0.3^k), and the apparent absorptance against the same cavity made of the
    quadric surfaces FiniteCylinder + RoundPlateGM, within Monte-Carlo error.
"""
import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')


def _engine(asm, accel='bvh', seed=9):
    from tracer_b200.tracer_engine import TracerEngine
    return TracerEngine(asm, rng='philox', seed=seed, device_accel=accel)


def test_mesh_object_traces_like_one_surface_per_face():
    out = []
    np.random.seed(11)
    bundle = None
    for per_face in (False, True):
        asm, source = scenes.c4_cavity_mesh(T, n_rays=50000, n_theta=40, n_z=10, per_face_objects=per_face)
        if bundle is None:
            bundle = source()
        eng = _engine(asm)
        eng.ray_tracer(bundle, reps=4, min_energy=1e-6, tree=True)
        out.append((asm, eng))
    (asm_m, m), (asm_o, o) = out
    assert m.stats['per_bounce'] == o.stats['per_bounce']
    assert len(m.stats['per_bounce']) == 4 and m.stats['per_bounce'][3][1] > 1000
    assert np.array_equal(m.hits_per_surface.cpu().numpy(), o.hits_per_surface.cpu().numpy())
    np.testing.assert_allclose(m.absorbed_per_surface.cpu().numpy(), o.absorbed_per_surface.cpu().numpy(), rtol=1e-12)
    for k in (1, 2, 3):
        assert np.array_equal(m.tree[k].get_parents(), o.tree[k].get_parents())
        np.testing.assert_allclose(m.tree[k].get_vertices(), o.tree[k].get_vertices(), atol=1e-12)
    mesh = asm_m.get_objects()[0]
    sm, so = asm_m.get_surfaces(), asm_o.get_surfaces()
    busiest = int(np.argmax(m.hits_per_surface.cpu().numpy()))
    for i in (0, busiest, len(sm) - 1):
        a_m, h_m = sm[i].get_optics_manager().get_all_hits()
        a_o, h_o = so[i].get_optics_manager().get_all_hits()
        np.testing.assert_allclose(a_m, a_o, rtol=1e-12)
        np.testing.assert_allclose(h_m, h_o, atol=1e-12)
    absorbed, hits = mesh.get_all_hits()
    assert hits.shape == (3, len(absorbed)) and len(absorbed) == sum(h for _, h, _ in m.stats['per_bounce'])
    np.testing.assert_allclose(mesh.absorbed_per_face(), m.absorbed_per_surface.cpu().numpy(), rtol=1e-10)
    asm_m.reset_all_optics()
    assert len(mesh.get_all_hits()[0]) == 0


N_THETA, N_Z = 1000, 500          # 1 001 000 faces


@pytest.fixture(scope='module')
def full_mesh():
    asm, _ = scenes.c4_cavity_mesh(T, n_theta=N_THETA, n_z=N_Z)
    return asm


def _disk(n, seed):
    return T.sources.disk_bundle_device(n, np.c_[[0., 0., -0.01]], np.r_[0., 0., 1.], 0.9, np.pi / 4., flux=1000.,
                                        seed=seed)


def test_full_size_walk_equals_brute(full_mesh):
    import torch
    src = _disk(100000, 3)
    res = []
    for accel in ('bvh', 'brute'):
        full_mesh.reset_all_optics()
        eng = _engine(full_mesh, accel)
        eng.ray_tracer(src, reps=2, min_energy=1e-9, tree=True)
        res.append((eng.stats['per_bounce'], eng.hits_per_surface.cpu().numpy(), eng.tree[1].get_parents().copy(),
                    eng.tree[2].get_parents().copy()))
        torch.cuda.synchronize()
    assert res[0][0] == res[1][0]
    assert np.array_equal(res[0][1], res[1][1])        # per-face hit counts of both bounces
    assert np.array_equal(res[0][2], res[1][2]) and np.array_equal(res[0][3], res[1][3])
    assert res[0][0][0][1] == 100000                   # every source ray enters the cavity and lands on a face


def test_full_size_energy_bookkeeping_and_smooth_cavity(full_mesh):
    n, reps, alpha = 20000000, 6, 0.7
    full_mesh.reset_all_optics()
    src = _disk(n, 5)
    e0 = float(src._props['energy'].pieces[0][0].item())
    eng = _engine(full_mesh)
    eng.kernel_events = []
    eng.ray_tracer(src, reps=reps, min_energy=1e-12, tree=False)
    pb = eng.stats['per_bounce']
    assert len(pb) == reps and pb[0][0] == n
    for k in range(1, reps):
        assert pb[k][0] == pb[k - 1][2] == pb[k - 1][1]        # every hit makes one child, none is culled
    hits = eng.hits_per_surface.cpu().numpy()
    assert hits.sum() == sum(h for _, h, _ in pb)
    expect = sum(h * e0 * (1. - alpha) ** k * alpha for k, (_, h, _) in enumerate(pb))
    got = float(eng.absorbed_per_surface.sum().item())
    assert abs(got - expect) <= 1e-9 * expect
    mesh = full_mesh.get_objects()[0]
    assert mesh.get_all('absorbed').shape == (hits.sum(),)
    print('C4 nearest-hit kernel ms per bounce:', [(b, nn, h, round(ms(), 3)) for b, nn, h, ms in eng.kernel_events])

    # the same cavity from quadric surfaces
    A, O, S, C, F = T.assembly, T.object, T.surface, T.cylinder, T.flat_surface
    opt = T.optics_callables.LambertianReceiver
    wall = S.Surface(C.FiniteCylinder(2., 2.), opt(alpha), location=np.r_[0., 0., 1.])
    cap = S.Surface(F.RoundPlateGM(1.), opt(alpha), location=np.r_[0., 0., 2.])
    smooth = A.Assembly(objects=[O.AssembledObject(surfs=[wall, cap])])
    eng2 = _engine(smooth, 'brute', seed=10)
    eng2.ray_tracer(_disk(n, 6), reps=reps, min_energy=1e-12, tree=False)
    frac_mesh = got / (n * e0)
    frac_smooth = float(eng2.absorbed_per_surface.sum().item()) / (n * e0)
    assert 0.8 < frac_smooth < 1.0
    # per-ray absorbed fraction lies in [0, 1]: sigma of the mean <= 0.5 / sqrt(n); 5 sigma of the difference
    assert abs(frac_mesh - frac_smooth) < 5. * np.sqrt(2.) * 0.5 / np.sqrt(n) + 2e-5, (frac_mesh, frac_smooth)
    cap_share_mesh = hits[-N_THETA:].sum() / float(hits.sum())
    h2 = eng2.hits_per_surface.cpu().numpy()
    assert abs(cap_share_mesh - h2[1] / float(h2.sum())) < 1e-3
