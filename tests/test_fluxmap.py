"""Flux maps: the meshes of the reference's get_fluxmap methods
(rectangle, annulus/disk, dish, cylinder sector, sphere) against golden vectors
recorded from the reference (tests/golden/make_golden.py:fluxmap_reference).

CPU: mesh edges/areas of tracer_b200.fluxmap._mesh + numpy.histogram2d.
GPU: trc_fluxmap_binned through tracer_b200.fluxmap.binned_energy."""
import os

import numpy as np
import pytest

from tests import scenes

T = scenes.NS('tracer_b200')
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'fluxmaps.npz'))
CASES = sorted(scenes.fluxmap_cases(T))


def _local_uv(coord, param, local):
    from tracer_b200 import _lib as L
    x, y, z = local
    phi = np.arctan2(y, x)
    phi[phi < 0.] += 2. * np.pi
    if coord == L.FLUX_XY:
        return x, y
    if coord == L.FLUX_R_ATAN_XY:
        a = np.arctan2(x, y)
        a[a < 0.] += 2. * np.pi
        return np.sqrt(x ** 2 + y ** 2), a
    if coord == L.FLUX_R_ATAN_YX:
        return np.sqrt(x ** 2 + y ** 2), phi
    if coord == L.FLUX_Z_PHI:
        return z, phi
    return np.arccos(z / param), phi


@pytest.mark.parametrize('name', CASES)
def test_mesh_and_areas_match_reference_cpu(name):
    from tracer_b200.fluxmap import _mesh_regions
    gm, frame, res, _ = scenes.fluxmap_cases(T)[name]
    proj = np.round(np.linalg.inv(frame), decimals=9)
    hits = GOLD[name + '_hits']
    local = np.dot(proj, np.vstack([hits, np.ones(hits.shape[1])]))[:3]
    maps = []
    for coord, param, eu, ev, areas in _mesh_regions(gm, res):
        u, v = _local_uv(coord, param, local)
        en = np.histogram2d(u, v, bins=[eu, ev], weights=GOLD[name + '_eners'])[0]
        with np.errstate(divide='ignore', invalid='ignore'):
            maps.append(np.hstack(en / areas))
    flux = maps[0] if len(maps) == 1 else np.array(maps)        # a multi-mesh class returns one map per mesh
    np.testing.assert_allclose(flux, GOLD[name + '_flux'], rtol=1e-12, atol=0., equal_nan=True)


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_device_binning_matches_reference(name):
    import torch
    from tracer_b200.fluxmap import binned_energy
    gm, frame, res, _ = scenes.fluxmap_cases(T)[name]
    surf = T.surface.Surface(gm, T.optics_callables.Reflective(0.))
    surf._temp_frame = frame
    hits = torch.from_numpy(GOLD[name + '_hits']).cuda()
    en = torch.from_numpy(GOLD[name + '_eners']).cuda()
    from tracer_b200.fluxmap import _mesh_regions
    maps = []
    for region in range(len(_mesh_regions(gm, res))):
        bins, areas = binned_energy(surf, hits, en, res, region=region)
        with np.errstate(divide='ignore', invalid='ignore'):
            maps.append(np.hstack(bins.cpu().numpy() / areas))
    flux = maps[0] if len(maps) == 1 else np.array(maps)
    np.testing.assert_allclose(flux, GOLD[name + '_flux'], rtol=1e-11, atol=0., equal_nan=True)
    # chunked input gives the same map; energy inside the mesh is conserved
    half = hits.shape[1] // 2
    b2, _ = binned_energy(surf, [hits[:, :half].contiguous(), hits[:, half:].contiguous()], [en[:half], en[half:]], res,
                          region=region)
    np.testing.assert_allclose(b2.cpu().numpy(), bins.cpu().numpy(), rtol=1e-12)


@pytest.mark.gpu
def test_surface_fluxmap_from_accountants():
    """End to end: trace C1, then the receiver's flux map from HBM-resident accountant data
    equals numpy.histogram2d on the same data read back."""
    from tracer_b200.fluxmap import surface_fluxmap
    from tracer_b200.tracer_engine import TracerEngine
    np.random.seed(3)
    asm, source = scenes.c1_flat_mirror(T, n_rays=20000)
    eng = TracerEngine(asm, rng='philox', seed=5)
    eng.ray_tracer(source(), reps=3, min_energy=1e-10, tree=True)
    recv = asm.get_surfaces()[1]
    flux = surface_fluxmap(recv, 10)
    absorbed, hits = recv.get_optics_manager().get_all_hits()
    local = np.dot(np.round(np.linalg.inv(recv._temp_frame), decimals=9), np.vstack([hits, np.ones(hits.shape[1])]))
    edges = np.linspace(-2., 2., 11)
    ref = np.hstack(np.histogram2d(local[0], local[1], bins=[edges, edges], weights=absorbed)[0] / 0.16)
    np.testing.assert_allclose(flux, ref, rtol=1e-11)
    assert flux.sum() > 0
