"""GPU: polychromatic rays and BiFacial managers with device draws (Philox).  The bundle-by-bundle comparison
with the reference is tests/test_gpu_parity.py (golden traces c10 / c11, injected draws); here the production mode
is held against it statistically and against the closed forms the optics imply."""
import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu

T = scenes.NS('tracer_b200')


def _absorbed(builder, rng, n_rays, reps, **kw):
    from tracer_b200.tracer_engine import TracerEngine
    np.random.seed(77)
    asm, source = builder(T, n_rays=n_rays, **kw)
    bundle = source()
    eng = TracerEngine(asm, rng=rng, seed=5)
    eng.ray_tracer(bundle, reps=reps, min_energy=1e-9, tree=True)
    return eng, asm, bundle


@pytest.mark.parametrize('builder,reps', [(scenes.c11_bifacial, 4), (scenes.c10_polychromatic, 3)])
def test_philox_agrees_with_injected_draws(builder, reps):
    """Per-surface absorbed power with device draws within 3 % (+ 0.2 % of the total) of the same trace with
    numpy.random draws injected in the reference's order: two independent samples of 2e5 rays."""
    a, _, _ = _absorbed(builder, 'numpy', 200000, reps)
    b, _, _ = _absorbed(builder, 'philox', 200000, reps)
    pa, pb = a.absorbed_per_surface.cpu().numpy(), b.absorbed_per_surface.cpu().numpy()
    assert np.array_equal(a.stats['per_bounce'][0][:2], b.stats['per_bounce'][0][:2])       # first bounce: no draws yet
    assert np.all(np.abs(pa - pb) <= 0.03 * np.abs(pa) + 2e-3 * np.abs(pa).sum()), (pa, pb)


def test_polychromatic_energy_is_the_integral_of_the_spectrum():
    """After a polychromatic surface the ray energy is trapz(spectra, wavelengths) (optics_callables.py:413), grey
    surfaces scale spectrum and energy alike (:137, :173), and the PolychromaticAccountant's lists add up."""
    eng, asm, bundle = _absorbed(scenes.c10_polychromatic, 'philox', 50000, 2)
    child = eng.tree[1]
    wl, sp, e = child.get_wavelengths(), child.get_spectra(), child.get_energy()
    assert wl.shape == sp.shape == (bundle.get_wavelengths().shape[0], child.get_num_rays())
    parents = child.get_parents()
    assert np.array_equal(wl, bundle.get_wavelengths()[:, parents])
    integral = np.sum(np.diff(wl, axis=0) * (sp[1:] + sp[:-1]) / 2., axis=0)
    hits = eng.hits_per_surface.cpu().numpy()
    assert hits[0] > 1000
    floor = asm.get_surfaces()[0].get_optics_manager()
    absorber, polyacc, loc = floor.accountants
    k = np.hstack(polyacc._wavelengths).shape[1]
    assert k > 1000 and np.hstack(polyacc._spectra).shape == (wl.shape[0], k)
    # the floor's children of bounce 1 are the first len(chunk 0) rays of the child bundle (pre-cull order = tree order
    # here: nothing is culled at 1e-9)
    k1 = np.asarray(polyacc._wavelengths[0]).shape[1]
    np.testing.assert_allclose(e[:k1], integral[:k1], rtol=1e-12)
    lost = np.asarray(polyacc._spectra[0])
    np.testing.assert_allclose(lost, bundle.get_spectra()[:, parents[:k1]] - sp[:, :k1], rtol=1e-12, atol=1e-9)
    np.testing.assert_allclose(np.asarray(absorber._absorbed[0]), bundle.get_energy()[parents[:k1]] - e[:k1], rtol=1e-12)


def test_bifacial_sides_see_every_hit_and_children_are_back_first():
    eng, asm, bundle = _absorbed(scenes.c11_bifacial, 'philox', 100000, 2)
    two_a = asm.get_surfaces()[0].get_optics_manager()
    front, back = two_a.get_all_hits()
    # both sides' Receiver accountants (absorbed, hits) saw the same hit points whenever both sides were hit
    assert front[1].shape[1] > 100 and back[1].shape[1] > 100
    # surface 2's managers have no accountants: ([], []) as in the reference
    assert asm.get_surfaces()[2].get_optics_manager().get_all_hits() == ([], [])
    # children of surface 0 at bounce 1: back side (rays travelling along +z of the plate) first
    child = eng.tree[1]
    k = np.asarray(two_a.OpticsCallable_front.accountants[0]._absorbed[0]).shape[0]
    parents = child.get_parents()[:k]
    d = bundle.get_directions()[:, parents]
    frame = np.asarray(asm.get_surfaces()[0]._temp_frame)
    is_back = np.around(np.sum(d * frame[:3, 2][:, None], axis=0), decimals=6) > 0.
    n_back = int(is_back.sum())
    assert 0 < n_back < k
    assert is_back[:n_back].all() and not is_back[n_back:].any()
    assert np.all(np.diff(parents[:n_back]) > 0) and np.all(np.diff(parents[n_back:]) > 0)
