"""Monte-Carlo view factors on the GPU.

Reference: emissive_losses/view_factors_3D.py.  There, ``RTVF`` holds the online
statistics of a view-factor matrix (weighted running mean, 3-sigma interval,
reciprocity and summation tests, :20-112) and geometry-specific subclasses
(``FONaR_RTVF``, ``Two_N_parameters_cavity_RTVF``) emit a Lambertian bundle from every
element, trace it for one bounce on black walls and bin the absorbed energy by
receiving element, pass after pass, until every row has converged.

Here ``RTVF`` keeps the same state and formulas, and ``AssemblyRTVF`` is the driver for
any assembly whose surfaces are grouped into elements: the one-bounce trace runs
through ``TracerEngine`` and the row of the matrix is read from the per-surface
absorbed-energy tallies that the bounce kernel leaves in HBM (one small device
reduction per pass; no per-hit data goes to the host)."""
import numpy as N

from .tracer_engine import TracerEngine


class RTVF(object):
    """Online statistics of a ray-traced view-factor matrix (view_factors_3D.py:20-112).

    num_rays - rays fired per bundle
    precision - interval threshold on every matrix entry and on the summation rule
    precision_rec - threshold on the reciprocity rule (defaults to ``precision``)"""

    def __init__(self, num_rays=10000, precision=0.01, precision_option='absolute', precision_rec=None):
        self.num_rays = num_rays
        self.precision = precision
        self.precision_rec = precision if precision_rec is None else precision_rec
        self.stdev = N.inf
        self.precision_option = precision_option

    def _init_stats(self, n, areas):
        self.areas = N.asarray(areas, float)
        self.VF = N.zeros((n, n))
        self.progress = N.ones((n, n), dtype=bool)
        self.VF_esperance = N.zeros((n, n))
        self.Qsum = N.zeros((n, n))
        self.stdev_VF = N.zeros((n, n))
        self.p = N.zeros(n)
        self.ray_counts = N.ones(n) * self.num_rays

    def reset_opt(self):
        for S in self.A.get_surfaces():
            if hasattr(S.get_optics_manager(), 'reset'):
                S.get_optics_manager().reset()

    def test_precision(self):
        """One pass has filled ``self.VF`` for the rows that cast ``self.ray_counts`` rays
        (``self.p`` = rays cast so far per row, this pass included).  Updates the weighted
        mean, the 3-sigma interval, the reciprocity matrix and ``self.progress`` (rows /
        entries that still need rays)."""
        r = N.vstack(self.ray_counts)
        p = N.vstack(self.p)
        p_1 = p - r
        Ai = N.ones(N.shape(self.VF_esperance)) * N.vstack(self.areas)
        with N.errstate(divide='ignore', invalid='ignore'):
            # online weighted standard deviation
            self.Qsum = self.Qsum + r * p_1 / p * (self.VF - self.VF_esperance) ** 2.
            self.stdev_VF = 3. * N.sqrt(self.Qsum / (p - 1.)) / N.sqrt(p)
            # weighted average
            self.VF_esperance = (self.VF_esperance * p_1 + self.VF * r) / p
            AiFij = self.VF_esperance * Ai
            AjFji = AiFij.T
            self.VF_reciprocity = N.abs(AiFij - AjFji)
            if self.precision_option == 'absolute':
                stdev_test = self.stdev_VF <= self.precision / 2.
                tas = self.stdev_VF * Ai
                reciprocity_test = (tas + (tas.T)) <= self.precision_rec
            elif self.precision_option == 'relative':
                rel_stdev = (self.stdev_VF / self.VF_esperance)
                rel_stdev[N.isnan(rel_stdev)] = 0.
                stdev_test = rel_stdev <= self.precision
                tas = Ai * self.stdev_VF
                rel_rec = (tas + (tas.T)) / AiFij
                rel_rec[N.isnan(rel_rec)] = 0.
                rel_rec[N.isinf(rel_rec)] = 0.
                reciprocity_precision = rel_rec <= self.precision_rec
                minimum_AF_test = AiFij < N.vstack((self.precision_rec * N.amax(AiFij, axis=1)))
                reciprocity_test = N.logical_or(reciprocity_precision, minimum_AF_test)
            else:
                raise ValueError("precision_option must be 'absolute' or 'relative'")
            summ_test = N.abs(N.sum(self.VF_esperance, axis=1) - 1.) < self.precision
        self.progress = N.logical_not(N.logical_and(summ_test, N.logical_and(stdev_test, reciprocity_test)))


class AssemblyRTVF(RTVF):
    """View-factor matrix between the elements of an assembly.

    assembly           - Assembly whose walls absorb everything (e.g. ``LambertianReceiver(1.)``)
    element_of_surface - for every surface of ``assembly.get_surfaces()`` the index of the element
                         it belongs to (-1: not part of any element, e.g. an open aperture cover)
    areas              - area of every element
    emit               - ``emit(i, num_rays) -> RayBundle`` with total energy 1 leaving element i
                         (``vf_cylinder_bundle``, ``vf_frustum_bundle``, ``disk_bundle`` ... or a
                         device source)
    min_passes         - passes every row gets before convergence may stop it (the reference's
                         ``stat_cond = 2``)"""

    def __init__(self, assembly, element_of_surface, areas, emit, num_rays=10000, precision=0.01,
                 precision_option='absolute', precision_rec=None, min_passes=2, seed=0, engine=None):
        RTVF.__init__(self, num_rays, precision, precision_option, precision_rec)
        self.A = assembly
        self.emit = emit
        self.element_of_surface = N.asarray(element_of_surface, dtype=N.int64)
        if len(self.element_of_surface) != len(assembly.get_surfaces()):
            raise ValueError("element_of_surface needs one entry per surface of the assembly")
        n = len(areas)
        if self.element_of_surface.max() >= n:
            raise ValueError("element index out of range")
        self._init_stats(n, areas)
        self.min_passes = min_passes
        self.engine = engine if engine is not None else TracerEngine(assembly, rng='philox', seed=seed)
        self.passes = 0

    def _row(self, i, bundle):
        """Trace one bundle for one bounce; energy absorbed by every element / energy emitted."""
        import torch
        eng = self.engine
        eng.ray_tracer(bundle, reps=1, min_energy=1e-15, tree=False)       # itmax = 1 (view_factors_3D.py:141-142)
        dev = eng.absorbed_per_surface.device
        if getattr(self, '_elem_dev', None) is None or self._elem_dev.device != dev:
            e = torch.from_numpy(self.element_of_surface).to(dev)
            self._elem_mask = e >= 0
            self._elem_dev = e.clamp(min=0)
        row = torch.zeros(self.VF.shape[0], dtype=torch.float64, device=dev)
        row.index_add_(0, self._elem_dev, eng.absorbed_per_surface * self._elem_mask)
        emitted = bundle._props['energy'] if hasattr(bundle, '_props') else None
        if emitted is not None and hasattr(emitted, 'pieces'):
            total = sum(float(p.sum().item()) for p in emitted.pieces)
        else:
            total = float(N.sum(bundle.get_energy()))
        return (row / total).cpu().numpy()

    def run(self, max_passes=1000):
        """Passes over all unconverged rows until every entry meets the precision targets.
        Returns (VF_esperance, stdev_VF)."""
        n = self.VF.shape[0]
        stable = N.zeros(n)
        while ((self.progress.any() or (stable < self.min_passes).any()) and self.passes < max_passes):
            self.ray_counts = N.zeros(n)
            for i in range(n):
                if self.progress[i].any() or stable[i] < self.min_passes:
                    self.VF[i] = self._row(i, self.emit(i, self.num_rays))
                    self.ray_counts[i] = self.num_rays
                    self.p[i] += self.num_rays
                    stable[i] += 1
                    self.reset_opt()
            self.test_precision()
            self.passes += 1
        return self.VF_esperance, self.stdev_VF
