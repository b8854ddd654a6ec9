"""Monte-Carlo convergence control: the weighted Welford estimator
and the "trace until the confidence interval is reached" loop of the reference
(ray_trace_utils/estimator.py:3-56), written for tallies that live on the GPU.

``Estimator`` keeps the same state and formulas (mean, M2, n, n2; batches weighted
by their sample count) but accepts NumPy arrays *or* torch tensors, so a flux map or
the per-surface absorbed-energy vector of TracerEngine can be accumulated without
leaving HBM; only the scalar decision (max CI vs target) is read by the host."""
import numpy as np


def _lib_of(x):
    try:
        import torch
        if torch.is_tensor(x):
            return torch
    except ImportError:
        pass
    return np


class Estimator(object):
    """Weighted Welford algorithm over batches (estimator.py:3-39)."""

    def __init__(self, n_sigmas=3., relative_CI=True):
        self.mean = np.array([0.])
        self.M2 = np.array([0.])
        self.n = 0.
        self.n2 = 0.
        self.n_sigmas = n_sigmas
        self.relative_CI = relative_CI

    def update(self, values, num_samples):
        delta = values - self.mean if self.n else values - 0.
        self.n += num_samples
        if self.n == num_samples:
            self.mean = num_samples * delta / self.n
            self.M2 = num_samples * delta * (values - self.mean)
        else:
            self.mean = self.mean + num_samples * delta / self.n
            self.M2 = self.M2 + num_samples * delta * (values - self.mean)
        self.n2 += num_samples ** 2.

    def get_CI(self):
        """Confidence interval (relative to the mean when ``relative_CI``); inf before
        two batches have been seen."""
        xp = _lib_of(self.mean)
        if self.n == 0:
            return np.inf * np.ones(np.shape(self.mean))
        denom = self.n - self.n2 / self.n
        if denom <= 0:
            return xp.full_like(self.mean, np.inf) if xp is not np else np.inf * np.ones(self.mean.shape)
        stdev = xp.sqrt(self.M2 / denom)
        CI = self.n_sigmas * stdev / np.sqrt(self.n ** 2 / self.n2)
        if self.relative_CI:
            CI = CI / self.mean
        CI[stdev == 0.] = 0.
        return CI


def MCRT_to_CI(fun, target_CI, num_samples, n_sigmas=3., *args, **kwargs):
    """Call ``fun(num_rays=num_samples, ...)`` until every component of the estimate has a
    relative confidence interval below ``target_CI`` (estimator.py:41-56).  ``fun`` may
    return a scalar, an array or a CUDA tensor."""
    estimator = Estimator(n_sigmas)
    while True:
        ci = estimator.get_CI()
        worst = float(ci.max()) if hasattr(ci, 'max') else float(ci)
        if not worst > target_CI:
            break
        samples = fun(num_rays=num_samples, *args, **kwargs)
        estimator.update(samples, num_samples=num_samples)
    return estimator
