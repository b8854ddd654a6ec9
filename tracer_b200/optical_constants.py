"""Optical materials for the spectral optics (reference: ray_trace_utils/optical_constants.py).

Only what the GPU path needs: a complex refractive index m(lambda) = n + i k tabulated
against wavelength and interpolated linearly, which is what the reference's file-based
materials do (``interp1d(lambdas, m)``, optical_constants.py:100-140).  The scene compiler
uploads the knots, so the kernel interpolates exactly like the host."""
import numpy as np


class optical_material(object):
    def __init__(self, l_min, l_max, force_bounds=False):
        self.l_min, self.l_max = l_min, l_max
        self.force_bounds = force_bounds


class OpticalMaterialTable(optical_material):
    """m(lambda) from knots: wavelengths (m, increasing), n, k."""

    def __init__(self, wavelengths, n, k):
        self.lambdas = np.asarray(wavelengths, float)
        self.n, self.k = np.asarray(n, float), np.asarray(k, float)
        if self.lambdas.ndim != 1 or len(self.lambdas) < 2 or np.any(np.diff(self.lambdas) <= 0):
            raise ValueError("wavelengths must be a strictly increasing 1D array")
        optical_material.__init__(self, self.lambdas.min(), self.lambdas.max())

    def m(self, lambdas):
        lambdas = np.asarray(lambdas, float)
        out = np.interp(lambdas, self.lambdas, self.n) + 1j * np.interp(lambdas, self.lambdas, self.k)
        bad = (lambdas < self.l_min) | (lambdas > self.l_max)
        if np.any(bad):
            out = np.where(bad, np.nan, out)
        return out


class OpticalMaterialFromFile(OpticalMaterialTable):
    """CSV with a header row: wavelength, n, k columns (optical_constants.py:114-140)."""

    def __init__(self, filename, wavelength_col=0, n_col=1, k_col=2, wavelength_unit='nm'):
        data = np.loadtxt(filename, skiprows=1, delimiter=',', usecols=(wavelength_col, n_col, k_col))
        try:
            factor = {'nm': 1e-9, 'um': 1e-6, 'm': 1}[wavelength_unit]
        except KeyError:
            raise ValueError("Invalid wavelength unit. Use 'nm', 'um', or 'm'.")
        OpticalMaterialTable.__init__(self, data[:, 0] * factor, data[:, 1], data[:, 2])


class Air(object):
    """Vacuum / air placeholder: m = 1 at every wavelength (optical_constants.py:143-155)."""

    def m(self, lambdas):
        return np.ones(len(lambdas))

