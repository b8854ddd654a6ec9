"""Optics laws on host arrays, with the names and arguments of tracer/optics.py.

During a trace these laws run inside the CUDA optics kernel (csrc/optics.cuh: ``reflect``,
``apply_refractive``, the TRC_O_FRESNEL_CONDUCTOR branch).  The functions here are the
reference's *utility* API for scripts that evaluate a reflectance or a refracted direction on a
handful of vectors outside a trace (tests/test_optics.py of the reference calls them that way);
the bounce loop never calls them.

All arrays are (3, n) for vectors and (n,) for scalars per ray, as in the reference.
"""
import numpy as np


def _cosines(ray_dirs, normals):
    return np.sum(np.asarray(normals, float) * np.asarray(ray_dirs, float), axis=0)


def reflections(ray_dirs, normals):
    """Mirror directions d - 2 (d.n) n  (optics.py:145-157)."""
    ray_dirs = np.asarray(ray_dirs, float)
    normals = np.asarray(normals, float)
    return ray_dirs - 2. * (_cosines(ray_dirs, normals) * normals)


def refractions(n1, n2, ray_dirs, normals):
    """Snell's law in vector form (optics.py:159-192).  Returns (mask of the rays that are not
    totally reflected, refracted directions of those rays)."""
    ray_dirs = np.asarray(ray_dirs, float)
    ratio = np.broadcast_to(np.asarray(n2, float) / np.asarray(n1, float), ray_dirs[0].shape)
    normals = np.broadcast_to(np.asarray(normals, float), ray_dirs.shape)
    cos1 = _cosines(ray_dirs, normals)
    refracted = cos1 ** 2 >= 1. - ratio ** 2
    cos1, ratio = cos1[refracted], ratio[refracted]
    d, n = ray_dirs[:, refracted], normals[:, refracted]
    out = (d - cos1 * n) / ratio
    cos2 = np.sqrt(1 - 1. / ratio ** 2 * (1. - cos1 ** 2))
    out = out + n * cos2 * np.where(cos1 < 0., -1, 1)
    return refracted, out


def fresnel(ray_dirs, normals, n1, n2):
    """Reflectance of unpolarised light at a dielectric interface, (R_s + R_p) / 2 (optics.py:13-39)."""
    theta_in = np.arccos(np.abs(_cosines(ray_dirs, normals)))
    c1 = np.cos(theta_in)
    c2 = np.sqrt(1 - (n1 / n2 * np.sin(theta_in)) ** 2)
    r_s = ((n1 * c1 - n2 * c2) / (n1 * c1 + n2 * c2)) ** 2
    r_p = ((n1 * c2 - n2 * c1) / (n1 * c2 + n2 * c1)) ** 2
    return (r_s + r_p) / 2


def fresnel_to_attenuating(n1, m2, theta1):
    """Perfect dielectric n1 onto an absorbing medium of complex index m2 (Modest, ch. 2;
    optics.py:63-81).  Returns (R_p, R_s, refraction angle)."""
    m2 = np.asarray(m2)
    s1, c1 = n1 * np.sin(theta1), n1 * np.cos(theta1)
    b = m2.real ** 2 - m2.imag ** 2 - s1 ** 2
    a = np.sqrt(b ** 2 + 4. * (m2.real * m2.imag) ** 2)
    p, q = np.sqrt(0.5 * (a + b)), np.sqrt(0.5 * (a - b))
    theta2 = np.arctan(s1 / p)
    r_s = ((c1 - p) ** 2 + q ** 2) / ((c1 + p) ** 2 + q ** 2)
    st = s1 * np.tan(theta1)
    r_p = ((p - st) ** 2 + q ** 2) / ((p + st) ** 2 + q ** 2) * r_s
    return r_p, r_s, theta2


def fresnel_conductor(ray_dirs, normals, lambdas, material, n1=1., m2=None):
    """Reflectances of rays arriving from a dielectric n1 on a conductor described by an optical
    material (``material.m(wavelengths)``) or a forced complex index m2 (optics.py:41-61)."""
    if m2 is None:
        m2 = material.m(lambdas)
    theta_in = np.arccos(np.abs(_cosines(ray_dirs, normals)))
    return fresnel_to_attenuating(n1, m2, theta_in)


def refr_idx_hartmann(wavelength, a, b, c, d, e):
    """Hartmann dispersion n(L) = a + b / (c - L) + d / (e - L)  (optics.py:194-203)."""
    return a + b / (c - wavelength) + d / (e - wavelength)


def attenuations(path_lengths, k, lambda_0, energy):
    """Beer-Lambert attenuation exp(-4 pi L k / lambda_0) of ``energy`` (optics.py:205-212)."""
    return np.exp(-4. * np.pi * path_lengths * k / lambda_0) * energy
