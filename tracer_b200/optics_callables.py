"""Optics-manager descriptions and accountants with the reference's names.

Reference: tracer/optics_callables.py.  There, each class is a callable doing
NumPy work per surface per bounce, and a factory (:1942-2092) generates ~20 000
``<Optics><Accountants...>`` classes at import.  Here the optics run inside the
CUDA optics kernel (csrc/optics.cuh); these classes hold parameters only, and
accountant-composed classes are synthesised on demand from their *name* by the
module-level ``__getattr__`` (PEP 562), using the reference's naming rules:

    <Optics> + shorthands in the order Absorber, Attenuator, Receptor,
    Scatterer, Polychromatic, Spectral, Location, Directional, Normal
    aliases: Receiver = Location+Absorber, Detector = Directional+Location+
    Absorber, Transmitter = Location+Scatterer   (optics_callables.py:2063-2074)

e.g. ``ReflectiveReceiver``, ``LambertianReceiver``, ``OneSidedReflectiveReceiver``,
``OneSidedRealReflectiveDetector``.  ``get_all_hits()`` returns the accountants'
data in the reference's order (energy, spectral, location, directions, normals).
The GPU engine appends one array per bounce to the accountants' lists
(``_absorbed``, ``_hits`` ...), exactly what Accountant.count would have stored.
"""
import sys
from copy import deepcopy

import numpy as np


def copy_optical_manager(opt):
    newopt = deepcopy(opt)
    if hasattr(newopt, 'reset'):
        newopt.reset()
    return newopt


class _OpticsDescription(object):
    """Common base: optics are evaluated on the GPU, never called on the host."""

    def __call__(self, geometry, rays, selector):
        raise TypeError(
            f"{type(self).__name__} is a description; optics run on the GPU through "
            "TracerEngine.ray_tracer (no CPU path)")


class Transparent(_OpticsDescription):
    """Passes rays through unchanged (optics_callables.py:93)."""

    def __init__(self):
        pass


class Reflective(_OpticsDescription):
    """Specular, opaque, grey absorptivity (optics_callables.py:116)."""

    def __init__(self, absorptivity):
        self._abs = absorptivity


class Lambertian(_OpticsDescription):
    """Ideal diffuse reflector within ``ang_range`` of the normal (:143)."""

    def __init__(self, absorptivity=0., ang_range=np.pi / 2.):
        self._abs = absorptivity
        self._ang_range = ang_range


class Reflective_spectral(_OpticsDescription):
    """Specular with absorptance interpolated in wavelength (:178)."""

    def __init__(self, absorptances, wavelengths):
        self._wavelengths = wavelengths
        self._absorptances = absorptances


class OneSidedReflective(Reflective):
    """Reflective from +z; rays arriving from behind lose their energy (:195)."""


class RealReflective(_OpticsDescription):
    """Specular with Gaussian slope error, bi-variate or conical (:214)."""

    def __init__(self, absorptivity, sigma, bi_var=False):
        self._abs = absorptivity
        self._sig = sigma
        self.bi_var = bi_var


class OneSidedRealReflective(RealReflective):
    """(:492)"""


class LambertianSpecular(_OpticsDescription):
    """Per-ray choice between specular and diffuse reflection (:553)."""

    def __init__(self, absorptivity=0., specularity=0.5):
        self._abs = absorptivity
        self.specularity = specularity


class Lambertian_directional_axisymmetric_piecewise(_OpticsDescription):
    """Diffuse; absorptance interpolated in the incidence angle (:331)."""

    def __init__(self, thetas, absorptance_th, specularity=0.):
        self.thetas = thetas
        self.abs_th = absorptance_th
        self.specularity = specularity


class LambertianSpecular_directional_axisymmetric_piecewise(_OpticsDescription):
    """Specular/diffuse mix; absorptance interpolated in the incidence angle (:427)."""

    def __init__(self, thetas, absorptance_th, specularity=0.):
        self.thetas = thetas
        self.abs_th = absorptance_th
        self.specularity = specularity


class Lambertian_directional_axisymmetric_piecewise_spectral(_OpticsDescription):
    """Diffuse; absorptance bilinear in (incidence angle, wavelength) (:359)."""

    def __init__(self, thetas, absorptance, wavelengths):
        self.thetas, self.wavelengths = np.unique(thetas), np.unique(wavelengths)
        self.absorptance = np.reshape(absorptance, (len(self.thetas), len(self.wavelengths)))


class Lambertian_directional_axisymmetric_piecewise_Polychromatic(_OpticsDescription):
    """Diffuse; every ray carries a spectrum (bundle properties ``wavelengths`` and ``spectra``, both (W, n)) whose
    W samples are each absorbed with the absorptance bilinear in (incidence angle, wavelength); the outgoing energy
    is the trapezoid integral of the outgoing spectrum over the ray's wavelengths (:393-425)."""

    def __init__(self, thetas, absorptance, wavelengths):
        self.thetas, self.wavelengths = np.unique(thetas), np.unique(wavelengths)
        self.absorptance = np.reshape(absorptance, (len(self.thetas), len(self.wavelengths)))


class BiFacial(object):
    """Two optics managers on one surface: rays travelling along the surface's +z ("up", geometry_manager.py:29-34)
    meet ``OpticsCallable_back``, the others ``OpticsCallable_front`` (:1877-1924).  As in the reference, each
    side's manager processes ALL the rays that hit the surface whenever one of them is on its side -- so its
    accountants record every hit with that side's optics -- and only its own side's children are kept, the back
    ones first.  Not part of the accountant class factory (:2086)."""

    def __init__(self, OpticsCallable_front, OpticsCallable_back):
        self.OpticsCallable_front = OpticsCallable_front
        self.OpticsCallable_back = OpticsCallable_back

    def __call__(self, geometry, rays, selector):
        raise TypeError("BiFacial is a description; optics run on the GPU through TracerEngine.ray_tracer (no CPU path)")

    def get_all_hits(self):
        out = []
        for side in (self.OpticsCallable_front, self.OpticsCallable_back):
            try:
                out.append(side.get_all_hits())
            except Exception:                       # the reference's bare except (:1905-1912)
                out.append([])
        return tuple(out)

    def reset(self):
        # the reference calls side.reset(self) -- one argument too many -- inside try/except (:1916-1924): the
        # sides are never reset through the BiFacial manager.  Reset them directly between traces.
        pass


class Lambertian_IAM(Lambertian):
    """Lambertian whose energy follows the Martin-Ruiz incidence-angle modifier (:303-318)."""

    def __init__(self, absorptivity, a_r, c=1):
        Lambertian.__init__(self, absorptivity)
        self.a_r = a_r
        self.c = c


class RealReflective_IAM(RealReflective):
    """RealReflective with the incidence-angle modifier, exponent c = 1 (:320-329)."""

    def __init__(self, absorptivity, a_r, sigma, bi_var=False):
        RealReflective.__init__(self, absorptivity, sigma, bi_var)
        self.a_r = a_r
        self.c = 1


class LambertianSpecular_IAM(_OpticsDescription):
    """Specular/diffuse mix (:586-627).  As written in the reference the modifier is
    evaluated on a zeroed direction array, so cos(AOI) = 0 and the energy is passed on
    unchanged whatever absorptivity and a_r are; that behaviour is what is reproduced."""

    def __init__(self, absorptivity=0., specularity=0.5, a_r=0.16):
        self._abs = absorptivity
        self.specularity = specularity
        self.a_r = a_r


class Lambertian_piecewise_Specular_directional_axisymmetric_piecewise(_OpticsDescription):
    """Specular/diffuse mix with absorptance and specularity both interpolated in the
    incidence angle (:457-487)."""

    def __init__(self, thetas, absorptance_th, specularity_th):
        self.thetas = thetas
        self.abs_th = absorptance_th
        self.spec_th = specularity_th


class LambertianAbsorbant(Lambertian):
    """Opaque diffuse wall of an absorbing volume (:891-906): the ray is attenuated over
    the path from its parent vertex (imposed attenuation coefficient, Absorbant :860-889),
    then absorbed.  ``attenuation_coefficient`` must be a one-element sequence, as in the
    reference (its ``len(self.a_c)`` rejects scalars)."""

    def __init__(self, absorptivity=0., attenuation_coefficient=0., ang_range=np.pi / 2., scaling=1.):
        self._absorptivity = absorptivity
        Lambertian.__init__(self, absorptivity=0., ang_range=ang_range)
        self._scaling = scaling
        self.a_c = attenuation_coefficient
        if self.a_c is not None:
            self.a_c = np.array(attenuation_coefficient)


class FresnelConductorHomogenous(_OpticsDescription):
    """Specular reflection off a conductor seen from a perfect dielectric of index ``n1``;
    reflectance from the Fresnel equations with the complex index of ``material`` at the
    ray's wavelength (:1523-1558)."""

    def __init__(self, n1, material):
        self._n1 = n1
        self._material = material


class PeriodicBoundary(_OpticsDescription):
    """Periodic boundary condition (:690-723): every hit leaves a zero-energy stub at the
    hit point and re-emits the ray, unchanged, from hit + period * normal."""

    def __init__(self, period):
        self.period = period


class RefractiveHomogenous(_OpticsDescription):
    """Interface between two homogenous media of refractive indices ``n1`` and ``n2``
    (:1186-1296).  The rays must carry a ``ref_index`` property; a ray travelling in one index
    enters the other.  ``single_ray=True`` picks reflection or refraction per ray from the
    Fresnel reflectance, ``False`` emits both children; ``sigma`` (None = smooth) perturbs the
    normal with a conical Gaussian error."""

    def __init__(self, n1, n2, single_ray=True, sigma=None):
        self._ref_idxs = (n1, n2)
        self._materials = [n1, n2]
        self._single_ray = single_ray
        self._sigma = sigma


class RefractiveTransmissiveHomogenous(RefractiveHomogenous):
    """RefractiveHomogenous followed by Absorbant.attenuate (:1323-1345, :874-889): the outgoing
    rays are attenuated over the path from the parent vertex with the imposed
    ``attenuation_coefficients`` (one value, or one per medium in the order n1, n2)."""

    def __init__(self, n1, n2, attenuation_coefficients, single_ray=True, sigma=None, scaling=1.):
        RefractiveHomogenous.__init__(self, n1, n2, single_ray, sigma)
        self._scaling = scaling
        self.a_c = attenuation_coefficients
        if self.a_c is not None:
            self.a_c = np.array(attenuation_coefficients)


class Refractive(_OpticsDescription):
    """Interface between two media whose complex refractive indices come from material objects of the
    optical_constants module (``material.m(wavelengths)`` = n + i k, :726-858).  The rays carry ``wavelengths`` and a
    (complex) ``ref_index``; a ray whose index equals material_1's at its wavelength enters material_2 and vice
    versa.  Snell's law uses the real parts, the Fresnel reflectance the complex indices (its real part is kept)."""

    def __init__(self, material_1, material_2, single_ray=True, sigma=None):
        self._materials = [material_1, material_2]
        self._single_ray = single_ray
        self._sigma = sigma


class RefractiveAbsorbant(Refractive):
    """Refractive followed by Absorbant.attenuate (:908-944, :874-881): the outgoing rays are attenuated over the path
    from the parent vertex by exp(-4 pi k path / lambda), k the imaginary part of their own index.  As in the
    reference the constructor's logic is inverted: that happens when an attenuation coefficient IS given (its value
    is discarded, :920-924); with none given the reference raises AttributeError on the first hit."""

    def __init__(self, material_1, material_2, single_ray=True, sigma=None, attenuation_coefficient_1=None,
                 attenuation_coefficient_2=None, scaling=1.):
        Refractive.__init__(self, material_1, material_2, single_ray, sigma)
        if (attenuation_coefficient_1 is None) and (attenuation_coefficient_2 is None):
            attenuation_coefficients = [attenuation_coefficient_1, attenuation_coefficient_2]
        else:
            attenuation_coefficients = None
        self._scaling = scaling
        self.a_c = attenuation_coefficients
        if self.a_c is not None:
            self.a_c = np.array(attenuation_coefficients)


perfect_mirror = Reflective(0)

_BASE_OPTICS = {c.__name__: c for c in (
    Refractive, RefractiveAbsorbant, RefractiveHomogenous, RefractiveTransmissiveHomogenous,
    PeriodicBoundary, FresnelConductorHomogenous,
    Lambertian_IAM, RealReflective_IAM, LambertianSpecular_IAM, LambertianAbsorbant,
    Lambertian_piecewise_Specular_directional_axisymmetric_piecewise,
    Transparent, Reflective, Lambertian, Reflective_spectral, OneSidedReflective, RealReflective,
    OneSidedRealReflective, LambertianSpecular, Lambertian_directional_axisymmetric_piecewise,
    LambertianSpecular_directional_axisymmetric_piecewise,
    Lambertian_directional_axisymmetric_piecewise_spectral,
    Lambertian_directional_axisymmetric_piecewise_Polychromatic)}

# Optics of the reference without a GPU counterpart (DESIGN.md 4.4 gives the reason for each: the
# scattering family, Reflective_IAM and SemiLambertian raise inside the reference itself; the others
# need complex indices, per-ray spectra or two optics per surface): constructing them says so
# instead of silently mis-tracing.
_UNSUPPORTED = (
    'RefractiveAbsorbantHomogenous', 'RefractiveScattering',
    'RefractiveScatteringHomogenous', 'Scattering', 'ScatteringPeriodicBoundary',
    'Absorbant', 'SemiLambertian', 'Reflective_IAM')


# ------------------------------------------------------------ accountants
class Accountant(object):
    """Per-surface result store (optics_callables.py:1577-1595).  ``field`` is
    the list attribute the engine appends per-bounce arrays to."""
    shorthand = None
    field = None
    vector = False

    def __init__(self):
        self.reset()

    def reset(self):
        setattr(self, self.field, [])

    def get_data(self):
        chunks = getattr(self, self.field)
        if self.vector:
            chunks = [c for c in chunks if c.shape[1]]
            return np.hstack(chunks) if chunks else np.array([]).reshape(3, 0)
        chunks = [c for c in chunks if len(c)]
        return np.hstack(chunks) if chunks else np.array([])


class AbsorptionAccountant(Accountant):
    shorthand, field = 'Absorber', '_absorbed'


class AttenuationAccountant(Accountant):
    shorthand, field = 'Attenuator', '_attenuated'


class ReceptionAccountant(Accountant):
    shorthand, field = 'Receptor', '_received'


class ScatteringAccountant(Accountant):
    shorthand, field = 'Scatterer', '_scattered'


class PolychromaticAccountant(Accountant):
    """Per bounce: the (W, h) wavelengths of the outgoing rays in ``_wavelengths`` and the spectral power lost at the
    hit (incoming minus outgoing spectra) in ``_spectra`` (:1825-1838).  ``get_data`` raises as the reference's does
    (it reads ``_absorbed``, which ``reset`` never creates, :1840-1848): read the two lists."""
    shorthand, field = 'Polychromatic', '_spectra'

    def reset(self):
        self._wavelengths = []
        self._spectra = []

    def get_data(self):
        raise AttributeError("'PolychromaticAccountant' object has no attribute '_absorbed'")


class SpectralAccountant(Accountant):
    """Per bounce: the wavelengths of the incoming rays that hit, in hit order (:1801-1811).  ``get_data`` raises as
    the reference's does (it tests ``self._absorbed``, which this accountant never creates, :1813-1823): read
    ``_wavelengths``."""
    shorthand, field = 'Spectral', '_wavelengths'

    def get_data(self):
        raise AttributeError("'SpectralAccountant' object has no attribute '_absorbed'")


class LocationAccountant(Accountant):
    shorthand, field, vector = 'Location', '_hits', True


class DirectionAccountant(Accountant):
    shorthand, field, vector = 'Directional', '_directions', True


class NormalAccountant(Accountant):
    shorthand, field, vector = 'Normal', '_normals', True


# canonical order = get_all_hits() order (optics_callables.py:2063-2068)
_ACCOUNTANTS = (AbsorptionAccountant, AttenuationAccountant, ReceptionAccountant, ScatteringAccountant,
                PolychromaticAccountant, SpectralAccountant, LocationAccountant, DirectionAccountant,
                NormalAccountant)
_ALIASES = {'Receiver': ('Location', 'Absorber'),
            'Detector': ('Directional', 'Location', 'Absorber'),
            'Transmitter': ('Location', 'Scatterer')}


class OpticsCallable(_OpticsDescription):
    """Wrapper pairing a base optics description (``_opt``) with accountants
    (``accountants``), as produced by make_mixed_accountant_class (:1942-1972)."""
    optics_class = None
    accountant_classes = ()

    def __init__(self, *args, **kwargs):
        self._opt = self.optics_class(*args, **kwargs)
        self.accountants = [a() for a in self.accountant_classes]

    def reset(self):
        for a in self.accountants:
            a.reset()

    def get_all_hits(self):
        return [a.get_data() for a in self.accountants]


def _parse_accountant_suffix(suffix):
    """'DirectionalReceiver' -> tuple of accountant classes in canonical order, or None."""
    wanted = set()
    for alias, parts in _ALIASES.items():
        if suffix.endswith(alias):
            wanted.update(parts)
            suffix = suffix[:-len(alias)]
            break
    pos = 0
    for acc in _ACCOUNTANTS:          # explicit shorthands must come in canonical order
        if suffix.startswith(acc.shorthand, pos):
            wanted.add(acc.shorthand)
            pos += len(acc.shorthand)
    if pos != len(suffix) or not wanted:
        return None
    return tuple(a for a in _ACCOUNTANTS if a.shorthand in wanted)


_generated = {}


def _make_class(name):
    best = None
    for base_name, base in _BASE_OPTICS.items():
        if name.startswith(base_name) and (best is None or len(base_name) > len(best[0])):
            accs = _parse_accountant_suffix(name[len(base_name):])
            if accs is not None:
                best = (base_name, base, accs)
    if best is None:
        return None
    _, base, accs = best
    return type(name, (OpticsCallable,), {'optics_class': base, 'accountant_classes': accs,
                                          '__module__': __name__})


def __getattr__(name):
    if name.startswith('__'):
        raise AttributeError(name)
    if name in _generated:
        return _generated[name]
    for bad in _UNSUPPORTED:
        if name.startswith(bad) and not any(name.startswith(b) and len(b) > len(bad) for b in _BASE_OPTICS):
            raise NotImplementedError(
                f"optics class {name!r} (base {bad!r}) is outside the GPU bounce-loop scope "
                "(SURVEY.md section 8f row N4); no CPU fallback is provided")
    cls = _make_class(name)
    if cls is None:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    _generated[name] = cls
    setattr(sys.modules[__name__], name, cls)
    return cls
