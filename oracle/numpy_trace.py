"""TEST INFRASTRUCTURE ONLY -- optics, accountants and the driver loop of the
NumPy restatement (see oracle/numpy_engine.py for the header and the parity
pin).  ``OracleEngine.ray_tracer`` restates tracer/tracer_engine.py:124-295.
"""
import numpy as N

from .numpy_engine import (_names, find_intersections, reflections, rotation_to_z, rotate_z_to_normal,
                           pillbox_directions)


class Bundle(object):
    """Minimal SoA bundle of the oracle (tracer/ray_bundle.py:6-63).  ``wavelengths`` is (n,) for monochromatic
    rays or (W, n) next to ``spectra`` (W, n) for polychromatic ones: the ray index is the last axis of every
    property (ray_bundle.py:84-90)."""

    def __init__(self, vertices, directions, energy, parents=None, wavelengths=None, ref_index=None, spectra=None):
        self.vertices, self.directions, self.energy = vertices, directions, energy
        self.parents, self.wavelengths, self.ref_index = parents, wavelengths, ref_index
        self.spectra = spectra

    @property
    def n(self):
        return self.vertices.shape[1]

    def take(self, idx):
        return Bundle(self.vertices[:, idx], self.directions[:, idx], self.energy[idx],
                      None if self.parents is None else self.parents[idx],
                      None if self.wavelengths is None else self.wavelengths[..., idx],
                      None if self.ref_index is None else self.ref_index[idx],
                      None if self.spectra is None else self.spectra[..., idx])

    @staticmethod
    def concat(bs):
        if not bs:
            return Bundle(N.zeros((3, 0)), N.zeros((3, 0)), N.array([]), N.array([], dtype=int))
        wl = None if bs[0].wavelengths is None else N.hstack([b.wavelengths for b in bs])
        ri = None if bs[0].ref_index is None else N.hstack([b.ref_index for b in bs])
        sp = None if bs[0].spectra is None else N.hstack([b.spectra for b in bs])
        return Bundle(N.hstack([b.vertices for b in bs]), N.hstack([b.directions for b in bs]),
                      N.hstack([b.energy for b in bs]), N.hstack([b.parents for b in bs]), wl, ri, sp)

    # reference-style getters so tests can treat oracle and engine bundles alike
    def get_vertices(self):
        return self.vertices

    def get_directions(self):
        return self.directions

    def get_energy(self):
        return self.energy

    def get_parents(self):
        return self.parents

    def get_ref_index(self):
        return self.ref_index

    def get_wavelengths(self):
        return self.wavelengths

    def get_spectra(self):
        return self.spectra

    def get_num_rays(self):
        return self.n


def _unwrap(opt):
    if hasattr(opt, 'accountants') and hasattr(opt, '_opt'):
        return opt._opt, list(opt.accountants)
    return opt, []


def _lambert(normals, ang_range, n):
    """Lambertian.__call__ direction part, optics_callables.py:163-165"""
    directs = pillbox_directions(n, ang_range)
    rot = rotation_to_z(normals.T)
    return N.sum(rot * directs.T[:, None, :], axis=2).T


def _theta_in(d, normals):
    vertical = N.sum(d * normals, axis=0) * normals
    return N.arccos(N.sqrt(N.sum(vertical ** 2, axis=0)))


_KNOWN = ('Transparent', 'OneSidedReflective', 'Reflective', 'OneSidedRealReflective', 'RealReflective',
          'Lambertian', 'LambertianSpecular', 'Lambertian_directional_axisymmetric_piecewise',
          'LambertianSpecular_directional_axisymmetric_piecewise',
          'Lambertian_directional_axisymmetric_piecewise_spectral', 'Reflective_spectral',
          'Lambertian_IAM', 'RealReflective_IAM', 'LambertianSpecular_IAM', 'LambertianAbsorbant',
          'Lambertian_piecewise_Specular_directional_axisymmetric_piecewise', 'PeriodicBoundary',
          'FresnelConductorHomogenous', 'Lambertian_directional_axisymmetric_piecewise_Polychromatic')


def fresnel(ray_dirs, normals, n1, n2):
    """optics.fresnel, optics.py:13-39"""
    theta_in = N.arccos(N.abs((normals * ray_dirs).sum(axis=0)))
    foo = N.cos(theta_in)
    bar = N.sqrt(1 - (n1 / n2 * N.sin(theta_in)) ** 2)
    Rs = ((n1 * foo - n2 * bar) / (n1 * foo + n2 * bar)) ** 2
    Rp = ((n1 * bar - n2 * foo) / (n1 * bar + n2 * foo)) ** 2
    return (Rs + Rp) / 2


def refractions(n1, n2, ray_dirs, normals):
    """optics.refractions, optics.py:159-192: (mask of rays not totally reflected, their new directions)"""
    n = N.broadcast_arrays(n2 / n1, ray_dirs[0])[0]
    cos1 = (normals * ray_dirs).sum(axis=0)
    refracted = cos1 ** 2 >= 1. - n ** 2
    cos1, ray_dirs, normals, n = cos1[refracted], ray_dirs[:, refracted], normals[:, refracted], n[refracted]
    refr_dirs = (ray_dirs - cos1 * normals) / n
    cos2 = N.sqrt(1 - 1. / n ** 2 * (1. - cos1 ** 2))
    refr_dirs += normals * cos2 * N.where(cos1 < 0., -1, 1)
    return refracted, refr_dirs


def refractive_children(base, hit, d, normals, energy, sel, wl, m1, v_parent, all_wl, all_ri):
    """RefractiveHomogenous.__call__ (optics_callables.py:842-858 through :1222-1296) and, for
    RefractiveTransmissiveHomogenous (:1342-1345), Absorbant.attenuate (:874-889) with imposed
    coefficients.  Returns the child Bundle in [reflected][refracted] order.

    In single_ray mode the reference inherits the columns it does not set (ref_index of the
    reflected rays, wavelengths of both kinds) through ``rays.inherit(sel_refl, ...)`` where
    sel_refl = argwhere(refl) counts positions WITHIN the surface's selection (:1257-1258) but
    indexes the whole incoming bundle; ``all_wl`` / ``all_ri`` are those whole-bundle columns."""
    h = len(sel)
    normals = normals.copy()
    if base._sigma is not None:                                    # :1229-1241
        th = N.random.normal(scale=base._sigma, size=h)
        phi = N.random.uniform(low=0., high=2. * N.pi, size=h)
        errors = N.vstack((N.sin(th) * N.cos(phi), N.sin(th) * N.sin(phi), N.cos(th)))
        rots = rotation_to_z(normals.T)
        if rots.ndim == 2:
            rots = [rots]
        for i in range(h):
            normals[:, i] = N.dot(rots[i], errors[:, i])
    tabulated = not hasattr(base, '_ref_idxs')      # Refractive / RefractiveAbsorbant: materials m(lambda), complex
    if tabulated:                                   # toggle_ref_idx :744-756, _refract_dirs :781-785
        mat_0 = base._materials[0].m(wl)
        n2 = N.where(m1 == mat_0, base._materials[1].m(wl), mat_0)
        refr, out_dirs = refractions(N.real(m1), N.real(n2), d, normals)
    else:
        n2 = N.where(m1 == base._ref_idxs[0], base._ref_idxs[1], base._ref_idxs[0])     # :1217-1218
        refr, out_dirs = refractions(m1, n2, d, normals)
    R = N.ones(h)
    import warnings as _w
    with _w.catch_warnings():                       # complex reflectance into the float array: the real part (:839)
        _w.simplefilter('ignore')
        R[refr] = fresnel(d[:, refr], normals[:, refr], m1[refr], n2[refr])         # :841-843
    take = (lambda a, m: None if a is None else a[m])
    if base._single_ray:                                           # :1252-1281
        refl = N.random.uniform(size=R.shape) <= R
        dirs_refr = N.zeros((3, h))
        dirs_refr[:, refr] = out_dirs
        parts = []
        sel_refl, sel_refr = N.argwhere(refl).flatten(), N.argwhere(~refl).flatten()
        if refl.any():
            parts.append(Bundle(hit[:, refl], reflections(d[:, refl], normals[:, refl]), energy[refl], sel[refl],
                                take(all_wl, sel_refl), all_ri[sel_refl]))
        if (~refl).any():
            parts.append(Bundle(hit[:, ~refl], dirs_refr[:, ~refl], energy[~refl], sel[~refl],
                                take(all_wl, sel_refr), n2[~refl]))
    else:                                                          # :1284-1294
        parts = [Bundle(hit, reflections(d, normals), energy * R, sel, wl, m1),
                 Bundle(hit[:, refr], out_dirs, energy[refr] * (1. - R[refr]), sel[refr], take(wl, refr), n2[refr])]
    child = Bundle.concat(parts)
    if tabulated and hasattr(base, 'a_c') and base.a_c is None:    # RefractiveAbsorbant: attenuate :874-881 with k
        pos = N.searchsorted(sel, child.parents)
        path = N.sqrt(N.sum((child.vertices - v_parent[:, pos]) ** 2, axis=0))
        if base._scaling != 1.:
            path = path * base._scaling
        child.energy = N.exp(-4. * N.pi * path * N.imag(child.ref_index) / child.wavelengths) * child.energy   # optics.py:205-212
    elif getattr(base, 'a_c', None) is not None:                   # attenuate :874-889
        # parents index the incoming bundle; v_parent / m1 are given per selected ray, so map back
        pos = N.searchsorted(sel, child.parents)
        path = N.sqrt(N.sum((child.vertices - v_parent[:, pos]) ** 2, axis=0))
        if base._scaling != 1.:
            path = path * base._scaling
        if len(base.a_c) > 1:
            a_c = base.a_c[N.array(m1[pos] == base._ref_idxs[1], dtype=int)]
        else:
            a_c = base.a_c
        child.energy = child.energy * N.exp(-a_c * path)
    return child


def _iam(base, d, normals, energy):
    """IAM.__call__, optics_callables.py:277-282"""
    vertical = N.sum(d * normals, axis=0) * normals
    cos_aoi = N.sqrt(N.sum(vertical ** 2, axis=0))
    return energy * (1. - base._abs) * (1. - N.exp(-cos_aoi ** base.c / base.a_r)) / (1. - N.exp(-1. / base.a_r))


def apply_optics(opt, d, normals, energy, wavelengths, up, v_parent=None, hit=None, spectra=None):
    """One optics callable on the selected rays of one surface.
    Returns (directions (3,h), energy (h,)); apply_optics.attenuations holds the
    attenuated energy of the last call (LambertianAbsorbant) or None, apply_optics.spectra_out the
    outgoing spectra of a polychromatic callable or None (then the child inherits the parent's)."""
    base, _ = _unwrap(opt)
    names = _names(base)
    h = d.shape[1]
    one_sided = False
    apply_optics.attenuations = None
    apply_optics.spectra_out = None
    for cls in type(base).__mro__:          # a subclass with its own __call__ is not its parent
        if cls.__name__ in _KNOWN:
            break
        if '__call__' in vars(cls) and cls.__name__ != '_OpticsDescription':
            raise NotImplementedError("oracle: optics %s" % type(base).__name__)
    for name in names:
        if name == 'Lambertian_IAM':                                # :303-318
            out = _lambert(normals, base._ang_range, h), _iam(base, d, normals, energy)
        elif name == 'RealReflective_IAM':                          # :320-329
            dirs, _ = apply_optics(_PlainRealReflective(base), d, normals, energy, wavelengths, up)
            out = dirs, _iam(base, d, normals, energy)
        elif name == 'LambertianSpecular_IAM':                      # :596-627
            directs = N.zeros(d.shape)
            vertical = N.sum(directs * normals, axis=0) * normals     # sic: the zeroed array (:606)
            cos_aoi = N.sqrt(N.sum(vertical ** 2, axis=0))
            specular = N.random.rand(h) < base.specularity
            directs[:, specular] = reflections(d[:, specular], normals[:, specular])
            directs[:, ~specular] = _lambert(normals[:, ~specular], N.pi / 2., int(N.sum(~specular)))
            out = directs, energy * (1. - base._abs * (1. - N.exp(-cos_aoi / base.a_r)) / (1. - N.exp(-1. / base.a_r)))
        elif name == 'Lambertian_piecewise_Specular_directional_axisymmetric_piecewise':   # :466-487
            th = _theta_in(d, normals)
            ab = N.interp(th, base.thetas, base.abs_th)
            spec = N.interp(th, base.thetas, base.spec_th)
            directs = N.zeros(d.shape)
            specular = N.random.rand(h) < spec
            directs[:, specular] = reflections(d[:, specular], normals[:, specular])
            directs[:, ~specular] = _lambert(normals[:, ~specular], N.pi / 2., int(N.sum(~specular)))
            out = directs, energy * (1. - ab)
        elif name == 'LambertianAbsorbant':                         # :901-906 with attenuate :874-889
            dirs = _lambert(normals, base._ang_range, h)
            e0 = energy * (1. - 0.)
            path = N.sqrt(N.sum((hit - v_parent) ** 2, axis=0))
            if base._scaling != 1.:
                path = path * base._scaling
            if len(base.a_c) > 1:
                raise NotImplementedError("oracle: per-medium attenuation coefficients")
            e_att = e0 * N.exp(-base.a_c * path)
            apply_optics.attenuations = energy - e_att
            out = dirs, e_att * (1. - base._absorptivity)
        elif name == 'FresnelConductorHomogenous':                  # :1536-1558, optics.py:41-81
            m2 = base._material.m(wavelengths)
            th1 = N.arccos(N.abs((normals * d).sum(axis=0)))
            n1 = base._n1
            b = (m2.real ** 2 - m2.imag ** 2 - (n1 * N.sin(th1)) ** 2)
            a = N.sqrt(b ** 2 + 4. * (m2.real * m2.imag) ** 2)
            p, q = N.sqrt(0.5 * (a + b)), N.sqrt(0.5 * (a - b))
            R_s = ((n1 * N.cos(th1) - p) ** 2 + q ** 2) / ((n1 * N.cos(th1) + p) ** 2 + q ** 2)
            R_p = ((p - n1 * N.sin(th1) * N.tan(th1)) ** 2 + q ** 2) / ((p + n1 * N.sin(th1) * N.tan(th1)) ** 2 + q ** 2) * R_s
            out = reflections(d, normals), energy * ((R_p + R_s) / 2.)
        elif name == 'Transparent':                                 # optics_callables.py:106-113
            out = d.copy(), energy.copy()
        elif name in ('Reflective', 'OneSidedReflective'):          # :130-140, :201-212
            out = reflections(d, normals), energy * (1. - base._abs)
        elif name in ('RealReflective', 'OneSidedRealReflective'):  # :231-269, :498-504
            real = normals
            if base._sig > 0.:
                if base.bi_var == True:  # noqa: E712
                    tanx = N.tan(N.random.normal(scale=base._sig, size=N.shape(normals[1])))
                    tany = N.tan(N.random.normal(scale=base._sig, size=N.shape(normals[1])))
                    ez = (1. / (1. + tanx ** 2. + tany ** 2.)) ** 0.5
                    ex, ey = tanx * ez, tany * ez
                else:
                    th = N.random.normal(scale=base._sig, size=N.shape(normals[1]))
                    phi = N.random.uniform(low=0., high=2. * N.pi, size=N.shape(normals[1]))
                    ez = N.cos(th)
                    ex, ey = N.sin(th) * N.cos(phi), N.sin(th) * N.sin(phi)
                real = rotate_z_to_normal(N.vstack((ex, ey, ez)), normals)
                real = real / N.sqrt(N.sum(real ** 2, axis=0))
            out = reflections(d, real), energy * (1 - base._abs)
        elif name == 'Lambertian':                                  # :154-176
            out = _lambert(normals, base._ang_range, h), energy * (1. - base._abs)
        elif name == 'LambertianSpecular':                          # :562-585
            directs = N.zeros(d.shape)
            specular = N.random.rand(h) < base.specularity
            directs[:, specular] = reflections(d[:, specular], normals[:, specular])
            directs[:, ~specular] = _lambert(normals[:, ~specular], N.pi / 2., int(N.sum(~specular)))
            out = directs, energy * (1. - base._abs)
        elif name == 'Lambertian_directional_axisymmetric_piecewise':       # :340-358
            ab = N.interp(_theta_in(d, normals), base.thetas, base.abs_th)
            out = _lambert(normals, N.pi / 2., h), energy * (1. - ab)
        elif name == 'LambertianSpecular_directional_axisymmetric_piecewise':   # :437-455
            ab = N.interp(_theta_in(d, normals), base.thetas, base.abs_th)
            directs = N.zeros(d.shape)
            specular = N.random.rand(h) < base.specularity
            directs[:, specular] = reflections(d[:, specular], normals[:, specular])
            directs[:, ~specular] = _lambert(normals[:, ~specular], N.pi / 2., int(N.sum(~specular)))
            out = directs, energy * (1. - ab)
        elif name == 'Lambertian_directional_axisymmetric_piecewise_spectral':  # :370-388
            th = _theta_in(d, normals)
            if hasattr(base, 'interpolator'):
                ab = base.interpolator(N.array([th, wavelengths]).T)
            else:
                from scipy.interpolate import RegularGridInterpolator
                ab = RegularGridInterpolator((base.thetas, base.wavelengths), base.absorptance)(
                    N.array([th, wavelengths]).T)
            out = _lambert(normals, N.pi / 2., h), energy * (1. - ab)
        elif name == 'Lambertian_directional_axisymmetric_piecewise_Polychromatic':   # :403-425
            th = _theta_in(d, normals)
            pts = N.array([N.tile(th, (wavelengths.shape[0], 1)), wavelengths]).T      # (h, W, 2)
            if hasattr(base, 'interpolator'):
                ab = base.interpolator(pts)
            else:
                from scipy.interpolate import RegularGridInterpolator
                ab = RegularGridInterpolator((base.thetas, base.wavelengths), base.absorptance)(pts)
            sp = spectra * (1. - ab.T)
            dx = N.diff(wavelengths, axis=0)                            # N.trapz(sp, wavelengths, axis=0)
            e_out = N.sum(dx * (sp[1:] + sp[:-1]) / 2.0, axis=0)
            apply_optics.spectra_out = sp
            out = _lambert(normals, N.pi / 2., h), e_out
        elif name == 'Reflective_spectral':                         # :183-193
            ab = N.interp(wavelengths, base._wavelengths, base._absorptances)
            out = reflections(d, normals), energy * (1. - ab)
        else:
            continue
        one_sided = any(nm in ('OneSidedReflective', 'OneSidedRealReflective') for nm in names)
        if spectra is not None and apply_optics.spectra_out is None:
            # polychromatic bundle through a grey optics: the classes below scale the spectra like the energy
            # (optics_callables.py:137, 173, 266, 358), the IAM classes are not restated, the others pass them on
            if name in ('Reflective', 'OneSidedReflective', 'RealReflective', 'OneSidedRealReflective', 'Lambertian'):
                apply_optics.spectra_out = spectra * (1. - base._abs)
            elif name == 'Lambertian_directional_axisymmetric_piecewise':
                apply_optics.spectra_out = spectra * (1. - N.interp(_theta_in(d, normals), base.thetas, base.abs_th))
            elif name.endswith('_IAM') or name in ('LambertianAbsorbant', 'FresnelConductorHomogenous',
                                                   'Reflective_spectral',
                                                   'Lambertian_directional_axisymmetric_piecewise_spectral'):
                raise NotImplementedError("oracle: %s with a polychromatic bundle" % name)
        break
    else:
        raise NotImplementedError("oracle: optics %s" % type(base).__name__)
    dirs, e = out
    if one_sided:
        e = e.copy()
        e[N.sum(d * up[:, None], axis=0) > 0] = 0
    return dirs, e


class _PlainRealReflective(object):
    """View of a RealReflective_IAM object as its RealReflective part."""

    def __init__(self, base):
        self._abs, self._sig, self.bi_var = base._abs, base._sig, base.bi_var


_PlainRealReflective.__name__ = 'RealReflective'


_ACC = {'Absorber': '_absorbed', 'Attenuator': '_attenuated', 'Receptor': '_received', 'Scatterer': '_scattered', 'Location': '_hits',
        'Directional': '_directions', 'Normal': '_normals', 'Spectral': '_wavelengths'}


class OracleEngine(object):
    """CPU restatement of TracerEngine (tracer/tracer_engine.py:11-295)."""

    def __init__(self, assembly, write_accountants=False):
        self._asm = assembly
        self.write_accountants = write_accountants

    def intersect_ray(self, bund, surfaces):
        """tracer_engine.py:27-64 with surf_relevancy all ones"""
        n = bund.n
        rays_mins = N.ones(n) * N.inf
        earliest = -1 * N.ones(n, dtype=int)
        per_surface = []
        for s_num, surf in enumerate(surfaces):
            stack, points, normals = find_intersections(surf, bund.vertices, bund.directions)
            stack = stack.copy()
            stack[stack == 0.] = N.inf
            earlier = stack < rays_mins
            if earlier.any():
                rays_mins[earlier] = stack[earlier]
                earliest[earlier] = s_num
            per_surface.append((points, normals))
        return earliest, per_surface

    def ray_tracer(self, bundle, reps=100, min_energy=1e-10, tree=True):
        """Returns (vertices, directions) of the last live bundle; sets
        self.tree (list of Bundle), self.front_surfs (per-bounce hit indices) and
        self.tallies[s] = dict(field -> list of per-bounce arrays)."""
        surfaces = self._asm.get_surfaces()
        wl = None
        if hasattr(bundle, 'has_property') and bundle.has_property('wavelengths'):
            try:
                wl = N.asarray(bundle.get_wavelengths())
            except (KeyError, AttributeError):
                wl = None
        ri = None
        if hasattr(bundle, 'has_property') and bundle.has_property('ref_index'):
            try:
                ri = N.asarray(bundle.get_ref_index())
                ri = ri if N.iscomplexobj(ri) else ri.astype(float)
            except (KeyError, AttributeError):
                ri = None
        sp = None
        if hasattr(bundle, 'has_property') and bundle.has_property('spectra'):
            try:
                sp = N.asarray(bundle.get_spectra(), float)
            except (KeyError, AttributeError):
                sp = None
        bund = Bundle(N.asarray(bundle.get_vertices(), float), N.asarray(bundle.get_directions(), float),
                      N.asarray(bundle.get_energy(), float), None, wl, ri, sp)
        self.tree = [bund] if tree else []
        self.front_surfs = []
        self.tallies = {}
        self.ray_bounces = 0
        record = []
        for i in range(reps):
            self.ray_bounces += bund.n
            front, per_surface = self.intersect_ray(bund, surfaces)
            self.front_surfs.append(front)
            outg, record, weak = [], [], []
            for s_idx, surf in enumerate(surfaces):             # tracer_engine.py:218-249
                sel = N.nonzero(front == s_idx)[0]
                if len(sel) == 0:
                    continue
                points, normals_fn = per_surface[s_idx]
                hit = points(sel)
                normals = normals_fn(sel)
                d_in = bund.directions[:, sel]
                e_in = bund.energy[sel]
                wl_in = None if bund.wavelengths is None else bund.wavelengths[..., sel]
                sp_in = None if bund.spectra is None else bund.spectra[..., sel]
                ri_sel = None if bund.ref_index is None else bund.ref_index[sel]
                if 'BiFacial' in _names(surf._opt):                 # optics_callables.py:1877-1924
                    child = self._bifacial(s_idx, surf, hit, normals, d_in, e_in, wl_in, sp_in, ri_sel, sel,
                                           bund.vertices[:, sel])
                    record.append(child)
                    delete = child.energy <= min_energy
                    weak.append(delete)
                    outg.append(child.take(N.nonzero(~delete)[0]) if delete.any() else child)
                    continue
                base_opt, _ = _unwrap(surf._opt)
                if 'PeriodicBoundary' in _names(base_opt):          # optics_callables.py:703-721
                    ri_in = None if bund.ref_index is None else bund.ref_index[sel]
                    stub = Bundle(hit, d_in.copy(), N.zeros(len(sel)), sel, wl_in, ri_in)
                    moved = Bundle(hit + base_opt.period * normals, d_in.copy(), e_in.copy(), sel, wl_in, ri_in)
                    child = Bundle.concat([stub, moved])
                    record.append(child)
                    delete = child.energy <= min_energy
                    weak.append(delete)
                    outg.append(child.take(N.nonzero(~delete)[0]) if delete.any() else child)
                    continue
                if 'RefractiveHomogenous' in _names(base_opt) or 'Refractive' in _names(base_opt):      # row O8
                    child = refractive_children(base_opt, hit, d_in, normals, e_in, sel, wl_in,
                                                bund.ref_index[sel], bund.vertices[:, sel],
                                                bund.wavelengths, bund.ref_index)
                    if _unwrap(surf._opt)[1]:
                        # Accountant.count on the child bundle AS RETURNED, [reflected][refracted]: one ray per hit in
                        # single_ray mode, subtracted in that order from the hit-ordered incoming energies (:1638-1643);
                        # otherwise the shapes differ and numpy raises ValueError, as in the reference
                        self._count(s_idx, surf._opt, hit, d_in, normals, e_in, child.energy, wl_in)
                    record.append(child)
                    delete = child.energy <= min_energy
                    weak.append(delete)
                    outg.append(child.take(N.nonzero(~delete)[0]) if delete.any() else child)
                    continue
                dirs, e_out = apply_optics(surf._opt, d_in, normals, e_in, wl_in, surf._temp_frame[:3, 2],
                                           v_parent=bund.vertices[:, sel], hit=hit, spectra=sp_in)
                sp_out = sp_in if apply_optics.spectra_out is None else apply_optics.spectra_out
                child = Bundle(hit, dirs, e_out, sel, wl_in, ri_sel, sp_out)
                self._count(s_idx, surf._opt, hit, d_in, normals, e_in, e_out, wl_in, apply_optics.attenuations,
                            sp_in, sp_out)
                record.append(child)
                delete = e_out <= min_energy
                weak.append(delete)
                outg.append(child.take(N.nonzero(~delete)[0]) if delete.any() else child)
            bund = Bundle.concat(outg)
            if tree:
                rec = Bundle.concat(record)
                if rec.n != 0:
                    weak = N.hstack(weak)
                    self.tree.append(Bundle.concat([bund, rec.take(N.nonzero(weak)[0])]))
            if bund.n == 0:
                break
        if not tree:
            self.tree.append(Bundle.concat(record))
        return bund.vertices, bund.directions

    def _bifacial(self, s_idx, surf, hit, normals, d_in, e_in, wl_in, sp_in, ri_sel, sel, v_parent):
        """BiFacial.__call__ (optics_callables.py:1885-1901): rays travelling along +z of the surface ("up",
        geometry_manager.py:29-34) hit the back.  Each side's callable runs on ALL the selected rays when at least
        one ray is on its side (its draws and its accountants see all of them) and keeps its own side's children;
        the back children come first."""
        opt = surf._opt
        up = surf._temp_frame[:3, 2]
        proj = N.around(N.sum(d_in * up[:, None], axis=0), decimals=6)
        back = proj > 0.
        parts = []
        for side, side_opt, mask in (('back', opt.OpticsCallable_back, back), ('front', opt.OpticsCallable_front, ~back)):
            if not mask.any():
                continue
            base, _ = _unwrap(side_opt)
            if any(nm in ('PeriodicBoundary', 'RefractiveHomogenous', 'BiFacial') for nm in _names(base)):
                raise NotImplementedError("oracle: two-children optics inside BiFacial")
            dirs, e_out = apply_optics(side_opt, d_in, normals, e_in, wl_in, up, v_parent=v_parent, hit=hit,
                                       spectra=sp_in)
            sp_out = sp_in if apply_optics.spectra_out is None else apply_optics.spectra_out
            self._count((s_idx, side), side_opt, hit, d_in, normals, e_in, e_out, wl_in, apply_optics.attenuations,
                        sp_in, sp_out)
            parts.append(Bundle(hit, dirs, e_out, sel, wl_in, ri_sel, sp_out).take(N.nonzero(mask)[0]))
        return Bundle.concat(parts)

    def _count(self, key, opt, hit, d_in, normals, e_in, e_out, wl_in, attenuated=None, sp_in=None, sp_out=None):
        """Accountant.count for each accountant (optics_callables.py:1597-1848)."""
        _, accs = _unwrap(opt)
        data = {'_absorbed': e_in - e_out, '_attenuated': attenuated, '_received': e_in, '_scattered': e_out, '_hits': hit,
                '_directions': d_in, '_normals': normals, '_wavelengths': wl_in}
        t = self.tallies.setdefault(key, {})
        for a in accs:
            sh = getattr(a, 'shorthand', None)
            if sh == 'Polychromatic':                   # :1825-1838: the child's wavelengths, spectra lost at the hit
                t.setdefault('_wavelengths', []).append(wl_in)
                t.setdefault('_spectra', []).append(sp_in - sp_out)
                continue
            field = _ACC.get(sh)
            if field is None:
                raise NotImplementedError("oracle: accountant %s" % type(a).__name__)
            t.setdefault(field, []).append(data[field])
            if self.write_accountants:
                getattr(a, field).append(data[field])

    def _hits_of(self, key, opt):
        _, accs = _unwrap(opt)
        out = []
        for a in accs:
            if a.shorthand == 'Polychromatic':
                # PolychromaticAccountant.get_data reads self._absorbed, which its reset() never creates (:1840-1848)
                raise AttributeError("'PolychromaticAccountant' object has no attribute '_absorbed'")
            if a.shorthand == 'Spectral':               # the same slip in SpectralAccountant.get_data (:1813-1823)
                raise AttributeError("'SpectralAccountant' object has no attribute '_absorbed'")
            field = _ACC[a.shorthand]
            chunks = self.tallies.get(key, {}).get(field, [])
            if field in ('_hits', '_directions', '_normals'):
                out.append(N.hstack(chunks) if chunks else N.zeros((3, 0)))
            else:
                out.append(N.hstack(chunks) if chunks else N.array([]))
        return out

    def get_all_hits(self, s_idx):
        """Same list the surface's optics manager would return from get_all_hits(); a BiFacial manager returns
        (front, back), each the list of that side's callable or [] when it has no accountants (:1903-1914)."""
        opt = self._asm.get_surfaces()[s_idx]._opt
        if 'BiFacial' in _names(opt):
            sides = []
            for side, side_opt in (('front', opt.OpticsCallable_front), ('back', opt.OpticsCallable_back)):
                try:                                    # the reference swallows every error here (:1905-1912)
                    sides.append(self._hits_of((s_idx, side), side_opt) if _unwrap(side_opt)[1] else [])
                except AttributeError:
                    sides.append([])
            return tuple(sides)
        return self._hits_of(s_idx, opt)

    def get_polychromatic(self, key):
        """(wavelengths (W, h), absorbed spectra (W, h)) a PolychromaticAccountant holds in its lists; ``key`` is a
        surface index or (surface index, 'front' | 'back')."""
        t = self.tallies.get(key, {})
        return N.hstack(t['_wavelengths']), N.hstack(t['_spectra'])
