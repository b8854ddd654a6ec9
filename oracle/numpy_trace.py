"""TEST INFRASTRUCTURE ONLY -- optics, accountants and the driver loop of the
NumPy restatement (see oracle/numpy_engine.py for the header and the parity
pin).  ``OracleEngine.ray_tracer`` restates tracer/tracer_engine.py:124-295.
"""
import numpy as N

from .numpy_engine import (_names, find_intersections, reflections, rotation_to_z, rotate_z_to_normal,
                           pillbox_directions)


class Bundle(object):
    """Minimal SoA bundle of the oracle (tracer/ray_bundle.py:6-63)."""

    def __init__(self, vertices, directions, energy, parents=None, wavelengths=None):
        self.vertices, self.directions, self.energy = vertices, directions, energy
        self.parents, self.wavelengths = parents, wavelengths

    @property
    def n(self):
        return self.vertices.shape[1]

    def take(self, idx):
        return Bundle(self.vertices[:, idx], self.directions[:, idx], self.energy[idx],
                      None if self.parents is None else self.parents[idx],
                      None if self.wavelengths is None else self.wavelengths[idx])

    @staticmethod
    def concat(bs):
        if not bs:
            return Bundle(N.zeros((3, 0)), N.zeros((3, 0)), N.array([]), N.array([], dtype=int))
        wl = None if bs[0].wavelengths is None else N.hstack([b.wavelengths for b in bs])
        return Bundle(N.hstack([b.vertices for b in bs]), N.hstack([b.directions for b in bs]),
                      N.hstack([b.energy for b in bs]), N.hstack([b.parents for b in bs]), wl)

    # reference-style getters so tests can treat oracle and engine bundles alike
    def get_vertices(self):
        return self.vertices

    def get_directions(self):
        return self.directions

    def get_energy(self):
        return self.energy

    def get_parents(self):
        return self.parents

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
          'FresnelConductorHomogenous')


def _iam(base, d, normals, energy):
    """IAM.__call__, optics_callables.py:277-282"""
    vertical = N.sum(d * normals, axis=0) * normals
    cos_aoi = N.sqrt(N.sum(vertical ** 2, axis=0))
    return energy * (1. - base._abs) * (1. - N.exp(-cos_aoi ** base.c / base.a_r)) / (1. - N.exp(-1. / base.a_r))


def apply_optics(opt, d, normals, energy, wavelengths, up, v_parent=None, hit=None):
    """One optics callable on the selected rays of one surface.
    Returns (directions (3,h), energy (h,)); apply_optics.attenuations holds the
    attenuated energy of the last call (LambertianAbsorbant) or None."""
    base, _ = _unwrap(opt)
    names = _names(base)
    h = d.shape[1]
    one_sided = False
    apply_optics.attenuations = None
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
        elif name == 'Reflective_spectral':                         # :183-193
            ab = N.interp(wavelengths, base._wavelengths, base._absorptances)
            out = reflections(d, normals), energy * (1. - ab)
        else:
            continue
        one_sided = any(nm in ('OneSidedReflective', 'OneSidedRealReflective') for nm in names)
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
        bund = Bundle(N.asarray(bundle.get_vertices(), float), N.asarray(bundle.get_directions(), float),
                      N.asarray(bundle.get_energy(), float), None, wl)
        self.tree = [bund] if tree else []
        self.front_surfs = []
        self.tallies = [dict() for _ in surfaces]
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
                wl_in = None if bund.wavelengths is None else bund.wavelengths[sel]
                base_opt, _ = _unwrap(surf._opt)
                if 'PeriodicBoundary' in _names(base_opt):          # optics_callables.py:703-721
                    stub = Bundle(hit, d_in.copy(), N.zeros(len(sel)), sel, wl_in)
                    moved = Bundle(hit + base_opt.period * normals, d_in.copy(), e_in.copy(), sel, wl_in)
                    child = Bundle.concat([stub, moved])
                    record.append(child)
                    delete = child.energy <= min_energy
                    weak.append(delete)
                    outg.append(child.take(N.nonzero(~delete)[0]) if delete.any() else child)
                    continue
                dirs, e_out = apply_optics(surf._opt, d_in, normals, e_in, wl_in, surf._temp_frame[:3, 2],
                                           v_parent=bund.vertices[:, sel], hit=hit)
                child = Bundle(hit, dirs, e_out, sel, wl_in)
                self._count(s_idx, surf, hit, d_in, normals, e_in, e_out, wl_in, apply_optics.attenuations)
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

    def _count(self, s_idx, surf, hit, d_in, normals, e_in, e_out, wl_in, attenuated=None):
        """Accountant.count for each accountant (optics_callables.py:1597-1790)."""
        _, accs = _unwrap(surf._opt)
        data = {'_absorbed': e_in - e_out, '_attenuated': attenuated, '_received': e_in, '_scattered': e_out, '_hits': hit,
                '_directions': d_in, '_normals': normals, '_wavelengths': wl_in}
        t = self.tallies[s_idx]
        for a in accs:
            field = _ACC.get(getattr(a, 'shorthand', None))
            if field is None:
                raise NotImplementedError("oracle: accountant %s" % type(a).__name__)
            t.setdefault(field, []).append(data[field])
            if self.write_accountants:
                getattr(a, field).append(data[field])

    def get_all_hits(self, s_idx):
        """Same list the surface's optics manager would return from get_all_hits()."""
        _, accs = _unwrap(self._asm.get_surfaces()[s_idx]._opt)
        out = []
        for a in accs:
            field = _ACC[a.shorthand]
            chunks = self.tallies[s_idx].get(field, [])
            if field in ('_hits', '_directions', '_normals'):
                out.append(N.hstack(chunks) if chunks else N.zeros((3, 0)))
            else:
                out.append(N.hstack(chunks) if chunks else N.array([]))
        return out
