"""Scene compiler: Assembly -> flat device tables.

Walks ``Assembly.get_surfaces()`` once (the reference does this per trace at
tracer_engine.py:165-166 and then loops over the Python objects twice per
bounce) and emits one TrcSurfaceRec per surface: global frame, host-inverted
frame, geometry kind + parameters, optics kind + parameters, flags; plus a
global AABB per bounded surface for the BVH.  Works on this package's
description classes *and* on the reference's own classes (dispatch is by class
name along the MRO and by the reference's attribute names), so a scene built
with ``tracer.*`` traces unchanged.

Unknown geometry/optics classes, overridden object relevance hooks and
accountants that have no device counterpart raise NotImplementedError naming
the class: there is no CPU fallback.
"""
import numpy as np

from . import _lib as L


def _mro_names(obj):
    return [c.__name__ for c in type(obj).__mro__]


def _scalar(x, what):
    a = np.asarray(x, dtype=float)
    if a.size != 1:
        raise NotImplementedError(f"{what} must be a scalar for the GPU engine (got shape {a.shape})")
    return float(a.reshape(-1)[0])


# --------------------------------------------------------------- geometry
def _geom_record(gm, surface):
    """-> (kind, gp list, local bbox (lo, hi) or None for unbounded)."""
    names = _mro_names(gm)
    name = names[0]
    inf = np.inf

    def box(x, y, z0, z1):
        return (np.array([-x, -y, z0], float), np.array([x, y, z1], float))

    def holes():
        c = np.asarray(gm.extr_centers, float).reshape(-1, 2)
        r = np.asarray(gm.extr_radii, float).reshape(-1)
        if len(c) != len(r):
            raise ValueError("extr_centers needs one (x, y) row per entry of extr_radii")
        return np.column_stack([c, r]).reshape(-1), len(r)

    for name in names:
        if name == 'PerforatedRectPlateGM':
            hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
            pool, n_holes = holes()
            return L.G_RECT_PERFORATED, ([hx, hy, 0., n_holes], pool, 2), box(hx, hy, 0., 0.)
        if name in ('PerforatedPolygonGM', 'FlatSimplePolygonGM'):
            prof = np.asarray(gm.profile, float)
            if prof.ndim != 2 or prof.shape[0] != 2 or prof.shape[1] < 3:
                raise ValueError("polygon profile must be a (2, n >= 3) array")
            closed = np.concatenate((prof, prof[:, :1]), axis=1)
            pool, n_holes = holes() if name == 'PerforatedPolygonGM' else (np.zeros(0), 0)
            lo, hi = prof.min(axis=1), prof.max(axis=1)
            return (L.G_POLYGON, ([0., prof.shape[1], n_holes], np.concatenate((closed[0], closed[1], pool)), 0),
                    (np.array([lo[0], lo[1], 0.]), np.array([hi[0], hi[1], 0.])))
        if name in ('EllipsoidGM', 'Ellipsoid'):
            semi = 1. / np.sqrt(np.array([gm.a, gm.b, gm.c], float))
            lims = []
            for k, lim in enumerate((getattr(gm, 'xlim', None), getattr(gm, 'ylim', None), getattr(gm, 'zlim', None))):
                lims += [-inf, inf] if lim is None else [float(lim[0]), float(lim[1])]
            lims = np.array(lims).reshape(3, 2)
            lo, hi = np.maximum(-semi, lims[:, 0]), np.minimum(semi, lims[:, 1])
            return L.G_ELLIPSOID, [gm.a, gm.b, gm.c] + list(lims.reshape(-1)), (lo, np.maximum(lo, hi))
        if name == 'ExtrudedRectPlateGM':
            hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
            cx, cy = np.asarray(gm.extr_center, float).reshape(-1)[:2]
            ex, ey = np.asarray(gm.extr_half_dims, float).reshape(-1)
            return L.G_RECT_EXTRUDED, [hx, hy, cx, cy, ex, ey], box(hx, hy, 0., 0.)
        if name == 'RectPlateGM':
            hx, hy = np.asarray(gm._half_dims, float).reshape(-1)
            return L.G_RECT_PLATE, [hx, hy], box(hx, hy, 0., 0.)
        if name in ('StraightCutRoundPlateGM', 'RoundPlateGM'):
            re = float(gm._Re)
            ri = -1.0 if gm._Ri is None else float(gm._Ri)
            xc = getattr(gm, 'x_cut', None)
            xc = inf if xc is None else float(xc)
            return L.G_ROUND_PLATE, [re, ri, xc], box(re, re, 0., 0.)
        if name == 'TriangularFace':
            v = np.asarray(gm._verts, float)
            if v.shape != (3, 2):
                raise ValueError("TriangularFace vertices must be a 3x2 array")
            pts = np.stack([np.zeros(3), v[:, 0], v[:, 1]], axis=1)
            return L.G_TRIANGLE, list(v[:, 0]) + list(v[:, 1]), (pts.min(axis=1), pts.max(axis=1))
        if name in ('FlatGeometryManager', 'FiniteFlatGM'):
            return L.G_FLAT_INFINITE, [], None
        if name == 'ParabolicDishGM':
            return L.G_PARAB_DISH, [gm.a, gm.b, gm._h], box(gm._R, gm._R, 0., gm._h)
        if name == 'HexagonalParabolicDishGM':
            r = float(gm._R)
            return L.G_PARAB_HEX, [gm.a, gm.b, r], box(r, r, 0., max(gm.a, gm.b) * 2 * r * r)
        if name == 'RectangularParabolicDishGM':
            if getattr(gm, '_off_axis_normal', None) is not None:
                break
            w, h = float(gm._w), float(gm._h)
            return L.G_PARAB_RECT, [gm.a, gm.b, w, h], box(w, h, 0., gm.a * w * w + gm.b * h * h)
        if name == 'ParabolicTroughGM':
            hl, h = gm._l / 2., float(gm._h)
            return L.G_PARAB_TROUGH, [gm.a, hl, h], box(np.sqrt(h / gm.a), hl, 0., h)
        if name == 'ParabolicCylinder':
            return L.G_PARAB_CYL, [gm.a], None
        if name == 'Paraboloid':
            return L.G_PARABOLOID, [gm.a, gm.b], None
        if name == 'SphericalRectFacet':
            r = float(gm._rad)
            return L.G_SPHERE_RECT, [r, gm.lx / 2., gm.ly / 2.], box(min(r, gm.lx / 2.), min(r, gm.ly / 2.), -r, 0.)
        if name == 'HemisphereGM':
            r = float(gm._rad)
            return L.G_HEMISPHERE, [r], box(r, r, -r, 0.)
        if name == 'CutSphereGM':
            r = float(gm._rad)
            bound = gm._bound
            if bound is None:
                return L.G_SPHERE, [r], box(r, r, -r, r)
            bnames = _mro_names(bound)
            if 'BoundaryPlane' in bnames:
                # CutSphereGM.find_intersections moves the bound into the surface's frame first (:184-186);
                # in_bounds is inv(frame)[2] . [p; 1] >= 0 (boundary_shape.py:161-164)
                row = np.linalg.inv(np.dot(np.asarray(surface._temp_frame, float), np.asarray(bound.get_transform(), float)))[2]
                return L.G_SPHERE_CUT, [r] + list(row) + [1.], box(r, r, -r, r)
            if 'BoundaryCylinder' in bnames:
                raise NotImplementedError(
                    "CutSphereGM bounded by a BoundaryCylinder: the reference itself raises ValueError there "
                    "(boundary_shape.py:146-147 stacks a (3, 2) array on 3 ones), there is no behaviour to reproduce")
            if 'BoundarySphere' not in bnames:
                raise NotImplementedError(
                    f"CutSphereGM bounding volume {type(bound).__name__} is not supported on the GPU "
                    "(BoundarySphere and BoundaryPlane are)")
            # BoundarySphere never moves its centre with the frame (boundary_shape.py:101)
            c = np.asarray(bound._temp_loc, float)[:3]
            rb = float(bound._radius)
            return L.G_SPHERE_CUT, [r] + list(c) + [rb, 0.], ('global', c - rb, c + rb)
        if name == 'SphericalGM':
            r = float(gm._rad)
            return L.G_SPHERE, [r], box(r, r, -r, r)
        if name == 'FiniteCone':
            return L.G_CONE_FINITE, [gm.c, gm.a, gm.h], box(gm.r, gm.r, 0., gm.h)
        if name == 'RectCutConicalFrustum':
            rm = max(gm.r1, gm.r2)
            hw, hh = np.asarray(gm.half_dims, float)
            return (L.G_FRUSTUM_RECT, [gm.c, gm.a, gm.zmin, gm.zmax, hw, hh],
                    (np.array([-min(rm, hw), -min(rm, hh), gm.zmin]), np.array([min(rm, hw), min(rm, hh), gm.zmax])))
        if name == 'ConicalFrustum':
            rm = max(gm.r1, gm.r2)
            return L.G_FRUSTUM, [gm.c, gm.a, gm.zmin, gm.zmax], box(rm, rm, gm.zmin, gm.zmax)
        if name == 'InfiniteCone':
            return L.G_CONE_INFINITE, [gm.c, gm.a], None
        if name == 'RectCutCylinder':
            r, hh_ = float(gm._R), float(gm._half_h)
            hw, hh = np.asarray(gm.half_dims, float)
            return L.G_CYL_RECT, [r, hh_, hw, hh], box(min(r, hw), min(r, hh), -hh_, hh_)
        if name == 'FiniteCylinder':
            r, hh_ = float(gm._R), float(gm._half_h)
            return L.G_CYL_FINITE, [r, hh_, float(gm._ang_range[0]), float(gm._ang_range[1])], box(r, r, -hh_, hh_)
        if name == 'InfiniteCylinder':
            return L.G_CYL_INFINITE, [float(gm._R)], None
        if name == 'RectFlatQuadricSurfaceGM':
            w, h = float(gm._w), float(gm._h)
            m = abs(gm.a) * w * w + abs(gm.b) * h * h + abs(gm.c) * w * h + abs(gm.d) * w + abs(gm.e) * h
            return (L.G_QUADRIC_RECT, [gm.a, gm.b, gm.c, gm.d, gm.e, gm.f, w, h], box(w, h, gm.f - m, gm.f + m))
        if name == 'FlatQuadricSurfaceGM':
            return L.G_QUADRIC_FLAT, [gm.a, gm.b, gm.c, gm.d, gm.e, gm.f], None
    raise NotImplementedError(
        f"geometry manager {type(gm).__name__} has no GPU kernel (no CPU fallback); "
        "see SURVEY.md section 8 for the supported classes")


def _global_aabb(frame, local_box):
    if local_box is None:
        return None
    if len(local_box) == 3 and isinstance(local_box[0], str):
        lo, hi = local_box[1], local_box[2]
    else:
        lo, hi = local_box
        c = np.ones((4, 8))
        for k in range(8):
            c[0, k] = hi[0] if k & 1 else lo[0]
            c[1, k] = hi[1] if k & 2 else lo[1]
            c[2, k] = hi[2] if k & 4 else lo[2]
        g = np.dot(frame, c)[:3]
        lo, hi = g.min(axis=1), g.max(axis=1)
    pad = 1e-6 * (1.0 + np.abs(np.concatenate([lo, hi])).max())
    return lo - pad, hi + pad


# ----------------------------------------------------------------- optics
_ACC_FIELDS = {  # shorthand -> (list attribute on the accountant, source)
    'Absorber': ('_absorbed', 'absorbed'),
    'Attenuator': ('_attenuated', 'attenuated'),
    'Receptor': ('_received', 'e_in'),
    'Scatterer': ('_scattered', 'e_out'),
    'Location': ('_hits', 'hits'),
    'Directional': ('_directions', 'dirs_in'),
    'Normal': ('_normals', 'normals'),
    'Spectral': ('_wavelengths', 'wavelengths'),
    # PolychromaticAccountant keeps two lists: the children's wavelengths and the spectra lost at the hit
    'Polychromatic': ('_spectra', 'lost_spectra'),
}


def _material_table(material):
    """(wavelengths, [n..., k...]) knots of a material's complex index.  Materials that
    interpolate a table (scipy interp1d in the reference, OpticalMaterialTable here) give
    their own knots, so the kernel reproduces them exactly; analytic models are sampled."""
    if type(material).__name__ == 'Air':           # m = 1 at every wavelength (optical_constants.py:143-155)
        return np.array([0., 1.]), np.array([1., 1., 0., 0.])
    if hasattr(material, 'lambdas') and hasattr(material, 'n') and hasattr(material, 'k'):
        wl, n, k = material.lambdas, material.n, material.k
    elif hasattr(material, 'm_func') and hasattr(material.m_func, 'x'):
        wl = np.asarray(material.m_func.x, float)
        m = np.asarray(material.m_func.y)
        n, k = m.real, m.imag
    else:
        wl = np.linspace(material.l_min, material.l_max, 4096)
        m = np.asarray(material.m(wl))
        n, k = m.real, m.imag
    wl = np.asarray(wl, float)
    order = np.argsort(wl)
    return wl[order], np.concatenate([np.asarray(n, float)[order], np.asarray(k, float)[order]])


def _check_refractive_accountants(base, accountants):
    """Accountants on a refractive surface.  Accountant.count works on the child bundle as the callable returns it,
    [reflected][refracted]: in ``single_ray`` mode it has one ray per hit and the reference subtracts it, in that order,
    from the hit-ordered incoming energies (optics_callables.py:1638-1643) -- reproduced as it is; when every hit
    emits two children the shapes differ and the reference raises ValueError."""
    if not accountants:
        return
    if not base._single_ray:
        raise NotImplementedError(
            "refractive optics with accountants and single_ray=False: the reference raises ValueError there "
            "(Accountant.count subtracts a child bundle of another length, optics_callables.py:1638-1643)")
    for _, _, src in accountants:
        if src == 'normals' and getattr(base, '_sigma', None) is not None:
            raise NotImplementedError(
                "NormalAccountant on a rough refractive surface: the reference's _refract_dirs perturbs the geometry "
                "manager's normals in place before the accountant reads them")
        if src not in ('absorbed', 'e_in', 'e_out', 'hits', 'dirs_in', 'normals', 'wavelengths'):
            raise NotImplementedError(f"accountant column {src!r} on a refractive surface")


def _optics_record(opt):
    """-> dict(okind, op, flags, table, accountants, draws); a BiFacial manager gives its front side's record
    with the back side's under 'back'."""
    if 'BiFacial' in _mro_names(opt):
        sides = []
        for side in (opt.OpticsCallable_front, opt.OpticsCallable_back):
            if 'BiFacial' in _mro_names(side):
                raise NotImplementedError("BiFacial inside BiFacial")
            r = _optics_record(side)
            if r['okind'] in (L.O_PERIODIC, L.O_REFRACTIVE):
                raise NotImplementedError("two-children optics (periodic, refractive) inside BiFacial")
            if r['okind'] == L.O_LAMB_PIECEWISE_SPEC:
                raise NotImplementedError(
                    "Lambertian_piecewise_Specular_directional_axisymmetric_piecewise inside BiFacial")
            sides.append(r)
        rec = dict(sides[0])
        rec['back'] = sides[1]
        rec['instance'] = opt
        return rec
    accountants = []
    base = opt
    if hasattr(opt, 'accountants') and hasattr(opt, '_opt'):
        base = opt._opt
        for a in opt.accountants:
            sh = getattr(a, 'shorthand', None)
            if sh not in _ACC_FIELDS:
                raise NotImplementedError(
                    f"accountant {type(a).__name__} of {type(opt).__name__} has no GPU counterpart")
            accountants.append((a,) + _ACC_FIELDS[sh])
            if sh == 'Polychromatic':
                accountants.append((a, '_wavelengths', 'wavelengths'))
    elif hasattr(opt, 'get_all_hits'):
        raise NotImplementedError(f"optics manager {type(opt).__name__}: unknown accounting wrapper")
    rec = dict(okind=None, op=[0.] * 8, flags=0, table=None, table_len=None, accountants=accountants,
               needs_wavelengths=False, instance=opt)
    # classes whose __call__ the kernels reproduce; anything deriving from one of them with
    # its own __call__ (e.g. Reflective_IAM, SemiLambertian, a user subclass) must not be
    # traced as its parent
    known = ('Transparent', 'OneSidedReflective', 'Reflective', 'OneSidedRealReflective', 'RealReflective',
             'Lambertian', 'LambertianSpecular', 'Lambertian_directional_axisymmetric_piecewise',
             'LambertianSpecular_directional_axisymmetric_piecewise',
             'Lambertian_directional_axisymmetric_piecewise_spectral', 'Reflective_spectral',
             'Lambertian_IAM', 'RealReflective_IAM', 'LambertianSpecular_IAM', 'LambertianAbsorbant',
             'Lambertian_piecewise_Specular_directional_axisymmetric_piecewise', 'PeriodicBoundary',
             'FresnelConductorHomogenous', 'RefractiveHomogenous', 'RefractiveTransmissiveHomogenous',
             'Lambertian_directional_axisymmetric_piecewise_Polychromatic', 'Refractive', 'RefractiveAbsorbant')
    for cls in type(base).__mro__:
        if cls.__name__ in known:
            break
        if '__call__' in vars(cls) and cls.__name__ != '_OpticsDescription':
            raise NotImplementedError(
                f"optics manager {type(base).__name__}: {cls.__name__}.__call__ has no GPU kernel "
                "(no CPU fallback); see SURVEY.md section 8 for the supported classes")
    for name in _mro_names(base):
        if name == 'Lambertian_IAM':
            rec['okind'] = L.O_LAMBERTIAN
            rec['op'][:2] = [_scalar(base._abs, 'absorptivity'), _scalar(base._ang_range, 'ang_range')]
            rec['op'][4:6] = [_scalar(base.a_r, 'a_r'), _scalar(base.c, 'c')]
            rec['flags'] |= L.F_IAM
        elif name == 'RealReflective_IAM':
            rec['okind'] = L.O_REAL_REFLECTIVE
            rec['op'][:3] = [_scalar(base._abs, 'absorptivity'), _scalar(base._sig, 'sigma'),
                             1.0 if base.bi_var == True else 0.0]  # noqa: E712
            rec['op'][4:6] = [_scalar(base.a_r, 'a_r'), _scalar(base.c, 'c')]
            rec['flags'] |= L.F_IAM
        elif name == 'LambertianSpecular_IAM':
            # optics_callables.py:606-607: the modifier sees a zeroed direction array, cos(AOI) = 0,
            # so the energy factor is exactly 1 whatever absorptivity and a_r are
            rec['okind'] = L.O_LAMBERTIAN_SPECULAR
            rec['op'][:2] = [0., _scalar(base.specularity, 'specularity')]
        elif name == 'Lambertian_piecewise_Specular_directional_axisymmetric_piecewise':
            rec['okind'] = L.O_LAMB_PIECEWISE_SPEC
            th = np.asarray(base.thetas, float)
            rec['table'] = (th, np.concatenate([np.asarray(base.abs_th, float), np.asarray(base.spec_th, float)]))
            rec['table_len'] = len(th)
        elif name == 'LambertianAbsorbant':
            a_c = base.a_c
            if a_c is None or np.ndim(a_c) != 1 or len(a_c) != 1:
                raise NotImplementedError(
                    "LambertianAbsorbant: only one imposed attenuation coefficient (a one-element sequence) "
                    "is supported; per-medium coefficients need the refractive family (SURVEY 8f N4)")
            rec['okind'] = L.O_LAMBERTIAN_ABSORBANT
            rec['op'][:4] = [_scalar(base._absorptivity, 'absorptivity'), _scalar(base._ang_range, 'ang_range'),
                             float(a_c[0]), _scalar(base._scaling, 'scaling')]
        elif name == 'FresnelConductorHomogenous':
            rec['okind'] = L.O_FRESNEL_CONDUCTOR
            rec['op'][0] = _scalar(base._n1, 'n1')
            rec['table'] = _material_table(base._material)
            rec['needs_wavelengths'] = True
        elif name in ('RefractiveAbsorbant', 'Refractive') and not hasattr(base, '_ref_idxs'):
            # media given by material tables, complex indices (:726-858, :908-944)
            _check_refractive_accountants(base, accountants)
            rec['okind'] = L.O_REFRACTIVE
            rec['flags'] |= L.F_TABULATED
            sigma = base._sigma
            if sigma is not None and sigma < 0.:
                raise ValueError("sigma must not be negative")
            rec['op'][2] = 1.0 if base._single_ray else 0.0
            rec['op'][3] = -1.0 if sigma is None else _scalar(sigma, 'sigma')
            wl_a, nk_a = _material_table(base._materials[0])
            wl_b, nk_b = _material_table(base._materials[1])
            rec['table'] = (wl_a, nk_a)                     # material 1: lambdas, then n, then k
            rec['table2'] = np.concatenate([wl_b, nk_b])    # material 2 goes to the pool too: op[0] = offset, op[1] = knots
            rec['table2_len'] = len(wl_b)
            rec['needs_wavelengths'] = True
            if name == 'RefractiveAbsorbant':
                if base.a_c is not None:
                    raise NotImplementedError(
                        "RefractiveAbsorbant without attenuation coefficients: the reference raises AttributeError on "
                        "the first hit (Absorbant.attenuate reads self._ref_idxs, optics_callables.py:883); give "
                        "attenuation_coefficient_1 (any value: the attenuation then comes from the imaginary indices)")
                rec['op'][4] = 1.0
                rec['op'][7] = _scalar(base._scaling, 'scaling')
        elif name in ('RefractiveTransmissiveHomogenous', 'RefractiveHomogenous'):
            _check_refractive_accountants(base, accountants)
            n_a, n_b = base._ref_idxs
            if np.iscomplexobj(np.asarray([n_a, n_b])):
                raise NotImplementedError("refractive optics: complex refractive indices (attenuation from the "
                                          "imaginary part) are outside the GPU path")
            rec['okind'] = L.O_REFRACTIVE
            sigma = base._sigma
            rec['op'][:4] = [_scalar(n_a, 'n1'), _scalar(n_b, 'n2'), 1.0 if base._single_ray else 0.0,
                             -1.0 if sigma is None else _scalar(sigma, 'sigma')]
            if sigma is not None and sigma < 0.:
                raise ValueError("sigma must not be negative")
            if name == 'RefractiveTransmissiveHomogenous':
                a_c = base.a_c
                if a_c is None or np.ndim(a_c) != 1 or len(a_c) not in (1, 2):
                    raise NotImplementedError(
                        "RefractiveTransmissiveHomogenous: one or two imposed attenuation coefficients are "
                        "supported (attenuation from complex indices and wavelengths is not)")
                rec['op'][4] = float(len(a_c))
                rec['op'][5] = float(a_c[0])
                rec['op'][6] = float(a_c[-1])
                rec['op'][7] = _scalar(base._scaling, 'scaling')
        elif name == 'PeriodicBoundary':
            if accountants:
                raise NotImplementedError(
                    "PeriodicBoundary with accountants: the reference's Accountant.count sees 2 children per "
                    "hit and fails on the shape mismatch; not reproduced")
            rec['okind'] = L.O_PERIODIC
            rec['op'][0] = _scalar(base.period, 'period')
        elif name == 'Transparent':
            rec['okind'] = L.O_TRANSPARENT
        elif name in ('OneSidedReflective', 'Reflective'):
            rec['okind'] = L.O_REFLECTIVE
            rec['op'][0] = _scalar(base._abs, 'absorptivity')
        elif name in ('OneSidedRealReflective', 'RealReflective'):
            rec['okind'] = L.O_REAL_REFLECTIVE
            rec['op'][:3] = [_scalar(base._abs, 'absorptivity'), _scalar(base._sig, 'sigma'),
                             1.0 if base.bi_var == True else 0.0]  # noqa: E712 (reference compares with ==)
        elif name == 'Lambertian':
            rec['okind'] = L.O_LAMBERTIAN
            rec['op'][:2] = [_scalar(base._abs, 'absorptivity'), _scalar(base._ang_range, 'ang_range')]
        elif name == 'LambertianSpecular':
            rec['okind'] = L.O_LAMBERTIAN_SPECULAR
            rec['op'][:2] = [_scalar(base._abs, 'absorptivity'), _scalar(base.specularity, 'specularity')]
        elif name == 'Lambertian_directional_axisymmetric_piecewise':
            rec['okind'] = L.O_LAMB_DIRECTIONAL
            rec['table'] = (np.asarray(base.thetas, float), np.asarray(base.abs_th, float))
        elif name == 'LambertianSpecular_directional_axisymmetric_piecewise':
            rec['okind'] = L.O_LAMBSPEC_DIRECTIONAL
            rec['op'][0] = _scalar(base.specularity, 'specularity')
            rec['table'] = (np.asarray(base.thetas, float), np.asarray(base.abs_th, float))
        elif name == 'Lambertian_directional_axisymmetric_piecewise_spectral':
            rec['okind'] = L.O_LAMB_DIRECTIONAL_SPECTRAL
            if hasattr(base, 'interpolator'):      # reference object: scipy RegularGridInterpolator
                th, wl = [np.asarray(g, float) for g in base.interpolator.grid]
                val = np.asarray(base.interpolator.values, float)
            else:
                th, wl, val = base.thetas, base.wavelengths, base.absorptance
            rec['op'][1] = float(len(wl))
            rec['table'] = (np.asarray(th, float), np.concatenate([np.asarray(wl, float), np.ravel(val)]))
            rec['needs_wavelengths'] = True
        elif name == 'Lambertian_directional_axisymmetric_piecewise_Polychromatic':
            # the optics kernel sees a Lambertian that absorbs nothing (same draws, :415-416); trc_spectra absorbs
            # the spectrum and integrates the energy
            rec['okind'] = L.O_LAMBERTIAN
            rec['op'][:2] = [0., np.pi / 2.]
            rec['flags'] |= L.F_POLY
            if hasattr(base, 'interpolator'):      # reference object: scipy RegularGridInterpolator
                th, wl = [np.asarray(g, float) for g in base.interpolator.grid]
                val = np.asarray(base.interpolator.values, float)
            else:
                th, wl, val = base.thetas, base.wavelengths, base.absorptance
            rec['op'][6] = float(len(wl))
            rec['table'] = (np.asarray(th, float), np.concatenate([np.asarray(wl, float), np.ravel(val)]))
            rec['needs_spectra'] = True
        elif name == 'Reflective_spectral':
            rec['okind'] = L.O_REFLECTIVE_SPECTRAL
            rec['table'] = (np.asarray(base._wavelengths, float), np.asarray(base._absorptances, float))
            rec['needs_wavelengths'] = True
        else:
            continue
        if name.startswith('OneSided') or 'OneSidedReflective' in _mro_names(base) \
                or 'OneSidedRealReflective' in _mro_names(base):
            rec['flags'] |= L.F_ONE_SIDED
        return rec
    raise NotImplementedError(
        f"optics manager {type(base).__name__} has no GPU kernel (no CPU fallback); "
        "see SURVEY.md section 8 for the supported classes")


# -------------------------------------------------------------------- BVH
def build_bvh(lo, hi, leaf_size=1):
    """Binned-SAH binary BVH over n AABBs, built by the native host builder
    (csrc/bvh_build.cu; no GPU needed).  Children of an inner node are adjacent
    (``first``, ``first + 1``).  Returns (nodes structured array, leaf order int32)."""
    import ctypes as C
    lib = L.load()
    lo = np.ascontiguousarray(lo, dtype=np.float64)
    hi = np.ascontiguousarray(hi, dtype=np.float64)
    n = lo.shape[0]
    nodes = np.zeros(2 * n, dtype=L.BVH_NODE_DTYPE)
    order = np.zeros(n, dtype=np.int32)
    n_nodes = C.c_int32(0)
    L.check(lib.trc_bvh_build(lo.ctypes.data_as(C.c_void_p), hi.ctypes.data_as(C.c_void_p), n,
                              int(min(max(leaf_size, 1), L.BVH_MAX_LEAF)),
                              nodes.ctypes.data_as(C.c_void_p), len(nodes), C.byref(n_nodes),
                              order.ctypes.data_as(C.c_void_p)))
    return nodes[:n_nodes.value].copy(), order


class CompiledScene(object):
    """Flat tables for one Assembly + the device handle."""

    def __init__(self, assembly, leaf_size=1, upload=True):
        self.assembly = assembly
        self.handle = None
        surfaces = assembly.get_surfaces()
        if len(surfaces) == 0:
            raise ValueError("assembly has no surfaces")
        self._check_objects(assembly)
        self.surfaces = surfaces
        S = len(surfaces)
        recs = np.zeros(S, dtype=L.SURFACE_DTYPE)
        self._back = {}                  # surface -> (okind, op, flags, table_off, table_len) of a BiFacial back side
        self._tables = []
        self._table_len = 0
        self.optics = []
        self.aabb_lo = np.full((S, 3), np.nan)
        self.aabb_hi = np.full((S, 3), np.nan)
        self.mesh_segments = []          # (first surface, one past the last, TriangleMeshObject)
        self._seen_opt = {}
        objs = assembly.get_objects()
        if any(getattr(o, 'is_triangle_mesh', False) for o in objs):
            i = 0
            for obj in objs:
                if getattr(obj, 'is_triangle_mesh', False):
                    n = obj.num_triangles()
                    self._fill_mesh(recs, i, obj)
                    self.mesh_segments.append((i, i + n, obj))
                    i += n
                else:
                    for s in obj.get_surfaces():
                        self._fill_one(recs, i, s)
                        i += 1
            assert i == S
        else:
            for i, s in enumerate(surfaces):
                self._fill_one(recs, i, s)
        tables = self._tables
        in_mesh = np.zeros(S, dtype=bool)
        for a, b, _ in self.mesh_segments:
            in_mesh[a:b] = True
        # surfaces whose accountants are fed one by one (TracerEngine._account); a mesh gets one chunk per bounce
        self.account_list = [i for i in np.nonzero(~in_mesh)[0]
                             if self.optics[i]['accountants'] or self.optics[i].get('back', {}).get('accountants')]
        self.records = recs
        # BiFacial managers: a second table of the same scene with the back sides' optics (same geometry, same pool)
        self.has_bifacial = bool(self._back)
        self.records_back = None
        self.handle_back = None
        if self.has_bifacial:
            self.records_back = recs.copy()
            self.bifacial_mask = np.zeros(S, dtype=bool)
            for i, (okind, op, flags, toff, tlen) in self._back.items():
                rb = self.records_back[i]
                rb['okind'], rb['op'], rb['table_off'], rb['table_len'] = okind, op, toff, tlen
                rb['flags'] = (int(recs[i]['flags']) & L.F_UNBOUNDED) | flags
                self.bifacial_mask[i] = True
        self.tables = np.concatenate(tables) if tables else np.zeros(0)
        uniq = list({id(o): o for o in self.optics}.values())        # a mesh shares one record
        uniq = uniq + [o['back'] for o in uniq if o.get('back')]
        self.needs_wavelengths = any(o['needs_wavelengths'] for o in uniq)
        self.needs_spectra = any(o.get('needs_spectra') for o in uniq)
        self.has_accountants = any(o['accountants'] for o in uniq)
        self.children_per_hit = np.where((recs['okind'] == L.O_PERIODIC) | (recs['okind'] == L.O_REFRACTIVE),
                                         2, 1).astype(np.int64)
        self.has_refractive = bool((recs['okind'] == L.O_REFRACTIVE).any())
        self.has_tabulated = bool(((recs['okind'] == L.O_REFRACTIVE) & ((recs['flags'] & L.F_TABULATED) != 0)).any())
        self.has_split = bool((self.children_per_hit > 1).any())
        if self.has_split and self.has_bifacial:
            raise NotImplementedError("BiFacial managers in a scene with two-children optics (periodic, refractive)")
        self.needs_theta_for_draws = any(o['okind'] == L.O_LAMB_PIECEWISE_SPEC for o in uniq)
        self.accountant_sources = set(src for o in uniq for (_, _, src) in o['accountants'])
        self.has_poly_accountant = 'lost_spectra' in self.accountant_sources
        self.unbounded = np.nonzero((recs['flags'] & L.F_UNBOUNDED) != 0)[0].astype(np.int32)
        self.bounded = np.nonzero((recs['flags'] & L.F_UNBOUNDED) == 0)[0].astype(np.int32)
        self.leaf_size = leaf_size
        self.bvh_nodes = None
        if len(self.bounded):
            self.bvh_nodes, order = build_bvh(self.aabb_lo[self.bounded], self.aabb_hi[self.bounded], leaf_size)
            self.bvh_leaf_surfaces = np.ascontiguousarray(self.bounded[order])
        else:
            self.bvh_nodes = np.zeros(0, dtype=L.BVH_NODE_DTYPE)
            self.bvh_leaf_surfaces = np.zeros(0, np.int32)
        if upload:
            self._upload()

    def _add_table(self, arr):
        off = self._table_len
        arr = np.asarray(arr, float)
        self._tables.append(arr)
        self._table_len += len(arr)
        return off

    def _fill_one(self, recs, i, s):
        frame = np.asarray(s._temp_frame, dtype=float)
        kind, gp, lbox = _geom_record(s._geom, s)
        orec = _optics_record(s._opt)
        seen_opt = self._seen_opt
        if orec['accountants'] and id(s._opt) in seen_opt:
            raise NotImplementedError(
                f"optics manager instance shared by surfaces {seen_opt[id(s._opt)]} and {i}: "
                "accountant data would be merged (optics_callables.py:82); use copy_optical_manager")
        seen_opt[id(s._opt)] = i
        r = recs[i]
        r['frame'] = frame[:3].reshape(-1)
        r['inv'] = np.linalg.inv(frame)[:3].reshape(-1)
        if isinstance(gp, tuple):                   # aperture data in the table pool: gp[slot] = its offset
            gp, pool, slot = gp
            gp[slot] = float(self._add_table(pool))
        r['gp'][:len(gp)] = gp
        r['op'] = orec['op']
        r['gkind'], r['okind'] = kind, orec['okind']
        flags = orec['flags']
        box = _global_aabb(frame, lbox)
        if box is None:
            flags |= L.F_UNBOUNDED
        else:
            self.aabb_lo[i], self.aabb_hi[i] = box
        r['flags'] = flags
        r['table_off'], r['table_len'] = -1, 0
        if orec['table'] is not None:
            x, y = orec['table']
            r['table_off'], r['table_len'] = self._add_table(x), len(x)
            self._add_table(y)
        if orec.get('table2') is not None:
            r['op'][0], r['op'][1] = float(self._add_table(orec['table2'])), float(orec['table2_len'])
        back = orec.get('back')
        if back is not None:
            toff, tlen = -1, 0
            if back['table'] is not None:
                x, y = back['table']
                toff, tlen = self._add_table(x), len(x)
                self._add_table(y)
            self._back[i] = (back['okind'], back['op'], back['flags'], toff, tlen)
            if back['accountants'] and id(back['instance']) in seen_opt:
                raise NotImplementedError("optics manager instance shared between BiFacial sides / surfaces")
            seen_opt[id(back['instance'])] = i
        self.optics.append(orec)

    def _fill_mesh(self, recs, i0, mesh):
        """Records of a whole TriangleMeshObject with array operations (same content as
        _fill_one on each of its faces: TRC_G_TRIANGLE, the mesh's optics on every face)."""
        frames = np.asarray(mesh.mesh_frames(), dtype=float)
        verts = np.asarray(mesh.mesh_vertices(), dtype=float)
        n = len(frames)
        orec = _optics_record(mesh.mesh_optics())
        if orec.get('back') is not None:
            raise NotImplementedError("BiFacial optics on a TriangleMeshObject")
        if orec['okind'] in (L.O_PERIODIC, L.O_REFRACTIVE):
            raise NotImplementedError("two-children optics (periodic, refractive) on a TriangleMeshObject")
        orec['mesh'] = mesh
        sl = slice(i0, i0 + n)
        recs['frame'][sl] = frames[:, :3, :].reshape(n, 12)
        recs['inv'][sl] = np.linalg.inv(frames)[:, :3, :].reshape(n, 12)
        recs['gp'][sl, 0:3] = verts[:, :, 0]
        recs['gp'][sl, 3:6] = verts[:, :, 1]
        recs['op'][sl] = np.asarray(orec['op'], float)
        recs['gkind'][sl] = L.G_TRIANGLE
        recs['okind'][sl] = orec['okind']
        recs['flags'][sl] = orec['flags']
        recs['table_off'][sl], recs['table_len'][sl] = -1, 0
        if orec['table'] is not None:               # one copy of the table serves every face
            x, y = orec['table']
            recs['table_off'][sl], recs['table_len'][sl] = self._add_table(x), len(x)
            self._add_table(y)
        rot, p0 = frames[:, :3, :3], frames[:, :3, 3]
        p1 = p0 + np.einsum('nij,nj->ni', rot, verts[:, :, 0])
        p2 = p0 + np.einsum('nij,nj->ni', rot, verts[:, :, 1])
        lo = np.minimum(np.minimum(p0, p1), p2)
        hi = np.maximum(np.maximum(p0, p1), p2)
        pad = 1e-6 * (1.0 + np.maximum(np.abs(lo).max(axis=1), np.abs(hi).max(axis=1)))[:, None]
        self.aabb_lo[sl], self.aabb_hi[sl] = lo - pad, hi + pad
        self.optics.extend([orec] * n)

    @staticmethod
    def _check_objects(assembly):
        for obj in assembly.get_objects():
            for hook in ('own_rays', 'surfaces_for_next_iteration'):
                fn = getattr(obj, hook, None)
                qual = getattr(getattr(fn, '__func__', fn), '__qualname__', '')
                if fn is not None and qual != 'AssembledObject.' + hook:
                    raise NotImplementedError(
                        f"{type(obj).__name__} overrides {hook}(): per-object relevance hooks "
                        "(object.py:81-112) are not honoured by the GPU engine")

    def _upload(self):
        import ctypes as C
        lib = L.load()
        h = C.c_void_p()
        tab = self.tables
        L.check(lib.trc_scene_create(self.records.ctypes.data_as(C.c_void_p), len(self.records),
                                     tab.ctypes.data_as(C.c_void_p) if tab.size else None, int(tab.size),
                                     C.byref(h)))
        self.handle = h
        self._upload_bvh()
        if self.has_bifacial:              # optics-only copy: trc_optics reads its surface table and the shared pool
            hb = C.c_void_p()
            L.check(lib.trc_scene_create(self.records_back.ctypes.data_as(C.c_void_p), len(self.records_back),
                                         tab.ctypes.data_as(C.c_void_p) if tab.size else None, int(tab.size),
                                         C.byref(hb)))
            self.handle_back = hb

    def _upload_bvh(self):
        import ctypes as C
        lib = L.load()
        nodes, leaf_surfs = self.bvh_nodes, self.bvh_leaf_surfaces
        unb = np.ascontiguousarray(self.unbounded)
        L.check(lib.trc_scene_set_bvh(
            self.handle, nodes.ctypes.data_as(C.c_void_p) if len(nodes) else None, len(nodes),
            leaf_surfs.ctypes.data_as(C.c_void_p) if len(leaf_surfs) else None, len(leaf_surfs),
            unb.ctypes.data_as(C.c_void_p) if len(unb) else None, len(unb)))

    GRID_MAX_SURFACES = 1 << 16       # larger scenes keep the tree walk (cells would list dozens of faces)

    def set_shadow_grid(self, direction, half_angle):
        """Shadow grid for bundles whose directions lie within ``half_angle`` of ``direction`` (the first bounce
        of a solar source): trc_scene_set_shadow_grid.  Cached per (direction, angle); returns whether the
        nearest-hit kernel may be called with use_bvh = 2."""
        import ctypes as C
        key = None if direction is None else (tuple(np.round(np.asarray(direction, float), 12)), float(half_angle))
        if key == getattr(self, '_grid_key', ('unset',)):
            return self._grid_on
        lib = L.load()
        on = C.c_int32(0)
        if key is None or len(self.bounded) == 0 or len(self.bounded) > self.GRID_MAX_SURFACES:
            L.check(lib.trc_scene_set_shadow_grid(self.handle, 0, None, None, None, None, None, -1., C.byref(on)))
        else:
            lo = np.ascontiguousarray(self.aabb_lo[self.bounded], dtype=np.float64)
            hi = np.ascontiguousarray(self.aabb_hi[self.bounded], dtype=np.float64)
            quads = np.ascontiguousarray(self.planar_quads()[self.bounded], dtype=np.float64)
            ids = np.ascontiguousarray(self.bounded, dtype=np.int32)
            d = np.ascontiguousarray(direction, dtype=np.float64)
            L.check(lib.trc_scene_set_shadow_grid(
                self.handle, len(ids), lo.ctypes.data_as(C.c_void_p), hi.ctypes.data_as(C.c_void_p),
                quads.ctypes.data_as(C.c_void_p), ids.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p),
                float(half_angle), C.byref(on)))
        self._grid_key, self._grid_on = key, bool(on.value)
        return self._grid_on

    def bifacial_tensors(self, device):
        """(up (S, 3): +z of every surface frame in global coordinates, mask (S,) of the BiFacial surfaces) on device."""
        import torch
        cached = getattr(self, '_bifacial_dev', None)
        if cached is None or cached[0].device != device:
            up = np.ascontiguousarray(self.records['frame'][:, [2, 6, 10]], dtype=np.float64)
            cached = (torch.from_numpy(up).to(device), torch.from_numpy(self.bifacial_mask.copy()).to(device))
            self._bifacial_dev = cached
        return cached

    def planar_quads(self):
        """(S, 4, 3) corners, global coordinates, of the quadrilateral every flat plate / triangle lies in (NaN rows for
        the other kinds): their shadow in the shadow grid is the hull of the corners' shadows, not the box's."""
        r = self.records
        S = len(r)
        out = np.full((S, 4, 3), np.nan)
        frame = r['frame'].reshape(S, 3, 4)
        gp = r['gp']
        kind = r['gkind']
        local = np.zeros((S, 4, 3))
        rect = np.isin(kind, (L.G_RECT_PLATE, L.G_RECT_EXTRUDED, L.G_RECT_PERFORATED))
        sx, sy = np.array([-1., 1., 1., -1.]), np.array([-1., -1., 1., 1.])
        local[rect, :, 0] = gp[rect, 0:1] * sx
        local[rect, :, 1] = gp[rect, 1:2] * sy
        rnd = kind == L.G_ROUND_PLATE
        local[rnd, :, 0] = gp[rnd, 0:1] * sx
        local[rnd, :, 1] = gp[rnd, 0:1] * sy
        tri = kind == L.G_TRIANGLE
        # TriangularFace: p = p0 + u e0 + v e1 with e = R * verts (gp[0:3] = verts[:, 0], gp[3:6] = verts[:, 1])
        local[tri, 1, :] = gp[tri, 0:3]
        local[tri, 2, :] = gp[tri, 3:6]
        local[tri, 3, :] = gp[tri, 3:6]
        sel = rect | rnd | tri
        out[sel] = np.einsum('sij,skj->ski', frame[sel][:, :, :3], local[sel]) + frame[sel][:, None, :, 3]
        return out

    def seg_entries_needed(self, device):
        """(host list, device index tensor) of the slot-offset entries the engine reads on the host after a
        bounce: both ends of every mesh and of every surface with accountants."""
        import torch
        cached = getattr(self, '_seg_needed', None)
        if cached is None or cached[1].device != device:
            idx = set()
            for i0, i1, _ in self.mesh_segments:
                idx.update((i0, i1))
            for s_ in self.account_list:
                idx.update((int(s_), int(s_) + 1))
            idx = sorted(idx)
            cached = (idx, torch.tensor(idx, dtype=torch.int64, device=device))
            self._seg_needed = cached
        return cached

    @property
    def num_surfaces(self):
        return len(self.records)

    def close(self):
        if self.handle is not None:
            L.load().trc_scene_destroy(self.handle)
            self.handle = None
        if getattr(self, 'handle_back', None) is not None:
            L.load().trc_scene_destroy(self.handle_back)
            self.handle_back = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
