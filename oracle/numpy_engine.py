"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy) of the reference's bounce
loop, used as the parity checker for the CUDA path and as the timed CPU
baseline of bench.py.  Nothing under tracer_b200/ imports this file.

Parity pin: tests/test_oracle_golden.py checks this restatement against golden
vectors recorded from the *real* reference (tests/golden/make_golden.py imports
/root/reference and runs tracer.tracer_engine.TracerEngine with a fixed
numpy.random seed).  With the same seed this file consumes numpy.random in the
same order as the reference, so bundles match bit for bit on those scenes.

Every function cites the reference lines it restates (paths relative to the
reference tree).  The structure follows the reference on purpose -- a Python
loop over surfaces with NumPy vectorised over rays -- because that is the
algorithm whose CPU time bench.py reports; it is *not* how the GPU path works.
Scene objects are read by duck typing (class name along the MRO + the
reference's attribute names), so both reference-built and tracer_b200-built
scenes run.
"""
import numpy as N


def _names(obj):
    return [c.__name__ for c in type(obj).__mro__]


# ====================================================================== math
def rotation_to_z(vecs):
    """tracer/spatial_geometry.py:24-48"""
    vecs = N.atleast_2d(vecs)
    perp = N.hstack((vecs[:, 1][:, None], -vecs[:, 0][:, None], N.zeros((vecs.shape[0], 1))))
    perp[N.all(perp == 0., axis=1)] = N.r_[1., 0., 0.]
    perp /= N.sqrt(N.sum(perp ** 2., axis=1))[:, None]
    return N.concatenate((perp[..., None], N.cross(vecs, perp)[..., None], vecs[..., None]), axis=2)


def rotate_z_to_normal(vecs, normals):
    """ray_trace_utils/vector_manipulations.py:56-74 with general_axis_rotation
    (spatial_geometry.py:8-22) vectorised over rays; same arithmetic per ray."""
    n = vecs.shape[1]
    plane = N.vstack((-normals[1], normals[0], N.zeros(n)))          # cross(z, n)
    with N.errstate(invalid='ignore', divide='ignore'):
        plane = plane / N.sqrt(N.sum(plane ** 2, axis=0))
    nans = N.isnan(plane[0])
    plane[:, nans] = N.array([[1., 0., 0.]]).T
    ang = N.arccos(normals[2])
    s = N.round(N.sin(ang), decimals=14)
    c = N.round(N.cos(ang), decimals=14)
    v = 1 - c
    ax, ay, az = plane
    out = N.empty_like(vecs)
    x, y, z = vecs
    out[0] = (ax * ax * v + c) * x + (ax * ay * v - az * s) * y + (ax * az * v + ay * s) * z
    out[1] = (ay * ax * v + az * s) * x + (ay * ay * v + c) * y + (ay * az * v - ax * s) * z
    out[2] = (az * ax * v - ay * s) * x + (az * ay * v + ax * s) * y + (az * az * v + c) * z
    same = ang == 0.
    out[:, same] = vecs[:, same]
    return out


def reflections(ray_dirs, normals):
    """tracer/optics.py:145-157"""
    vertical = N.sum(ray_dirs * normals, axis=0) * normals
    return ray_dirs - 2. * vertical


def pillbox_directions(num_rays, ang_range):
    """tracer/sources.py:88-117 (draw order: phi, then the polar variate)"""
    xi1 = N.random.uniform(low=0., high=2. * N.pi, size=num_rays)
    if ang_range == 0.:
        dirs = N.zeros((3, num_rays))
        dirs[2] = 1.
        return dirs
    xi2 = N.random.uniform(size=num_rays)
    sinsqrt = N.sin(ang_range) * N.sqrt(xi2)
    return N.vstack((N.cos(xi1) * sinsqrt, N.sin(xi1) * sinsqrt, N.sqrt(1. - sinsqrt ** 2.)))


def _to_local(frame, pts):
    return N.dot(N.linalg.inv(frame), N.vstack((pts, N.ones(pts.shape[1]))))


# ================================================================== geometry
class Hit(object):
    """What the reference GM keeps between find_intersections and get_outgoing."""
    __slots__ = ('params', 'points', 'normal_fn')


def _flat(frame, v, d):
    """FlatGeometryManager.find_intersections, flat_surface.py:16-62"""
    n = v.shape[1]
    vv = v - frame[:3, 3][:, None]
    dt = N.dot(d.T, frame[:3, 2])
    unparallel = N.abs(dt) > 1e-7
    params = N.empty(n)
    params.fill(N.inf)
    vt = N.dot(frame[:3, 2], vv[:, unparallel])
    params[unparallel] = -vt / dt[unparallel]
    params[params < 1e-7] = N.inf
    return params, dt > 0.


def find_intersections(surface, v, d):
    """-> (params (n,), points (3,n) global hit points, normals(idx) callable)."""
    gm, frame = surface._geom, surface._temp_frame
    names = _names(gm)
    for name in names:
        if name in _FLAT:
            return _find_flat(name, gm, frame, v, d)
        if name in _QUADRIC:
            return _find_quadric(name, gm, frame, v, d)
    raise NotImplementedError("oracle: geometry manager %s" % type(gm).__name__)


_FLAT = ('PerforatedRectPlateGM', 'PerforatedPolygonGM', 'FlatSimplePolygonGM', 'ExtrudedRectPlateGM', 'RectPlateGM', 'StraightCutRoundPlateGM', 'RoundPlateGM', 'TriangularFace',
         'FlatGeometryManager', 'FiniteFlatGM')


def _in_a_hole(gm, local):
    """Strictly inside one of the circular perforations (centres (n,2), radii (n,))."""
    c = N.asarray(gm.extr_centers, float).reshape(-1, 2)
    dist = N.sqrt((local[0][:, None] - c[:, 0]) ** 2 + (local[1][:, None] - c[:, 1]) ** 2)
    return N.any(dist < N.asarray(gm.extr_radii, float).reshape(-1), axis=1)


def _in_polygon(profile, px, py):
    """Crossing number of FlatSimplePolygonGM.in_poly / intersect (polygon.py:31-63), edge by edge."""
    xs = N.append(profile[0], profile[0, 0])
    ys = N.append(profile[1], profile[1, 0])
    crossings = N.zeros(len(px), dtype=N.int64)
    for k in range(len(xs) - 1):
        xp0, xp1 = px <= xs[k], px <= xs[k + 1]
        yp0, yp1 = py <= ys[k], py <= ys[k + 1]
        across_y = yp0 ^ yp1
        hit = xp0 & xp1 & across_y
        doubt = (xp0 ^ xp1) & across_y
        if doubt.any():
            a = (ys[k + 1] - ys[k]) / (xs[k + 1] - xs[k])
            y0 = ys[k] - a * xs[k]
            hit = hit | (doubt & ((py - y0) / a >= px))
        crossings += hit
    return (crossings % 2).astype(bool)


def _find_flat(name, gm, frame, v, d):
    params, backside = _flat(frame, v, d)
    with N.errstate(invalid='ignore'):
        glob = v + params[None, :] * d                      # flat_surface.py:142-150
    if name not in ('FlatGeometryManager', 'FiniteFlatGM'):
        with N.errstate(invalid='ignore'):
            local = _to_local(frame, glob)                  # :160-164
            if name in ('RectPlateGM', 'ExtrudedRectPlateGM', 'PerforatedRectPlateGM'):
                params[N.any(N.abs(local[:2]) > gm._half_dims, axis=0)] = N.inf          # :209
                if name == 'PerforatedRectPlateGM':
                    params[_in_a_hole(gm, local)] = N.inf                                # :372-373
                if name == 'ExtrudedRectPlateGM':
                    params[N.all(N.abs(local[:2] - gm.extr_center) < gm.extr_half_dims, axis=0)] = N.inf   # :272
            elif name in ('FlatSimplePolygonGM', 'PerforatedPolygonGM'):                 # polygon.py:20-29, :150-160
                params[~_in_polygon(N.asarray(gm.profile, float), local[0], local[1])] = N.inf
                if name == 'PerforatedPolygonGM':
                    params[_in_a_hole(gm, local)] = N.inf
            elif name in ('RoundPlateGM', 'StraightCutRoundPlateGM'):
                r2 = N.sum(local[:2] ** 2., axis=0)
                params[r2 > gm._Re ** 2.] = N.inf                                        # :488
                if gm._Ri is not None:
                    params[r2 < gm._Ri ** 2.] = N.inf
                if name == 'StraightCutRoundPlateGM':
                    params[local[0] > gm.x_cut] = N.inf                                  # :558
            elif name == 'TriangularFace':                  # triangular_face.py:59-72
                glob_verts = N.dot(frame, N.vstack((gm._verts, N.array([1, 1]))))
                rel_glob = glob_verts[:3].T - frame[:3, 3]
                w = glob.T - frame[:3, 3]
                uv = N.dot(gm._verts[:, 0], gm._verts[:, 1])
                rel_dots = N.dot(w, rel_glob.T)
                norms_sq = N.sum(gm._verts ** 2, axis=0)
                bc = (uv * rel_dots[:, ::-1] - norms_sq[::-1] * rel_dots) / (uv ** 2 - norms_sq[0] * norms_sq[1])
                params[N.any(bc < 0, axis=1) | (bc.sum(axis=1) > 1.)] = N.inf

    def normals(idx):                                       # flat_surface.py:84-91
        norms = N.tile(frame[:3, 2].copy()[:, None], (1, len(idx)))
        norms[:, N.nonzero(backside[idx])[0]] *= -1
        return norms

    def points(idx):                                        # select_rays, :64-82 / :168-179
        return v[:, idx] + params[idx][None, :] * d[:, idx]
    return params, points, normals


_QUADRIC = ('ParabolicDishGM', 'HexagonalParabolicDishGM', 'RectangularParabolicDishGM', 'ParabolicTroughGM',
            'ParabolicCylinder', 'Paraboloid', 'SphericalRectFacet', 'HemisphereGM', 'CutSphereGM', 'SphericalGM',
            'FiniteCone', 'RectCutConicalFrustum', 'ConicalFrustum', 'InfiniteCone', 'RectCutCylinder',
            'FiniteCylinder', 'InfiniteCylinder', 'RectFlatQuadricSurfaceGM', 'FlatQuadricSurfaceGM',
            'EllipsoidGM', 'Ellipsoid')
_PARAB = ('ParabolicDishGM', 'HexagonalParabolicDishGM', 'RectangularParabolicDishGM', 'Paraboloid')
_PCYL = ('ParabolicTroughGM', 'ParabolicCylinder')
_SPH = ('SphericalRectFacet', 'HemisphereGM', 'CutSphereGM', 'SphericalGM')
_CONE = ('FiniteCone', 'RectCutConicalFrustum', 'ConicalFrustum', 'InfiniteCone')
_CYL = ('RectCutCylinder', 'FiniteCylinder', 'InfiniteCylinder')


def _abc(name, gm, frame, v, d):
    if name in _SPH:                                        # sphere_surface.py:52-68
        c = frame[:3, 3]
        A = (d ** 2).sum(axis=0)
        B = 2. * (d * (v - c[:, None])).sum(axis=0)
        C = ((v - c[:, None]) ** 2).sum(axis=0) - gm.get_radius() ** 2
        return A, B, C
    dl = N.dot(frame[:3, :3].T, d)
    vl = _to_local(frame, v)[:3]
    if name in _PARAB:                                      # paraboloid.py:28-42
        A = gm.a * dl[0] ** 2 + gm.b * dl[1] ** 2
        B = 2 * gm.a * dl[0] * vl[0] + 2 * gm.b * dl[1] * vl[1] - dl[2]
        C = gm.a * vl[0] ** 2 + gm.b * vl[1] ** 2 - vl[2]
    elif name in _PCYL:                                     # paraboloid.py:343-356
        A = gm.a * dl[0] ** 2
        B = 2 * gm.a * dl[0] * vl[0] - dl[2]
        C = gm.a * vl[0] ** 2 - vl[2]
    elif name in _CONE:                                     # cone.py:59-72
        A = dl[0] ** 2. + dl[1] ** 2. - (gm.c * dl[2]) ** 2.
        B = 2. * (vl[0] * dl[0] + vl[1] * dl[1] - gm.c ** 2. * (vl[2] - gm.a) * dl[2])
        C = vl[0] ** 2. + vl[1] ** 2. - (gm.c * (vl[2] - gm.a)) ** 2.
    elif name in _CYL:                                      # cylinder.py:32-57
        A = N.sum(dl[:2] ** 2, axis=0)
        B = 2. * N.sum(dl[:2] * vl[:2], axis=0)
        C = N.sum(vl[:2] ** 2, axis=0) - gm._R ** 2
    elif name in ('EllipsoidGM', 'Ellipsoid'):              # ellipsoid.py:20-34
        A = gm.a * dl[0] ** 2 + gm.b * dl[1] ** 2 + gm.c * dl[2] ** 2
        B = 2 * gm.a * dl[0] * vl[0] + 2 * gm.b * dl[1] * vl[1] + 2 * gm.c * dl[2] * vl[2]
        C = gm.a * vl[0] ** 2 + gm.b * vl[1] ** 2 + gm.c * vl[2] ** 2 - 1
    else:                                                   # quadratic_surface.py:46-61
        A = gm.a * dl[0] ** 2. + gm.b * dl[1] ** 2. + gm.c * dl[0] * dl[1]
        B = 2. * gm.a * dl[0] * vl[0] + 2. * gm.b * dl[1] * vl[1] + gm.c * (vl[0] * dl[1] + vl[1] * dl[0]) \
            + gm.d * dl[0] + gm.e * dl[1] - dl[2]
        C = gm.a * vl[0] ** 2 + gm.b * vl[1] ** 2 + gm.c * vl[0] * vl[1] + gm.d * vl[0] + gm.e * vl[1] + gm.f - vl[2]
    return A, B, C


def _default_select(prm):
    """QuadricGM._select_coords, quadric.py:115-144"""
    is_positive = prm >= 1e-6
    select = N.empty(prm.shape[1])
    select.fill(N.nan)
    select[N.logical_and(*is_positive)] = 1
    one_pos = N.logical_xor(*is_positive)
    select[one_pos] = N.nonzero(is_positive.T[one_pos, :])[1]
    return select


def _hitting_select(hitting):
    select = N.empty(hitting.shape[1])
    select.fill(N.nan)
    select[N.logical_and(*hitting)] = 1
    one = N.logical_xor(*hitting)
    select[one] = N.nonzero(hitting.T[one, :])[1]
    return select


# sphere_surface.py:138, :203, :227 pick the single valid root with
# ``N.nonzero(mask[:, one_hit])[0]``, which returns row indices sorted by ROW, not
# per ray: when the bundle mixes rays whose valid root is #0 with rays whose valid
# root is #1 the choices are assigned to the wrong rays (even negative params come
# out).  That is a bundle-order-dependent reference bug (DESIGN.md "excluded
# quirks"); the oracle and the GPU use the per-ray choice the code intends.  The
# flag reproduces the bug so the golden test can still pin every mask bit for bit.
REPRODUCE_SPHERE_SELECT_BUG = False


def _refine_select(select, inside, sphere_family=False):
    select[~N.logical_or(*inside)] = N.nan
    one = N.logical_xor(*inside)
    if sphere_family and REPRODUCE_SPHERE_SELECT_BUG:
        select[one] = N.nonzero(inside[:, one])[0]
    else:
        select[one] = N.nonzero(inside.T[one, :])[1]
    return select


def _select(name, gm, frame, coords, prm):
    """The per-class _select_coords overrides; coords (2,3,n) global, prm (2,n)."""
    if name in ('Paraboloid', 'SphericalGM', 'InfiniteCone', 'InfiniteCylinder', 'FlatQuadricSurfaceGM',
                'ParabolicCylinder', 'Ellipsoid'):
        return _default_select(prm)
    if name == 'CutSphereGM':                               # sphere_surface.py:188-204
        select = _default_select(prm)
        if gm._bound is None:
            return select
        gm._bound.transform_frame(frame)
        in_bd = N.array([gm._bound.in_bounds(coords[..., c]) for c in range(prm.shape[1])]).T.reshape(2, -1)
        in_bd = in_bd & (prm > 1e-6)
        return _refine_select(select, in_bd, True)
    inv = N.linalg.inv(frame)
    local = N.stack([N.dot(inv[:3], N.vstack((coords[i], N.ones(coords.shape[2])))) for i in range(2)])
    x, y, z = local[:, 0], local[:, 1], local[:, 2]
    if name == 'ParabolicDishGM':                           # paraboloid.py:89-119
        return _hitting_select((z <= gm._h) & (z >= 0) & (prm > 1e-6))
    if name == 'HexagonalParabolicDishGM':                  # paraboloid.py:193-223
        outside = N.abs(x) > N.sqrt(3) * gm._R / 2.
        outside |= N.abs(y) > gm._R - N.tan(N.pi / 6.) * N.abs(x)
        return _refine_select(_default_select(prm), (~outside) & (prm > 0))
    if name == 'RectangularParabolicDishGM':                # paraboloid.py:257-294
        outside = (N.abs(x) > gm._w) | (N.abs(y) > gm._h)
        return _refine_select(_default_select(prm), (~outside) & (prm > 1e-6))
    if name == 'ParabolicTroughGM':                         # paraboloid.py:411-443
        return _hitting_select((N.abs(y) <= gm._l / 2.) & (z <= gm._h) & (z >= 0.) & (prm > 1e-6))
    if name == 'HemisphereGM':                              # sphere_surface.py:122-139
        return _refine_select(_default_select(prm), (z <= 0) & (prm > 1e-6), True)
    if name == 'SphericalRectFacet':                        # sphere_surface.py:215-228
        good = (z <= 0) & (prm > 1e-6) & (N.abs(x) <= gm.lx / 2.) & (N.abs(y) <= gm.ly / 2.)
        return _refine_select(_default_select(prm), good, True)
    if name == 'FiniteCone':                                # cone.py:94-120
        return _hitting_select((z >= 0) & (z <= gm.h) & (prm > 1e-9))
    if name == 'ConicalFrustum':                            # cone.py:285-320
        return _hitting_select((gm.zmin <= z) & (z <= gm.zmax) & (prm > 1e-6))
    if name == 'RectCutConicalFrustum':                     # cone.py:365-395
        return _hitting_select((gm.zmin <= z) & (z <= gm.zmax) & (N.abs(x) <= gm.half_dims[0])
                               & (N.abs(y) <= gm.half_dims[1]) & (prm > 1e-6))
    if name == 'FiniteCylinder':                            # cylinder.py:71-110
        angs = N.arctan2(y, x)
        angs[angs < 0] = 2 * N.pi + angs[angs < 0]
        inside = (abs(z) <= gm._half_h) & N.logical_and(angs >= gm._ang_range[0], angs <= gm._ang_range[1])
        return _hitting_select(inside & (prm > 1e-6))
    if name == 'RectCutCylinder':                           # cylinder.py:171-199
        inside = (-gm._half_h <= z) & (z <= gm._half_h) & (N.abs(x) <= gm.half_dims[0]) & (N.abs(y) <= gm.half_dims[1])
        return _hitting_select(inside & (prm > 1e-6))
    if name == 'EllipsoidGM':                               # ellipsoid.py:77-116
        good = prm > 1e-7
        for coord, lim in ((x, gm.xlim), (y, gm.ylim), (z, gm.zlim)):
            if gm.truncated and lim is not None:
                good = good & (coord >= lim[0]) & (coord <= lim[1])
        return _hitting_select(good)
    if name == 'RectFlatQuadricSurfaceGM':                  # quadratic_surface.py:75-105
        outside = (N.abs(x) > gm._w) | (N.abs(y) > gm._h)
        return _refine_select(_default_select(prm), (~outside) & (prm > 1e-6))
    raise NotImplementedError(name)


def _quadric_normals(name, gm, frame, hits, directs):
    """hits (n,3) global, directs (n,3).  paraboloid.py:44-69, :358-384,
    sphere_surface.py:35-50, cone.py:28-57, cylinder.py:20-30, quadratic_surface.py:22-44"""
    if name in _SPH:
        c = frame[:3, 3]
        sides = N.sum((c - hits) * directs, axis=1)
        normal = (hits - c).T
        normal[:, sides < 0.] *= -1
        return normal / N.sqrt(N.sum(normal ** 2, axis=0))
    hit = _to_local(frame, hits.T)
    dir_loc = N.dot(frame[:3, :3].T, directs.T)
    n = hits.shape[0]
    if name in _CYL:
        local_norm = N.vstack((hit[:2], N.zeros(n)))
        local_norm /= gm._R
        local_norm[:, N.sum(local_norm[:2] * dir_loc[:2], axis=0) > 0.] *= -1.
        return N.dot(frame[:3, :3], local_norm)
    eps, apex = 0., None
    if name in _PARAB:
        ln = N.vstack((2 * hit[0] * gm.a, 2 * hit[1] * gm.b, -1 * N.ones(n)))
    elif name in _PCYL:
        ln = N.vstack((2 * hit[0] * gm.a, N.zeros(n), -1 * N.ones(n)))
    elif name in _CONE:
        ln = N.vstack((2. * hit[0], 2. * hit[1], -2. * (hit[2] - gm.a) * gm.c ** 2.))
        eps, apex = 1e-9, (hit[2] == gm.a)
    elif name in ('EllipsoidGM', 'Ellipsoid'):              # ellipsoid.py:36-61
        ln = N.vstack((2 * hit[0] * gm.a, 2 * hit[1] * gm.b, 2 * hit[2] * gm.c))
    else:
        ln = N.vstack((2. * hit[0] * gm.a + gm.c * hit[1] + gm.d, 2. * hit[1] * gm.b + gm.c * hit[0] + gm.e,
                       -1 * N.ones(n)))
    with N.errstate(invalid='ignore', divide='ignore'):
        lu = ln / N.sqrt(N.sum(ln ** 2, axis=0))
    lu[:, N.sum(dir_loc * lu, axis=0) > eps] *= -1
    if apex is not None:
        lu[:, apex] = N.vstack((0., 0., -1.))
    return N.dot(frame[:3, :3], lu)


def _find_quadric(name, gm, frame, v, d):
    """QuadricGM.find_intersections, quadric.py:28-113"""
    n = v.shape[1]
    A, B, C = _abc(name, gm, frame, v, d)
    delta = B ** 2. - 4. * A * C
    any_inters = delta >= 1e-6
    params = N.empty(n)
    params.fill(N.inf)
    vertices = N.empty((3, n))
    if any_inters.sum() != 0:
        A, B, C = A[any_inters], B[any_inters], C[any_inters]
        delta = N.sqrt(B ** 2. - 4. * A * C)
        hits = N.empty((2, len(A)))
        hits.fill(N.nan)
        is_linear = A == 0
        is_Bnull = B == 0
        m = is_linear & ~is_Bnull
        hits[:, m] = N.tile(-C[m] / B[m], (2, 1))
        m = ~is_linear & is_Bnull
        hits[0, m] = -N.sqrt(-C[m] / A[m])
        hits[1, m] = N.sqrt(-C[m] / A[m])
        with N.errstate(invalid='ignore', divide='ignore'):
            q = -0.5 * (B + N.sign(B) * delta)
            m = ~is_linear & ~is_Bnull
            hits[0, m] = q[m] / A[m]
            hits[1, m] = C[m] / q[m]
        inters_coords = v[:, any_inters] + d[:, any_inters] * hits.reshape(2, 1, -1)
        with N.errstate(invalid='ignore'):
            select = _select(name, gm, frame, inters_coords, hits)
        not_missed = ~N.isnan(select)
        any_inters[any_inters] = not_missed
        select = N.array(select[not_missed], dtype=N.int_)
        params[any_inters] = N.choose(select, hits[:, not_missed])
        vertices[:, any_inters] = N.choose(select, inters_coords[..., not_missed])

    def points(idx):
        return vertices[:, idx]

    def normals(idx):
        return _quadric_normals(name, gm, frame, vertices[:, idx].T, d[:, idx].T)
    return params, points, normals
