"""Ray sources with the reference's function names (tracer/sources.py).

Two families:

* host generators (``disk_bundle``, ``rect_bundle``, ``buie_sunshape``,
  ``rect_buie_sunshape``, ``single_ray_source`` ...) draw from ``numpy.random``
  in the reference's order (sources.py:88-117, :175-264, :330-515), so with
  the same ``numpy.random.seed`` they return the same bundle as the reference.
  They are scene-setup code, run once per trace.
* device generators (``disk_bundle_device``, ``buie_sunshape_device``) fill a
  bundle directly in HBM with the Philox kernels of csrc/sources.cu; this is
  the path for 1e7-1e9 ray sources (north star: "the Buie and pillbox sources").
"""
import ctypes as C

import numpy as np
from numpy import random

from .ray_bundle import RayBundle, _DeviceColumns
from .spatial_geometry import rotation_to_z

THETA_DNI = 4.65e-3   # rad, edge of the solar disk
THETA_TOT = 43.6e-3   # rad, edge of the aureole
BUIE_SEGMENTS = 210


def single_ray_source(position, direction, flux=None):
    d = np.array(direction, dtype=float).reshape(3, 1)
    d /= np.sqrt(np.sum(d ** 2, axis=0))
    b = RayBundle(vertices=position, directions=d)
    b.set_energy(flux * np.ones(1))
    return b


def Lambertian_directions(num_rays, ang_range, normals=None):
    """Directions within ``ang_range`` of +z with Lambertian weighting (sources.py:88-100)."""
    if normals is not None:
        raise NotImplementedError("Lambertian_directions(normals=...) is host-side reference code")
    phi = random.uniform(low=0., high=2. * np.pi, size=num_rays)
    if ang_range == 0.:
        dirs = np.zeros((3, num_rays))
        dirs[2] = 1.
        return dirs
    xi = random.uniform(size=num_rays)
    s = np.sin(ang_range) * np.sqrt(xi)
    return np.vstack((np.cos(phi) * s, np.sin(phi) * s, np.sqrt(1. - s ** 2.)))


def pillbox_sunshape_directions(num_rays, ang_range):
    return Lambertian_directions(num_rays, ang_range)


def _apply(rot, local):
    return np.sum(rot[..., None] * local[None, ...], axis=1)


def disk_bundle(num_rays, center, direction, radius, ang_range, flux=None, radius_in=0.,
                angular_span=[0., 2. * np.pi], x_cut=None, procs=1, rays_direction=None):
    """Pillbox source on a disk / annulus sector (sources.py:175-239)."""
    if x_cut is not None:
        raise NotImplementedError("disk_bundle(x_cut=...) is not provided")
    radius, radius_in = float(radius), float(radius_in)
    a = pillbox_sunshape_directions(num_rays, ang_range)
    if rays_direction is None:
        rays_direction = direction
    rot = rotation_to_z(rays_direction)
    directions = _apply(rot, a)
    xi1 = random.uniform(size=num_rays)
    thetas = random.uniform(low=angular_span[0], high=angular_span[1], size=num_rays)
    rs = np.sqrt(radius_in ** 2. + xi1 * (radius ** 2. - radius_in ** 2.))
    local = np.vstack((rs * np.cos(thetas), rs * np.sin(thetas), np.zeros(num_rays)))
    b = RayBundle(vertices=np.dot(rot, local) + center, directions=directions)
    if flux is not None:
        ang = 2. * np.sin(np.sqrt(np.sum((rays_direction - direction) ** 2)) / 2.)
        b.set_energy(np.pi * (radius ** 2. - radius_in ** 2.) / num_rays * flux * np.ones(num_rays) * np.cos(ang))
    else:
        b.set_energy(np.ones(num_rays) / float(num_rays) / procs)
    return b


def rect_bundle(num_rays, center, direction, x, y, ang_range, flux=None, procs=1):
    """Pillbox source on a rectangle (sources.py:241-264)."""
    a = pillbox_sunshape_directions(num_rays, ang_range)
    rot = rotation_to_z(direction)
    directions = _apply(rot, a)
    xs = random.uniform(low=-x / 2., high=x / 2., size=num_rays)
    ys = random.uniform(low=-y / 2., high=y / 2., size=num_rays)
    if (direction == np.array([0, 0, -1])).all():
        xs, ys = ys, xs
    local = np.vstack((ys, xs, np.zeros(num_rays)))
    b = RayBundle(vertices=np.dot(rot, local) + center, directions=directions)
    if flux is not None:
        b.set_energy(x * y / num_rays * flux * np.ones(num_rays))
    else:
        b.set_energy(np.ones(num_rays) / float(num_rays) / procs)
    return b


def oblique_solar_rect_bundle(num_rays, center, source_direction, rays_direction, x, y, ang_range, flux=None,
                              procs=1, wavelength=None, ref_index=None):
    """Rectangular pillbox source whose rays leave along ``rays_direction`` while the source
    plane faces ``source_direction`` (sources.py:268-302)."""
    a = pillbox_sunshape_directions(num_rays, ang_range)
    directions = _apply(rotation_to_z(rays_direction), a)
    xs = np.random.uniform(low=-x / 2., high=x / 2., size=num_rays)
    ys = np.random.uniform(low=-y / 2., high=y / 2., size=num_rays)
    if (np.asarray(source_direction) == np.array([0, 0, -1])).all():
        xs, ys = ys, xs
    vertices_local = np.vstack((ys, xs, np.zeros(num_rays)))
    vertices_global = np.dot(rotation_to_z(source_direction), vertices_local)
    kw = {}
    if wavelength is not None:
        kw['wavelengths'] = np.ones(num_rays) * wavelength
    if ref_index is not None:
        kw['ref_index'] = np.ones(num_rays) * ref_index
    rayb = RayBundle(vertices=vertices_global + center, directions=directions, **kw)
    if flux is not None:
        cosangle = 2. * np.arcsin(0.5 * np.sqrt(np.sum((np.asarray(rays_direction) - np.asarray(source_direction)) ** 2)))
        rayb.set_energy(x * y / num_rays * flux * np.ones(num_rays) * np.cos(cosangle))
    else:
        rayb.set_energy(np.ones(num_rays) / float(num_rays) / procs)
    return rayb


def _rotz_each(phi, v, then=None):
    """rotz(phi[t])[:3, :3] applied to column t of ``v`` for every t at once (the reference loops over the
    rays, sources.py:676-678, 741-743): batched 3x3 products, bit-identical to the per-ray numpy.dot.
    ``then``: a fixed 3x3 matrix folded in first as (rotz . then) . v, the association of vf_cylinder_bundle."""
    c, s = np.cos(phi), np.sin(phi)
    R = np.zeros((len(phi), 3, 3))
    R[:, 0, 0], R[:, 0, 1], R[:, 1, 0], R[:, 1, 1], R[:, 2, 2] = c, -s, s, c, 1.
    if then is not None:
        R = np.matmul(R, then)
    return np.matmul(R, v.T[:, :, None])[:, :, 0].T


def vf_frustum_bundle(num_rays, r0, r1, depth, center, direction, flux=None, rays_in=True,
                      angular_span=[0., 2. * np.pi], angular_range=np.pi / 2.):
    """Lambertian emission from the lateral surface of a conical frustum, for view-factor
    ray traces (sources.py:644-714).  Draw order: directions, radius variate, azimuth."""
    from .spatial_geometry import roty, rotz
    r0, r1, depth, num_rays = float(r0), float(r1), float(depth), int(num_rays)
    dir_flat = pillbox_sunshape_directions(num_rays, angular_range)
    c = (r1 - r0) / depth
    R = np.random.uniform(size=num_rays)
    rs = np.sqrt((r1 ** 2. - r0 ** 2.) * R + r0 ** 2.)
    zs = (rs - r0) / ((r1 - r0) / depth)
    phi_s = np.random.uniform(low=angular_span[0], high=angular_span[1], size=num_rays)
    xs, ys = rs * np.cos(phi_s), rs * np.sin(phi_s)
    yrot = roty(-np.pi / 2. + np.arctan(c))[:3, :3]
    local_unit = _rotz_each(phi_s, np.dot(yrot, dir_flat))
    if rays_in == False:  # noqa: E712 (reference spelling)
        local_unit = -1. * local_unit
    perp_rot = rotation_to_z(direction)
    vertices_global = np.dot(perp_rot, np.vstack((xs, ys, zs)))
    directions = np.dot(perp_rot, local_unit)
    if flux is None:
        energy = np.ones(num_rays) / float(num_rays)
    else:
        area = (angular_span[1] - angular_span[0]) * (r1 + r0) / 2. * np.sqrt(abs(r1 - r0) ** 2. + depth ** 2.)
        energy = np.ones(num_rays) * flux * area / float(num_rays)
    return RayBundle(vertices=vertices_global + center, directions=directions, energy=energy)


def vf_cylinder_bundle(num_rays, rc, lc, center, direction, flux=None, rays_in=True,
                       angular_span=[0., 2. * np.pi], ang_range=np.pi / 2.):
    """Lambertian emission from the lateral surface of a cylinder (sources.py:716-767).
    Draw order: height, azimuth, directions."""
    from .spatial_geometry import roty, rotz
    rc, lc, num_rays = float(rc), float(lc), int(num_rays)
    zs = lc * np.random.uniform(size=num_rays) - lc / 2.
    phi_s = np.random.uniform(low=angular_span[0], high=angular_span[1], size=num_rays)
    xs, ys = rc * np.cos(phi_s), rc * np.sin(phi_s)
    dir_flat = pillbox_sunshape_directions(num_rays, ang_range)
    yrot = roty(-np.pi / 2.)[:3, :3]
    local_unit = _rotz_each(phi_s, dir_flat, then=yrot)
    if rays_in == False:  # noqa: E712
        local_unit = -local_unit
    perp_rot = rotation_to_z(direction)
    vertices_global = np.dot(perp_rot, np.vstack((xs, ys, zs)))
    directions = np.dot(perp_rot, local_unit)
    if flux is None:
        energy = np.ones(num_rays) / float(num_rays)
    else:
        area = rc * (angular_span[1] - angular_span[0]) * lc
        energy = np.ones(num_rays) * flux * area / float(num_rays)
    return RayBundle(vertices=vertices_global + center, directions=directions, energy=energy)


def buie_tables(CSR, pre_process_CSR=True):
    """Piecewise-linear solar-disk CDF + aureole constants of Buie et al. (2003)
    as used by Buie_directions_sampling (sources.py:330-362).  Returns a dict."""
    theta = np.linspace(0., THETA_DNI, BUIE_SEGMENTS + 1)
    phi = np.cos(0.326 * theta * 1e3) / np.cos(0.308 * theta * 1e3)
    w = phi * np.cos(theta) * np.sin(theta)
    integ = 0.5 * (w[:-1] + w[1:]) * (theta[1:] - theta[:-1])
    gamma = kappa = 0.
    if CSR == 0.:
        total = np.sum(integ)
    else:
        if pre_process_CSR:
            if CSR <= 0.1:
                CSR = -2.245e+03 * CSR ** 4. + 5.207e+02 * CSR ** 3. - 3.939e+01 * CSR ** 2. + 1.891e+00 * CSR + 8e-03
            else:
                CSR = 1.973 * CSR ** 4. - 2.481 * CSR ** 3. + 0.607 * CSR ** 2. + 1.151 * CSR - 0.020
        kappa = 0.9 * np.log(13.5 * CSR) * CSR ** (-0.3)
        gamma = 2.2 * np.log(0.52 * CSR) * CSR ** (0.43) - 0.1
        aureole = 1e-6 * np.exp(kappa) / (gamma + 2.) * ((THETA_TOT * 1000.) ** (gamma + 2.) - (THETA_DNI * 1000.) ** (gamma + 2.))
        total = np.sum(integ) + aureole
    cdf = np.add.accumulate(np.hstack(([0], integ / total)))
    return dict(theta=theta, w=w, cdf=cdf, sum_integ=np.sum(integ), gamma=gamma, kappa=kappa, CSR=CSR)


def Buie_directions_sampling(num_rays, CSR, pre_process_CSR=True):
    """Directions around +z distributed as the Buie sunshape (sources.py:330-384).
    One uniform for the polar angle (segment-wise closed-form inversion), one for
    the azimuth -- same draw order as the reference."""
    tb = buie_tables(CSR, pre_process_CSR)
    theta, w, cdf, tot = tb['theta'], tb['w'], tb['cdf'], tb['sum_integ']
    gamma, kappa = tb['gamma'], tb['kappa']
    r = np.random.uniform(size=num_rays)
    thetas = np.zeros(num_rays)
    seg = np.searchsorted(cdf, r, side='right') - 1
    disk = r < cdf[-1]
    i = seg[disk]
    A, B = w[i], w[i + 1]
    Cq = 2. * tot * (r[disk] - cdf[i]) * (theta[i + 1] - theta[i])
    thetas[disk] = -(-A * theta[i + 1] + B * theta[i] + np.sqrt(((theta[i] - theta[i + 1]) * A) ** 2. + Cq * (B - A))) / (A - B)
    if tb['CSR'] > 0.:
        ra = r[~disk]
        thetas[~disk] = ((ra - 1.) * ((gamma + 2.) / (10. ** (3. * gamma) * np.exp(kappa)) * tot - THETA_DNI ** (gamma + 2.))
                         + ra * THETA_TOT ** (gamma + 2.)) ** (1. / (gamma + 2.))
    xi = random.uniform(high=2. * np.pi, size=num_rays)
    st = np.sin(thetas)
    return np.vstack((np.cos(xi) * st, np.sin(xi) * st, np.cos(thetas)))


def _buie_bundle(local_xy, area, num_rays, center, direction, CSR, flux, pre_process_CSR, rays_direction):
    if rays_direction is None:
        rays_direction = direction
    ang = 2. * np.sin(np.sqrt(np.sum((rays_direction - direction) ** 2)) / 2.)
    energy = np.ones(num_rays) * flux * area / num_rays * np.cos(ang)
    a = Buie_directions_sampling(num_rays, CSR, pre_process_CSR)
    directions = _apply(rotation_to_z(rays_direction), a)
    local = np.vstack((local_xy[0], local_xy[1], np.zeros(num_rays)))
    vertices = np.dot(rotation_to_z(direction), local) + center
    return RayBundle(vertices=vertices, directions=directions, energy=energy)


def buie_sunshape(num_rays, center, direction, radius, CSR, flux=None, pre_process_CSR=True, rays_direction=None):
    """Buie-sunshape source on a disk (sources.py:412-464)."""
    xv = random.uniform(size=num_rays)
    phiv = random.uniform(high=2. * np.pi, size=num_rays)
    rs = radius * np.sqrt(xv)
    return _buie_bundle((rs * np.cos(phiv), rs * np.sin(phiv)), np.pi * radius ** 2., num_rays, center, direction,
                        CSR, flux, pre_process_CSR, rays_direction)


def rect_buie_sunshape(num_rays, center, direction, width, height, CSR, flux=None, pre_process_CSR=True,
                       rays_direction=None):
    """Buie-sunshape source on a rectangle (sources.py:466-515)."""
    xs = width * (random.uniform(size=num_rays) - 0.5)
    ys = height * (random.uniform(size=num_rays) - 0.5)
    return _buie_bundle((xs, ys), width * height, num_rays, center, direction, CSR, flux, pre_process_CSR,
                        rays_direction)


# ----------------------------------------------------------------- device
def _device_bundle_buffers(num_rays, device):
    import torch
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    V = torch.empty((3, num_rays), dtype=torch.float64, device=dev)
    D = torch.empty((3, num_rays), dtype=torch.float64, device=dev)
    E = torch.empty(num_rays, dtype=torch.float64, device=dev)
    return dev, V, D, E


MAX_CONE_HINT = 0.25     # rad: wider sources are not worth a shadow grid


def _wrap(V, D, E, first_ray=0, cone=None):
    """Device-backed bundle.  ``_ray_id_base`` is the global index of its first ray: the engine keys the
    optics draws of every descendant by (seed, global source-ray index, bounce), so the shards of one source
    traced on several GPUs draw exactly what the whole source draws."""
    b = RayBundle(vertices=_DeviceColumns([V]), directions=_DeviceColumns([D]), energy=_DeviceColumns([E]))
    b.__dict__['_ray_id_base'] = int(first_ray)
    if cone is not None and cone[1] <= MAX_CONE_HINT:
        # every direction lies within cone[1] radians of cone[0]: lets the engine trace the first bounce through the
        # scene's shadow grid (scene.py set_shadow_grid); the kernel re-checks each ray, so the hint only prunes
        b.__dict__['_dir_cone'] = (np.asarray(cone[0], float).reshape(3).copy(), float(cone[1]) * (1. + 1e-9) + 1e-12)
    return b


def _need_flux(flux, who):
    if flux is None:
        raise ValueError(who + ": flux (W/m2) is required")


def disk_bundle_device(num_rays, center, direction, radius, ang_range, flux=None, radius_in=0.,
                       angular_span=[0., 2. * np.pi], procs=1, rays_direction=None, seed=0, first_ray=0,
                       total_rays=None, device=None):
    """``disk_bundle`` generated in HBM by the Philox pillbox kernel.  ``seed`` and
    ``first_ray`` (global index of this shard's first ray) fix the bundle; a source of ``total_rays``
    sharded over GPUs passes the common total so that the per-ray energy is that of the whole source."""
    import torch
    from . import _lib as L
    lib = L.load()
    direction = np.asarray(direction, float)
    if rays_direction is None:
        rays_direction = direction
    rot = rotation_to_z(rays_direction)
    if flux is not None:
        ang = 2. * np.sin(np.sqrt(np.sum((rays_direction - direction) ** 2)) / 2.)
        e = np.pi * (float(radius) ** 2. - float(radius_in) ** 2.) / (total_rays or num_rays) * flux * np.cos(ang)
    else:
        e = 1. / float(total_rays or num_rays) / procs
    dev, V, D, E = _device_bundle_buffers(num_rays, device)
    _, c = L.host_doubles(center, 3)
    _, rd = L.host_doubles(rot, 9)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        L.check(lib.trc_source_pillbox_disk(num_rays, first_ray, seed, c, rd, rd, float(radius_in), float(radius),
                                            float(angular_span[0]), float(angular_span[1]), float(ang_range),
                                            float(e), L.dptr(V), L.dptr(D), L.dptr(E), num_rays, st))
    return _wrap(V, D, E, first_ray, cone=(np.asarray(rot)[:, 2], float(ang_range)))


def buie_sunshape_device(num_rays, center, direction, radius, CSR, flux=None, pre_process_CSR=True,
                         rays_direction=None, seed=0, first_ray=0, total_rays=None, device=None):
    """``buie_sunshape`` generated in HBM by the Philox Buie kernel.  When a source
    of ``total_rays`` is sharded over GPUs each shard passes its ``first_ray`` and
    the common ``total_rays`` so the per-ray energy is flux*area/total_rays."""
    import torch
    from . import _lib as L
    lib = L.load()
    _need_flux(flux, 'buie_sunshape_device')
    direction = np.asarray(direction, float)
    if rays_direction is None:
        rays_direction = direction
    tb = buie_tables(CSR, pre_process_CSR)
    t = L.BuieTables()
    t.n_seg, t.has_aureole = BUIE_SEGMENTS, int(tb['CSR'] > 0.)
    for k in range(BUIE_SEGMENTS + 1):
        t.cdf[k], t.theta[k], t.w[k] = tb['cdf'][k], tb['theta'][k], tb['w'][k]
    t.sum_integ_dni, t.gamma, t.kappa = tb['sum_integ'], tb['gamma'], tb['kappa']
    t.theta_dni, t.theta_tot = THETA_DNI, THETA_TOT
    ang = 2. * np.sin(np.sqrt(np.sum((rays_direction - direction) ** 2)) / 2.)
    e = flux * np.pi * radius ** 2. / (total_rays or num_rays) * np.cos(ang)
    dev, V, D, E = _device_bundle_buffers(num_rays, device)
    _, c = L.host_doubles(center, 3)
    _, rd = L.host_doubles(rotation_to_z(rays_direction), 9)
    _, rp = L.host_doubles(rotation_to_z(direction), 9)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        L.check(lib.trc_source_buie_disk(num_rays, first_ray, seed, C.byref(t), c, rd, rp, float(radius), float(e),
                                         L.dptr(V), L.dptr(D), L.dptr(E), num_rays, st))
    return _wrap(V, D, E, first_ray, cone=(rotation_to_z(rays_direction)[:, 2], THETA_TOT))


def _rect_local_half_extents(x, y, plane_normal):
    """Half extents along the local axes of the source plane: the reference lays the points
    out as (ys, xs, 0) and swaps the two when the plane faces straight down
    (sources.py:251-257, 278-284)."""
    if (np.asarray(plane_normal) == np.array([0, 0, -1])).all():
        return x / 2., y / 2.
    return y / 2., x / 2.


def rect_bundle_device(num_rays, center, direction, x, y, ang_range, flux=None, procs=1, seed=0, first_ray=0,
                       total_rays=None, device=None, rays_direction=None):
    """``rect_bundle`` (and, with ``rays_direction``, ``oblique_solar_rect_bundle``) generated in
    HBM by the Philox pillbox kernel."""
    import torch
    from . import _lib as L
    lib = L.load()
    direction = np.asarray(direction, float)
    if rays_direction is None:
        rays_direction = direction
    rays_direction = np.asarray(rays_direction, float)
    total = total_rays or num_rays
    if flux is not None:
        cosangle = 2. * np.arcsin(0.5 * np.sqrt(np.sum((rays_direction - direction) ** 2)))
        e = x * y / total * flux * np.cos(cosangle)
    else:
        e = 1. / float(total) / procs
    hx, hy = _rect_local_half_extents(x, y, direction)
    dev, V, D, E = _device_bundle_buffers(num_rays, device)
    _, c = L.host_doubles(center, 3)
    _, rd = L.host_doubles(rotation_to_z(rays_direction), 9)
    _, rp = L.host_doubles(rotation_to_z(direction), 9)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        L.check(lib.trc_source_pillbox_rect(num_rays, first_ray, seed, c, rd, rp, float(hx), float(hy),
                                            float(ang_range), float(e), L.dptr(V), L.dptr(D), L.dptr(E), num_rays, st))
    return _wrap(V, D, E, first_ray, cone=(rotation_to_z(rays_direction)[:, 2], float(ang_range)))


def oblique_solar_rect_bundle_device(num_rays, center, source_direction, rays_direction, x, y, ang_range, flux=None,
                                     procs=1, seed=0, first_ray=0, total_rays=None, device=None):
    return rect_bundle_device(num_rays, center, source_direction, x, y, ang_range, flux=flux, procs=procs, seed=seed,
                              first_ray=first_ray, total_rays=total_rays, device=device,
                              rays_direction=rays_direction)


def rect_buie_sunshape_device(num_rays, center, direction, width, height, CSR, flux=None, pre_process_CSR=True,
                              rays_direction=None, seed=0, first_ray=0, total_rays=None, device=None):
    """``rect_buie_sunshape`` (sources.py:466-515) generated in HBM by the Philox Buie kernel."""
    import torch
    from . import _lib as L
    lib = L.load()
    _need_flux(flux, 'rect_buie_sunshape_device')
    direction = np.asarray(direction, float)
    if rays_direction is None:
        rays_direction = direction
    tb = buie_tables(CSR, pre_process_CSR)
    t = L.BuieTables()
    t.n_seg, t.has_aureole = BUIE_SEGMENTS, int(tb['CSR'] > 0.)
    for k in range(BUIE_SEGMENTS + 1):
        t.cdf[k], t.theta[k], t.w[k] = tb['cdf'][k], tb['theta'][k], tb['w'][k]
    t.sum_integ_dni, t.gamma, t.kappa = tb['sum_integ'], tb['gamma'], tb['kappa']
    t.theta_dni, t.theta_tot = THETA_DNI, THETA_TOT
    ang = 2. * np.sin(np.sqrt(np.sum((rays_direction - direction) ** 2)) / 2.)
    e = flux * width * height / (total_rays or num_rays) * np.cos(ang)
    dev, V, D, E = _device_bundle_buffers(num_rays, device)
    _, c = L.host_doubles(center, 3)
    _, rd = L.host_doubles(rotation_to_z(rays_direction), 9)
    _, rp = L.host_doubles(rotation_to_z(direction), 9)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        L.check(lib.trc_source_buie_rect(num_rays, first_ray, seed, C.byref(t), c, rd, rp, float(width) / 2.,
                                         float(height) / 2., float(e), L.dptr(V), L.dptr(D), L.dptr(E), num_rays, st))
    return _wrap(V, D, E, first_ray, cone=(rotation_to_z(rays_direction)[:, 2], THETA_TOT))
