"""Heliostat field model (reference tracer/models/heliostat_field.py:19-251).

Tower at the origin, +x east, +y north.  Every heliostat is an assembly
  hstat(location = position) -> [primary RotationAxis] + facet assembly
  facet(location = axes_pos[0]) -> [mirror object, secondary RotationAxis]
so object order per heliostat is (mirror, secondary axis, primary axis) and the
mirror surfaces come out of ``get_surfaces()`` in heliostat order.
"""
import numpy as np

from ..scenegraph import Assembly, AssembledObject, BoundaryBox
from ..spatial_geometry import general_axis_rotation, rotx, roty
from .one_sided_mirror import rect_one_sided_mirror, rect_para_one_sided_mirror, flat_quad_one_sided_mirror


class RotationAxis(AssembledObject):
    """Surfaceless object that carries a tracking axis through the assembly transforms."""

    def __init__(self, axis=None):
        self.axis = axis
        AssembledObject.__init__(self)

    def get_rotation_axis(self):
        return np.dot(self.get_rotation()[:3, :3], self.axis)


def solar_vector(azimuth, zenith):
    """Unit vector towards the sun; azimuth from north towards east, zenith from +z (radians)."""
    az = np.pi / 2. - azimuth
    if az < 0.:
        az += 2 * np.pi
    return np.r_[np.sin(zenith) * np.cos(az), np.sin(zenith) * np.sin(az), np.cos(zenith)]


class HeliostatField(Assembly):
    def __init__(self, positions, width, height, absorptivity, sigma, bi_var=True, focal_lengths=None,
                 quad_params=None, MCRT_option='fast',
                 rotation_axes_pos=np.array([[0., 0., 0.], [0., 0., 0.]]),
                 rotation_axes_vec=np.array([[0., 0., 1.], [1., 0., 0.]])):
        n = positions.shape[0]
        self._pos = positions
        focal_lengths = [None] * n if focal_lengths is None else focal_lengths
        quad_params = [None] * n if quad_params is None else quad_params
        if not hasattr(absorptivity, '__len__'):
            absorptivity = np.ones(n) * absorptivity
        self.rotation_axes_pos = rotation_axes_pos
        axes_offset = rotation_axes_pos[1] - rotation_axes_pos[0]
        self._heliostats = []
        for p in range(n):
            f, q = focal_lengths[p], quad_params[p]
            assert not ((f is not None) and (q is not None))
            bounds = BoundaryBox(np.array([[-width / 2., width / 2.], [-height / 2., height / 2.], [-1e-6, 1e-6]]).T)
            if f is None and q is None:
                mirror = rect_one_sided_mirror(width, height, absorptivity[p], sigma, bi_var, MCRT_option, bounds=bounds)
            elif f is not None:
                mirror = rect_para_one_sided_mirror(width, height, f, absorptivity[p], sigma, bi_var, MCRT_option,
                                                    bounds=bounds)
            else:
                mirror = flat_quad_one_sided_mirror(width, height, q, absorptivity[p], sigma, bi_var, MCRT_option,
                                                    bounds=bounds)
            mirror.set_location(axes_offset)
            facet = Assembly(objects=[mirror, RotationAxis(axis=rotation_axes_vec[1])], location=rotation_axes_pos[0])
            hstat = Assembly(objects=[RotationAxis(axis=rotation_axes_vec[0])], subassemblies=[facet],
                             location=positions[p])
            self._heliostats.append(hstat)
        Assembly.__init__(self, subassemblies=self._heliostats)

    def get_heliostats(self):
        return self._heliostats

    def set_aim_height(self, h):
        self._th = h

    def track_sun(self, azimuth, zenith, aim_points=None, aim_vectors=None, tracking='azimuth_elevation',
                  tracking_error=None, tracking_limits_primary_axis=None, tracking_limits_secondary_axis=None):
        """Orient every heliostat so that it reflects the sun towards its aim
        point (or along its aim vector).  Like the reference (:104-112)
        ``aim_points`` / ``aim_vectors`` are modified in place."""
        sun = solar_vector(azimuth, zenith)
        if aim_points is None:
            if aim_vectors is None:
                raise ValueError('aim-points or aiming vectors have to be set')
            aim = aim_vectors
        else:
            aim_points -= self._pos + np.sum(self.rotation_axes_pos, axis=0)
            aim = aim_points
        aim /= np.sqrt(np.sum(aim ** 2, axis=1)[:, None])
        trac = sun + aim
        trac /= np.sqrt(np.sum(trac ** 2, axis=1)[:, None])
        lim1 = [-np.pi, np.pi] if tracking_limits_primary_axis is None else tracking_limits_primary_axis
        lim2 = [-np.pi, np.pi] if tracking_limits_secondary_axis is None else tracking_limits_secondary_axis
        err1 = err2 = 0.
        if tracking == 'azimuth_elevation':
            az_all = np.arctan2(trac[:, 1], trac[:, 0])
            ze_all = np.arccos(trac[:, 2])
            for h, hstat in enumerate(self._heliostats):
                if tracking_error is not None:
                    err1 = np.random.normal(scale=tracking_error)
                    err2 = np.random.normal(scale=tracking_error)
                ang_az, ang_ze = az_all[h] + err1, ze_all[h] + err2
                if ang_az < -np.pi:
                    ang_az += np.pi
                if ang_az > np.pi:
                    ang_az -= np.pi
                if not (lim1[0] <= ang_az <= lim1[1]):
                    print(ang_az, 'is outside of tracking limits')
                    continue
                if not (lim2[0] <= ang_ze <= lim2[1]):
                    print(ang_ze, 'is outside of tracking limits')
                    continue
                facet = hstat.get_assemblies()[0]
                primary = hstat.get_local_objects()[0]
                facet.set_rotation(general_axis_rotation(primary.get_rotation_axis(), np.pi / 2. + ang_az))
                mirror, secondary = facet.get_objects()
                mirror.set_rotation(general_axis_rotation(secondary.get_rotation_axis(), ang_ze))
        elif tracking == 'tilt_roll':
            tilt_all = np.arctan2(trac[:, 1], trac[:, 2])
            roll_all = np.arcsin(trac[:, 0])
            for h, hstat in enumerate(self._heliostats):
                if tracking_error is not None:
                    err1 = np.random.normal(scale=tracking_error)
                    err2 = np.random.normal(scale=tracking_error)
                tilt, roll = tilt_all[h] + err1, roll_all[h] + err2
                if not (lim1[0] <= tilt <= lim1[1]) or not (lim2[0] <= roll <= lim2[1]):
                    continue
                hstat.set_rotation(np.dot(rotx(-tilt)[:3, :3], roty(roll)[:3, :3]))
        # re-resolve every global frame from the root
        Assembly.__init__(self, subassemblies=self._heliostats)

    def get_tracking_vectors(self):
        return [np.dot(h.get_rotation(), np.vstack([0., 0., 1.])) for h in self.get_heliostats()]


def radial_stagger(start_ang, end_ang, az_space, rmin, rmax, r_space):
    """(n,2) x,y positions of a radial-stagger layout (:219-251)."""
    rs = np.r_[rmin:rmax:r_space]
    angs = np.r_[start_ang:end_ang:az_space / 2]
    x = np.r_[np.outer(rs[::2], np.cos(angs[::2])).flatten(), np.outer(rs[1::2], np.cos(angs[1::2])).flatten()]
    y = np.r_[np.outer(rs[::2], np.sin(angs[::2])).flatten(), np.outer(rs[1::2], np.sin(angs[1::2])).flatten()]
    return np.vstack((x, y)).T
