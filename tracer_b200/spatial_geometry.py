"""Rigid-transform helpers with the reference's conventions.

Mirrors the public functions of the reference's ``tracer/spatial_geometry.py``
(general_axis_rotation :8-22, rotation_to_z :24-48, generate_transform :50-61,
rotx/roty/rotz/translate :63-99) so scene scripts run unchanged.  Host-side
only: these build 4x4 frames once per scene; the device code has its own
per-ray versions (csrc/optics.cuh).
"""
import math

import numpy as np


def general_axis_rotation(axis, ang):
    """3x3 right-handed rotation by ``ang`` radians about the unit vector ``axis``.

    Like the reference (spatial_geometry.py:18) the sine and cosine are rounded
    to 14 decimals, which makes quarter turns exact.
    """
    ax = np.asarray(axis, dtype=float)
    s = np.round(math.sin(ang), decimals=14)
    c = np.round(math.cos(ang), decimals=14)
    cross = np.array([[0.0, -ax[2], ax[1]],
                      [ax[2], 0.0, -ax[0]],
                      [-ax[1], ax[0], 0.0]])
    return np.multiply.outer(ax, ax) * (1 - c) + np.eye(3) * c + cross * s


def rotation_to_z(vecs):
    """Frame(s) whose z axis is ``vecs``; x lies in the original xy plane.

    Returns a (3,3) array (or (n,3,3) for n vectors) whose columns are the new
    x, y, z axes (spatial_geometry.py:24-48).  For vecs parallel to z the
    x axis is kept.
    """
    v = np.atleast_2d(np.asarray(vecs, dtype=float))
    x_new = np.stack([v[:, 1], -v[:, 0], np.zeros(len(v))], axis=1)
    degenerate = np.all(x_new == 0.0, axis=1)
    x_new[degenerate] = (1.0, 0.0, 0.0)
    x_new /= np.sqrt((x_new ** 2.0).sum(axis=1))[:, None]
    y_new = np.cross(v, x_new)
    return np.squeeze(np.stack([x_new, y_new, v], axis=2))


def generate_transform(axis, angle, translation):
    """4x4 homogeneous transform: rotation about ``axis`` then ``translation`` (3x1 column)."""
    out = np.eye(4)
    out[:3, :3] = general_axis_rotation(axis, angle)
    out[:3, 3] = np.asarray(translation, dtype=float).reshape(3)
    return out


def _homog(r3):
    out = np.eye(4)
    out[:3, :3] = r3
    return out


def rotx(ang):
    s, c = np.sin(ang), np.cos(ang)
    return _homog([[1.0, 0, 0], [0, c, -s], [0, s, c]])


def roty(ang):
    s, c = np.sin(ang), np.cos(ang)
    return _homog([[c, 0, s], [0, 1.0, 0], [-s, 0, c]])


def rotz(ang):
    s, c = np.sin(ang), np.cos(ang)
    return _homog([[c, -s, 0], [s, c, 0], [0, 0, 1.0]])


def translate(x=0, y=0, z=0):
    out = np.eye(4)
    out[:3, 3] = (x, y, z)
    return out
