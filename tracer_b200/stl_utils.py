"""Triangle meshes as tracer objects: the reference's ray_trace_utils/stl_utils.py
(stl_to_tracer_geom :178-211, make_stl_tracer_object :213-227) re-stated for meshes of ~1e6 faces.

The reference builds one Surface + geometry manager + optics manager + BoundaryBox per
triangle in a Python loop.  (As shipped the module does not import: it asks
vector_manipulations for ``get_plane_normal``, which only exists as ``get_plane_normals``,
:9 -- so there is no live reference output to pin against; the frame convention below is
the one its code states: location = vertex A, rotation = rotz(phi_n) . roty(theta_n) of the
face normal n = (B-A) x (C-B), TriangularFace vertices = the images of B-A and C-A in that
frame.)  Here the per-triangle frames and vertices are computed for the whole mesh at once
and held as arrays by ONE ``TriangleMeshObject``; the scene compiler turns them into surface
records with array operations and the engine keeps the mesh's accountant data as one chunk
per bounce.  A ``Surface`` (with its own optics manager, filled with that surface's share of
the accountant data) is only created when somebody indexes ``mesh.get_surfaces()[i]``.
"""
import numpy as np

from .scenegraph import AssembledObject, Surface
from .geometry import TriangularFace


def triangle_frames(triangles):
    """(n,3,3) triangles [face, vertex, xyz] -> (local frames (n,4,4), verts (n,3,2)).
    Vectorised stl_to_tracer_geom(option='triangle') (stl_utils.py:178-211)."""
    tri = np.asarray(triangles, dtype=float)
    if tri.ndim != 3 or tri.shape[1:] != (3, 3):
        raise ValueError("triangles must be an (n, 3, 3) array: face, vertex, xyz")
    A, B, C = tri[:, 0], tri[:, 1], tri[:, 2]
    normal = np.cross(B - A, C - B)
    length = np.sqrt(np.sum(normal ** 2, axis=1))
    if not np.all(length > 0.):
        raise ValueError("degenerate (zero-area) triangle %d" % int(np.argmin(length > 0.)))
    normal /= length[:, None]
    phi = np.arctan2(normal[:, 1], normal[:, 0])
    th = np.arccos(np.clip(normal[:, 2], -1., 1.))
    cp, sp, ct, st = np.cos(phi), np.sin(phi), np.cos(th), np.sin(th)
    zero = np.zeros_like(cp)
    # rotz(phi) . roty(theta)
    rot = np.stack([np.stack([cp * ct, -sp, cp * st], axis=1),
                    np.stack([sp * ct, cp, sp * st], axis=1),
                    np.stack([-st, zero, ct], axis=1)], axis=1)
    back = np.transpose(rot, (0, 2, 1))                     # roty(-theta) . rotz(-phi)
    Bp = np.einsum('nij,nj->ni', back, B - A)
    Cp = np.einsum('nij,nj->ni', back, C - A)
    verts = np.zeros((len(tri), 3, 2))
    verts[:, :2, 0] = Bp[:, :2]
    verts[:, :2, 1] = Cp[:, :2]
    frames = np.zeros((len(tri), 4, 4))
    frames[:, :3, :3] = rot
    frames[:, :3, 3] = A
    frames[:, 3, 3] = 1.
    return frames, verts


def stl_to_tracer_geom(triangles, option='triangle'):
    """-> (geoms, locs, rots), one entry per triangle (stl_utils.py:178-211)."""
    frames, verts = triangle_frames(triangles)
    if option == 'triangle':
        geoms = [TriangularFace(v) for v in verts]
    elif option == 'polygon':
        from .geometry import FlatSimplePolygonGM
        geoms = [FlatSimplePolygonGM(np.array([[0., v[0, 0], v[0, 1]], [0., v[1, 0], v[1, 1]]])) for v in verts]
    else:
        raise ValueError("option must be 'triangle' or 'polygon'")
    return geoms, [f[:3, 3] for f in frames], [f[:3, :3] for f in frames]


class _MeshSurfaces(object):
    """Read-only sequence of the mesh's surfaces; a Surface is made when indexed."""

    def __init__(self, mesh):
        self._mesh = mesh

    def __len__(self):
        return self._mesh.num_triangles()

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._mesh.make_surface(k) for k in range(*i.indices(len(self)))]
        n = len(self)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError("surface index out of range")
        return self._mesh.make_surface(int(i))

    def __iter__(self):
        for k in range(len(self)):
            yield self._mesh.make_surface(k)


class TriangleMeshObject(AssembledObject):
    """All the triangles of a mesh as one assembled object (what make_stl_tracer_object
    returns).  ``optics`` is an optics-manager class or factory and ``optics_args`` its
    keyword arguments, as in stl_utils.py:213; every face gets that optics."""

    is_triangle_mesh = True

    def __init__(self, triangles, optics, optics_args=None, location=None, rotation=None, transform=None):
        self._local_frames, self._verts = triangle_frames(triangles)
        self._optics_factory = optics
        self._optics_args = dict(optics_args or {})
        self._template_opt = optics(**self._optics_args)
        self._global_frames = self._local_frames
        self._chunks = []            # accountant data: one entry per bounce that hit the mesh
        AssembledObject.__init__(self, surfs=_MeshSurfaces(self), bounds=[], location=location, rotation=rotation,
                                 transform=transform)

    def num_triangles(self):
        return len(self._local_frames)

    def add_surface(self, surface):
        raise TypeError("a TriangleMeshObject holds only its triangles; put other surfaces in another object")

    def transform_children(self, assembly_transform=np.eye(4)):
        mine = np.dot(assembly_transform, self.get_transform())
        self._global_frames = np.matmul(mine, self._local_frames)     # a NEW array: the engine's change detector
        for child in self.boundaries:
            child.transform_frame(mine)

    # -- what the scene compiler reads
    def mesh_frames(self):
        return self._global_frames

    def mesh_vertices(self):
        return self._verts

    def mesh_optics(self):
        return self._template_opt

    # -- accountant data (filled by TracerEngine._account, one chunk per bounce)
    def _add_chunk(self, seg, src):
        self._chunks.append((seg, src))

    def reset_optics(self):
        self._chunks = []

    def make_surface(self, i):
        """Surface i with an optics manager of its own holding that face's accountant rows."""
        from .scene import _optics_record
        from .tracer_engine import _LazyRows
        opt = self._optics_factory(**self._optics_args)
        lf = self._local_frames[i]
        surf = Surface(TriangularFace(self._verts[i].copy()), opt, location=lf[:3, 3].copy(), rotation=lf[:3, :3].copy())
        surf._temp_frame = self._global_frames[i]
        accountants = _optics_record(opt)['accountants']
        for seg, src in self._chunks:
            a, b = int(seg[i] - seg[0]), int(seg[i + 1] - seg[0])
            if b > a:
                for acc, field, name in accountants:
                    getattr(acc, field).append(_LazyRows(src[name][..., a:b]))
        return surf

    def get_all(self, name):
        """Accountant column ``name`` ('absorbed', 'hits', 'e_in', ...: the sources of
        scene._ACC_FIELDS) of every face, bounce after bounce, faces in index order inside a
        bounce -- what concatenating the per-surface accountant lists would give bounce-major."""
        cols = [src[name].cpu().numpy() for _, src in self._chunks if name in src]
        if not cols:
            return np.zeros((3, 0)) if name in ('hits', 'normals', 'dirs_in') else np.zeros(0)
        return np.concatenate(cols, axis=-1)

    def get_all_hits(self):
        """(absorbed energies (H,), hit points (3,H)) over the whole mesh."""
        return self.get_all('absorbed'), self.get_all('hits')

    def face_of_hits(self):
        """Face index of every column of get_all()."""
        out = [np.repeat(np.arange(len(seg) - 1), np.diff(seg)) for seg, _ in self._chunks]
        return np.concatenate(out) if out else np.zeros(0, dtype=np.int64)

    def absorbed_per_face(self):
        tot = np.zeros(self.num_triangles())
        if self._chunks:
            np.add.at(tot, self.face_of_hits(), self.get_all('absorbed'))
        return tot


def make_stl_tracer_object(triangles, optics, optics_args=None, option='triangle'):
    """Triangle vertices + an optics class and its arguments -> one assembled object
    (stl_utils.py:213-227).  Only ``option='triangle'`` has a mesh object; the reference's
    default 'polygon' makes FlatSimplePolygonGM faces, identical in results for triangles
    (use stl_to_tracer_geom to build those one by one)."""
    if option != 'triangle':
        raise NotImplementedError("make_stl_tracer_object: only option='triangle' is built as a mesh object")
    return TriangleMeshObject(triangles, optics, optics_args)


def load_stl(stl_file):
    """Binary or ASCII STL -> (n,3,3) triangles (stl_utils.py:156-165, without numpy-stl)."""
    with open(stl_file, 'rb') as f:
        data = f.read()
    if len(data) >= 84:
        n = int(np.frombuffer(data[80:84], dtype='<u4')[0])
        if len(data) == 84 + 50 * n:
            rec = np.frombuffer(data[84:], dtype=np.dtype([('n', '<f4', 3), ('v', '<f4', (3, 3)), ('a', '<u2')]))
            return rec['v'].astype(float)
    tris = [float(x) for line in data.decode('ascii', 'replace').splitlines()
            if line.strip().startswith('vertex') for x in line.split()[1:4]]
    return np.array(tris).reshape(-1, 3, 3)


def load_stl_into_tracer(stl_file, optics, optics_args=None, option='triangle'):
    return make_stl_tracer_object(load_stl(stl_file), optics, optics_args, option)
