"""Scene graph with the reference's public API: frames, assemblies, objects,
surfaces and boundary boxes.

These are *descriptions* only.  The reference's classes (tracer/has_frame.py,
assembly.py, object.py, surface.py, boundary_shape.py) also carry the per-bounce
protocol state (register_incoming/select_rays/get_outgoing/done,
surface.py:55-112); here that work happens on the GPU, so a Surface just pairs a
geometry description with an optics description and knows its global frame.
The scene compiler (tracer_b200/scene.py) reads ``Assembly.get_surfaces()``,
``Surface._temp_frame``, ``Surface._geom`` and ``Surface._opt`` -- the same
attribute names the reference uses, so reference-built scenes compile too.
"""
import numpy as np


class HasFrame(object):
    """Location + rotation relative to the parent, and the resolved global frame.

    API of has_frame.py:3-83: get/set_location, get/set_rotation,
    get/set_transform, transform_frame; ``_temp_frame`` is the 4x4 global frame.
    """

    def __init__(self, location=None, rotation=None):
        self._transform = np.eye(4)
        self.set_location(np.zeros(3) if location is None else location)
        self.set_rotation(np.eye(3) if rotation is None else rotation)
        self._temp_frame = self._transform

    def get_location(self):
        return self._loc

    def get_rotation(self):
        return self._rot

    def get_transform(self):
        return self._transform

    def set_location(self, location):
        location = np.asarray(location)
        if location.shape != (3,):
            raise ValueError("location must be a 1D 3-component array")
        self._loc = location
        self._transform[:3, 3] = location

    def set_rotation(self, rotation):
        rotation = np.asarray(rotation)
        if rotation.shape != (3, 3):
            raise ValueError("rotation must be a 3x3 array")
        self._rot = rotation
        self._transform[:3, :3] = rotation

    def set_transform(self, transform):
        self._transform = transform
        self._loc = transform[:3, 3]
        self._rot = transform[:3, :3]

    def transform_frame(self, transform):
        """Resolve the global frame given the parent's global transform."""
        self._temp_frame = np.dot(transform, self._transform)

    def _to_local(self, points):
        # the reference rounds the inverse to 9 decimals (surface.py:125, assembly.py:44)
        proj = np.round(np.linalg.inv(self._temp_frame), decimals=9)
        return np.dot(proj, np.vstack((points, np.ones(points.shape[1]))))


class Surface(HasFrame):
    """A geometry manager + an optics manager placed in an object (surface.py:12-53)."""

    def __init__(self, geometry, optics, location=None, rotation=None, fixed_color=False):
        HasFrame.__init__(self, location, rotation)
        self._geom = geometry
        self._opt = optics
        self._fixed_color = fixed_color[:3] if fixed_color else fixed_color
        self._transparency = fixed_color[-1] if (fixed_color and len(fixed_color) == 4) else 0

    def get_optics_manager(self):
        return self._opt

    def get_geometry_manager(self):
        return self._geom

    def global_to_local(self, points):
        """(3,n) global points -> (4,n) homogeneous local coordinates (surface.py:114-126)."""
        return self._to_local(points)

    def get_fluxmap(self, eners, local_coords, resolution):
        return self._geom.get_fluxmap(eners, local_coords, resolution)

    # -- the per-surface trace protocol of the reference (surface.py:55-112), served by the same
    # -- CUDA kernels: each call runs a one-surface, one-bounce trace on the GPU (no host arithmetic).
    def _single_surface_trace(self, bundle, optics, rng='philox'):
        from .tracer_engine import TracerEngine
        twin = Surface(self._geom, optics)
        twin._temp_frame = self._temp_frame
        holder = _ProtocolScene(twin)
        eng = TracerEngine(holder, device_accel='brute', rng=rng)
        eng.ray_tracer(bundle, reps=1, min_energy=-np.inf, tree=True)
        return eng.tree[1] if eng.tree.num_bunds() > 1 else None

    def register_incoming(self, ray_bundle):
        """Records the bundle and returns the ray parameter of its intersection with this surface,
        +inf for rays that miss (surface.py:55-70)."""
        from .optics_callables import Transparent
        from .ray_bundle import RayBundle
        self._current_bundle = ray_bundle
        v = np.asarray(ray_bundle.get_vertices(), float)
        d = np.asarray(ray_bundle.get_directions(), float)
        params = np.full(v.shape[1], np.inf)
        child = self._single_surface_trace(
            RayBundle(vertices=v, directions=d, energy=np.ones(v.shape[1])), Transparent())
        if child is not None and child.get_num_rays():
            hit = np.asarray(child.get_parents(), int)
            params[hit] = np.sum((child.get_vertices() - v[:, hit]) * d[:, hit], axis=0) / np.sum(d[:, hit] ** 2, axis=0)
        return params

    def select_rays(self, idxs):
        self._selected = np.asarray(idxs, int)

    def update_current_bundle(self, bundle):
        self._current_bundle = bundle

    def get_outgoing(self):
        """The optics of this surface applied to the selected rays of the registered bundle
        (surface.py:84-94); draws come from numpy.random in the reference's order."""
        from .ray_bundle import RayBundle
        sel = self._selected
        picked = self._current_bundle.inherit(sel)
        child = self._single_surface_trace(picked, self._opt, rng='numpy')
        if child is None:
            return RayBundle.empty_bund()
        local = np.asarray(child.get_parents(), int)
        if len(local) != len(sel) and len(np.unique(local)) == len(local) and len(local) < len(sel):
            raise ValueError("get_outgoing: a selected ray does not intersect this surface")
        return child.inherit(parents=sel[local])

    def done(self):
        if hasattr(self, '_current_bundle'):
            del self._current_bundle


class _ProtocolScene(object):
    """The smallest thing the scene compiler accepts: one surface already carrying its global frame."""

    def __init__(self, surface):
        self._surface = surface

    def get_surfaces(self):
        return [self._surface]

    def get_objects(self):
        return [self]

    def get_boundaries(self):
        return []


class BoundaryShape(HasFrame):
    def in_bounds(self, points):
        raise TypeError("Virtual function in_bounds() called. Implement this in a derived class")


class BoundaryBox(BoundaryShape):
    """Axis-aligned box given in the owner's local frame; its global AABB
    (``_minpoint``/``_maxpoint``) is refreshed on every transform
    (boundary_shape.py:58-87).  Feeds the acceleration structure."""

    def __init__(self, aabb, location=None, rotation=None):
        BoundaryShape.__init__(self, location, rotation)
        self._aabb = np.array(aabb, dtype=float)
        self._AABB = np.array(aabb, dtype=float)

    def update_AABB(self):
        lo, hi = self._aabb
        corners = np.ones((4, 8))
        for k in range(8):
            corners[0, k] = hi[0] if k & 1 else lo[0]
            corners[1, k] = hi[1] if k & 2 else lo[1]
            corners[2, k] = hi[2] if k & 4 else lo[2]
        glob = np.dot(self._temp_frame, corners)[:3]
        self._minpoint, self._maxpoint = glob.min(axis=1), glob.max(axis=1)
        self._AABB = np.array([self._minpoint, self._maxpoint])

    def in_bounds(self, bund_vertices):
        lo, hi = self._minpoint, self._maxpoint
        if bund_vertices.ndim == 2:
            lo, hi = lo[:, None], hi[:, None]
        return np.logical_and(bund_vertices > lo, bund_vertices < hi).all(axis=0)

    def transform_frame(self, transform):
        BoundaryShape.transform_frame(self, transform)
        self.update_AABB()


class BoundarySphere(BoundaryShape):
    """Sphere used by CutSphereGM to keep only part of a spherical surface
    (boundary_shape.py:89-111).  As in the reference the centre is the location
    given at construction and does not follow later frame transforms."""

    def __init__(self, location=None, radius=1.):
        BoundaryShape.__init__(self, location, None)
        self._temp_loc = self._loc
        self._radius = radius

    def in_bounds(self, bund_vertices):
        return self._radius ** 2 >= ((bund_vertices - self._temp_loc[:3]) ** 2).sum(axis=1)


class BoundaryCylinder(BoundaryShape):
    """Infinite cylinder along the local z axis (boundary_shape.py:130-149).  Kept as the reference writes it:
    ``in_bounds`` projects with the frame itself, and stacks ``vertices.T`` on ``vertices.shape[1]`` ones, which
    only works for a (3, 3) argument -- CutSphereGM passes (2, 3) and the reference raises ValueError there, so
    the GPU engine refuses a CutSphereGM bounded by a cylinder instead of inventing a behaviour."""

    def __init__(self, diameter=1., location=None, rotation=None):
        self._R = diameter / 2.
        BoundaryShape.__init__(self, location, None)

    def in_bounds(self, vertices):
        local_xy = np.dot(self._temp_frame[:2], np.vstack((vertices.T, np.ones(vertices.shape[1]))))
        return np.sum(local_xy ** 2, axis=0) <= self._R ** 2


class BoundaryPlane(BoundaryShape):
    """Half space on the positive local z side of the frame's xy plane (boundary_shape.py:151-165)."""

    def in_bounds(self, vertices):
        local_z = np.dot(np.linalg.inv(self._temp_frame)[2], np.vstack((vertices.T, np.ones(vertices.shape[0]))))
        return local_z >= 0


# Bumped by every structural change made through the API of any Assembly / AssembledObject;
# get_surfaces() results are cached against it (a heliostat field is ~650 objects and the
# engine asks for the surface list at the start of every trace).
_STRUCTURE_VERSION = [0]


class SurfaceSeq(object):
    """Read-only concatenation of per-object surface sequences, some of which make their
    Surface objects on demand; indexable and iterable like the list get_surfaces() returns."""

    def __init__(self, parts):
        self.parts = parts
        self._starts = np.concatenate(([0], np.cumsum([len(p) for p in parts]))).astype(np.int64)

    def __len__(self):
        return int(self._starts[-1])

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(len(self)))]
        n = len(self)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError("surface index out of range")
        k = int(np.searchsorted(self._starts, i, side='right')) - 1
        return self.parts[k][int(i - self._starts[k])]

    def __iter__(self):
        for p in self.parts:
            for s in p:
                yield s


class Assembly(HasFrame):
    """Tree of objects and sub-assemblies (assembly.py:7-151)."""

    def __init__(self, objects=None, subassemblies=None, location=None, rotation=None):
        self._objects = [] if objects is None else objects
        self._assemblies = [] if subassemblies is None else subassemblies
        self._surfaces_cache = None
        _STRUCTURE_VERSION[0] += 1
        HasFrame.__init__(self, location, rotation)

    # -- queries
    def get_local_objects(self):
        return self._objects

    def get_assemblies(self):
        return self._assemblies

    def get_objects(self):
        """Objects of the sub-assemblies first (depth first), then the local ones
        (same order as assembly.py:60-65, which defines surface indices)."""
        out = []
        for asm in self._assemblies:
            out.extend(asm.get_objects())
        out.extend(self._objects)
        return out

    def get_surfaces(self):
        key = (_STRUCTURE_VERSION[0], len(self._objects), len(self._assemblies))
        cache = self.__dict__.get('_surfaces_cache')
        if cache is None or cache[0] != key:
            objs = self.get_objects()
            if any(getattr(obj, 'is_triangle_mesh', False) for obj in objs):
                # a mesh object's surfaces only exist when indexed (stl_utils.TriangleMeshObject)
                cache = (key, SurfaceSeq([obj.get_surfaces() for obj in objs]))
            else:
                cache = (key, [s for obj in objs for s in obj.get_surfaces()])
            self._surfaces_cache = cache
        return cache[1] if isinstance(cache[1], SurfaceSeq) else list(cache[1])

    def global_to_local(self, points):
        return self._to_local(points)

    # -- mutation
    def add_object(self, object, transform=None):
        _STRUCTURE_VERSION[0] += 1
        self._objects.append(object)
        if transform is not None:
            object.set_transform(transform)
        self.transform_children()

    def add_assembly(self, assembly, transform=None):
        _STRUCTURE_VERSION[0] += 1
        self._assemblies.append(assembly)
        if transform is not None:
            assembly.set_transform(transform)
        self.transform_children()

    def set_rotation(self, rotation):
        HasFrame.set_rotation(self, rotation)
        self.transform_children()

    def set_location(self, location):
        HasFrame.set_location(self, location)
        self.transform_children()

    def set_transform(self, transform):
        HasFrame.set_transform(self, transform)
        self.transform_children()

    def transform_children(self, assembly_transform=np.eye(4)):
        mine = np.dot(assembly_transform, self.get_transform())
        for child in self._assemblies + self._objects:
            child.transform_children(mine)

    def reset_all_optics(self):
        self.get_surfaces()                           # fills / validates the cached surface list
        cached = self._surfaces_cache[1]
        if not isinstance(cached, SurfaceSeq):        # plain scenes: no walk over the object tree per trace
            for s in cached:
                reset = getattr(s._opt, 'reset', None)
                if reset is not None:
                    reset()
            return
        for obj in self.get_objects():
            if getattr(obj, 'is_triangle_mesh', False):
                obj.reset_optics()
                continue
            for s in obj.get_surfaces():
                if hasattr(s.get_optics_manager(), 'reset'):
                    s.get_optics_manager().reset()


class AssembledObject(Assembly):
    """A rigid group of surfaces (+ optional boundary boxes) (object.py:8-79)."""

    def __init__(self, surfs=None, bounds=None, location=None, rotation=None, transform=None):
        _STRUCTURE_VERSION[0] += 1
        self.surfaces = [] if surfs is None else surfs
        self.boundaries = [] if bounds is None else bounds
        if not hasattr(self.surfaces, '__len__'):
            self.surfaces = [self.surfaces]
        if not hasattr(self.boundaries, '__len__'):
            self.boundaries = [self.boundaries]
        if transform is None:
            transform = np.eye(4)
            if location is not None:
                transform[:3, 3] = location
            if rotation is not None:
                transform[:3, :3] = rotation
        self.set_transform(transform)

    def get_objects(self):
        return [self]

    def get_surfaces(self):
        return self.surfaces

    def get_boundaries(self):
        if self.boundaries is None:
            self.boundaries = []
        return self.boundaries

    def add_surface(self, surface):
        _STRUCTURE_VERSION[0] += 1
        self.surfaces.append(surface)
        self.transform_children()

    def add_boundary(self, boundary):
        self.boundaries.append(boundary)
        self.transform_children()

    def transform_children(self, assembly_transform=np.eye(4)):
        mine = np.dot(assembly_transform, self.get_transform())
        for child in self.surfaces + self.boundaries:
            child.transform_frame(mine)

    # Legacy relevance hooks (object.py:81-112).  Only the defaults are honoured
    # by the GPU engine; overriding them raises at compile time.
    def own_rays(self, rays, surface_id):
        return np.zeros(rays.get_num_rays(), dtype=bool)

    def surfaces_for_next_iteration(self, rays, surface_id):
        return np.ones((len(self.surfaces), rays.get_num_rays()), dtype=bool)
