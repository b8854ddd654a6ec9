"""Acceleration structure with the reference's name (tracer/accel_tree.py:16-365).

In the reference ``KdTree(assembly, ...)`` builds a SAH kd-tree over the objects' boundary boxes
and ``traversal(bundle)`` produces an (S, N) boolean relevance matrix that limits which
``find_intersections`` calls see which rays; hit results are identical with or without it
(SURVEY F10).  Here the structure the nearest-hit kernel walks is a binned-SAH binary BVH over
per-surface boxes (native builder ``trc_bvh_build``) that belongs to the compiled
scene, and no relevance matrix is ever materialised.  ``KdTree`` is therefore a handle: it
compiles the scene tables on the host (no GPU needed), exposes the tree for inspection and, when
passed as ``ray_tracer(..., Kd_Tree=tree)`` or built by ``accel=True``, selects the tree walk.
Constructor arguments keep the reference's names; the SAH cost constants of the kd-tree
(``t_trav``, ``t_isec``, ``empty_bonus``, ``split_threshold``, ``fast``) are accepted and recorded
but the binned builder has its own fixed costs.
"""
import logging

import numpy as np


class KdTree(object):
    def __init__(self, assembly, max_depth=np.inf, min_leaf=1, loglevel=logging.DEBUG, debug=False, fast=False,
                 t_trav=1., t_isec=1000., empty_bonus=0.2, split_threshold=None):
        from .scene import CompiledScene
        from . import _lib as L
        self.loglevel = loglevel
        self.t_trav, self.t_isec, self.empty_bonus = t_trav, t_isec, empty_bonus
        self.split_threshold, self.fast, self.debug = split_threshold, fast, False
        self.min_leaf, self.max_depth = min_leaf, max_depth
        self.objects = [assembly] if not hasattr(assembly, 'get_objects') else assembly.get_objects()
        self.n_surfs = len(assembly.get_surfaces())
        logging.log(self.loglevel, 'Building tree')
        leaf = int(min(max(min_leaf, 1), L.BVH_MAX_LEAF))
        scene = CompiledScene(assembly, leaf_size=leaf, upload=False)
        if len(scene.bounded) == 0:
            # accel_tree.py:47-48
            raise Exception('No boundary defined in the assembly, please revert to non-accelerated ray-tracing')
        self.leaf_size = leaf
        self.nodes = scene.bvh_nodes                       # structured array: lo[3], hi[3], first, count
        self.leaf_surfaces = scene.bvh_leaf_surfaces       # surface index of every leaf entry (tree order)
        self.always_relevant = [int(s) for s in scene.unbounded]   # surfaces without a finite box (:63-64)
        self.minpoint = scene.aabb_lo[scene.bounded].min(axis=0)
        self.maxpoint = scene.aabb_hi[scene.bounded].max(axis=0)
        logging.log(self.loglevel, 'Tree built: %d nodes over %d bounded surfaces' % (len(self.nodes), len(scene.bounded)))

    def depth(self):
        """Number of levels of the tree (root = 1)."""
        def walk(i):
            nd = self.nodes[i]
            if nd['count'] > 0:
                return 1
            return 1 + max(walk(int(nd['first'])), walk(int(nd['first']) + 1))
        return walk(0) if len(self.nodes) else 0

    def intersect_bounds(self, poss, dirs, inv_dirs=None, bounds=None):
        """Rays against a box (accel_tree.py:314-330; default: the root box): (hit mask, t_mins,
        t_maxs).  Same slab rule: entry parameters start at 0, exits at +inf, planes picked by the
        sign of the direction, IEEE infinities for axis-parallel rays, hit iff t_max > 0 and
        t_min <= t_max."""
        poss, dirs = np.asarray(poss, float), np.asarray(dirs, float)
        if bounds is None:
            bounds = np.vstack((self.minpoint, self.maxpoint))
        with np.errstate(divide='ignore', invalid='ignore'):
            if inv_dirs is None:
                inv_dirs = 1. / dirs
            neg = np.array(dirs < 0, dtype=int)
            t_mins = np.zeros(poss.shape[1])
            t_maxs = np.full(poss.shape[1], np.inf)
            for i in range(3):
                near = (bounds[neg[i], i] - poss[i]) * inv_dirs[i]
                far = (bounds[1 - neg[i], i] - poss[i]) * inv_dirs[i]
                lo, hi = np.where(near > far, far, near), np.where(near > far, near, far)
                t_mins = np.maximum(t_mins, lo)
                t_maxs = np.minimum(t_maxs, hi)
        inters = t_maxs > 0
        inters[t_mins > t_maxs] = False
        return inters, t_mins, t_maxs

    def traversal(self, bundle, lightweight=False):
        raise NotImplementedError(
            "KdTree.traversal: the (surfaces x rays) relevance matrix of accel_tree.py:213-312 is not built; "
            "the nearest-hit kernel walks the tree per ray on the GPU (TracerEngine.ray_tracer(accel=True))")
