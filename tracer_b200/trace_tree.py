"""RayTree: the per-bounce record of a trace (reference tracer/trace_tree.py:6-55).

Bundle k's ``parents`` index bundle k-1 (survivors-first order), so a ray's
ancestry is recovered by walking the parent arrays backwards.
"""
import numpy as np


class RayTree(object):
    def __init__(self):
        self._bunds = []

    def __getitem__(self, level):
        return self._bunds[level]

    def __len__(self):
        return len(self._bunds)

    def num_bunds(self):
        return len(self._bunds)

    def append(self, bund):
        self._bunds.append(bund)

    def ordered_parents(self):
        """Parent arrays from the first hit of the source bundle to the last bundle."""
        return [b.get_parents() for b in self._bunds[1:]]

    def ray_history(self, ray_index, level=None):
        """Indices of a ray and its ancestors, newest first (trace_tree.py:34-55)."""
        if level is None:
            level = self.num_bunds()
        chain = np.empty(level, dtype=int)
        chain[0] = ray_index
        for back in range(1, level):
            chain[back] = self._bunds[level - back].get_parents()[chain[back - 1]]
        return chain
