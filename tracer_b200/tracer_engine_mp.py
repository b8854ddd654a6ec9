"""TracerEngineMP: data parallelism over independent source bundles
(reference tracer/tracer_engine_mp.py:19-123), one process per GPU.

The reference forks a pathos pool, traces one source per process and pickles
whole engines back.  Rays never interact, so here each rank (torchrun, NCCL)
traces its own source shard with zero communication during the bounce loop;
the only exchange is one all-reduce(sum) of a packed tally vector
[absorbed energy per surface | hit count per surface | flux-map bins] at the
end (``reduce_tallies``).  Per-hit lists and tree bundles stay sharded on the
rank that produced them; ``merge_trees`` applies the reference's parent-offset
rule (tracer_engine_mp.py:54-62) when bundles from several traces are joined,
which is also what a single-process ``multi_ray_sim`` over several sources does.
"""
import numpy as np

from .ray_bundle import concatenate_rays
from .trace_tree import RayTree
from .tracer_engine import TracerEngine


def merge_trees(trees):
    """Join RayTrees of independent traces level by level.  Parents of level b
    (b >= 1) are shifted by the number of rays already merged into level b-1
    (tracer_engine_mp.py:54-62)."""
    out = RayTree()
    for tree in trees:
        shift = 0
        for b in range(tree.num_bunds()):
            bund = tree[b]
            if b > 0 and bund.get_num_rays():
                bund = bund.inherit(parents=np.asarray(bund.get_parents()) + shift)
            if b == out.num_bunds():
                shift = 0
                out.append(bund)
            else:
                shift = out[b].get_num_rays()
                out._bunds[b] = concatenate_rays([out[b], bund])
    return out


def pack_tallies(absorbed, hits, bins=None):
    parts = [np.asarray(absorbed, float).ravel(), np.asarray(hits, float).ravel()]
    if bins is not None:
        parts.append(np.asarray(bins, float).ravel())
    return np.concatenate(parts)


class TracerEngineMP(TracerEngine):
    def multi_ray_sim(self, sources, procs=1, minener=1e-10, reps=1000, tree=True):
        """Trace ``sources`` (a list of RayBundles, one per worker).

        * under torch.distributed with world_size == procs: this rank traces
          ``sources[rank]`` (or ``sources[0]`` when only the local shard is given);
        * otherwise the sources are traced one after the other on this GPU and the
          trees are merged as the reference does.
        """
        import torch.distributed as dist
        self.minener, self.reps, self.tree_switch = minener, reps, tree
        distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() == procs and procs > 1
        if distributed:
            if len(sources) not in (1, procs):
                raise Exception('Number of sources and processors do not agree')
            src = sources[dist.get_rank()] if len(sources) == procs else sources[0]
            return self.ray_tracer(src, reps, minener, tree)
        if len(sources) != procs:
            raise Exception('Number of sources and processors do not agree')
        trees, out = [], None
        absorbed = hits = None
        for src in sources:
            out = self.ray_tracer(src, reps, minener, tree)
            trees.append(self.tree)
            absorbed = self.absorbed_per_surface if absorbed is None else absorbed + self.absorbed_per_surface
            hits = self.hits_per_surface if hits is None else hits + self.hits_per_surface
        self.absorbed_per_surface, self.hits_per_surface = absorbed, hits
        if tree:
            self.tree = merge_trees(trees)
        return out

    def reduce_tallies(self, bins=None, group=None, async_op=False):
        """All-reduce(sum) [absorbed per surface | hits per surface | bins] over the
        ranks; returns (absorbed (S,), hits (S,), bins or None) as tensors on the
        device they were on.  Without an initialised process group it is a no-op.
        With ``async_op`` the collective is only enqueued (on the process group's own stream) and a handle is
        returned whose ``result()`` gives the same triple: the caller's stream -- the next trace -- does not wait
        for the slowest rank until somebody reads the sums."""
        import torch
        import torch.distributed as dist
        S = self.absorbed_per_surface.shape[0]
        if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            out = (self.absorbed_per_surface, self.hits_per_surface, bins)
            return _PendingTallies(None, out) if async_op else out
        parts = [self.absorbed_per_surface, self.hits_per_surface.to(torch.float64)]
        if bins is not None:
            parts.append(bins.reshape(-1))
        packed = torch.cat(parts)
        work = dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group, async_op=async_op)

        def unpack():
            absorbed, hits = packed[:S], packed[S:2 * S].round().to(torch.int64)
            out_bins = packed[2 * S:].reshape(bins.shape) if bins is not None else None
            return absorbed, hits, out_bins
        return _PendingTallies(work, unpack) if async_op else unpack()


class _PendingTallies(object):
    """An enqueued all-reduce of the tallies (TracerEngineMP.reduce_tallies(async_op=True))."""

    def __init__(self, work, out):
        self._work, self._out = work, out

    def result(self):
        if self._work is not None:
            self._work.wait()                   # the current stream waits for the collective, the host does not
            self._work = None
        if callable(self._out):
            self._out = self._out()
        return self._out
