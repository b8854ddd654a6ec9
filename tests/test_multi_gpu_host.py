"""CPU (gloo, world_size 2): the N>1 path of TracerEngineMP -- every rank packs
[absorbed per surface | hits per surface | flux-map bins], one all-reduce(sum)
merges them (reference: the accountant merge of tracer_engine_mp.py:64-119
reduced to its sums), trees merge with the parent-offset rule (:54-62)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from tracer_b200.tracer_engine_mp import TracerEngineMP
    eng = TracerEngineMP.__new__(TracerEngineMP)        # no GPU here: exercise reduce_tallies only
    S = 5
    eng.absorbed_per_surface = torch.arange(S, dtype=torch.float64) * (rank + 1)
    eng.hits_per_surface = torch.arange(S, dtype=torch.int64) + 10 * rank
    bins = torch.full((4, 4), float(rank + 1), dtype=torch.float64)
    absorbed, hits, b = eng.reduce_tallies(bins)
    # the enqueue-only form gives the same sums when its handle is read
    pend = eng.reduce_tallies(bins, async_op=True)
    a2, h2, b2 = pend.result()
    assert torch.equal(a2, absorbed) and torch.equal(h2, hits) and torch.equal(b2, b)
    assert pend.result()[0] is a2
    out[rank] = (absorbed.numpy().copy(), hits.numpy().copy(), b.numpy().copy())
    dist.destroy_process_group()


def test_reduce_tallies_world_size_2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    S = 5
    want_abs = np.arange(S) * 1. + np.arange(S) * 2.
    want_hits = 2 * np.arange(S) + 10
    for rank in range(world):
        absorbed, hits, bins = out[rank]
        np.testing.assert_allclose(absorbed, want_abs)
        np.testing.assert_array_equal(hits, want_hits)
        np.testing.assert_allclose(bins, np.full((4, 4), 3.))


def test_merge_trees_parent_offsets():
    from tracer_b200.ray_bundle import RayBundle
    from tracer_b200.trace_tree import RayTree
    from tracer_b200.tracer_engine_mp import merge_trees

    def tree(n0, parents1):
        t = RayTree()
        t.append(RayBundle(vertices=np.zeros((3, n0)), directions=np.zeros((3, n0)), energy=np.ones(n0),
                           parents=np.arange(n0)))
        n1 = len(parents1)
        t.append(RayBundle(vertices=np.zeros((3, n1)), directions=np.zeros((3, n1)), energy=np.ones(n1),
                           parents=np.asarray(parents1)))
        return t
    merged = merge_trees([tree(3, [2, 0]), tree(4, [1, 3, 3])])
    assert merged[0].get_num_rays() == 7
    np.testing.assert_array_equal(merged[1].get_parents(), [2, 0, 4, 6, 6])
