"""CPU estimate of the tree-walk work on the NSTTF scene: pair visits and leaf candidates per ray
(ordered traversal without pruning by the best hit = upper bound).  Tuning aid, not a test."""
import sys
sys.path.insert(0, '/root/repo')
import numpy as np
from tests import scenes
from bench import nsttf_source_geometry
from tracer_b200.scene import CompiledScene

T = scenes.NS('tracer_b200')
asm, _ = scenes.c3_nsttf(T, n_rays=10)
leaf = int(sys.argv[1]) if len(sys.argv) > 1 else 1
sc = CompiledScene(asm, leaf_size=leaf, upload=False)
nodes = sc.bvh_nodes
centre, direction, radius = nsttf_source_geometry()
n = 200000
np.random.seed(1)
src = T.sources.buie_sunshape(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False)
V, D = src.get_vertices(), src.get_directions()


def slab(lo, hi, V, D):
    with np.errstate(divide='ignore', invalid='ignore'):
        inv = 1. / D
        t1 = (lo[:, None] - V) * inv
        t2 = (hi[:, None] - V) * inv
    tmin = np.nanmax(np.minimum(t1, t2), axis=0)
    tmax = np.nanmin(np.maximum(t1, t2), axis=0)
    return (tmin <= tmax) & (tmax >= 0)


alive = {0: np.nonzero(slab(nodes['lo'][0], nodes['hi'][0], V, D))[0]}
print('nodes', len(nodes), 'root pass %.3f' % (len(alive[0]) / n))
visits = np.zeros(n, int)
cands = np.zeros(n, int)
depth = 0
while alive:
    nxt = {}
    for node, rays in alive.items():
        nd = nodes[node]
        if nd['count'] > 0:
            cands[rays] += nd['count']
            continue
        visits[rays] += 1
        for c in (nd['first'], nd['first'] + 1):
            ok = slab(nodes['lo'][c], nodes['hi'][c], V[:, rays], D[:, rays])
            if ok.any():
                nxt[c] = rays[ok]
    alive = nxt
    depth += 1
print('depth', depth)
print('pair visits per ray (all rays) %.2f, per surviving ray %.2f, max %d' % (
    visits.mean(), visits[visits > 0].mean(), visits.max()))
print('leaf candidates per ray %.3f; rays with >=1 candidate %.3f' % (cands.mean(), (cands > 0).mean()))
print('visit histogram', np.bincount(visits)[:40])
