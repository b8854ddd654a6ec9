// libtracer_b200.so -- sm_100a kernels and C ABI for the Tracer bounce loop.
// See include/tracer_b200.h for the boundary and DESIGN.md for the data layout.
//
// Pipeline of one bounce (all on the caller's stream, SoA float64 buffers):
//   1. nearest_hit_*   : per ray, nearest surface index (brute force over a
//                        shared-memory staged surface table, or BVH walk)
//   2. sort (radix)    : stable (surface, ray) ordering of the hits
//   3. segment_offsets : per-surface slot ranges
//   4. optics          : per slot, hit point + normal + optics + tallies
//   5. sort + gather   : survivors first, culled rays at the back
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <new>

#include "common.cuh"
#include "geom.cuh"
#include "optics.cuh"
#include "philox.cuh"

using namespace trc;

// ------------------------------------------------------------------ errors
static thread_local char g_err[TRC_ERR_BUF] = "";
char *trc_err_buf() { return g_err; }

#include <atomic>
static std::atomic<unsigned long long> g_launches{0};
void trc_note_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
extern "C" uint64_t trc_launch_count(void) { return (uint64_t)g_launches.load(std::memory_order_relaxed); }

extern "C" int trc_abi_version(void) { return TRC_ABI_VERSION; }
extern "C" const char *trc_last_error(void) { return g_err; }

// ------------------------------------------------------------------- scene
extern "C" int trc_scene_create(const TrcSurfaceRec *surfaces, int32_t n_surfaces,
                                const double *tables, int32_t n_table, TrcScene **out) {
    if (!surfaces || n_surfaces <= 0 || !out) return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_create: bad argument%s");
    for (int32_t i = 0; i < n_surfaces; ++i) {
        if (surfaces[i].gkind < 0 || surfaces[i].gkind >= TRC_G_KIND_COUNT)
            return fail(TRC_ERR_UNSUPPORTED_KIND, "trc_scene_create: unknown geometry kind%s");
        if (surfaces[i].okind < 0 || surfaces[i].okind >= TRC_O_KIND_COUNT)
            return fail(TRC_ERR_UNSUPPORTED_KIND, "trc_scene_create: unknown optics kind%s");
    }
    TrcScene *sc = new (std::nothrow) TrcScene();
    if (!sc) return fail(TRC_ERR_CAPACITY, "trc_scene_create: out of host memory%s");
    sc->n_surfaces = n_surfaces;
    cudaError_t e = cudaMalloc(&sc->d_surfaces, sizeof(TrcSurfaceRec) * (size_t)n_surfaces);
    if (e == cudaSuccess)
        e = cudaMemcpy(sc->d_surfaces, surfaces, sizeof(TrcSurfaceRec) * (size_t)n_surfaces, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n_table > 0 && tables) {
        sc->n_table = n_table;
        e = cudaMalloc(&sc->d_tables, sizeof(double) * (size_t)n_table);
        if (e == cudaSuccess) e = cudaMemcpy(sc->d_tables, tables, sizeof(double) * (size_t)n_table, cudaMemcpyHostToDevice);
    }
    int dev = 0;
    if (e == cudaSuccess) e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sc->num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) {
        trc_scene_destroy(sc);
        return fail(TRC_ERR_CUDA, "trc_scene_create: %s", cudaGetErrorString(e));
    }
    *out = sc;
    return TRC_OK;
}

extern "C" int trc_scene_destroy(TrcScene *sc) {
    if (!sc) return TRC_OK;
    cudaFree(sc->d_surfaces);
    cudaFree(sc->d_tables);
    cudaFree(sc->d_nodes);
    cudaFree(sc->d_leaf_surfaces);
    cudaFree(sc->d_unbounded);
    delete sc;
    return TRC_OK;
}

extern "C" int32_t trc_scene_num_surfaces(const TrcScene *sc) { return sc ? sc->n_surfaces : 0; }

__global__ void update_frames_kernel(TrcSurfaceRec *s, const double *fr, const double *iv, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 12) return;
    int k = i / 12, j = i % 12;
    s[k].frame[j] = fr[i];
    s[k].inv[j] = iv[i];
}

extern "C" int trc_scene_update_frames(TrcScene *sc, const double *frames12, const double *inv12, int32_t n) {
    if (!sc || !frames12 || !inv12 || n != sc->n_surfaces)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_update_frames: bad argument%s");
    double *tmp = nullptr;
    size_t bytes = sizeof(double) * 12 * (size_t)n;
    TRC_CUDA(cudaMalloc(&tmp, 2 * bytes));
    cudaError_t e = cudaMemcpy(tmp, frames12, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(tmp + 12 * (size_t)n, inv12, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        update_frames_kernel<<<(n * 12 + 255) / 256, 256>>>(sc->d_surfaces, tmp, tmp + 12 * (size_t)n, n);
        e = cudaDeviceSynchronize();
    }
    cudaFree(tmp);
    if (e != cudaSuccess) return fail(TRC_ERR_CUDA, "trc_scene_update_frames: %s", cudaGetErrorString(e));
    return TRC_OK;
}

extern "C" int trc_scene_set_bvh(TrcScene *sc, const TrcBvhNode *nodes, int32_t n_nodes,
                                 const int32_t *leaf_surfaces, int32_t n_leaf_surfaces,
                                 const int32_t *unbounded, int32_t n_unbounded) {
    if (!sc || n_nodes < 0 || n_leaf_surfaces < 0 || n_unbounded < 0)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: bad argument%s");
    cudaFree(sc->d_nodes); sc->d_nodes = nullptr; sc->n_nodes = 0;
    cudaFree(sc->d_leaf_surfaces); sc->d_leaf_surfaces = nullptr; sc->n_leaf_surfaces = 0;
    cudaFree(sc->d_unbounded); sc->d_unbounded = nullptr; sc->n_unbounded = 0;
    if (n_nodes > 0) {
        TRC_CUDA(cudaMalloc(&sc->d_nodes, sizeof(TrcBvhNode) * (size_t)n_nodes));
        TRC_CUDA(cudaMemcpy(sc->d_nodes, nodes, sizeof(TrcBvhNode) * (size_t)n_nodes, cudaMemcpyHostToDevice));
        sc->n_nodes = n_nodes;
    }
    if (n_leaf_surfaces > 0) {
        TRC_CUDA(cudaMalloc(&sc->d_leaf_surfaces, sizeof(int32_t) * (size_t)n_leaf_surfaces));
        TRC_CUDA(cudaMemcpy(sc->d_leaf_surfaces, leaf_surfaces, sizeof(int32_t) * (size_t)n_leaf_surfaces, cudaMemcpyHostToDevice));
        sc->n_leaf_surfaces = n_leaf_surfaces;
    }
    if (n_unbounded > 0) {
        TRC_CUDA(cudaMalloc(&sc->d_unbounded, sizeof(int32_t) * (size_t)n_unbounded));
        TRC_CUDA(cudaMemcpy(sc->d_unbounded, unbounded, sizeof(int32_t) * (size_t)n_unbounded, cudaMemcpyHostToDevice));
        sc->n_unbounded = n_unbounded;
    }
    return TRC_OK;
}

// ------------------------------------------------------------- nearest hit
// Replaces TracerEngine.intersect_ray (tracer_engine.py:27-64).  Candidate
// accepted iff t < best, or t == best from a lower surface index, which is the
// reference's strict '<' over ascending surface order made order-independent
// (needed because the BVH visits surfaces out of index order).
__device__ __forceinline__ void consider(const TrcSurfaceRec *s, int idx, V3 v, V3 d, double &best_t, int &best_s) {
    double t = intersect(s, v, d);
    if (t < best_t || (t == best_t && idx < best_s && t != TRC_INF)) { best_t = t; best_s = idx; }
}

// Geometry part of a surface record (frame, inv, gp, kinds): 304 of 352 bytes.
#define TRC_STAGE_SURFS 96

// Brute force: every ray against every surface, surface table staged through
// shared memory in chunks so all 256 threads broadcast-read the same record.
__global__ void __launch_bounds__(256)
nearest_hit_brute_kernel(const TrcSurfaceRec *__restrict__ surfs, int n_surf, int64_t n,
                         const double *__restrict__ vert, const double *__restrict__ dir, int64_t ld,
                         int32_t *__restrict__ hit_surf) {
    __shared__ TrcSurfaceRec stage[TRC_STAGE_SURFS];
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool live = i < n;
    V3 v = mk(0, 0, 0), d = mk(0, 0, 1);
    if (live) {
        v = mk(vert[i], vert[ld + i], vert[2 * ld + i]);
        d = mk(dir[i], dir[ld + i], dir[2 * ld + i]);
    }
    double best_t = TRC_INF;
    int best_s = -1;
    for (int base = 0; base < n_surf; base += TRC_STAGE_SURFS) {
        int cnt = min(TRC_STAGE_SURFS, n_surf - base);
        __syncthreads();
        {
            const uint4 *src = reinterpret_cast<const uint4 *>(surfs + base);
            uint4 *dst = reinterpret_cast<uint4 *>(stage);
            int n16 = cnt * (int)(sizeof(TrcSurfaceRec) / 16);
            for (int k = threadIdx.x; k < n16; k += blockDim.x) dst[k] = src[k];
        }
        __syncthreads();
        if (live) {
            for (int k = 0; k < cnt; ++k) consider(&stage[k], base + k, v, d, best_t, best_s);
        }
    }
    if (live) hit_surf[i] = best_s;
}

// Slab test of a ray against an AABB; returns entry distance or +inf.
// Conservative: boxes are padded on the host, NaNs from 0*inf fall out of
// fmin/fmax (CUDA returns the non-NaN operand).
__device__ __forceinline__ double slab_entry(const TrcBvhNode *nd, V3 v, V3 inv_d, double t_far) {
    double lo0 = __ldg(&nd->lo[0]), lo1 = __ldg(&nd->lo[1]), lo2 = __ldg(&nd->lo[2]);
    double hi0 = __ldg(&nd->hi[0]), hi1 = __ldg(&nd->hi[1]), hi2 = __ldg(&nd->hi[2]);
    double t1 = (lo0 - v.x) * inv_d.x, t2 = (hi0 - v.x) * inv_d.x;
    double tmin = fmin(t1, t2), tmax = fmax(t1, t2);
    t1 = (lo1 - v.y) * inv_d.y; t2 = (hi1 - v.y) * inv_d.y;
    tmin = fmax(tmin, fmin(t1, t2)); tmax = fmin(tmax, fmax(t1, t2));
    t1 = (lo2 - v.z) * inv_d.z; t2 = (hi2 - v.z) * inv_d.z;
    tmin = fmax(tmin, fmin(t1, t2)); tmax = fmin(tmax, fmax(t1, t2));
    tmin = fmax(tmin, 0.0);
    // widen by a relative epsilon so rounding in the slab arithmetic can never
    // cull a box the exact test would accept
    if (tmax * (1.0 + 1e-12) + 1e-12 >= tmin && tmin <= t_far) return tmin;
    return TRC_INF;
}

#define TRC_BVH_STACK 64

// BVH walk: replaces KdTree.traversal + the relevance-masked surface loop
// (accel_tree.py:213-312, tracer_engine.py:45-63).  Ordered descent (near child
// first) with pruning against the current best hit; unbounded surfaces are
// tested first for every ray (accel_tree.py:59-73 "always_relevant").
__global__ void __launch_bounds__(256)
nearest_hit_bvh_kernel(const TrcSurfaceRec *__restrict__ surfs,
                       const TrcBvhNode *__restrict__ nodes, int n_nodes,
                       const int32_t *__restrict__ leaf_surfs,
                       const int32_t *__restrict__ unbounded, int n_unbounded,
                       int64_t n, const double *__restrict__ vert, const double *__restrict__ dir, int64_t ld,
                       int32_t *__restrict__ hit_surf) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 v = mk(vert[i], vert[ld + i], vert[2 * ld + i]);
    V3 d = mk(dir[i], dir[ld + i], dir[2 * ld + i]);
    double best_t = TRC_INF;
    int best_s = -1;
    for (int k = 0; k < n_unbounded; ++k) {
        int s = __ldg(&unbounded[k]);
        consider(&surfs[s], s, v, d, best_t, best_s);
    }
    if (n_nodes > 0) {
        V3 inv_d = mk(1.0 / d.x, 1.0 / d.y, 1.0 / d.z);
        int stack[TRC_BVH_STACK];
        int sp = 0;
        int cur = 0;
        if (slab_entry(&nodes[0], v, inv_d, best_t) == TRC_INF) cur = -1;
        while (cur >= 0) {
            const TrcBvhNode *nd = &nodes[cur];
            int first = __ldg(&nd->first), count = __ldg(&nd->count);
            if (count != 0) {   // leaf (count < 0 marks an empty leaf of the balanced tree)
                for (int k = 0; k < count; ++k) {
                    int s = __ldg(&leaf_surfs[first + k]);
                    consider(&surfs[s], s, v, d, best_t, best_s);
                }
                cur = -1;
            } else {
                double ta = slab_entry(&nodes[first], v, inv_d, best_t);
                double tb = slab_entry(&nodes[first + 1], v, inv_d, best_t);
                int na = first, nb = first + 1;
                if (tb < ta) { double tt = ta; ta = tb; tb = tt; na = first + 1; nb = first; }
                if (ta == TRC_INF) cur = -1;
                else {
                    cur = na;
                    if (tb != TRC_INF && sp < TRC_BVH_STACK) stack[sp++] = nb;
                }
            }
            while (cur < 0 && sp > 0) {
                int cand = stack[--sp];
                // re-test against the (possibly closer) current best
                if (slab_entry(&nodes[cand], v, inv_d, best_t) != TRC_INF) cur = cand;
            }
        }
    }
    hit_surf[i] = best_s;
}

extern "C" int trc_nearest_hit(const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir,
                               int64_t ld, int32_t use_bvh, int32_t *d_hit_surf, void *stream) {
    if (!sc || n < 0 || (n > 0 && (!d_vert || !d_dir || !d_hit_surf)) || ld < n)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: bad argument%s");
    if (n == 0) return TRC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 0x7fffffffLL) return fail(TRC_ERR_CAPACITY, "trc_nearest_hit: bundle too large, chunk it%s");
    if (use_bvh) {
        if (sc->n_nodes == 0 && sc->n_unbounded == 0)
            return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: use_bvh set but no BVH uploaded%s");
        nearest_hit_bvh_kernel<<<(unsigned)blocks, 256, 0, st>>>(sc->d_surfaces, sc->d_nodes, sc->n_nodes,
                                                                 sc->d_leaf_surfaces, sc->d_unbounded, sc->n_unbounded,
                                                                 n, d_vert, d_dir, ld, d_hit_surf);
    } else {
        nearest_hit_brute_kernel<<<(unsigned)blocks, 256, 0, st>>>(sc->d_surfaces, sc->n_surfaces, n, d_vert, d_dir,
                                                                   ld, d_hit_surf);
    }
    return check_launch("nearest_hit");
}
