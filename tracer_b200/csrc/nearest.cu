// Nearest hit across all surfaces for every ray of a bundle.
//
// Replaces TracerEngine.intersect_ray (tracer_engine.py:27-64) together with
// every GeometryManager.find_intersections it calls and, for the tree walk,
// KdTree.traversal (accel_tree.py:213-312).  Candidate accepted iff t < best,
// or t == best from a lower surface index: the reference's strict '<' over
// ascending surface order, made independent of the visiting order.
//
// Two kernels:
//   nearest_hit_brute_kernel : few surfaces; one thread per ray, surface table
//                              staged through shared memory (broadcast reads).
//   nearest_hit_walk_kernel  : persistent warps over chunks of consecutive rays:
//       stage A (dense)  - load 32 rays, float32 set-up, root-box test; survivors
//                          go to the warp's active list in shared memory;
//       stage B (refill) - lanes walk the float32 BVH; a lane whose ray is done
//                          takes the next survivor from the list at once, so
//                          rays that die early do not idle their lane;
//       stage C (dense)  - (ray, surface) candidates found at leaves are queued
//                          in shared memory and tested 32 at a time in float64
//                          with the reference's arithmetic (geom.cuh); the
//                          per-ray best (t, index) lives in shared memory and is
//                          read back by the walk for pruning.
//   The float32 tree only prunes; every accepted hit is a float64 test, so hit
//   indices stay bit-exact against the brute-force kernel and the oracle.
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include "common.cuh"
#include "geom.cuh"

using namespace trc;

namespace {

__device__ __forceinline__ void consider(const TrcSurfaceRec *s, int idx, V3 v, V3 d, double &best_t, int &best_s) {
    double t = intersect(s, v, d);
    if (t < best_t || (t == best_t && idx < best_s && t != TRC_INF)) { best_t = t; best_s = idx; }
}

#define TRC_STAGE_SURFS 96

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

// ------------------------------------------------------------------ tree walk
constexpr int kWalkWarps = 4;                 // warps per CTA
constexpr int kChunk = 256;                   // consecutive rays owned by a warp at a time
constexpr int kListCap = 64;                  // active list: filled while < 32, by <= 32 at a time
constexpr int kQueueCap = 31 + 32 * 2 * TRC_BVH_MAX_LEAF;   // leftover + one visit of every lane
constexpr int kStack = 64;
constexpr int kListFields = 10;               // inv[3], ovl[3], ovh[3], slot

struct WalkShared {
    unsigned long long best_t[kChunk];        // bits of the best float64 t (+inf = none)
    int best_s[kChunk];                       // its surface index (INT_MAX = none)
    float list[kListFields][kListCap];        // SoA: lanes read consecutive entries conflict-free
    int2 queue[kQueueCap];                    // (slot, surface)
    int qn;                                   // queue fill
};

struct WalkParams {
    const TrcSurfaceRec *surfs;
    const TrcNode32 *nodes;
    const int32_t *leaf_surfs;
    const int32_t *unbounded;
    int n_unbounded;
    int has_tree;
    float extent[3];
    int64_t n;
    const double *vert, *dir;
    int64_t ld;
    int32_t *hit_surf;
    unsigned long long *chunk_counter;
};

#define TRC_INF_BITS 0x7ff0000000000000ULL

// Per-ray float32 walk state: t(box plane) = fma(plane, inv, nov).  The box is
// widened per ray by pad_k = 1e-6 (|v_k| + extent_k): that covers the rounding
// of the origin and of 1/d to float32 and of the two FMAs, all of which are
// <= a few 2^-24 (|v_k| + |plane|) in position units.  `lo` planes use the
// origin moved by +pad, `hi` planes by -pad, which widens the slab for either
// sign of d.  1/d is clamped to +-1e20 so an axis-parallel ray gives a huge
// finite interval instead of 0 * inf.
struct RayF {
    float inv[3], novl[3], novh[3];
};

__device__ __forceinline__ void ray_setup(V3 v, V3 d, const float *extent, RayF &r) {
    const float dv[3] = {(float)d.x, (float)d.y, (float)d.z};
    const float ov[3] = {(float)v.x, (float)v.y, (float)v.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float inv = fabsf(dv[k]) < 1e-20f ? copysignf(1e20f, dv[k]) : __fdividef(1.0f, dv[k]);
        float pad = 1e-6f * (fabsf(ov[k]) + extent[k]);
        r.inv[k] = inv;
        r.novl[k] = -(ov[k] + pad) * inv;
        r.novh[k] = -(ov[k] - pad) * inv;
    }
}

#define TRC_FINF (__int_as_float(0x7f800000))

// Conservative entry distance of the ray into a node box, +inf when the box is
// missed or lies beyond tbest.  a = (lo0, lo1, lo2, hi0), b = (hi1, hi2, first, count).
__device__ __forceinline__ float slab32(float4 a, float4 b, const RayF &r, float tbest) {
    float t1 = fmaf(a.x, r.inv[0], r.novl[0]), t2 = fmaf(a.w, r.inv[0], r.novh[0]);
    float tmin = fminf(t1, t2), tmax = fmaxf(t1, t2);
    t1 = fmaf(a.y, r.inv[1], r.novl[1]); t2 = fmaf(b.x, r.inv[1], r.novh[1]);
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = fmaf(a.z, r.inv[2], r.novl[2]); t2 = fmaf(b.y, r.inv[2], r.novh[2]);
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    tmin = fmaxf(tmin, 0.0f);
    return (tmin <= tmax && tmin <= tbest) ? tmin : TRC_FINF;
}

__device__ __forceinline__ float best_as_float(unsigned long long bits) {
    return __double2float_ru(__longlong_as_double((long long)bits));
}

// Stage C: test queued (slot, surface) candidates in float64, 32 at a time, and
// fold them into the per-slot best (t, surface).  `all` drains the queue.
__device__ __forceinline__ void flush_queue(WalkShared &sh, const WalkParams &p, int64_t chunk_base, int lane,
                                            bool all) {
    int qn = sh.qn;
    while (qn >= 32 || (all && qn > 0)) {
        int take = min(32, qn);
        int base = qn - take;
        bool has = lane < take;
        int2 e = has ? sh.queue[base + lane] : make_int2(0, 0);
        double t = TRC_INF;
        if (has) {
            int64_t ray = chunk_base + e.x;
            V3 v = mk(p.vert[ray], p.vert[p.ld + ray], p.vert[2 * p.ld + ray]);
            V3 d = mk(p.dir[ray], p.dir[p.ld + ray], p.dir[2 * p.ld + ray]);
            t = intersect(&p.surfs[e.y], v, d);
        }
        unsigned long long tb = (unsigned long long)__double_as_longlong(t);
        bool cand = has && tb < TRC_INF_BITS;
        unsigned long long before = cand ? sh.best_t[e.x] : 0ULL;
        __syncwarp();
        if (cand) atomicMin(&sh.best_t[e.x], tb);
        __syncwarp();
        bool win = cand && sh.best_t[e.x] == tb;
        if (win && tb < before) sh.best_s[e.x] = INT_MAX;     // the old index belonged to a larger t
        __syncwarp();
        if (win) atomicMin(&sh.best_s[e.x], e.y);              // ties: lowest surface index
        __syncwarp();
        qn = base;
    }
    if (lane == 0) sh.qn = qn;
    __syncwarp();
}

__global__ void __launch_bounds__(kWalkWarps * 32)
nearest_hit_walk_kernel(WalkParams p) {
    __shared__ WalkShared shared[kWalkWarps];
    WalkShared &sh = shared[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int64_t n_chunks = (p.n + kChunk - 1) / kChunk;

    // root of the tree (uniform loads)
    float4 root_a = make_float4(0.f, 0.f, 0.f, 0.f), root_b = root_a;
    int root_first = 0, root_count = -1;
    if (p.has_tree) {
        const float4 *np = reinterpret_cast<const float4 *>(p.nodes);
        root_a = __ldg(np); root_b = __ldg(np + 1);
        root_first = __float_as_int(root_b.z); root_count = __float_as_int(root_b.w);
    }
    int stack[kStack];

    for (;;) {
        unsigned long long c64 = 0;
        if (lane == 0) c64 = atomicAdd(p.chunk_counter, 1ULL);
        c64 = __shfl_sync(full, c64, 0);
        if ((int64_t)c64 >= n_chunks) break;
        const int64_t chunk_base = (int64_t)c64 * kChunk;
        const int chunk_n = (int)min((int64_t)kChunk, p.n - chunk_base);

        for (int i = lane; i < kChunk; i += 32) { sh.best_t[i] = TRC_INF_BITS; sh.best_s[i] = INT_MAX; }
        if (lane == 0) sh.qn = 0;
        __syncwarp();

        int next = 0;          // rays of the chunk already through stage A   (warp-uniform)
        int list_n = 0;        // active-list fill                              (warp-uniform)
        bool active = false;   // this lane is walking a ray
        RayF r;
        int slot = 0, cur = 0, sp = 0;
        float tbest = TRC_FINF;

        for (;;) {
            // ---------------- stage A: produce survivors until the list can feed a full refill
            while (list_n < 32 && next < chunk_n) {
                int j = next + lane;
                bool valid = j < chunk_n;
                next += 32;
                RayF nr;
                bool surv = false;
                if (valid) {
                    int64_t ray = chunk_base + j;
                    V3 v = mk(p.vert[ray], p.vert[p.ld + ray], p.vert[2 * p.ld + ray]);
                    V3 d = mk(p.dir[ray], p.dir[p.ld + ray], p.dir[2 * p.ld + ray]);
                    ray_setup(v, d, p.extent, nr);
                    surv = p.has_tree && slab32(root_a, root_b, nr, TRC_FINF) < TRC_FINF;
                }
                if (root_count > 0) {
                    // the whole tree is one leaf: its surfaces are candidates of every surviving ray
                    for (int k = 0; k < root_count; ++k) {
                        unsigned m = __ballot_sync(full, surv);
                        int qn = sh.qn;
                        if (surv) sh.queue[qn + __popc(m & lt_mask)] = make_int2(j, __ldg(&p.leaf_surfs[root_first + k]));
                        __syncwarp();
                        if (lane == 0) sh.qn = qn + __popc(m);
                        __syncwarp();
                        flush_queue(sh, p, chunk_base, lane, false);
                    }
                } else {
                    unsigned m = __ballot_sync(full, surv);
                    if (surv) {
                        int e = list_n + __popc(m & lt_mask);
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            sh.list[k][e] = nr.inv[k]; sh.list[3 + k][e] = nr.novl[k]; sh.list[6 + k][e] = nr.novh[k];
                        }
                        sh.list[9][e] = __int_as_float(j);
                    }
                    list_n += __popc(m);
                }
                // surfaces without a finite box are candidates of every ray (accel_tree.py:59-73)
                for (int k = 0; k < p.n_unbounded; ++k) {
                    unsigned m = __ballot_sync(full, valid);
                    int qn = sh.qn;
                    if (valid) sh.queue[qn + __popc(m & lt_mask)] = make_int2(j, __ldg(&p.unbounded[k]));
                    __syncwarp();
                    if (lane == 0) sh.qn = qn + __popc(m);
                    __syncwarp();
                    flush_queue(sh, p, chunk_base, lane, false);
                }
                __syncwarp();
            }
            // ---------------- refill idle lanes from the list
            {
                unsigned idle = __ballot_sync(full, !active);
                int rk = __popc(idle & lt_mask);
                if (!active && rk < list_n) {
                    int e = list_n - 1 - rk;
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        r.inv[k] = sh.list[k][e]; r.novl[k] = sh.list[3 + k][e]; r.novh[k] = sh.list[6 + k][e];
                    }
                    slot = __float_as_int(sh.list[9][e]);
                    cur = root_first; sp = 0; active = true;
                    tbest = best_as_float(sh.best_t[slot]);
                }
                list_n -= min(__popc(idle), list_n);
                __syncwarp();
            }
            if (__ballot_sync(full, active) == 0u) break;   // list empty and chunk exhausted

            // ---------------- stage B: walk until enough lanes are idle to be worth a refill
            for (;;) {
                if (active) {
                    const float4 *np = reinterpret_cast<const float4 *>(p.nodes + cur);
                    float4 a0 = __ldg(np), b0 = __ldg(np + 1), a1 = __ldg(np + 2), b1 = __ldg(np + 3);
                    float t0 = slab32(a0, b0, r, tbest), t1 = slab32(a1, b1, r, tbest);
                    int f0 = __float_as_int(b0.z), c0 = __float_as_int(b0.w);
                    int f1 = __float_as_int(b1.z), c1 = __float_as_int(b1.w);
                    bool h0 = t0 < TRC_FINF && c0 >= 0, h1 = t1 < TRC_FINF && c1 >= 0;
                    int m0 = h0 ? c0 : 0, m1 = h1 ? c1 : 0;
                    if (m0 + m1) {
                        int pos = atomicAdd(&sh.qn, m0 + m1);
                        for (int k = 0; k < m0; ++k) sh.queue[pos++] = make_int2(slot, __ldg(&p.leaf_surfs[f0 + k]));
                        for (int k = 0; k < m1; ++k) sh.queue[pos++] = make_int2(slot, __ldg(&p.leaf_surfs[f1 + k]));
                    }
                    bool i0 = h0 && c0 == 0, i1 = h1 && c1 == 0;
                    if (i0 && i1) {
                        bool swap = t1 < t0;
                        cur = swap ? f1 : f0;
                        stack[sp++] = swap ? f0 : f1;
                    } else if (i0) cur = f0;
                    else if (i1) cur = f1;
                    else if (sp > 0) cur = stack[--sp];
                    else active = false;
                }
                __syncwarp();
                if (sh.qn >= 32) {
                    flush_queue(sh, p, chunk_base, lane, false);
                    if (active) tbest = best_as_float(sh.best_t[slot]);
                }
                int n_active = __popc(__ballot_sync(full, active));
                if (n_active == 0) break;
                if (n_active <= 28 && (list_n > 0 || next < chunk_n)) break;
            }
        }
        flush_queue(sh, p, chunk_base, lane, true);
        for (int i = lane; i < chunk_n; i += 32)
            p.hit_surf[chunk_base + i] = sh.best_t[i] == TRC_INF_BITS ? -1 : sh.best_s[i];
        __syncwarp();
    }
}

}  // namespace

extern "C" int trc_nearest_hit(const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir,
                               int64_t ld, int32_t use_bvh, int32_t *d_hit_surf, void *stream) {
    if (!sc || n < 0 || (n > 0 && (!d_vert || !d_dir || !d_hit_surf)) || ld < n)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: bad argument%s");
    if (n == 0) return TRC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (!use_bvh) {
        int64_t blocks = (n + 255) / 256;
        if (blocks > 0x7fffffffLL) return fail(TRC_ERR_CAPACITY, "trc_nearest_hit: bundle too large, chunk it%s");
        nearest_hit_brute_kernel<<<(unsigned)blocks, 256, 0, st>>>(sc->d_surfaces, sc->n_surfaces, n, d_vert, d_dir,
                                                                   ld, d_hit_surf);
        return check_launch("nearest_hit_brute");
    }
    if ((sc->n_nodes == 0 && sc->n_unbounded == 0) || !sc->d_chunk_counter)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: use_bvh set but no BVH uploaded%s");
    static int occ = 0;
    if (occ == 0) {
        TRC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, nearest_hit_walk_kernel, kWalkWarps * 32, 0));
        if (occ < 1) occ = 1;
    }
    WalkParams p;
    p.surfs = sc->d_surfaces; p.nodes = sc->d_nodes32; p.leaf_surfs = sc->d_leaf_surfaces;
    p.unbounded = sc->d_unbounded; p.n_unbounded = sc->n_unbounded;
    p.has_tree = sc->n_nodes > 0 ? 1 : 0;
    for (int k = 0; k < 3; ++k) p.extent[k] = sc->extent[k];
    p.n = n; p.vert = d_vert; p.dir = d_dir; p.ld = ld; p.hit_surf = d_hit_surf;
    p.chunk_counter = sc->d_chunk_counter;
    int64_t n_chunks = (n + kChunk - 1) / kChunk;
    int64_t blocks = (n_chunks + kWalkWarps - 1) / kWalkWarps;
    int64_t resident = (int64_t)sc->num_sms * occ;
    if (blocks > resident) blocks = resident;
    TRC_CUDA(cudaMemsetAsync(sc->d_chunk_counter, 0, sizeof(unsigned long long), st));
    nearest_hit_walk_kernel<<<(unsigned)blocks, kWalkWarps * 32, 0, st>>>(p);
    return check_launch("nearest_hit_walk");
}
