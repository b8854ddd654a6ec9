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
#include <stdlib.h>

#include "common.cuh"
#include "geom.cuh"

using namespace trc;

namespace {

__device__ __forceinline__ void consider(const TrcSurfaceRec *s, int idx, V3 v, V3 d, const double *pool,
                                         double &best_t, int &best_s) {
    double t = intersect(s, v, d, pool);
    if (t < best_t || (t == best_t && idx < best_s && t != TRC_INF)) { best_t = t; best_s = idx; }
}

#define TRC_STAGE_SURFS 96

// Optional outputs of both kernels: the int32 hit index per ray and/or the hit
// list (trc_trace_hits): 64-byte records appended chunk by chunk.
struct HitSink {
    int32_t *hit_surf;                 // nullable
    const double *energy;              // parent energies (read for hits only)
    TrcHitRecord *records;             // nullable
    int64_t capacity;
    uint32_t *chunk_off, *chunk_cnt;
    unsigned long long *total;
    int32_t ray_base;                  // added to the ray index stored in a record (slices of a larger bundle)
    int32_t *keys;                     // nullable: surface index of record k at keys[k] (4-byte sort keys, so that
                                       // the ordering kernels do not fetch a 32-byte sector per hit for them)
};

__device__ __forceinline__ void store_record(TrcHitRecord *dst, V3 v, V3 d, double e, int ray, int surf) {
    double2 *o = reinterpret_cast<double2 *>(dst);
    o[0] = make_double2(v.x, v.y);
    o[1] = make_double2(v.z, d.x);
    o[2] = make_double2(d.y, d.z);
    double tail;
    {
        long long bits = ((long long)(unsigned)surf << 32) | (long long)(unsigned)ray;   // little endian: ray, surf
        tail = __longlong_as_double(bits);
    }
    o[3] = make_double2(e, tail);
}

static_assert(TRC_HIT_CHUNK == 256, "one brute-force block = one chunk");

__global__ void __launch_bounds__(256)
nearest_hit_brute_kernel(const TrcSurfaceRec *__restrict__ surfs, const double *__restrict__ pool, int n_surf, int64_t n,
                         const TrcBounceCtl *__restrict__ ctl,
                         const double *__restrict__ vert, const double *__restrict__ dir, int64_t ld,
                         HitSink sink) {
    __shared__ TrcSurfaceRec stage[TRC_STAGE_SURFS];
    __shared__ unsigned warp_cnt[8];
    __shared__ unsigned long long block_base;
    pdl_wait();
    pdl_trigger();
    if (ctl) {                         // device-driven trace: the bundle is described in device memory
        n = ctl->n_in; vert = ctl->in.V; dir = ctl->in.D; ld = ctl->in.ld; sink.energy = ctl->in.E;
    }
    if ((int64_t)blockIdx.x * blockDim.x >= n) return;      // grid sized for the capacity of the input bundle
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
            for (int k = 0; k < cnt; ++k) consider(&stage[k], base + k, v, d, pool, best_t, best_s);
        }
    }
    if (live && sink.hit_surf) sink.hit_surf[i] = best_s;
    if (sink.records) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        bool hit = live && best_s >= 0;
        unsigned m = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) warp_cnt[warp] = __popc(m);
        __syncthreads();
        unsigned before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { unsigned c = warp_cnt[w]; if (w < warp) before += c; total += c; }
        if (threadIdx.x == 0) {
            block_base = atomicAdd(sink.total, (unsigned long long)total);
            sink.chunk_off[blockIdx.x] = (uint32_t)block_base;
            sink.chunk_cnt[blockIdx.x] = total;
        }
        __syncthreads();
        if (hit) {
            unsigned long long pos = block_base + before + __popc(m & ((1u << lane) - 1u));
            if ((int64_t)pos < sink.capacity) {
                store_record(sink.records + pos, v, d, sink.energy[i], sink.ray_base + (int)i, best_s);
                if (sink.keys) sink.keys[pos] = best_s;
            }
        }
    }
}

// ------------------------------------------------------------------ tree walk
#ifndef TRC_WALK_MIN_CTAS
#define TRC_WALK_MIN_CTAS 2
#endif
#ifndef TRC_WALK_MIN_CTAS_FLAT
#define TRC_WALK_MIN_CTAS_FLAT 2
#endif
#ifndef TRC_WALK_WARPS
#define TRC_WALK_WARPS 8
#endif
constexpr int kWalkWarps = TRC_WALK_WARPS;    // warps per CTA
constexpr int kChunk = 256;                   // consecutive rays owned by a warp at a time (slot fits 8 bits)
constexpr int kListCap = 64;                  // active list: filled while < 32, by <= 32 at a time
constexpr int kQueueCap = 32 + 32 * 2;        // leftover + one visit (two leaf children) of every lane
constexpr int kStack = 64;
#ifndef TRC_WALK_STEPS
#define TRC_WALK_STEPS 3
#endif
constexpr int kWalkSteps = TRC_WALK_STEPS;                  // visits per lane between two warp-level collections
constexpr unsigned kNoEntry = 0xffffffffu;    // (first < 2^22 - 1, so no real entry is all ones)
// Ray ring of a warp, two interchangeable transports (TRC_WALK_TMA):
//   1: batches of 64 rays, six 1-D bulk copies (cp.async.bulk -> SASS UBLKCP) issued by one lane, completion on an
//      mbarrier; the rays bypass L1, so the float64 stage and the record emission re-read them from L2;
//   0: batches of 32 rays, six 8-byte cp.async.ca per lane (LDGSTS), which also leave the rays in L1.
#ifndef TRC_WALK_TMA
#define TRC_WALK_TMA 1
#endif
#if TRC_WALK_TMA
#ifndef TRC_WALK_BATCH
#define TRC_WALK_BATCH 64
#endif
#ifndef TRC_WALK_AHEAD
#define TRC_WALK_AHEAD 1
#endif
constexpr int kBatch = TRC_WALK_BATCH;        // rays per batch: six rows of 8 * kBatch bytes, one cp.async.bulk each
constexpr int kAhead = TRC_WALK_AHEAD;        // batches landing while one is consumed
constexpr int kRing = kAhead < 2 ? 2 : 4;     // slots (a power of two)
static_assert(kAhead >= 1 && kAhead <= 3 && (kBatch == 32 || kBatch == 64), "ring shape");
#else
constexpr int kRing = 4;                      // three batches in flight (a power of two keeps the slot index cheap)
constexpr int kBatch = 32;
#endif
constexpr int kListFields = 10;               // A[3], Bn[3], Bf[3], slot | direction signs << 8
#ifndef TRC_WALK_SMEM_PAIRS
#define TRC_WALK_SMEM_PAIRS 512
#endif
#ifndef TRC_GRID_QUEUE
#define TRC_GRID_QUEUE 512
#endif
constexpr int kGridQueue = TRC_GRID_QUEUE;               // candidate queue of the shadow-grid kernel (entries per warp)
constexpr int kGridMaxWords = 6144;           // occupancy bits of a shadow grid in shared memory: at most 24 KB
constexpr int kSmemPairs = TRC_WALK_SMEM_PAIRS;               // top of the tree (breadth-first) cached per CTA: 16 KB
// queue entry = slot << 24 | (count - 1) << 22 | first candidate-list index: one entry per leaf
constexpr unsigned kCandMask = 0x003fffffu;

struct WalkWarpShared {
#ifdef TRC_WALK_EXPERIMENT_NO_F64
    unsigned long long best_t[1];
    int best_s[2];
#else
    unsigned long long best_t[kChunk];        // bits of the best float64 t (+inf = none)
    int best_s[kChunk];                       // its surface index (INT_MAX = none)
#endif
    float list[kListFields][kListCap];        // SoA: lanes read consecutive entries conflict-free
    unsigned queue[kQueueCap];                // packed (slot, leaf)
    unsigned queue2[64];                      // candidates that passed the float32 triangle pre-test
    int q2n;                                  // its fill (warp-uniform, kept here to spare a register)
    long long fifo[4];                        // chunk ids in consumption order (written by the prefetcher)
    alignas(16) double ring[kRing][6][kBatch];   // bulk-copy landing zone (16-byte aligned rows): vertex + direction rows
    unsigned long long mbar[kRing];           // one mbarrier per landing slot (transaction bytes of its six copies)
    const double *vert, *dir, *energy;        // the bundle (kernel parameters, or the control block of a device-driven
    long long ld;                             // trace): read from here so that they cost no registers in the walk loop
};

// Dynamic shared memory of a CTA: the per-warp blocks, then the cached top of the tree,
// pair_a[n_cached] followed by pair_b[n_cached] (TrcPairHalf arrays, see common.cuh) with
// n_cached = min(n_pairs, kSmemPairs) rounded up to 8.  Only what the tree needs is
// allocated: every KB not carved out for shared memory stays L1 for the float64 stage.
struct WalkShared {
    WalkWarpShared w[kWalkWarps];
    long long n_rays;                         // ray count (read from device memory in a device-driven trace); kept
    long long pad_;                           // here so that it costs no registers in the walk loop
};

// Per-warp shared memory of the shadow-grid kernel (nearest_hit_grid_kernel): the per-slot best (t, surface) of a
// chunk and the candidate queues; same member names as WalkWarpShared where the float64 stage reads them.
#ifndef TRC_GRID_AHEAD
#define TRC_GRID_AHEAD 1
#endif
constexpr int kGridAhead = TRC_GRID_AHEAD;    // groups of 32 rays a warp has in flight ahead of the one it keys
constexpr int kGridRing = kGridAhead + 1;
struct GridWarpShared {
    double ring[kGridRing][6][32];            // ray groups arriving by cp.async: rows (vn, va, vb, dn, da, db) x lane
    unsigned long long best_t[kChunk];
    int best_s[kChunk];
    unsigned queue[kGridQueue];               // packed (slot, surface); lends its storage to the hit-slot list
    unsigned queue2[64];
    int q2n;
    const double *vert, *dir, *energy;
    long long ld;
};
__device__ __forceinline__ unsigned char *slot_scratch(WalkWarpShared &sh) {
    static_assert(sizeof(float) * kListFields * kListCap >= kChunk, "slot list borrows the active list");
    return reinterpret_cast<unsigned char *>(&sh.list[0][0]);
}
__device__ __forceinline__ unsigned char *slot_scratch(GridWarpShared &sh) {
    static_assert(sizeof(unsigned) * kGridQueue >= kChunk, "slot list borrows the candidate queue");
    return reinterpret_cast<unsigned char *>(&sh.queue[0]);
}

typedef TrcShadowGrid GridParams;             // shadow grid of a directionally coherent bundle (common.cuh)

struct WalkParams {
    GridParams g;                             // shadow-grid kernel only
    uint32_t *dirty_list;                     // shadow-grid kernel: chunks it leaves to the walk (some ray outside the
    unsigned long long *dirty_count;          //   grid's slope box) ...
    const uint32_t *chunk_list;               // ... and the walk kernel launched after it: the chunk ids to process
    const unsigned long long *chunk_list_n;   //   (null: every chunk of the bundle)
    const TrcSurfaceRec *surfs;
    const double *pool;                       // scene table pool (apertures of the pool-based flat kinds)
    const float4 *tri32;                      // float32 triangles for the pre-test, or null (see tri_prefilter)
    const double *tri64;                      // compact float64 triangle table (geom.cuh), or null
    const uint4 *pair_a, *pair_b;
    const uint4 *pair_ab;                     // interleaved copy for the walk outside the shared-memory cache
    int n_pairs;
    unsigned root_meta;
    const int32_t *cand;                      // candidate list: leaf surfaces, then unbounded surfaces
    int unbounded_first, n_unbounded;
    int has_tree;
    int debug;                                // TRC_WALK_DEBUG experiment bits (timing only)
    int refill_at;                            // leave the walk loop to refill when <= this many lanes are active
    int n_cached;                             // pairs cached in shared memory
    float root_lo[3], root_hi[3], qstep[3], extent[3];
    int64_t n;
    const TrcBounceCtl *ctl;                  // non-null: bundle and ray count are read from device memory (device-driven trace)
    const double *vert, *dir;
    int64_t ld;
    HitSink sink;
    unsigned long long *chunk_counter;
};

#define TRC_INF_BITS 0x7ff0000000000000ULL
#define TRC_FINF (__int_as_float(0x7f800000))

// Per-ray float32 walk state.  A box plane sits at root_lo + q * qstep (q a
// 16-bit integer); its ray parameter is
//     t = (root_lo + q qstep - v) / d = q A + B,   A = qstep / d,  B = (root_lo - v) / d.
// q enters the FMA as the float 2^23 + q built with one byte-permute
// (0x4B000000 | q), so B carries the -2^23 A term.  The box is widened per ray by
//     pad_k = 1e-6 (|v_k| + extent_k) + 1.5 qstep_k
// which covers the rounding of v and 1/d to float32 and of the FMA (a few 2^-24
// of |v_k| + |plane|) and the half-step error of folding 2^23 A into B.  `lo`
// planes use the origin moved by +pad, `hi` planes by -pad, which widens the
// slab for either sign of d.  1/d is clamped to +-1e20 so an axis-parallel ray
// gives a huge finite interval instead of 0 * inf.
struct RayF {
    float A[3], Bn[3], Bf[3];      // near / far plane: t = fma(2^23 + q, A, B)
    unsigned sn[3], sf[3];         // byte-permute selectors of the near / far plane of each axis
};

__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// selector words: 2^23 + (low / high 16 bits) as a float = prmt(w, 0x4B000000, sel)
#define TRC_SEL_LO 0x7610u
#define TRC_SEL_HI 0x7632u
__device__ __forceinline__ void set_selectors(RayF &r, unsigned signs) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        r.sn[k] = ((signs >> k) & 1u) ? TRC_SEL_HI : TRC_SEL_LO;    // d < 0: the ray enters through the hi plane
        r.sf[k] = r.sn[k] ^ 0x0022u;
    }
}

// Float32 set-up of one ray; returns whether the (widened) root box is hit.
// signs: bit k set when the direction component k is negative.
__device__ __forceinline__ bool ray_setup(V3 v, V3 d, const WalkParams &p, RayF &r, unsigned &signs) {
    const float dv[3] = {(float)d.x, (float)d.y, (float)d.z};
    const float ov[3] = {(float)v.x, (float)v.y, (float)v.z};
    float tmin = 0.0f, tmax = TRC_FINF;
    signs = 0u;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float inv = fabsf(dv[k]) < 1e-20f ? copysignf(1e20f, dv[k]) : rcp_approx(dv[k]);
        bool neg = inv < 0.0f;
        float pad = fmaf(1e-6f, fabsf(ov[k]) + p.extent[k], 1.5f * p.qstep[k]);
        float a = p.qstep[k] * inv;
        float shift = 8388608.0f * a;
        // lo planes see the origin moved by +pad, hi planes by -pad: the slab widens for either sign
        float bl = (p.root_lo[k] - (ov[k] + pad)) * inv, bh = (p.root_lo[k] - (ov[k] - pad)) * inv;
        r.A[k] = a;
        r.Bn[k] = (neg ? bh : bl) - shift;
        r.Bf[k] = (neg ? bl : bh) - shift;
        signs |= neg ? (1u << k) : 0u;
        // root box planes: lo = grid origin, hi = root_hi
        float t2 = (p.root_hi[k] - (ov[k] - pad)) * inv;
        tmin = fmaxf(tmin, neg ? t2 : bl); tmax = fminf(tmax, neg ? bl : t2);
    }
    return tmin <= tmax;
}

__device__ __forceinline__ float plane(unsigned w, unsigned sel) {
    unsigned o;          // raw prmt: __byte_perm would mask the (already clean) selector first
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(o) : "r"(w), "r"(0x4B000000u), "r"(sel));
    return __uint_as_float(o);
}
__device__ __forceinline__ uint4 lds128(unsigned addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// Conservative entry distance into the child box whose axis words are w0, w1, w2;
// hit = entered before tbest.
#ifndef TRC_WALK_FFMA2
#define TRC_WALK_FFMA2 0
#endif
// two float32 FMAs in one instruction (fma.rn.f32x2, sm_100+ -> SASS FFMA2): halves the FMA issue slots of a visit
__device__ __forceinline__ void fma2(float a0, float a1, float b0, float b1, float c0, float c1, float &d0, float &d1) {
    unsigned long long a, b, c, d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}
__device__ __forceinline__ float slab16(unsigned w0, unsigned w1, unsigned w2, const RayF &r, float tbest, bool &hit) {
#if TRC_WALK_FFMA2
    float n0, n1, n2, f0, f1, f2;
    fma2(plane(w0, r.sn[0]), plane(w1, r.sn[1]), r.A[0], r.A[1], r.Bn[0], r.Bn[1], n0, n1);
    fma2(plane(w0, r.sf[0]), plane(w1, r.sf[1]), r.A[0], r.A[1], r.Bf[0], r.Bf[1], f0, f1);
    fma2(plane(w2, r.sn[2]), plane(w2, r.sf[2]), r.A[2], r.A[2], r.Bn[2], r.Bf[2], n2, f2);
#else
    float n0 = fmaf(plane(w0, r.sn[0]), r.A[0], r.Bn[0]);
    float n1 = fmaf(plane(w1, r.sn[1]), r.A[1], r.Bn[1]);
    float n2 = fmaf(plane(w2, r.sn[2]), r.A[2], r.Bn[2]);
    float f0 = fmaf(plane(w0, r.sf[0]), r.A[0], r.Bf[0]);
    float f1 = fmaf(plane(w1, r.sf[1]), r.A[1], r.Bf[1]);
    float f2 = fmaf(plane(w2, r.sf[2]), r.A[2], r.Bf[2]);
#endif
    float tmin = fmaxf(fmaxf(n0, n1), fmaxf(n2, 0.0f));
    float tmax = fminf(fminf(f0, f1), fminf(f2, tbest));
    hit = tmin <= tmax;
    return tmin;
}

__device__ __forceinline__ float best_as_float(unsigned long long bits) {
    return __double2float_ru(__longlong_as_double((long long)bits));
}

// Stage C: test queued (slot, leaf) candidates in float64, 32 at a time, and
// fold them into the per-slot best (t, surface).  `all` drains the queue.
template <int kFlatOnly, class Sh>     // 0 generic scene, 1 flat-family surfaces only, 2 the same with the compact triangle table
__device__ __forceinline__ int flush_queue_body(Sh &sh, const unsigned *queue, const WalkParams &p,
                                                int64_t chunk_base, int lane, int qn, bool all) {
    __syncwarp();
    if (p.debug & 1) qn = 0;
    while (qn >= 32 || (all && qn > 0)) {
        int take = min(32, qn);
        int base = qn - take;
        bool has = lane < take;
        unsigned e = has ? queue[base + lane] : 0u;
        const int slot = (int)(e >> 24);
        const int first = (int)(e & kCandMask);
        const int cnt = has ? (int)((e >> 22) & 3u) + 1 : 0;
        // nearest of the leaf's surfaces for this ray (ties: lowest surface index)
        double t = TRC_INF;
        int surf = INT_MAX;
        if (has) {
            int64_t ray = chunk_base + slot;
            const double *vp = sh.vert + ray, *dp = sh.dir + ray;
            const long long ld = sh.ld;
            V3 v = mk(vp[0], vp[ld], vp[2 * ld]);
            V3 d = mk(dp[0], dp[ld], dp[2 * ld]);
            for (int k = 0; k < cnt; ++k) {
                int s = cnt == 1 ? first : __ldg(&p.cand[first + k]);     // single-surface entries hold the index
                const TrcSurfaceRec *sr = &p.surfs[s];
                double tk;
                if (kFlatOnly) {              // scene of flat-family surfaces only: no quadric code, few registers
                    const double2 *tr = reinterpret_cast<const double2 *>(p.tri64 + (size_t)TRC_TRI_REC_DOUBLES * s);
                    double2 h7;
                    if (kFlatOnly == 2 && (h7 = __ldg(tr + 7)).y != 0.0) {
                        // a triangle: one 128-byte line instead of the 384-byte surface record
                        const double2 h0 = __ldg(tr), h1 = __ldg(tr + 1), h2 = __ldg(tr + 2), h3 = __ldg(tr + 3);
                        const double2 h4 = __ldg(tr + 4), h5 = __ldg(tr + 5), h6 = __ldg(tr + 6);
                        const double4 q0 = make_double4(h0.x, h0.y, h1.x, h1.y), q1 = make_double4(h2.x, h2.y, h3.x, h3.y);
                        const double4 q2 = make_double4(h4.x, h4.y, h5.x, h5.y), q3 = make_double4(h6.x, h6.y, h7.x, h7.y);
                        TriTerms tt;
                        tt.e0 = mk(q1.z, q1.w, q2.x); tt.e1 = mk(q2.y, q2.z, q2.w);
                        tt.uv = q3.x; tt.n0 = q3.y; tt.n1 = q3.z; tt.den = q3.w;
                        tk = isect_triangle(mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), tt, v, d);
                    } else {
                        tk = isect_flat(sr, sr->gkind, v, d);
                    }
                    if (tk == 0.0) tk = TRC_INF;
                } else {
                    tk = intersect(sr, v, d, p.pool);
                }
                if (tk < t || (tk == t && s < surf && tk != TRC_INF)) { t = tk; surf = s; }
            }
        }
        unsigned long long tb = (unsigned long long)__double_as_longlong(t);
        bool cand = has && tb < TRC_INF_BITS;
        unsigned long long before = cand ? sh.best_t[slot] : 0ULL;
        // a slot's first hit: its energy will be wanted by the record emission, and only hits read that row
        if (cand && before == TRC_INF_BITS && p.sink.records)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(sh.energy + chunk_base + slot));
        __syncwarp();
        if (cand) atomicMin(&sh.best_t[slot], tb);
        __syncwarp();
        bool win = cand && sh.best_t[slot] == tb;
        if (win && tb < before) sh.best_s[slot] = INT_MAX;     // the old index belonged to a larger t
        __syncwarp();
        if (win) atomicMin(&sh.best_s[slot], surf);            // ties: lowest surface index
        __syncwarp();
        qn = base;
    }
    return qn;
}

// Generic scenes: the float64 tests of all 24 geometry kinds need ~75 registers; kept out of
// line so that they are not live together with the walk state.  Flat-only scenes (heliostat
// fields, triangle meshes) inline the short flat-family test instead and run at higher occupancy.
__device__ __noinline__ int flush_queue_generic(WalkWarpShared &sh, const WalkParams &p, int64_t chunk_base, int lane,
                                                int qn, bool all) {
    return flush_queue_body<false>(sh, sh.queue, p, chunk_base, lane, qn, all);
}
__device__ __noinline__ int flush_queue_generic(GridWarpShared &sh, const WalkParams &p, int64_t chunk_base, int lane,
                                                int qn, bool all) {
    return flush_queue_body<false>(sh, sh.queue, p, chunk_base, lane, qn, all);
}

// Conservative float32 ray/triangle test (Moeller-Trumbore with running error bounds): false only
// when the ray certainly misses the triangle, so that the float64 test of the reference
// (triangular_face.py:59-72) would give inf as well.  With eps = 2^-24, A = |o|_1 + |p0|_1 and
// |x|_1 the 1-norm: the numerators of u and v carry an absolute error below
// K eps A |e1|_1 |d|_1 and K eps A |e0|_1 |d|_1, the determinant below K eps |e0|_1 |e1|_1 |d|_1
// (rounding of origin, vertices and direction to float32, of the differences and of the
// products; K = 16 leaves a factor > 2 over the first-order sums).  A determinant inside its
// own error band decides nothing.  The plane distance t is not used: the float64 stage rules on it.
__device__ __forceinline__ bool tri_prefilter(const float4 *tri, V3 v, V3 d) {
    const float4 a = __ldg(tri), b = __ldg(tri + 1), c = __ldg(tri + 2);
    if (c.y == 0.f) return true;                                 // neither a triangle nor a plate: not pre-tested
    const float ox = (float)v.x, oy = (float)v.y, oz = (float)v.z;
    const float dx = (float)d.x, dy = (float)d.y, dz = (float)d.z;
    const float e0x = a.w, e0y = b.x, e0z = b.y, e1x = b.z, e1y = b.w, e1z = c.x;
    const float tx = ox - a.x, ty = oy - a.y, tz = oz - a.z;
    const float px = dy * e1z - dz * e1y, py = dz * e1x - dx * e1z, pz = dx * e1y - dy * e1x;
    const float det = e0x * px + e0y * py + e0z * pz;
    const float qx = ty * e0z - tz * e0y, qy = tz * e0x - tx * e0z, qz = tx * e0y - ty * e0x;
    const float sgn = det < 0.f ? -1.f : 1.f;
    const float nu = sgn * (tx * px + ty * py + tz * pz);
    const float nv = sgn * (dx * qx + dy * qy + dz * qz);
    const float ad = fabsf(det);
    const float keps = 16.f * 5.9604645e-8f;
    const float A = fabsf(ox) + fabsf(oy) + fabsf(oz) + fabsf(a.x) + fabsf(a.y) + fabsf(a.z);
    const float n0 = fabsf(e0x) + fabsf(e0y) + fabsf(e0z), n1 = fabsf(e1x) + fabsf(e1y) + fabsf(e1z);
    const float nd = fabsf(dx) + fabsf(dy) + fabsf(dz);
    const float err_u = keps * A * n1 * nd, err_v = keps * A * n0 * nd, err_d = keps * n0 * n1 * nd;
    if (!(ad > 2.f * err_d)) return true;                        // grazing / degenerate / NaN: undecided
    if (c.y == 2.f) {                                            // centred parallelogram: |u| <= 1 and |v| <= 1
        if (fabsf(nu) - ad > err_u + err_d) return false;
        if (fabsf(nv) - ad > err_v + err_d) return false;
        return true;
    }
    if (nu < -err_u || nv < -err_v) return false;
    if (nu + nv - ad > err_u + err_v + err_d) return false;
    return true;
}

// Flat-only scenes with triangles: queue entries first take the float32 pre-test, 32 at a time;
// the few that pass wait in queue2 for the float64 stage.
template <int kFlatOnly, class Sh>
__device__ __forceinline__ int prefilter_queue(Sh &sh, const WalkParams &p, int64_t chunk_base, int lane,
                                               int qn, bool all) {
    const unsigned lt_mask = (1u << lane) - 1u;
    __syncwarp();
    int q2n = sh.q2n;
    while (qn >= 32 || (all && qn > 0)) {
        int take = min(32, qn);
        int base = qn - take;
        bool has = lane < take;
        unsigned e = has ? sh.queue[base + lane] : 0u;
        bool pass = has;
        if (has && ((e >> 22) & 3u) == 0u) {                     // single-surface entry holds the surface index
            int64_t ray = chunk_base + (int)(e >> 24);
            const double *vp = sh.vert + ray, *dp = sh.dir + ray;
            const long long ld = sh.ld;
            V3 v = mk(vp[0], vp[ld], vp[2 * ld]);
            V3 d = mk(dp[0], dp[ld], dp[2 * ld]);
            pass = tri_prefilter(p.tri32 + 3 * (size_t)(e & kCandMask), v, d);
        }
        if (q2n >= 32)                                           // leaves < 32: room for 32 more
            q2n = flush_queue_body<kFlatOnly>(sh, sh.queue2, p, chunk_base, lane, q2n, false);
        unsigned m = __ballot_sync(0xffffffffu, pass);
        if (pass) sh.queue2[q2n + __popc(m & lt_mask)] = e;
        q2n += __popc(m);
        __syncwarp();
        qn = base;
    }
    if (q2n >= 32 || (all && q2n > 0)) q2n = flush_queue_body<kFlatOnly>(sh, sh.queue2, p, chunk_base, lane, q2n, all);
    if (lane == 0) sh.q2n = q2n;
    __syncwarp();
    return qn;
}

template <int kFlatOnly, class Sh>     // 0 generic scene, 1 flat-family surfaces only, 2 the same with the compact triangle table
__device__ __forceinline__ int flush_queue(Sh &sh, const WalkParams &p, int64_t chunk_base, int lane,
                                           int qn, bool all) {
#ifdef TRC_WALK_EXPERIMENT_NO_F64      // timing experiment only: what the walk costs without the float64 stage
    if (kFlatOnly) return (qn >= 32 || all) ? 0 : qn;
#endif
    if (kFlatOnly) {
        if (p.tri32) return prefilter_queue<kFlatOnly>(sh, p, chunk_base, lane, qn, all);
        return flush_queue_body<kFlatOnly>(sh, sh.queue, p, chunk_base, lane, qn, all);
    }
    return flush_queue_generic(sh, p, chunk_base, lane, qn, all);
}

// Append one chunk's hits (ascending ray index) to the hit list of trc_trace_hits.
// The hit slots are first compacted into shared memory (the active list is empty at this
// point and lends its storage), so the dependent global loads of the records run once for
// 32 hits at a time instead of once per 32 slots.
template <class Sh>
__device__ __forceinline__ void emit_chunk_hits(Sh &sh, const WalkParams &p, int64_t c_cur,
                                             int64_t chunk_base, int lane) {
    const unsigned full = 0xffffffffu;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned char *slots = slot_scratch(sh);
    int h_c = 0;
#pragma unroll
    for (int k = 0; k < kChunk / 32; ++k) {
        int i = lane + 32 * k;
        bool hit = sh.best_t[i] != TRC_INF_BITS;
        unsigned m = __ballot_sync(full, hit);
        if (hit) slots[h_c + __popc(m & lt_mask)] = (unsigned char)i;
        h_c += __popc(m);
    }
    __syncwarp();
    // the first 32 hits' rays are requested BEFORE the reservation of record space: the atomic's round trip and the
    // loads (the energy comes from DRAM: it is read for hits only) then overlap instead of adding up, once per chunk
    const long long ld = sh.ld;
    V3 v = mk(0.0, 0.0, 0.0), d = v;
    double en = 0.0;
    int i = 0;
    if (lane < h_c) {
        i = slots[lane];
        const int64_t ray = chunk_base + i;
        const double *vp = sh.vert + ray, *dp = sh.dir + ray;
        v = mk(vp[0], vp[ld], vp[2 * ld]);
        d = mk(dp[0], dp[ld], dp[2 * ld]);
        en = sh.energy[ray];
    }
    unsigned long long base = 0;
    if (lane == 0) {
        if (h_c) base = atomicAdd(p.sink.total, (unsigned long long)h_c);    // one address: not for nothing
        p.sink.chunk_off[c_cur] = (uint32_t)base;
        p.sink.chunk_cnt[c_cur] = (uint32_t)h_c;
    }
    base = __shfl_sync(full, base, 0);
    for (int e = lane; e < h_c; e += 32) {
        unsigned long long pos = base + e;
        if ((int64_t)pos >= p.sink.capacity) break;       // overflow: the host re-runs with the exact size
        if (e >= 32) {
            i = slots[e];
            const int64_t ray = chunk_base + i;
            const double *vp = sh.vert + ray, *dp = sh.dir + ray;
            v = mk(vp[0], vp[ld], vp[2 * ld]);
            d = mk(dp[0], dp[ld], dp[2 * ld]);
            en = sh.energy[ray];
        }
        const int surf = sh.best_s[i];
        store_record(p.sink.records + pos, v, d, en, p.sink.ray_base + (int)(chunk_base + i), surf);
        if (p.sink.keys) p.sink.keys[pos] = surf;
    }
    __syncwarp();
}

// ---- 1-D bulk copies (TMA engine, sm_90+) into the per-warp ray ring, completion through an mbarrier
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gmem_src) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// kAllShared: the whole tree fits the shared-memory cache (no global-memory path in the walk)
template <bool kAllShared, int kFlatOnly>
__global__ void __launch_bounds__(kWalkWarps * 32, kFlatOnly ? TRC_WALK_MIN_CTAS_FLAT : TRC_WALK_MIN_CTAS)
nearest_hit_walk_kernel(WalkParams p) {
    extern __shared__ __align__(16) unsigned char walk_smem[];
    WalkShared &S = *reinterpret_cast<WalkShared *>(walk_smem);
    WalkWarpShared &sh = S.w[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    const unsigned lt_mask = (1u << lane) - 1u;

    // top of the tree -> shared memory (breadth-first order: the first kSmemPairs pairs).  The tree is
    // constant data, so this prologue runs before the grid dependency resolves (PDL).
    {
        uint4 *cache = reinterpret_cast<uint4 *>(walk_smem + sizeof(WalkShared));
        int np = min(p.n_pairs, p.n_cached);
        for (int k = threadIdx.x; k < np; k += blockDim.x) {
            cache[k] = __ldg(p.pair_a + k);
            cache[p.n_cached + k] = __ldg(p.pair_b + k);
        }
    }
    pdl_wait();
    pdl_trigger();
    if (threadIdx.x == 0) S.n_rays = p.ctl ? p.ctl->n_in : (long long)p.n;
    if (lane == 0) {
        if (p.ctl) { sh.vert = p.ctl->in.V; sh.dir = p.ctl->in.D; sh.energy = p.ctl->in.E; sh.ld = p.ctl->in.ld; }
        else { sh.vert = p.vert; sh.dir = p.dir; sh.energy = p.sink.energy; sh.ld = p.ld; }
    }
    __syncthreads();
    unsigned pair_a_smem = (unsigned)__cvta_generic_to_shared(walk_smem + sizeof(WalkShared));
    unsigned pair_b_smem = pair_a_smem + 16u * (unsigned)p.n_cached;             // pair_b follows pair_a
    asm volatile("" : "+r"(pair_a_smem), "+r"(pair_b_smem));   // keep them in registers (else recomputed per visit)
    const int root_count = p.has_tree ? (int)TRC_META_COUNT(p.root_meta) : 7;   // 0 inner, 1..4 leaf, 7 nothing
    // entry payload of a tree that is a single leaf: list position, or the surface index itself for one surface
    const unsigned root_first = root_count == 1 ? (unsigned)__ldg(&p.cand[TRC_META_INDEX(p.root_meta)])
                                                : (unsigned)TRC_META_INDEX(p.root_meta);
    int stack[kStack];

    // ---- ray prefetcher: the warp streams batches of kBatch rays (chunk after chunk, chunk ids from the
    // global counter) into a two-slot shared-memory ring, one batch ahead of stage A: six 1-D bulk copies
    // (cp.async.bulk, one per coordinate row) issued by ONE lane and tracked by the slot's mbarrier, so a ray
    // is read from HBM exactly once, stage A does not wait for DRAM, and no lane computes addresses.
    int64_t pc;            // chunk being prefetched           (warp-uniform)
    int pc_n = 0;          // rays in that chunk (0: past the end of the bundle)
    int poff = 0;          // next ray offset in that chunk
    int fifo_w = 0, fifo_r = 0, ring_w = 0, ring_r = 0;
#if TRC_WALK_TMA
    if (lane == 0) {
        for (int k = 0; k < kRing; ++k) mbar_init(smem_u32(&sh.mbar[k]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    // bulk copies need 16-byte aligned rows: otherwise (odd row stride, sliced bundle) lanes copy by hand
    const bool bulk_ok = ((((unsigned long long)sh.vert | (unsigned long long)sh.dir) & 15ULL) == 0ULL) && ((sh.ld & 1LL) == 0LL);
#endif
    auto fetch_chunk = [&]() -> int64_t {
        unsigned long long c = 0;
        if (lane == 0) {
            c = atomicAdd(p.chunk_counter, 1ULL);
            // after the shadow-grid kernel: only the chunks it left over, by id (past the list: past the bundle)
            if (p.chunk_list) c = c < *p.chunk_list_n ? (unsigned long long)p.chunk_list[c] : (1ULL << 40);
        }
        c = __shfl_sync(full, c, 0);
        if (lane == 0) sh.fifo[fifo_w & 3] = (long long)c;
        ++fifo_w;
        __syncwarp();
        long long rem = S.n_rays - (long long)c * kChunk;
        pc_n = rem <= 0 ? 0 : (rem < kChunk ? (int)rem : kChunk);
        return (int64_t)c;
    };
#if !TRC_WALK_TMA
    auto issue_batch = [&]() {
        if (pc_n > 0) {
            int64_t ray = pc * kChunk + poff + lane;
            if (poff + lane < pc_n) {
                double *dst = &sh.ring[ring_w & (kRing - 1)][0][lane];
                const double *vp = sh.vert + ray, *dp = sh.dir + ray;
                const long long ld = sh.ld;
                cp_async8(dst, vp);
                cp_async8(dst + kBatch, vp + ld);
                cp_async8(dst + 2 * kBatch, vp + 2 * ld);
                cp_async8(dst + 3 * kBatch, dp);
                cp_async8(dst + 4 * kBatch, dp + ld);
                cp_async8(dst + 5 * kBatch, dp + 2 * ld);
            }
            poff += 32;
            if (poff >= pc_n) { pc = fetch_chunk(); poff = 0; }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");      // possibly empty: keeps the group count uniform
        ++ring_w;
    };
    pc = fetch_chunk();
    issue_batch();
    issue_batch();
#else
    auto issue_batch = [&]() {
        if (pc_n <= 0) return;                    // past the end of the bundle: nobody will wait for this batch
        const int cnt = min(kBatch, pc_n - poff);
        const int slot = ring_w & (kRing - 1);
        const int64_t ray0 = pc * kChunk + poff;
        if (bulk_ok && (cnt & 1) == 0) {
            if (lane == 0) {
                const unsigned bar = smem_u32(&sh.mbar[slot]);
                const unsigned dst = smem_u32(&sh.ring[slot][0][0]);
                const unsigned bytes = 8u * (unsigned)cnt;
                const double *vp = sh.vert + ray0, *dp = sh.dir + ray0;
                const long long ld = sh.ld;
                mbar_arrive_expect_tx(bar, 6u * bytes);
                bulk_g2s(dst, vp, bytes, bar);
                bulk_g2s(dst + 8u * kBatch, vp + ld, bytes, bar);
                bulk_g2s(dst + 16u * kBatch, vp + 2 * ld, bytes, bar);
                bulk_g2s(dst + 24u * kBatch, dp, bytes, bar);
                bulk_g2s(dst + 32u * kBatch, dp + ld, bytes, bar);
                bulk_g2s(dst + 40u * kBatch, dp + 2 * ld, bytes, bar);
            }
        } else {
            for (int j = lane; j < cnt; j += 32) {
                const double *vp = sh.vert + ray0 + j, *dp = sh.dir + ray0 + j;
                const long long ld = sh.ld;
                double *dst = &sh.ring[slot][0][j];
                dst[0] = vp[0]; dst[kBatch] = vp[ld]; dst[2 * kBatch] = vp[2 * ld];
                dst[3 * kBatch] = dp[0]; dst[4 * kBatch] = dp[ld]; dst[5 * kBatch] = dp[2 * ld];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sh.mbar[slot]));
        }
        poff += cnt;
        if (poff >= pc_n) { pc = fetch_chunk(); poff = 0; }
        ++ring_w;
    };
    pc = fetch_chunk();
#pragma unroll
    for (int k = 0; k < kAhead; ++k) issue_batch();
#endif

    for (;;) {
        const int64_t c_cur = (int64_t)sh.fifo[fifo_r & 3];
        ++fifo_r;
        const int64_t chunk_base = c_cur * kChunk;
        if (chunk_base >= S.n_rays) break;
        const int chunk_n = (int)min((int64_t)kChunk, (int64_t)S.n_rays - chunk_base);

#ifndef TRC_WALK_EXPERIMENT_NO_F64
        for (int i = lane; i < kChunk; i += 32) { sh.best_t[i] = TRC_INF_BITS; sh.best_s[i] = INT_MAX; }
#endif
        if (lane == 0) sh.q2n = 0;
        __syncwarp();

        int next = 0;          // rays of the chunk already through stage A   (warp-uniform)
        int list_n = 0;        // active-list fill                              (warp-uniform)
        int qn = 0;            // candidate-queue fill                          (warp-uniform)
        RayF r;
        int cur = -1, sp = 0;  // cur = pair index this lane visits next, -1 = the lane is idle
        unsigned pend0 = kNoEntry, pend1 = kNoEntry;   // leaves reached but not yet queued
        unsigned tag = 0;      // slot << 24
        float tbest = TRC_FINF;

        for (;;) {
            // ---------------- stage A: produce survivors until the list can feed a full refill
#pragma unroll 1
            while (list_n < 32 && next < chunk_n) {
                int j = next + lane;
                bool valid = j < chunk_n;
#if TRC_WALK_TMA
                if ((next & (kBatch - 1)) == 0) {
                    // a new batch: start the copy of the one after it, then wait for this one to have landed
                    __syncwarp();
                    issue_batch();
                    mbar_wait(smem_u32(&sh.mbar[ring_r & (kRing - 1)]), (unsigned)(ring_r / kRing) & 1u);
                    ++ring_r;
                }
#else
                issue_batch();
                asm volatile("cp.async.wait_group 2;" ::: "memory");
                ++ring_r;
#endif
                RayF nr;
                unsigned nsigns = 0u;
                bool surv = false;
                if (valid) {
                    const double *src = &sh.ring[(ring_r - 1) & (kRing - 1)][0][(next & (kBatch - 1)) + lane];
                    V3 v = mk(src[0], src[kBatch], src[2 * kBatch]);
                    V3 d = mk(src[3 * kBatch], src[4 * kBatch], src[5 * kBatch]);
                    surv = ray_setup(v, d, p, nr, nsigns) && root_count != 7 && !(p.debug & 2);
                }
                next += 32;
                unsigned m = __ballot_sync(full, surv);
                if (root_count > 0) {
                    // the whole tree is one leaf: its surfaces are candidates of every surviving ray
                    if (root_count != 7) {
                        if (surv) sh.queue[qn + __popc(m & lt_mask)] =
                            ((unsigned)j << 24) | ((unsigned)(root_count - 1) << 22) | root_first;
                        qn += __popc(m);
                        if (qn >= 32) qn = flush_queue<kFlatOnly>(sh, p, chunk_base, lane, qn, false);
                    }
                } else {
                    if (surv) {
                        int e = list_n + __popc(m & lt_mask);
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            sh.list[k][e] = nr.A[k]; sh.list[3 + k][e] = nr.Bn[k]; sh.list[6 + k][e] = nr.Bf[k];
                        }
                        sh.list[9][e] = __int_as_float(j | (int)(nsigns << 8));
                    }
                    list_n += __popc(m);
                }
                // surfaces without a finite box are candidates of every ray (accel_tree.py:59-73)
                if (p.n_unbounded) {
                    unsigned mv = __ballot_sync(full, valid);
                    for (int k = 0; k < p.n_unbounded; ++k) {
                        const unsigned s_unb = (unsigned)__ldg(&p.cand[p.unbounded_first + k]);   // entry = surface index
                        if (valid) sh.queue[qn + __popc(mv & lt_mask)] = ((unsigned)j << 24) | s_unb;
                        qn += __popc(mv);
                        if (qn >= 32) qn = flush_queue<kFlatOnly>(sh, p, chunk_base, lane, qn, false);
                    }
                }
                __syncwarp();
            }
            // ---------------- refill idle lanes from the list
            {
                unsigned idle = __ballot_sync(full, cur < 0);
                int rk = __popc(idle & lt_mask);
                if (cur < 0 && rk < list_n) {
                    int e = list_n - 1 - rk;
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        r.A[k] = sh.list[k][e]; r.Bn[k] = sh.list[3 + k][e]; r.Bf[k] = sh.list[6 + k][e];
                    }
                    int word = __float_as_int(sh.list[9][e]);
                    int slot = word & 0xff;
                    set_selectors(r, (unsigned)word >> 8);
                    tag = (unsigned)slot << 24;
                    cur = 0; sp = 0;
#ifdef TRC_WALK_EXPERIMENT_NO_F64
                    tbest = TRC_FINF;
#else
                    tbest = best_as_float(sh.best_t[slot]);
#endif
                }
                list_n -= min(__popc(idle), list_n);
                __syncwarp();
            }
            if (__ballot_sync(full, cur >= 0) == 0u) break;   // list empty and chunk exhausted

            // ---------------- stage B: walk.  kWalkSteps visits per lane without any warp-level
            // coordination (a lane that reaches a leaf parks the leaf in `pend` and waits), then
            // the warp collects the parked leaves into the candidate queue, flushes it when 32 are
            // waiting, and leaves the loop when enough lanes are idle to be worth a refill.
#pragma unroll 1
            for (;;) {
#pragma unroll 1
                for (int it = 0; it < kWalkSteps; ++it) {
                    if (cur >= 0 && pend0 == kNoEntry) {
                        uint4 a, b;
                        if (kAllShared || cur < p.n_cached) {
                            a = lds128(pair_a_smem + 16u * (unsigned)cur);
                            b = lds128(pair_b_smem + 16u * (unsigned)cur);
                        }
                        else { a = __ldg(p.pair_ab + 2 * cur); b = __ldg(p.pair_ab + 2 * cur + 1); }
                        bool h0, h1;
                        float t0 = slab16(a.x, a.y, a.z, r, tbest, h0);
                        float t1 = slab16(a.w, b.x, b.y, r, tbest, h1);
                        bool i0 = h0 && (int)b.z >= 0, i1 = h1 && (int)b.w >= 0;     // hit inner children
                        // hit leaves: child word -> queue entry (slot << 24 | (count - 1) << 22 | first)
                        if (h1 && !i1) pend0 = tag | (b.w & TRC_CHILD_ENTRY_MASK);
                        if (h0 && !i0) { pend1 = pend0; pend0 = tag | (b.z & TRC_CHILD_ENTRY_MASK); }
                        int f0 = (int)b.z, f1 = (int)b.w;        // inner: meta == pair index
                        if (i0 && i1) {
                            bool swap = t1 < t0;
                            cur = swap ? f1 : f0;
                            stack[sp++] = swap ? f0 : f1;
                        } else if (i0 || i1) {
                            cur = i0 ? f0 : f1;
                        } else if (sp > 0) {
                            cur = stack[--sp];
                        } else {
                            cur = -1;
                        }
                    }
                }
                // parked leaves -> candidate queue (positions from ballots, no atomics)
                unsigned m0 = __ballot_sync(full, pend0 != kNoEntry);
                if (m0) {
                    unsigned m1 = __ballot_sync(full, pend1 != kNoEntry);
                    if (pend0 != kNoEntry) sh.queue[qn + __popc(m0 & lt_mask)] = pend0;
                    qn += __popc(m0);
                    if (pend1 != kNoEntry) sh.queue[qn + __popc(m1 & lt_mask)] = pend1;
                    qn += __popc(m1);
                    pend0 = pend1 = kNoEntry;
                    if (qn >= 32) {
                        qn = flush_queue<kFlatOnly>(sh, p, chunk_base, lane, qn, false);
#ifndef TRC_WALK_EXPERIMENT_NO_F64
                        if (cur >= 0) tbest = best_as_float(sh.best_t[tag >> 24]);
#endif
                    }
                }
                int n_active = __popc(__ballot_sync(full, cur >= 0));
                if (n_active == 0) break;
                if (n_active <= p.refill_at && (list_n > 0 || next < chunk_n)) break;
            }
        }
        flush_queue<kFlatOnly>(sh, p, chunk_base, lane, qn, true);
        __syncwarp();
        if (p.sink.hit_surf) {
            for (int i = lane; i < chunk_n; i += 32)
#ifndef TRC_WALK_EXPERIMENT_NO_F64
                p.sink.hit_surf[chunk_base + i] = sh.best_t[i] == TRC_INF_BITS ? -1 : sh.best_s[i];
#else
                p.sink.hit_surf[chunk_base + i] = -1;
#endif
        }
#ifdef TRC_WALK_EXPERIMENT_NO_F64
        if (!kFlatOnly)
#endif
        if (p.sink.records) emit_chunk_hits(sh, p, c_cur, chunk_base, lane);
        __syncwarp();
    }
#if !TRC_WALK_TMA
    asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}


// ------------------------------------------------------------------ shadow grid (coherent bundles)
// Nearest hit of a bundle whose rays lie in the slope box of the scene's shadow grid (common.cuh TrcShadowGrid,
// built by shadow_grid.cu): no tree, no per-ray walk state.  A warp owns a chunk of 256 consecutive rays; a lane
// keys its ray by the point where it crosses the grid's reference plane, tests the cell's bit in shared memory
// (two thirds of the rays of a heliostat field end here), fetches the cell's candidates and queues them; the
// queue is tested 32 at a time in float64 by the same code as the tree walk's stage C, and the chunk's hits are
// emitted in ascending ray order.  A chunk with a ray outside the slope box (or too far out for the float32 key) is
// left untouched and listed for the tree-walk kernel, which is launched right after over that list -- so a wrong
// or partial hint costs time, never correctness.
#ifndef TRC_GRID_MIN_CTAS
#define TRC_GRID_MIN_CTAS 1
#endif
#ifndef TRC_GRID_WARPS
#define TRC_GRID_WARPS 16
#endif
constexpr int kGridWarps = TRC_GRID_WARPS;    // warps per CTA of the shadow-grid kernel
struct GridShared {
    GridWarpShared w[kGridWarps];
    long long n_rays;
    long long pad_;
};

static_assert(sizeof(GridShared) + 4 * kGridMaxWords <= (TRC_GRID_MIN_CTAS == 1 ? 232448 : 232448 / TRC_GRID_MIN_CTAS - 1024),
              "shadow-grid kernel: per-warp state + occupancy bits exceed the shared memory of its CTAs per SM");

// the float64 stage as ONE out-of-line call site: the key / lookup loop below stays small and keeps its registers
template <int kFlatOnly>
__device__ __noinline__ int grid_flush(GridWarpShared &sh, const WalkParams &p, int64_t chunk_base, int lane, int qn,
                                       bool all) {
    return flush_queue<kFlatOnly>(sh, p, chunk_base, lane, qn, all);
}

// record emission out of line as well: the pipelined loop then shares its registers with nothing
__device__ __noinline__ void grid_emit(GridWarpShared &sh, const WalkParams &p, int64_t c_cur, int64_t chunk_base, int lane) {
    emit_chunk_hits(sh, p, c_cur, chunk_base, lane);
}

struct GridRay { double vn, va, vb, dn, da, db; };

template <int kFlatOnly>
__global__ void __launch_bounds__(kGridWarps * 32, TRC_GRID_MIN_CTAS)
nearest_hit_grid_kernel(const __grid_constant__ WalkParams p) {
    extern __shared__ __align__(16) unsigned char grid_smem[];
    GridShared &S = *reinterpret_cast<GridShared *>(grid_smem);
    GridWarpShared &sh = S.w[threadIdx.x >> 5];
    const uint32_t *bits = reinterpret_cast<const uint32_t *>(grid_smem + sizeof(GridShared));
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    const unsigned lt_mask = (1u << lane) - 1u;
    {
        uint32_t *dst = reinterpret_cast<uint32_t *>(grid_smem + sizeof(GridShared));
        for (int k = threadIdx.x; k < p.g.n_words; k += blockDim.x) dst[k] = __ldg(p.g.bits + k);   // constant data
    }
    pdl_wait();
    pdl_trigger();
    if (threadIdx.x == 0) S.n_rays = p.ctl ? p.ctl->n_in : (long long)p.n;
    if (lane == 0) {
        if (p.ctl) { sh.vert = p.ctl->in.V; sh.dir = p.ctl->in.D; sh.energy = p.ctl->in.E; sh.ld = p.ctl->in.ld; }
        else { sh.vert = p.vert; sh.dir = p.dir; sh.energy = p.sink.energy; sh.ld = p.ld; }
    }
    __syncthreads();
    const long long n_rays = S.n_rays;
    // the six rows in the grid's axis order (normal, a, b)
    const double *vn_p = sh.vert + p.g.axis_n * sh.ld, *va_p = sh.vert + p.g.axis_a * sh.ld, *vb_p = sh.vert + p.g.axis_b * sh.ld;
    const double *dn_p = sh.dir + p.g.axis_n * sh.ld, *da_p = sh.dir + p.g.axis_a * sh.ld, *db_p = sh.dir + p.g.axis_b * sh.ld;
    // Chunks are dealt round robin (warp w of the grid takes chunks w, w + W, ...), so a warp's rays form one known
    // sequence of groups of 32; the groups arrive in a shared-memory ring by cp.async, kGridAhead of them in flight
    // per warp ahead of the one being keyed -- across the float64 stage, the record emission and chunk boundaries.
    // (Registers cannot hold that much: one group in flight per warp streams at half the HBM rate at 16 warps per SM.)
    const unsigned long long c_stride = (unsigned long long)gridDim.x * kGridWarps;
    unsigned long long c_next = (unsigned long long)blockIdx.x * kGridWarps + (threadIdx.x >> 5);
    unsigned long long c_iss = c_next;        // chunk, group and ring slot of the next group to request
    int g_iss = 0, slot_iss = 0, slot_con = 0;
    auto issue = [&]() {
        const long long ray = (long long)c_iss * kChunk + 32 * g_iss + lane;
        if (ray < n_rays) {
            double (*dst)[32] = sh.ring[slot_iss];
            cp_async8(&dst[0][lane], vn_p + ray); cp_async8(&dst[1][lane], va_p + ray); cp_async8(&dst[2][lane], vb_p + ray);
            cp_async8(&dst[3][lane], dn_p + ray); cp_async8(&dst[4][lane], da_p + ray); cp_async8(&dst[5][lane], db_p + ray);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (++g_iss == kChunk / 32) { g_iss = 0; c_iss += c_stride; }
        slot_iss = slot_iss + 1 == kGridRing ? 0 : slot_iss + 1;
    };
#pragma unroll
    for (int k = 0; k < kGridAhead; ++k) issue();
    for (;;) {
        const unsigned long long c = c_next;
        const int64_t chunk_base = (int64_t)c * kChunk;
        if (chunk_base >= n_rays) break;
        const int chunk_n = (int)min((int64_t)kChunk, (int64_t)n_rays - chunk_base);
        c_next = c + c_stride;
        for (int i = lane; i < kChunk; i += 32) { sh.best_t[i] = TRC_INF_BITS; sh.best_s[i] = INT_MAX; }
        if (lane == 0) sh.q2n = 0;
        __syncwarp();
        int qn = 0, done_g = 0;           // done_g: groups of this chunk taken from the ring
        bool dirty = false;
        unsigned ent_prev = 0u;            // cell entry fetched for the previous group's ray of this lane
        int next = 0;
        // software pipeline over the chunk's groups of 32 rays: [load rays of g+1] [key + cell fetch of g] [queue g-1].
        // The float64 stage is called from OUTSIDE the pipelined loop (which leaves when the queue cannot take another
        // group), so that the loop carries no call and keeps its state in registers.
        for (;;) {
            bool need_flush = false;
#pragma unroll 1
            for (; next < chunk_n + 32; next += 32) {
                // room for the previous group's candidates (and the unbounded surfaces of this one)?
                const int cnt = (int)(ent_prev & 255u);
                int incl = cnt;
                const unsigned mq = __ballot_sync(full, cnt != 0);
                const bool simple = __all_sync(full, cnt <= 1);
                if (mq && !simple) {
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int x = __shfl_up_sync(full, incl, o); if (lane >= o) incl += x; }
                }
                const int tot = !mq ? 0 : (simple ? __popc(mq) : __shfl_sync(full, incl, 31));
                const int want = tot + 32 * p.n_unbounded;
                if (want > kGridQueue - 64) { dirty = true; break; }   // absurdly crowded cells: leave it to the walk
                if (qn + want > kGridQueue - 32) { need_flush = true; break; }
                const int j = next + lane;
                const bool valid = j < chunk_n;
                GridRay cur = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
                if (next < chunk_n) {
                    issue();
                    asm volatile("cp.async.wait_group %0;" ::"n"(kGridAhead) : "memory");
                    if (valid) {
                        const double (*src)[32] = sh.ring[slot_con];
                        cur.vn = src[0][lane]; cur.va = src[1][lane]; cur.vb = src[2][lane];
                        cur.dn = src[3][lane]; cur.da = src[4][lane]; cur.db = src[5][lane];
                    }
                    slot_con = slot_con + 1 == kGridRing ? 0 : slot_con + 1;
                    ++done_g;
                }
                unsigned ent = 0u;
                bool out = false;
                if (p.debug & 4) {      // experiment (TRC_WALK_DEBUG=4): stream the rays, do nothing with them
                    if (valid && cur.vn + cur.va + cur.vb + cur.dn + cur.da + cur.db == 1.2345e300) ent = 1u;
                } else
                if (valid) {
                    const float vn = (float)cur.vn, va = (float)cur.va, vb = (float)cur.vb;
                    const float dn = (float)cur.dn, da = (float)cur.da, db = (float)cur.db;
                    const float inv = rcp_approx(dn);
                    const bool in_box = dn * p.g.sgn > 0.0f && fabsf(fmaf(da, inv, -p.g.t0a)) <= p.g.da &&
                                        fabsf(fmaf(db, inv, -p.g.t0b)) <= p.g.db &&
                                        fmaxf(fmaxf(fabsf(va), fabsf(vb)), fabsf(vn)) <= p.g.rmax;
                    out = !in_box;
                    const float t = (p.g.h0 - vn) * inv;
                    const float ca = (fmaf(t, da, va) - p.g.ga0) * p.g.inv_ca;
                    const float cb = (fmaf(t, db, vb) - p.g.gb0) * p.g.inv_cb;
                    // float -> int saturates and NaN gives 0: a ray outside the grid fails the unsigned compare
                    const unsigned ia = (unsigned)__float2int_rd(ca), ib = (unsigned)__float2int_rd(cb);
                    if (in_box && ia < (unsigned)p.g.nx && ib < (unsigned)p.g.ny) {
                        const unsigned cell = ib * (unsigned)p.g.nx + ia;
                        if (((bits[cell >> 5] >> (cell & 31u)) & 1u) && !(p.debug & 32)) ent = __ldg(p.g.cells + cell);
                    }
                }
                if (__any_sync(full, out)) { dirty = true; break; }
                // ---- queue the candidates of the previous group (its cell entries have landed by now)
                if (mq) {
                    const unsigned tag = (unsigned)(j - 32) << 24;
                    const unsigned pay = ent_prev >> 8;
                    const int pos = qn + (simple ? __popc(mq & lt_mask) : incl - cnt);
                    if (cnt == 1) sh.queue[pos] = tag | pay;                 // the surface index itself
                    else {
#pragma unroll 1
                        for (int k = 0; k < cnt; ++k) sh.queue[pos + k] = tag | (unsigned)__ldg(p.g.list + pay + k);
                    }
                    qn += tot;
                }
                // surfaces without a finite box are candidates of every ray (accel_tree.py:59-73)
                if (p.n_unbounded && next < chunk_n) {
                    const unsigned mv = __ballot_sync(full, valid);
#pragma unroll 1
                    for (int k = 0; k < p.n_unbounded; ++k) {
                        const unsigned s_unb = (unsigned)__ldg(&p.cand[p.unbounded_first + k]);
                        if (valid) sh.queue[qn + __popc(mv & lt_mask)] = ((unsigned)j << 24) | s_unb;
                        qn += __popc(mv);
                    }
                }
                ent_prev = ent;
            }
            if (!need_flush) break;
            qn = grid_flush<kFlatOnly>(sh, p, chunk_base, lane, qn, false);     // whole batches of 32; the rest stays queued
        }
        if (dirty) {
            // leave the whole chunk to the tree walk (its kernel follows on the stream and writes the chunk's tables)
            if (lane == 0) p.dirty_list[atomicAdd(p.dirty_count, 1ULL)] = (uint32_t)c;
            // step over the groups of this chunk that were requested or due but not keyed
            for (int k = done_g; k < kChunk / 32; ++k) {
                issue();
                slot_con = slot_con + 1 == kGridRing ? 0 : slot_con + 1;
            }
            __syncwarp();
            continue;
        }
        if (!(p.debug & 16)) grid_flush<kFlatOnly>(sh, p, chunk_base, lane, qn, true);
        __syncwarp();
        if (p.sink.hit_surf)
            for (int i = lane; i < chunk_n; i += 32)
                p.sink.hit_surf[chunk_base + i] = sh.best_t[i] == TRC_INF_BITS ? -1 : sh.best_s[i];
        if (p.sink.records && !(p.debug & 8)) grid_emit(sh, p, (int64_t)c, chunk_base, lane);
        __syncwarp();
    }
}

}  // namespace

// 32-bit words of shared memory the shadow-grid kernel can give to a grid's occupancy bits
int trc_walk_grid_word_budget(const TrcScene *) { return kGridMaxWords; }

// dyn != nullptr: device-driven trace (trace.cu).  `n` is then the CAPACITY of the input bundle (it
// sizes the grid), the real count is read from dyn->n_ptr by the kernels, the work counter and the
// hit total live in the trace's control block (already zeroed) and the launch carries the PDL attribute.
static int launch_nearest(const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir, int64_t ld,
                          int32_t use_bvh, HitSink sink, cudaStream_t st, bool zero_total = true,
                          TrcBounceCtl *dyn = nullptr) {
    if (!dyn && sink.records && zero_total) TRC_CUDA(cudaMemsetAsync(sink.total, 0, sizeof(unsigned long long), st));
    if (n == 0) return TRC_OK;
    if (n >= 0x7fffffffLL) return fail(TRC_ERR_CAPACITY, "trc_nearest_hit: more than 2^31-1 rays, chunk the bundle%s");
    if (!use_bvh) {
        int64_t blocks = (n + 255) / 256;
        if (dyn) {
            TRC_CUDA(launch_pdl(nearest_hit_brute_kernel, dim3((unsigned)blocks), dim3(256), 0, st, sc->d_surfaces,
                                sc->d_tables, sc->n_surfaces, n, dyn, d_vert, d_dir, ld, sink));
        } else {
            nearest_hit_brute_kernel<<<(unsigned)blocks, 256, 0, st>>>(sc->d_surfaces, sc->d_tables, sc->n_surfaces, n,
                                                                       (const TrcBounceCtl *)nullptr, d_vert, d_dir, ld, sink);
        }
        return check_launch("nearest_hit_brute");
    }
    if ((sc->n_nodes == 0 && sc->n_unbounded == 0) || !sc->d_chunk_counter || !sc->d_cand)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: use_bvh set but no BVH uploaded%s");
    typedef void (*WalkKernel)(WalkParams);
    static const WalkKernel kernels[6] = {nearest_hit_walk_kernel<false, 0>, nearest_hit_walk_kernel<true, 0>,
                                          nearest_hit_walk_kernel<false, 1>, nearest_hit_walk_kernel<true, 1>,
                                          nearest_hit_walk_kernel<false, 2>, nearest_hit_walk_kernel<true, 2>};
    static int g_refill_at = -1;
    static bool attr_set[64] = {};       // cudaFuncSetAttribute is per device
    if (sc->device < 0 || sc->device >= 64) return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: device index out of range%s");
    if (!attr_set[sc->device]) {
        const char *env = getenv("TRC_WALK_REFILL_AT");      // tuning knob (0..31)
        g_refill_at = 24;
        if (env) { int v = atoi(env); if (v >= 0 && v < 32) g_refill_at = v; }
        for (int k = 0; k < 6; ++k)
            TRC_CUDA(cudaFuncSetAttribute(kernels[k], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(sizeof(WalkShared) + 32 * kSmemPairs)));
        attr_set[sc->device] = true;
    }
    // a tree that fits is cached whole; of a larger one only the top 256 pairs, so that two CTAs still fit an SM
    int n_cached = sc->n_pairs <= kSmemPairs ? sc->n_pairs : kSmemPairs / 2;
    n_cached = (n_cached + 7) & ~7;
    const size_t smem = sizeof(WalkShared) + 32 * (size_t)n_cached;
    const int variant = (sc->n_pairs <= kSmemPairs ? 1 : 0) | (sc->flat_only ? (sc->d_tri64 ? 4 : 2) : 0);
    WalkParams p;
    p.g = TrcShadowGrid();
    p.dirty_list = nullptr; p.dirty_count = nullptr; p.chunk_list = nullptr; p.chunk_list_n = nullptr;
    p.surfs = sc->d_surfaces;
    p.pool = sc->d_tables;
    p.tri32 = sc->d_tri32;
    p.tri64 = sc->d_tri64;
    p.pair_a = reinterpret_cast<const uint4 *>(sc->d_pair_a);
    p.pair_b = reinterpret_cast<const uint4 *>(sc->d_pair_b);
    p.pair_ab = reinterpret_cast<const uint4 *>(sc->d_pair_ab);
    p.n_pairs = sc->n_pairs; p.root_meta = sc->root_meta;
    p.cand = sc->d_cand; p.unbounded_first = sc->n_leaf_surfaces; p.n_unbounded = sc->n_unbounded;
    p.has_tree = sc->n_nodes > 0 ? 1 : 0;
    p.refill_at = g_refill_at;
    p.n_cached = n_cached;
    { const char *dbg = getenv("TRC_WALK_DEBUG"); p.debug = dbg ? atoi(dbg) : 0; }
    for (int k = 0; k < 3; ++k) {
        p.root_lo[k] = sc->root_lo[k]; p.root_hi[k] = sc->root_hi[k];
        p.qstep[k] = sc->qstep[k]; p.extent[k] = sc->extent[k];
    }
    p.n = n; p.ctl = dyn; p.vert = d_vert; p.dir = d_dir; p.ld = ld; p.sink = sink;
    p.chunk_counter = dyn ? &dyn->chunk_counter : sc->d_chunk_counter;
    int64_t n_chunks = (n + kChunk - 1) / kChunk;
    int64_t blocks = (n_chunks + kWalkWarps - 1) / kWalkWarps;
    if (use_bvh == 2 && sc->has_grid) {
        // ---- coherent bundle: the shadow-grid kernel takes every chunk whose rays all lie in the grid's slope box and
        // lists the others; the tree walk below then runs over that list only (a request without a grid set is an
        // ordinary walk)
        typedef void (*GridKernel)(WalkParams);
        static const GridKernel gkernels[3] = {nearest_hit_grid_kernel<0>, nearest_hit_grid_kernel<1>,
                                               nearest_hit_grid_kernel<2>};
        static bool gattr_set[64] = {};
        if (!gattr_set[sc->device]) {
            for (int k = 0; k < 3; ++k)
                TRC_CUDA(cudaFuncSetAttribute(gkernels[k], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)(sizeof(GridShared) + 4 * kGridMaxWords)));
            gattr_set[sc->device] = true;
        }
        TrcScene *scw = const_cast<TrcScene *>(sc);
        if (!scw->d_dirty_counters) TRC_CUDA(cudaMalloc(&scw->d_dirty_counters, 2 * sizeof(unsigned long long)));
        if (scw->dirty_cap < n_chunks) {
            TRC_CUDA(cudaStreamSynchronize(st));             // an earlier launch on this stream may still use the list
            cudaFree(scw->d_dirty);
            scw->d_dirty = nullptr; scw->dirty_cap = 0;
            TRC_CUDA(cudaMalloc(&scw->d_dirty, (size_t)n_chunks * sizeof(uint32_t)));
            scw->dirty_cap = n_chunks;
        }
        TRC_CUDA(cudaMemsetAsync(scw->d_dirty_counters, 0, 2 * sizeof(unsigned long long), st));
        const int gvariant = sc->flat_only ? (sc->d_tri64 ? 2 : 1) : 0;
        const size_t gsmem = sizeof(GridShared) + 4 * (size_t)sc->grid.n_words;
        int gocc = sc->grid_occ[gvariant];
        if (gocc <= 0) {
            TRC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&gocc, gkernels[gvariant], kGridWarps * 32, gsmem));
            if (gocc < 1) gocc = 1;
            scw->grid_occ[gvariant] = gocc;
        }
        WalkParams g = p;
        g.g = sc->grid;
        g.dirty_list = scw->d_dirty; g.dirty_count = scw->d_dirty_counters;
        int64_t gblocks = std::min<int64_t>((n_chunks + kGridWarps - 1) / kGridWarps, (int64_t)sc->num_sms * gocc);
        if (dyn) {
            TRC_CUDA(launch_pdl(gkernels[gvariant], dim3((unsigned)gblocks), dim3(kGridWarps * 32), gsmem, st, g));
        } else {
            TRC_CUDA(cudaMemsetAsync(sc->d_chunk_counter, 0, sizeof(unsigned long long), st));
            gkernels[gvariant]<<<(unsigned)gblocks, kGridWarps * 32, gsmem, st>>>(g);
        }
        int rc = check_launch("nearest_hit_grid");
        if (rc != TRC_OK) return rc;
        // the walk that follows: chunk ids from the list, its own work counter
        p.chunk_list = scw->d_dirty; p.chunk_list_n = scw->d_dirty_counters;
        p.chunk_counter = scw->d_dirty_counters + 1;
        int occ2 = sc->walk_occ[variant];
        if (occ2 <= 0) {
            TRC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, kernels[variant], kWalkWarps * 32, smem));
            if (occ2 < 1) occ2 = 1;
            scw->walk_occ[variant] = occ2;
        }
        int64_t wblocks = std::min<int64_t>(blocks, (int64_t)sc->num_sms * occ2);
        TRC_CUDA(launch_pdl(kernels[variant], dim3((unsigned)wblocks), dim3(kWalkWarps * 32), smem, st, p));
        return check_launch("nearest_hit_walk");
    }
    int occ = sc->walk_occ[variant];
    if (occ <= 0) {
        TRC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernels[variant], kWalkWarps * 32, smem));
        if (occ < 1) occ = 1;
        const_cast<TrcScene *>(sc)->walk_occ[variant] = occ;     // per scene: depends on the tree's shared-memory share
    }
    int64_t resident = (int64_t)sc->num_sms * occ;
    if (blocks > resident) blocks = resident;
    if (dyn) {
        TRC_CUDA(launch_pdl(kernels[variant], dim3((unsigned)blocks), dim3(kWalkWarps * 32), smem, st, p));
    } else {
        TRC_CUDA(cudaMemsetAsync(sc->d_chunk_counter, 0, sizeof(unsigned long long), st));
        kernels[variant]<<<(unsigned)blocks, kWalkWarps * 32, smem, st>>>(p);
    }
    return check_launch("nearest_hit_walk");
}

int trc_nearest_dyn(const TrcScene *sc, int64_t n_cap, TrcBounceCtl *ctl, int32_t use_bvh, TrcHitRecord *d_records,
                    int32_t *d_keys, int64_t capacity, uint32_t *d_chunk_off, uint32_t *d_chunk_cnt, cudaStream_t st) {
    HitSink sink = {nullptr, nullptr, d_records, capacity, d_chunk_off, d_chunk_cnt, &ctl->hits, 0, d_keys};
    return launch_nearest(sc, n_cap, nullptr, nullptr, n_cap, use_bvh, sink, st, false, ctl);
}

extern "C" int trc_nearest_hit(const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir,
                               int64_t ld, int32_t use_bvh, int32_t *d_hit_surf, void *stream) {
    if (!sc || n < 0 || (n > 0 && (!d_vert || !d_dir || !d_hit_surf)) || ld < n)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_nearest_hit: bad argument%s");
    HitSink sink = {d_hit_surf, nullptr, nullptr, 0, nullptr, nullptr, nullptr, 0, nullptr};
    return launch_nearest(sc, n, d_vert, d_dir, ld, use_bvh, sink, (cudaStream_t)stream);
}

static int trace_hits_impl(const char *who, const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir,
                           const double *d_energy, int64_t ld, int32_t use_bvh, int32_t *d_hit_surf,
                           TrcHitRecord *d_records, int32_t *d_keys, int64_t capacity, uint32_t *d_chunk_off,
                           uint32_t *d_chunk_cnt, int64_t *d_total, int64_t ray_base, bool zero_total, void *stream) {
    if (!sc || n < 0 || capacity < 0 || !d_total || ld < n || ray_base < 0 || ray_base % TRC_HIT_CHUNK != 0 ||
        ray_base + n >= 0x7fffffffLL ||
        (n > 0 && (!d_vert || !d_dir || !d_energy || !d_chunk_off || !d_chunk_cnt || (capacity > 0 && !d_records))))
        return fail(TRC_ERR_BAD_ARGUMENT, "%s: bad argument", who);
    static TrcHitRecord dummy_rec;   // never written: capacity 0
    HitSink sink = {d_hit_surf, d_energy, d_records ? d_records : &dummy_rec, d_records ? capacity : 0, d_chunk_off,
                    d_chunk_cnt, reinterpret_cast<unsigned long long *>(d_total), (int32_t)ray_base,
                    d_records ? d_keys : nullptr};
    return launch_nearest(sc, n, d_vert, d_dir, ld, use_bvh, sink, (cudaStream_t)stream, zero_total);
}

extern "C" int trc_trace_hits(const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir,
                              const double *d_energy, int64_t ld, int32_t use_bvh, int32_t *d_hit_surf,
                              TrcHitRecord *d_records, int64_t capacity, uint32_t *d_chunk_off,
                              uint32_t *d_chunk_cnt, int64_t *d_total, void *stream) {
    return trace_hits_impl("trc_trace_hits", sc, n, d_vert, d_dir, d_energy, ld, use_bvh, d_hit_surf, d_records,
                           nullptr, capacity, d_chunk_off, d_chunk_cnt, d_total, 0, true, stream);
}

extern "C" int trc_trace_hits_slice(const TrcScene *sc, int64_t n, const double *d_vert, const double *d_dir,
                                    const double *d_energy, int64_t ld, int32_t use_bvh, int32_t *d_hit_surf,
                                    TrcHitRecord *d_records, int32_t *d_keys, int64_t capacity, uint32_t *d_chunk_off,
                                    uint32_t *d_chunk_cnt, int64_t *d_total, int64_t ray_base, void *stream) {
    return trace_hits_impl("trc_trace_hits_slice", sc, n, d_vert, d_dir, d_energy, ld, use_bvh, d_hit_surf,
                           d_records, d_keys, capacity, d_chunk_off, d_chunk_cnt, d_total, ray_base, false, stream);
}
