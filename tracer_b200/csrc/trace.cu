// trc_trace: several iterations of the reference loop body (tracer_engine.py:193-274) enqueued
// back to back with NO host synchronisation in between.  Every count a later kernel needs -- rays
// entering a bounce, hits, survivors -- lives in a device-side control block (TrcBounceCtl); grids are
// sized from host-side capacities; an overflowing bounce raises a flag, does nothing else and feeds
// nothing forward.  The caller synchronises once, after the final asynchronous copy of the counts.
//
// Kernels of one bounce (each starts with griddepcontrol.wait and is launched with the programmatic
// stream-serialisation attribute, so its prologue overlaps the tail of the one before):
//   1. nearest_hit_walk / nearest_hit_brute (nearest.cu)  ray count read from the control block
//   2. order_hist      per (tile, surface) hit counts from the 4-byte key array written next to the hit records
//      order_scan      one block: scans them over the tiles -> slot offsets per surface (seg), hit totals
//   3. order_scatter   stable (surface-major, ray-minor) slot -> record index
//      [scenes with more than 256 surfaces: passes 1.. of the LSD radix sort of sort.cu + segment offsets]
//   4. optics          (bounce.cu) children in slot order, accountant columns, per-surface absorbed sums
//      bounce_close    one block: survivors, reorder flag, the next bounce's ray count and bundle
//   5. partition       only acts when some but not all children are weak: a survivors-first copy of the
//                      children (tracer_engine.py:242-245, 268-274), which the next bounce then reads -- the
//                      choice is recorded in the next control block, so no pointer passes through the host
#include <new>

#include "common.cuh"

int trc_sort_tail_passes(int64_t n_cap, const int64_t *d_n, int32_t n_keys, int first_shift, int32_t *k_a,
                         int32_t *v_a, int32_t *k_b, int32_t *v_b, void *d_work, cudaStream_t st, bool *in_a);

namespace {

inline int64_t up256(int64_t x) { return (x + 255) & ~(int64_t)255; }

constexpr int kOrderWarps = 8;
constexpr int kMaxBins = 256;
#ifndef TRC_ORDER_MIN_CHUNKS
#define TRC_ORDER_MIN_CHUNKS 1
#endif
#ifndef TRC_ORDER_TILES
#define TRC_ORDER_TILES 592
#endif
constexpr int kTargetTiles = TRC_ORDER_TILES;   // 4 CTAs per SM: bounds the (tile, surface) table the scan block walks

struct OrderParams {
    TrcBounceCtl *ctl;
    int64_t capacity;               // hit capacity: offsets beyond it were never written
    const uint32_t *chunk_off, *chunk_cnt;
    const int32_t *keys;            // surface index per record
    int chunks_per_tile;            // chunks of 256 rays handled by one CTA, in ray order (an eighth per warp)
    int n_tiles;                    // tiles covering the capacity of the input bundle
    int nbins;                      // n_keys when <= 256 (single pass), else 256 (low digit of a radix sort)
    int n_keys;
    uint32_t *hist;                 // [n_tiles][nbins]; exclusive-scanned over the tiles by the last block
    int64_t *bin_off;               // [nbins + 1]: == seg (n_keys + 1) for a single pass
    int64_t *hits_per_surface;      // single pass: += per-surface counts (nullable)
    int64_t *h_eff;                 // clamped hit count for the kernels that follow
    int32_t *out_keys, *out_vals;   // scatter targets: slot -> surface, slot -> record index
};

// Chunk range [w0, w1) of this warp: the tile's chunks split into contiguous eighths, so that
// (tile, warp, position) order is ray order.
__device__ __forceinline__ void warp_chunk_range(const OrderParams &p, long long n_chunks, long long &w0, long long &w1) {
    const int warp = threadIdx.x >> 5;
    const long long t0 = (long long)blockIdx.x * p.chunks_per_tile;
    long long t1 = t0 + p.chunks_per_tile;
    if (t1 > n_chunks) t1 = n_chunks;
    const int per_warp = (p.chunks_per_tile + kOrderWarps - 1) / kOrderWarps;
    w0 = t0 + (long long)warp * per_warp;
    w1 = w0 + per_warp;
    if (w1 > t1) w1 = t1;
}

// Walk the hits of the chunks [c0, c1) in ray order: chunk after chunk, records of a chunk in ascending ray
// index.  f(key, record index, valid) is called with all 32 lanes converged.
template <typename F>
__device__ __forceinline__ void for_each_hit(const OrderParams &p, long long c0, long long c1, F f) {
    const int lane = threadIdx.x & 31;
    for (long long cb = c0; cb < c1; cb += 32) {
        long long c = cb + lane;
        uint32_t off = 0, cnt = 0;
        if (c < c1) { off = p.chunk_off[c]; cnt = p.chunk_cnt[c]; }
        if ((long long)off >= p.capacity) cnt = 0;
        else if ((long long)off + cnt > p.capacity) cnt = (uint32_t)(p.capacity - off);
        const int nc = (int)((c1 - cb) < 32 ? (c1 - cb) : 32);
        for (int k = 0; k < nc; ++k) {
            const uint32_t o = __shfl_sync(0xffffffffu, off, k), n = __shfl_sync(0xffffffffu, cnt, k);
            for (uint32_t i = 0; i < n; i += 32) {
                const bool valid = i + lane < n;
                const int32_t key = valid ? p.keys[o + i + lane] : -1;
                f(key, (int32_t)(o + i + lane), valid);
            }
        }
    }
}

// Per-warp digit counts of this warp's chunk range into cnt[warp][*] (zeroed here).
__device__ __forceinline__ void count_digits(const OrderParams &p, long long n_chunks, uint32_t (*cnt)[kMaxBins]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool single = p.n_keys <= kMaxBins;
    for (int b = lane; b < p.nbins; b += 32) cnt[warp][b] = 0;
    __syncwarp();
    long long w0, w1;
    warp_chunk_range(p, n_chunks, w0, w1);
    for_each_hit(p, w0, w1, [&](int32_t key, int32_t, bool valid) {
        const int dg = valid ? (single ? key : (key & 255)) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, dg);
        if (valid && (peers & ((1u << lane) - 1u)) == 0u) cnt[warp][dg] += __popc(peers);
        __syncwarp();
    });
}

__global__ void __launch_bounds__(kOrderWarps * 32)
order_hist_kernel(OrderParams p) {
    __shared__ uint32_t cnt[kOrderWarps][kMaxBins];
    pdl_wait();
    pdl_trigger();
    const long long hraw = (long long)p.ctl->hits;
    const bool overflow = hraw > p.capacity;
    const long long n = overflow ? 0 : p.ctl->n_in;
    const long long n_chunks = (n + TRC_HIT_CHUNK - 1) / TRC_HIT_CHUNK;
    count_digits(p, n_chunks, cnt);
    __syncthreads();
    if ((int)threadIdx.x < p.nbins) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < kOrderWarps; ++w) t += cnt[w][threadIdx.x];
        p.hist[(size_t)blockIdx.x * p.nbins + threadIdx.x] = t;
    }
}

// One block of kScanWarps warps: exclusive scan of every bin's column of the (tile, bin) table over the tiles, then
// of the bin totals.  Warp w owns a contiguous band of tiles; lanes own bins lane, lane + 32, ... (coalesced rows).
// Pass 1 sums the bands, pass 2 rewrites them as running prefixes starting from the sum of the bands before.
// 32 warps: the bands are short (19 rows of 592), the kernel's ~50 us at 8 warps were a serial row walk.
constexpr int kScanWarps = 32;
__global__ void __launch_bounds__(kScanWarps * 32)
order_scan_kernel(OrderParams p) {
    __shared__ uint32_t cnt[kScanWarps][kMaxBins];
    __shared__ long long s_tot[kMaxBins];
    pdl_wait();
    pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool single = p.n_keys <= kMaxBins;
    const int rows = p.n_tiles;
    const int band = (rows + kScanWarps - 1) / kScanWarps;
    const int r0 = min(rows, warp * band), r1 = min(rows, r0 + band);
    uint32_t acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0;
    for (int r = r0; r < r1; ++r) {
        const uint32_t *row = p.hist + (size_t)r * p.nbins;
#pragma unroll
        for (int k = 0; k < 8; ++k) { int b = lane + 32 * k; if (b < p.nbins) acc[k] += row[b]; }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) cnt[warp][lane + 32 * k] = acc[k];
    __syncthreads();
    uint32_t run[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        uint32_t base = 0;
        for (int w = 0; w < warp; ++w) base += cnt[w][lane + 32 * k];
        run[k] = base;
    }
    for (int r = r0; r < r1; r += 4) {
        uint32_t v[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t *row = p.hist + (size_t)(r + q) * p.nbins;
#pragma unroll
            for (int k = 0; k < 8; ++k) { int b = lane + 32 * k; v[q][k] = (r + q < r1 && b < p.nbins) ? row[b] : 0u; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t *row = p.hist + (size_t)(r + q) * p.nbins;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                int b = lane + 32 * k;
                if (r + q < r1 && b < p.nbins) row[b] = run[k];
                run[k] += v[q][k];
            }
        }
    }
    // bin totals and their exclusive scan (= slot offsets per surface for a single pass): the first 256 threads
    const bool low = threadIdx.x < kMaxBins;
    long long tot = 0;
    if (low && (int)threadIdx.x < p.nbins) {
#pragma unroll 8
        for (int w = 0; w < kScanWarps; ++w) tot += cnt[w][threadIdx.x];
    }
    if (low) s_tot[threadIdx.x] = tot;
    __syncthreads();
    for (int o = 1; o < kMaxBins; o <<= 1) {          // inclusive Hillis-Steele scan over the 256 bins
        long long add = (low && (int)threadIdx.x >= o) ? s_tot[threadIdx.x - o] : 0;
        __syncthreads();
        if (low) s_tot[threadIdx.x] += add;
        __syncthreads();
    }
    if (low && (int)threadIdx.x < p.nbins) {
        p.bin_off[threadIdx.x] = s_tot[threadIdx.x] - tot;
        if (single && p.hits_per_surface) p.hits_per_surface[threadIdx.x] += tot;
    }
    if (threadIdx.x == kMaxBins - 1) {
        p.bin_off[p.nbins] = s_tot[threadIdx.x];
        *p.h_eff = s_tot[threadIdx.x];
    }
}

__global__ void __launch_bounds__(kOrderWarps * 32)
order_scatter_kernel(OrderParams p) {
    __shared__ uint32_t pos[kOrderWarps][kMaxBins];
    pdl_wait();
    pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long hraw = (long long)p.ctl->hits;
    if (hraw > p.capacity) return;
    const long long n = p.ctl->n_in;
    const long long n_chunks = (n + TRC_HIT_CHUNK - 1) / TRC_HIT_CHUNK;
    if ((long long)blockIdx.x * p.chunks_per_tile >= n_chunks) return;
    const bool single = p.n_keys <= kMaxBins;
    // counts per warp again (4 bytes per hit out of L2), then this warp's first slot of every bin
    count_digits(p, n_chunks, pos);
    __syncthreads();
    if ((int)threadIdx.x < p.nbins) {
        uint32_t base = (uint32_t)p.bin_off[threadIdx.x] + p.hist[(size_t)blockIdx.x * p.nbins + threadIdx.x];
#pragma unroll
        for (int w = 0; w < kOrderWarps; ++w) { uint32_t c = pos[w][threadIdx.x]; pos[w][threadIdx.x] = base; base += c; }
    }
    __syncthreads();
    long long w0, w1;
    warp_chunk_range(p, n_chunks, w0, w1);
    for_each_hit(p, w0, w1, [&](int32_t key, int32_t rec, bool valid) {
        const int dg = valid ? (single ? key : (key & 255)) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, dg);
        const unsigned before = __popc(peers & ((1u << lane) - 1u));
        uint32_t base = 0;
        if (valid) base = pos[warp][dg];
        __syncwarp();
        if (valid && before == 0) pos[warp][dg] = base + __popc(peers);
        __syncwarp();
        if (valid) {
            p.out_keys[base + before] = key;
            p.out_vals[base + before] = rec;
        }
    });
}

__global__ void add_counts_dyn_kernel(const int64_t *__restrict__ seg, int n_surfaces, int64_t *__restrict__ hits) {
    pdl_wait();
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_surfaces) hits[s] += seg[s + 1] - seg[s];
}

// ---- survivors first, weak rays to the back (tracer_engine.py:242-245, 268-274)
struct Columns {
    double *v, *d, *e;
    int64_t *par, *ids;
    double *w;                     // nullable
    int64_t ld;
};
struct PartParams {
    const TrcBounceCtl *ctl;
    int64_t capacity;
    const int32_t *cull;
    const uint32_t *cta_surv;      // exclusive scan of the survivors per 256-slot block (optics finalize)
    Columns a, b;                  // a: slot order (written by the optics kernel); b: the other copy
};

__global__ void __launch_bounds__(256)
partition_kernel(PartParams p) {
    __shared__ unsigned s_w[8];
    pdl_wait();
    pdl_trigger();
    if (!p.ctl->reorder) return;
    const long long hraw = (long long)p.ctl->hits;
    const long long h = hraw < p.capacity ? hraw : p.capacity;
    const long long blk0 = (long long)blockIdx.x * 256;
    if (blk0 >= h) return;
    const long long q = blk0 + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool live = q < h;
    const bool surv = live && p.cull[q] == 0;
    const unsigned m = __ballot_sync(0xffffffffu, surv);
    if (lane == 0) s_w[warp] = __popc(m);
    __syncthreads();
    unsigned before = p.cta_surv[blockIdx.x] + __popc(m & ((1u << lane) - 1u));
    for (int w = 0; w < warp; ++w) before += s_w[w];
    if (!live) return;
    const long long n_surv = p.ctl->n_surv;
    const long long dst = surv ? (long long)before : n_surv + (q - (long long)before);
    const Columns &a = p.a, &b = p.b;
    double v0 = a.v[q], v1 = a.v[a.ld + q], v2 = a.v[2 * a.ld + q];
    double d0 = a.d[q], d1 = a.d[a.ld + q], d2 = a.d[2 * a.ld + q];
    double en = a.e[q];
    int64_t pp = a.par[q], id = a.ids[q];
    b.v[dst] = v0; b.v[b.ld + dst] = v1; b.v[2 * b.ld + dst] = v2;
    b.d[dst] = d0; b.d[b.ld + dst] = d1; b.d[2 * b.ld + dst] = d2;
    b.e[dst] = en; b.par[dst] = pp; b.ids[dst] = id;
    if (a.w) b.w[dst] = a.w[q];
}

__global__ void trace_init_kernel(TrcBounceCtl *ctl, int n_ctl, long long n0, TrcBundleDesc in0, int64_t *counts,
                                  int n_counts, double *absorbed, int64_t *hits, int n_tally) {
    pdl_wait();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int k = i; k < n_tally; k += gridDim.x * blockDim.x) { absorbed[k] = 0.0; hits[k] = 0; }
    long long *w = reinterpret_cast<long long *>(ctl);
    const int words = n_ctl * (int)(sizeof(TrcBounceCtl) / 8);
    const int w_in = (int)(offsetof(TrcBounceCtl, in) / 8), n_in_words = (int)(sizeof(TrcBundleDesc) / 8);
    if (i < words) {
        long long v = 0;
        if (i == 0) v = n0;                                            // ctl[0].n_in
        else if (i >= w_in && i < w_in + n_in_words) v = reinterpret_cast<const long long *>(&in0)[i - w_in];
        w[i] = v;
    }
    if (i < n_counts) counts[i] = 0;
}

// ---- memory plan
struct KeepOne {
    int64_t off[21];
    int64_t total;
};

void bounce_layout(int64_t base, int64_t cap, uint32_t columns, int has_w, int32_t n_surfaces, TrcTraceBounceLayout *o,
                   int64_t *end) {
    int64_t off = base;
    auto take = [&](int64_t bytes) { int64_t r = off; off = up256(off + bytes); return r; };
    o->ld = cap;
    o->off_vertices = take(24 * cap); o->off_directions = take(24 * cap); o->off_energy = take(8 * cap);
    o->off_parents = take(8 * cap); o->off_ray_ids = take(8 * cap);
    o->off_wavelengths = has_w ? take(8 * cap) : -1;
    o->off_alt_vertices = take(24 * cap); o->off_alt_directions = take(24 * cap); o->off_alt_energy = take(8 * cap);
    o->off_alt_parents = take(8 * cap); o->off_alt_ray_ids = take(8 * cap);
    o->off_alt_wavelengths = has_w ? take(8 * cap) : -1;
    o->off_absorbed = take(8 * cap);
    o->off_normals = (columns & TRC_COL_NORMALS) ? take(24 * cap) : -1;
    o->off_e_in = (columns & TRC_COL_E_IN) ? take(8 * cap) : -1;
    o->off_dirs_in = (columns & TRC_COL_DIRS_IN) ? take(24 * cap) : -1;
    o->off_attenuated = (columns & TRC_COL_ATTENUATED) ? take(8 * cap) : -1;
    o->off_seg = take(8 * ((int64_t)n_surfaces + 1));
    *end = off;
}

struct ScratchPlan {
    int64_t off_ctl, off_counts, off_heff, off_records, off_keys, off_chunk_off, off_chunk_cnt, off_slot_surf,
        off_slot_rec, off_alt_k, off_alt_v, off_hist, off_bin_off, off_sort_work, off_cull, off_cta_surv, total;
    int64_t cap_max, in_max;
};

int tiles_for(int64_t n_in_cap, int *chunks_per_tile) {
    int64_t n_chunks = (n_in_cap + TRC_HIT_CHUNK - 1) / TRC_HIT_CHUNK;
    if (n_chunks < 1) n_chunks = 1;
    int64_t g = (n_chunks + kTargetTiles - 1) / kTargetTiles;
    // at least TRC_ORDER_MIN_CHUNKS chunks per warp: small bundles get few tiles (short scan, fuller CTAs)
    if (g < kOrderWarps * TRC_ORDER_MIN_CHUNKS) g = kOrderWarps * TRC_ORDER_MIN_CHUNKS;
    *chunks_per_tile = (int)g;
    return (int)((n_chunks + g - 1) / g);
}

// Words of the read-back per bounce: [rays in, hits, survivors, reorder] and, when asked for, seg -- gathered
// next to the counts on the device (bounce_close) so that ONE copy brings everything back.  Scenes with many
// surfaces (meshes) keep seg where it is and copy it separately.
constexpr int kInlineSegMax = 4096;
inline bool seg_inline(const TrcScene *sc, const TrcTraceArgs *a) {
    return (a->flags & TRC_TRACE_SEG) && sc->n_surfaces + 1 <= kInlineSegMax;
}
inline int64_t counts_per_bounce(const TrcScene *sc, const TrcTraceArgs *a) {
    return 4 + (seg_inline(sc, a) ? sc->n_surfaces + 1 : 0);
}

int validate(const TrcScene *sc, const TrcTraceArgs *a, const char *who) {
    if (!sc || !a || a->n < 0 || a->n_bounces < 1 || a->n_bounces > TRC_TRACE_MAX_BOUNCES || a->ld < a->n)
        return fail(TRC_ERR_BAD_ARGUMENT, "%s: bad argument", who);
    if (sc->has_split)
        return fail(TRC_ERR_UNSUPPORTED_KIND, "%s: scenes with two-children optics (periodic, refractive) use the step-by-step calls", who);
    if (a->n >= 0x7fffffffLL) return fail(TRC_ERR_CAPACITY, "%s: more than 2^31-1 rays, chunk the bundle", who);
    for (int j = 0; j < a->n_bounces; ++j)
        if (a->capacity[j] < 256 || a->capacity[j] % 256 != 0 || a->capacity[j] >= 0x7fffffffLL)
            return fail(TRC_ERR_BAD_ARGUMENT, "%s: capacities must be positive multiples of 256", who);
    return TRC_OK;
}

ScratchPlan scratch_plan(const TrcScene *sc, const TrcTraceArgs *a) {
    ScratchPlan s;
    const int32_t S = sc->n_surfaces;
    int64_t cap_max = 0, in_max = a->n;
    for (int j = 0; j < a->n_bounces; ++j) {
        if (a->capacity[j] > cap_max) cap_max = a->capacity[j];
        if (j + 1 < a->n_bounces && a->capacity[j] > in_max) in_max = a->capacity[j];
    }
    s.cap_max = cap_max; s.in_max = in_max;
    const int64_t n_chunks = (in_max + TRC_HIT_CHUNK - 1) / TRC_HIT_CHUNK + 1;
    int g;
    const int64_t tiles = tiles_for(in_max, &g);
    const int nbins = S <= kMaxBins ? S : kMaxBins;
    const int64_t nb = (cap_max + 255) / 256;
    int64_t off = 0;
    auto take = [&](int64_t bytes) { int64_t r = off; off = up256(off + bytes); return r; };
    s.off_ctl = take((int64_t)sizeof(TrcBounceCtl) * (TRC_TRACE_MAX_BOUNCES + 1));
    s.off_counts = take(8 * counts_per_bounce(sc, a) * TRC_TRACE_MAX_BOUNCES);
    s.off_heff = take(8 * TRC_TRACE_MAX_BOUNCES);
    s.off_records = take(64 * cap_max);
    s.off_keys = take(4 * cap_max);
    s.off_chunk_off = take(4 * n_chunks);
    s.off_chunk_cnt = take(4 * n_chunks);
    s.off_slot_surf = take(4 * cap_max);
    s.off_slot_rec = take(4 * cap_max);
    const bool multi = S > kMaxBins;
    s.off_alt_k = take(multi ? 4 * cap_max : 0);
    s.off_alt_v = take(multi ? 4 * cap_max : 0);
    (void)tiles;
    s.off_hist = take(4 * (int64_t)kTargetTiles * nbins);      // tiles_for never exceeds kTargetTiles, but is not monotonic in n
    s.off_bin_off = take(8 * (nbins + 1));
    s.off_sort_work = take(multi ? trc_sort_work_bytes(cap_max, S) : 0);
    s.off_cull = take(4 * cap_max);
    s.off_cta_surv = take(4 * nb);
    s.total = off;
    return s;
}

}  // namespace

extern "C" int trc_trace_plan(const TrcScene *sc, const TrcTraceArgs *a, TrcTracePlan *plan) {
    if (!plan) return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace_plan: bad argument%s");
    int rc = validate(sc, a, "trc_trace_plan");
    if (rc != TRC_OK) return rc;
    memset(plan, 0, sizeof(*plan));
    const int has_w = a->d_wavelengths ? 1 : 0;
    int64_t off = 0;
    for (int j = 0; j < a->n_bounces; ++j) bounce_layout(off, a->capacity[j], a->columns, has_w, sc->n_surfaces, &plan->bounce[j], &off);
    plan->keep_bytes = off > 0 ? off : 256;
    ScratchPlan s = scratch_plan(sc, a);
    plan->scratch_bytes = s.total;
    plan->counts_words = (int64_t)a->n_bounces * (4 + ((a->flags & TRC_TRACE_SEG) ? sc->n_surfaces + 1 : 0));
    plan->off_records = s.off_records; plan->off_keys = s.off_keys; plan->off_chunk_off = s.off_chunk_off;
    plan->off_chunk_cnt = s.off_chunk_cnt;
    plan->off_total = s.off_ctl + (int64_t)offsetof(TrcBounceCtl, hits);
    return TRC_OK;
}

// Zeroes the control blocks of a trace (and sets the first ray count).  trc_trace does this itself unless the
// caller runs the first nearest-hit pass (TRC_TRACE_WALK_DONE): then it goes first, before the slices.
extern "C" int trc_trace_init(const TrcScene *sc, const TrcTraceArgs *a, void *stream) {
    int rc = validate(sc, a, "trc_trace_init");
    if (rc != TRC_OK) return rc;
    if (!a->d_scratch) return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace_init: no scratch arena%s");
    ScratchPlan s = scratch_plan(sc, a);
    if (a->scratch_bytes < s.total) return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace_init: scratch arena smaller than planned%s");
    char *scr = (char *)a->d_scratch;
    const int words = (TRC_TRACE_MAX_BOUNCES + 1) * (int)(sizeof(TrcBounceCtl) / 8);
    TrcBundleDesc in0 = {a->d_vertices, a->d_directions, a->d_energy, a->d_wavelengths,
                         (const long long *)a->d_ray_ids, (long long)a->ray_id_base, (long long)a->ld};
    const int n_tally = (a->flags & TRC_TRACE_ZERO_TALLIES) ? sc->n_surfaces : 0;
    if (n_tally && (!a->d_absorbed_per_surface || !a->d_hits_per_surface))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace_init: TRC_TRACE_ZERO_TALLIES without tally arrays%s");
    int blocks = (words + 255) / 256;
    if ((n_tally + 255) / 256 > blocks) blocks = (n_tally + 255) / 256;
    if (blocks > 1024) blocks = 1024;
    TRC_CUDA(launch_pdl(trace_init_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream,
                        (TrcBounceCtl *)(scr + s.off_ctl), TRC_TRACE_MAX_BOUNCES + 1, (long long)a->n, in0,
                        (int64_t *)(scr + s.off_counts), (int)(counts_per_bounce(sc, a) * TRC_TRACE_MAX_BOUNCES), a->d_absorbed_per_surface,
                        a->d_hits_per_surface, n_tally));
    return check_launch("trace_init");
}

extern "C" int trc_trace(const TrcScene *sc, const TrcTraceArgs *a, void *stream) {
    int rc = validate(sc, a, "trc_trace");
    if (rc != TRC_OK) return rc;
    if (!a->d_keep || !a->d_scratch || !a->d_absorbed_per_surface || !a->d_hits_per_surface || !a->h_counts ||
        (a->n > 0 && (!a->d_vertices || !a->d_directions || !a->d_energy)))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace: null buffer%s");
    cudaStream_t st = (cudaStream_t)stream;
    const int32_t S = sc->n_surfaces;
    const int has_w = a->d_wavelengths ? 1 : 0;
    TrcTracePlan plan;
    rc = trc_trace_plan(sc, a, &plan);
    if (rc != TRC_OK) return rc;
    ScratchPlan sp = scratch_plan(sc, a);
    if (a->keep_bytes < plan.keep_bytes || a->scratch_bytes < plan.scratch_bytes)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace: arena smaller than trc_trace_plan asked for%s");
    char *scr = (char *)a->d_scratch, *keep = (char *)a->d_keep;
    TrcBounceCtl *ctl = (TrcBounceCtl *)(scr + sp.off_ctl);
    int64_t *d_counts = (int64_t *)(scr + sp.off_counts);
    const int64_t cpb = counts_per_bounce(sc, a);
    int64_t *d_heff = (int64_t *)(scr + sp.off_heff);
    TrcHitRecord *records = (TrcHitRecord *)(scr + sp.off_records);
    int32_t *keys = (int32_t *)(scr + sp.off_keys);
    uint32_t *chunk_off = (uint32_t *)(scr + sp.off_chunk_off), *chunk_cnt = (uint32_t *)(scr + sp.off_chunk_cnt);
    int32_t *slot_surf = (int32_t *)(scr + sp.off_slot_surf), *slot_rec = (int32_t *)(scr + sp.off_slot_rec);
    int32_t *alt_k = (int32_t *)(scr + sp.off_alt_k), *alt_v = (int32_t *)(scr + sp.off_alt_v);
    const int nbins = S <= kMaxBins ? S : kMaxBins;
    const bool multi = S > kMaxBins;

    if (!(a->flags & TRC_TRACE_WALK_DONE)) {
        rc = trc_trace_init(sc, a, stream);
        if (rc != TRC_OK) return rc;
    }
    int64_t in_cap = a->n;            // capacity of the bundle entering the current bounce (it sizes the grids)
    for (int j = 0; j < a->n_bounces; ++j) {
        const int64_t cap = a->capacity[j];
        const TrcTraceBounceLayout &L = plan.bounce[j];
        TrcBounceCtl *c = ctl + j;
        // 1. nearest hit + hit records
        if (!(j == 0 && (a->flags & TRC_TRACE_WALK_DONE)) && in_cap > 0) {
            const bool timed = j == 0 && (a->flags & TRC_TRACE_TIME_WALK);
            TrcScene *scw = const_cast<TrcScene *>(sc);
            if (timed) {
                if (!scw->ev_walk[0]) {
                    TRC_CUDA(cudaEventCreate(&scw->ev_walk[0]));
                    TRC_CUDA(cudaEventCreate(&scw->ev_walk[1]));
                }
                TRC_CUDA(cudaEventRecord(scw->ev_walk[0], st));
            }
            // use_bvh == 2 (shadow grid of a coherent bundle) concerns the bundle the caller passed: bounce j == 0
            rc = trc_nearest_dyn(sc, in_cap, c, (j == 0 || a->use_bvh < 2) ? a->use_bvh : 1, records, keys, cap,
                                 chunk_off, chunk_cnt, st);
            if (rc != TRC_OK) return rc;
            if (timed) TRC_CUDA(cudaEventRecord(scw->ev_walk[1], st));
        }
        // 2./3. child order
        int64_t *seg = (int64_t *)(keep + L.off_seg);
        OrderParams op;
        op.ctl = c; op.capacity = cap; op.chunk_off = chunk_off; op.chunk_cnt = chunk_cnt; op.keys = keys;
        op.n_tiles = tiles_for(in_cap > 0 ? in_cap : 1, &op.chunks_per_tile);
        op.nbins = nbins; op.n_keys = S;
        op.hist = (uint32_t *)(scr + sp.off_hist);
        op.bin_off = multi ? (int64_t *)(scr + sp.off_bin_off) : seg;
        op.hits_per_surface = multi ? nullptr : a->d_hits_per_surface;
        op.h_eff = d_heff + j;
        // the low digit lands where an even number of further passes brings the result back to the slot arrays
        int key_bits = 1;
        while (key_bits < 31 && (1LL << key_bits) < (long long)S) ++key_bits;
        const int tail_passes = multi ? (key_bits - 8 + 7) / 8 : 0;
        const bool first_to_alt = (tail_passes & 1) != 0;
        op.out_keys = first_to_alt ? alt_k : slot_surf;
        op.out_vals = first_to_alt ? alt_v : slot_rec;
        const unsigned oblocks = (unsigned)op.n_tiles;
        TRC_CUDA(launch_pdl(order_hist_kernel, dim3(oblocks), dim3(kOrderWarps * 32), 0, st, op));
        TRC_CUDA(launch_pdl(order_scan_kernel, dim3(1), dim3(kScanWarps * 32), 0, st, op));
        TRC_CUDA(launch_pdl(order_scatter_kernel, dim3(oblocks), dim3(kOrderWarps * 32), 0, st, op));
        rc = check_launch("order", 3);
        if (rc != TRC_OK) return rc;
        if (multi) {
            bool in_first = true;
            rc = trc_sort_tail_passes(cap, d_heff + j, S, 8, op.out_keys, op.out_vals, first_to_alt ? slot_surf : alt_k,
                                      first_to_alt ? slot_rec : alt_v, scr + sp.off_sort_work, st, &in_first);
            if (rc != TRC_OK) return rc;
            if (in_first != !first_to_alt) return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace: radix pass parity%s");
            rc = trc_segment_offsets_dyn(d_heff + j, cap, S, slot_surf, seg, stream);
            if (rc != TRC_OK) return rc;
            add_counts_dyn_kernel<<<(S + 255) / 256, 256, 0, st>>>(seg, S, a->d_hits_per_surface);
            rc = check_launch("add_counts");
            if (rc != TRC_OK) return rc;
        }
        // 4. optics, accountant columns, per-surface sums; its last block closes the bounce
        TrcOpticsArgs o;
        memset(&o, 0, sizeof(o));
        o.h = cap; o.d_records = records; o.d_slot_rec = slot_rec;       // wavelengths / ids: from the control block
        o.d_out_vertices = (double *)(keep + L.off_vertices); o.d_out_directions = (double *)(keep + L.off_directions);
        o.d_out_energy = (double *)(keep + L.off_energy); o.d_out_parents = (int64_t *)(keep + L.off_parents);
        o.d_out_ray_ids = (int64_t *)(keep + L.off_ray_ids);
        o.ld_out = cap;
        o.seed = a->seed; o.bounce = a->first_bounce + (uint32_t)j; o.min_energy = a->min_energy;
        o.d_absorbed = (double *)(keep + L.off_absorbed);
        o.d_normals = L.off_normals >= 0 ? (double *)(keep + L.off_normals) : nullptr;
        o.d_e_in = L.off_e_in >= 0 ? (double *)(keep + L.off_e_in) : nullptr;
        o.d_dirs_in = L.off_dirs_in >= 0 ? (double *)(keep + L.off_dirs_in) : nullptr;
        o.d_attenuated = L.off_attenuated >= 0 ? (double *)(keep + L.off_attenuated) : nullptr;
        o.d_out_wavelengths = has_w ? (double *)(keep + L.off_wavelengths) : nullptr;
        o.d_cull = (int32_t *)(scr + sp.off_cull);
        TrcOpticsDyn od;
        od.ctl = c; od.ctl_next = c + 1;
        od.children = {o.d_out_vertices, o.d_out_directions, o.d_out_energy, o.d_out_wavelengths,
                       (const long long *)o.d_out_ray_ids, 0, (long long)cap};
        od.children_sorted = {(const double *)(keep + L.off_alt_vertices), (const double *)(keep + L.off_alt_directions),
                              (const double *)(keep + L.off_alt_energy),
                              has_w ? (const double *)(keep + L.off_alt_wavelengths) : nullptr,
                              (const long long *)(keep + L.off_alt_ray_ids), 0, (long long)cap};
        od.cta_surv = (uint32_t *)(scr + sp.off_cta_surv);
        od.absorbed_per_surface = a->d_absorbed_per_surface;
        od.d_counts = d_counts + cpb * j;
        od.seg = seg_inline(sc, a) ? seg : nullptr;
        od.n_seg = S + 1;
        rc = trc_optics_dyn(sc, &o, &od, stream);
        if (rc != TRC_OK) return rc;
        // 5. survivors-first copy (acts only when some but not all children are weak)
        PartParams pp;
        pp.ctl = c; pp.capacity = cap; pp.cull = o.d_cull; pp.cta_surv = od.cta_surv;
        pp.a = {o.d_out_vertices, o.d_out_directions, o.d_out_energy, o.d_out_parents, o.d_out_ray_ids,
                o.d_out_wavelengths, cap};
        pp.b = {(double *)(keep + L.off_alt_vertices), (double *)(keep + L.off_alt_directions),
                (double *)(keep + L.off_alt_energy), (int64_t *)(keep + L.off_alt_parents),
                (int64_t *)(keep + L.off_alt_ray_ids), has_w ? (double *)(keep + L.off_alt_wavelengths) : nullptr, cap};
        const unsigned pblocks = (unsigned)((cap + 255) / 256);
        TRC_CUDA(launch_pdl(partition_kernel, dim3(pblocks), dim3(256), 0, st, pp));
        rc = check_launch("partition");
        if (rc != TRC_OK) return rc;
        in_cap = cap;      // the children are the next bundle; which copy, the device decided (TrcBounceCtl::in)
    }
    // one read-back for the whole trace
    const int64_t per = 4 + ((a->flags & TRC_TRACE_SEG) ? S + 1 : 0);
    if (per == cpb) {
        TRC_CUDA(cudaMemcpyAsync(a->h_counts, d_counts, sizeof(int64_t) * (size_t)(per * a->n_bounces), cudaMemcpyDeviceToHost, st));
    } else {
        for (int j = 0; j < a->n_bounces; ++j) {
            TRC_CUDA(cudaMemcpyAsync(a->h_counts + per * j, d_counts + cpb * j, sizeof(int64_t) * 4, cudaMemcpyDeviceToHost, st));
            TRC_CUDA(cudaMemcpyAsync(a->h_counts + per * j + 4, keep + plan.bounce[j].off_seg,
                                     sizeof(int64_t) * (size_t)(S + 1), cudaMemcpyDeviceToHost, st));
        }
    }
    return TRC_OK;
}

// Device time of the first bounce's nearest-hit kernel of the last trc_trace that asked for it
// (TRC_TRACE_TIME_WALK); waits for that kernel.  The event pair belongs to the scene and is reused.
extern "C" int trc_trace_walk_ms(const TrcScene *sc, float *ms) {
    if (!sc || !ms || !sc->ev_walk[0]) return fail(TRC_ERR_BAD_ARGUMENT, "trc_trace_walk_ms: no timed trace on this scene%s");
    TRC_CUDA(cudaEventSynchronize(sc->ev_walk[1]));
    TRC_CUDA(cudaEventElapsedTime(ms, sc->ev_walk[0], sc->ev_walk[1]));
    return TRC_OK;
}
