// Stable LSD radix sort of (key, value) int32 pairs, dropping negative keys.
//
// This is the compaction/ordering step of the bounce loop.  The reference
// builds the child bundle by visiting surfaces in ascending index and, inside
// each surface, hit rays in ascending parent index, then hstack-ing the pieces
// (tracer_engine.py:218-266, ray_bundle.py:197-223).  That order is exactly a
// stable sort of the hit rays by surface index; "survivors first, weak rays at
// the back" (tracer_engine.py:268-274) is a stable sort by the 1-bit cull flag.
// Stability comes from warp ballots (__match_any_sync) inside a tile and an
// exclusive scan of the (digit, tile) histogram across tiles -- no atomics
// decide an output position.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 8;                       // rounds of 32 per warp
constexpr int kTile = kThreads * kItems;        // 2048 pairs per block
constexpr int kMaxBins = 256;
constexpr int kScanPerBlock = 4096;

__device__ __forceinline__ int64_t load_count(const int64_t *p, int64_t fallback) { return p ? *p : fallback; }

// One tile: per-digit counts (hist mode) or stable scatter (scatter mode).
template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
radix_tile_kernel(const int32_t *__restrict__ keys_in, const int32_t *__restrict__ vals_in,
                  int64_t n_static, const int64_t *__restrict__ n_dynamic,
                  int shift, int nbins, bool drop_negative, int64_t n_tiles,
                  uint32_t *__restrict__ hist /* [nbins][n_tiles] */,
                  int32_t *__restrict__ keys_out, int32_t *__restrict__ vals_out) {
    __shared__ uint32_t wcnt[kWarps][kMaxBins];
    __shared__ uint32_t base[kMaxBins];
    const int64_t n = load_count(n_dynamic, n_static);
    const int64_t tile = blockIdx.x;
    const int64_t tile_start = tile * kTile;
    if (tile_start >= n) {
        if (!kScatter) for (int b = threadIdx.x; b < nbins; b += kThreads) hist[(int64_t)b * n_tiles + tile] = 0;
        return;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int b = threadIdx.x; b < kWarps * kMaxBins; b += kThreads) (&wcnt[0][0])[b] = 0;
    __syncthreads();

    const uint32_t mask = (uint32_t)nbins - 1u;
    int32_t key[kItems], val[kItems];
    int digit[kItems];
    uint32_t rank[kItems];
    const int64_t warp_start = tile_start + (int64_t)warp * (32 * kItems);
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        int64_t i = warp_start + r * 32 + lane;
        bool valid = i < n;
        int32_t k = valid ? keys_in[i] : -1;
        if (valid && !drop_negative && k < 0) k = 0;   // later passes never see negatives
        valid = valid && k >= 0;
        key[r] = k;
        if (kScatter) val[r] = valid ? (vals_in ? vals_in[i] : (int32_t)i) : 0;
        int dg = valid ? (int)(((uint32_t)k >> shift) & mask) : -1;
        digit[r] = dg;
        unsigned peers = __match_any_sync(0xffffffffu, dg);
        uint32_t before = __popc(peers & ((1u << lane) - 1u));
        uint32_t running = 0;
        if (dg >= 0) running = wcnt[warp][dg];
        __syncwarp();
        if (dg >= 0 && before == 0) wcnt[warp][dg] = running + __popc(peers);
        __syncwarp();
        rank[r] = running + before;
    }
    __syncthreads();
    // per-digit exclusive prefix over warps; tile total per digit
    for (int b = threadIdx.x; b < nbins; b += kThreads) {
        uint32_t acc = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { uint32_t c = wcnt[w][b]; wcnt[w][b] = acc; acc += c; }
        if (kScatter) base[b] = hist[(int64_t)b * n_tiles + tile];
        else hist[(int64_t)b * n_tiles + tile] = acc;
    }
    if (!kScatter) return;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        int dg = digit[r];
        if (dg < 0) continue;
        int64_t pos = (int64_t)base[dg] + wcnt[warp][dg] + rank[r];
        keys_out[pos] = key[r];
        vals_out[pos] = val[r];
    }
}

// ---- exclusive scan of a uint32 array (three small kernels) ----
__global__ void __launch_bounds__(kThreads)
scan_reduce_kernel(const uint32_t *__restrict__ in, int64_t n, uint32_t *__restrict__ block_sums) {
    __shared__ uint32_t red[kWarps];
    int64_t start = (int64_t)blockIdx.x * kScanPerBlock;
    uint32_t acc = 0;
    for (int k = threadIdx.x; k < kScanPerBlock; k += kThreads) {
        int64_t i = start + k;
        if (i < n) acc += in[i];
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < kWarps; ++w) t += red[w];
        block_sums[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(kThreads)
scan_sums_kernel(uint32_t *__restrict__ block_sums, int64_t n_blocks, int64_t *__restrict__ total_out) {
    __shared__ uint32_t sh[kThreads];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t start = 0; start < n_blocks; start += kThreads) {
        int64_t i = start + threadIdx.x;
        uint32_t v = i < n_blocks ? block_sums[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < kThreads; o <<= 1) {
            uint32_t add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += add;
            __syncthreads();
        }
        uint32_t incl = sh[threadIdx.x];
        if (i < n_blocks) block_sums[i] = carry + incl - v;
        __syncthreads();
        if (threadIdx.x == kThreads - 1) carry += incl;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total_out) *total_out = (int64_t)carry;
}

__global__ void __launch_bounds__(kThreads)
scan_apply_kernel(uint32_t *__restrict__ data, int64_t n, const uint32_t *__restrict__ block_sums) {
    __shared__ uint32_t sh[kThreads];
    constexpr int per = kScanPerBlock / kThreads;   // 16 consecutive per thread
    int64_t start = (int64_t)blockIdx.x * kScanPerBlock + (int64_t)threadIdx.x * per;
    uint32_t v[per];
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < per; ++k) {
        int64_t i = start + k;
        v[k] = i < n ? data[i] : 0;
        acc += v[k];
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 1; o < kThreads; o <<= 1) {
        uint32_t add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
        __syncthreads();
        sh[threadIdx.x] += add;
        __syncthreads();
    }
    uint32_t run = block_sums[blockIdx.x] + sh[threadIdx.x] - acc;
#pragma unroll
    for (int k = 0; k < per; ++k) {
        int64_t i = start + k;
        if (i < n) data[i] = run;
        run += v[k];
    }
}

int key_bits(int32_t n_keys) {
    int bits = 1;
    while (bits < 31 && (1LL << bits) < (long long)n_keys) ++bits;
    return bits;
}

struct WorkLayout {
    int64_t n_tiles, hist_elems, n_scan_blocks;
    size_t off_hist, off_sums, off_k0, off_v0, off_k1, off_v1, total;
};

WorkLayout layout(int64_t n, int32_t n_keys) {
    WorkLayout w;
    w.n_tiles = (n + kTile - 1) / kTile;
    if (w.n_tiles < 1) w.n_tiles = 1;
    int bits = key_bits(n_keys);
    int nb = 1 << (bits < 8 ? bits : 8);
    w.hist_elems = (int64_t)nb * w.n_tiles;
    w.n_scan_blocks = (w.hist_elems + kScanPerBlock - 1) / kScanPerBlock;
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t off = 0;
    w.off_hist = off; off = align(off + sizeof(uint32_t) * (size_t)w.hist_elems);
    w.off_sums = off; off = align(off + sizeof(uint32_t) * (size_t)w.n_scan_blocks);
    bool multi = bits > 8;
    size_t pair = multi ? sizeof(int32_t) * (size_t)n : 0;
    w.off_k0 = off; off = align(off + pair);
    w.off_v0 = off; off = align(off + pair);
    w.off_k1 = off; off = align(off + pair);
    w.off_v1 = off; off = align(off + pair);
    w.total = off + 256;
    return w;
}

}  // namespace

extern "C" int64_t trc_sort_work_bytes(int64_t n, int32_t n_keys) {
    if (n < 0 || n_keys < 1) return -1;
    return (int64_t)layout(n, n_keys).total;
}

extern "C" int trc_sort_pairs(int64_t n, int32_t n_keys, const int32_t *d_keys_in, const int32_t *d_vals_in,
                              int32_t *d_keys_out, int32_t *d_vals_out, int64_t *d_count, void *d_work,
                              void *stream) {
    if (n < 0 || n_keys < 1 || !d_count || (n > 0 && (!d_keys_in || !d_keys_out || !d_vals_out || !d_work)))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_sort_pairs: bad argument%s");
    if (n >= 0x7fffffffLL) return fail(TRC_ERR_CAPACITY, "trc_sort_pairs: more than 2^31-1 pairs, chunk the bundle%s");
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        TRC_CUDA(cudaMemsetAsync(d_count, 0, sizeof(int64_t), st));
        return TRC_OK;
    }
    WorkLayout w = layout(n, n_keys);
    char *wk = (char *)d_work;
    uint32_t *hist = (uint32_t *)(wk + w.off_hist);
    uint32_t *sums = (uint32_t *)(wk + w.off_sums);
    int32_t *tk[2] = {(int32_t *)(wk + w.off_k0), (int32_t *)(wk + w.off_k1)};
    int32_t *tv[2] = {(int32_t *)(wk + w.off_v0), (int32_t *)(wk + w.off_v1)};
    int bits = key_bits(n_keys);
    int passes = (bits + 7) / 8;
    const int32_t *ki = d_keys_in, *vi = d_vals_in;
    for (int p = 0; p < passes; ++p) {
        int shift = 8 * p;
        int pb = bits - shift < 8 ? bits - shift : 8;
        int nbins = 1 << pb;
        bool last = p == passes - 1;
        int32_t *ko = last ? d_keys_out : tk[p & 1];
        int32_t *vo = last ? d_vals_out : tv[p & 1];
        const int64_t *dyn = p == 0 ? nullptr : d_count;
        int64_t hist_elems = (int64_t)nbins * w.n_tiles;
        int64_t scan_blocks = (hist_elems + kScanPerBlock - 1) / kScanPerBlock;
        radix_tile_kernel<false><<<(unsigned)w.n_tiles, kThreads, 0, st>>>(ki, vi, n, dyn, shift, nbins, p == 0,
                                                                           w.n_tiles, hist, nullptr, nullptr);
        scan_reduce_kernel<<<(unsigned)scan_blocks, kThreads, 0, st>>>(hist, hist_elems, sums);
        scan_sums_kernel<<<1, kThreads, 0, st>>>(sums, scan_blocks, p == 0 ? d_count : nullptr);
        scan_apply_kernel<<<(unsigned)scan_blocks, kThreads, 0, st>>>(hist, hist_elems, sums);
        radix_tile_kernel<true><<<(unsigned)w.n_tiles, kThreads, 0, st>>>(ki, vi, n, dyn, shift, nbins, p == 0,
                                                                          w.n_tiles, hist, ko, vo);
        ki = ko; vi = vo;
    }
    return check_launch("sort_pairs", 5 * passes);
}

// Passes first_shift, first_shift + 8, ... of the same LSD radix sort over *d_n (<= n_cap) pairs whose lower
// digits are already sorted (the device-driven trace sorts the low digit itself, trace.cu).  Ping-pongs between
// (k_a, v_a) -- the input -- and (k_b, v_b); *in_a tells where the result ended up.
int trc_sort_tail_passes(int64_t n_cap, const int64_t *d_n, int32_t n_keys, int first_shift, int32_t *k_a,
                         int32_t *v_a, int32_t *k_b, int32_t *v_b, void *d_work, cudaStream_t st, bool *in_a) {
    WorkLayout w = layout(n_cap, n_keys);
    char *wk = (char *)d_work;
    uint32_t *hist = (uint32_t *)(wk + w.off_hist);
    uint32_t *sums = (uint32_t *)(wk + w.off_sums);
    const int bits = key_bits(n_keys);
    int32_t *ki = k_a, *vi = v_a, *ko = k_b, *vo = v_b;
    bool a_holds = true;
    int launches = 0;
    for (int shift = first_shift; shift < bits; shift += 8) {
        int pb = bits - shift < 8 ? bits - shift : 8;
        int nbins = 1 << pb;
        int64_t hist_elems = (int64_t)nbins * w.n_tiles;
        int64_t scan_blocks = (hist_elems + kScanPerBlock - 1) / kScanPerBlock;
        radix_tile_kernel<false><<<(unsigned)w.n_tiles, kThreads, 0, st>>>(ki, vi, n_cap, d_n, shift, nbins, false,
                                                                           w.n_tiles, hist, nullptr, nullptr);
        scan_reduce_kernel<<<(unsigned)scan_blocks, kThreads, 0, st>>>(hist, hist_elems, sums);
        scan_sums_kernel<<<1, kThreads, 0, st>>>(sums, scan_blocks, nullptr);
        scan_apply_kernel<<<(unsigned)scan_blocks, kThreads, 0, st>>>(hist, hist_elems, sums);
        radix_tile_kernel<true><<<(unsigned)w.n_tiles, kThreads, 0, st>>>(ki, vi, n_cap, d_n, shift, nbins, false,
                                                                          w.n_tiles, hist, ko, vo);
        int32_t *t;
        t = ki; ki = ko; ko = t;
        t = vi; vi = vo; vo = t;
        a_holds = !a_holds;
        launches += 5;
    }
    *in_a = a_holds;
    return launches ? check_launch("sort_tail_passes", launches) : TRC_OK;
}

// ------------------------------------------------------- segment offsets
namespace {
__global__ void segment_offsets_kernel(const int64_t *count, int32_t n_keys, const int32_t *__restrict__ keys,
                                       int64_t *__restrict__ off) {
    int64_t n = *count;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q <= n; q += stride) {
        int32_t prev = q == 0 ? -1 : keys[q - 1];
        int32_t cur = q == n ? n_keys : keys[q];
        for (int32_t k = prev + 1; k <= cur; ++k) off[k] = q;
    }
}
}  // namespace

// d_sorted_keys holds *d_count valid entries (d_count from trc_sort_pairs).
extern "C" int trc_segment_offsets_dyn(const int64_t *d_count, int64_t n_max, int32_t n_keys,
                                       const int32_t *d_sorted_keys, int64_t *d_offsets, void *stream) {
    if (!d_count || n_keys < 1 || !d_offsets || n_max < 0)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_segment_offsets: bad argument%s");
    int64_t blocks = (n_max + 1 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    segment_offsets_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_count, n_keys, d_sorted_keys, d_offsets);
    return check_launch("segment_offsets");
}

// ------------------------------------------------------- order the hit list
// trc_trace_hits appends chunks of hits in arbitrary order; the reference
// order (surface-major, parent-ray-minor) is recovered by (1) an exclusive scan
// of the per-chunk counts in chunk (= ray) order, (2) writing (surface, record
// index) pairs at those positions, (3) one stable radix sort by surface.
namespace {

__global__ void __launch_bounds__(kThreads)
copy_u32_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i];
}

// one warp per chunk
__global__ void __launch_bounds__(kThreads)
hit_pairs_kernel(int64_t n_chunks, const uint32_t *__restrict__ chunk_off, const uint32_t *__restrict__ chunk_cnt,
                 const uint32_t *__restrict__ ord_off, const TrcHitRecord *__restrict__ records, int64_t h,
                 int32_t *__restrict__ keys, int32_t *__restrict__ vals) {
    int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (c >= n_chunks) return;
    uint32_t cnt = chunk_cnt[c], src = chunk_off[c], dst = ord_off[c];
    for (uint32_t i = lane; i < cnt; i += 32) {
        int64_t q = (int64_t)dst + i;
        if (q < h) {
            keys[q] = records[src + i].surf;
            vals[q] = (int32_t)(src + i);
        }
    }
}

struct OrderLayout {
    int64_t scan_blocks;
    size_t off_ord, off_sums, off_keys, off_vals, off_sort, total;
};

OrderLayout order_layout(int64_t n_chunks, int64_t h, int32_t n_keys) {
    OrderLayout w;
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    w.scan_blocks = (n_chunks + kScanPerBlock - 1) / kScanPerBlock;
    if (w.scan_blocks < 1) w.scan_blocks = 1;
    size_t off = 0;
    w.off_ord = off; off = align(off + sizeof(uint32_t) * (size_t)(n_chunks + 1));
    w.off_sums = off; off = align(off + sizeof(uint32_t) * (size_t)w.scan_blocks);
    w.off_keys = off; off = align(off + sizeof(int32_t) * (size_t)h);
    w.off_vals = off; off = align(off + sizeof(int32_t) * (size_t)h);
    w.off_sort = off; off = align(off + (size_t)layout(h > 0 ? h : 1, n_keys).total);
    w.total = off + 256;
    return w;
}

}  // namespace

extern "C" int64_t trc_order_work_bytes(int64_t n_chunks, int64_t h, int32_t n_keys) {
    if (n_chunks < 0 || h < 0 || n_keys < 1) return -1;
    return (int64_t)order_layout(n_chunks, h, n_keys).total;
}

extern "C" int trc_order_hits(int64_t n_chunks, const uint32_t *d_chunk_off, const uint32_t *d_chunk_cnt,
                              const TrcHitRecord *d_records, int64_t h, int32_t n_keys, int32_t *d_slot_surf,
                              int32_t *d_slot_rec, int64_t *d_offsets, void *d_work, void *stream) {
    if (n_chunks < 0 || h < 0 || n_keys < 1 || !d_offsets || !d_work ||
        (h > 0 && (!d_chunk_off || !d_chunk_cnt || !d_records || !d_slot_surf || !d_slot_rec || n_chunks < 1)))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_order_hits: bad argument%s");
    if (h >= 0x7fffffffLL) return fail(TRC_ERR_CAPACITY, "trc_order_hits: more than 2^31-1 hits%s");
    cudaStream_t st = (cudaStream_t)stream;
    OrderLayout w = order_layout(n_chunks, h, n_keys);
    char *wk = (char *)d_work;
    uint32_t *ord = (uint32_t *)(wk + w.off_ord);
    uint32_t *sums = (uint32_t *)(wk + w.off_sums);
    int32_t *keys = (int32_t *)(wk + w.off_keys);
    int32_t *vals = (int32_t *)(wk + w.off_vals);
    int64_t *count = (int64_t *)(wk + w.total - 256);
    if (h == 0) {
        TRC_CUDA(cudaMemsetAsync(d_offsets, 0, sizeof(int64_t) * (size_t)(n_keys + 1), st));
        return TRC_OK;
    }
    copy_u32_kernel<<<(unsigned)((n_chunks + kThreads - 1) / kThreads), kThreads, 0, st>>>(d_chunk_cnt, ord, n_chunks);
    scan_reduce_kernel<<<(unsigned)w.scan_blocks, kThreads, 0, st>>>(ord, n_chunks, sums);
    scan_sums_kernel<<<1, kThreads, 0, st>>>(sums, w.scan_blocks, nullptr);
    scan_apply_kernel<<<(unsigned)w.scan_blocks, kThreads, 0, st>>>(ord, n_chunks, sums);
    int64_t warps_per_block = kThreads / 32;
    hit_pairs_kernel<<<(unsigned)((n_chunks + warps_per_block - 1) / warps_per_block), kThreads, 0, st>>>(
        n_chunks, d_chunk_off, d_chunk_cnt, ord, d_records, h, keys, vals);
    int rc = check_launch("order_hits", 5);
    if (rc != TRC_OK) return rc;
    rc = trc_sort_pairs(h, n_keys, keys, vals, d_slot_surf, d_slot_rec, count, wk + w.off_sort, stream);
    if (rc != TRC_OK) return rc;
    return trc_segment_offsets_dyn(count, h, n_keys, d_slot_surf, d_offsets, stream);
}

// ------------------------------------------------------- allocation-free variant
// trc_sort_pairs for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the work area is a file-scope device array instead of a caller-provided buffer.  No counterpart in
// the reference.  Limits: n <= TRC_SORT_STATIC_MAX_PAIRS and
// n_keys <= TRC_SORT_STATIC_MAX_KEYS (one radix pass, so no ping-pong pair arrays are needed).  Every GPU that
// loads the library holds the array (about 2 MB) until exit.  All calls share it: calls that overlap on the GPU
// (different streams, or different host threads) race.  Nothing in it has to start at zero -- the histogram
// kernel writes every (bin, tile) cell, empty tiles included, and the scan kernels write every block sum before
// it is read -- so the contents a previous call left behind are harmless.
namespace {
constexpr size_t kStaticScratchBytes = 2u * 1024u * 1024u + 4096u;
constexpr int kStaticMaxDevices = 64;
}  // namespace
static __device__ __align__(256) unsigned char g_sort_static_scratch[kStaticScratchBytes];
static void *g_sort_static_ptr[kStaticMaxDevices];      // device address of the array, per device (host code must
                                                         // not pass the symbol itself: it is a host-side stand-in)

extern "C" int trc_sort_pairs_static(int64_t n, int32_t n_keys, const int32_t *d_keys_in, const int32_t *d_vals_in,
                                     int32_t *d_keys_out, int32_t *d_vals_out, int64_t *d_count, void *stream) {
    if (n < 0 || n_keys < 1) return fail(TRC_ERR_BAD_ARGUMENT, "trc_sort_pairs_static: bad argument%s");
    if (n > TRC_SORT_STATIC_MAX_PAIRS)
        return fail(TRC_ERR_CAPACITY, "trc_sort_pairs_static: more than TRC_SORT_STATIC_MAX_PAIRS (4194304) pairs; use trc_sort_pairs%s");
    if (n_keys > TRC_SORT_STATIC_MAX_KEYS)
        return fail(TRC_ERR_CAPACITY, "trc_sort_pairs_static: more than TRC_SORT_STATIC_MAX_KEYS (256) keys; use trc_sort_pairs%s");
    if (layout(n > 0 ? n : 1, n_keys).total > kStaticScratchBytes)
        return fail(TRC_ERR_CAPACITY, "trc_sort_pairs_static: work area larger than the static scratch%s");
    int dev = 0;
    TRC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kStaticMaxDevices) return fail(TRC_ERR_BAD_ARGUMENT, "trc_sort_pairs_static: device index out of range%s");
    if (!g_sort_static_ptr[dev]) {          // set-up, once per device, with that device current
        void *ptr = nullptr;
        TRC_CUDA(cudaGetSymbolAddress(&ptr, g_sort_static_scratch));
        g_sort_static_ptr[dev] = ptr;
    }
    return trc_sort_pairs(n, n_keys, d_keys_in, d_vals_in, d_keys_out, d_vals_out, d_count, g_sort_static_ptr[dev], stream);
}
