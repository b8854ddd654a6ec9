// trc_bounce: one whole iteration of the reference loop body (tracer_engine.py:193-274)
// sequenced on the host side in C++ -- nearest hit + hit records, hit count read-back,
// child order, optics + accountant columns, per-surface tallies, cull count read-back and
// (when some but not all children are weak) the survivors-first reorder -- so that the
// Python engine pays one foreign call and two arena allocations per bounce instead of
// ~12 calls and ~25 allocations.  Same kernels and the same results as the step-by-step
// entry points; used for Philox draws and scenes without two-children optics.
#include <new>

#include "common.cuh"

namespace {

inline int64_t up256(int64_t x) { return (x + 255) & ~(int64_t)255; }

struct KeepLayout {
    int64_t off_v, off_d, off_e, off_p, off_w;            // slot-ordered children (pre-cull order)
    int64_t off_sv, off_sd, off_se, off_sp, off_sw;       // survivors-first copies (only when reordered)
    int64_t off_abs, off_nrm, off_ein, off_din, off_att;  // accountant columns, slot order
    int64_t off_seg;                                      // int64 (n_surfaces + 1)
    int64_t total;
};

KeepLayout keep_layout(int64_t h, uint32_t columns, int has_w, int32_t n_surfaces) {
    KeepLayout k;
    int64_t off = 0;
    auto take = [&](int64_t bytes) { int64_t o = off; off = up256(off + bytes); return o; };
    k.off_v = take(24 * h); k.off_d = take(24 * h); k.off_e = take(8 * h); k.off_p = take(8 * h);
    k.off_w = has_w ? take(8 * h) : -1;
    k.off_sv = take(24 * h); k.off_sd = take(24 * h); k.off_se = take(8 * h); k.off_sp = take(8 * h);
    k.off_sw = has_w ? take(8 * h) : -1;
    k.off_abs = take(8 * h);
    k.off_nrm = (columns & TRC_COL_NORMALS) ? take(24 * h) : -1;
    k.off_ein = (columns & TRC_COL_E_IN) ? take(8 * h) : -1;
    k.off_din = (columns & TRC_COL_DIRS_IN) ? take(24 * h) : -1;
    k.off_att = (columns & TRC_COL_ATTENUATED) ? take(8 * h) : -1;
    k.off_seg = take(8 * ((int64_t)n_surfaces + 1));
    k.total = off;
    return k;
}

struct ScratchLayout {
    int64_t off_records, off_chunk_off, off_chunk_cnt, off_counters, off_slot_surf, off_slot_rec;
    int64_t off_cull, off_k2, off_perm, off_order_work, off_sort_work, total;
};

ScratchLayout scratch_layout(int64_t n, int64_t cap, int32_t n_surfaces) {
    ScratchLayout s;
    int64_t off = 0;
    auto take = [&](int64_t bytes) { int64_t o = off; off = up256(off + bytes); return o; };
    const int64_t n_chunks = (n + TRC_HIT_CHUNK - 1) / TRC_HIT_CHUNK;
    s.off_records = take(64 * cap);
    s.off_chunk_off = take(4 * n_chunks);
    s.off_chunk_cnt = take(4 * n_chunks);
    s.off_counters = take(32);
    s.off_slot_surf = take(4 * cap);
    s.off_slot_rec = take(4 * cap);
    s.off_cull = take(4 * cap);
    s.off_k2 = take(4 * cap);
    s.off_perm = take(4 * cap);
    s.off_order_work = take(trc_order_work_bytes(n_chunks, cap, n_surfaces));
    s.off_sort_work = take(trc_sort_work_bytes(cap > 0 ? cap : 1, 2));
    s.total = off;
    return s;
}

__global__ void add_counts_kernel(const int64_t *__restrict__ seg, int n_surfaces, int64_t *__restrict__ hits) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_surfaces) hits[s] += seg[s + 1] - seg[s];
}

// Survivors-first reorder of a whole child bundle in one pass: vertices, directions, energy and
// parents of child perm[j] move to position j (the three trc_gather_rows calls of the step-by-step
// path read the permutation three times and cost three launches).
__global__ void __launch_bounds__(256)
gather_bundle_kernel(int64_t h, const int32_t *__restrict__ perm, const double *__restrict__ v,
                     const double *__restrict__ d, const double *__restrict__ e, const int64_t *__restrict__ par,
                     double *__restrict__ ov, double *__restrict__ od, double *__restrict__ oe,
                     int64_t *__restrict__ opar) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= h) return;
    int64_t src = perm[j];
    double v0 = v[src], v1 = v[h + src], v2 = v[2 * h + src];
    double d0 = d[src], d1 = d[h + src], d2 = d[2 * h + src];
    double en = e[src];
    int64_t pp = par[src];
    ov[j] = v0; ov[h + j] = v1; ov[2 * h + j] = v2;
    od[j] = d0; od[h + j] = d1; od[2 * h + j] = d2;
    oe[j] = en; opar[j] = pp;
}

}  // namespace

extern "C" int64_t trc_bounce_keep_bytes(int64_t capacity, uint32_t columns, int32_t has_wavelengths,
                                         int32_t n_surfaces) {
    if (capacity < 0 || n_surfaces < 1) return -1;
    return keep_layout(capacity, columns, has_wavelengths, n_surfaces).total;
}

extern "C" int64_t trc_bounce_scratch_bytes(int64_t n_rays, int64_t capacity, int32_t n_surfaces) {
    if (n_rays < 0 || capacity < 0 || n_surfaces < 1) return -1;
    return scratch_layout(n_rays, capacity, n_surfaces).total;
}

extern "C" int trc_bounce(const TrcScene *sc, const TrcBounceArgs *a, TrcBounceResult *r, void *stream) {
    if (!sc || !a || !r || a->n < 0 || a->capacity < 0 || !a->d_keep || !a->d_scratch ||
        !a->d_absorbed_per_surface || !a->d_hits_per_surface)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_bounce: bad argument%s");
    if (sc->has_split)
        return fail(TRC_ERR_UNSUPPORTED_KIND,
                    "trc_bounce: scenes with two-children optics (periodic, refractive) use the step-by-step calls%s");
    cudaStream_t st = (cudaStream_t)stream;
    const int32_t S = sc->n_surfaces;
    const int64_t n = a->n, cap = a->capacity;
    const int has_w = a->d_wavelengths ? 1 : 0;
    memset(r, 0, sizeof(*r));
    if (a->scratch_bytes < scratch_layout(n, cap, S).total || a->keep_bytes < keep_layout(cap, a->columns, has_w, S).total)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_bounce: arena smaller than trc_bounce_*_bytes(capacity)%s");
    char *scr = (char *)a->d_scratch, *keep = (char *)a->d_keep;
    ScratchLayout sl = scratch_layout(n, cap, S);
    TrcHitRecord *records = (TrcHitRecord *)(scr + sl.off_records);
    uint32_t *chunk_off = (uint32_t *)(scr + sl.off_chunk_off), *chunk_cnt = (uint32_t *)(scr + sl.off_chunk_cnt);
    int64_t *counters = (int64_t *)(scr + sl.off_counters);          // [hits, culled, sort count]
    const int64_t n_chunks = (n + TRC_HIT_CHUNK - 1) / TRC_HIT_CHUNK;

    TRC_CUDA(cudaMemsetAsync(counters, 0, 32, st));
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (a->time_trace) {
        TRC_CUDA(cudaEventCreate(&ev0)); TRC_CUDA(cudaEventCreate(&ev1));
        TRC_CUDA(cudaEventRecord(ev0, st));
    }
    int rc = trc_trace_hits(sc, n, a->d_vertices, a->d_directions, a->d_energy, a->ld, a->use_bvh, nullptr, records,
                            cap, chunk_off, chunk_cnt, counters, stream);
    if (rc != TRC_OK) return rc;
    if (a->time_trace) {
        TRC_CUDA(cudaEventRecord(ev1, st));
        r->ev_start = ev0; r->ev_stop = ev1;
    }
    // counts come back through page-locked words of the scene (a pageable destination makes the driver
    // stage the copy: ~60 us of idle GPU per read-back on the timeline)
    int64_t *h_counts = sc->h_pinned;
    h_counts[0] = h_counts[1] = 0;
    TRC_CUDA(cudaMemcpyAsync(&h_counts[0], counters, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    TRC_CUDA(cudaStreamSynchronize(st));
    const int64_t H = h_counts[0];
    r->hits = H;
    if (H > cap) return fail(TRC_ERR_CAPACITY, "trc_bounce: more hits than the arenas were sized for%s");
    KeepLayout kl = keep_layout(H, a->columns, has_w, S);
    r->ld = H;
    r->off_seg = kl.off_seg;
    int64_t *seg = (int64_t *)(keep + kl.off_seg);
    if (H == 0) {
        TRC_CUDA(cudaMemsetAsync(seg, 0, sizeof(int64_t) * (size_t)(S + 1), st));
        if (a->h_seg) memset(a->h_seg, 0, sizeof(int64_t) * (size_t)(S + 1));
        return TRC_OK;
    }
    int32_t *slot_surf = (int32_t *)(scr + sl.off_slot_surf), *slot_rec = (int32_t *)(scr + sl.off_slot_rec);
    rc = trc_order_hits(n_chunks, chunk_off, chunk_cnt, records, H, S, slot_surf, slot_rec, seg,
                        scr + sl.off_order_work, stream);
    if (rc != TRC_OK) return rc;

    TrcOpticsArgs o;
    memset(&o, 0, sizeof(o));
    o.h = H; o.d_records = records; o.d_slot_rec = slot_rec; o.d_wavelengths = a->d_wavelengths;
    o.d_out_vertices = (double *)(keep + kl.off_v); o.d_out_directions = (double *)(keep + kl.off_d);
    o.d_out_energy = (double *)(keep + kl.off_e); o.d_out_parents = (int64_t *)(keep + kl.off_p);
    o.ld_out = H;
    o.seed = a->seed; o.bounce = a->bounce; o.min_energy = a->min_energy;
    o.d_absorbed = (double *)(keep + kl.off_abs);
    o.d_normals = kl.off_nrm >= 0 ? (double *)(keep + kl.off_nrm) : nullptr;
    o.d_e_in = kl.off_ein >= 0 ? (double *)(keep + kl.off_ein) : nullptr;
    o.d_dirs_in = kl.off_din >= 0 ? (double *)(keep + kl.off_din) : nullptr;
    o.d_attenuated = kl.off_att >= 0 ? (double *)(keep + kl.off_att) : nullptr;
    o.d_out_wavelengths = has_w ? (double *)(keep + kl.off_w) : nullptr;
    int32_t *cull = (int32_t *)(scr + sl.off_cull);
    o.d_cull = cull;
    o.d_n_cull = counters + 1;
    rc = trc_optics(sc, &o, stream);
    if (rc != TRC_OK) return rc;
    rc = trc_segment_sums(H, slot_surf, o.d_absorbed, a->d_absorbed_per_surface, stream);
    if (rc != TRC_OK) return rc;
    add_counts_kernel<<<(S + 255) / 256, 256, 0, st>>>(seg, S, a->d_hits_per_surface);
    rc = check_launch("add_counts");
    if (rc != TRC_OK) return rc;

    TRC_CUDA(cudaMemcpyAsync(&h_counts[1], counters + 1, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    if (a->h_seg)
        TRC_CUDA(cudaMemcpyAsync(h_counts + 2, seg, sizeof(int64_t) * (size_t)(S + 1), cudaMemcpyDeviceToHost, st));
    TRC_CUDA(cudaStreamSynchronize(st));
    if (a->h_seg) memcpy(a->h_seg, h_counts + 2, sizeof(int64_t) * (size_t)(S + 1));
    const int64_t n_cull = h_counts[1];
    r->culled = n_cull;
    r->off_vertices = kl.off_v; r->off_directions = kl.off_d; r->off_energy = kl.off_e; r->off_parents = kl.off_p;
    r->off_wavelengths = kl.off_w;
    r->off_absorbed = kl.off_abs; r->off_normals = kl.off_nrm; r->off_e_in = kl.off_ein;
    r->off_dirs_in = kl.off_din; r->off_attenuated = kl.off_att;
    r->off_sorted_vertices = r->off_sorted_directions = r->off_sorted_energy = r->off_sorted_parents = -1;
    r->off_sorted_wavelengths = -1;
    if (n_cull > 0 && n_cull < H) {
        // survivors first, weak rays to the back (tracer_engine.py:242-245, 268-274): stable sort on the flag
        int32_t *k2 = (int32_t *)(scr + sl.off_k2), *perm = (int32_t *)(scr + sl.off_perm);
        rc = trc_sort_pairs(H, 2, cull, nullptr, k2, perm, counters + 2, scr + sl.off_sort_work, stream);
        if (rc != TRC_OK) return rc;
        gather_bundle_kernel<<<(unsigned)((H + 255) / 256), 256, 0, st>>>(
            H, perm, o.d_out_vertices, o.d_out_directions, o.d_out_energy, o.d_out_parents,
            (double *)(keep + kl.off_sv), (double *)(keep + kl.off_sd), (double *)(keep + kl.off_se),
            (int64_t *)(keep + kl.off_sp));
        rc = check_launch("gather_bundle");
        if (rc != TRC_OK) return rc;
        if (has_w) {
            rc = trc_gather_rows(H, perm, 1, o.d_out_wavelengths, H, (double *)(keep + kl.off_sw), H, nullptr, nullptr,
                                 stream);
            if (rc != TRC_OK) return rc;
        }
        r->off_sorted_vertices = kl.off_sv; r->off_sorted_directions = kl.off_sd; r->off_sorted_energy = kl.off_se;
        r->off_sorted_parents = kl.off_sp; r->off_sorted_wavelengths = kl.off_sw;
    }
    return TRC_OK;
}

extern "C" int trc_event_elapsed_ms(void *ev_start, void *ev_stop, float *ms) {
    if (!ev_start || !ev_stop || !ms) return fail(TRC_ERR_BAD_ARGUMENT, "trc_event_elapsed_ms: bad argument%s");
    TRC_CUDA(cudaEventSynchronize((cudaEvent_t)ev_stop));
    TRC_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)ev_start, (cudaEvent_t)ev_stop));
    cudaEventDestroy((cudaEvent_t)ev_start);
    cudaEventDestroy((cudaEvent_t)ev_stop);
    return TRC_OK;
}
