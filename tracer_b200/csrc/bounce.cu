// Optics + accounting kernel, row gather, per-surface sums, flux-map binning.
#include "common.cuh"
#include "geom.cuh"
#include "optics.cuh"
#include "philox.cuh"

using namespace trc;

namespace {

struct OpticsParams {
    int64_t h;
    const TrcHitRecord *records;
    const int32_t *slot_rec;
    const double *wavelengths;
    double *o_vert, *o_dir, *o_energy;
    int64_t *o_parents;
    int64_t ld_out;
    const double *draws[3];
    uint64_t seed;
    uint32_t bounce;
    double min_energy;
    double *absorbed;
    double *normals;
    double *e_in;
    double *dirs_in;
    double *o_wavelengths;
    int32_t *cull;
    unsigned long long *n_cull;
    double *attenuated;
    const int32_t *slot_surf;
    const int64_t *seg, *child_off;
    const double *ref_index;
    double *o_ref_index;
    unsigned long long *n_void;
    // source-ray identity (Philox key): id of parent ray r = ids_in ? ids_in[r] : id_base + r; children inherit it
    const int64_t *ids_in;
    int64_t id_base;
    int64_t *o_ids;
    const double *ref_index_im;    // imaginary parts of ref_index (scenes with tabulated refractive media), nullable
    double *o_ref_index_im;
    // device-driven trace (trace.cu): hit count and bookkeeping live in device memory
    TrcBounceCtl *ctl;            // nullable; when set h = min(ctl->hits, capacity)
    int64_t capacity;
    uint32_t *cta_surv;           // survivors per CTA -> exclusive scan (last block), read by the partition kernel
    double *absorbed_per_surface;
};

// Philox draws in the units the reference's numpy calls produce
// (normal(scale=sigma), uniform(0, 2pi), uniform(0, 1)) for one slot.
__device__ __forceinline__ void philox_draws(const TrcSurfaceRec *s, uint64_t seed, uint32_t bounce, uint64_t q,
                                             double *u) {
    const double two_pi = 2.0 * M_PI;
    switch (s->okind) {
    case TRC_O_REAL_REFLECTIVE: {
        if (!(s->op[1] > 0.0)) break;
        double a, b;
        normal2(seed, q, bounce, 0, a, b);
        if (s->op[2] != 0.0) { u[0] = a * s->op[1]; u[1] = b * s->op[1]; }
        else {
            double c, e;
            uniform2(seed, q, bounce, 1, c, e);
            u[0] = a * s->op[1]; u[1] = two_pi * c;
        }
        break;
    }
    case TRC_O_LAMBERTIAN: case TRC_O_LAMB_DIRECTIONAL: case TRC_O_LAMB_DIRECTIONAL_SPECTRAL:
    case TRC_O_LAMBERTIAN_ABSORBANT: {
        double a, b;
        uniform2(seed, q, bounce, 0, a, b);
        u[0] = two_pi * a; u[1] = b;
        break;
    }
    case TRC_O_LAMBERTIAN_SPECULAR: case TRC_O_LAMBSPEC_DIRECTIONAL: case TRC_O_LAMB_PIECEWISE_SPEC: {
        double a, b, c, e;
        uniform2(seed, q, bounce, 0, a, b);
        uniform2(seed, q, bounce, 1, c, e);
        u[0] = a; u[1] = two_pi * b; u[2] = c;
        break;
    }
    case TRC_O_REFRACTIVE: {
        double a, b, c, e;
        normal2(seed, q, bounce, 0, a, b);
        uniform2(seed, q, bounce, 1, c, e);
        u[0] = a * (s->op[3] > 0.0 ? s->op[3] : 0.0); u[1] = two_pi * c; u[2] = e;
        break;
    }
    default: break;
    }
}

// TRC_O_REFRACTIVE: two child slots per hit (block A reflected, block B refracted), either of
// which may be void (d_cull = 2).  Out of line: keeps the common optics path at its register count.
// TRC_F_TABULATED: the two media are material tables, the indices complex (optics.cuh apply_refractive_tab).
__device__ __noinline__ void refractive_tab_slot(const OpticsParams &p, const TrcSurfaceRec *s, const double *tables,
                                                 int64_t q, int sidx, int64_t r, V3 v, V3 d, double e_in, double wl,
                                                 V3 hit, V3 n, const double *u) {
    const Cx m1 = cx(p.ref_index[r], p.ref_index_im ? p.ref_index_im[r] : 0.0);
    RefractTabOut ro = apply_refractive_tab(s, tables, d, n, e_in, m1, wl, u);
    const int64_t first = p.seg[sidx];
    // single_ray mode inherits the columns it does not set (the reflected ray's ref_index, both rays' wavelengths)
    // from bundle position = rank of the hit among the surface's hits (rays.inherit(argwhere(refl), ...), :797-815)
    Cx m_inherit = m1;
    if (s->op[2] != 0.0) {
        m_inherit = cx(p.ref_index[q - first], p.ref_index_im ? p.ref_index_im[q - first] : 0.0);
        wl = p.wavelengths[q - first];
    }
    double att_path = 0.0;
    if (s->op[4] != 0.0) {          // Absorbant.attenuate :874-881: path from the parent vertex, scaled
        double dx = hit.x - v.x, dy = hit.y - v.y, dz = hit.z - v.z;
        att_path = sqrt(dx * dx + dy * dy + dz * dz);
        if (s->op[7] != 1.0) att_path *= s->op[7];
    }
    const int64_t ca = p.child_off[sidx] + (q - first), cb = ca + (p.seg[sidx + 1] - first);
    int n_weak = 0, n_gone = 0;
    double e_out = 0.0;
    for (int side = 0; side < 2; ++side) {
        const int64_t c = side ? cb : ca;
        const bool has = side ? ro.has_refr : ro.has_refl;
        const V3 od = side ? ro.dir_refr : ro.dir_refl;
        const Cx mc = side ? ro.m2 : m_inherit;
        double oe = has ? (side ? ro.e_refr : ro.e_refl) : 0.0;
        if (s->op[4] != 0.0) oe = exp(-4.0 * M_PI * att_path * mc.im / wl) * oe;      // optics.attenuations, optics.py:205-212
        p.o_vert[c] = hit.x; p.o_vert[p.ld_out + c] = hit.y; p.o_vert[2 * p.ld_out + c] = hit.z;
        p.o_dir[c] = od.x; p.o_dir[p.ld_out + c] = od.y; p.o_dir[2 * p.ld_out + c] = od.z;
        p.o_energy[c] = oe;
        p.o_parents[c] = r;
        p.o_ref_index[c] = mc.re;
        if (p.o_ref_index_im) p.o_ref_index_im[c] = mc.im;
        if (p.o_ids) {
            uint64_t id = (uint64_t)(p.ids_in ? p.ids_in[r] : p.id_base + r);
            p.o_ids[c] = (int64_t)((id << 1) | (uint64_t)side);
        }
        if (p.o_wavelengths) p.o_wavelengths[c] = wl;
        const bool weak = has && oe <= p.min_energy;
        if (p.cull) p.cull[c] = has ? (weak ? 1 : 0) : 2;
        n_weak += weak ? 1 : 0;
        n_gone += has ? 0 : 1;
        e_out += has ? oe : 0.0;
    }
    if (p.absorbed) p.absorbed[q] = e_in - e_out;
    if (p.normals) { p.normals[q] = n.x; p.normals[p.ld_out + q] = n.y; p.normals[2 * p.ld_out + q] = n.z; }
    if (p.e_in) p.e_in[q] = e_in;
    if (p.attenuated) p.attenuated[q] = 0.0;
    if (p.dirs_in) { p.dirs_in[q] = d.x; p.dirs_in[p.ld_out + q] = d.y; p.dirs_in[2 * p.ld_out + q] = d.z; }
    if (n_weak && p.n_cull) atomicAdd(p.n_cull, (unsigned long long)n_weak);
    if (n_gone && p.n_void) atomicAdd(p.n_void, (unsigned long long)n_gone);
}

__device__ __noinline__ void refractive_slot(const OpticsParams &p, const TrcSurfaceRec *s, int64_t q, int sidx,
                                             int64_t r, V3 v, V3 d, double e_in, double wl, V3 hit, V3 n,
                                             const double *u) {
    double m1 = p.ref_index[r];
    RefractOut ro = apply_refractive(s, d, n, e_in, m1, u, v, hit);
    int64_t first = p.seg[sidx];
    // single_ray mode inherits the columns it does not set from bundle position (rank of the hit
    // among the surface's hits), not from the parent ray: rays.inherit(argwhere(refl), ...) at
    // optics_callables.py:1257-1278 -- reproduced as the reference behaves
    double m_inherit = m1;
    if (s->op[2] != 0.0) {
        m_inherit = p.ref_index[q - first];
        if (p.wavelengths) wl = p.wavelengths[q - first];
    }
    int64_t ca = p.child_off[sidx] + (q - first), cb = ca + (p.seg[sidx + 1] - first);
    int n_weak = 0, n_gone = 0;
    for (int side = 0; side < 2; ++side) {
        int64_t c = side ? cb : ca;
        bool has = side ? ro.has_refr : ro.has_refl;
        V3 od = side ? ro.dir_refr : ro.dir_refl;
        double oe = has ? (side ? ro.e_refr : ro.e_refl) : 0.0;
        p.o_vert[c] = hit.x; p.o_vert[p.ld_out + c] = hit.y; p.o_vert[2 * p.ld_out + c] = hit.z;
        p.o_dir[c] = od.x; p.o_dir[p.ld_out + c] = od.y; p.o_dir[2 * p.ld_out + c] = od.z;
        p.o_energy[c] = oe;
        p.o_parents[c] = r;
        p.o_ref_index[c] = side ? ro.m2 : m_inherit;
        if (p.o_ref_index_im) p.o_ref_index_im[c] = 0.0;
        if (p.o_ids) {                  // the two children of a split get distinct ids: (id << 1) | side
            uint64_t id = (uint64_t)(p.ids_in ? p.ids_in[r] : p.id_base + r);
            p.o_ids[c] = (int64_t)((id << 1) | (uint64_t)side);
        }
        if (p.o_wavelengths) p.o_wavelengths[c] = wl;
        bool weak = has && oe <= p.min_energy;
        if (p.cull) p.cull[c] = has ? (weak ? 1 : 0) : 2;
        n_weak += weak ? 1 : 0;
        n_gone += has ? 0 : 1;
    }
    if (p.absorbed) p.absorbed[q] = e_in - ((ro.has_refl ? ro.e_refl : 0.0) + (ro.has_refr ? ro.e_refr : 0.0));
    if (p.normals) { p.normals[q] = n.x; p.normals[p.ld_out + q] = n.y; p.normals[2 * p.ld_out + q] = n.z; }
    if (p.e_in) p.e_in[q] = e_in;
    if (p.attenuated) p.attenuated[q] = 0.0;
    if (p.dirs_in) { p.dirs_in[q] = d.x; p.dirs_in[p.ld_out + q] = d.y; p.dirs_in[2 * p.ld_out + q] = d.z; }
    if (n_weak && p.n_cull) atomicAdd(p.n_cull, (unsigned long long)n_weak);
    if (n_gone && p.n_void) atomicAdd(p.n_void, (unsigned long long)n_gone);
}

// One thread per child slot q (surface-major, parent-minor order).  Reads the
// 64-byte hit record (parent ray + surface) written by the nearest-hit kernel,
// recomputes the hit on the known surface (same arithmetic as that kernel, hence
// the same bits), applies the optics and writes the child coalesced.  Replaces
// Surface.select_rays/get_outgoing (surface.py:72-94) for every surface at once.
#ifndef TRC_OPTICS_MIN_CTAS
#define TRC_OPTICS_MIN_CTAS 3
#endif
// kRefractive: the scene has TRC_O_REFRACTIVE surfaces (the common instantiation does not carry that code).
template <bool kRefractive>
__global__ void __launch_bounds__(256, TRC_OPTICS_MIN_CTAS)
optics_kernel(const TrcSurfaceRec *__restrict__ surfs, const double *__restrict__ tables, OpticsParams p) {
    pdl_wait();
    pdl_trigger();
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p.ctl) {                       // device-driven: the hit count never visited the host
        long long hraw = (long long)p.ctl->hits;
        p.h = hraw <= p.capacity ? hraw : 0;       // an overflowed bounce does nothing: the host re-runs it
        p.wavelengths = p.ctl->in.W; p.ids_in = (const int64_t *)p.ctl->in.ids; p.id_base = p.ctl->in.id_base;
    }
    bool live = q < p.h;
    bool culled = false;
    double a_val = 0.0;                // absorbed energy of this slot and its surface (per-surface sums)
    int a_key = -1;
    if (live) {
        const double2 *rec = reinterpret_cast<const double2 *>(p.records + p.slot_rec[q]);
        double2 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3);
        V3 v = mk(r0.x, r0.y, r1.x);
        V3 d = mk(r1.y, r2.x, r2.y);
        double e_in = r3.x;
        long long tail = __double_as_longlong(r3.y);
        int64_t r = (int64_t)(unsigned)(tail & 0xffffffffLL);
        int sidx = (int)(tail >> 32);
        const TrcSurfaceRec *s = &surfs[sidx];
        double wl = p.wavelengths ? p.wavelengths[r] : 0.0;
        double t = intersect(s, v, d, tables);
        V3 hit = axpy(v, t, d);
        V3 n = surface_normal(s, hit, d);
        double u[3] = {0.0, 0.0, 0.0};
        if (p.draws[0]) {
            u[0] = p.draws[0][q];
            if (p.draws[1]) u[1] = p.draws[1][q];
            if (p.draws[2]) u[2] = p.draws[2][q];
        } else {
            // keyed by the SOURCE ray, not by the child slot: a sharded trace draws what the whole trace draws
            philox_draws(s, p.seed, p.bounce, (uint64_t)(p.ids_in ? p.ids_in[r] : p.id_base + r), u);
        }
        if (kRefractive && s->okind == TRC_O_REFRACTIVE) {
            if (s->flags & TRC_F_TABULATED) refractive_tab_slot(p, s, tables, q, sidx, r, v, d, e_in, wl, hit, n, u);
            else refractive_slot(p, s, q, sidx, r, v, d, e_in, wl, hit, n, u);
        } else {
        OpticsOut o = apply_optics(s, tables, d, n, e_in, wl, u, v, hit);
        if (s->flags & TRC_F_ONE_SIDED) {
            // OneSided*: rays arriving along +z of the surface lose all energy
            // (optics_callables.py:201-212, :498-504)
            const double *f = s->frame;
            double proj = d.x * f[2] + d.y * f[6] + d.z * f[10];
            if (proj > 0.0) o.energy = 0.0;
        }
        // child index: the slot itself, unless some surface of the scene emits two children per hit
        int64_t c = q, c2 = -1;
        if (p.child_off) {
            int64_t first = p.seg[sidx];
            c = p.child_off[sidx] + (q - first);
            if (s->okind == TRC_O_PERIODIC) c2 = c + (p.seg[sidx + 1] - first);
        }
        double e_first = o.energy;
        if (c2 >= 0) {
            // PeriodicBoundary (:703-721): block A = zero-energy stubs at the hit, block B = the
            // rays themselves moved by period * normal
            e_first = 0.0;
            V3 moved = mk(hit.x + s->op[0] * n.x, hit.y + s->op[0] * n.y, hit.z + s->op[0] * n.z);
            p.o_vert[c2] = moved.x; p.o_vert[p.ld_out + c2] = moved.y; p.o_vert[2 * p.ld_out + c2] = moved.z;
            p.o_dir[c2] = d.x; p.o_dir[p.ld_out + c2] = d.y; p.o_dir[2 * p.ld_out + c2] = d.z;
            p.o_energy[c2] = e_in;
            p.o_parents[c2] = r;
            if (p.o_ref_index) p.o_ref_index[c2] = p.ref_index[r];
            if (p.o_wavelengths) p.o_wavelengths[c2] = wl;
            bool weak2 = e_in <= p.min_energy;
            if (p.cull) p.cull[c2] = weak2 ? 1 : 0;
            if (weak2 && p.n_cull) atomicAdd(p.n_cull, 1ULL);
        }
        p.o_vert[c] = hit.x; p.o_vert[p.ld_out + c] = hit.y; p.o_vert[2 * p.ld_out + c] = hit.z;
        p.o_dir[c] = o.dir.x; p.o_dir[p.ld_out + c] = o.dir.y; p.o_dir[2 * p.ld_out + c] = o.dir.z;
        p.o_energy[c] = e_first;
        p.o_parents[c] = r;
        if (p.o_ref_index) p.o_ref_index[c] = p.ref_index[r];
        if (kRefractive && p.o_ref_index_im) {
            const double im = p.ref_index_im ? p.ref_index_im[r] : 0.0;
            p.o_ref_index_im[c] = im;
            if (c2 >= 0) p.o_ref_index_im[c2] = im;
        }
#ifndef TRC_OPT_NO_IDS
        if (p.o_ids) {
            int64_t id = p.ids_in ? p.ids_in[r] : p.id_base + r;
            p.o_ids[c] = id;
            if (c2 >= 0) p.o_ids[c2] = id;
        }
#endif
        a_val = e_in - o.energy; a_key = sidx;
        if (p.absorbed) p.absorbed[q] = e_in - o.energy;
        if (p.normals) { p.normals[q] = n.x; p.normals[p.ld_out + q] = n.y; p.normals[2 * p.ld_out + q] = n.z; }
        if (p.e_in) p.e_in[q] = e_in;
        if (p.attenuated) p.attenuated[q] = o.attenuated;
        if (p.dirs_in) { p.dirs_in[q] = d.x; p.dirs_in[p.ld_out + q] = d.y; p.dirs_in[2 * p.ld_out + q] = d.z; }
        if (p.o_wavelengths) p.o_wavelengths[c] = wl;
        culled = e_first <= p.min_energy;
        if (p.cull) p.cull[c] = culled ? 1 : 0;
        }
    }
    if (p.n_cull) {
        unsigned m = __ballot_sync(0xffffffffu, culled);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(p.n_cull, (unsigned long long)__popc(m));
    }
    if (!p.ctl) return;
    // ---- device-driven trace: survivors per CTA and per-surface absorbed sums with ONE block barrier;
    // bounce_close_kernel (next launch) turns the per-CTA counts into the bounce's totals.
    __shared__ unsigned s_cnt[8];
    __shared__ double s_sum[8];
    __shared__ int s_key[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t blk0 = (int64_t)blockIdx.x * blockDim.x;
    if (blk0 < p.h) {
        const unsigned ms = __ballot_sync(0xffffffffu, live && !culled);
        // warp-level segmented sum of a_val by surface (keys are non-decreasing with the lane: slot order;
        // dead lanes at the end of the last block carry key -1)
        double acc = a_val;
        const int key = a_key;
        for (int o = 1; o < 32; o <<= 1) {
            double other = __shfl_down_sync(0xffffffffu, acc, o);
            int okey = __shfl_down_sync(0xffffffffu, key, o);
            if (lane + o < 32 && okey == key) acc += other;
        }
        const int prev = __shfl_up_sync(0xffffffffu, key, 1);
        const bool head = key >= 0 && (lane == 0 || prev != key);
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        // a warp whose live slots lie on ONE surface hands its sum to thread 0; otherwise one atomic per segment
        const bool one = __popc(heads) == 1;
        if (head) {
            if (one) { s_sum[warp] = acc; s_key[warp] = key; }
            else atomicAdd(&p.absorbed_per_surface[key], acc);
        }
        if (lane == 0) {
            s_cnt[warp] = __popc(ms);
            if (!one) s_key[warp] = -1;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned c = 0;
            int k = -1;
            double t = 0.0;
            for (int w = 0; w < 8; ++w) {                 // runs of warps on the same surface: one atomic per run
                c += s_cnt[w];
                const int kw = s_key[w];
                if (kw != k) {
                    if (k >= 0) atomicAdd(&p.absorbed_per_surface[k], t);
                    k = kw; t = 0.0;
                }
                if (kw >= 0) t += s_sum[w];
            }
            if (k >= 0) atomicAdd(&p.absorbed_per_surface[k], t);
            p.cta_surv[blockIdx.x] = c;
        }
    }
}

// Closes a bounce of the device-driven trace (one block, launched right after the optics kernel): survivor
// total from the per-CTA counts, the reorder decision, the exclusive scan the partition kernel needs when it
// will run, and the next bounce's ray count and bundle.  A separate launch rather than a "last block" ticket
// inside the optics kernel: a ticket costs every CTA an atomic round trip before it can retire.
struct CloseParams {
    TrcBounceCtl *ctl, *ctl_next;
    TrcBundleDesc children, children_sorted;
    int64_t capacity;
    uint32_t *cta_surv;
    int64_t *d_counts;
    const int64_t *seg;
    int n_seg;
};
__global__ void __launch_bounds__(1024)
bounce_close_kernel(CloseParams p) {
    __shared__ unsigned long long s_part[32];
    __shared__ unsigned s_scan[32];
    __shared__ unsigned s_carry;
    pdl_wait();
    pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long hraw = (long long)p.ctl->hits;
    const unsigned overflow = hraw > p.capacity ? 1u : 0u;
    const long long h = overflow ? 0 : hraw;
    const int64_t nb = (h + 255) / 256;
    if (p.seg)                         // the slot offsets travel with the counts: one device-to-host copy per trace
        for (int k = threadIdx.x; k < p.n_seg; k += 1024) p.d_counts[4 + k] = p.seg[k];
    unsigned long long t = 0;
    for (int64_t i = threadIdx.x; i < nb; i += 1024) t += p.cta_surv[i];
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) s_part[warp] = t;
    __syncthreads();
    unsigned long long tot = 0;
    for (int w = 0; w < 32; ++w) tot += s_part[w];
    const long long n_surv = (long long)tot;
    const unsigned reorder = (n_surv > 0 && n_surv < h) ? 1u : 0u;
    if (reorder) {
        if (threadIdx.x == 0) s_carry = 0;
        __syncthreads();
        for (int64_t start = 0; start < nb; start += 4096) {
            const int64_t i0 = start + 4 * (int64_t)threadIdx.x;
            unsigned v[4], s4 = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) { v[k] = i0 + k < nb ? p.cta_surv[i0 + k] : 0u; s4 += v[k]; }
            unsigned incl = s4;
            for (int o = 1; o < 32; o <<= 1) { unsigned x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
            if (lane == 31) s_scan[warp] = incl;
            __syncthreads();
            unsigned wbase = 0;
            for (int w = 0; w < warp; ++w) wbase += s_scan[w];
            unsigned run = s_carry + wbase + incl - s4;
#pragma unroll
            for (int k = 0; k < 4; ++k) { if (i0 + k < nb) p.cta_surv[i0 + k] = run; run += v[k]; }
            __syncthreads();
            if (threadIdx.x == 1023) s_carry = run;
            __syncthreads();
        }
    }
    if (threadIdx.x == 0) {
        p.ctl->n_surv = n_surv;
        p.ctl->reorder = reorder;
        p.ctl->overflow = overflow;
        if (p.ctl_next) {
            p.ctl_next->n_in = overflow ? 0 : n_surv;     // an overflowed bounce feeds nothing forward
            p.ctl_next->in = reorder ? p.children_sorted : p.children;
        }
        p.d_counts[0] = p.ctl->n_in; p.d_counts[1] = hraw; p.d_counts[2] = n_surv; p.d_counts[3] = reorder;
    }
}

// Polychromatic rays (trc_spectra): one thread per child slot, W samples each; see the header for the rules.
__global__ void __launch_bounds__(256)
spectra_kernel(const TrcSurfaceRec *__restrict__ surfs, const double *__restrict__ tables, TrcSpectraArgs a) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool culled = false;
    if (q < a.h) {
        const TrcSurfaceRec *s = &surfs[a.d_slot_surf[q]];
        const int64_t r = a.d_parents[q];
        const int W = a.n_samples;
        double factor = 1.0;
        bool poly = (s->flags & TRC_F_POLY) != 0;
        double th = 0.0;
        if (poly || s->okind == TRC_O_LAMB_DIRECTIONAL) {
            // incidence angle as the optics compute it: acos(|(d.n) n|)  (optics_callables.py:341-345, :405-409)
            V3 n = mk(a.d_normals[q], a.d_normals[a.ld_cols + q], a.d_normals[2 * a.ld_cols + q]);
            V3 d = mk(a.d_dirs_in[q], a.d_dirs_in[a.ld_cols + q], a.d_dirs_in[2 * a.ld_cols + q]);
            double k = dot(d, n);
            V3 vert = mk(k * n.x, k * n.y, k * n.z);
            th = acos(sqrt(vert.x * vert.x + vert.y * vert.y + vert.z * vert.z));
        }
        if (!poly && !(s->flags & TRC_F_IAM)) {
            if (s->okind == TRC_O_REFLECTIVE || s->okind == TRC_O_REAL_REFLECTIVE || s->okind == TRC_O_LAMBERTIAN)
                factor = 1.0 - s->op[0];
            else if (s->okind == TRC_O_LAMB_DIRECTIONAL)
                factor = 1.0 - interp1(th, tables + s->table_off, tables + s->table_off + s->table_len, s->table_len);
        }
        double energy = 0.0, wl_prev = 0.0, sp_prev = 0.0;
        for (int w = 0; w < W; ++w) {
            const double wl = a.d_in_wavelengths[(int64_t)w * a.ld_in + r];
            const double sp_in = a.d_in_spectra[(int64_t)w * a.ld_in + r];
            double sp = sp_in * factor;
            if (poly) sp = sp_in * (1.0 - interp2(th, wl, tables + s->table_off, s->table_len, (int)s->op[6]));
            a.d_out_wavelengths[(int64_t)w * a.ld_out + q] = wl;
            a.d_out_spectra[(int64_t)w * a.ld_out + q] = sp;
            if (a.d_lost_spectra) a.d_lost_spectra[(int64_t)w * a.ld_out + q] = sp_in - sp;
            if (w > 0) energy += (wl - wl_prev) * (sp + sp_prev) / 2.0;      // numpy.trapz: sum(diff(x) * (y[1:] + y[:-1]) / 2)
            wl_prev = wl; sp_prev = sp;
        }
        double e = a.d_energy[q];
        if (poly) {
            if (a.d_absorbed) a.d_absorbed[q] = e - energy;
            a.d_energy[q] = energy;
            e = energy;
        }
        culled = e <= a.min_energy;
        a.d_cull[q] = culled ? 1 : 0;
    }
    const unsigned m = __ballot_sync(0xffffffffu, culled);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(reinterpret_cast<unsigned long long *>(a.d_n_cull), (unsigned long long)__popc(m));
}

__global__ void __launch_bounds__(256)
gather_rows_kernel(int64_t n_out, const int32_t *__restrict__ idx, int n_rows,
                   const double *__restrict__ in, int64_t ld_in, double *__restrict__ out, int64_t ld_out,
                   const int64_t *__restrict__ in_i64, int64_t *__restrict__ out_i64) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_out) return;
    int64_t src = idx[j];
    for (int r = 0; r < n_rows; ++r) out[(int64_t)r * ld_out + j] = in[(int64_t)r * ld_in + src];
    if (in_i64) out_i64[j] = in_i64[src];
}

// Sum of `values` per key over a key-sorted slot array.  Each thread folds 8
// consecutive slots (flushing on a key change), then the warp does a segmented
// shuffle reduction so that one atomicAdd per (warp, key) reaches memory -- the
// work is spread over all slots even when every hit is on one surface.
constexpr int kSumItems = 8;
__global__ void __launch_bounds__(256)
segment_sums_kernel(int64_t h, const int32_t *__restrict__ keys, const double *__restrict__ values,
                    double *__restrict__ totals) {
    const int lane = threadIdx.x & 31;
    int64_t first = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * kSumItems;
    int key = -1;
    double acc = 0.0;
    for (int k = 0; k < kSumItems; ++k) {
        int64_t i = first + k;
        if (i >= h) break;
        int kk = keys[i];
        if (kk != key) {
            if (key >= 0) atomicAdd(&totals[key], acc);
            key = kk; acc = 0.0;
        }
        acc += values[i];
    }
    // segmented reduction across the warp (keys are non-decreasing with the lane)
    for (int o = 1; o < 32; o <<= 1) {
        double other = __shfl_down_sync(0xffffffffu, acc, o);
        int okey = __shfl_down_sync(0xffffffffu, key, o);
        if (lane + o < 32 && okey == key) acc += other;
    }
    int prev = __shfl_up_sync(0xffffffffu, key, 1);
    if (key >= 0 && (lane == 0 || prev != key)) atomicAdd(&totals[key], acc);
}

struct Proj { double m[12]; };

// histogram2d semantics: bins [e_k, e_{k+1}) with the last bin closed, edges
// from linspace(-h, h, n+1) (flat_surface.py:241-244).
__device__ __forceinline__ int bin_index(double x, double half, int nb) {
    if (x < -half || x > half) return -1;
    double w = 2.0 * half / nb;
    int k = (int)floor((x + half) / w);
    if (k >= nb) k = nb - 1;
    if (k < 0) k = 0;
    return k;
}

// Each CTA bins into a private shared-memory copy of the map (all hits of a
// receiver land in a few thousand bins: global atomics would serialise) and
// adds its non-empty bins to the global map once at the end.
constexpr int kFluxSmemBins = 4096;
__global__ void __launch_bounds__(256)
fluxmap_rect_kernel(int64_t n, const double *__restrict__ hits, int64_t ld, const double *__restrict__ energy,
                    Proj pr, double hx, double hy, int nx, int ny, double *__restrict__ bins) {
    __shared__ double local[kFluxSmemBins];
    const int nb = nx * ny;
    const bool priv = nb <= kFluxSmemBins;
    if (priv) {
        for (int k = threadIdx.x; k < nb; k += blockDim.x) local[k] = 0.0;
        __syncthreads();
    }
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        V3 g = mk(hits[i], hits[ld + i], hits[2 * ld + i]);
        double lx = pr.m[0] * g.x + pr.m[1] * g.y + pr.m[2] * g.z + pr.m[3];
        double ly = pr.m[4] * g.x + pr.m[5] * g.y + pr.m[6] * g.z + pr.m[7];
        int ix = bin_index(lx, hx, nx), iy = bin_index(ly, hy, ny);
        if (ix < 0 || iy < 0) continue;
        if (priv) atomicAdd(&local[ix * ny + iy], energy[i]);
        else atomicAdd(&bins[ix * ny + iy], energy[i]);
    }
    if (priv) {
        __syncthreads();
        for (int k = threadIdx.x; k < nb; k += blockDim.x) {
            double v = local[k];
            if (v != 0.0) atomicAdd(&bins[k], v);
        }
    }
}

// numpy.histogramdd's bin of x on explicit edges e[0..n]: searchsorted(e, x, 'right') - 1,
// values equal to the last edge fall in the last bin, anything else outside -> -1.
__device__ __forceinline__ int edge_bin(double x, const double *e, int n) {
    if (!(x >= e[0]) || !(x <= e[n])) return -1;
    if (x == e[n]) return n - 1;
    int lo = 0, hi = n;              // invariant: e[lo] <= x < e[hi]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (e[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// Flux-map binning in the surface's own mesh coordinates (the get_fluxmap methods of the
// geometry managers): local = proj . [hit; 1], (u, v) per TrcFluxCoord, weights summed per
// (u, v) bin on explicit edges.  Edges are staged in shared memory; bins are privatised per
// block when they fit.
constexpr int kFluxMaxEdges = 1000;   // static shared memory: 32 KB of bins + 2 x 8 KB of edges < 48 KB
__global__ void __launch_bounds__(256)
fluxmap_binned_kernel(int coord, int64_t n, const double *__restrict__ hits, int64_t ld,
                      const double *__restrict__ energy, Proj pr, double param,
                      const double *__restrict__ edges_u, int nu, const double *__restrict__ edges_v, int nv,
                      double *__restrict__ bins) {
    __shared__ double local[kFluxSmemBins];
    __shared__ double eu[kFluxMaxEdges], ev[kFluxMaxEdges];
    const int nb = nu * nv;
    const bool priv = nb <= kFluxSmemBins;
    for (int k = threadIdx.x; k <= nu; k += blockDim.x) eu[k] = edges_u[k];
    for (int k = threadIdx.x; k <= nv; k += blockDim.x) ev[k] = edges_v[k];
    if (priv)
        for (int k = threadIdx.x; k < nb; k += blockDim.x) local[k] = 0.0;
    __syncthreads();
    const double two_pi = 2.0 * M_PI;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        V3 g = mk(hits[i], hits[ld + i], hits[2 * ld + i]);
        double lx = pr.m[0] * g.x + pr.m[1] * g.y + pr.m[2] * g.z + pr.m[3];
        double ly = pr.m[4] * g.x + pr.m[5] * g.y + pr.m[6] * g.z + pr.m[7];
        double lz = pr.m[8] * g.x + pr.m[9] * g.y + pr.m[10] * g.z + pr.m[11];
        double u, v;
        switch (coord) {
        case TRC_FLUX_XY: u = lx; v = ly; break;
        case TRC_FLUX_R_ATAN_XY: u = sqrt(lx * lx + ly * ly); v = atan2(lx, ly); if (v < 0.0) v += two_pi; break;
        case TRC_FLUX_R_ATAN_YX: u = sqrt(lx * lx + ly * ly); v = atan2(ly, lx); if (v < 0.0) v += two_pi; break;
        case TRC_FLUX_Z_PHI: u = lz; v = atan2(ly, lx); if (v < 0.0) v += two_pi; break;
        default: u = acos(lz / param); v = atan2(ly, lx); if (v < 0.0) v += two_pi; break;   // TRC_FLUX_THETA_PHI
        }
        int iu = edge_bin(u, eu, nu), iv = edge_bin(v, ev, nv);
        if (iu < 0 || iv < 0) continue;
        if (priv) atomicAdd(&local[iu * nv + iv], energy[i]);
        else atomicAdd(&bins[iu * nv + iv], energy[i]);
    }
    if (priv) {
        __syncthreads();
        for (int k = threadIdx.x; k < nb; k += blockDim.x) {
            double w = local[k];
            if (w != 0.0) atomicAdd(&bins[k], w);
        }
    }
}

}  // namespace

// dyn != nullptr: launch of the device-driven trace (trace.cu): a->h is the hit CAPACITY (grid size), the
// count is read from dyn->ctl->hits, and the last block closes the bounce.
int trc_optics_dyn(const TrcScene *sc, const TrcOpticsArgs *a, const TrcOpticsDyn *dyn, void *stream) {
    if (!sc || !a || a->h < 0) return fail(TRC_ERR_BAD_ARGUMENT, "trc_optics: bad argument%s");
    if (a->h == 0) return TRC_OK;
    if (!a->d_records || !a->d_slot_rec || !a->d_out_vertices || !a->d_out_directions || !a->d_out_energy ||
        !a->d_out_parents || a->ld_out < a->h)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_optics: null buffer or bad stride%s");
    OpticsParams p;
    p.h = a->h; p.records = a->d_records; p.slot_rec = a->d_slot_rec; p.wavelengths = a->d_wavelengths;
    p.o_vert = a->d_out_vertices; p.o_dir = a->d_out_directions; p.o_energy = a->d_out_energy;
    p.o_parents = a->d_out_parents; p.ld_out = a->ld_out;
    for (int k = 0; k < 3; ++k) p.draws[k] = a->d_draws[k];
    p.seed = a->seed; p.bounce = a->bounce; p.min_energy = a->min_energy;
    p.absorbed = a->d_absorbed; p.normals = a->d_normals; p.e_in = a->d_e_in; p.dirs_in = a->d_dirs_in;
    p.o_wavelengths = a->d_out_wavelengths; p.cull = a->d_cull;
    p.n_cull = reinterpret_cast<unsigned long long *>(a->d_n_cull);
    p.attenuated = a->d_attenuated;
    p.slot_surf = a->d_slot_surf; p.seg = a->d_seg; p.child_off = a->d_child_off;
    p.ref_index = a->d_ref_index; p.o_ref_index = a->d_out_ref_index;
    p.n_void = reinterpret_cast<unsigned long long *>(a->d_n_void);
    p.ids_in = a->d_ray_ids; p.id_base = a->ray_id_base; p.o_ids = a->d_out_ray_ids;
    p.ref_index_im = a->d_ref_index_im; p.o_ref_index_im = a->d_out_ref_index_im;
    p.ctl = nullptr; p.capacity = a->h; p.cta_surv = nullptr; p.absorbed_per_surface = nullptr;
    if (dyn) {
        if (sc->has_split) return fail(TRC_ERR_UNSUPPORTED_KIND, "trc_optics: device-driven trace with two-children optics%s");
        p.ctl = dyn->ctl; p.cta_surv = dyn->cta_surv;
        p.absorbed_per_surface = dyn->absorbed_per_surface;
    }
    if ((p.ref_index == nullptr) != (p.o_ref_index == nullptr))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_optics: d_ref_index and d_out_ref_index go together%s");
    if (sc->has_refractive && (!p.ref_index || !p.child_off))
        return fail(TRC_ERR_BAD_ARGUMENT,
                    "trc_optics: a scene with TRC_O_REFRACTIVE surfaces needs d_ref_index and d_child_off%s");
    if (p.child_off && !p.seg) return fail(TRC_ERR_BAD_ARGUMENT, "trc_optics: d_child_off needs d_seg%s");
    if (sc->has_tabulated && (!p.wavelengths || !p.o_ref_index_im))
        return fail(TRC_ERR_BAD_ARGUMENT,
                    "trc_optics: refractive media from material tables need d_wavelengths and d_out_ref_index_im%s");
    int64_t blocks = (a->h + 255) / 256;
    if (dyn) {
        TRC_CUDA(launch_pdl(optics_kernel<false>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream,
                            (const TrcSurfaceRec *)sc->d_surfaces, (const double *)sc->d_tables, p));
        CloseParams c;
        c.ctl = dyn->ctl; c.ctl_next = dyn->ctl_next; c.children = dyn->children;
        c.children_sorted = dyn->children_sorted; c.capacity = a->h; c.cta_surv = dyn->cta_surv;
        c.d_counts = dyn->d_counts; c.seg = dyn->seg; c.n_seg = dyn->n_seg;
        TRC_CUDA(launch_pdl(bounce_close_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream, c));
        return check_launch("optics", 2);
    } else if (sc->has_refractive)
        optics_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(sc->d_surfaces, sc->d_tables, p);
    else
        optics_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(sc->d_surfaces, sc->d_tables, p);
    return check_launch("optics");
}

extern "C" int trc_optics(const TrcScene *sc, const TrcOpticsArgs *a, void *stream) {
    return trc_optics_dyn(sc, a, nullptr, stream);
}

extern "C" int trc_spectra(const TrcScene *sc, const TrcSpectraArgs *a, void *stream) {
    if (!sc || !a || a->h < 0 || a->n_samples < 1) return fail(TRC_ERR_BAD_ARGUMENT, "trc_spectra: bad argument%s");
    if (a->h == 0) return TRC_OK;
    if (!a->d_slot_surf || !a->d_parents || !a->d_in_wavelengths || !a->d_in_spectra || !a->d_normals || !a->d_dirs_in ||
        !a->d_out_wavelengths || !a->d_out_spectra || !a->d_energy || !a->d_cull || !a->d_n_cull ||
        a->ld_out < a->h || a->ld_cols < a->h)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_spectra: null buffer or bad stride%s");
    if (sc->has_split) return fail(TRC_ERR_UNSUPPORTED_KIND, "trc_spectra: scene with two-children optics%s");
    spectra_kernel<<<(unsigned)((a->h + 255) / 256), 256, 0, (cudaStream_t)stream>>>(sc->d_surfaces, sc->d_tables, *a);
    return check_launch("spectra");
}

extern "C" int trc_gather_rows(int64_t n_out, const int32_t *d_idx, int32_t n_rows, const double *d_in,
                               int64_t ld_in, double *d_out, int64_t ld_out, const int64_t *d_in_i64,
                               int64_t *d_out_i64, void *stream) {
    if (n_out < 0 || n_rows < 0 || (n_out > 0 && (!d_idx || (n_rows > 0 && (!d_in || !d_out)))))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_gather_rows: bad argument%s");
    if (n_out == 0) return TRC_OK;
    int64_t blocks = (n_out + 255) / 256;
    gather_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(n_out, d_idx, n_rows, d_in, ld_in, d_out,
                                                                           ld_out, d_in_i64, d_out_i64);
    return check_launch("gather_rows");
}

extern "C" int trc_segment_sums(int64_t h, const int32_t *d_sorted_keys, const double *d_values,
                                double *d_totals, void *stream) {
    if (h < 0 || !d_totals || (h > 0 && (!d_sorted_keys || !d_values)))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_segment_sums: bad argument%s");
    if (h == 0) return TRC_OK;
    int64_t threads = (h + kSumItems - 1) / kSumItems;
    segment_sums_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(h, d_sorted_keys,
                                                                                             d_values, d_totals);
    return check_launch("segment_sums");
}

extern "C" int trc_fluxmap_rect(int64_t n, const double *d_hits, int64_t ld, const double *d_energy,
                                const double *proj12, double hx, double hy, int32_t nx, int32_t ny,
                                double *d_bins, void *stream) {
    if (n < 0 || !proj12 || nx < 1 || ny < 1 || !d_bins || (n > 0 && (!d_hits || !d_energy)))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_fluxmap_rect: bad argument%s");
    if (n == 0) return TRC_OK;
    Proj pr;
    for (int k = 0; k < 12; ++k) pr.m[k] = proj12[k];
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 2) blocks = 148 * 2;
    fluxmap_rect_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(n, d_hits, ld, d_energy, pr, hx, hy, nx,
                                                                            ny, d_bins);
    return check_launch("fluxmap_rect");
}

extern "C" int trc_fluxmap_binned(int32_t coord, int64_t n, const double *d_hits, int64_t ld, const double *d_energy,
                                  const double *proj12, double param, const double *d_edges_u, int32_t nu,
                                  const double *d_edges_v, int32_t nv, double *d_bins, void *stream) {
    if (coord < 0 || coord >= TRC_FLUX_COORD_COUNT || n < 0 || !proj12 || !d_edges_u || !d_edges_v || !d_bins ||
        nu < 1 || nv < 1 || nu >= kFluxMaxEdges || nv >= kFluxMaxEdges || (n > 0 && (!d_hits || !d_energy || ld < n)))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_fluxmap_binned: bad argument%s");
    if (n == 0) return TRC_OK;
    Proj pr;
    for (int k = 0; k < 12; ++k) pr.m[k] = proj12[k];
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 2) blocks = 148 * 2;
    fluxmap_binned_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(coord, n, d_hits, ld, d_energy, pr,
                                                                              param, d_edges_u, nu, d_edges_v, nv,
                                                                              d_bins);
    return check_launch("fluxmap_binned");
}
