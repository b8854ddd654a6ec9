// Shared host-side helpers of the C ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../include/tracer_b200.h"

// thread-local last-error text, defined in tracer_b200.cu
char *trc_err_buf();
static const int TRC_ERR_BUF = 512;

static inline int fail(int code, const char *fmt, const char *detail = "") {
    snprintf(trc_err_buf(), TRC_ERR_BUF, fmt, detail);
    return code;
}
#define TRC_CUDA(call)                                                         \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) return fail(TRC_ERR_CUDA, #call ": %s", cudaGetErrorString(e_)); \
    } while (0)

// process-wide count of kernels launched by this library (bench.py's gpu_launches)
void trc_note_launches(int n);

static inline int check_launch(const char *what, int n_kernels = 1) {
    trc_note_launches(n_kernels);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(trc_err_buf(), TRC_ERR_BUF, "%s launch: %s", what, cudaGetErrorString(e));
        return TRC_ERR_CUDA;
    }
    return TRC_OK;
}

// ---- programmatic dependent launch (sm_90+): every kernel of the device-driven trace starts with
// pdl_wait() in ALL its threads (before any early exit: "kernel B complete" must imply "kernel A
// complete" for the kernel after B) and may call pdl_trigger() early so that the next kernel's
// prologue overlaps this one's tail.  Launched without the attribute both are no-ops.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args... args) {
    static const bool no_pdl = getenv("TRC_NO_PDL") != nullptr;   // profiling: kernels then neither overlap nor wait
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = no_pdl ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

// ---- device-side control block of the device-driven trace (trace.cu).  One per bounce; the counts
// never visit the host between the first launch of a trace and its final read-back.
struct TrcBundleDesc {               // where a bundle lives: SoA rows of stride ld (ray_bundle.py:6-63)
    const double *V, *D, *E, *W;       // W (wavelengths) nullable
    const long long *ids;              // source-ray ids, nullable: id = id_base + index
    long long id_base;
    long long ld;
};
struct TrcBounceCtl {
    long long n_in;                    // rays entering the bounce (bounce 0: host; later: survivors of the previous one)
    unsigned long long chunk_counter;  // dynamic work distribution of the tree walk
    unsigned long long hits;           // raw hit count of the nearest-hit kernel (may exceed the capacity)
    long long n_surv;                  // children with energy > min_energy (optics finalize)
    unsigned int ticket_order;         // "last block" tickets
    unsigned int ticket_optics;
    unsigned int reorder;              // 1: some but not all children are weak -> the survivors-first copy is made
    unsigned int overflow;             // hits > capacity: results of this and later bounces are void
    TrcBundleDesc in;                  // the bundle entering the bounce: the children of the bounce before, in slot
                                       // order or (reorder) the survivors-first copy -- chosen on the device
    long long pad_[3];
};                                     // 128 bytes
static_assert(sizeof(TrcBounceCtl) == 128, "TrcBounceCtl layout");

// The tree as the walk reads it: one 32-byte record per INNER node holding the
// boxes of its two children quantised to 16 bits on the grid of the root box
// (rounded outward), split into two 16-byte halves kept in separate arrays so
// that random per-lane fetches spread over all shared-memory bank groups.  The
// two planes of an axis share a word (lo | hi << 16), so the walk picks the
// near / far plane of a ray with one byte-permute whose selector depends only
// on the sign of the ray direction:
//   a = { c0.axis0, c0.axis1, c0.axis2, c1.axis0 }
//   b = { c1.axis1, c1.axis2, child0, child1 }
// child word: inner node -> its pair index (bit 31 clear);
//             leaf       -> bit 31 | (surface count - 1) << 22 | first entry in the
//                           candidate list, i.e. the low 24 bits are a queue entry
//                           without its ray slot; a leaf of ONE surface carries the
//                           surface index itself instead of a list position.
// Pairs are numbered breadth-first from the root (pair 0 = the root's children).
// root_meta (host side only): bits 28..30 = leaf surface count of the root itself
// (0 = inner node), bits 0..27 = first candidate.
struct TrcPairHalf { uint32_t w[4]; };
#define TRC_META_COUNT(m) (((m) >> 28) & 7u)
#define TRC_META_INDEX(m) ((m) & 0x0fffffffu)
#define TRC_CHILD_LEAF 0x80000000u
#define TRC_CHILD_ENTRY_MASK 0x00ffffffu

// Shadow grid of a directionally coherent bundle (built by trc_scene_set_shadow_grid, shadow_grid.cu; read by the
// tree-walk kernel's stage A).  A ray whose slopes (d_a / d_n, d_b / d_n) lie in the slope box is keyed by the point
// where it crosses the reference plane x_n = h0; the cell of that point lists every surface whose box the ray can
// touch (the builder widened each shadow by the slope box and by the float32 rounding of the kernel's arithmetic),
// so such a ray needs no tree walk.  Rays outside the slope box, too far out (|v| > rmax) or with d_n of the other
// sign take the walk: the grid only prunes, results are those of the walk bit for bit.
struct TrcShadowGrid {
    const uint32_t *bits;          // one bit per cell: some surface's shadow covers it (copied to shared memory)
    const uint32_t *cells;         // payload << 8 | count; count == 1: payload = the surface index, else list offset
    const int32_t *list;           // surface indices of the cells with more than one candidate
    int n_words;                   // 32-bit words of `bits`
    int axis_n, axis_a, axis_b;
    int nx, ny;
    float h0, sgn;                 // reference plane, sign of d_n inside the slope box
    float t0a, t0b, da, db;        // slope box centre and half widths
    float ga0, gb0, inv_ca, inv_cb;   // grid origin and 1 / cell size
    float rmax;
};

// Device-side scene tables; owned by the library (trc_scene_create/destroy).
struct TrcScene {
    int32_t n_surfaces = 0;
    TrcSurfaceRec *d_surfaces = nullptr;
    double *d_tables = nullptr;
    int32_t n_table = 0;
    TrcBvhNode *d_nodes = nullptr;     // float64 boxes as uploaded (kept for inspection)
    int32_t n_nodes = 0;
    TrcPairHalf *d_pair_a = nullptr;   // quantised children boxes, first halves  (n_pairs)
    TrcPairHalf *d_pair_b = nullptr;   // second halves + child metadata          (n_pairs)
    TrcPairHalf *d_pair_ab = nullptr;  // the same pairs interleaved (a0 b0 a1 b1 ...): one 32 B sector per visit for the
                                       // part of the tree that is walked from global memory / L2
    int32_t n_pairs = 0;
    uint32_t root_meta = 0;            // meta word of the root node itself
    float root_lo[3] = {0.f, 0.f, 0.f};   // root box (float32, rounded outward) = quantisation grid origin
    float root_hi[3] = {0.f, 0.f, 0.f};
    float qstep[3] = {0.f, 0.f, 0.f};  // grid step per axis: plane = root_lo + q * qstep
    float extent[3] = {0.f, 0.f, 0.f}; // max |coordinate| of the root box per axis
    unsigned long long *d_chunk_counter = nullptr;   // dynamic work distribution of the tree walk
    int32_t *d_leaf_surfaces = nullptr;
    int32_t n_leaf_surfaces = 0;
    int32_t *d_unbounded = nullptr;
    int32_t n_unbounded = 0;
    int32_t *d_cand = nullptr;         // leaf surfaces followed by unbounded surfaces (the walk's candidate list)
    int num_sms = 148;
    int device = 0;                    // CUDA device the tables live on
    int walk_occ[6] = {0, 0, 0, 0, 0, 0};    // resident CTAs per SM of each tree-walk instantiation (cached query)
    bool flat_only = false;            // every surface is of the flat family (kinds <= TRC_G_TRIANGLE)
    float4 *d_tri32 = nullptr;         // float32 copies of the triangles (3 x float4 per surface) for the walk's
                                       // conservative pre-test; null unless the scene is flat-only and has triangles
    int32_t n_triangles = 0;
    double *d_tri64 = nullptr;         // compact float64 triangle table (TRC_TRI_REC_DOUBLES per surface) for the walk's
                                       // float64 stage; null unless the scene is flat-only and has triangles
    bool has_refractive = false;       // some surface is TRC_O_REFRACTIVE: the optics need the ref_index column
    bool has_split = false;
    bool has_tabulated = false;        // some TRC_O_REFRACTIVE surface takes its indices from material tables
    bool has_grid = false;             // a shadow grid is set (use_bvh == 2 routes coherent rays through it)
    TrcShadowGrid grid = {};
    uint32_t *d_grid_bits = nullptr, *d_grid_cells = nullptr;
    int32_t *d_grid_list = nullptr;
    int grid_occ[3] = {0, 0, 0};       // resident CTAs per SM of the shadow-grid kernel's instantiations (cached query)
    uint32_t *d_dirty = nullptr;       // chunks the shadow-grid kernel leaves to the tree walk (dirty_cap entries)
    int64_t dirty_cap = 0;
    unsigned long long *d_dirty_counters = nullptr;   // [entries of d_dirty, work counter of the walk over them]
    int64_t *h_pinned = nullptr;       // page-locked host words for the count read-backs of trc_bounce: [hits, culled, seg(S+1)]
    cudaEvent_t ev_walk[2] = {nullptr, nullptr};   // reusable event pair around the bounce-0 nearest-hit kernel (time_trace)
};

// Optics launch of the device-driven trace (bounce.cu).
struct TrcOpticsDyn {
    TrcBounceCtl *ctl, *ctl_next;
    TrcBundleDesc children, children_sorted;   // where this bounce's children are written: slot order / survivors first
    uint32_t *cta_surv;
    double *absorbed_per_surface;
    int64_t *d_counts;
    const int64_t *seg;            // nullable: per-surface slot offsets to append to d_counts (n_seg words)
    int n_seg;
};
int trc_optics_dyn(const TrcScene *sc, const TrcOpticsArgs *a, const TrcOpticsDyn *dyn, void *stream);

// 32-bit words of shared memory the tree-walk kernel can spare for a shadow grid's occupancy bits (nearest.cu)
int trc_walk_grid_word_budget(const TrcScene *sc);
void trc_scene_free_shadow_grid(TrcScene *sc);

// Nearest-hit launch of the device-driven trace (nearest.cu): the bundle (pointers and ray count), the work
// counter and the hit total live in the control block; n_cap sizes the grid.
int trc_nearest_dyn(const TrcScene *sc, int64_t n_cap, TrcBounceCtl *ctl, int32_t use_bvh, TrcHitRecord *d_records,
                    int32_t *d_keys, int64_t capacity, uint32_t *d_chunk_off, uint32_t *d_chunk_cnt, cudaStream_t st);
