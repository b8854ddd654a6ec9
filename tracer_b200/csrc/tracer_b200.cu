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
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <new>
#include <vector>

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
static cudaError_t build_tri32(TrcScene *sc);
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
    sc->flat_only = true;
    for (int32_t i = 0; i < n_surfaces; ++i) {
        if (surfaces[i].gkind > TRC_G_TRIANGLE) sc->flat_only = false;
        if (surfaces[i].okind == TRC_O_REFRACTIVE) sc->has_refractive = sc->has_split = true;
        if (surfaces[i].okind == TRC_O_REFRACTIVE && (surfaces[i].flags & TRC_F_TABULATED)) sc->has_tabulated = true;
        if (surfaces[i].okind == TRC_O_PERIODIC) sc->has_split = true;
    }
    cudaError_t e = cudaMalloc(&sc->d_surfaces, sizeof(TrcSurfaceRec) * (size_t)n_surfaces);
    if (e == cudaSuccess)
        e = cudaMemcpy(sc->d_surfaces, surfaces, sizeof(TrcSurfaceRec) * (size_t)n_surfaces, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n_table > 0 && tables) {
        sc->n_table = n_table;
        e = cudaMalloc(&sc->d_tables, sizeof(double) * (size_t)n_table);
        if (e == cudaSuccess) e = cudaMemcpy(sc->d_tables, tables, sizeof(double) * (size_t)n_table, cudaMemcpyHostToDevice);
    }
    int32_t n_prefilter = 0;           // surfaces the walk can pre-test in float32: triangles, rect and round plates
    for (int32_t i = 0; i < n_surfaces; ++i) {
        const int k = surfaces[i].gkind;
        if (k == TRC_G_TRIANGLE) ++sc->n_triangles;
        if (k == TRC_G_TRIANGLE || k == TRC_G_RECT_PLATE || k == TRC_G_RECT_EXTRUDED || k == TRC_G_ROUND_PLATE) ++n_prefilter;
    }
    if (e == cudaSuccess && sc->flat_only) {
        if (n_prefilter > 0 && !getenv("TRC_NO_TRI_PREFILTER"))
            e = cudaMalloc(&sc->d_tri32, sizeof(float4) * 3 * (size_t)n_surfaces);
        if (e == cudaSuccess && sc->n_triangles > 0)      // the walk of a flat-only scene with triangles relies on it
            e = cudaMalloc(&sc->d_tri64, sizeof(double) * TRC_TRI_REC_DOUBLES * (size_t)n_surfaces);
        if (e == cudaSuccess) e = build_tri32(sc);
    }
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&sc->h_pinned, sizeof(int64_t) * (size_t)(n_surfaces + 3), cudaHostAllocDefault);
    int dev = 0;
    if (e == cudaSuccess) e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sc->num_sms, cudaDevAttrMultiProcessorCount, dev);
    sc->device = dev;
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
    cudaFree(sc->d_tri32);
    cudaFree(sc->d_tri64);
    cudaFree(sc->d_nodes);
    cudaFree(sc->d_pair_a);
    cudaFree(sc->d_pair_b);
    cudaFree(sc->d_pair_ab);
    cudaFree(sc->d_cand);
    cudaFree(sc->d_chunk_counter);
    cudaFree(sc->d_leaf_surfaces);
    cudaFree(sc->d_unbounded);
    trc_scene_free_shadow_grid(sc);
    cudaFree(sc->d_dirty);
    cudaFree(sc->d_dirty_counters);
    if (sc->h_pinned) cudaFreeHost(sc->h_pinned);
    if (sc->ev_walk[0]) cudaEventDestroy(sc->ev_walk[0]);
    if (sc->ev_walk[1]) cudaEventDestroy(sc->ev_walk[1]);
    delete sc;
    return TRC_OK;
}

extern "C" int32_t trc_scene_num_surfaces(const TrcScene *sc) { return sc ? sc->n_surfaces : 0; }

// Float32 copy of every triangle / plate for the tree walk's pre-test (nearest.cu, tri_prefilter):
// [p0.x p0.y p0.z e0.x] [e0.y e0.z e1.x e1.y] [e1.z flag - -], p0 = frame origin.
//   flag 1: triangle p0 + u e0 + v e1, u, v >= 0, u + v <= 1, e = R * verts (triangular_face.py:60-61);
//   flag 2: centred parallelogram |u|, |v| <= 1 -- a rect plate (e = half sides along the frame's x, y;
//           flat_surface.py:202-211; the hole of an extruded plate is left to the float64 test) or the
//           bounding square of a round plate (e = Re along x, y; :484-492);
//   flag 0: any other kind (never pre-tested).
__global__ void build_tri32_kernel(const TrcSurfaceRec *s, float4 *out, double *out64, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *f = s[i].frame, *gp = s[i].gp;
    if (out64) {
        double *o = out64 + (size_t)TRC_TRI_REC_DOUBLES * i;
        if (s[i].gkind == TRC_G_TRIANGLE) {
            TriTerms tt = tri_terms(f, gp);
            o[0] = f[3]; o[1] = f[7]; o[2] = f[11]; o[3] = f[2]; o[4] = f[6]; o[5] = f[10];
            o[6] = tt.e0.x; o[7] = tt.e0.y; o[8] = tt.e0.z; o[9] = tt.e1.x; o[10] = tt.e1.y; o[11] = tt.e1.z;
            o[12] = tt.uv; o[13] = tt.n0; o[14] = tt.n1; o[15] = tt.den;
        } else {
            for (int k = 0; k < TRC_TRI_REC_DOUBLES; ++k) o[k] = 0.0;
        }
    }
    if (!out) return;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
    if (s[i].gkind == TRC_G_TRIANGLE) {
        double e0[3], e1[3];
        for (int k = 0; k < 3; ++k) {
            e0[k] = f[4 * k] * gp[0] + f[4 * k + 1] * gp[1] + f[4 * k + 2] * gp[2];
            e1[k] = f[4 * k] * gp[3] + f[4 * k + 1] * gp[4] + f[4 * k + 2] * gp[5];
        }
        a = make_float4((float)f[3], (float)f[7], (float)f[11], (float)e0[0]);
        b = make_float4((float)e0[1], (float)e0[2], (float)e1[0], (float)e1[1]);
        c = make_float4((float)e1[2], 1.f, 0.f, 0.f);
    } else if (s[i].gkind == TRC_G_RECT_PLATE || s[i].gkind == TRC_G_RECT_EXTRUDED || s[i].gkind == TRC_G_ROUND_PLATE) {
        const double hx = gp[0], hy = s[i].gkind == TRC_G_ROUND_PLATE ? gp[0] : gp[1];
        a = make_float4((float)f[3], (float)f[7], (float)f[11], (float)(f[0] * hx));
        b = make_float4((float)(f[4] * hx), (float)(f[8] * hx), (float)(f[1] * hy), (float)(f[5] * hy));
        c = make_float4((float)(f[9] * hy), 2.f, 0.f, 0.f);
    }
    out[3 * (size_t)i] = a; out[3 * (size_t)i + 1] = b; out[3 * (size_t)i + 2] = c;
}

static cudaError_t build_tri32(TrcScene *sc) {
    if (!sc->d_tri32 && !sc->d_tri64) return cudaSuccess;
    build_tri32_kernel<<<(sc->n_surfaces + 255) / 256, 256>>>(sc->d_surfaces, sc->d_tri32, sc->d_tri64, sc->n_surfaces);
    return cudaDeviceSynchronize();
}

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
        if (e == cudaSuccess) e = build_tri32(sc);
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
    cudaFree(sc->d_pair_a); sc->d_pair_a = nullptr;
    cudaFree(sc->d_pair_b); sc->d_pair_b = nullptr; sc->n_pairs = 0;
    cudaFree(sc->d_pair_ab); sc->d_pair_ab = nullptr;
    cudaFree(sc->d_leaf_surfaces); sc->d_leaf_surfaces = nullptr; sc->n_leaf_surfaces = 0;
    cudaFree(sc->d_unbounded); sc->d_unbounded = nullptr; sc->n_unbounded = 0;
    if (n_nodes > 0) {
        TRC_CUDA(cudaMalloc(&sc->d_nodes, sizeof(TrcBvhNode) * (size_t)n_nodes));
        TRC_CUDA(cudaMemcpy(sc->d_nodes, nodes, sizeof(TrcBvhNode) * (size_t)n_nodes, cudaMemcpyHostToDevice));
        sc->n_nodes = n_nodes;
        // Walk copy: inner nodes renumbered breadth-first as "pairs" (the top of the
        // tree is what the kernel caches in shared memory); children boxes quantised
        // to 16 bits on the grid of the root box, rounded outward.
        if (nodes[0].count > TRC_BVH_MAX_LEAF)
            return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: leaf larger than TRC_BVH_MAX_LEAF%s");
        double step[3];
        for (int k = 0; k < 3; ++k) {
            float l = (float)nodes[0].lo[k], hh = (float)nodes[0].hi[k];
            if ((double)l > nodes[0].lo[k]) l = nextafterf(l, -INFINITY);
            if ((double)hh < nodes[0].hi[k]) hh = nextafterf(hh, INFINITY);
            sc->root_lo[k] = l; sc->root_hi[k] = hh;
            float st = (float)(((double)hh - (double)l) / 65535.0 * (1.0 + 1e-6));
            st = nextafterf(st, INFINITY);
            if (!(st > 1e-30f)) st = 1e-30f;
            sc->qstep[k] = st; step[k] = (double)st;
            sc->extent[k] = fmaxf(fabsf(l), fabsf(hh));
        }
        auto meta_of = [&](const TrcBvhNode &nd) -> uint32_t {      // root word, host side only
            if (nd.count < 0) return 7u << 28;                       // empty tree
            if (nd.count == 0) return 0u;                            // inner
            return ((uint32_t)nd.count << 28) | (uint32_t)nd.first;  // the whole tree is one leaf
        };
        auto child_word = [&](const TrcBvhNode &nd, int32_t pair_index) -> uint32_t {
            if (nd.count == 0) return (uint32_t)pair_index;                                       // inner
            if (nd.count < 0) return TRC_CHILD_LEAF;      // empty leaf: inverted box, harmless candidate 0
            if (nd.count == 1)                  // one surface: its index itself, no candidate-list lookup
                return TRC_CHILD_LEAF | (uint32_t)leaf_surfaces[nd.first];
            return TRC_CHILD_LEAF | ((uint32_t)(nd.count - 1) << 22) | (uint32_t)nd.first;        // leaf
        };
        std::vector<int32_t> pair_of((size_t)n_nodes, -1), order;   // inner node -> pair index
        order.reserve((size_t)n_nodes);
        if (nodes[0].count == 0) { pair_of[0] = 0; order.push_back(0); }
        for (size_t h = 0; h < order.size(); ++h) {
            const TrcBvhNode &nd = nodes[order[h]];
            if (nd.first < 1 || nd.first + 1 >= n_nodes)
                return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: child index out of range%s");
            for (int c = 0; c < 2; ++c) {
                int32_t ci = nd.first + c;
                const TrcBvhNode &ch = nodes[ci];
                if (ch.count > TRC_BVH_MAX_LEAF)
                    return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: leaf larger than TRC_BVH_MAX_LEAF%s");
                if (ch.count > 0 && (ch.first < 0 || ch.first + ch.count > n_leaf_surfaces))
                    return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: leaf range out of bounds%s");
                if (ch.count == 0) {
                    if (pair_of[ci] >= 0 || order.size() >= (size_t)n_nodes)
                        return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: node referenced twice%s");
                    pair_of[ci] = (int32_t)order.size();
                    order.push_back(ci);
                }
            }
        }
        if (order.size() >= (1u << 28)) return fail(TRC_ERR_CAPACITY, "trc_scene_set_bvh: too many nodes%s");
        sc->n_pairs = (int32_t)order.size();
        sc->root_meta = meta_of(nodes[0]);
        std::vector<TrcPairHalf> pa(order.size() + 1), pb(order.size() + 1);
        for (size_t h = 0; h < order.size(); ++h) {
            const TrcBvhNode &nd = nodes[order[h]];
            uint32_t ax[6];          // per child and axis: lo | hi << 16
            uint32_t word[2];
            for (int c = 0; c < 2; ++c) {
                const TrcBvhNode &ch = nodes[nd.first + c];
                word[c] = child_word(ch, ch.count == 0 ? pair_of[nd.first + c] : 0);
                for (int k = 0; k < 3; ++k) {
                    double ql = floor((ch.lo[k] - (double)sc->root_lo[k]) / step[k]);
                    double qh = ceil((ch.hi[k] - (double)sc->root_lo[k]) / step[k]);
                    if (!(ql >= 0.0)) ql = 0.0;
                    if (ql > 65535.0) ql = 65535.0;
                    if (!(qh <= 65535.0)) qh = 65535.0;
                    if (qh < 0.0) qh = 0.0;
                    if (ch.count < 0) { ql = 65535.0; qh = 0.0; }      // empty leaf: never entered
                    ax[3 * c + k] = (uint32_t)ql | ((uint32_t)qh << 16);
                }
            }
            pa[h].w[0] = ax[0]; pa[h].w[1] = ax[1]; pa[h].w[2] = ax[2]; pa[h].w[3] = ax[3];
            pb[h].w[0] = ax[4]; pb[h].w[1] = ax[5]; pb[h].w[2] = word[0]; pb[h].w[3] = word[1];
        }
        TRC_CUDA(cudaMalloc(&sc->d_pair_a, sizeof(TrcPairHalf) * pa.size()));
        TRC_CUDA(cudaMalloc(&sc->d_pair_b, sizeof(TrcPairHalf) * pb.size()));
        TRC_CUDA(cudaMemcpy(sc->d_pair_a, pa.data(), sizeof(TrcPairHalf) * pa.size(), cudaMemcpyHostToDevice));
        TRC_CUDA(cudaMemcpy(sc->d_pair_b, pb.data(), sizeof(TrcPairHalf) * pb.size(), cudaMemcpyHostToDevice));
        {
            std::vector<TrcPairHalf> ab(2 * pa.size());
            for (size_t k = 0; k < pa.size(); ++k) { ab[2 * k] = pa[k]; ab[2 * k + 1] = pb[k]; }
            TRC_CUDA(cudaMalloc(&sc->d_pair_ab, sizeof(TrcPairHalf) * ab.size()));
            TRC_CUDA(cudaMemcpy(sc->d_pair_ab, ab.data(), sizeof(TrcPairHalf) * ab.size(), cudaMemcpyHostToDevice));
        }
    }
    if ((int64_t)n_leaf_surfaces + n_unbounded >= (1 << 22) || sc->n_surfaces >= (1 << 22))
        return fail(TRC_ERR_CAPACITY, "trc_scene_set_bvh: more than 2^22 candidate entries%s");
    cudaFree(sc->d_cand); sc->d_cand = nullptr;
    {
        std::vector<int32_t> cand;
        cand.reserve((size_t)n_leaf_surfaces + (size_t)n_unbounded + 1);
        for (int32_t i = 0; i < n_leaf_surfaces; ++i) cand.push_back(leaf_surfaces[i]);
        for (int32_t i = 0; i < n_unbounded; ++i) cand.push_back(unbounded[i]);
        for (int32_t v : cand)
            if (v < 0 || v >= sc->n_surfaces) return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: surface index out of range%s");
        if (cand.empty()) cand.push_back(0);
        TRC_CUDA(cudaMalloc(&sc->d_cand, sizeof(int32_t) * cand.size()));
        TRC_CUDA(cudaMemcpy(sc->d_cand, cand.data(), sizeof(int32_t) * cand.size(), cudaMemcpyHostToDevice));
    }
    if (!sc->d_chunk_counter) TRC_CUDA(cudaMalloc(&sc->d_chunk_counter, sizeof(unsigned long long)));
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


// ------------------------------------------------------------ pipe peaks
// FMA micro-benchmark (MEASURED_PEAKS.json has no scalar-pipe figure): every thread runs eight
// independent chains of dependent fused multiply-adds; returns the sustained rate in TFLOP/s (2 flop per FMA).
template <typename T>
__global__ void __launch_bounds__(256)
fma_peak_kernel(T *out, int iters, T a, T b) {
    T x0 = (T)threadIdx.x, x1 = x0 + (T)1, x2 = x0 + (T)2, x3 = x0 + (T)3;
    T x4 = x0 + (T)4, x5 = x0 + (T)5, x6 = x0 + (T)6, x7 = x0 + (T)7;
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
        x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
    }
    T s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == (T)123456789) out[0] = s;          // never true in practice: keeps the chains alive
}

extern "C" int trc_fma_peak(int32_t fp64, double *tflops) {
    if (!tflops) return fail(TRC_ERR_BAD_ARGUMENT, "trc_fma_peak: bad argument%s");
    int dev = 0, sms = 0;
    TRC_CUDA(cudaGetDevice(&dev));
    TRC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    void *out = nullptr;
    TRC_CUDA(cudaMalloc(&out, 64));
    cudaEvent_t e0, e1;
    TRC_CUDA(cudaEventCreate(&e0));
    TRC_CUDA(cudaEventCreate(&e1));
    const int blocks = sms * 8, iters = fp64 ? 1 << 14 : 1 << 16;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {           // first repetition warms up
        cudaEventRecord(e0);
        if (fp64) fma_peak_kernel<double><<<blocks, 256>>>((double *)out, iters, 1.0000001, 1e-9);
        else fma_peak_kernel<float><<<blocks, 256>>>((float *)out, iters, 1.0000001f, 1e-9f);
        cudaEventRecord(e1);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { cudaFree(out); return fail(TRC_ERR_CUDA, "trc_fma_peak: %s", cudaGetErrorString(e)); }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    trc_note_launches(4);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    const double flop = 2.0 * 8.0 * (double)iters * 256.0 * (double)blocks;
    *tflops = flop / ((double)best * 1e-3) / 1e12;
    return TRC_OK;
}
