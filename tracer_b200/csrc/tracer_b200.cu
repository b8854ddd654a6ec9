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
    cudaFree(sc->d_nodes32);
    cudaFree(sc->d_chunk_counter);
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
    cudaFree(sc->d_nodes32); sc->d_nodes32 = nullptr;
    cudaFree(sc->d_leaf_surfaces); sc->d_leaf_surfaces = nullptr; sc->n_leaf_surfaces = 0;
    cudaFree(sc->d_unbounded); sc->d_unbounded = nullptr; sc->n_unbounded = 0;
    if (n_nodes > 0) {
        TRC_CUDA(cudaMalloc(&sc->d_nodes, sizeof(TrcBvhNode) * (size_t)n_nodes));
        TRC_CUDA(cudaMemcpy(sc->d_nodes, nodes, sizeof(TrcBvhNode) * (size_t)n_nodes, cudaMemcpyHostToDevice));
        sc->n_nodes = n_nodes;
        // float32 walk copy: boxes rounded outward, children pairs shifted to even slots
        std::vector<TrcNode32> n32((size_t)n_nodes + 1);
        memset(n32.data(), 0, sizeof(TrcNode32) * n32.size());
        for (int k = 0; k < 3; ++k) sc->extent[k] = 0.f;
        for (int32_t i = 0; i < n_nodes; ++i) {
            TrcNode32 &o = n32[i == 0 ? 0 : (size_t)i + 1];
            const TrcBvhNode &nd = nodes[i];
            if (nd.count < 0 || (nd.count > 0 && nd.count > TRC_BVH_MAX_LEAF)) {
                if (nd.count > 0) return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: leaf larger than TRC_BVH_MAX_LEAF%s");
                for (int k = 0; k < 3; ++k) { o.lo[k] = INFINITY; o.hi[k] = -INFINITY; }
                o.first = 0; o.count = 1;      // empty leaf: the inverted box is never entered
                continue;
            }
            for (int k = 0; k < 3; ++k) {
                float l = (float)nd.lo[k], h = (float)nd.hi[k];
                if ((double)l > nd.lo[k]) l = nextafterf(l, -INFINITY);
                if ((double)h < nd.hi[k]) h = nextafterf(h, INFINITY);
                o.lo[k] = l; o.hi[k] = h;
            }
            o.count = nd.count;
            if (nd.count == 0) {
                if (nd.first < 1 || nd.first + 1 >= n_nodes)
                    return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: child index out of range%s");
                o.first = nd.first + 1;
            } else {
                if (nd.first < 0 || nd.first + nd.count > n_leaf_surfaces)
                    return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_bvh: leaf range out of bounds%s");
                o.first = nd.first;
            }
        }
        for (int k = 0; k < 3; ++k)
            sc->extent[k] = fmaxf(fabsf(n32[0].lo[k]), fabsf(n32[0].hi[k]));
        TRC_CUDA(cudaMalloc(&sc->d_nodes32, sizeof(TrcNode32) * n32.size()));
        TRC_CUDA(cudaMemcpy(sc->d_nodes32, n32.data(), sizeof(TrcNode32) * n32.size(), cudaMemcpyHostToDevice));
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

