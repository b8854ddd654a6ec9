// Shared host-side helpers of the C ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
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

// Tree node as the walk reads it: 32 bytes, the two children of an inner node
// are adjacent and 64-byte aligned (root at 0, slot 1 unused, pairs from 2).
// Leaf: count > 0, `first` indexes the leaf-surface list; inner: count == 0,
// children at `first`, `first + 1`.  Empty leaf: inverted box.
struct TrcNode32 {
    float lo[3];
    float hi[3];
    int32_t first;
    int32_t count;
};

// Device-side scene tables; owned by the library (trc_scene_create/destroy).
struct TrcScene {
    int32_t n_surfaces = 0;
    TrcSurfaceRec *d_surfaces = nullptr;
    double *d_tables = nullptr;
    int32_t n_table = 0;
    TrcBvhNode *d_nodes = nullptr;     // float64 boxes as uploaded (kept for inspection)
    int32_t n_nodes = 0;
    TrcNode32 *d_nodes32 = nullptr;    // float32 outward-rounded boxes walked by the kernel (n_nodes + 1 entries)
    float extent[3] = {0.f, 0.f, 0.f}; // max |coordinate| of the root box per axis
    unsigned long long *d_chunk_counter = nullptr;   // dynamic work distribution of the tree walk
    int32_t *d_leaf_surfaces = nullptr;
    int32_t n_leaf_surfaces = 0;
    int32_t *d_unbounded = nullptr;
    int32_t n_unbounded = 0;
    int num_sms = 148;
};
