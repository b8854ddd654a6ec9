// Instrumentation: how fast can persistent warps stream a ray bundle (six float64 rows, SoA) the way the shadow-grid
// kernel does -- a warp owns a chunk of 256 consecutive rays and reads it in groups of 32 rays per row -- as a
// function of the groups kept in flight per warp, the warps per SM and the width of a lane's load?  No counterpart in
// the reference; used by scripts/stream_probe.py to choose the kernel's load structure.
#include "common.cuh"

namespace {

struct Ray6 { double a[6]; };

// DEPTH groups in flight per warp (register prefetch, statically indexed: the 8 groups of a chunk are unrolled)
template <int DEPTH, bool STATIC>
__global__ void __launch_bounds__(256, 2)
stream_probe_kernel(const double *__restrict__ V, const double *__restrict__ D, long long ld, long long n,
                    unsigned long long *counter, double *out) {
    const int lane = threadIdx.x & 31;
    double sum = 0.0;
    unsigned long long cs = (unsigned long long)blockIdx.x * 8 + (threadIdx.x >> 5);      // chunks dealt round robin
    for (;; cs += (unsigned long long)gridDim.x * 8) {
        unsigned long long c = cs;
        if (!STATIC) {
            if (lane == 0) c = atomicAdd(counter, 1ULL);
            c = __shfl_sync(0xffffffffu, c, 0);
        }
        const long long base = (long long)c * 256;
        if (base >= n) break;
        Ray6 buf[DEPTH];
#pragma unroll
        for (int g = 0; g < DEPTH; ++g) {
            const long long r = base + 32 * g + lane;
#pragma unroll
            for (int k = 0; k < 3; ++k) { buf[g].a[k] = r < n ? V[k * ld + r] : 0.0; buf[g].a[3 + k] = r < n ? D[k * ld + r] : 0.0; }
        }
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            Ray6 cur = buf[g % DEPTH];
            if (g + DEPTH < 8) {
                const long long r = base + 32 * (g + DEPTH) + lane;
#pragma unroll
                for (int k = 0; k < 3; ++k) { buf[g % DEPTH].a[k] = r < n ? V[k * ld + r] : 0.0; buf[g % DEPTH].a[3 + k] = r < n ? D[k * ld + r] : 0.0; }
            }
            // a little dependent work per group, as the kernel's key arithmetic
            double s = cur.a[0] * cur.a[3] + cur.a[1] * cur.a[4] + cur.a[2] * cur.a[5];
            sum += s;
            if (__any_sync(0xffffffffu, s == 1.2345e300)) sum = 0.0;
        }
    }
    if (sum == 1.2345e300) out[0] = sum;
}

// the same with one cp.async.bulk per row and chunk (2 KB each) into shared memory, two chunks of buffer per warp
__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
stream_probe_bulk_kernel(const double *__restrict__ V, const double *__restrict__ D, long long ld, long long n,
                         unsigned long long *counter, double *out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double (*ring)[6][256] = reinterpret_cast<double (*)[6][256]>(smem + (size_t)warp * 2 * 6 * 256 * 8);
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem + (size_t)WARPS * 2 * 6 * 256 * 8) + 2 * warp;
    if (lane == 0) {
        for (int k = 0; k < 2; ++k)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&bars[k])), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    auto fetch = [&]() -> long long {
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(counter, 1ULL);
        return (long long)__shfl_sync(0xffffffffu, c, 0) * 256;
    };
    auto issue = [&](long long base, int slot) {
        if (base + 256 > n) return;                       // whole chunks only in this probe
        if (lane == 0) {
            const unsigned bar = smem_addr(&bars[slot]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(6u * 2048u) : "memory");
            for (int k = 0; k < 6; ++k) {
                const double *src = (k < 3 ? V + k * ld : D + (k - 3) * ld) + base;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_addr(&ring[slot][k][0])), "l"(src), "r"(2048u), "r"(bar) : "memory");
            }
        }
    };
    double sum = 0.0;
    long long cur = fetch();
    issue(cur, 0);
    int it = 0;
    for (;;) {
        if (cur + 256 > n) break;
        const long long nxt = fetch();
        issue(nxt, (it + 1) & 1);
        const int slot = it & 1;
        unsigned done;
        do {
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                         : "=r"(done) : "r"(smem_addr(&bars[slot])), "r"((unsigned)(it >> 1) & 1u) : "memory");
        } while (!done);
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            const int j = 32 * g + lane;
            double s = ring[slot][0][j] * ring[slot][3][j] + ring[slot][1][j] * ring[slot][4][j] + ring[slot][2][j] * ring[slot][5][j];
            sum += s;
            if (__any_sync(0xffffffffu, s == 1.2345e300)) sum = 0.0;
        }
        __syncwarp();
        cur = nxt;
        ++it;
    }
    if (sum == 1.2345e300) out[0] = sum;
}

}  // namespace

// variant: 1, 2, 4 = groups in flight per warp (register prefetch, 2 CTAs x 8 warps per SM), chunk ids from an atomic
//          counter; 11, 12, 14 = the same with the chunks dealt round robin (no atomic);
//          108 = bulk copies of whole chunk rows, 8 warps per SM (one CTA), atomic counter
extern "C" int trc_stream_probe(int32_t variant, const double *d_vertices, const double *d_directions, int64_t ld, int64_t n,
                                float *ms_out) {
    if (!d_vertices || !d_directions || !ms_out || n <= 0 || ld < n) return fail(TRC_ERR_BAD_ARGUMENT, "trc_stream_probe: bad argument%s");
    int dev = 0, sms = 0;
    TRC_CUDA(cudaGetDevice(&dev));
    TRC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    unsigned long long *counter = nullptr;
    double *out = nullptr;
    TRC_CUDA(cudaMalloc(&counter, 8));
    TRC_CUDA(cudaMalloc(&out, 8));
    cudaEvent_t e0, e1;
    TRC_CUDA(cudaEventCreate(&e0));
    TRC_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaMemset(counter, 0, 8);
        cudaEventRecord(e0);
        switch (variant) {
        case 1: stream_probe_kernel<1, false><<<2 * sms, 256>>>(d_vertices, d_directions, ld, n, counter, out); break;
        case 2: stream_probe_kernel<2, false><<<2 * sms, 256>>>(d_vertices, d_directions, ld, n, counter, out); break;
        case 4: stream_probe_kernel<4, false><<<2 * sms, 256>>>(d_vertices, d_directions, ld, n, counter, out); break;
        case 11: stream_probe_kernel<1, true><<<2 * sms, 256>>>(d_vertices, d_directions, ld, n, counter, out); break;
        case 12: stream_probe_kernel<2, true><<<2 * sms, 256>>>(d_vertices, d_directions, ld, n, counter, out); break;
        case 14: stream_probe_kernel<4, true><<<2 * sms, 256>>>(d_vertices, d_directions, ld, n, counter, out); break;
        case 108: {
            const int warps = variant - 100;
            const size_t smem = (size_t)warps * (2 * 6 * 256 * 8 + 16);
            cudaFuncSetAttribute(stream_probe_bulk_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            stream_probe_bulk_kernel<8><<<sms, 256, smem>>>(d_vertices, d_directions, ld, n, counter, out);
            (void)warps;
            break;
        }
        default:
            cudaFree(counter); cudaFree(out);
            return fail(TRC_ERR_BAD_ARGUMENT, "trc_stream_probe: unknown variant%s");
        }
        cudaEventRecord(e1);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { cudaFree(counter); cudaFree(out); return fail(TRC_ERR_CUDA, "trc_stream_probe: %s", cudaGetErrorString(e)); }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    trc_note_launches(4);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(counter); cudaFree(out);
    *ms_out = best;
    return TRC_OK;
}
