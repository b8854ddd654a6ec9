// Counter-based RNG for the bounce kernels: Philox4x32-10 (Salmon et al. 2011).
// The reference draws from numpy's global MT19937 stream in (bounce, surface,
// draw-kind, hit) order; reproducing that stream on a
// GPU is pointless, so production runs key Philox by (seed, bounce, slot) and
// parity runs inject the reference's recorded draws instead.
#pragma once
#include <stdint.h>

namespace trc {

struct Philox4 { uint32_t x, y, z, w; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    Philox4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}

// 53-bit uniform in [0, 1) from two 32-bit words (same bit budget as
// numpy's random_sample: 27 + 26 bits).
__device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}

// Two independent uniforms per Philox call.
__device__ __forceinline__ void uniform2(uint64_t seed, uint64_t index, uint32_t stream, uint32_t slot,
                                         double &u0, double &u1) {
    Philox4 r = philox4x32_10((uint32_t)index, (uint32_t)(index >> 32), stream, slot,
                              (uint32_t)seed, (uint32_t)(seed >> 32));
    u0 = u53(r.x, r.y);
    u1 = u53(r.z, r.w);
}

// Two independent standard normals (Box-Muller on (0,1] x [0,1)).  These are the production draws (slope
// errors, conical scattering): they have no parity constraint, only their distribution matters.  The bulk is
// evaluated in float32 (relative error 1e-7 of a draw whose scale is a milliradian); the far tail, where
// 1 - u0 would lose its resolution in float32, takes the float64 path so that it is not truncated.
__device__ __forceinline__ void normal2(uint64_t seed, uint64_t index, uint32_t stream, uint32_t slot,
                                        double &n0, double &n1) {
    double u0, u1;
    uniform2(seed, index, stream, slot, u0, u1);
    const double q = 1.0 - u0;                 // in (0, 1]
    if (q > 1e-3) {
        float r = sqrtf(-2.0f * logf((float)q));
        float s, c;
        sincospif(2.0f * (float)u1, &s, &c);
        n0 = (double)(r * c);
        n1 = (double)(r * s);
    } else {
        double r = sqrt(-2.0 * log(q));
        double s, c;
        sincospi(2.0 * u1, &s, &c);
        n0 = r * c;
        n1 = r * s;
    }
}

}  // namespace trc
