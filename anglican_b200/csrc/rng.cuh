// rng.cuh — counter-based RNG of the engine: Philox4x32-10, keyed by the ctx
// seed, counter = (slot, step, stream|kind|draw).  Replaces the reference's
// single globally locked Well19937c (src/anglican/runtime.cljc:142-147), whose
// streams are unseeded and therefore only distributionally comparable.
// Draws cost zero bytes of memory traffic and are independent of how the
// population is sharded over ranks (counters use the GLOBAL slot id).
#pragma once
#include <stdint.h>

namespace apop {

enum { STREAM_INIT = 0, STREAM_TRANSITION = 1, STREAM_RESAMPLE = 2, STREAM_USER = 3,
       STREAM_AUX = 4 };

struct U4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                      uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
#if defined(__CUDA_ARCH__)
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
#else
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    U4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}

__host__ __device__ __forceinline__ uint32_t ctr3(uint32_t stream, uint32_t kind, uint32_t draw) {
    return (stream << 24) | (kind << 20) | (draw & 0xFFFFFu);
}
// 53-bit uniform strictly inside (0,1)
__host__ __device__ __forceinline__ double u01_open(uint32_t hi, uint32_t lo) {
    uint64_t bits = ((uint64_t)(hi >> 5) << 26) | (uint64_t)(lo >> 6);
    return ((double)bits + 0.5) * (1.0 / 9007199254740992.0);
}
// 53-bit uniform in [0,1)
__host__ __device__ __forceinline__ double u01_half(uint32_t hi, uint32_t lo) {
    uint64_t bits = ((uint64_t)(hi >> 5) << 26) | (uint64_t)(lo >> 6);
    return (double)bits * (1.0 / 9007199254740992.0);
}

__host__ __device__ __forceinline__ double uniform_at(uint64_t seed, uint64_t slot, uint32_t t,
                                                       uint32_t stream, uint32_t draw) {
    U4 o = philox4x32_10((uint32_t)slot, (uint32_t)(slot >> 32), t, ctr3(stream, 1, draw),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
    return u01_half(o.x, o.y);
}

#if defined(__CUDACC__)
} // namespace apop
#include "fastmath.cuh"
namespace apop {
// Gaussian draws: fp64 Box-Muller on one Philox block per slot pair (the reference's
// nextGaussian runs on an unseeded Well19937c, runtime.cljc:142-147, 403-409: only
// distributional parity is meaningful).  A table-driven inverse CDF (piecewise degree-4
// polynomial, 24 KB in shared memory) was built and measured in round 2: fewer
// instructions (132 vs 150 per particle) but k_propagate went from 992 to 1341 us -
// the per-lane table reads stall on the shared-memory scoreboard and leave the issue
// slots 36 % busy (profiles/r02_table_normal_experiment.md).  Box-Muller stays.
// Box-Muller pair for the slot pair (2p, 2p+1): z0 -> even slot, z1 -> odd.
__device__ __forceinline__ void normal_pair_at(uint64_t seed, uint64_t pair, uint32_t t,
                                               uint32_t stream, uint32_t draw, double& z0,
                                               double& z1) {
    U4 o = philox4x32_10((uint32_t)pair, (uint32_t)(pair >> 32), t, ctr3(stream, 0, draw),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
    double u1 = u01_open(o.x, o.y);
    double u2 = u01_half(o.z, o.w);
    double r = fm::sqrt_fast(-2.0 * fm::log_core(u1));     // u1 in [2^-54, 1): normal, < 1
    double s, c;
    fm::sincospi_core(u2 + u2, s, c);
    z0 = r * c;
    z1 = r * s;
}
// four consecutive pairs (slots 8p .. 8p+7 when pair0 = 4p): vectorised Box-Muller
template <int NP>
__device__ __forceinline__ void normal_pairs_at(uint64_t seed, uint64_t pair0, uint32_t t,
                                                uint32_t stream, uint32_t draw,
                                                double (&z0)[NP], double (&z1)[NP]) {
    double u1[NP], u2[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const uint64_t pair = pair0 + i;
        U4 o = philox4x32_10((uint32_t)pair, (uint32_t)(pair >> 32), t, ctr3(stream, 0, draw),
                             (uint32_t)seed, (uint32_t)(seed >> 32));
        u1[i] = u01_open(o.x, o.y);
        u2[i] = u01_half(o.z, o.w);
    }
    fm::box_muller_v<NP>(u1, u2, z0, z1);
}
__device__ __forceinline__ double normal_at(uint64_t seed, uint64_t slot, uint32_t t,
                                            uint32_t stream, uint32_t draw) {
    double z0, z1;
    normal_pair_at(seed, slot >> 1, t, stream, draw, z0, z1);
    return (slot & 1) ? z1 : z0;
}
#endif

} // namespace apop
