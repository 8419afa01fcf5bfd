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
#include "normal_table.inc"
namespace apop {

// ---- Gaussian draws: table-driven inverse CDF --------------------------------
// The reference draws normals with commons-math nextGaussian on an unseeded
// Well19937c (runtime.cljc:142-147, 403-409), so only distributional parity is
// meaningful.  64 random bits -> z = sign * Q(v), Q = -Phi^-1, v uniform on
// [0, 1/2) built from 52 of the bits: the exponent field and the top NT_LOG2K
// mantissa bits of v select a row of NT_TABLE (tools/gen_normal_table.py), a
// degree-4 polynomial in d = v - (row centre), |fit - Q| < 5e-12.  6 fp64
// instructions and three 16-byte table reads per draw, against 28 fp64
// instructions for fp64 Box-Muller.  v < 2^-(NT_OCT+1) (2^-NT_OCT of the draws)
// takes normal_tail.  The oracle restates the same map from its own copy of the
// table, so whole sweeps compare number for number.
static __device__ const double g_ntab[NT_ROWS * NT_STRIDE] = {NT_TABLE};

// far tail: v extended by 11 more random bits, Q(v) by Newton on erfc from the
// Abramowitz-Stegun 26.2.23 start (3 steps: 4.5e-4 -> 1e-6 -> 1e-12 -> fp64)
static __device__ __noinline__ double normal_tail(uint32_t hi, uint32_t lo) {
    const double v = __hiloint2double((int)((hi & 0x000fffffu) | 0x3fe00000u), (int)lo) - 0.5;
    const double ve = v + ((double)((hi >> 20) & 0x7ffu) + 0.5) * 5.42101086242752217e-20;   // 2^-64
    const double t = sqrt(-2.0 * log(ve));
    double z = t - (2.515517 + t * (0.802853 + t * 0.010328)) /
                       (1.0 + t * (1.432788 + t * (0.189269 + t * 0.001308)));
#pragma unroll 1
    for (int it = 0; it < 3; ++it) {
        const double tailp = 0.5 * erfc(z * 0.70710678118654752440);
        const double dens = exp(-0.5 * z * z) * 0.39894228040143267794;
        z = z + (tailp - ve) / dens;
    }
    return (hi & 0x80000000u) ? -z : z;
}

// tab: NT_TABLE in global memory (g_ntab, COPIES = 1) or replicated in shared memory
// (hot kernels).  Shared-memory layout: the 16-byte word k of row r, copy c, sits at
// double2 index (3 r + k) * COPIES + c; lane l reads copy l % COPIES, so the lanes
// of a quarter warp that share a bank group are 1 in COPIES / 8 instead of all
// (random rows: a 24 KB table read by 32 lanes at random is bank-conflict bound
// otherwise - measured, k_propagate 992 -> 1368 us with a single copy).
// idx < 0: the draw belongs to the tail and the returned value must be replaced by
// normal_tail(hi, lo).
template <int COPIES>
__device__ __forceinline__ double normal_core(const double* __restrict__ tab, uint32_t hi,
                                              uint32_t lo, int& idx, int copy = 0) {
    const double v = __hiloint2double((int)((hi & 0x000fffffu) | 0x3fe00000u), (int)lo) - 0.5;
    const int vh = __double2hiint(v);
    idx = (vh >> (20 - NT_LOG2K)) - ((1022 - NT_OCT) << NT_LOG2K);
    const double mid = __hiloint2double((vh & ~((1 << (20 - NT_LOG2K)) - 1)) | (1 << (19 - NT_LOG2K)), 0);
    const double d = v - mid;                                   // exact
    const double2* row = reinterpret_cast<const double2*>(tab) + (max(idx, 0) * 3 * COPIES + copy);
    const double2 c01 = row[0];
    const double2 c23 = row[COPIES];
    const double c4 = row[2 * COPIES].x;
    double p = fma(c4, d, c23.y);
    p = fma(p, d, c23.x);
    p = fma(p, d, c01.y);
    p = fma(p, d, c01.x);
    return __hiloint2double(__double2hiint(p) ^ (int)(hi & 0x80000000u), __double2loint(p));
}
__device__ __forceinline__ double normal_from_bits(const double* __restrict__ tab, uint32_t hi,
                                                   uint32_t lo) {
    int idx;
    const double z = normal_core<1>(tab, hi, lo, idx);
    return (idx < 0) ? normal_tail(hi, lo) : z;
}

// cooperative copy of the table into shared memory, COPIES times (caller synchronises)
template <int COPIES>
__device__ __forceinline__ void normal_table_to_smem(double* s_tab) {
    const double2* src = reinterpret_cast<const double2*>(g_ntab);
    double2* dst = reinterpret_cast<double2*>(s_tab);
    for (int i = threadIdx.x; i < NT_ROWS * 3; i += blockDim.x) {
        const double2 w = src[i];
#pragma unroll
        for (int c = 0; c < COPIES; ++c) dst[i * COPIES + c] = w;
    }
}

// normals of the slot pair (2p, 2p+1) come from ONE Philox block: words (x, y)
// -> even slot, (z, w) -> odd slot
__device__ __forceinline__ void normal_pair_at(const double* __restrict__ tab, uint64_t seed,
                                               uint64_t pair, uint32_t t, uint32_t stream,
                                               uint32_t draw, double& z0, double& z1) {
    U4 o = philox4x32_10((uint32_t)pair, (uint32_t)(pair >> 32), t, ctr3(stream, 0, draw),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
    z0 = normal_from_bits(tab, o.x, o.y);
    z1 = normal_from_bits(tab, o.z, o.w);
}
__device__ __forceinline__ void normal_pair_at(uint64_t seed, uint64_t pair, uint32_t t,
                                               uint32_t stream, uint32_t draw, double& z0,
                                               double& z1) {
    normal_pair_at(g_ntab, seed, pair, t, stream, draw, z0, z1);
}
// NP consecutive pairs (slots 2 pair0 .. 2 pair0 + 2 NP - 1) from a table replicated
// COPIES times in shared memory.  One tail test for the whole batch: the fix-up
// (2^-16 of the draws) recomputes the Philox block.
template <int NP, int COPIES>
__device__ __forceinline__ void normal_pairs_at(const double* __restrict__ tab, uint64_t seed,
                                                uint64_t pair0, uint32_t t, uint32_t stream,
                                                uint32_t draw, double (&z0)[NP], double (&z1)[NP]) {
    int any = 0;
    const int copy = (int)(threadIdx.x & (COPIES - 1));
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const uint64_t pair = pair0 + i;
        U4 o = philox4x32_10((uint32_t)pair, (uint32_t)(pair >> 32), t, ctr3(stream, 0, draw),
                             (uint32_t)seed, (uint32_t)(seed >> 32));
        int i0, i1;
        z0[i] = normal_core<COPIES>(tab, o.x, o.y, i0, copy);
        z1[i] = normal_core<COPIES>(tab, o.z, o.w, i1, copy);
        any |= i0 | i1;
    }
    if (any < 0) {
#pragma unroll 1
        for (int i = 0; i < NP; ++i) {
            double a, b;
            normal_pair_at(g_ntab, seed, pair0 + i, t, stream, draw, a, b);
            z0[i] = a; z1[i] = b;
        }
    }
}
__device__ __forceinline__ double normal_at(uint64_t seed, uint64_t slot, uint32_t t,
                                            uint32_t stream, uint32_t draw) {
    const uint64_t pair = slot >> 1;
    U4 o = philox4x32_10((uint32_t)pair, (uint32_t)(pair >> 32), t, ctr3(stream, 0, draw),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
    return (slot & 1) ? normal_from_bits(g_ntab, o.z, o.w) : normal_from_bits(g_ntab, o.x, o.y);
}
#endif

} // namespace apop
