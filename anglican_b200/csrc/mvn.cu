// mvn.cu — multivariate-normal kernels on the fp64 tensor pipe (DMMA).
//
// Reference: `mvn` of anglican.runtime (src/anglican/runtime.cljc:349-368)
//   observe* = -1/2 |iL (x - mu)|^2 - Z        (364-366; iL = inverse Cholesky factor)
//   sample*  = mu + L z                         (360-361)
// evaluated per particle with core.matrix `mmul`.  Here a warp owns 8 particles
// and multiplies their 8 x d row block by the d x d factor with
// mma.sync.m8n8k4.f64 (the only tensor-core shape that computes in fp64; the
// reference arithmetic is fp64 and parity is 1e-9 — tcgen05 has no fp64 kind).
// The factors live in shared memory for the whole kernel; triangular factors
// skip the k-steps above the diagonal.
//
// k_propagate_mvn is K1 of the fused sweep for the d-dimensional linear-Gaussian
// state-space program (BASELINE config 5):
//     x' = A x[parent] + Lq z,   lw = base + log N(y; x', R)
// with the ancestor gather folded into the parent-row read.
#include <math.h>

#include "apop_internal.cuh"
#include "rng.cuh"

namespace apop {

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// acc[j][0..1] += T (8 x D, row-major in smem, leading dim LD) times M^T, where M
// (D x D, row-major in smem, leading dim LD) — i.e. out[p][r] = sum_c T[p][c] M[r][c].
// LOWER: M is lower triangular, k-steps beyond the diagonal block are skipped.
// Fragment layout of m8n8k4: A a[lane/4][lane%4], B b[k=lane%4][n=lane/4],
// C c[lane/4][2*(lane%4) + {0,1}].
template <int D, int LD, bool LOWER>
__device__ __forceinline__ void warp_gemm_8xD(const double* __restrict__ T,
                                              const double* __restrict__ M,
                                              double (&acc)[D / 8][2]) {
    const int lane = threadIdx.x & 31;
    const int r = lane >> 2, c = lane & 3;
#pragma unroll
    for (int ks = 0; ks < D / 4; ++ks) {
        const double a = T[r * LD + 4 * ks + c];
#pragma unroll
        for (int j = 0; j < D / 8; ++j) {
            if (LOWER && 4 * ks > 8 * j + 7) continue;          // above the diagonal: zeros
            const double b = M[(8 * j + r) * LD + 4 * ks + c];
            dmma_m8n8k4(acc[j][0], acc[j][1], a, b);
        }
    }
}

template <int D>
struct MvnSmem {
    static constexpr int LD = D + 4;          // rows shifted by 8 banks: conflict-free fragments
};

// one warp = 8 particles per iteration; persistent CTAs
template <int D>
__global__ void __launch_bounds__(512, 1) k_propagate_mvn(MvnSweepArgs a) {
    constexpr int LD = MvnSmem<D>::LD;
    constexpr int NW = 16;
    extern __shared__ double sh[];
    double* sA = sh;                       // D x LD
    double* sLq = sA + D * LD;
    double* siL = sLq + D * LD;
    double* sy = siL + D * LD;             // D
    double* tiles = sy + D;                // NW x (8 x LD)
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < D * D; i += blockDim.x) {
        const int r = i / D, c = i - r * D;
        sA[r * LD + c] = a.A[i];
        sLq[r * LD + c] = a.Lq[i];
        siL[r * LD + c] = a.iL[i];
    }
    for (int i = threadIdx.x; i < D; i += blockDim.x) sy[i] = a.y[i];
    __syncthreads();
    double* T = tiles + wid * (8 * LD);
    const double wbase = a.sd ? a.sd->base : 0.0;
    const int64_t ngroups = (a.n + 7) / 8;
    const int r8 = lane >> 2, c4 = lane & 3;
    for (int64_t g = (int64_t)blockIdx.x * NW + wid; g < ngroups; g += (int64_t)gridDim.x * NW) {
        const int64_t k0 = g * 8;
        // ---- parent rows -> tile (ancestor gather folded into the read)
        // In the sorted order consecutive slots often share their parent (always,
        // when the weights are degenerate): a row equal to the previous one is
        // copied inside the tile instead of being fetched again.  Rows of a peer's
        // shard come over NVLink with ordinary (L1-cacheable) loads: they were
        // written by an earlier kernel and the fabric barrier kernel separates
        // the two, so a kernel-lifetime L1 line can never be stale.
        int64_t par_prev = -1;
#pragma unroll 1
        for (int p = 0; p < 8; ++p) {
            const int64_t k = k0 + p;
            int64_t par = 0;
            if (k < a.n) par = a.first_step ? k : (int64_t)a.anc[k];
            if (p > 0 && par == par_prev && k < a.n) {
                for (int cidx = lane; cidx < D; cidx += 32) T[p * LD + cidx] = T[(p - 1) * LD + cidx];
            } else if (a.world > 1 && !a.first_step) {
                // global parent id: the shard that owns it, local or a peer's heap
                const int q = (int)(par / a.n);
                const double* src = (q == (int)(a.g0 / a.n)) ? a.xp + (par - a.g0) * D
                                                             : a.xpeer.p[q] + (par - (int64_t)q * a.n) * D;
                for (int cidx = lane; cidx < D; cidx += 32)
                    T[p * LD + cidx] = (k < a.n) ? src[cidx] : 0.0;
            } else {
                const double* src = a.xp + par * D;
                for (int cidx = lane; cidx < D; cidx += 32)
                    T[p * LD + cidx] = (k < a.n) ? src[cidx] : 0.0;
            }
            par_prev = (k < a.n) ? par : -1;
        }
        __syncwarp();
        double acc[D / 8][2];
#pragma unroll
        for (int j = 0; j < D / 8; ++j) { acc[j][0] = 0.0; acc[j][1] = 0.0; }
        warp_gemm_8xD<D, LD, false>(T, sA, acc);                       // A x
        __syncwarp();
        // ---- noise: component c of particles (2q, 2q+1) share one Philox block
        //      (normal_at, rng.cuh); 4 pairs x D components per warp
        for (int i = lane; i < 4 * D; i += 32) {
            const int q = i / D, cidx = i - q * D;
            double z0, z1;
            normal_pair_at(a.seed, (uint64_t)(a.g0 + k0 + 2 * q) >> 1, a.t, STREAM_TRANSITION,
                           (uint32_t)cidx, z0, z1);
            T[(2 * q) * LD + cidx] = z0;
            T[(2 * q + 1) * LD + cidx] = z1;
        }
        __syncwarp();
        warp_gemm_8xD<D, LD, true>(T, sLq, acc);                       // + Lq z  (runtime.cljc:360-361)
        __syncwarp();
        // ---- x' to the tile (row-major), conditional slot 0, coalesced store
#pragma unroll
        for (int j = 0; j < D / 8; ++j) {
            T[r8 * LD + 8 * j + 2 * c4] = acc[j][0];
            T[r8 * LD + 8 * j + 2 * c4 + 1] = acc[j][1];
        }
        __syncwarp();
        if (a.retained && a.g0 + k0 == 0)
            for (int cidx = lane; cidx < D; cidx += 32) T[cidx] = a.retained[cidx];
        __syncwarp();
#pragma unroll 1
        for (int p = 0; p < 8; ++p) {
            const int64_t k = k0 + p;
            if (k >= a.n) break;
            double* dst = a.xn + k * D;
            for (int cidx = lane; cidx < D; cidx += 32) {
                const double v = T[p * LD + cidx];
                dst[cidx] = v;
                T[p * LD + cidx] = sy[cidx] - v;                       // y - x'  (observe, mean = x')
            }
        }
        __syncwarp();
        // ---- whitened residual and log-density (runtime.cljc:364-366)
#pragma unroll
        for (int j = 0; j < D / 8; ++j) { acc[j][0] = 0.0; acc[j][1] = 0.0; }
        warp_gemm_8xD<D, LD, true>(T, siL, acc);
        double q = 0.0;
#pragma unroll
        for (int j = 0; j < D / 8; ++j) q += acc[j][0] * acc[j][0] + acc[j][1] * acc[j][1];
        q += __shfl_xor_sync(0xffffffffu, q, 1);
        q += __shfl_xor_sync(0xffffffffu, q, 2);
        if (c4 == 0 && k0 + r8 < a.n) a.lw[k0 + r8] = wbase + (-0.5 * q - a.Z);
        __syncwarp();
    }
}

// x0[k][i] = s0 * normal_at(seed, k, 0, STREAM_INIT, i)
__global__ void k_init_mvn(double* x, int64_t n, int d, double s0, int64_t g0, uint64_t seed,
                           const double* retained) {
    const int64_t total = n * (int64_t)d;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / d;
        const int c = (int)(i - k * d);
        double v = s0 * normal_at(seed, (uint64_t)(g0 + k), 0, STREAM_INIT, (uint32_t)c);
        if (retained && g0 + k == 0) v = retained[c];
        x[i] = v;
    }
}

// column-level observe: lw[k] += -1/2 |iL (val_k - mean_k)|^2 - Z, tensor-core path
template <int D>
__global__ void __launch_bounds__(256) k_observe_mvn_dmma(const double* __restrict__ iL, double Z,
                                                          const double* __restrict__ mean,
                                                          int mstride,
                                                          const double* __restrict__ val,
                                                          int vstride, double* lw, int64_t n) {
    constexpr int LD = MvnSmem<D>::LD;
    constexpr int NW = 8;
    extern __shared__ double sh[];
    double* siL = sh;
    double* tiles = siL + D * LD;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < D * D; i += blockDim.x) {
        const int r = i / D, c = i - r * D;
        siL[r * LD + c] = iL[i];
    }
    __syncthreads();
    double* T = tiles + wid * (8 * LD);
    const int r8 = lane >> 2, c4 = lane & 3;
    const int64_t ngroups = (n + 7) / 8;
    for (int64_t g = (int64_t)blockIdx.x * NW + wid; g < ngroups; g += (int64_t)gridDim.x * NW) {
        const int64_t k0 = g * 8;
#pragma unroll 1
        for (int p = 0; p < 8; ++p) {
            const int64_t k = k0 + p;
            for (int cidx = lane; cidx < D; cidx += 32)
                T[p * LD + cidx] = (k < n) ? val[(int64_t)vstride * k + cidx] -
                                                 mean[(int64_t)mstride * k + cidx]
                                           : 0.0;
        }
        __syncwarp();
        double acc[D / 8][2];
#pragma unroll
        for (int j = 0; j < D / 8; ++j) { acc[j][0] = 0.0; acc[j][1] = 0.0; }
        warp_gemm_8xD<D, LD, true>(T, siL, acc);
        double q = 0.0;
#pragma unroll
        for (int j = 0; j < D / 8; ++j) q += acc[j][0] * acc[j][0] + acc[j][1] * acc[j][1];
        q += __shfl_xor_sync(0xffffffffu, q, 1);
        q += __shfl_xor_sync(0xffffffffu, q, 2);
        if (c4 == 0 && k0 + r8 < n) lw[k0 + r8] += -0.5 * q - Z;
        __syncwarp();
    }
}

// ------------------------------------------------------------------ host side
template <int D>
static int launch_propagate(apop_ctx* c, const MvnSweepArgs& a) {
    constexpr int LD = MvnSmem<D>::LD;
    const size_t smem = sizeof(double) * (size_t)(3 * D * LD + D + 16 * 8 * LD);
    static bool configured[64] = {};              // the attribute is per device
    if (c->device < 0 || c->device >= 64 || !configured[c->device]) {
        APOP_CUDA(c, cudaFuncSetAttribute(k_propagate_mvn<D>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (c->device >= 0 && c->device < 64) configured[c->device] = true;
    }
    const int sms = c->sm_count;
    const int64_t groups = (a.n + 7) / 8;
    int grid = (int)((groups + 15) / 16);
    if (grid > sms) grid = sms;                   // persistent: one CTA (16 warps) per SM
    k_propagate_mvn<D><<<grid, 512, smem, c->stream>>>(a);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

bool mvn_dim_supported(int d) { return d == 8 || d == 16 || d == 32 || d == 64; }

// resolve the kernels of a sharded mvn sweep before the first launch (lazy module
// loading would wait for an idle device while a peer's kernel spins, ssm.cu)
void mvn_preload_kernels(int d) {
    cudaFuncAttributes a;
    cudaFuncGetAttributes(&a, k_init_mvn);
    switch (d) {
    case 8: cudaFuncGetAttributes(&a, k_propagate_mvn<8>); break;
    case 16: cudaFuncGetAttributes(&a, k_propagate_mvn<16>); break;
    case 32: cudaFuncGetAttributes(&a, k_propagate_mvn<32>); break;
    case 64: cudaFuncGetAttributes(&a, k_propagate_mvn<64>); break;
    default: break;
    }
    cudaGetLastError();
}

int mvn_propagate_enqueue(apop_ctx* c, int d, const MvnSweepArgs& a) {
    switch (d) {
    case 8: return launch_propagate<8>(c, a);
    case 16: return launch_propagate<16>(c, a);
    case 32: return launch_propagate<32>(c, a);
    case 64: return launch_propagate<64>(c, a);
    default:
        return set_err(c, APOP_ERR_UNSUPPORTED,
                       "fused mvn sweep: state dimension must be 8, 16, 32 or 64 (got %d)", d);
    }
}

int mvn_init_enqueue(apop_ctx* c, double* x, int d, double s0, const double* retained) {
    const int64_t total = c->n * (int64_t)d;
    int blocks = (int)((total + 255) / 256);
    if (blocks > c->sm_count * 16) blocks = c->sm_count * 16;
    k_init_mvn<<<blocks, 256, 0, c->stream>>>(x, c->n, d, s0, c->g0, c->seed, retained);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

template <int D>
static int launch_observe(apop_ctx* c, const double* iL, double Z, const double* mean, int ms,
                          const double* val, int vs, double* lw) {
    constexpr int LD = MvnSmem<D>::LD;
    const size_t smem = sizeof(double) * (size_t)(D * LD + 8 * 8 * LD);
    static bool configured = false;
    if (!configured) {
        APOP_CUDA(c, cudaFuncSetAttribute(k_observe_mvn_dmma<D>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    const int64_t groups = (c->n + 7) / 8;
    int grid = (int)((groups + 7) / 8);
    if (grid > c->sm_count * 2) grid = c->sm_count * 2;
    k_observe_mvn_dmma<D><<<grid, 256, smem, c->stream>>>(iL, Z, mean, ms, val, vs, lw, c->n);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

// tensor-core observe for the supported dimensions; false -> caller uses the generic kernel
int mvn_observe_dmma_enqueue(apop_ctx* c, int d, const double* iL, double Z, const double* mean,
                             int ms, const double* val, int vs, double* lw) {
    switch (d) {
    case 8: return launch_observe<8>(c, iL, Z, mean, ms, val, vs, lw);
    case 16: return launch_observe<16>(c, iL, Z, mean, ms, val, vs, lw);
    case 32: return launch_observe<32>(c, iL, Z, mean, ms, val, vs, lw);
    case 64: return launch_observe<64>(c, iL, Z, mean, ms, val, vs, lw);
    default: return APOP_ERR_UNSUPPORTED;
    }
}

} // namespace apop
