// gamma_draw.cuh — the gamma sampler shared by the distribution kernels
// (gamma, beta, dirichlet, chi-squared, student-t).
#pragma once
#include <math.h>
#include <stdint.h>

#include "rng.cuh"

namespace apop {

// Marsaglia-Tsang (shape >= 1) with the U^(1/shape) boost for shape < 1;
// distributional parity with commons-math's sampler.
__device__ inline double gamma_draw(double shape, double rate, uint64_t seed, uint64_t slot, uint32_t t,
                             uint32_t stream, uint32_t draw_base) {
    double boost = 1.0;
    uint32_t nu = draw_base, nn = draw_base;
    if (shape < 1.0) {
        double u = uniform_at(seed, slot, t, stream, nu++);
        if (u <= 0.0) u = 1.1102230246251565e-16;
        boost = pow(u, 1.0 / shape);
        shape += 1.0;
    }
    const double d = shape - 1.0 / 3.0;
    const double c = 1.0 / (3.0 * sqrt(d));
    for (int it = 0; it < 64; ++it) {
        double x = normal_at(seed, slot * 2, t, stream, nn++);
        double v = 1.0 + c * x;
        v = v * v * v;
        if (v <= 0) continue;
        double x2 = x * x;
        double u = uniform_at(seed, slot, t, stream, nu++);
        if (u < 1.0 - 0.0331 * x2 * x2 || log(u) < 0.5 * x2 + d * (1.0 - v + log(v)))
            return boost * d * v / rate;
    }
    return boost * d / rate;
}


} // namespace apop
