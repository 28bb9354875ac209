// Euler-Maruyama simulation of dx = f(x) dt + L dB, L = chol(q)  (markovflow/sde/sde_utils.py:36-96), batched:
//   out[b, 0] = x0[b],  out[b, k+1] = out[b, k] + f(out[b, k]) dt_k + sqrt(dt_k) L xi_k,
//   dt_0 = time_grid[0] - 0,  dt_k = time_grid[k] - time_grid[k-1]   (the reference starts its clock at t = 0
//   and returns [x0, first n-1 states], n = len(time_grid)).
// One thread per sequence (the recursion is sequential in time; B = 4096 sequences keep 32 warps per SM busy).
// xi_k: standard normals from the caller (`noise` [B, n-1, D], reproduces a given draw exactly) or from
// Philox4x32-10 (curand device API), sub-sequence = sequence index, so a sequence's path does not depend on B.
#pragma once
#include <curand_kernel.h>

#include "common.cuh"
#include "drifts.cuh"

namespace vdp {

template <int DRIFT, int D>
__global__ void k_euler_maruyama(DriftParams P, Mat<D> cholq, const double* __restrict__ x0,
                                 const double* __restrict__ time_grid, const double* __restrict__ noise,
                                 unsigned long long seed, double* __restrict__ out, int64_t B, int64_t n) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    curandStatePhilox4_32_10_t rng;
    if (!noise) curand_init(seed, (unsigned long long)b, 0ULL, &rng);
    Vec<D> x = vec_load<D>(x0 + b * D);
    double* o = out + b * n * D;
    vec_store<D>(o, x);
    double t = 0.0;
    for (int64_t k = 0; k + 1 < n; ++k) {
        const double tk = time_grid[k];
        const double dt = tk - t;
        t = tk;
        Vec<D> xi;
        if (noise) {
            xi = vec_load<D>(noise + (b * (n - 1) + k) * D);
        } else {
#pragma unroll
            for (int i = 0; i < D; i += 2) {
                const double2 z = curand_normal2_double(&rng);
                xi(i) = z.x;
                if (i + 1 < D) xi(i + 1) = z.y;
            }
        }
        Vec<D> f;
        Mat<D> J;
        Drift<DRIFT, D>::eval(P, x, f, J);
        const double sdt = sqrt(dt);
        Mat<D> Ls;
#pragma unroll
        for (int e = 0; e < D * D; ++e) Ls.a[e] = cholq.a[e] * sdt;
        const Vec<D> dW = mv(Ls, xi);
#pragma unroll
        for (int i = 0; i < D; ++i) x(i) = (x(i) + f(i) * dt) + dW(i);
        vec_store<D>(o + (k + 1) * D, x);
    }
}

}  // namespace vdp
