// Gauss-Hermite quadrature core (gpflow.quadrature.mvnquad restated; call sites sde.py:109,128,515 and
// sde_utils.py:247,354).  Nodes are the Cartesian product of the 1-D hermgauss(H) rule, last
// coordinate fastest (itertools.product order); node weight = ((w_i w_j) w_l) pi^(-D/2);
// X = sqrt(2) (chol(S) xi) + m.  Tables live in constant memory.
//
// A quadrature over one grid point is shared by a group of G lanes (G = 1: one thread per grid
// point, used for D = 1; G = 32: one warp per grid point, lanes stride over the H^D nodes, used for
// D >= 2).  Partial sums are combined with a shuffle butterfly so every lane ends with the total.
#pragma once
#include "common.cuh"
#include "gh_tables.h"

namespace vdp {

static __constant__ double c_gh_x10[10] = VDP_GH_X_10;
static __constant__ double c_gh_w10[10] = VDP_GH_W_10;
static __constant__ double c_gh_x20[20] = VDP_GH_X_20;
static __constant__ double c_gh_w20[20] = VDP_GH_W_20;
// low-order rules: an H-point rule is exact up to degree 2H - 1, so for a polynomial drift of degree p every
// integrand of the KL term and of its adjoint (degree <= 2p) gets from the (p + 1)-point rule the value the
// reference's 20-point rule gives (DriftPoly in drifts.cuh)
static __constant__ double c_gh_x2[2] = VDP_GH_X_2;
static __constant__ double c_gh_w2[2] = VDP_GH_W_2;
static __constant__ double c_gh_x3[3] = VDP_GH_X_3;
static __constant__ double c_gh_w3[3] = VDP_GH_W_3;
static __constant__ double c_gh_x4[4] = VDP_GH_X_4;
static __constant__ double c_gh_w4[4] = VDP_GH_W_4;

template <int H> __device__ __forceinline__ double gh_x(int i) {
    static_assert(H == 2 || H == 3 || H == 4 || H == 10 || H == 20, "no Gauss-Hermite table of this order");
    if constexpr (H == 2) return c_gh_x2[i];
    else if constexpr (H == 3) return c_gh_x3[i];
    else if constexpr (H == 4) return c_gh_x4[i];
    else if constexpr (H == 10) return c_gh_x10[i];
    else return c_gh_x20[i];
}
template <int H> __device__ __forceinline__ double gh_w(int i) {
    if constexpr (H == 2) return c_gh_w2[i];
    else if constexpr (H == 3) return c_gh_w3[i];
    else if constexpr (H == 4) return c_gh_w4[i];
    else if constexpr (H == 10) return c_gh_w10[i];
    else return c_gh_w20[i];
}

// The same tables in global memory for lookups whose index differs from lane to lane (D >= 2: the lanes of a
// group hold different nodes).  A constant-bank read is serialised once per distinct address in the warp (up
// to 20 replays per lookup here); a read-only global load of a 160-byte table is one or two L1 wavefronts.
static __device__ const double d_gh_x10[10] = VDP_GH_X_10;
static __device__ const double d_gh_w10[10] = VDP_GH_W_10;
static __device__ const double d_gh_x20[20] = VDP_GH_X_20;
static __device__ const double d_gh_w20[20] = VDP_GH_W_20;
template <int H> __device__ __forceinline__ double gh_x_div(int i) {
    if constexpr (H <= 4) return gh_x<H>(i);      // (2-4 entries: the constant bank replays are few)
    else return __ldg(H == 10 ? &d_gh_x10[i] : &d_gh_x20[i]);
}
template <int H> __device__ __forceinline__ double gh_w_div(int i) {
    if constexpr (H <= 4) return gh_w<H>(i);
    else return __ldg(H == 10 ? &d_gh_w10[i] : &d_gh_w20[i]);
}

template <int H, int D> struct NodeCount { static constexpr int value = H * NodeCount<H, D - 1>::value; };
template <int H> struct NodeCount<H, 0> { static constexpr int value = 1; };

// node p -> xi (standard-normal-scaled abscissae) and product weight (without the pi factor)
template <int H, int D>
__device__ __forceinline__ void gh_node(int p, Vec<D>& xi, double& w) {
    int idx[D];
    int r = p;
#pragma unroll
    for (int d = D - 1; d >= 0; --d) { idx[d] = r % H; r /= H; }
    if constexpr (D == 1) {   // every lane of the warp asks for the same entry: constant-bank broadcast
        w = gh_w<H>(idx[0]);
        xi(0) = gh_x<H>(idx[0]);
    } else {
        w = gh_w_div<H>(idx[0]);
        xi(0) = gh_x_div<H>(idx[0]);
#pragma unroll
        for (int d = 1; d < D; ++d) { w *= gh_w_div<H>(idx[d]); xi(d) = gh_x_div<H>(idx[d]); }
    }
}

// X = sqrt(2) * (L xi) + m
template <int D>
__device__ __forceinline__ Vec<D> gh_point(const Mat<D>& L, const Vec<D>& m, const Vec<D>& xi) {
    const Vec<D> t = mv(tril(L), xi);
    Vec<D> X;
#pragma unroll
    for (int i = 0; i < D; ++i) X(i) = 1.4142135623730951 * t(i) + m(i);
    return X;
}

// (groups narrower than a warp: only the lanes of the group take part -- neighbouring groups of the same warp may
// be on another branch, e.g. the group of a sequence's last grid point, which has no transition term)
template <int G> __device__ __forceinline__ double group_sum(double v) {
    if (G > 1) {
        const unsigned mask = (G >= 32) ? 0xffffffffu : (((1u << (G & 31)) - 1u) << ((threadIdx.x & 31u) & ~(unsigned)(G - 1)));
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, G);
    }
    return v;
}
template <int G, int D> __device__ __forceinline__ void group_sum(Vec<D>& v) {
#pragma unroll
    for (int i = 0; i < D; ++i) v(i) = group_sum<G>(v(i));
}
template <int G, int D> __device__ __forceinline__ void group_sum(Mat<D>& v) {
#pragma unroll
    for (int i = 0; i < D * D; ++i) v.a[i] = group_sum<G>(v.a[i]);
}

// E[f], E[J] under N(m, S)   (SDE.expected_drift / expected_gradient_drift, H = 10 in the reference)
template <int DRIFT, int D, int H, int G>
__device__ __forceinline__ void gh_expect_point(const DriftParams& P, const Vec<D>& m, const Mat<D>& S, int lane,
                                                double pi_factor, Vec<D>& Ef, Mat<D>& EJ, bool& ok) {
    const Mat<D> L = chol(S, ok);
    Ef = vec_zero<D>();
    EJ = mat_zero<D>();
    constexpr int P_ = NodeCount<H, D>::value;
    for (int p = lane; p < P_; p += G) {
        Vec<D> xi, f;
        Mat<D> J;
        double w;
        gh_node<H, D>(p, xi, w);
        w *= pi_factor;
        Drift<DRIFT, D>::eval(P, gh_point(L, m, xi), f, J);
#pragma unroll
        for (int i = 0; i < D; ++i) Ef(i) += f(i) * w;
#pragma unroll
        for (int i = 0; i < D * D; ++i) EJ.a[i] += J.a[i] * w;
    }
    group_sum<G>(Ef);
    group_sum<G>(EJ);
}

// vdp_gh_expect: one group per (m, S) pair.  EJ output is [D, D] for full-Jacobian drifts, else [D]
// holding E[d(sum_i f_i)/dx_j] (column sums), the reference's tape semantics (sde.py:86-90).
template <int DRIFT, int D, int H, int G>
__global__ void k_gh_expect(DriftParams P, const double* __restrict__ m, const double* __restrict__ S,
                            double pi_factor, int full_jac, double* __restrict__ Ef, double* __restrict__ EJ,
                            int64_t count, int64_t* status) {
    const int64_t gid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / G;
    const int lane = threadIdx.x % G;
    if (gid >= count) return;
    constexpr int DD = D * D;
    bool ok = true;
    Vec<D> ef;
    Mat<D> ej;
    gh_expect_point<DRIFT, D, H, G>(P, vec_load<D>(m + gid * D), mat_load<D>(S + gid * DD), lane, pi_factor, ef, ej,
                                    ok);
    if (lane == 0) {
        if (!ok) report_not_pd(status, gid);
        if (Ef) vec_store<D>(Ef + gid * D, ef);
        if (EJ) {
            if (full_jac) {
                mat_store<D>(EJ + gid * DD, ej);
            } else {
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    double s = ej(0, j);
#pragma unroll
                    for (int i = 1; i < D; ++i) s += ej(i, j);
                    EJ[gid * D + j] = s;
                }
            }
        }
    }
}

// linearize_sde + LinearDrift.to_ssm (sde_utils.py:119-179, drift.py:64-116) + optional clip
// (variational_cvi_sde.py:417-430).  One group per transition; path point = grid point k+1.
template <int DRIFT, int D, int H, int G>
__global__ void k_linearize(DriftParams P, Mat<D> q, const double* __restrict__ time_grid,
                            const double* __restrict__ m, const double* __restrict__ S, double pi_factor,
                            int full_jac, double* __restrict__ A, double* __restrict__ bo,
                            double* __restrict__ cholQ, double* __restrict__ A_lin, double* __restrict__ b_lin,
                            int clip, double lo, double hi, int64_t B, int64_t N, int64_t* status) {
    const int64_t t = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / G;
    const int lane = threadIdx.x % G;
    if (t >= B * N) return;
    constexpr int DD = D * D;
    const int64_t k = t % N;
    bool ok = true;
    Vec<D> ef;
    Mat<D> ej;
    const Vec<D> mk = vec_load<D>(m + t * D);
    gh_expect_point<DRIFT, D, H, G>(P, mk, mat_load<D>(S + t * DD), lane, pi_factor, ef, ej, ok);
    if (lane != 0) return;
    if (!ok) report_not_pd(status, t);
    // (element-wise drifts: the reference's A is the [D] tape gradient, only usable for D = 1, sde_utils.py:162-164;
    // for D > 1 the diagonal Jacobian it stands for is kept)
    (void)full_jac;
    Vec<D> bl = ef - mv(ej, mk);
    const double delta = time_grid[k + 1] - time_grid[k];
    // sde.diffusion = chol(q); q passed to to_ssm is chol_q @ chol_q (sde_utils.py:168-171)
    bool okq = true;
    const Mat<D> cq = chol(q, okq);
    const Mat<D> qq = mm(cq, cq);
    Mat<D> At = delta * ej + mat_eye<D>();
    Vec<D> bt = delta * bl;
    const Mat<D> cQ = chol(delta * qq, okq);
    if (!okq) report_not_pd(status, t);   // spectral density (or a grid step <= 0) that is not positive definite
    if (A_lin) mat_store<D>(A_lin + t * DD, At);
    if (b_lin) vec_store<D>(b_lin + t * D, bt);
    if (clip) {
#pragma unroll
        for (int i = 0; i < DD; ++i) At.a[i] = fmin(fmax(At.a[i], lo), hi);
#pragma unroll
        for (int i = 0; i < D; ++i) bt(i) = fmin(fmax(bt(i), lo), hi);
    }
    mat_store<D>(A + t * DD, At);
    vec_store<D>(bo + t * D, bt);
    mat_store<D>(cholQ + t * DD, cQ);
}

}  // namespace vdp
