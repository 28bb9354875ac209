// KL[q || p_SDE] along the Gaussian path and its reverse-mode gradient in expectation parameters.
//
// Reference: SDE_SSM_KL_with_grads_wrt_exp_params (sde_utils.py:473-547) -> expectations_to_ssm_params
// (ssm_gaussian_transformations.py:93-178) -> SSM_KL_along_Gaussian_path (sde_utils.py:262-359) ->
// mvnquad(H = 20); the gradient is what tf.GradientTape returns at sde_utils.py:545, i.e. the adjoint
// of the DISCRETISED quadrature and of the D x D Choleskys (derivation: oracle/sde_np.py
// sde_ssm_kl_and_grads, checked there against torch.autograd).
//
// Work decomposition: one group of G lanes per grid point i = (b, k).  The group evaluates transition
// term k (if k < N): its quadrature, constants and adjoint.  A term touches nodes k ("left") and k+1
// ("right"); the right part is handed to the neighbouring group through shared memory, the first
// group of a block is a halo that only produces the right part for the block's first node.
#pragma once
#include "pointwise_kernels.cuh"
#include "quad_core.cuh"

namespace vdp {

template <int D>
struct KlNodePart {   // adjoint contributions of one transition term to one of its two nodes
    Vec<D> gm;        // d term / d m        (before the -2 g_S m fold-in)
    Mat<D> gS;        // d term / d S        (symmetric)
};

template <int D>
struct KlTermOut {
    double value;         // 0.5 (quad + C)
    KlNodePart<D> left, right;
    Mat<D> ges;           // d term / d eta_s[k]
};

// Gradient of one scalar transition term for a polynomial drift (closed-form moments), from the transition A_k, the
// means and variances of its two grid points and 1 / S_k:  d term / d (m_k, S_k) -> (Lgm, LgS), d term / d (m_{k+1},
// S_{k+1}) -> (Rgm, RgS), d term / d eta_s[k] -> gC  (the adjoint chain of kl_term_scalar below, same statements).
template <int DRIFT>
__device__ __forceinline__ void kl_grad_closed_core(const DriftParams& P, double Qp, double dt, double A, double iSk,
                                                    double mk, double mn, double Sk, double Sn, double& Lgm,
                                                    double& LgS, double& Rgm, double& RgS, double& gC, bool& ok) {
    const double b = mn - A * mk;
    const double Qc = Sn - A * (Sk * A);
    if (!(Qc > 0.0) || !(Qp > 0.0)) ok = false;
    const double iQc = rcp_fast(Qc), Qpinv = rcp_fast(Qp);
    const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, dt, A, b);
    double c0, c1, c2, c3;
    DriftStep1<DRIFT>::taylor(kst, mk, c0, c1, c2, c3);
    CubicMoments cm;
    cm.template eval<true>(c0, c1, c2, c3, Sk);
    const double gb = -Qpinv * cm.Ed;
    double gA = -Qpinv * fma(mk, cm.Ed, cm.Ede);
    Lgm = Qpinv * cm.Ejd;
    LgS = 0.5 * ((Qpinv * cm.Ejde) * iSk);             // adjoint of chol(S_k)
    const double gQc = 0.5 * (Qpinv - iQc);
    Rgm = gb;                                          // b = m_{k+1} - A m_k
    gA -= gb * mk;
    Lgm -= A * gb;
    gA -= 2.0 * (gQc * (A * Sk));                      // Qc = S_{k+1} - A S_k A
    LgS -= A * (gQc * A);
    gC = gA * iSk;                                     // A = C / S_k
    LgS -= gC * A;
    Lgm -= gC * mn;                                    // C = eta_s - m_k m_{k+1}
    Rgm -= gC * mk;
    RgS = gQc;
}

// D = 1, one thread per transition: the same term with every block a scalar.  Per-step invariants are
// hoisted out of the node loop, pi^(-1/2) and sqrt(2) are factored out of the sums, and each of the
// three divisors (chol S_k, chol Q_c, chol Q_p) is inverted once (3 sqrt + 3 div + 1 log per step
// instead of ~20 div/sqrt/log).  Results differ from the generic path in the last bits only.
template <int DRIFT, int H, bool CLOSED = false>
__device__ __forceinline__ void kl_term_scalar(const DriftParams& P, double Qp, double dt, double pi_factor,
                                               double mk, double mn, double edk, double edn, double esk,
                                               KlTermOut<1>& out, bool& ok, bool want_value = true) {
    // expectations_to_ssm_params (ssm_gaussian_transformations.py:93-178)
    const double Sk = edk - mk * mk;
    const double Sn = edn - mn * mn;
    const double Cs = esk - mk * mn;
    if (!(Sk > 0.0)) ok = false;
    if constexpr (CLOSED) {
        if (!want_value) {
            // Gradient only, closed-form moments: every divisor enters as a reciprocal of S_k, Q_c or Q_p (the
            // Cholesky factors only serve the node placement of the quadrature loops and the log-determinants of the
            // value), so three reciprocals replace the three coupled square-root / reciprocal-square-root
            // evaluations (~27 float64 instructions per grid point less; same values to rounding).
            const double iSk = rcp_fast(Sk);
            double Lgm, LgS, Rgm, gQc, gC;
            kl_grad_closed_core<DRIFT>(P, Qp, dt, Cs * iSk, iSk, mk, mn, Sk, Sn, Lgm, LgS, Rgm, gQc, gC, ok);
            out.value = 0.0;
            out.ges(0, 0) = gC;
            out.left.gm(0) = Lgm;
            out.left.gS(0, 0) = LgS;
            out.right.gm(0) = Rgm;
            out.right.gS(0, 0) = gQc;
            return;
        }
    }
    double LS, iLS;
    sqrt_rsqrt_fast(Sk, LS, iLS);
    const double A = (Cs * iLS) * iLS;
    const double b = mn - A * mk;
    const double Qc = Sn - A * (Sk * A);
    if (!(Qc > 0.0) || !(Qp > 0.0)) ok = false;
    double LQ, iLQ, cQp, iQp;
    sqrt_rsqrt_fast(Qc, LQ, iLQ);
    const double Qq = LQ * LQ;
    sqrt_rsqrt_fast(Qp, cQp, iQp);
    const double Qpinv = iQp * iQp;
    // C = -(logdet Qq - logdet Qp) - 1 + Qp^-1 Qq   (sde_utils.py:317-330); chol(Qq) = LQ exactly
    // (skipped when only the gradient is wanted: the fused site update takes the KL value from the refreshed
    // posterior later, classic_elbo, and never reads this one)
    const double Cterm = want_value ? -2.0 * log(LQ * iQp) - 1.0 + Qpinv * Qq : 0.0;
    const double sL = 1.4142135623730951 * LS;
    // d = x + dt f(x) - (A x + b) and its slope j (DriftStep1);  Q_p^-1 and pi^(-1/2) are factored out of the sums
    const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, dt, A, b);
    double quad = 0.0, gb = 0.0, gmq = 0.0, gA = 0.0, lsb = 0.0;
    if constexpr (CLOSED) {
        // polynomial drift: the Gauss-Hermite sums in closed form (CubicMoments)
        double c0, c1, c2, c3;
        DriftStep1<DRIFT>::taylor(kst, mk, c0, c1, c2, c3);
        CubicMoments cm;
        cm.template eval<true>(c0, c1, c2, c3, Sk);
        quad = Qpinv * cm.Ed2;
        gb = -Qpinv * cm.Ed;
        gmq = Qpinv * cm.Ejd;
        gA = -Qpinv * fma(mk, cm.Ed, cm.Ede);
        lsb = (Qpinv * cm.Ejde) * iLS;
    } else {
#pragma unroll
    for (int p = 0; p < H; ++p) {
        const double xi = gh_x<H>(p), w = gh_w<H>(p);
        const double X = fma(sL, xi, mk);
        double d, j;
        DriftStep1<DRIFT>::eval(P, kst, X, d, j);
        const double wd = w * d;
        quad = fma(wd, d, quad);
        const double Xb = j * wd;
        gb -= wd;
        gmq += Xb;
        gA = fma(-wd, X, gA);
        lsb = fma(Xb, xi, lsb);
    }
    const double sc = pi_factor * Qpinv;
    quad *= sc;
    gb *= sc;
    gmq *= sc;
    gA *= sc;
    lsb *= 1.4142135623730951 * sc;
    }
    out.value = 0.5 * (quad + Cterm);
    // reverse through the pointwise algebra
    double Lgm = gmq;
    double LgS = 0.5 * lsb * iLS;                 // adjoint of chol(S_k)
    const double gQc = 0.5 * (Qpinv - iLQ * iLQ);
    double Rgm = gb;                              // b = m_{k+1} - A m_k
    gA -= gb * mk;
    Lgm -= A * gb;
    const double RgS = gQc;                       // Qc = S_{k+1} - A S_k A
    gA -= 2.0 * (gQc * (A * Sk));
    LgS -= A * (gQc * A);
    const double gC = (gA * iLS) * iLS;           // A = C / S_k
    LgS -= gC * A;
    Lgm -= gC * mn;                               // C = eta_s - m_k m_{k+1}
    Rgm -= gC * mk;
    out.ges(0, 0) = gC;
    out.left.gm(0) = Lgm;
    out.left.gS(0, 0) = LgS;
    out.right.gm(0) = Rgm;
    out.right.gS(0, 0) = RgS;
}

// D = 3 quadrature over the tensor-product nodes as nested loops: the lanes of the warp stride over the 400
// (i0, i1) pairs, the inner loop runs over the last coordinate with a warp-uniform table index.  chol(S) is lower
// triangular, so only X_2 changes in the inner loop; everything that depends on (X_0, X_1) alone -- the partial
// residual c = X - (A X + b), the j = 0, 1 columns of the accumulators -- is hoisted per pair.  Same sums as the
// flat loop of kl_term (other association of the partial sums), about half the float64 operations per node.
template <int DRIFT, int H, bool GRAD>
__device__ __forceinline__ void kl_quad_nested3(const DriftParams& P, const Mat<3>& LS, const Vec<3>& m,
                                                const Mat<3>& A, const Vec<3>& b, const Mat<3>& Qs, double dt,
                                                int lane, double& quad, Vec<3>& gb, Vec<3>& gmq, Mat<3>& gA,
                                                Mat<3>& LSb) {
    constexpr double R2 = 1.4142135623730951;
    quad = 0.0;
    gb = vec_zero<3>(); gmq = vec_zero<3>(); gA = mat_zero<3>(); LSb = mat_zero<3>();
    const double s22 = R2 * LS(2, 2);
    Mat<3> IA;   // I - A
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) IA(i, j) = ((i == j) ? 1.0 : 0.0) - A(i, j);
    for (int o = lane; o < H * H; o += 32) {
        const int i0 = o / H, i1 = o - i0 * H;
        const double x0 = gh_x_div<H>(i0), x1 = gh_x_div<H>(i1);
        const double w01 = gh_w_div<H>(i0) * gh_w_div<H>(i1);
        const double X0 = R2 * (LS(0, 0) * x0) + m(0);
        const double X1 = R2 * (LS(1, 0) * x0 + LS(1, 1) * x1) + m(1);
        const double X2base = R2 * (LS(2, 0) * x0 + LS(2, 1) * x1) + m(2);
        // c_i = (X_i [i < 2]) - (A_i0 X0 + A_i1 X1 + b_i)
        const double c0 = X0 - (A(0, 0) * X0 + A(0, 1) * X1 + b(0));
        const double c1 = X1 - (A(1, 0) * X0 + A(1, 1) * X1 + b(1));
        const double c2 = -(A(2, 0) * X0 + A(2, 1) * X1 + b(2));
        double q_o = 0.0, rb0 = 0.0, rb1 = 0.0, rb2 = 0.0, xb0 = 0.0, xb1 = 0.0, xb2 = 0.0;
        double ra0 = 0.0, ra1 = 0.0, ra2 = 0.0, xs2 = 0.0;
#pragma unroll 4
        for (int i2 = 0; i2 < H; ++i2) {
            const double x2 = gh_x<H>(i2), w2 = gh_w<H>(i2);
            Vec<3> X, f;
            Mat<3> J;
            X(0) = X0; X(1) = X1; X(2) = fma(s22, x2, X2base);
            Drift<DRIFT, 3>::eval(P, X, f, J);
            const double d0 = fma(dt, f(0), fma(-A(0, 2), X(2), c0));
            const double d1 = fma(dt, f(1), fma(-A(1, 2), X(2), c1));
            const double d2 = fma(dt, f(2), fma(IA(2, 2), X(2), c2));
            const double u0 = w2 * (Qs(0, 0) * d0 + Qs(0, 1) * d1 + Qs(0, 2) * d2);
            const double u1 = w2 * (Qs(1, 0) * d0 + Qs(1, 1) * d1 + Qs(1, 2) * d2);
            const double u2 = w2 * (Qs(2, 0) * d0 + Qs(2, 1) * d1 + Qs(2, 2) * d2);
            q_o += d0 * u0 + d1 * u1 + d2 * u2;
            if (GRAD) {
                // Xb = ((I + dt J) - A)^T u
                const double v0 = fma(dt, J(0, 0) * u0 + J(1, 0) * u1 + J(2, 0) * u2,
                                      IA(0, 0) * u0 + IA(1, 0) * u1 + IA(2, 0) * u2);
                const double v1 = fma(dt, J(0, 1) * u0 + J(1, 1) * u1 + J(2, 1) * u2,
                                      IA(0, 1) * u0 + IA(1, 1) * u1 + IA(2, 1) * u2);
                const double v2 = fma(dt, J(0, 2) * u0 + J(1, 2) * u1 + J(2, 2) * u2,
                                      IA(0, 2) * u0 + IA(1, 2) * u1 + IA(2, 2) * u2);
                rb0 += u0; rb1 += u1; rb2 += u2;
                xb0 += v0; xb1 += v1; xb2 += v2;
                ra0 = fma(u0, X(2), ra0); ra1 = fma(u1, X(2), ra1); ra2 = fma(u2, X(2), ra2);
                xs2 = fma(v2, x2, xs2);
            }
        }
        quad = fma(w01, q_o, quad);
        if (GRAD) {
            rb0 *= w01; rb1 *= w01; rb2 *= w01; xb0 *= w01; xb1 *= w01; xb2 *= w01;
            gb(0) -= rb0; gb(1) -= rb1; gb(2) -= rb2;
            gmq(0) += xb0; gmq(1) += xb1; gmq(2) += xb2;
            gA(0, 0) -= rb0 * X0; gA(0, 1) -= rb0 * X1; gA(0, 2) -= w01 * ra0;
            gA(1, 0) -= rb1 * X0; gA(1, 1) -= rb1 * X1; gA(1, 2) -= w01 * ra1;
            gA(2, 0) -= rb2 * X0; gA(2, 1) -= rb2 * X1; gA(2, 2) -= w01 * ra2;
            LSb(0, 0) += xb0 * x0;
            LSb(1, 0) += xb1 * x0; LSb(1, 1) += xb1 * x1;
            LSb(2, 0) += xb2 * x0; LSb(2, 1) += xb2 * x1; LSb(2, 2) += w01 * xs2;
        }
    }
}

// One transition term.  Inputs are the expectation parameters rebuilt from (A_q, m, S) exactly as
// ssm_to_expectations does (ssm_gaussian_transformations.py:32-89).
template <int DRIFT, int D, int H, int G, bool CLOSED = false>
__device__ __forceinline__ void kl_term(const DriftParams& P, const Mat<D>& Qp, double dt, double pi_factor,
                                        const Vec<D>& e1k, const Vec<D>& e1n, const Mat<D>& edk, const Mat<D>& edn,
                                        const Mat<D>& esk, int lane, KlTermOut<D>& out, bool& ok,
                                        bool want_value = true) {
    if constexpr (D == 1 && G == 1) {
        kl_term_scalar<DRIFT, H, CLOSED>(P, Qp(0, 0), dt, pi_factor, e1k(0), e1n(0), edk(0, 0), edn(0, 0), esk(0, 0), out, ok,
                                 want_value);
        return;
    }
    ExpToSsm<D> r;
    exp_to_ssm_step(e1k, e1n, edk, edn, esk, r, ok);
    const Mat<D> Qq = mmT(r.LQ, r.LQ);
    const Mat<D> cQp = chol(Qp, ok);
    const Mat<D> Qpinv = chol_inverse(cQp);
    const Mat<D> Qpinv_s = sym(Qpinv);
    // constants  C = -(logdet Qq - logdet Qp) - D + sum_i (Qp^-1)_ii (Qq)_ii   (sde_utils.py:317-330)
    const Mat<D> cQq = chol(Qq, ok);
    double tr = Qpinv(0, 0) * Qq(0, 0);
#pragma unroll
    for (int i = 1; i < D; ++i) tr += Qpinv(i, i) * Qq(i, i);
    const double Cterm = -(logdet_from_chol(cQq) - logdet_from_chol(cQp)) - (double)D + tr;

    // quadrature and its adjoint accumulators
    double quad = 0.0;
    Vec<D> gb = vec_zero<D>(), gmq = vec_zero<D>();
    Mat<D> gA = mat_zero<D>(), LSb = mat_zero<D>();
    constexpr int NP = NodeCount<H, D>::value;
    constexpr bool kNested = (D == 3 && G == 32);
    if constexpr (kNested) {
        kl_quad_nested3<DRIFT, H, true>(P, r.LS, r.mk, r.A, r.b, Qpinv_s, dt, lane, quad, gb, gmq, gA, LSb);
        quad *= pi_factor;
#pragma unroll
        for (int i = 0; i < D; ++i) { gb(i) *= pi_factor; gmq(i) *= pi_factor; }
#pragma unroll
        for (int e = 0; e < D * D; ++e) { gA.a[e] *= pi_factor; LSb.a[e] *= 1.4142135623730951 * pi_factor; }
    }
    for (int p = lane; p < (kNested ? 0 : NP); p += G) {
        Vec<D> xi, f;
        Mat<D> J;
        double w;
        gh_node<H, D>(p, xi, w);
        w *= pi_factor;
        const Vec<D> X = gh_point(r.LS, r.mk, xi);
        Drift<DRIFT, D>::eval(P, X, f, J);
        Vec<D> d;
        const Vec<D> AX = mv(r.A, X);
#pragma unroll
        for (int i = 0; i < D; ++i) d(i) = (X(i) + dt * f(i)) - (AX(i) + r.b(i));
        quad += w * dot(d, mv(Qpinv, d));
        const Vec<D> rb = w * mv(Qpinv_s, d);
        // Jd = (I + dt J) - A ;  Xb = Jd^T rb
        Vec<D> Xb;
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double s = 0.0;
#pragma unroll
            for (int j = 0; j < D; ++j) s += ((((i == j) ? 1.0 : 0.0) + dt * J(j, i)) - r.A(j, i)) * rb(j);
            Xb(i) = s;
        }
#pragma unroll
        for (int i = 0; i < D; ++i) {
            gb(i) -= rb(i);
            gmq(i) += Xb(i);
#pragma unroll
            for (int j = 0; j < D; ++j) {
                gA(i, j) -= rb(i) * X(j);
                if (j <= i) LSb(i, j) += 1.4142135623730951 * (Xb(i) * xi(j));
            }
        }
    }
    quad = group_sum<G>(quad);
    group_sum<G>(gb);
    group_sum<G>(gmq);
    group_sum<G>(gA);
    group_sum<G>(LSb);

    out.value = 0.5 * (quad + Cterm);
    // ---- reverse through the pointwise algebra (same statement order as the oracle)
    KlNodePart<D>& Lf = out.left;
    KlNodePart<D>& Rt = out.right;
    Lf.gm = gmq;
    Lf.gS = chol_adjoint(r.LS, LSb);
    Mat<D> diagQp = mat_zero<D>();
#pragma unroll
    for (int i = 0; i < D; ++i) diagQp(i, i) = Qpinv(i, i);
    const Mat<D> gQc = sym(0.5 * (diagQp - chol_inverse(r.LQ)));
    // b = m_{k+1} - A m_k
    Rt.gm = gb;
    gA = gA - outer(gb, r.mk);
    Lf.gm = Lf.gm - Tmv(r.A, gb);
    // Qc = S_{k+1} - A S_k A^T
    Rt.gS = gQc;
    gA = gA - 2.0 * mm(gQc, mm(r.A, r.Sk));
    Lf.gS = Lf.gS - Tmm(r.A, mm(gQc, r.A));
    // A = (S_k^-1 C)^T
    const Mat<D> gC = chol_solve(r.LS, transpose(gA));
    Lf.gS = Lf.gS + sym(-1.0 * mm(gC, r.A));
    // C = eta_s^T - m_k m_{k+1}^T
    out.ges = transpose(gC);
    Lf.gm = Lf.gm - mv(gC, r.mn);
    Rt.gm = Rt.gm - Tmv(gC, r.mk);
}

// Optional epilogue of the girsanov site update: the natural parameters of the posterior that follows
// (prior naturals + updated sites + scattered data sites), so that the rebuild after update_girsanov_sites
// (variational_cvi_sde.py:299) starts from them without another pass over the sites.
struct KlNaturalsOut {
    const double *pA = nullptr, *pb = nullptr, *pcholQ = nullptr, *pmu0 = nullptr, *pcholP0 = nullptr;
    double *th1 = nullptr, *thd = nullptr, *ths = nullptr;
    // optional per-sequence step sizes held on the device (vdp_cvi_problem.lr_girsanov_dev / active_dev)
    const double* lr_dev = nullptr;
    const int32_t* active = nullptr;
};

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// shared-memory record handed from the group of node i to the group of node i+1
template <int D>
struct KlHandoff { double gm[D]; double gS[D * D]; };

// Outputs per node i = (b,k): g1, gd (and gs for k < N) = d KL / d eta, plus the per-transition value
// terms[i] (0 for k = N; KL_0 added at k = 0) for the deterministic reduction.
// If `sites` update pointers are given (upd != 0) the girsanov site update of
// variational_cvi_sde.py:286-296 is applied in place instead of writing the gradients:
//   g1 += lr (scatter(d1) - dKL/deta1),  gd += lr (scatter(d2) - dKL/deta_d),  gs -= lr dKL/deta_s
// where scatter(d.) has been pre-scattered into sc1/scd by the caller.
// register budget: the scalar kernel (one thread per grid point, blocks of 128) is bound by the latency of dependent
// float64 instructions, so resident warps count for more than a few spilled words (measured on B200, C4 shape:
// 56 registers 0.79 ms, 48: 0.76, 40 + 48 bytes of stack: 0.73)
#ifndef VDP_KL_MINBLOCKS
#define VDP_KL_MINBLOCKS 12
#endif
// warp-per-grid-point kernels (D >= 2): 2 blocks of 256 threads per SM = 128 registers (measured on B200: 255
// registers / 1 block 14.1 ms (D = 2, 4096 x 500) and 9.44 ms (D = 3, 2 x 10^5); 128 / 2 blocks 10.6 and 8.86 ms; 80 / 3
// blocks spills: 11.9 and 12.5 ms)
#ifndef VDP_KLG_MINBLOCKS
#define VDP_KLG_MINBLOCKS 2
#endif
template <int G> struct KlBounds {
    static constexpr int threads = (G == 1) ? 128 : 256, blocks = (G == 1) ? VDP_KL_MINBLOCKS : VDP_KLG_MINBLOCKS;
};
// D >= 2 with the exact low-order rule of a polynomial drift (ExactOrder: 4 .. 64 nodes per grid point): every lane of
// a group repeats the pointwise algebra around the quadrature, which now outweighs the node loop, so one thread per
// grid point does best (measured on B200, kl_grad phase: D = 2 VdP 4096 x 500: 1 lane 0.26 ms, 2: 0.45, 4: 0.79;
// D = 3 OU 2 x 10^5: 1 lane 0.088 ms, 2: 0.118, 4: 0.194, 8: 0.365); the register budget is for the D x D blocks
#ifndef VDP_KLX_G2
#define VDP_KLX_G2 1
#endif
#ifndef VDP_KLX_G3
#define VDP_KLX_G3 1
#endif
#ifndef VDP_KLX_MINBLOCKS2
#define VDP_KLX_MINBLOCKS2 4
#endif
#ifndef VDP_KLX_MINBLOCKS3
#define VDP_KLX_MINBLOCKS3 2
#endif
template <int D> struct KlExactGroup { static constexpr int value = (D == 2) ? VDP_KLX_G2 : VDP_KLX_G3; };
// launch shape of k_kl_grad by (D, H, G): threads per block and resident blocks the register budget is cut for
template <int D, int H, int G> struct KlShape {
    static constexpr bool exact = (D >= 2 && H <= 4);
    static constexpr int threads = exact ? 128 : KlBounds<G>::threads;
    static constexpr int blocks = exact ? (D == 2 ? VDP_KLX_MINBLOCKS2 : VDP_KLX_MINBLOCKS3) : KlBounds<G>::blocks;
};
// UPDT >= 0: the update mode is a compile-time constant (prunes the other modes' code from the scalar kernel)
// NOVAL: the caller passes terms == nullptr (gradient only), known at compile time
template <int DRIFT, int D, int H, int G, bool CLOSED = false, int UPDT = -1, bool NOVAL = false>
__global__ void __launch_bounds__(KlShape<D, H, G>::threads, KlShape<D, H, G>::blocks) k_kl_grad(DriftParams P, Mat<D> q, double dt, double pi_factor,
                          const double* __restrict__ time_grid, Vec<D> x0_mu, Mat<D> x0_cov,
                          const double* __restrict__ Aq, const double* __restrict__ m,
                          const double* __restrict__ Sd, double* __restrict__ terms, double* __restrict__ g1,
                          double* __restrict__ gd, double* __restrict__ gs, int upd, double lr,
                          const double* __restrict__ sc1, const double* __restrict__ scd, int64_t B, int64_t N,
                          int64_t* status, int64_t node0 = 0, int64_t node1 = -1, KlNaturalsOut nat = KlNaturalsOut()) {
    if (UPDT >= 0) upd = UPDT;
    if (NOVAL) terms = nullptr;
    // nodes [node0, node1) of the flat (b, k) index are processed (default: all); a rank of the sharded
    // quadrature passes its own range, the halo group recomputes the term that straddles the range start
    extern __shared__ unsigned char smem_raw[];
    KlHandoff<D>* hand = reinterpret_cast<KlHandoff<D>*>(smem_raw);
    constexpr int DD = D * D;
    const int groups = blockDim.x / G;          // groups per block, first one is the halo
    const int gidx = threadIdx.x / G;
    const int lane = threadIdx.x % G;
    const int64_t n = N + 1;
    const int64_t total = (node1 < 0) ? B * n : node1;
    const int64_t i = node0 + (int64_t)blockIdx.x * (groups - 1) + gidx - 1;   // flat node index of this group
    const bool in_range = (i >= 0) && (i < total);
    int64_t b = 0, k = 0;
    if (in_range) split_index(i, n, b, k);
    const bool has_term = in_range && (k < N);
    bool ok = true;
    KlTermOut<D> t;
    Vec<D> mk = vec_zero<D>();
    Mat<D> Sk = mat_eye<D>();
    if (in_range) { mk = vec_load<D>(m + i * D); Sk = mat_load<D>(Sd + i * DD); }
    if (upd && in_range && lane == 0 && gidx != 0) {
        // the site update at the end of the kernel reads the sites, the scattered data sites and (for the fused
        // naturals) the prior: start those lines on their way now, the quadrature hides the latency
        const int64_t tt = b * N + k;
        if (upd == 2) {   // sites held in theta
            prefetch_l2(nat.th1 + i * D);
            prefetch_l2(nat.thd + i * DD);
            if (k < N) prefetch_l2(nat.ths + tt * DD);
        } else {
            prefetch_l2(g1 + i * D);
            prefetch_l2(gd + i * DD);
            if (k < N) prefetch_l2(gs + tt * DD);
        }
        if (sc1) { prefetch_l2(sc1 + i * D); prefetch_l2(scd + i * DD); }
        if (nat.pA && k < N) {
            prefetch_l2(nat.pA + tt * DD);
            prefetch_l2(nat.pb + tt * D);
            prefetch_l2(nat.pcholQ + tt * DD);
        }
    }
    if (has_term) {
        const Vec<D> mn = vec_load<D>(m + (i + 1) * D);
        const Mat<D> Sn = mat_load<D>(Sd + (i + 1) * DD);
        const Mat<D> Ak = mat_load<D>(Aq + (b * N + k) * DD);
        const Mat<D> edk = Sk + outer(mk, mk);
        const Mat<D> edn = Sn + outer(mn, mn);
        const Mat<D> esk = mm(Ak, Sk) + outer(mn, mk);
        const double delta = time_grid[k + 1] - time_grid[k];
        kl_term<DRIFT, D, H, G, CLOSED>(P, delta * q, dt, pi_factor, mk, mn, edk, edn, esk, lane, t, ok, terms != nullptr);
        if (!ok && lane == 0) report_not_pd(status, i);
        if (lane == 0) {
#pragma unroll
            for (int e = 0; e < D; ++e) hand[gidx].gm[e] = t.right.gm(e);
#pragma unroll
            for (int e = 0; e < DD; ++e) hand[gidx].gS[e] = t.right.gS.a[e];
        }
    }
    __syncthreads();
    if (!in_range || gidx == 0 || lane != 0) return;
    // ---- node i: left part of term k + right part of term k-1 (+ KL_0 at k = 0)
    Vec<D> gm = vec_zero<D>();
    Mat<D> gS = mat_zero<D>();
    double value = 0.0;
    if (has_term) { gm = t.left.gm; gS = t.left.gS; value = t.value; }
    if (k > 0) {
#pragma unroll
        for (int e = 0; e < D; ++e) gm(e) += hand[gidx - 1].gm[e];
#pragma unroll
        for (int e = 0; e < DD; ++e) gS.a[e] += hand[gidx - 1].gS[e];
    } else {
        // KL[q(x0) || p(x0)]  (sde_utils.py:536-541), q(x0) = N(m_0, S_0) through its Cholesky factor
        bool ok0 = true;
        const Mat<D> Lp = chol(x0_cov, ok0);
        const Mat<D> Lq = chol(Sk, ok0);
        const Mat<D> Ppinv = chol_inverse(Lp);
        const Mat<D> Mq = trsm(Lp, Lq);
        const Vec<D> diff = trsv(Lp, x0_mu - mk);
        double fro = 0.0;
#pragma unroll
        for (int e = 0; e < DD; ++e) fro += Mq.a[e] * Mq.a[e];
        value += 0.5 * (fro + dot(diff, diff) - (double)D + logdet_from_chol(Lp) - logdet_from_chol(Lq));
        gm = gm + mv(Ppinv, mk - x0_mu);
        gS = gS + 0.5 * (Ppinv - chol_inverse(Lq));
    }
    // S = eta_d - m m^T
    gm = gm - 2.0 * mv(gS, mk);
    if (nat.lr_dev) lr = (!nat.active || nat.active[b]) ? nat.lr_dev[b] : 0.0;
    if (terms) terms[i] = value;
    if (!upd) {
        vec_store<D>(g1 + i * D, gm);
        mat_store<D>(gd + i * DD, gS);
        if (has_term) mat_store<D>(gs + (b * N + k) * DD, t.ges);
    } else if (upd == 2) {
        // VDP_FLAG_SITES_IN_NATURALS: the sites live in theta = nat(p) + g + scatter(data sites); the update
        // g += lr (scatter - dKL/deta) (variational_cvi_sde.py:290-292) moves theta by the same amount.  No read of
        // the prior or of g, no write of g: 88 instead of 136 bytes per grid step (D = 1).
        Vec<D> s1 = vec_zero<D>();
        Mat<D> s2 = mat_zero<D>();
        if (sc1) { s1 = vec_load<D>(sc1 + i * D); s2 = mat_load<D>(scd + i * DD); }
        vec_store<D>(nat.th1 + i * D, vec_load<D>(nat.th1 + i * D) + lr * (s1 - gm));
        mat_store<D>(nat.thd + i * DD, mat_load<D>(nat.thd + i * DD) + lr * (s2 - gS));
        if (has_term) {
            const int64_t tt = b * N + k;
            mat_store<D>(nat.ths + tt * DD, mat_load<D>(nat.ths + tt * DD) - lr * t.ges);
        }
    } else {
        Vec<D> s1 = vec_zero<D>();
        Mat<D> s2 = mat_zero<D>();
        if (sc1) { s1 = vec_load<D>(sc1 + i * D); s2 = mat_load<D>(scd + i * DD); }
        const Vec<D> n1 = vec_load<D>(g1 + i * D) + lr * (s1 - gm);
        const Mat<D> nd = mat_load<D>(gd + i * DD) + lr * (s2 - gS);
        vec_store<D>(g1 + i * D, n1);
        mat_store<D>(gd + i * DD, nd);
        Mat<D> ns = mat_zero<D>();
        const int64_t tt = b * N + k;
        if (has_term) {
            ns = mat_load<D>(gs + tt * DD) - lr * t.ges;
            mat_store<D>(gs + tt * DD, ns);
        }
        if (nat.th1) {
            // full_sites of the posterior that follows (variational_cvi_sde.py:161-174): naturals of the linearised
            // prior (same statements as k_ssm_to_naturals) + the updated girsanov sites + the scattered data sites
            const Mat<D> ck = (k == 0) ? mat_load<D>(nat.pcholP0 + b * DD) : mat_load<D>(nat.pcholQ + (tt - 1) * DD);
            const Vec<D> ok0 = (k == 0) ? vec_load<D>(nat.pmu0 + b * D) : vec_load<D>(nat.pb + (tt - 1) * D);
            Mat<D> Ak = mat_zero<D>(), cn = mat_eye<D>(), td, ts;
            Vec<D> bn = vec_zero<D>(), t1;
            if (has_term) {
                Ak = mat_load<D>(nat.pA + tt * DD);
                cn = mat_load<D>(nat.pcholQ + tt * DD);
                bn = vec_load<D>(nat.pb + tt * D);
            }
            ssm_node_naturals<D>(ck, ok0, has_term, Ak, cn, bn, t1, td, ts);
            if (has_term) mat_store<D>(nat.ths + tt * DD, ts + ns);
            t1 = t1 + n1;
            td = td + nd;
            vec_store<D>(nat.th1 + i * D, t1 + s1);
            mat_store<D>(nat.thd + i * DD, td + s2);
        }
    }
}

// D = 1, polynomial drift: the classic_elbo KL term of one transition from scalars (the statements of k_kl_value's
// scalar path with the closed-form moments) and the KL of the initial state (variational_cvi_sde.py:478-484), for the
// mean sweep that adds the terms up on its way (k_mean_tiled, tile_sweeps.cuh).
// The log-determinant part, -2 log(|chol Q_q| / chol Q_p), is left to the caller as the factor `ratio` of a running
// product: one log per chunk of transitions instead of one per transition.
template <int DRIFT>
__device__ __forceinline__ double kl_value_scalar_closed(const DriftParams& P, double Qp, double dt, double Ak,
                                                         double bk, double cq, double mk, double Sk, bool& ok,
                                                         double& ratio) {
    const double Qq = cq * cq;
    if (!(Qp > 0.0) || !(Sk > 0.0) || !(Qq > 0.0)) ok = false;
    double cQp, iQp;
    sqrt_rsqrt_fast(Qp, cQp, iQp);
    const double Qpinv = iQp * iQp;
    ratio = fabs(cq) * iQp;
    const double Cterm = -1.0 + Qpinv * Qq;
    const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, dt, Ak, bk);
    double c0, c1, c2, c3;
    DriftStep1<DRIFT>::taylor(kst, mk, c0, c1, c2, c3);
    CubicMoments cm;
    cm.template eval<false>(c0, c1, c2, c3, Sk);
    return 0.5 * (Qpinv * cm.Ed2 + Cterm);
}

template <int D>
__device__ __forceinline__ double kl_initial_state(const double* __restrict__ qmu0, const double* __restrict__ qcholP0,
                                                   const double* __restrict__ pmu0,
                                                   const double* __restrict__ pcholP0, int64_t b, bool& ok) {
    constexpr int DD = D * D;
    const Mat<D> cq0 = mat_load<D>(qcholP0 + b * DD), cp0 = mat_load<D>(pcholP0 + b * DD);
    const Mat<D> Lq = chol(mmT(cq0, cq0), ok), Lp = chol(mmT(cp0, cp0), ok);
    const Mat<D> Mq = trsm(Lp, Lq);
    const Vec<D> diff = trsv(Lp, vec_load<D>(pmu0 + b * D) - vec_load<D>(qmu0 + b * D));
    double fro = 0.0;
#pragma unroll
    for (int e = 0; e < DD; ++e) fro += Mq.a[e] * Mq.a[e];
    return 0.5 * (fro + dot(diff, diff) - (double)D + logdet_from_chol(Lp) - logdet_from_chol(Lq));
}

// KL_q_p of classic_elbo (variational_cvi_sde.py:446-486): value only, with the reference's
// element-wise square of the Cholesky factor (:452) standing in for Q_q.
template <int DRIFT, int D, int H, int G, bool CLOSED = false>
__global__ void k_kl_value(DriftParams P, Mat<D> q, double dt, double pi_factor,
                           const double* __restrict__ Aq, const double* __restrict__ bq,
                           const double* __restrict__ cholQq, const double* __restrict__ m,
                           const double* __restrict__ Sd, const double* __restrict__ qmu0,
                           const double* __restrict__ qcholP0, const double* __restrict__ pmu0,
                           const double* __restrict__ pcholP0, double* __restrict__ terms, int64_t B, int64_t N,
                           int64_t* status, int64_t node0 = 0, int64_t node1 = -1) {
    const int64_t i = node0 + (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / G;
    const int lane = threadIdx.x % G;
    const int64_t n = N + 1;
    if (i >= ((node1 < 0) ? B * n : node1)) return;
    constexpr int DD = D * D;
    int64_t b, k;
    split_index(i, n, b, k);
    if (k == N) { if (lane == 0) terms[i] = 0.0; return; }
    const int64_t tt = b * N + k;
    bool ok = true;
    const Mat<D> Ak = mat_load<D>(Aq + tt * DD);
    const Vec<D> bk = vec_load<D>(bq + tt * D);
    const Mat<D> cq = mat_load<D>(cholQq + tt * DD);
    Mat<D> Qq;
#pragma unroll
    for (int e = 0; e < DD; ++e) Qq.a[e] = cq.a[e] * cq.a[e];
    const Mat<D> Qp = dt * q;
    const Vec<D> mk = vec_load<D>(m + i * D);
    Mat<D> cQp, Qpinv, LS;
    double Cterm;
    if constexpr (D == 1 && G == 1) {
        // scalar: one reciprocal, one log (chol(cq^2) = |cq| exactly)
        const double Sk = Sd[i];
        if (!(Qp(0, 0) > 0.0) || !(Sk > 0.0) || !(Qq(0, 0) > 0.0)) ok = false;
        double iQp;
        sqrt_rsqrt_fast(Qp(0, 0), cQp(0, 0), iQp);
        Qpinv(0, 0) = iQp * iQp;
        Cterm = -2.0 * log(fabs(cq(0, 0)) * iQp) - 1.0 + Qpinv(0, 0) * Qq(0, 0);
        LS(0, 0) = sqrt_fast(Sk);
    } else {
        cQp = chol(Qp, ok);
        Qpinv = chol_inverse(cQp);
        const Mat<D> cQq = chol(Qq, ok);   // tf.linalg.logdet reads the lower triangle
        double tr = Qpinv(0, 0) * Qq(0, 0);
#pragma unroll
        for (int e = 1; e < D; ++e) tr += Qpinv(e, e) * Qq(e, e);
        Cterm = -(logdet_from_chol(cQq) - logdet_from_chol(cQp)) - (double)D + tr;
        LS = chol(mat_load<D>(Sd + i * DD), ok);
    }
    double quad = 0.0;
    if constexpr (D == 1 && G == 1) {
        const double sL = 1.4142135623730951 * LS(0, 0), m1 = mk(0);
        const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, dt, Ak(0, 0), bk(0));
        if constexpr (CLOSED) {
            double c0, c1, c2, c3;
            DriftStep1<DRIFT>::taylor(kst, m1, c0, c1, c2, c3);
            CubicMoments cm;
            cm.template eval<false>(c0, c1, c2, c3, Sd[i]);
            quad = Qpinv(0, 0) * cm.Ed2;
        } else {
        double quad_ = 0.0;
#pragma unroll
        for (int p = 0; p < H; ++p) {
            double d, j;
            DriftStep1<DRIFT>::eval(P, kst, fma(sL, gh_x<H>(p), m1), d, j);   // x + dt f(x) - (A x + b)
            quad_ = fma(gh_w<H>(p) * d, d, quad_);
        }
        quad = (pi_factor * Qpinv(0, 0)) * quad_;
        }
    } else {
    constexpr int NP = NodeCount<H, D>::value;
    constexpr bool kNested = (D == 3 && G == 32);
    if constexpr (kNested) {
        Vec<3> t1, t2;
        Mat<3> t3, t4;
        kl_quad_nested3<DRIFT, H, false>(P, LS, mk, Ak, bk, sym(Qpinv), dt, lane, quad, t1, t2, t3, t4);
        quad *= pi_factor;
    }
    for (int p = lane; p < (kNested ? 0 : NP); p += G) {
        Vec<D> xi, f, d;
        Mat<D> J;
        double w;
        gh_node<H, D>(p, xi, w);
        w *= pi_factor;
        const Vec<D> X = gh_point(LS, mk, xi);
        Drift<DRIFT, D>::eval(P, X, f, J);
        const Vec<D> AX = mv(Ak, X);
#pragma unroll
        for (int e = 0; e < D; ++e) d(e) = (X(e) + dt * f(e)) - (AX(e) + bk(e));
        quad += w * dot(d, mv(Qpinv, d));
    }
    }
    quad = group_sum<G>(quad);
    if (lane == 0) {
        double value = 0.5 * (quad + Cterm);
        if (k == 0) {
            // kl_0 (variational_cvi_sde.py:478-484): both covariances are rebuilt as chol chol^T
            const Mat<D> cq0 = mat_load<D>(qcholP0 + b * DD), cp0 = mat_load<D>(pcholP0 + b * DD);
            const Mat<D> Lq = chol(mmT(cq0, cq0), ok), Lp = chol(mmT(cp0, cp0), ok);
            const Mat<D> Mq = trsm(Lp, Lq);
            const Vec<D> diff = trsv(Lp, vec_load<D>(pmu0 + b * D) - vec_load<D>(qmu0 + b * D));
            double fro = 0.0;
#pragma unroll
            for (int e = 0; e < DD; ++e) fro += Mq.a[e] * Mq.a[e];
            value += 0.5 * (fro + dot(diff, diff) - (double)D + logdet_from_chol(Lp) - logdet_from_chol(Lq));
        }
        if (!ok) report_not_pd(status, i);
        terms[i] = value;
    }
}

// deterministic per-sequence sum of terms[b, 0..n) (+ optional extra[b]) : one block per sequence
static __global__ void k_reduce_terms(const double* __restrict__ terms, int64_t n, double* __restrict__ out) {
    __shared__ double sh[256];
    const int64_t b = blockIdx.x;
    double acc = 0.0;
    for (int64_t k = threadIdx.x; k < n; k += blockDim.x) acc += terms[b * n + k];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[b] = sh[0];
}

// Long sequences (one chain of 10^7 grid points is ONE block above): first the sum of every chunk of kReduceChunk terms
// into the chunk's first slot -- a block reads and writes its own chunk only -- then the strided sum of those.  The
// order of the additions is fixed by (n, chunk), so the result is reproducible.  `terms` is scratch and is overwritten.
constexpr int64_t kReduceChunk = 8192;
static __global__ void k_reduce_chunks(double* __restrict__ terms, int64_t n, int64_t nchunks) {
    __shared__ double sh[256];
    const int64_t b = blockIdx.x / nchunks, c = blockIdx.x % nchunks;
    double* t = terms + b * n + c * kReduceChunk;
    const int64_t cnt = min(kReduceChunk, n - c * kReduceChunk);
    double acc = 0.0;
    for (int64_t k = threadIdx.x; k < cnt; k += blockDim.x) acc += t[k];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) t[0] = sh[0];
}
static __global__ void k_reduce_strided(const double* __restrict__ terms, int64_t n, int64_t nchunks,
                                        double* __restrict__ out) {
    __shared__ double sh[256];
    const int64_t b = blockIdx.x;
    double acc = 0.0;
    for (int64_t c = threadIdx.x; c < nchunks; c += blockDim.x) acc += terms[b * n + c * kReduceChunk];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[b] = sh[0];
}
// out[b] = sum_k terms[b, k]; terms may be overwritten
static inline void launch_reduce_terms(double* terms, int64_t n, double* out, int64_t B, cudaStream_t st) {
    if (n >= 8 * kReduceChunk && B <= 1024) {
        const int64_t nchunks = (n + kReduceChunk - 1) / kReduceChunk;
        k_reduce_chunks<<<(unsigned)(B * nchunks), 256, 0, st>>>(terms, n, nchunks);
        k_reduce_strided<<<(unsigned)B, 256, 0, st>>>(terms, n, nchunks, out);
    } else {
        k_reduce_terms<<<(unsigned)B, 256, 0, st>>>(terms, n, out);
    }
}

}  // namespace vdp
