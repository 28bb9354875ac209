// Sequential-in-time kernels of the block-tridiagonal layer, one thread per sequence.
// (Round-1 baseline mapping: independent sequences across threads.  The chunked parallel-in-time
// mapping reuses the same step functions from sweeps.cuh.)
#pragma once
#include "common.cuh"
#include "sweeps.cuh"

namespace vdp {

// K_kk = scaleD * Xd[k], K_{k+1,k} = scaleS * Xs[k]; (scaleD, scaleS) = (-2, -1) reads natural
// parameters directly (ssm_gaussian_transformations.py:440), (1, 1) a plain precision.
template <int D>
__global__ void k_chol_fwd(const double* __restrict__ Xd, const double* __restrict__ Xs, double scaleD,
                           double scaleS, double* __restrict__ Ld, double* __restrict__ Ls, int64_t B,
                           int64_t n, int64_t* status) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    const double* xd = Xd + b * n * DD;
    const double* xs = Xs ? Xs + b * (n - 1) * DD : nullptr;
    double* ld = Ld + b * n * DD;
    double* ls = Ls ? Ls + b * (n - 1) * DD : nullptr;
    bool ok = true;
    Mat<D> L = chol_first(scaleD * mat_load<D>(xd), ok);
    if (!ok) { report_not_pd(status, b * n); ok = true; }
    mat_store<D>(ld, L);
    for (int64_t k = 0; k + 1 < n; ++k) {
        const Mat<D> Knext = scaleD * mat_load<D>(xd + (k + 1) * DD);
        if (xs) {
            Mat<D> Lsub, Lnext;
            chol_step(L, scaleS * mat_load<D>(xs + k * DD), Knext, Lsub, Lnext, ok);
            mat_store<D>(ls + k * DD, Lsub);
            L = Lnext;
        } else {
            L = chol(Knext, ok);
        }
        if (!ok) { report_not_pd(status, b * n + k + 1); ok = true; }
        mat_store<D>(ld + (k + 1) * DD, L);
    }
}

// plain inverse band: Sd (always) and Ss (optional)
template <int D>
__global__ void k_invband_bwd(const double* __restrict__ Ld, const double* __restrict__ Ls,
                              double* __restrict__ Sd, double* __restrict__ Ss, int64_t B, int64_t n) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    const double* ld = Ld + b * n * DD;
    const double* ls = Ls ? Ls + b * (n - 1) * DD : nullptr;
    double* sd = Sd + b * n * DD;
    double* ss = Ss ? Ss + b * (n - 1) * DD : nullptr;
    Mat<D> S = invband_last(mat_load<D>(ld + (n - 1) * DD));
    mat_store<D>(sd + (n - 1) * DD, S);
    for (int64_t k = n - 2; k >= 0; --k) {
        const Mat<D> Lkk = mat_load<D>(ld + k * DD);
        if (ls) {
            Mat<D> Ssub, Skk;
            invband_step(Lkk, mat_load<D>(ls + k * DD), S, Ssub, Skk);
            if (ss) mat_store<D>(ss + k * DD, Ssub);
            S = Skk;
        } else {
            S = invband_last(Lkk);
        }
        mat_store<D>(sd + k * DD, S);
    }
}

// Directional derivative of (Cholesky factor ->) inverse band with respect to the factorised matrix: for K = L L^T
// block tridiagonal and a symmetric block-tridiagonal direction dK = (dKd, dKs), the diagonal and sub-diagonal blocks of
// d(K^-1) = -K^-1 dK K^-1.  Forward-mode differentiation of the two recursions the layer runs (cholesky_band,
// block_tri_diag.py:428-440; inverse_from_cholesky_band, :318-337):
//   forward   dS_k = dKd_k - dLs_{k-1} Ls_{k-1}^T - Ls_{k-1} dLs_{k-1}^T,   dL_kk = L_kk Phi(L_kk^-1 dS_k L_kk^-T),
//             dLs_k = (dKs_k - Ls_k dL_kk^T) L_kk^-T            (Phi: lower triangle, diagonal halved)
//   backward  U = L_kk^-1, dU = -U dL_kk U, G = Ls_k U, dG = dLs_k U + Ls_k dU,
//             dSig_{k+1,k} = -dSig_{k+1,k+1} G - Sig_{k+1,k+1} dG,
//             dSig_kk = dU^T U + U^T dU - dSig_{k+1,k}^T G - Sig_{k+1,k}^T dG.
// The output arrays hold dL between the two passes.  One thread per sequence (O(n) blocks each): this serves the
// Fisher-norm term of SSMNaturalGradient(momentum=True) (ssm_natgrad.py:163-190), not a hot path.
template <int D>
__global__ void k_invband_jvp(const double* __restrict__ Ld, const double* __restrict__ Ls,
                              const double* __restrict__ dKd, const double* __restrict__ dKs,
                              double* __restrict__ dSd, double* __restrict__ dSs, int64_t B, int64_t n) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    const double* ld = Ld + b * n * DD;
    const double* ls = Ls + b * (n - 1) * DD;
    const double* kd = dKd + b * n * DD;
    const double* ks = dKs + b * (n - 1) * DD;
    double* od = dSd + b * n * DD;
    double* os = dSs + b * (n - 1) * DD;
    // ---- forward: tangent of the factor
    Mat<D> dLs_prev = mat_zero<D>(), Ls_prev = mat_zero<D>();
    for (int64_t k = 0; k < n; ++k) {
        const Mat<D> Lkk = mat_load<D>(ld + k * DD);
        Mat<D> dS = mirror_lower(mat_load<D>(kd + k * DD));
        if (k > 0) dS = dS - mmT(dLs_prev, Ls_prev) - mmT(Ls_prev, dLs_prev);
        // Phi(L^-1 dS L^-T)
        Mat<D> X = trsm(Lkk, transpose(trsm(Lkk, dS)));      // L^-1 (L^-1 dS)^T = L^-1 dS^T L^-T, dS symmetric
        Mat<D> P = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i)
#pragma unroll
            for (int j = 0; j <= i; ++j) P(i, j) = (i == j) ? 0.5 * X(i, j) : X(i, j);
        const Mat<D> dLkk = mm(Lkk, P);
        mat_store<D>(od + k * DD, dLkk);
        if (k < n - 1) {
            const Mat<D> Lsk = mat_load<D>(ls + k * DD);
            const Mat<D> R = mat_load<D>(ks + k * DD) - mmT(Lsk, dLkk);     // dKs_k - Ls_k dL_kk^T
            const Mat<D> dLsk = transpose(trsm(Lkk, transpose(R)));         // R L_kk^-T
            mat_store<D>(os + k * DD, dLsk);
            dLs_prev = dLsk;
            Ls_prev = Lsk;
        }
    }
    // ---- backward: tangent of the inverse band
    Mat<D> S, dS;
    {
        const Mat<D> U = trsm(mat_load<D>(ld + (n - 1) * DD), mat_eye<D>());
        const Mat<D> dU = -1.0 * mm(U, mm(mat_load<D>(od + (n - 1) * DD), U));
        S = Tmm(U, U);
        dS = Tmm(dU, U) + Tmm(U, dU);
        mat_store<D>(od + (n - 1) * DD, dS);
    }
    for (int64_t k = n - 2; k >= 0; --k) {
        const Mat<D> U = trsm(mat_load<D>(ld + k * DD), mat_eye<D>());
        const Mat<D> dU = -1.0 * mm(U, mm(mat_load<D>(od + k * DD), U));
        const Mat<D> Lsk = mat_load<D>(ls + k * DD);
        const Mat<D> G = mm(Lsk, U);
        const Mat<D> dG = mm(mat_load<D>(os + k * DD), U) + mm(Lsk, dU);
        const Mat<D> Ssub = -1.0 * mm(S, G);
        const Mat<D> dSsub = -1.0 * (mm(dS, G) + mm(S, dG));
        const Mat<D> Skk = Tmm(U, U) - Tmm(Ssub, G);
        const Mat<D> dSkk = (Tmm(dU, U) + Tmm(U, dU)) - (Tmm(dSsub, G) + Tmm(Ssub, dG));
        mat_store<D>(os + k * DD, dSsub);
        mat_store<D>(od + k * DD, dSkk);
        S = mirror_lower(Skk);
        dS = sym(dSkk);
    }
}

// naturals_to_ssm_params, backward half (ssm_gaussian_transformations.py:443-504) fused in one sweep:
// inverse band -> A_k -> conditional precision -> covariance, chol -> (A^-T)-solve of theta1 -> offsets.
template <int D>
__global__ void k_nat2ssm_bwd(const double* __restrict__ Ld, const double* __restrict__ Ls,
                              const double* __restrict__ th1, const double* __restrict__ thd,
                              const double* __restrict__ ths, double* __restrict__ A, double* __restrict__ bo,
                              double* __restrict__ cholQ, double* __restrict__ mu0, double* __restrict__ cholP0,
                              int64_t B, int64_t N, int64_t* status) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    const int64_t n = N + 1;
    const double* ld = Ld + b * n * DD;
    const double* ls = Ls + b * N * DD;
    const double* t1 = th1 + b * n * D;
    const double* td = thd + b * n * DD;
    const double* ts = ths + b * N * DD;
    bool ok = true;
    Mat<D> S = invband_last(mat_load<D>(ld + N * DD));
    Vec<D> y = vec_load<D>(t1 + N * D);
    for (int64_t k = N; k >= 0; --k) {
        Mat<D> prec = -2.0 * mat_load<D>(td + k * DD);
        if (k < N) {
            Mat<D> Ssub, Skk;
            invband_step(mat_load<D>(ld + k * DD), mat_load<D>(ls + k * DD), S, Ssub, Skk);
            S = Skk;
            const Mat<D> Ak = transition_from_cov(Skk, Ssub);
            mat_store<D>(A + (b * N + k) * DD, Ak);
            prec = prec + Tmm(Ak, -1.0 * mat_load<D>(ts + k * DD));
            y = vec_load<D>(t1 + k * D) + Tmv(Ak, y);
        }
        prec = mirror_lower(prec);
        Mat<D> cov, cc;
        cov_from_precision(prec, cov, cc, ok);
        if (!ok) { report_not_pd(status, b * n + k); ok = true; }
        const Vec<D> off = mv(cov, y);
        if (k == 0) {
            vec_store<D>(mu0 + b * D, off);
            mat_store<D>(cholP0 + b * DD, cc);
        } else {
            vec_store<D>(bo + (b * N + k - 1) * D, off);
            mat_store<D>(cholQ + (b * N + k - 1) * DD, cc);
        }
    }
}

// StateSpaceModel.marginal_means (state_space_model.py:251) + _build_precision (:431-483) + cholesky_band,
// one forward sweep.  Outputs m [B,N+1,D] and the Cholesky factor of the rebuilt precision.
template <int D>
__global__ void k_marg_fwd(const double* __restrict__ A, const double* __restrict__ bo,
                           const double* __restrict__ cholQ, const double* __restrict__ mu0,
                           const double* __restrict__ cholP0, double* __restrict__ m, double* __restrict__ Ld,
                           double* __restrict__ Ls, int64_t B, int64_t N, int64_t* status) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    const int64_t n = N + 1;
    const double* a = A + b * N * DD;
    const double* bb = bo + b * N * D;
    const double* cq = cholQ + b * N * DD;
    double* mm_ = m + b * n * D;
    double* ld = Ld + b * n * DD;
    double* ls = Ls + b * N * DD;
    bool ok = true;
    Vec<D> mk = vec_load<D>(mu0 + b * D);
    vec_store<D>(mm_, mk);
    Mat<D> ck = mat_load<D>(cholP0 + b * DD);
    Mat<D> L, subPrev;
    for (int64_t k = 0; k <= N; ++k) {
        const bool has_next = k < N;
        Mat<D> Ak = mat_zero<D>(), cnext = mat_eye<D>(), diag, sub = mat_zero<D>();
        if (has_next) { Ak = mat_load<D>(a + k * DD); cnext = mat_load<D>(cq + k * DD); }
        precision_blocks(ck, has_next, Ak, cnext, diag, sub);
        if (k == 0) {
            L = chol_first(diag, ok);
        } else {
            Mat<D> Lsub, Lnext;
            chol_step(L, subPrev, diag, Lsub, Lnext, ok);
            mat_store<D>(ls + (k - 1) * DD, Lsub);
            L = Lnext;
        }
        if (!ok) { report_not_pd(status, b * n + k); ok = true; }
        mat_store<D>(ld + k * DD, L);
        if (has_next) {
            mk = vec_load<D>(bb + k * D) + mv(Ak, mk);
            vec_store<D>(mm_ + (k + 1) * D, mk);
            ck = cnext;
            subPrev = sub;
        }
    }
}

// solve_triang_mat: x = L^-1 rhs or L^-T rhs
template <int D>
__global__ void k_btd_solve(const double* __restrict__ Ld, const double* __restrict__ Ls,
                            const double* __restrict__ rhs, double* __restrict__ x, int transpose_, int64_t B,
                            int64_t n) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    const double* ld = Ld + b * n * DD;
    const double* ls = Ls ? Ls + b * (n - 1) * DD : nullptr;
    const double* r = rhs + b * n * D;
    double* xo = x + b * n * D;
    Vec<D> prev = vec_zero<D>();
    if (!transpose_) {
        for (int64_t k = 0; k < n; ++k) {
            Vec<D> rk = vec_load<D>(r + k * D);
            if (ls && k > 0) rk = rk - mv(mat_load<D>(ls + (k - 1) * DD), prev);
            prev = trsv(mat_load<D>(ld + k * DD), rk);
            vec_store<D>(xo + k * D, prev);
        }
    } else {
        for (int64_t k = n - 1; k >= 0; --k) {
            Vec<D> rk = vec_load<D>(r + k * D);
            if (ls && k < n - 1) rk = rk - Tmv(mat_load<D>(ls + k * DD), prev);
            prev = trsvT(mat_load<D>(ld + k * DD), rk);
            vec_store<D>(xo + k * D, prev);
        }
    }
}

// product_band_mat: one thread per (sequence, block row)
template <int D>
__global__ void k_btd_matvec(const double* __restrict__ Md, const double* __restrict__ Ms,
                             const double* __restrict__ x, double* __restrict__ y, int symmetric, int transpose_,
                             int64_t B, int64_t n) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n) return;
    constexpr int DD = D * D;
    const int64_t b = i / n, k = i % n;
    const double* md = Md + b * n * DD;
    const double* ms = Ms ? Ms + b * (n - 1) * DD : nullptr;
    const double* xb = x + b * n * D;
    Mat<D> Mk = mat_load<D>(md + k * DD);
    if (!symmetric) Mk = tril(Mk);
    const Vec<D> xk = vec_load<D>(xb + k * D);
    Vec<D> yk = (!symmetric && transpose_) ? Tmv(Mk, xk) : mv(Mk, xk);
    if (ms) {
        const bool lower_part = symmetric || !transpose_;   // M_{k,k-1} x_{k-1}
        const bool upper_part = symmetric || transpose_;    // M_{k+1,k}^T x_{k+1}
        if (lower_part && k > 0) yk = yk + mv(mat_load<D>(ms + (k - 1) * DD), vec_load<D>(xb + (k - 1) * D));
        if (upper_part && k < n - 1) yk = yk + Tmv(mat_load<D>(ms + k * DD), vec_load<D>(xb + (k + 1) * D));
    }
    vec_store<D>(y + i * D, yk);
}

// abs_log_det: one thread per sequence (0.5 sum log(L_ii^2), block_tri_diag.py:364-366)
template <int D>
__global__ void k_btd_logdet(const double* __restrict__ Ld, double* __restrict__ out, int64_t B, int64_t n) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    double s = 0.0;
    for (int64_t k = 0; k < n; ++k)
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double v = Ld[(b * n + k) * DD + i * D + i];
            s += log(v * v);
        }
    out[b] = 0.5 * s;
}

}  // namespace vdp

namespace vdp {

// Lower-band storage of banded_matrices (block_tri_diag.py:206-237 -> block_to_band): band[r, j] = M[j + r, j],
// row 0 = main diagonal, R = 2 D rows when there is a sub-diagonal (lower bandwidth 2 D - 1), else D rows.
// One thread per band entry (b, r, j), j = c D + i0.
template <int D>
__global__ void k_block_to_band(const double* __restrict__ Md, const double* __restrict__ Ms, double* __restrict__ band,
                                int64_t B, int64_t n) {
    constexpr int DD = D * D;
    const int R = Ms ? 2 * D : D;
    const int64_t cols = n * D;
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * R * cols) return;
    const int64_t b = i / (R * cols), rem = i % (R * cols);
    const int r = (int)(rem / cols);
    const int64_t j = rem % cols, c = j / D;
    const int i0 = (int)(j % D), row = i0 + r;
    double v = 0.0;
    if (row < D) v = Md[(b * n + c) * DD + row * D + i0];
    else if (row < 2 * D && Ms && c < n - 1) v = Ms[(b * (n - 1) + c) * DD + (row - D) * D + i0];
    band[i] = v;
}

// band_to_block: the block diagonal (upper triangles mirrored when symmetric, zero otherwise) and sub-diagonal
template <int D>
__global__ void k_band_to_block(const double* __restrict__ band, double* __restrict__ Md, double* __restrict__ Ms,
                                int symmetric, int64_t B, int64_t n) {
    constexpr int DD = D * D;
    const int R = Ms ? 2 * D : D;
    const int64_t cols = n * D;
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n) return;
    const int64_t b = i / n, c = i % n;
    const double* bd = band + b * R * cols;
#pragma unroll
    for (int row = 0; row < D; ++row)
#pragma unroll
        for (int col = 0; col < D; ++col) {
            double v = 0.0;
            if (row >= col) v = bd[(row - col) * cols + c * D + col];
            else if (symmetric) v = bd[(col - row) * cols + c * D + row];
            Md[i * DD + row * D + col] = v;
        }
    if (Ms && c < n - 1) {
#pragma unroll
        for (int row = 0; row < D; ++row)
#pragma unroll
            for (int col = 0; col < D; ++col)
                Ms[(b * (n - 1) + c) * DD + row * D + col] = bd[(D + row - col) * cols + c * D + col];
    }
}

// Block diagonal of  A_inv^-T P  with A_inv = LowerTriangularBlockTriDiagonal(I, -As) and P the symmetric
// block-tridiagonal (Kd, Ks): the solve_triang_band(..., transpose_left=True) of
// ssm_gaussian_transformations.py:470-481 followed by band_to_block(...)[..., :D, :] -- the conditional
// precisions  Kd_k + A_k^T Ks_k  (k < n-1),  Kd_{n-1}.
template <int D>
__global__ void k_conditional_precisions(const double* __restrict__ As, const double* __restrict__ Kd,
                                         const double* __restrict__ Ks, double* __restrict__ out, int64_t B,
                                         int64_t n) {
    constexpr int DD = D * D;
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n) return;
    const int64_t b = i / n, c = i % n;
    Mat<D> p = mat_load<D>(Kd + i * DD);
    if (c < n - 1) {
        const int64_t t = b * (n - 1) + c;
        p = p + Tmm(mat_load<D>(As + t * DD), mat_load<D>(Ks + t * DD));
    }
    mat_store<D>(out + i * DD, p);
}

}  // namespace vdp
