// Time recursions of the block-tridiagonal layer as per-step device functions.  A "sweep" walks a
// contiguous range of grid points carrying a small state (the carry) from one step to the next; the
// kernels in btd_kernels.cu / cvi_update.cu decide how ranges are assigned to threads.
//
// Reference operations restated here (block form of the banded_matrices kernels, same operation
// order as oracle/gauss_markov_np.py):
//   chol_step        cholesky_band                block_tri_diag.py:428-440
//   invband_step     inverse_from_cholesky_band   block_tri_diag.py:318-337, ssm_gaussian_transformations.py:443-445
//   ssm_from_band    ssm_gaussian_transformations.py:462-504 (A_k, conditional precisions, chols, offsets)
//   precision_blocks state_space_model.py:431-483 (_build_precision)
#pragma once
#include "smallmat.cuh"

namespace vdp {

// ---- forward Cholesky:  carry = L_kk ------------------------------------------------------------
// first block
template <int D>
__device__ __forceinline__ Mat<D> chol_first(const Mat<D>& K0, bool& ok) { return chol(K0, ok); }

// given L_kk, K_{k+1,k} and K_{k+1,k+1}: L_{k+1,k} = K_{k+1,k} L_kk^-T, L_{k+1,k+1} = chol(K - L L^T)
template <int D>
__device__ __forceinline__ void chol_step(const Mat<D>& Lkk, const Mat<D>& Ksub, const Mat<D>& Knext,
                                          Mat<D>& Lsub, Mat<D>& Lnext, bool& ok) {
    Lsub = transpose(trsm(Lkk, transpose(Ksub)));
    const Mat<D> S = Knext - mmT(Lsub, Lsub);
    Lnext = chol(S, ok);
}

// ---- backward inverse band (Takahashi):  carry = Sigma_{k+1,k+1} --------------------------------
template <int D>
__device__ __forceinline__ Mat<D> invband_last(const Mat<D>& Lnn) {
    const Mat<D> U = trsm(Lnn, mat_eye<D>());
    return Tmm(U, U);
}
// Sigma_{k+1,k} = -Sigma_{k+1,k+1} G,  Sigma_kk = U^T U - Sigma_{k+1,k}^T G,  G = L_{k+1,k} L_kk^-1
template <int D>
__device__ __forceinline__ void invband_step(const Mat<D>& Lkk, const Mat<D>& Lsub, const Mat<D>& Snext,
                                             Mat<D>& Ssub, Mat<D>& Skk) {
    const Mat<D> U = trsm(Lkk, mat_eye<D>());
    const Mat<D> G = mm(Lsub, U);
    Ssub = -1.0 * mm(Snext, G);
    Skk = mirror_lower(Tmm(U, U) - Tmm(Ssub, G));
}

// ---- pointwise pieces of naturals_to_ssm_params ------------------------------------------------
// A_k = (Sigma_kk^-1 Sigma_{k,k+1})^T  with Sigma_{k,k+1} = Sigma_{k+1,k}^T   (tf.linalg.solve = LU)
template <int D>
__device__ __forceinline__ Mat<D> transition_from_cov(const Mat<D>& Skk, const Mat<D>& Ssub) {
    return transpose(lu_solve(Skk, transpose(Ssub)));
}
// conditional precision -> covariance and its Cholesky factor (chol, cholesky_solve(I), chol)
template <int D>
__device__ __forceinline__ void cov_from_precision(const Mat<D>& prec, Mat<D>& cov, Mat<D>& cholcov, bool& ok) {
    if constexpr (D == 1) {
        // scalar: chol(1 / p) = p^(-1/2) and 1 / p = (p^(-1/2))^2 from one reciprocal square root
        const double p = prec.a[0];
        if (!(p > 0.0)) ok = false;
        double sp, rs;
        sqrt_rsqrt_fast(p, sp, rs);
        cholcov.a[0] = rs;
        cov.a[0] = rs * rs;
        return;
    }
    const Mat<D> c = chol(prec, ok);
    cov = chol_inverse(c);
    cholcov = chol(cov, ok);
}

// ---- _build_precision blocks --------------------------------------------------------------------
// node k with its own Cholesky factor c_k ([cholP0, cholQ...]) and, if k < N, the outgoing transition
// A_k with cholQ_k:  diag = c_k^-T c_k^-1 (+ A^T Q^-1 A),  sub = -Q^-1 A
template <int D>
__device__ __forceinline__ void precision_blocks(const Mat<D>& ck, bool has_next, const Mat<D>& A,
                                                 const Mat<D>& cholQnext, Mat<D>& diag, Mat<D>& sub) {
    diag = chol_inverse(ck);
    if (has_next) {
        const Mat<D> inv_q_a = chol_solve(cholQnext, A);
        diag = diag + Tmm(A, inv_q_a);
        sub = -1.0 * inv_q_a;
    }
}

}  // namespace vdp
