// StateSpaceModel.kl_divergence / log_pdf / sample (markovflow/state_space_model.py:528-593, 485-526, 298-324):
// the Gauss-Markov densities the linear-prior KL (CVISitesSSM.KL_q_p, tests/unit/test_sde.py:149-197) and the
// reference's sampling tests are built on.  Included by vdp_all.inl (translation unit "dist").
#include <curand_kernel.h>

namespace vdp {

// Per grid point i = (b, k) the contribution to
//   KL(dist1 || dist2) = 0.5 (tr(P2 P1^-1) + (mu2 - mu1)^T P2 (mu2 - mu1) - n D - log|P2| + log|P1|)
// in the reference's order of operations (state_space_model.py:562-591):
//   trace:  sum(P2_kk o S1_kk) + 2 sum(P2_{k+1,k} o (A1_k S1_kk))          (subsequent_covariances, :326-341)
//   Mahalanobis: || (L2^T d)_k ||^2, (L2^T d)_k = L2_kk^T d_k + L2_{k+1,k}^T d_{k+1}, L2 = cholesky_band(P2)
//   log-determinants from the Cholesky factors of the two state-space models (:343-373)
// P2 = (-2 thd2, -ths2) are the blocks ssm_to_naturals builds (the same statements as _build_precision, :431-483).
template <int D>
__global__ void k_ssm_kl_terms(const double* __restrict__ A1, const double* __restrict__ m1,
                               const double* __restrict__ S1, const double* __restrict__ c1_0,
                               const double* __restrict__ c1_q, const double* __restrict__ m2,
                               const double* __restrict__ thd2, const double* __restrict__ ths2,
                               const double* __restrict__ Ld2, const double* __restrict__ Ls2,
                               const double* __restrict__ c2_0, const double* __restrict__ c2_q,
                               double* __restrict__ terms, int64_t B, int64_t N) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t n = N + 1;
    if (i >= B * n) return;
    constexpr int DD = D * D;
    int64_t b, k;
    split_index(i, n, b, k);
    const Mat<D> S = mat_load<D>(S1 + i * DD);
    const Mat<D> Pd = -2.0 * mat_load<D>(thd2 + i * DD);
    double trace = 0.0;
#pragma unroll
    for (int e = 0; e < DD; ++e) trace += Pd.a[e] * S.a[e];
    const Vec<D> d = vec_load<D>(m2 + i * D) - vec_load<D>(m1 + i * D);
    Vec<D> v = Tmv(mat_load<D>(Ld2 + i * DD), d);
    if (k < N) {
        const int64_t t = b * N + k;
        const Mat<D> Ps = -1.0 * mat_load<D>(ths2 + t * DD);
        const Mat<D> sub = mm(mat_load<D>(A1 + t * DD), S);
        double tr2 = 0.0;
#pragma unroll
        for (int e = 0; e < DD; ++e) tr2 += Ps.a[e] * sub.a[e];
        trace += 2.0 * tr2;
        const Vec<D> dn = vec_load<D>(m2 + (i + 1) * D) - vec_load<D>(m1 + (i + 1) * D);
        v = v + Tmv(mat_load<D>(Ls2 + t * DD), dn);
    }
    const double* c1 = (k == 0) ? c1_0 + b * DD : c1_q + (b * N + k - 1) * DD;
    const double* c2 = (k == 0) ? c2_0 + b * DD : c2_q + (b * N + k - 1) * DD;
    double ld = 0.0;   // -log|P2| + log|P1| of this node = sum log c2_ii^2 - sum log c1_ii^2
#pragma unroll
    for (int j = 0; j < D; ++j) ld += log(c2[j * D + j] * c2[j * D + j]) - log(c1[j * D + j] * c1[j * D + j]);
    terms[i] = 0.5 * (trace + dot(v, v) - (double)D + ld);
}

// log p(x_0) and log p(x_{k+1} | x_k) of a trajectory (state_space_model.py:485-512; tfp MultivariateNormalTriL
// log_prob restated): states [S, B, N+1, D], terms [S, B, N+1]
template <int D>
__global__ void k_ssm_logpdf_terms(const double* __restrict__ A, const double* __restrict__ bo,
                                   const double* __restrict__ cholQ, const double* __restrict__ mu0,
                                   const double* __restrict__ cholP0, const double* __restrict__ x,
                                   double* __restrict__ terms, int64_t S, int64_t B, int64_t N) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t n = N + 1;
    if (i >= S * B * n) return;
    constexpr int DD = D * D;
    const int64_t k = i % n, b = (i / n) % B;
    const Vec<D> xk = vec_load<D>(x + i * D);
    Vec<D> loc;
    Mat<D> tril;
    if (k == 0) {
        loc = vec_load<D>(mu0 + b * D);
        tril = mat_load<D>(cholP0 + b * DD);
    } else {
        const int64_t t = b * N + k - 1;
        loc = mv(mat_load<D>(A + t * DD), vec_load<D>(x + (i - 1) * D)) + vec_load<D>(bo + t * D);
        tril = mat_load<D>(cholQ + t * DD);
    }
    const Vec<D> z = trsv(tril, xk - loc);
    double logdet = 0.0;
#pragma unroll
    for (int j = 0; j < D; ++j) logdet += log(fabs(tril(j, j)));
    terms[i] = -0.5 * dot(z, z) - logdet - 0.5 * (double)D * 1.8378770664093453;   // log(2 pi)
}

// StateSpaceModel.sample (state_space_model.py:298-324): x = A_inv_block^-1 (b_concat + chol_concat eps), i.e.
//   x_0 = mu_0 + chol P_0 eps_0,   x_{k+1} = A_k x_k + b_k + chol Q_k eps_{k+1},
// one thread per (sample, sequence).  eps [S, B, N+1, D] from the caller, or Philox4x32-10 (sub-sequence = s B + b).
template <int D>
__global__ void k_ssm_sample(const double* __restrict__ A, const double* __restrict__ bo,
                             const double* __restrict__ cholQ, const double* __restrict__ mu0,
                             const double* __restrict__ cholP0, const double* __restrict__ eps,
                             unsigned long long seed, double* __restrict__ out, int64_t S, int64_t B, int64_t N) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= S * B) return;
    constexpr int DD = D * D;
    const int64_t b = i % B, n = N + 1;
    curandStatePhilox4_32_10_t rng;
    if (!eps) curand_init(seed, (unsigned long long)i, 0ULL, &rng);
    Vec<D> x = vec_zero<D>();
    for (int64_t k = 0; k < n; ++k) {
        Vec<D> e;
        if (eps) {
            e = vec_load<D>(eps + (i * n + k) * D);
        } else {
#pragma unroll
            for (int j = 0; j < D; j += 2) {
                const double2 z = curand_normal2_double(&rng);
                e(j) = z.x;
                if (j + 1 < D) e(j + 1) = z.y;
            }
        }
        if (k == 0) {
            x = vec_load<D>(mu0 + b * D) + mv(mat_load<D>(cholP0 + b * DD), e);
        } else {
            const int64_t t = b * N + k - 1;
            x = mv(mat_load<D>(A + t * DD), x) + (vec_load<D>(bo + t * D) + mv(mat_load<D>(cholQ + t * DD), e));
        }
        vec_store<D>(out + (i * n + k) * D, x);
    }
}

}  // namespace vdp

extern "C" {

size_t vdp_ssm_kl_divergence_workspace(int64_t B, int64_t N, int D) {
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    return 3 * padded(n * D) + 3 * padded(n * DD) + padded(t * DD) + padded(n * DD) + padded(t * DD) + padded(n) +
           kSegScratchBytes;
}

int vdp_ssm_kl_divergence(const double* A1, const double* b1, const double* cholQ1, const double* mu01,
                          const double* cholP01, const double* A2, const double* b2, const double* cholQ2,
                          const double* mu02, const double* cholP02, double* kl, void* workspace,
                          size_t workspace_bytes, int64_t B, int64_t N, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(A1 && b1 && cholQ1 && mu01 && cholP01 && A2 && b2 && cholQ2 && mu02 && cholP02 && kl,
                "pointers");
    VDP_REQUIRE(B >= 0 && N >= 1, "B, N");
    if (B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    double* m1 = ar.take(n * D);
    double* m2 = ar.take(n * D);
    double* th1 = ar.take(n * D);
    double* S1 = ar.take(n * DD);
    double* S2 = ar.take(n * DD);
    double* thd = ar.take(n * DD);
    double* ths = ar.take(t * DD);
    double* Ld = ar.take(n * DD);
    double* Ls = ar.take(t * DD);
    double* terms = ar.take(n);
    void* seg = ar.take_bytes(kSegScratchBytes);
    if (!workspace || !seg) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    const SegScratch sc(seg, kSegScratchBytes);
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(D, {
        // marginals of dist1 (the factor scratch is reused), then of dist2: its means and the Cholesky factor of its
        // precision (Ld, Ls) are what the Mahalanobis term needs
        int rc = run_marginals<kD>(A1, b1, cholQ1, mu01, cholP01, m1, S1, Ld, Ls, B, N, status, st, sc);
        if (rc) return rc;
        rc = run_marginals<kD>(A2, b2, cholQ2, mu02, cholP02, m2, S2, Ld, Ls, B, N, status, st, sc);
        if (rc) return rc;
        k_ssm_to_naturals<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(
            A2, b2, cholQ2, mu02, cholP02, nullptr, nullptr, nullptr, th1, thd, ths, B, N);
        VDP_CHECK_LAUNCH();
        k_ssm_kl_terms<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(
            A1, m1, S1, cholP01, cholQ1, m2, thd, ths, Ld, Ls, cholP02, cholQ2, terms, B, N);
        VDP_CHECK_LAUNCH();
    });
    launch_reduce_terms(terms, N + 1, kl, B, st);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

size_t vdp_ssm_log_pdf_workspace(int64_t S, int64_t B, int64_t N) { return padded((size_t)S * B * (N + 1)); }

int vdp_ssm_log_pdf(const double* A, const double* b, const double* cholQ, const double* mu0, const double* cholP0,
                    const double* states, double* out, void* workspace, size_t workspace_bytes, int64_t S, int64_t B,
                    int64_t N, int D, void* stream) {
    VDP_REQUIRE(A && b && cholQ && mu0 && cholP0 && states && out && S >= 0 && B >= 0 && N >= 1, "pointers, S, B, N");
    if (S * B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    double* terms = workspace ? ar.take((size_t)S * B * (N + 1)) : nullptr;
    if (!terms) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(D, (k_ssm_logpdf_terms<kD><<<grid_for(S * B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(
                          A, b, cholQ, mu0, cholP0, states, terms, S, B, N)));
    VDP_CHECK_LAUNCH();
    k_reduce_terms<<<(unsigned)(S * B), 256, 0, st>>>(terms, N + 1, out);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_ssm_sample(const double* A, const double* b, const double* cholQ, const double* mu0, const double* cholP0,
                   const double* eps, unsigned long long seed, double* out, int64_t S, int64_t B, int64_t N, int D,
                   void* stream) {
    VDP_REQUIRE(A && b && cholQ && mu0 && cholP0 && out && S >= 0 && B >= 0 && N >= 1, "pointers, S, B, N");
    if (S * B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_ssm_sample<kD><<<grid_for(S * B, 64), 64, 0, as_stream(stream)>>>(
                          A, b, cholQ, mu0, cholP0, eps, seed, out, S, B, N)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

}  // extern "C"
