// Pointwise-in-time kernels: SSM <-> natural / expectation parameters, site assembly, data sites.
// One thread per (sequence, grid point).
#pragma once
#include "common.cuh"

namespace vdp {

// Natural parameters one node of a state-space model contributes (ssm_gaussian_transformations.py:182-253):
//   t1 = C_k^-1 o_k - A_k^T Q_k^-1 b_k,   td = -1/2 (C_k^-1 + A_k^T Q_k^-1 A_k),   ts = Q_k^-1 A_k
// with C_k = ck ck^T the covariance the node's own factor stands for (chol P0 for node 0, chol Q_{k-1} after),
// o_k its offset (mu0 / b_{k-1}) and (A_k, cn = chol Q_k, bn = b_k) the transition that leaves the node (has_next).
// D = 1 takes two reciprocals instead of the triangular-solve chain.
template <int D>
__device__ __forceinline__ void ssm_node_naturals(const Mat<D>& ck, const Vec<D>& ok, bool has_next, const Mat<D>& Ak,
                                                  const Mat<D>& cn, const Vec<D>& bn, Vec<D>& t1, Mat<D>& td,
                                                  Mat<D>& ts) {
    if constexpr (D == 1) {
        const double r = rcp_fast(ck.a[0]);
        const double ir2 = r * r;
        double v1 = ok.v[0] * ir2, vd = ir2;
        ts.a[0] = 0.0;
        if (has_next) {
            const double rn = rcp_fast(cn.a[0]);
            const double in2 = rn * rn;
            const double la = Ak.a[0] * rn;
            ts.a[0] = Ak.a[0] * in2;
            v1 = fma(-Ak.a[0], bn.v[0] * in2, v1);
            vd = fma(la, la, vd);
        }
        t1.v[0] = v1;
        td.a[0] = -0.5 * vd;
    } else {
        t1 = chol_solve(ck, ok);
        td = chol_inverse(ck);
        ts = mat_zero<D>();
        if (has_next) {
            const Vec<D> tn = chol_solve(cn, bn);
            const Mat<D> LinvA = trsm(cn, Ak);
            ts = trsmT(cn, LinvA);
            t1 = t1 - Tmv(Ak, tn);
            td = td + Tmm(LinvA, LinvA);
        }
        td = -0.5 * td;
    }
}

// ssm_to_naturals (ssm_gaussian_transformations.py:182-253) for grid point k of sequence b, optionally
// summed with the girsanov sites: theta = nat(p) + g  (variational_cvi_sde.py:161-174; the data sites
// are scattered on top by k_scatter_data, in the reference's left-to-right order).
template <int D>
__global__ void k_ssm_to_naturals(const double* __restrict__ A, const double* __restrict__ bo,
                                  const double* __restrict__ cholQ, const double* __restrict__ mu0,
                                  const double* __restrict__ cholP0, const double* __restrict__ g1,
                                  const double* __restrict__ gd, const double* __restrict__ gs,
                                  double* __restrict__ th1, double* __restrict__ thd, double* __restrict__ ths,
                                  int64_t B, int64_t N) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t n = N + 1;
    if (i >= B * n) return;
    constexpr int DD = D * D;
    int64_t b, k;
    split_index(i, n, b, k);
    const Mat<D> ck = (k == 0) ? mat_load<D>(cholP0 + b * DD) : mat_load<D>(cholQ + (b * N + k - 1) * DD);
    const Vec<D> ok = (k == 0) ? vec_load<D>(mu0 + b * D) : vec_load<D>(bo + (b * N + k - 1) * D);
    Mat<D> Ak = mat_zero<D>(), cn = mat_eye<D>(), td, ts;
    Vec<D> bn = vec_zero<D>(), t1;
    if (k < N) {
        Ak = mat_load<D>(A + (b * N + k) * DD);
        cn = mat_load<D>(cholQ + (b * N + k) * DD);
        bn = vec_load<D>(bo + (b * N + k) * D);
    }
    ssm_node_naturals<D>(ck, ok, k < N, Ak, cn, bn, t1, td, ts);
    if (k < N) {
        if (gs) ts = ts + mat_load<D>(gs + (b * N + k) * DD);
        mat_store<D>(ths + (b * N + k) * DD, ts);
    }
    if (g1) t1 = t1 + vec_load<D>(g1 + i * D);
    if (gd) td = td + mat_load<D>(gd + i * DD);
    vec_store<D>(th1 + i * D, t1);
    mat_store<D>(thd + i * DD, td);
}

// tf.scatter_nd of the data sites onto the grid, added to theta (duplicates accumulate)
template <int D>
__global__ void k_scatter_data(const double* __restrict__ d1, const double* __restrict__ d2,
                               const int64_t* __restrict__ obs_idx, double* __restrict__ th1,
                               double* __restrict__ thd, int64_t B, int64_t N, int64_t n_obs) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_obs) return;
    constexpr int DD = D * D;
    const int64_t b = i / n_obs;
    const int64_t k = obs_idx[i];
    if (k < 0 || k > N) return;
    const int64_t node = b * (N + 1) + k;
#pragma unroll
    for (int e = 0; e < D; ++e) atomicAdd(th1 + node * D + e, d1[i * D + e]);
#pragma unroll
    for (int e = 0; e < DD; ++e) atomicAdd(thd + node * DD + e, d2[i * DD + e]);
}

// The naturals theta of the previous site update are still in the workspace (VDP_FLAG_NATURALS_CARRIED) and only the
// data sites have changed since: rewrite theta_1 / theta_d at the observed grid points, in the statement order of the
// site-update kernel's epilogue ((nat(p) + g) + sum of the data sites at the node), and likewise the dense copy of the
// data sites (sc1 / scd, zero away from the observations) that kernel reads.  One thread per observation entry; of
// several entries on one grid point the first does the work.
template <int D>
__global__ void k_patch_naturals(const double* __restrict__ A, const double* __restrict__ bo,
                                 const double* __restrict__ cholQ, const double* __restrict__ mu0,
                                 const double* __restrict__ cholP0, const double* __restrict__ g1,
                                 const double* __restrict__ gd, const double* __restrict__ d1,
                                 const double* __restrict__ d2, const int64_t* __restrict__ obs_idx,
                                 double* __restrict__ th1, double* __restrict__ thd, double* __restrict__ sc1,
                                 double* __restrict__ scd, int64_t B, int64_t N, int64_t n_obs) {
    const int64_t o = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (o >= B * n_obs) return;
    constexpr int DD = D * D;
    const int64_t b = o / n_obs, j = o - b * n_obs;
    const int64_t k = obs_idx[o];
    if (k < 0 || k > N) return;
    const int64_t* row = obs_idx + b * n_obs;
    Vec<D> s1 = vec_zero<D>();
    Mat<D> s2 = mat_zero<D>();
    for (int64_t jj = 0; jj < n_obs; ++jj) {
        if (row[jj] != k) continue;
        if (jj < j) return;
        s1 = s1 + vec_load<D>(d1 + (b * n_obs + jj) * D);
        s2 = s2 + mat_load<D>(d2 + (b * n_obs + jj) * DD);
    }
    const int64_t i = b * (N + 1) + k, tt = b * N + k;
    const Mat<D> ck = (k == 0) ? mat_load<D>(cholP0 + b * DD) : mat_load<D>(cholQ + (tt - 1) * DD);
    const Vec<D> ok = (k == 0) ? vec_load<D>(mu0 + b * D) : vec_load<D>(bo + (tt - 1) * D);
    Mat<D> Ak = mat_zero<D>(), cn = mat_eye<D>(), td, ts;
    Vec<D> bn = vec_zero<D>(), t1;
    if (k < N) {
        Ak = mat_load<D>(A + tt * DD);
        cn = mat_load<D>(cholQ + tt * DD);
        bn = vec_load<D>(bo + tt * D);
    }
    ssm_node_naturals<D>(ck, ok, k < N, Ak, cn, bn, t1, td, ts);
    t1 = t1 + vec_load<D>(g1 + i * D);
    td = td + mat_load<D>(gd + i * DD);
    vec_store<D>(th1 + i * D, t1 + s1);
    mat_store<D>(thd + i * DD, td + s2);
    vec_store<D>(sc1 + i * D, s1);
    mat_store<D>(scd + i * DD, s2);
}

// girsanov sites from the natural parameters they are part of (variational_cvi_sde.py:161-174 read backwards):
//   g = theta - nat(p) - scatter(data sites),   sub-diagonal: g_s = theta_s - nat_s(p)
// sc1 / scd = the dense copy of the data sites (zero away from the observations).
template <int D>
__global__ void k_sites_from_naturals(const double* __restrict__ A, const double* __restrict__ bo,
                                      const double* __restrict__ cholQ, const double* __restrict__ mu0,
                                      const double* __restrict__ cholP0, const double* __restrict__ th1,
                                      const double* __restrict__ thd, const double* __restrict__ ths,
                                      const double* __restrict__ sc1, const double* __restrict__ scd,
                                      double* __restrict__ g1, double* __restrict__ gd, double* __restrict__ gs,
                                      int64_t B, int64_t N) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t n = N + 1;
    if (i >= B * n) return;
    constexpr int DD = D * D;
    int64_t b, k;
    split_index(i, n, b, k);
    const Mat<D> ck = (k == 0) ? mat_load<D>(cholP0 + b * DD) : mat_load<D>(cholQ + (b * N + k - 1) * DD);
    const Vec<D> ok = (k == 0) ? vec_load<D>(mu0 + b * D) : vec_load<D>(bo + (b * N + k - 1) * D);
    Mat<D> Ak = mat_zero<D>(), cn = mat_eye<D>(), td, ts;
    Vec<D> bn = vec_zero<D>(), t1;
    if (k < N) {
        Ak = mat_load<D>(A + (b * N + k) * DD);
        cn = mat_load<D>(cholQ + (b * N + k) * DD);
        bn = vec_load<D>(bo + (b * N + k) * D);
    }
    ssm_node_naturals<D>(ck, ok, k < N, Ak, cn, bn, t1, td, ts);
    if (k < N) mat_store<D>(gs + (b * N + k) * DD, mat_load<D>(ths + (b * N + k) * DD) - ts);
    vec_store<D>(g1 + i * D, (vec_load<D>(th1 + i * D) - vec_load<D>(sc1 + i * D)) - t1);
    mat_store<D>(gd + i * DD, (mat_load<D>(thd + i * DD) - mat_load<D>(scd + i * DD)) - td);
}

// ssm_to_expectations (ssm_gaussian_transformations.py:32-89)
template <int D>
__global__ void k_ssm_to_expectations(const double* __restrict__ A, const double* __restrict__ m,
                                      const double* __restrict__ Sd, double* __restrict__ e1,
                                      double* __restrict__ ed, double* __restrict__ es, int64_t B, int64_t N) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t n = N + 1;
    if (i >= B * n) return;
    constexpr int DD = D * D;
    int64_t b, k;
    split_index(i, n, b, k);
    const Vec<D> mk = vec_load<D>(m + i * D);
    const Mat<D> Sk = mat_load<D>(Sd + i * DD);
    vec_store<D>(e1 + i * D, mk);
    mat_store<D>(ed + i * DD, Sk + outer(mk, mk));
    if (k < N) {
        const Mat<D> Ak = mat_load<D>(A + (b * N + k) * DD);
        const Vec<D> mn = vec_load<D>(m + (i + 1) * D);
        mat_store<D>(es + (b * N + k) * DD, mm(Ak, Sk) + outer(mn, mk));
    }
}

// expectations_to_ssm_params (ssm_gaussian_transformations.py:93-178), one thread per transition
// (thread k = N of each sequence writes nothing; thread 0 also writes mu0 / cholP0)
template <int D>
struct ExpToSsm {
    Mat<D> Sk, Sn, Csub, LS, A, Qc, LQ;
    Vec<D> mk, mn, b;
};
template <int D>
__device__ __forceinline__ void exp_to_ssm_step(const Vec<D>& e1k, const Vec<D>& e1n, const Mat<D>& edk,
                                                const Mat<D>& edn, const Mat<D>& esk, ExpToSsm<D>& r, bool& ok) {
    r.mk = e1k;
    r.mn = e1n;
    r.Sk = edk - outer(e1k, e1k);
    r.Sn = edn - outer(e1n, e1n);
    r.Csub = transpose(esk) - outer(e1k, e1n);
    r.LS = chol(r.Sk, ok);
    r.A = transpose(chol_solve(r.LS, r.Csub));
    r.b = e1n - mv(r.A, e1k);
    r.Qc = r.Sn - mm(r.A, mmT(r.Sk, r.A));
    r.LQ = chol(r.Qc, ok);
}

template <int D>
__global__ void k_expectations_to_ssm(const double* __restrict__ e1, const double* __restrict__ ed,
                                      const double* __restrict__ es, double* __restrict__ A,
                                      double* __restrict__ bo, double* __restrict__ cholQ,
                                      double* __restrict__ mu0, double* __restrict__ cholP0, int64_t B, int64_t N,
                                      int64_t* status) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= B * N) return;
    constexpr int DD = D * D;
    const int64_t b = t / N, k = t % N;
    const int64_t i = b * (N + 1) + k;
    ExpToSsm<D> r;
    bool ok = true;
    exp_to_ssm_step(vec_load<D>(e1 + i * D), vec_load<D>(e1 + (i + 1) * D), mat_load<D>(ed + i * DD),
                    mat_load<D>(ed + (i + 1) * DD), mat_load<D>(es + t * DD), r, ok);
    if (!ok) report_not_pd(status, i);
    mat_store<D>(A + t * DD, r.A);
    vec_store<D>(bo + t * D, r.b);
    mat_store<D>(cholQ + t * DD, r.LQ);
    if (k == 0) {
        vec_store<D>(mu0 + b * D, r.mk);
        mat_store<D>(cholP0 + b * DD, r.LS);
    }
}

// update_data_sites (variational_cvi_sde.py:301-314) for the Gaussian likelihood
// (multivariate_gaussian.py:82-116, variational_cvi.py:448-462):
//   dVE/dmu = Sigma^-1 (y - mu),  dVE/dS = -1/2 Sigma^-1,  grads wrt (mu, S + mu mu^T) = (g_mu - 2 g_S mu, g_S)
template <int D>
__global__ void k_data_site_update(const double* __restrict__ fx_mus, const int64_t* __restrict__ obs_idx,
                                   const double* __restrict__ obs_y, Mat<D> lik_chol, double lr,
                                   double* __restrict__ d1, double* __restrict__ d2, int64_t B, int64_t N,
                                   int64_t n_obs, const double* __restrict__ lr_dev = nullptr,
                                   const int32_t* __restrict__ active = nullptr) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_obs) return;
    constexpr int DD = D * D;
    const int64_t b = i / n_obs;
    if (lr_dev) lr = (!active || active[b]) ? lr_dev[b] : 0.0;   // per-sequence step size held on the device
    const int64_t k = obs_idx[i];
    const Vec<D> mu = vec_load<D>(fx_mus + (b * (N + 1) + k) * D);
    const Mat<D> inv = chol_inverse(lik_chol);
    const Vec<D> gmu = mv(sym(inv), vec_load<D>(obs_y + i * D) - mu);
    const Mat<D> gS = -0.5 * inv;
    const Vec<D> g1 = gmu - 2.0 * mv(gS, mu);
    const Vec<D> o1 = vec_load<D>(d1 + i * D);
    const Mat<D> o2 = mat_load<D>(d2 + i * DD);
    vec_store<D>(d1 + i * D, (1.0 - lr) * o1 + lr * g1);
    mat_store<D>(d2 + i * DD, (1.0 - lr) * o2 + lr * gS);
}

// Steady state of the carried route with the girsanov sites held implicitly in theta (VDP_FLAG_SITES_IN_NATURALS:
// theta is the state, g = theta - nat(p) - scatter(data sites) is formed on request): update_data_sites
// (variational_cvi_sde.py:301-314, the statements of k_data_site_update) and the patch of theta in ONE launch with one
// block per sequence.  theta changes by the change of the data sites at the observed grid points, whose previous sum
// is the dense copy's (sc1 / scd) current entry; of several entries on one grid point the first does the work.  The
// sequence's grid indices sit in shared memory for that duplicate scan, and the data sites of all its observations
// are updated before any of them is summed.  n_obs <= kObsMaxShared.
constexpr int kObsBlock = 64;
constexpr int kObsMaxShared = 128;
template <int D>
__global__ void __launch_bounds__(kObsBlock) k_obs_update_patch(
    const double* __restrict__ fx_mus, const int64_t* __restrict__ obs_idx, const double* __restrict__ obs_y,
    Mat<D> lik_chol, double lr, const double* __restrict__ lr_dev, const int32_t* __restrict__ active,
    int skip_data_update, double* __restrict__ d1, double* __restrict__ d2, double* __restrict__ th1,
    double* __restrict__ thd, double* __restrict__ sc1, double* __restrict__ scd, int64_t N, int64_t n_obs) {
    __shared__ int64_t row[kObsMaxShared];
    constexpr int DD = D * D;
    const int64_t b = blockIdx.x;
    if (lr_dev) lr = (!active || active[b]) ? lr_dev[b] : 0.0;
    for (int64_t j = threadIdx.x; j < n_obs; j += kObsBlock) {
        const int64_t i = b * n_obs + j;
        const int64_t k = obs_idx[i];
        row[j] = k;
        if (skip_data_update) continue;
        const Vec<D> mu = vec_load<D>(fx_mus + (b * (N + 1) + k) * D);
        const Mat<D> inv = chol_inverse(lik_chol);
        const Vec<D> gmu = mv(sym(inv), vec_load<D>(obs_y + i * D) - mu);
        const Mat<D> gS = -0.5 * inv;
        const Vec<D> g1 = gmu - 2.0 * mv(gS, mu);
        const Vec<D> o1 = vec_load<D>(d1 + i * D);
        const Mat<D> o2 = mat_load<D>(d2 + i * DD);
        vec_store<D>(d1 + i * D, (1.0 - lr) * o1 + lr * g1);
        mat_store<D>(d2 + i * DD, (1.0 - lr) * o2 + lr * gS);
    }
    __syncthreads();
    for (int64_t j = threadIdx.x; j < n_obs; j += kObsBlock) {
        const int64_t k = row[j];
        if (k < 0 || k > N) continue;
        Vec<D> s1 = vec_zero<D>();
        Mat<D> s2 = mat_zero<D>();
        bool mine = true;
        for (int64_t jj = 0; jj < n_obs; ++jj) {
            if (row[jj] != k) continue;
            if (jj < j) { mine = false; break; }
            s1 = s1 + vec_load<D>(d1 + (b * n_obs + jj) * D);
            s2 = s2 + mat_load<D>(d2 + (b * n_obs + jj) * DD);
        }
        if (!mine) continue;
        const int64_t i = b * (N + 1) + k;
        vec_store<D>(th1 + i * D, vec_load<D>(th1 + i * D) + (s1 - vec_load<D>(sc1 + i * D)));
        mat_store<D>(thd + i * DD, mat_load<D>(thd + i * DD) + (s2 - mat_load<D>(scd + i * DD)));
        vec_store<D>(sc1 + i * D, s1);
        mat_store<D>(scd + i * DD, s2);
    }
}
// The data-site part of the girsanov update when the sites are held in theta: the update g += lr (scatter(data sites)
// - dKL/d eta) (variational_cvi_sde.py:286-292) moves theta by lr * scatter(data sites) at the observed grid points --
// applied here, one block per sequence (the first of several entries on one grid point does the work, from the dense
// copy's summed entry) -- and by -lr * dKL/d eta everywhere, which the forward pass that follows applies.
template <int D>
__global__ void __launch_bounds__(kObsBlock) k_add_data_step_seq(
    const int64_t* __restrict__ obs_idx, const double* __restrict__ sc1, const double* __restrict__ scd,
    double* __restrict__ th1, double* __restrict__ thd, double lr, const double* __restrict__ lr_dev,
    const int32_t* __restrict__ active, int64_t N, int64_t n_obs) {
    __shared__ int64_t row[kObsMaxShared];
    constexpr int DD = D * D;
    const int64_t b = blockIdx.x;
    for (int64_t j = threadIdx.x; j < n_obs; j += kObsBlock) row[j] = obs_idx[b * n_obs + j];
    if (lr_dev) lr = (!active || active[b]) ? lr_dev[b] : 0.0;
    __syncthreads();
    for (int64_t j = threadIdx.x; j < n_obs; j += kObsBlock) {
        const int64_t k = row[j];
        if (k < 0 || k > N) continue;
        bool mine = true;
        for (int64_t jj = 0; jj < j; ++jj)
            if (row[jj] == k) { mine = false; break; }
        if (!mine) continue;
        const int64_t i = b * (N + 1) + k;
        vec_store<D>(th1 + i * D, vec_load<D>(th1 + i * D) + lr * vec_load<D>(sc1 + i * D));
        mat_store<D>(thd + i * DD, mat_load<D>(thd + i * DD) + lr * mat_load<D>(scd + i * DD));
    }
}

// variational_expectation (variational_cvi_sde.py:319-337): one warp per sequence, the lanes stride over the
// observations and their partial sums are added in a fixed shuffle tree (deterministic)
template <int D>
__global__ void k_variational_expectation(const double* __restrict__ fx_mus, const double* __restrict__ fx_covs,
                                          const int64_t* __restrict__ obs_idx, const double* __restrict__ obs_y,
                                          Mat<D> lik_chol, double* __restrict__ ve, int64_t B, int64_t N,
                                          int64_t n_obs) {
    const int64_t b = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / 32;
    const int lane = threadIdx.x % 32;
    if (b >= B) return;
    constexpr int DD = D * D;
    const Mat<D> inv = chol_inverse(lik_chol);
    double logdet = 0.0;
#pragma unroll
    for (int e = 0; e < D; ++e) logdet += log(fabs(lik_chol(e, e)));
    double acc = 0.0;
    for (int64_t j = lane; j < n_obs; j += 32) {
        const int64_t i = b * n_obs + j;
        const int64_t node = b * (N + 1) + obs_idx[i];
        const Vec<D> z = trsv(lik_chol, vec_load<D>(obs_y + i * D) - vec_load<D>(fx_mus + node * D));
        const Mat<D> S = mat_load<D>(fx_covs + node * DD);
        double tr = 0.0;
#pragma unroll
        for (int e = 0; e < DD; ++e) tr += inv.a[e] * S.a[e];
        acc += -0.5 * tr + (-0.5 * dot(z, z) - logdet - 0.5 * D * 1.8378770664093453);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if (lane == 0) ve[b] = acc;
}


// The same sum for few sequences with many observations (one long chain): one block per sequence, the threads stride
// over the observations, fixed tree over the block.
template <int D>
__global__ void __launch_bounds__(256) k_variational_expectation_block(
    const double* __restrict__ fx_mus, const double* __restrict__ fx_covs, const int64_t* __restrict__ obs_idx,
    const double* __restrict__ obs_y, Mat<D> lik_chol, double* __restrict__ ve, int64_t B, int64_t N, int64_t n_obs) {
    __shared__ double sh[256];
    const int64_t b = blockIdx.x;
    constexpr int DD = D * D;
    const Mat<D> inv = chol_inverse(lik_chol);
    double logdet = 0.0;
#pragma unroll
    for (int e = 0; e < D; ++e) logdet += log(fabs(lik_chol(e, e)));
    double acc = 0.0;
    for (int64_t j = threadIdx.x; j < n_obs; j += 256) {
        const int64_t i = b * n_obs + j;
        const int64_t node = b * (N + 1) + obs_idx[i];
        const Vec<D> z = trsv(lik_chol, vec_load<D>(obs_y + i * D) - vec_load<D>(fx_mus + node * D));
        const Mat<D> S = mat_load<D>(fx_covs + node * DD);
        double tr = 0.0;
#pragma unroll
        for (int e = 0; e < DD; ++e) tr += inv.a[e] * S.a[e];
        acc += -0.5 * tr + (-0.5 * dot(z, z) - logdet - 0.5 * D * 1.8378770664093453);
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) ve[b] = sh[0];
}

// Vector-Jacobian product of expectations_to_ssm_params (what tape.gradient(ssm_params, etas,
// output_gradients=dL_dssm) returns in ssm_natgrad.py:168), pointwise in time.  Transition k touches nodes k
// and k+1; thread i = (b, k) recomputes the two transitions next to its node and keeps the parts that land
// on it, so no neighbour exchange is needed.
template <int D>
struct SsmCotangent { Mat<D> gA, gLQ; Vec<D> gb; };

// adjoint of one transition: contributions to (m_k, S_k) ["left"], (m_{k+1}, S_{k+1}) ["right"] and eta_s[k]
template <int D>
__device__ __forceinline__ void exp_to_ssm_step_vjp(const ExpToSsm<D>& r, const SsmCotangent<D>& c, Vec<D>& Lgm,
                                                    Mat<D>& LgS, Vec<D>& Rgm, Mat<D>& RgS, Mat<D>& ges) {
    const Mat<D> gQc = chol_adjoint(r.LQ, tril(c.gLQ));
    Mat<D> gA = c.gA;
    // b = m_{k+1} - A m_k
    Rgm = c.gb;
    gA = gA - outer(c.gb, r.mk);
    Lgm = -1.0 * Tmv(r.A, c.gb);
    // Qc = S_{k+1} - A S_k A^T
    RgS = gQc;
    gA = gA - 2.0 * mm(gQc, mm(r.A, r.Sk));
    LgS = -1.0 * Tmm(r.A, mm(gQc, r.A));
    // A = (S_k^-1 C)^T
    const Mat<D> gC = chol_solve(r.LS, transpose(gA));
    LgS = LgS + sym(-1.0 * mm(gC, r.A));
    // C = eta_s^T - m_k m_{k+1}^T
    ges = transpose(gC);
    Lgm = Lgm - mv(gC, r.mn);
    Rgm = Rgm - Tmv(gC, r.mk);
}

// out = theta - gamma * dL/d eta when th* are given (natural-gradient step), else out = dL/d eta
template <int D>
__global__ void k_expectations_to_ssm_vjp(const double* __restrict__ e1, const double* __restrict__ ed,
                                          const double* __restrict__ es, const double* __restrict__ gA,
                                          const double* __restrict__ gb, const double* __restrict__ gcholP0,
                                          const double* __restrict__ gcholQ, const double* __restrict__ gmu0,
                                          const double* __restrict__ th1, const double* __restrict__ thd,
                                          const double* __restrict__ ths, double gamma, double* __restrict__ o1,
                                          double* __restrict__ od, double* __restrict__ os, int64_t B, int64_t N,
                                          int64_t* status) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t n = N + 1;
    if (i >= B * n) return;
    constexpr int DD = D * D;
    int64_t b, k;
    split_index(i, n, b, k);
    bool ok = true;
    Vec<D> gm = vec_zero<D>();
    Mat<D> gS = mat_zero<D>();
    const Vec<D> mk = vec_load<D>(e1 + i * D);
    const Mat<D> edk = mat_load<D>(ed + i * DD);
    if (k < N) {   // transition k: left part and eta_s
        const int64_t t = b * N + k;
        ExpToSsm<D> r;
        exp_to_ssm_step(mk, vec_load<D>(e1 + (i + 1) * D), edk, mat_load<D>(ed + (i + 1) * DD),
                        mat_load<D>(es + t * DD), r, ok);
        SsmCotangent<D> c;
        c.gA = mat_load<D>(gA + t * DD);
        c.gb = vec_load<D>(gb + t * D);
        c.gLQ = mat_load<D>(gcholQ + t * DD);
        Vec<D> Lgm, Rgm;
        Mat<D> LgS, RgS, ges;
        exp_to_ssm_step_vjp(r, c, Lgm, LgS, Rgm, RgS, ges);
        gm = gm + Lgm;
        gS = gS + LgS;
        if (k == 0) {   // mu0 = eta1_0, cholP0 = chol(S_0)
            gm = gm + vec_load<D>(gmu0 + b * D);
            gS = gS + chol_adjoint(r.LS, tril(mat_load<D>(gcholP0 + b * DD)));
        }
        if (ths) ges = mat_load<D>(ths + t * DD) - gamma * ges;
        mat_store<D>(os + t * DD, ges);
    }
    if (k > 0) {   // transition k-1: right part
        const int64_t t = b * N + k - 1;
        ExpToSsm<D> r;
        exp_to_ssm_step(vec_load<D>(e1 + (i - 1) * D), mk, mat_load<D>(ed + (i - 1) * DD), edk,
                        mat_load<D>(es + t * DD), r, ok);
        SsmCotangent<D> c;
        c.gA = mat_load<D>(gA + t * DD);
        c.gb = vec_load<D>(gb + t * D);
        c.gLQ = mat_load<D>(gcholQ + t * DD);
        Vec<D> Lgm, Rgm;
        Mat<D> LgS, RgS, ges;
        exp_to_ssm_step_vjp(r, c, Lgm, LgS, Rgm, RgS, ges);
        gm = gm + Rgm;
        gS = gS + RgS;
    }
    if (!ok) report_not_pd(status, i);
    // S = eta_d - m m^T
    gm = gm - 2.0 * mv(sym(gS), mk);
    if (th1) {
        gm = vec_load<D>(th1 + i * D) - gamma * gm;
        gS = mat_load<D>(thd + i * DD) - gamma * gS;
    }
    vec_store<D>(o1 + i * D, gm);
    mat_store<D>(od + i * DD, gS);
}

}  // namespace vdp
