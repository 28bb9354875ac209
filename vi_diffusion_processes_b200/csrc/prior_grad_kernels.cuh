// Prior-parameter learning: gradients of the ELBO terms with respect to the drift parameters.
//
// Reference (tf.GradientTape watching prior_sde.trainable_variables):
//   CVISitesSDE.grad_KL_wrt_prior_params     variational_cvi_sde.py:495-504   (through KL_q_p :446-486)
//   CVISitesSDE.grad_VE_wrt_prior_params     variational_cvi_sde.py:506-518   (through set_linearized_prior ->
//                                            dist_q.marginals -> variational_expectation)
//   VariationalMarkovGP.grad_prior_sde_params vi_sde.py:457-470               (through E_sde)
// Derivations: oracle/models_np.py (checked there against torch.autograd and central differences).
//
// Quadrature kernels use one group of G lanes per transition (G = 1 for D = 1, one warp for D >= 2), grid =
// (blocks per sequence, B); the per-parameter sums are reduced in a fixed order: lanes (shuffle butterfly) ->
// groups of the block (sequential in shared memory) -> blocks of the sequence (k_reduce_partials).
// The sensitivity sweeps are affine recursions over the chain; they run as three passes (chunk summaries,
// sequential carry over the chunk summaries of a sequence, chunk replay), one thread per chunk.
#pragma once
#include "pointwise_kernels.cuh"
#include "quad_core.cuh"

namespace vdp {

constexpr int kMaxParams = VDP_MAX_DRIFT_PARAMS;

// fixed-order reduction of per-group sums to one partial per (block, parameter)
template <int NP>
__device__ __forceinline__ void block_reduce_params(const double (&acc)[NP], bool active, int gidx, int groups,
                                                    int lane, double* sh, double* __restrict__ partials_block) {
    if (lane == 0) {
#pragma unroll
        for (int p = 0; p < NP; ++p) sh[gidx * NP + p] = active ? acc[p] : 0.0;
    }
    __syncthreads();
    if ((int)threadIdx.x < NP) {
        double s = 0.0;
        for (int g = 0; g < groups; ++g) s += sh[g * NP + threadIdx.x];
        partials_block[threadIdx.x] = s;
    }
}

// sum_t E_{N(m_t, S_t)}[ r(x)^T W df/dp(x) ],  r(x) = c_x x + c_f f(x) - (a_sign A_t x + b_t),  H nodes per dimension.
//   KL_q_p gradient (variational_cvi_sde.py:495-504): c_x = 1, c_f = dt, a_sign = 1, W = sym((dt q)^-1), scale dt
//   E_sde gradient  (vi_sde.py:457-470):              c_x = 0, c_f = 1,  a_sign = -1, W = sym(q^-1),     scale dt
// A [B,N,D,D], b [B,N,D]; m, S: the Gaussian of transition t sits at m + (b * pts + t) * D (pts grid points per
// sequence in the arrays, so a caller can pass marginals [B,N+1,..] with or without an offset of one point).
template <int DRIFT, int D, int H, int G>
__global__ void __launch_bounds__(256)
k_drift_param_grad(DriftParams P, Mat<D> W, double c_x, double c_f, double a_sign, double pi_factor,
                   const double* __restrict__ A, const double* __restrict__ bo, const double* __restrict__ m,
                   const double* __restrict__ S, int64_t pts, double* __restrict__ partials, int64_t N,
                   int64_t* status) {
    using DG = DriftParamGrad<DRIFT, D>;
    constexpr int NP = DG::NP;
    constexpr int DD = D * D;
    extern __shared__ double sh_pg[];
    const int groups = blockDim.x / G;
    const int gidx = threadIdx.x / G, lane = threadIdx.x % G;
    const int64_t b = blockIdx.y;
    const int64_t k = (int64_t)blockIdx.x * groups + gidx;
    const bool active = k < N;
    double acc[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) acc[p] = 0.0;
    if (active) {
        const int64_t t = b * N + k, node = b * pts + k;
        bool ok = true;
        const Mat<D> Ak = a_sign * mat_load<D>(A + t * DD);
        const Vec<D> bk = vec_load<D>(bo + t * D);
        const Vec<D> mk = vec_load<D>(m + node * D);
        const Mat<D> LS = chol(mat_load<D>(S + node * DD), ok);
        if (!ok && lane == 0) report_not_pd(status, node);
        constexpr int NODES = NodeCount<H, D>::value;
        for (int pnode = lane; pnode < NODES; pnode += G) {
            Vec<D> xi, f;
            Mat<D> J;
            double w;
            gh_node<H, D>(pnode, xi, w);
            const Vec<D> X = gh_point(LS, mk, xi);
            Drift<DRIFT, D>::eval(P, X, f, J);
            Vec<D> df[NP];
            Mat<D> dJ[NP];
            DG::eval(P, X, df, dJ);
            const Vec<D> AX = mv(Ak, X);
            Vec<D> r;
#pragma unroll
            for (int i = 0; i < D; ++i) r(i) = (c_x * X(i) + c_f * f(i)) - (AX(i) + bk(i));
            const Vec<D> Wr = w * mv(W, r);
#pragma unroll
            for (int p = 0; p < NP; ++p) acc[p] += dot(Wr, df[p]);
        }
#pragma unroll
        for (int p = 0; p < NP; ++p) acc[p] = pi_factor * group_sum<G>(acc[p]);
    }
    block_reduce_params<NP>(acc, active, gidx, groups, lane, sh_pg,
                            partials + (b * gridDim.x + blockIdx.x) * NP);
}

// grads[b, p] = scale * sum over the blocks of sequence b (fixed order); entries p >= np are set to 0
static __global__ void k_reduce_partials(const double* __restrict__ partials, int64_t nbx, int np, double scale,
                                  double* __restrict__ grads) {
    __shared__ double sh[256];
    const int64_t b = blockIdx.x;
    for (int p = 0; p < kMaxParams; ++p) {
        double acc = 0.0;
        if (p < np)
            for (int64_t j = threadIdx.x; j < nbx; j += blockDim.x) acc += partials[(b * nbx + j) * np + p];
        sh[threadIdx.x] = acc;
        __syncthreads();
        for (int s = blockDim.x / 2; s > 0; s >>= 1) {
            if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
            __syncthreads();
        }
        if (threadIdx.x == 0) grads[b * kMaxParams + p] = (p < np) ? scale * sh[0] : 0.0;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// grad_VE_wrt_prior_params: cotangents of L = -VE at the observed grid points, scattered onto the grid
//   gbar_j = dL/dm_j = -sym(R^-1)(y_j - m_j),   G_j = dL/dSigma_jj = 1/2 sym(R^-1)       (multivariate_gaussian.py:82-116)
// ---------------------------------------------------------------------------------------------------
template <int D>
__global__ void k_ve_cotangents(const double* __restrict__ fx_mus, const int64_t* __restrict__ obs_idx,
                                const double* __restrict__ obs_y, Mat<D> lik_chol, double* __restrict__ gbar,
                                double* __restrict__ Gbar, int64_t B, int64_t N, int64_t n_obs) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_obs) return;
    constexpr int DD = D * D;
    const int64_t b = i / n_obs;
    const int64_t k = obs_idx[i];
    if (k < 0 || k > N) return;
    const int64_t node = b * (N + 1) + k;
    const Mat<D> Rinv = sym(chol_inverse(lik_chol));
    const Vec<D> g = mv(Rinv, vec_load<D>(fx_mus + node * D) - vec_load<D>(obs_y + i * D));
#pragma unroll
    for (int e = 0; e < D; ++e) atomicAdd(gbar + node * D + e, g(e));
#pragma unroll
    for (int e = 0; e < DD; ++e) atomicAdd(Gbar + node * DD + e, 0.5 * Rinv.a[e]);
}

// ---- sensitivity sweeps: u = Sigma gbar and the band of V = Sigma Gbar Sigma for a Gauss-Markov chain with
// transitions A_k and marginal covariances Sigma_kk (Sigma[k, j] = A_{k-1} .. A_j Sigma_jj for k >= j):
//   forward   f_k = A_{k-1} f_{k-1} + Sigma_kk gbar_k,        F_k = A_{k-1} F_{k-1} A_{k-1}^T + Sigma_kk G_k Sigma_kk
//   backward  r_k = A_k^T (r_{k+1} + gbar_{k+1}),             R_k = A_k^T (R_{k+1} + G_{k+1}) A_k
//   u_k = f_k + Sigma_kk r_k,   V_kk = F_k + Sigma_kk R_k Sigma_kk,
//   V_{k+1,k} = A_k F_k + Sigma_{k+1,k+1} (R_{k+1} + G_{k+1}) A_k Sigma_kk
template <int D> struct SensElem { Mat<D> M; Vec<D> v; Mat<D> C; };   // x -> M x + v,  X -> M X M^T + C
template <int D> struct SensState { Vec<D> x; Mat<D> X; };
enum { SENS_FWD = 0, SENS_BWD = 1 };

// step of position j (0-based along the sweep direction) of sequence b: FWD node k = j, BWD node k = N-1-j
template <int D, int DIR>
__device__ __forceinline__ void sens_step_inputs(int64_t b, int64_t j, int64_t N, const double* __restrict__ qA,
                                                 const double* __restrict__ Sd, const double* __restrict__ gbar,
                                                 const double* __restrict__ Gbar, Mat<D>& Ms, Vec<D>& s, Mat<D>& T) {
    constexpr int DD = D * D;
    if (DIR == SENS_FWD) {
        const int64_t node = b * (N + 1) + j;
        Ms = (j > 0) ? mat_load<D>(qA + (b * N + j - 1) * DD) : mat_zero<D>();
        const Mat<D> Sk = mat_load<D>(Sd + node * DD);
        s = mv(Sk, vec_load<D>(gbar + node * D));
        T = mm(Sk, mm(mat_load<D>(Gbar + node * DD), Sk));
    } else {
        const int64_t k = N - 1 - j, node = b * (N + 1) + k + 1;
        const Mat<D> Ak = mat_load<D>(qA + (b * N + k) * DD);
        Ms = transpose(Ak);
        s = Tmv(Ak, vec_load<D>(gbar + node * D));
        T = Tmm(Ak, mm(mat_load<D>(Gbar + node * DD), Ak));
    }
}

template <int D, int DIR>
__global__ void k_sens_summary(const double* __restrict__ qA, const double* __restrict__ Sd,
                               const double* __restrict__ gbar, const double* __restrict__ Gbar,
                               SensElem<D>* __restrict__ elems, int64_t B, int64_t N, int64_t Lc, int64_t nc) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * nc) return;
    const int64_t b = i / nc, c = i % nc;
    const int64_t len = (DIR == SENS_FWD) ? N + 1 : N;
    const int64_t j0 = c * Lc, j1 = (j0 + Lc < len) ? j0 + Lc : len;
    SensElem<D> e;
    e.M = mat_eye<D>();
    e.v = vec_zero<D>();
    e.C = mat_zero<D>();
    for (int64_t j = j0; j < j1; ++j) {
        Mat<D> Ms, T;
        Vec<D> s;
        sens_step_inputs<D, DIR>(b, j, N, qA, Sd, gbar, Gbar, Ms, s, T);
        e.M = mm(Ms, e.M);
        e.v = mv(Ms, e.v) + s;
        e.C = mm(Ms, mmT(e.C, Ms)) + T;
    }
    elems[i] = e;
}

template <int D>
__global__ void k_sens_carry(const SensElem<D>* __restrict__ elems, SensState<D>* __restrict__ carry, int64_t B,
                             int64_t nc) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    SensState<D> st;
    st.x = vec_zero<D>();
    st.X = mat_zero<D>();
    for (int64_t c = 0; c < nc; ++c) {
        carry[b * nc + c] = st;
        const SensElem<D> e = elems[b * nc + c];
        st.x = mv(e.M, st.x) + e.v;
        st.X = mm(e.M, mmT(st.X, e.M)) + e.C;
    }
}

// FWD: u <- f, Vd <- F.   BWD (after FWD): u <- f + Sigma r, Vs <- V_{k+1,k}, Vd <- F + Sigma R Sigma.
template <int D, int DIR>
__global__ void k_sens_apply(const double* __restrict__ qA, const double* __restrict__ Sd,
                             const double* __restrict__ gbar, const double* __restrict__ Gbar,
                             const SensState<D>* __restrict__ carry, double* __restrict__ u, double* __restrict__ Vd,
                             double* __restrict__ Vs, int64_t B, int64_t N, int64_t Lc, int64_t nc) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * nc) return;
    constexpr int DD = D * D;
    const int64_t b = i / nc, c = i % nc;
    const int64_t len = (DIR == SENS_FWD) ? N + 1 : N;
    const int64_t j0 = c * Lc, j1 = (j0 + Lc < len) ? j0 + Lc : len;
    SensState<D> st = carry[i];
    for (int64_t j = j0; j < j1; ++j) {
        Mat<D> Ms, T;
        Vec<D> s;
        sens_step_inputs<D, DIR>(b, j, N, qA, Sd, gbar, Gbar, Ms, s, T);
        if (DIR == SENS_FWD) {
            st.x = mv(Ms, st.x) + s;
            st.X = mm(Ms, mmT(st.X, Ms)) + T;
            const int64_t node = b * (N + 1) + j;
            vec_store<D>(u + node * D, st.x);
            mat_store<D>(Vd + node * DD, st.X);
        } else {
            const int64_t k = N - 1 - j, node = b * (N + 1) + k;
            const Mat<D> Ak = transpose(Ms);
            const Mat<D> Sk = mat_load<D>(Sd + node * DD), Sn = mat_load<D>(Sd + (node + 1) * DD);
            const Mat<D> Fk = mat_load<D>(Vd + node * DD);
            const Mat<D> RG = st.X + mat_load<D>(Gbar + (node + 1) * DD);   // R_{k+1} + G_{k+1}
            mat_store<D>(Vs + (b * N + k) * DD, mm(Ak, Fk) + mm(Sn, mm(RG, mm(Ak, Sk))));
            st.x = mv(Ms, st.x) + s;
            st.X = mm(Ms, mmT(st.X, Ms)) + T;
            vec_store<D>(u + node * D, vec_load<D>(u + node * D) + mv(Sk, st.x));
            mat_store<D>(Vd + node * DD, Fk + mm(Sk, mm(st.X, Sk)));
        }
    }
}

// Contribution of transition k to d(-VE)/dp:  with dA = dA_k/dp, db = db_k/dp of the (clipped) linearised prior
// (sde_utils.py:156-166, drift.py:102-108, variational_cvi_sde.py:417-430), W = Q_k^-1, Y = W dA, z = W db,
//   c = (z + Y m_k) . (u_{k+1} - A u_k) - (Y u_k) . (b + A m_k - m_{k+1}) + 2 <V_{k+1,k} - A V_kk, Y>_F
// (A, b: the prior's transition; m, u, V: posterior means and the sensitivities above).
template <int DRIFT, int D, int H, int G>
__global__ void __launch_bounds__(256)
k_ve_param_grad(DriftParams P, double pi_factor, const double* __restrict__ time_grid,
                const double* __restrict__ path_m, const double* __restrict__ path_S, int clip, double lo, double hi,
                const double* __restrict__ pA, const double* __restrict__ pb, const double* __restrict__ pcholQ,
                const double* __restrict__ m, const double* __restrict__ u, const double* __restrict__ Vd,
                const double* __restrict__ Vs, double* __restrict__ partials, int64_t N, int64_t* status) {
    using DG = DriftParamGrad<DRIFT, D>;
    constexpr int NP = DG::NP;
    constexpr int DD = D * D;
    extern __shared__ double sh_pg[];
    const int groups = blockDim.x / G;
    const int gidx = threadIdx.x / G, lane = threadIdx.x % G;
    const int64_t b = blockIdx.y;
    const int64_t k = (int64_t)blockIdx.x * groups + gidx;
    const bool active = k < N;
    double acc[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) acc[p] = 0.0;
    if (active) {
        const int64_t t = b * N + k, node = b * (N + 1) + k;
        bool ok = true;
        // linearisation point = grid point k+1 of the stored posterior path (variational_cvi_sde.py:412)
        const Vec<D> mp = vec_load<D>(path_m + (node + 1) * D);
        const Mat<D> LS = chol(mat_load<D>(path_S + (node + 1) * DD), ok);
        if (!ok && lane == 0) report_not_pd(status, node + 1);
        Vec<D> Ef = vec_zero<D>(), Edf[NP];
        Mat<D> EJ = mat_zero<D>(), EdJ[NP];
#pragma unroll
        for (int p = 0; p < NP; ++p) { Edf[p] = vec_zero<D>(); EdJ[p] = mat_zero<D>(); }
        constexpr int NODES = NodeCount<H, D>::value;
        for (int pnode = lane; pnode < NODES; pnode += G) {
            Vec<D> xi, f, df[NP];
            Mat<D> J, dJ[NP];
            double w;
            gh_node<H, D>(pnode, xi, w);
            w *= pi_factor;
            const Vec<D> X = gh_point(LS, mp, xi);
            Drift<DRIFT, D>::eval(P, X, f, J);
            DG::eval(P, X, df, dJ);
#pragma unroll
            for (int i = 0; i < D; ++i) Ef(i) += f(i) * w;
#pragma unroll
            for (int i = 0; i < DD; ++i) EJ.a[i] += J.a[i] * w;
#pragma unroll
            for (int p = 0; p < NP; ++p) {
#pragma unroll
                for (int i = 0; i < D; ++i) Edf[p](i) += df[p](i) * w;
#pragma unroll
                for (int i = 0; i < DD; ++i) EdJ[p].a[i] += dJ[p].a[i] * w;
            }
        }
        group_sum<G>(Ef);
        group_sum<G>(EJ);
#pragma unroll
        for (int p = 0; p < NP; ++p) { group_sum<G>(Edf[p]); group_sum<G>(EdJ[p]); }
        if (lane == 0) {
            const double delta = time_grid[k + 1] - time_grid[k];
            // un-clipped linearisation decides where tf.clip_by_value passes the gradient (lo <= x <= hi)
            const Mat<D> A_lin = delta * EJ + mat_eye<D>();
            const Vec<D> b_lin = delta * (Ef - mv(EJ, mp));
            const Mat<D> Ak = mat_load<D>(pA + t * DD);
            const Vec<D> bk = vec_load<D>(pb + t * D);
            const Mat<D> W = chol_inverse(mat_load<D>(pcholQ + t * DD));
            const Vec<D> mk = vec_load<D>(m + node * D), mn = vec_load<D>(m + (node + 1) * D);
            const Vec<D> uk = vec_load<D>(u + node * D), un = vec_load<D>(u + (node + 1) * D);
            const Vec<D> e = un - mv(Ak, uk);
            const Vec<D> rho = (bk + mv(Ak, mk)) - mn;
            const Mat<D> Xi = mat_load<D>(Vs + t * DD) - mm(Ak, mat_load<D>(Vd + node * DD));
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                Mat<D> dA = delta * EdJ[p];
                Vec<D> db = delta * (Edf[p] - mv(EdJ[p], mp));
                if (clip) {
#pragma unroll
                    for (int i = 0; i < DD; ++i)
                        if (!(A_lin.a[i] >= lo && A_lin.a[i] <= hi)) dA.a[i] = 0.0;
#pragma unroll
                    for (int i = 0; i < D; ++i)
                        if (!(b_lin(i) >= lo && b_lin(i) <= hi)) db(i) = 0.0;
                }
                const Mat<D> Y = mm(W, dA);
                const Vec<D> z = mv(W, db);
                double fro = 0.0;
#pragma unroll
                for (int i = 0; i < DD; ++i) fro += Xi.a[i] * Y.a[i];
                acc[p] = dot(z + mv(Y, mk), e) - dot(mv(Y, uk), rho) + 2.0 * fro;
            }
        }
    }
    block_reduce_params<NP>(acc, active, gidx, groups, lane, sh_pg,
                            partials + (b * gridDim.x + blockIdx.x) * NP);
}

}  // namespace vdp
