/*
 * vdp_b200 -- C ABI of the B200 (sm_100a, float64) implementation of markovflow's VDP / CVI-DP
 * site-update hot path.  Reference: AaltoML/vi-diffusion-processes (paths below are relative to it).
 *
 * The reference has no FFI of its own for this path; its only native seam is the set of TensorFlow
 * custom ops it imports from banded_matrices (markovflow/block_tri_diag.py:22-31,
 * markovflow/ssm_gaussian_transformations.py:23).  The entry points below are what a binding for
 * this path replaces, op level first (one per banded op, block layout), then the fused level
 * (one per Python function on the path).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; the caller owns every buffer,
 *     the library never allocates outputs;
 *   - float64, row-major, reference tensor layout with a leading batch axis:
 *       vectors   [B, n, D]      block diagonals [B, n, D, D]      block sub-diagonals [B, n-1, D, D]
 *     (n = N + 1 grid points, N transitions, D = state dimension, 1..3);
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it;
 *   - return value: 0 = launched, negative = VDP_E_* (nothing launched);
 *   - `status` (optional, may be NULL) is a device int64[2]: {code, flat grid index b*n+k of the first
 *     block whose Cholesky pivot was not positive / not finite}; code 0 = clean.  Read it with
 *     vdp_read_status (synchronises the stream).
 */
#ifndef VDP_B200_H
#define VDP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VDP_OK 0
#define VDP_E_BADARG (-1)   /* unsupported D, drift/D combination, null pointer, negative size */
#define VDP_E_CUDA (-2)     /* a CUDA runtime call failed; see vdp_last_error() */
#define VDP_E_WORKSPACE (-3) /* workspace too small */
#define VDP_STATUS_NOT_PD 1

/* drift plugin ids: markovflow/sde/sde.py classes; 7 and 8 are extensions (no 3-D SDE in the reference) */
enum {
    VDP_DRIFT_OU = 0,       /* params: decay                      sde.py:134-176 */
    VDP_DRIFT_DW = 1,       /* params: scale, c                   sde.py:179-224 */
    VDP_DRIFT_BENES = 2,    /* params: theta                      sde.py:227-268 */
    VDP_DRIFT_SINE = 3,     /* params: theta                      sde.py:271-312 */
    VDP_DRIFT_SQRT = 4,     /* params: theta                      sde.py:315-356 */
    VDP_DRIFT_VDP = 5,      /* params: a, tau      (D = 2)        sde.py:433-518 */
    VDP_DRIFT_MLP131 = 6,   /* params: w1[3], b1[3], w2[3], b2    sde.py:359-430 */
    VDP_DRIFT_LORENZ63 = 7, /* params: sigma, rho, beta (D = 3)   extension */
    VDP_DRIFT_OU_ND = 8     /* params: decay, full Jacobian       extension */
};

int vdp_version(void);
/* Process-wide tuning switches: "SWEEPS" (0 heuristic, 1 thread per sequence, 2 parallel-in-time tiles), "NSEG" (time
 * segments per sequence of the tiled sweeps, 0 = heuristic), and the route switches "GH_QUADRATURE",
 * "REFACTOR_PRECISION", "NO_CARRIED_NATURALS", "NO_FUSED_NATURALS", "SEPARATE_KL_VALUE", "NO_FUSED_FORWARD" (0 / 1; the per-call
 * VDP_FLAG_* of the same name are the re-entrant way to choose a route).  Their defaults are read from the
 * environment variables VDP_<KEY> once, when the library is loaded; no entry point reads the environment. */
int vdp_config_set(const char* key, int value);
int vdp_config_get(const char* key, int* value);
const char* vdp_last_error(void);
/* copies status[2] to the host after synchronising `stream`; returns status[0] */
int64_t vdp_read_status(const int64_t* status, int64_t* out_host2, void* stream);

/* ------------------------------------------------------------------------------------------------
 * op level -- banded_matrices replacements in block layout
 * ---------------------------------------------------------------------------------------------- */

/* cholesky_band  (block_tri_diag.py:440  SymmetricBlockTriDiagonal.cholesky)
 * Kd [B,n,D,D] symmetric, Ks [B,n-1,D,D] = K_{k+1,k} (NULL: block diagonal only) -> Ld, Ls */
int vdp_btd_cholesky(const double* Kd, const double* Ks, double* Ld, double* Ls,
                     int64_t B, int64_t n, int D, int64_t* status, void* stream);

/* inverse_from_cholesky_band  (block_tri_diag.py:318-337 block_diagonal_of_inverse;
 * ssm_gaussian_transformations.py:443-445): Sd = blocks Sigma_kk, Ss = Sigma_{k+1,k} (may be NULL) */
int vdp_btd_inverse_band(const double* Ld, const double* Ls, double* Sd, double* Ss,
                         int64_t B, int64_t n, int D, void* stream);

/* Directional derivative of the inverse band: for K = L L^T and a symmetric block-tridiagonal direction
 * dK = (dKd [B,n,D,D] read through its lower triangle, dKs [B,n-1,D,D] = dK_{k+1,k}) the blocks dSd = d(Sigma_kk),
 * dSs = d(Sigma_{k+1,k}) of d(K^-1) = -K^-1 dK K^-1 (forward-mode differentiation of cholesky_band and
 * inverse_from_cholesky_band, block_tri_diag.py:428-440, :318-337).  n >= 2.  Serves the Fisher-norm term of
 * SSMNaturalGradient(momentum=True): tape.gradient(naturals_to_ssm_params(theta), theta, output_gradients=dL_dssm)
 * contracted with dL/d eta (ssm_natgrad.py:163-190) is a directional derivative of theta -> eta. */
int vdp_btd_inverse_band_jvp(const double* Ld, const double* Ls, const double* dKd, const double* dKs,
                             double* dSd, double* dSs, int64_t B, int64_t n, int D, void* stream);

/* solve_triang_mat  (block_tri_diag.py:339-351 LowerTriangularBlockTriDiagonal.solve)
 * x = L^-1 rhs (transpose = 0) or L^-T rhs (transpose = 1); rhs, x: [B,n,D]; Ls may be NULL */
int vdp_btd_solve(const double* Ld, const double* Ls, const double* rhs, double* x, int transpose,
                  int64_t B, int64_t n, int D, void* stream);

/* product_band_mat  (block_tri_diag.py:175-199 dense_mult): y = M x, M symmetric block-tridiagonal
 * (symmetric = 1) or lower block-bidiagonal (symmetric = 0, optional transpose) */
int vdp_btd_matvec(const double* Md, const double* Ms, const double* x, double* y, int symmetric,
                   int transpose, int64_t B, int64_t n, int D, void* stream);

/* abs_log_det  (block_tri_diag.py:353-366): out[b] = sum_k sum_i log |Ld[b,k,i,i]| */
int vdp_btd_logdet(const double* Ld, double* out, int64_t B, int64_t n, int D, void* stream);

/* Band storage of banded_matrices for callers that hold `.as_band` tensors (block_tri_diag.py:206-237,
 * block_to_band / band_to_block): band [B, R, n*D] row-major with band[r, j] = M[j + r, j] (row 0 = main
 * diagonal), R = 2 D with a sub-diagonal (Ms != NULL), else D.  band_to_block mirrors the upper triangles of
 * the diagonal blocks when symmetric != 0 and zeroes them otherwise. */
int vdp_btd_block_to_band(const double* Md, const double* Ms, double* band, int64_t B, int64_t n, int D,
                          void* stream);
int vdp_btd_band_to_block(const double* band, double* Md, double* Ms, int symmetric, int64_t B, int64_t n, int D,
                          void* stream);
/* The triangular band solve of naturals_to_ssm_params (ssm_gaussian_transformations.py:470-483):
 * block diagonal of A_inv^-T P, A_inv = LowerTriangularBlockTriDiagonal(I, -As), P = (Kd, Ks) symmetric:
 * out_k = Kd_k + As_k^T Ks_k (k < n-1), out_{n-1} = Kd_{n-1}.   As, Ks [B,n-1,D,D]; Kd, out [B,n,D,D]. */
int vdp_btd_conditional_precisions(const double* As, const double* Kd, const double* Ks, double* out, int64_t B,
                                   int64_t n, int D, void* stream);

/* ------------------------------------------------------------------------------------------------
 * fused level
 * ---------------------------------------------------------------------------------------------- */

/* SDE.expected_drift / expected_gradient_drift  (sde.py:92-131, 501-518; gpflow mvnquad, H nodes/dim)
 * m [count,D], S [count,D,D] -> Ef [count,D]; EJ [count,D,D] when the drift has a full Jacobian
 * (VDP, LORENZ63, OU_ND), else [count,D] = E[d(sum_i f_i)/dx]  (either output may be NULL) */
int vdp_gh_expect(int drift_id, const double* params_host, const double* m, const double* S, int H,
                  double* Ef, double* EJ, int64_t count, int D, int64_t* status, void* stream);

/* linearize_sde + LinearDrift.to_ssm  (sde_utils.py:119-179, drift.py:64-116) with the optional
 * clip of variational_cvi_sde.py:417-430.  m,S are the linearisation path on grid points 1..N
 * ([B,N,D],[B,N,D,D]); time_grid [N+1]; q_host [D,D]; outputs A [B,N,D,D], b [B,N,D], cholQ [B,N,D,D].
 * clip = 1: A and b are clipped to [clip_lo, clip_hi] after the un-clipped values have been written to
 * A_lin, b_lin (which may be NULL). */
int vdp_linearize_to_ssm(int drift_id, const double* params_host, const double* q_host,
                         const double* time_grid, const double* m, const double* S,
                         double* A, double* b, double* cholQ, double* A_lin, double* b_lin,
                         int clip, double clip_lo, double clip_hi,
                         int64_t B, int64_t N, int D, int64_t* status, void* stream);

/* ssm_to_naturals  (ssm_gaussian_transformations.py:182-253):
 * A [B,N,D,D], b [B,N,D], cholQ [B,N,D,D], mu0 [B,D], cholP0 [B,D,D] -> th1 [B,N+1,D], thd, ths */
int vdp_ssm_to_naturals(const double* A, const double* b, const double* cholQ, const double* mu0,
                        const double* cholP0, double* th1, double* thd, double* ths,
                        int64_t B, int64_t N, int D, void* stream);

/* naturals_to_ssm_params  (ssm_gaussian_transformations.py:333-511).
 * workspace: vdp_naturals_to_ssm_workspace(B,N,D) bytes */
size_t vdp_naturals_to_ssm_workspace(int64_t B, int64_t N, int D);
int vdp_naturals_to_ssm(const double* th1, const double* thd, const double* ths,
                        double* A, double* b, double* cholQ, double* mu0, double* cholP0,
                        void* workspace, size_t workspace_bytes,
                        int64_t B, int64_t N, int D, int64_t* status, void* stream);

/* StateSpaceModel.marginals  (state_space_model.py:231-262, 431-483; gauss_markov.py:107-117)
 * -> m [B,N+1,D], Sd [B,N+1,D,D].  workspace: vdp_ssm_marginals_workspace(B,N,D) bytes */
size_t vdp_ssm_marginals_workspace(int64_t B, int64_t N, int D);
int vdp_ssm_marginals(const double* A, const double* b, const double* cholQ, const double* mu0,
                      const double* cholP0, double* m, double* Sd,
                      void* workspace, size_t workspace_bytes,
                      int64_t B, int64_t N, int D, int64_t* status, void* stream);

/* StateSpaceModel.kl_divergence  (state_space_model.py:528-593): kl[b] = KL(dist1 || dist2) of two state-space
 * models with the same shapes, by the reference's route: marginals of dist1, precision of dist2 and its banded
 * Cholesky factor, trace over the block tri-diagonal band, log-determinants from the models' Cholesky factors.
 * workspace: vdp_ssm_kl_divergence_workspace(B,N,D) bytes */
size_t vdp_ssm_kl_divergence_workspace(int64_t B, int64_t N, int D);
int vdp_ssm_kl_divergence(const double* A1, const double* b1, const double* cholQ1, const double* mu01,
                          const double* cholP01, const double* A2, const double* b2, const double* cholQ2,
                          const double* mu02, const double* cholP02, double* kl,
                          void* workspace, size_t workspace_bytes,
                          int64_t B, int64_t N, int D, int64_t* status, void* stream);

/* StateSpaceModel.log_pdf  (state_space_model.py:485-526): states [S,B,N+1,D] (S sample trajectories per
 * sequence) -> out [S,B] = log p(x_0) + sum_k log p(x_{k+1} | x_k).  workspace: vdp_ssm_log_pdf_workspace bytes */
size_t vdp_ssm_log_pdf_workspace(int64_t S, int64_t B, int64_t N);
int vdp_ssm_log_pdf(const double* A, const double* b, const double* cholQ, const double* mu0, const double* cholP0,
                    const double* states, double* out, void* workspace, size_t workspace_bytes,
                    int64_t S, int64_t B, int64_t N, int D, void* stream);

/* StateSpaceModel.sample  (state_space_model.py:298-324): out [S,B,N+1,D] = A_inv_block^-1 (offsets + chol eps),
 * i.e. x_0 = mu0 + cholP0 eps_0, x_{k+1} = A_k x_k + b_k + cholQ_k eps_{k+1}.  eps: standard normals [S,B,N+1,D]
 * used in place of the generator (NULL = Philox4x32-10 with `seed`, sub-sequence = s B + b). */
int vdp_ssm_sample(const double* A, const double* b, const double* cholQ, const double* mu0, const double* cholP0,
                   const double* eps, unsigned long long seed, double* out,
                   int64_t S, int64_t B, int64_t N, int D, void* stream);

/* Allocation-free variant of vdp_ssm_marginals.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs, outputs and results as vdp_ssm_marginals, but the scratch (the Cholesky
 * factor of the rebuilt precision and the segment carries) is a file-scope `static __device__` array of
 * VDP_STATIC_SCRATCH_BYTES (16 MiB) inside the library; every GPU that loads the library holds it until exit.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   Limit: 8 (2 N + 1) B D^2 <= VDP_STATIC_FACTOR_BYTES (8 MiB; each of the two factor arrays rounded up to
 *   256 B), e.g. B (N + 1) <= 524 288 grid points at D = 1.  Larger problems return VDP_E_WORKSPACE
 *   (vdp_last_error() names the limit); use vdp_ssm_marginals with a caller-owned workspace for those.
 *   All calls on one GPU share the array: calls that overlap on the GPU (different streams, or several host
 *   threads) RACE on it -- serialise them.  The array keeps the last call's contents; nothing in it has to start
 *   at zero (every word a kernel reads is written earlier in the same call). */
#define VDP_STATIC_FACTOR_BYTES ((size_t)8 << 20)
#define VDP_STATIC_SCRATCH_BYTES ((size_t)16 << 20)
size_t vdp_static_scratch_bytes(void);
int vdp_ssm_marginals_static(const double* A, const double* b, const double* cholQ, const double* mu0,
                             const double* cholP0, double* m, double* Sd,
                             int64_t B, int64_t N, int D, int64_t* status, void* stream);

/* ssm_to_expectations / expectations_to_ssm_params  (ssm_gaussian_transformations.py:32-178) */
int vdp_ssm_to_expectations(const double* A, const double* m, const double* Sd,
                            double* e1, double* ed, double* es, int64_t B, int64_t N, int D, void* stream);
int vdp_expectations_to_ssm(const double* e1, const double* ed, const double* es,
                            double* A, double* b, double* cholQ, double* mu0, double* cholP0,
                            int64_t B, int64_t N, int D, int64_t* status, void* stream);

/* Vector-Jacobian product of expectations_to_ssm_params at (e1, ed, es): the chain-rule step of
 * SSMNaturalGradient (ssm_natgrad.py:142-172, tape.gradient(ssm_params, etas, output_gradients=dL_dssm)).
 * Cotangents gA [B,N,D,D], gb [B,N,D], gcholP0 [B,D,D], gcholQ [B,N,D,D] (lower triangles), gmu0 [B,D]. */
int vdp_expectations_to_ssm_vjp(const double* e1, const double* ed, const double* es, const double* gA,
                                const double* gb, const double* gcholP0, const double* gcholQ, const double* gmu0,
                                double* g1, double* gd, double* gs, int64_t B, int64_t N, int D, int64_t* status,
                                void* stream);

/* SSMNaturalGradient._natgrad_step, momentum=False (ssm_natgrad.py:121-218): the SSM parameters are updated
 * in place from the Euclidean gradients dL/d(A, b, chol_P0, chol_Q, mu0). */
size_t vdp_ssm_natgrad_workspace(int64_t B, int64_t N, int D);
int vdp_ssm_natgrad_step(double* A, double* b, double* cholP0, double* cholQ, double* mu0, const double* dL_dA,
                         const double* dL_db, const double* dL_dcholP0, const double* dL_dcholQ,
                         const double* dL_dmu0, double gamma, void* workspace, size_t workspace_bytes, int64_t B,
                         int64_t N, int D, int64_t* status, void* stream);

/* SDE_SSM_KL_with_grads_wrt_exp_params  (sde_utils.py:473-547), H = 20:
 * q given by its transitions A [B,N,D,D] and marginals m, Sd; prior x0 ~ N(x0_mu_host, x0_cov_host);
 * dt = float(grid[1]-grid[0]) (used in f_p), time_grid gives the per-step Q_p = delta_k q.
 * -> kl [B], g1 [B,N+1,D], gd [B,N+1,D,D], gs [B,N,D,D].  workspace: vdp_cvi_kl_workspace bytes */
int vdp_cvi_kl_and_grad(int drift_id, const double* params_host, const double* q_host, double dt,
                        const double* time_grid, const double* x0_mu_host, const double* x0_cov_host,
                        const double* A, const double* m, const double* Sd,
                        double* kl, double* g1, double* gd, double* gs,
                        void* workspace, size_t workspace_bytes,
                        int64_t B, int64_t N, int D, int64_t* status, void* stream);
size_t vdp_cvi_kl_workspace(int64_t B, int64_t N, int D);

/* One CVI-DP site update = cvi_dp_trainer.py:71-77 once:
 *   CVISitesSSM.update_data_sites      variational_cvi_sde.py:301-317
 *   CVISitesSSM.update_girsanov_sites  variational_cvi_sde.py:279-299  (+ CVISitesSDE.grad_kl_wrt_exp_param :488-493)
 *   CVISitesSSM.classic_elbo           variational_cvi_sde.py:339-352  (+ CVISitesSDE.KL_q_p :446-486)
 * In/out: girsanov sites g1,gd,gs; data sites d1,d2; fx_mus,fx_covs (posterior marginals).
 * Out: q as an SSM, elbo/ve/kl [B]. */
typedef struct vdp_cvi_problem {
    int64_t B, N, n_obs;
    int32_t D, drift_id;
    double drift_params[16];
    double q[9];               /* spectral density, row-major D x D */
    double dt;                 /* float(time_grid[1] - time_grid[0]) */
    double x0_mu[3], x0_cov[9];/* prior on the initial state (Gaussian) */
    double lik_chol[9];        /* Cholesky factor of the Gaussian likelihood covariance */
    double lr_data, lr_girsanov;
    const double* time_grid;   /* [N+1] */
    const double *pA, *pb, *pcholQ, *pmu0, *pcholP0; /* linearised prior p: [B,N,D,D] [B,N,D] [B,N,D,D] [B,D] [B,D,D] */
    double *g1, *gd, *gs;      /* girsanov sites   [B,N+1,D] [B,N+1,D,D] [B,N,D,D] */
    double *d1, *d2;           /* data sites       [B,n_obs,D] [B,n_obs,D,D] */
    const int64_t* obs_idx;    /* [B,n_obs] grid index of each observation */
    const double* obs_y;       /* [B,n_obs,D] */
    double *fx_mus, *fx_covs;  /* [B,N+1,D] [B,N+1,D,D] */
    double *qA, *qb, *qcholQ, *qmu0, *qcholP0;       /* posterior q as an SSM (out) */
    double *elbo, *ve, *kl;    /* [B] (out) */
    void* workspace;
    size_t workspace_bytes;    /* >= vdp_cvi_workspace(B,N,D,n_obs) */
    int64_t* status;           /* device int64[2] or NULL */
    int32_t flags;             /* VDP_FLAG_* */
    int32_t reserved;
    /* optional per-sequence step sizes in DEVICE memory (all NULL: lr_data / lr_girsanov above apply to every
     * sequence): sequence b is updated with lr_data_dev[b] / lr_girsanov_dev[b], or left as it is where
     * active_dev[b] == 0.  This is what lets a trainer loop decay its step sizes and stop sequences without the host
     * (vdp_cvi_optimize_sites). */
    const double *lr_data_dev, *lr_girsanov_dev;   /* [B] */
    const int32_t* active_dev;                     /* [B] */
} vdp_cvi_problem;

#define VDP_FLAG_DEFAULT 0
#define VDP_FLAG_SKIP_DATA_UPDATE 1   /* leave the data sites as they are */
#define VDP_FLAG_SKIP_ELBO 2          /* do not evaluate classic_elbo */
#define VDP_FLAG_SKIP_GIRSANOV 4      /* update_data_sites only (variational_cvi_sde.py:301-317) */
#define VDP_FLAG_POSTERIOR_SSM_ONLY 8 /* vdp_cvi_elbo: only rebuild q as an SSM from the sites (dist_q, :176-192) */
#define VDP_FLAG_NATURALS_CARRIED 16  /* vdp_cvi_site_update: the caller's promise that (1) the previous call that used
                                         this workspace was a whole-chain vdp_cvi_site_update of this problem that ran
                                         the girsanov update (no VDP_FLAG_SKIP_GIRSANOV) or itself carried, and (2)
                                         prior, girsanov sites, data sites and observation indices are as that call
                                         left them.  The natural parameters and the dense copy of the data sites it
                                         left in the workspace are then rewritten at the observed grid points only
                                         instead of over the whole grid (same resulting state; honoured for
                                         n_obs <= 128, ignored otherwise) */

#define VDP_FLAG_SITES_IN_NATURALS 32  /* vdp_cvi_site_update, honoured only together with VDP_FLAG_NATURALS_CARRIED: the
                                         girsanov sites are held implicitly in the natural parameters theta = nat(p)
                                         + g + scatter(data sites) the workspace carries.  The girsanov update
                                         (variational_cvi_sde.py:290-296) then moves theta by lr (scatter - dKL/deta)
                                         and g1 / gd / gs are NEITHER READ NOR WRITTEN by the call (no pass over the
                                         sites and the prior: 88 instead of 136 bytes per grid step, D = 1).  Call
                                         vdp_cvi_materialize_sites before reading g1 / gd / gs, and before any call
                                         on this problem that does not carry the naturals.  A re-basing of the sites
                                         onto a new linearised prior (tranform_girsanov_sites, sde_utils.py:550-568:
                                         g + nat(p_old) - nat(p_new)) leaves theta unchanged. */

/* Route switches (same results up to rounding; each held to the default route by a test): */
#define VDP_FLAG_GH_QUADRATURE 64       /* polynomial scalar drifts: run the 20-node Gauss-Hermite loops of the reference
                                           instead of their closed form (Gaussian moments) */
#define VDP_FLAG_REFACTOR_PRECISION 128 /* marginals through a second factorisation of the rebuilt precision, as the
                                           reference does (four sweeps per rebuild instead of three) */
#define VDP_FLAG_NO_FUSED_NATURALS 256  /* rebuild the naturals from the sites after the girsanov update */
#define VDP_FLAG_SEPARATE_KL_VALUE 512  /* KL terms of classic_elbo in their own pass, not inside the mean sweep */
#define VDP_FLAG_NO_FUSED_FORWARD 1024  /* mean sweep, KL gradient / site update and the factorisation of the next
                                          * rebuild as three kernels instead of the one forward pass (D = 1) */

size_t vdp_cvi_workspace(int64_t B, int64_t N, int D, int64_t n_obs);
int vdp_cvi_site_update(const vdp_cvi_problem* problem_host, void* stream);
/* g1, gd, gs <- theta - nat(p) - scatter(data sites) from the natural parameters and the dense data-site copy the
 * last vdp_cvi_site_update left in the workspace (see VDP_FLAG_SITES_IN_NATURALS; full_sites of
 * variational_cvi_sde.py:161-174 read backwards).  p = the prior of the problem. */
int vdp_cvi_materialize_sites(const vdp_cvi_problem* problem_host, void* stream);
/* Sharded quadrature (one long sequence over several GPUs, BASELINE config 5, where the H^D-node quadrature is
 * > 95 % of the update): every rank holds the whole chain and runs the cheap rebuilds on it, the quadrature and the
 * girsanov site update only on its own range [node_begin, node_end) of the flat (sequence, grid point) index.
 *   phase 0: update_data_sites, rebuild, KL gradient + site update of the range   -> g1, gd, gs (range only)
 *            ... the caller exchanges the updated site ranges between ranks (broadcast / all-gather) ...
 *   phase 1: rebuild, variational expectation (whole chain), KL value of the range -> ve [B], kl [B] = the
 *            range's partial sum (KL_0 with the range holding grid point 0), elbo = ve - partial kl
 *            ... the caller sums the partial kl over ranks: ELBO = ve - sum_r kl_r ...
 * With the full range [0, B (N+1)) the two phases together equal vdp_cvi_site_update. */
int vdp_cvi_site_update_sharded(const vdp_cvi_problem* problem_host, int phase, int64_t node_begin,
                                int64_t node_end, void* stream);
/* The steps of a site update one at a time, WITHOUT the posterior rebuild between them -- for callers that rebuild
 * the posterior themselves, e.g. on a time-sharded chain (vdp_dist_sweep_*, BASELINE config 5) where `problem` is a
 * rank's local stretch of the chain extended by one halo grid point at either end:
 *   vdp_cvi_update_data_sites            update_data_sites, variational_cvi_sde.py:301-314 (the gradient step only;
 *                                        reads fx_mus at the observed grid points, writes d1, d2)
 *   vdp_cvi_full_sites                   full_sites, :161-174: theta = nat(p) + g + scatter(data sites)
 *                                        -> th1 [B,N+1,D], thd [B,N+1,D,D], ths [B,N,D,D]
 *   vdp_cvi_update_girsanov_sites_range  update_girsanov_sites, :279-296, on the flat node range [begin, end): reads
 *                                        qA, fx_mus, fx_covs (valid on the range and one grid point either side),
 *                                        updates g1, gd on the range and gs on its transitions; KL_0 enters at grid
 *                                        point 0 of the problem if the range holds it
 *   vdp_cvi_elbo_terms_range             the terms of classic_elbo, :339-352, from q*, fx_mus, fx_covs as they are:
 *                                        ve [B] over the problem's observations, kl [B] = the range's partial sum,
 *                                        elbo = ve - kl */
int vdp_cvi_update_data_sites(const vdp_cvi_problem* problem_host, void* stream);
int vdp_cvi_full_sites(const vdp_cvi_problem* problem_host, double* th1, double* thd, double* ths, void* stream);
int vdp_cvi_update_girsanov_sites_range(const vdp_cvi_problem* problem_host, int64_t node_begin, int64_t node_end,
                                        void* stream);
int vdp_cvi_elbo_terms_range(const vdp_cvi_problem* problem_host, int64_t node_begin, int64_t node_end, void* stream);

/* CVISitesTrainer._optimize_sites_under_stable_prior (docs/diffusion_processes/cvi_dp_trainer.py:63-101) ON THE DEVICE:
 * up to max_itr site updates (cvi_dp_trainer.py:71-77 each), the ELBO-decrease => step sizes / 10 rule (:84-87) and
 * the |ELBO change| < optim_tol stop (:89-91) evaluated per sequence by a kernel after each update, the loop itself a
 * CUDA-graph WHILE node whose condition that kernel sets -- no host synchronisation between updates.  The call
 * enqueues the whole loop on `stream` and returns; the host reads n_itr / elbo_trace / test_* once afterwards.
 *   in:     elbo_trace row 0 = the ELBO before the loop (classic_elbo); lr_data, lr_girsanov [B] current step sizes
 *   in/out: the problem's state as vdp_cvi_site_update; lr_data, lr_girsanov (decayed values persist, as the
 *           reference keeps them on the trainer object)
 *   out:    n_itr [B] updates a sequence took, elbo_trace rows 1..n_itr (later rows repeat the last value), and the
 *           marginals at the test grid points after every update (for NLPD / RMSE, exp_dp_utils.py:189-224)
 * The first update runs with problem->flags, the following ones with later_flags (typically | NATURALS_CARRIED
 * | SITES_IN_NATURALS).  A sequence that has stopped keeps its state (its step sizes are taken as 0). */
typedef struct vdp_cvi_loop {
    int32_t max_itr, n_test;
    int32_t later_flags, reserved;
    double optim_tol;
    double *lr_data, *lr_girsanov;   /* [B] */
    int32_t *active, *n_itr;         /* [B] (active: scratch) */
    double* elbo_trace;              /* [max_itr + 1, B] */
    const int64_t* test_idx;         /* [B, n_test] or NULL */
    double *test_mus, *test_covs;    /* [max_itr, B, n_test, D], [max_itr, B, n_test, D, D] or NULL */
    int32_t* counters;               /* device int32[2] scratch: {updates done, sequences still active} */
} vdp_cvi_loop;
int vdp_cvi_optimize_sites(const vdp_cvi_problem* problem_host, const vdp_cvi_loop* loop_host, void* stream);

/* classic_elbo only (no site is modified); fills elbo/ve/kl, q* and fx_* */
int vdp_cvi_elbo(const vdp_cvi_problem* problem_host, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Prior-parameter learning (cvi_dp_trainer.py:207-250, vi_markov_gp_trainer.py:163-201): gradients with respect
 * to the drift parameters, i.e. what the reference's tape returns for prior_sde.trainable_variables.  Every entry
 * of the drift's parameter block gets a gradient (the caller keeps the trainable ones); grads is [B, 16], entries
 * beyond vdp_drift_num_params(drift_id) are set to 0.  Sums over the grid are reduced in a fixed order.
 * ---------------------------------------------------------------------------------------------- */
int vdp_drift_num_params(int drift_id);

/* CVISitesSDE.grad_KL_wrt_prior_params (variational_cvi_sde.py:495-504): d KL_q_p / d params.  dist_q is rebuilt
 * from the sites as KL_q_p does (:450-462): q*, fx_mus, fx_covs of the problem are outputs (pass scratch buffers
 * for fx_* to leave the model's stored marginals alone, as the reference does); H = 20. */
int vdp_cvi_grad_kl_prior_params(const vdp_cvi_problem* problem_host, double* grads, void* stream);

/* CVISitesSDE.grad_VE_wrt_prior_params (variational_cvi_sde.py:506-518): d(-VE)/d params with the tape running
 * through set_linearized_prior -> dist_q.marginals -> variational_expectation.  The caller has re-linearised the
 * prior (vdp_linearize_to_ssm; pA, pb = the clipped SSM) around the stored posterior path path_m [B,N+1,D],
 * path_S [B,N+1,D,D] (grid points 1..N are used, :412); clip / clip_lo / clip_hi = the stabilize_ssm window
 * (:417-430; saturated entries pass no gradient).  q*, fx_mus, fx_covs of the problem are outputs (the marginals
 * of dist_q under the new prior) and must not alias the path. */
int vdp_cvi_grad_ve_prior_params(const vdp_cvi_problem* problem_host, const double* path_m, const double* path_S,
                                 int clip, double clip_lo, double clip_hi, double* grads, void* stream);

/* The quadrature both KL-type gradients share, H = 20:
 *   grads[b, p] = out_scale sum_t E_{N(m_t, S_t)}[ r(x)^T W df/dp(x) ],   W = sym((q_scale q)^-1),
 *   r(x) = c_x x + c_f f(x) - (a_sign A_t x + b_t),     A [B,N,D,D], b [B,N,D];
 * the Gaussian of transition t of sequence b is read at m + (b * points_per_sequence + t) * D (same for S), so
 * marginals [B,N+1,..] can be passed with or without an offset of one grid point.
 *   VariationalMarkovGP.grad_prior_sde_params (vi_sde.py:457-470): c_x = 0, c_f = 1, a_sign = -1, q_scale = 1,
 *   out_scale = dt, (m, S) = forward_pass[1:] (the reference pairs grid points 1..N with A[0..N-1] here). */
size_t vdp_drift_param_grad_workspace(int64_t B, int64_t N, int D);
int vdp_drift_param_grad(int drift_id, const double* params_host, const double* q_host, double q_scale, double c_x,
                         double c_f, double a_sign, double out_scale, const double* A, const double* b,
                         const double* m, const double* S, int64_t points_per_sequence, double* grads,
                         void* workspace, size_t workspace_bytes, int64_t B, int64_t N, int D, int64_t* status,
                         void* stream);

/* euler_maruyama (markovflow/sde/sde_utils.py:36-96), batched: x0 [B,D], time_grid [n] (device) -> out [B,n,D]
 * with out[:,0] = x0 and the reference's time convention (clock starts at 0; returns [x0, first n-1 states]).
 * noise: standard normals [B,n-1,D] used in place of the generator (NULL = Philox4x32-10 with `seed`,
 * sub-sequence = sequence index).  Input generator for BASELINE configs 4 and 5. */
int vdp_euler_maruyama(int drift_id, const double* params_host, const double* q_host, const double* x0,
                       const double* time_grid, const double* noise, unsigned long long seed, double* out,
                       int64_t B, int64_t n, int D, int64_t* status, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Host-buffer variant of vdp_cvi_site_update: the call a host program makes when its arrays live in HOST
 * memory (what the reference's Python loop holds, cvi_dp_trainer.py:71-77).  Every array pointer of `problem`
 * addresses host memory (pinned memory gives full PCIe speed; pageable memory works, slower); workspace,
 * workspace_bytes and status of the struct are ignored.  The B sequences are cut into chunks of `chunk`
 * sequences; chunk c is uploaded, updated and downloaded on stream c % slots, so with slots >= 3 the upload of
 * one chunk, the kernels of the previous and the download of the one before overlap.  Results are identical to
 * vdp_cvi_site_update on the whole batch (sequences are independent).  The call returns when all results are
 * in the host buffers.
 *   in/out (host): g1, gd, gs, d1, d2, fx_mus (of the input only the observed grid points are read and uploaded)
 *   in (host): time_grid, obs_idx, obs_y, prior
 *   out (host): fx_covs;  out (host, each optional = NULL): qA, qb, qcholQ, qmu0, qcholP0, elbo, ve, kl
 * The linearised prior changes only at a re-linearisation (cvi_dp_trainer.py:110-134), so it can be kept on the
 * device: upload it once with vdp_cvi_host_set_prior and pass pA = NULL in the site updates that follow.
 * status_host2 (optional) receives the device status word; a reported grid index is relative to its chunk.
 * ---------------------------------------------------------------------------------------------- */
typedef struct vdp_cvi_host_ctx vdp_cvi_host_ctx;
int vdp_cvi_host_create(vdp_cvi_host_ctx** ctx, int64_t chunk, int64_t N, int D, int64_t n_obs, int slots,
                        int device);
int vdp_cvi_host_set_prior(vdp_cvi_host_ctx* ctx, int64_t B, const double* pA, const double* pb,
                           const double* pcholQ, const double* pmu0, const double* pcholP0);
int vdp_cvi_host_site_update(vdp_cvi_host_ctx* ctx, const vdp_cvi_problem* problem, int64_t* status_host2);
int vdp_cvi_host_destroy(vdp_cvi_host_ctx* ctx);

/* Resident-state session of the host-buffer interface: the state of the B sequences (sites, linearised prior,
 * posterior, marginals, carried naturals) is uploaded ONCE and stays on the device between calls, the way the
 * reference's model object keeps its tf.Variables between trainer iterations (cvi_dp_trainer.py:71-77); per site
 * update the host hands over that update's inputs -- the observation values, host memory -- and receives its results
 * in host memory: ELBO terms [B] and the posterior marginals fx_mus [B,N+1,D], fx_covs [B,N+1,D,D] (each optional =
 * NULL).  Sequences run in chunks of `chunk` on `streams` rotating streams, so the download of one chunk overlaps the
 * kernels of the next; consecutive updates carry the naturals and keep the girsanov sites in theta.
 *   vdp_cvi_session_create          every array pointer of `problem` addresses HOST memory; obs_idx is fixed for the
 *                                   session (the observation times), obs_y is the initial set of values
 *   vdp_cvi_session_site_update     obs_y_host [B,n_obs,D] (NULL: unchanged); flags: VDP_FLAG_GH_QUADRATURE,
 *                                   VDP_FLAG_REFACTOR_PRECISION, VDP_FLAG_SEPARATE_KL_VALUE; returns when the results
 *                                   are in the host buffers
 *   vdp_cvi_session_download_state  the whole state back into the host arrays of `problem` (girsanov sites
 *                                   materialised from theta); NULL pointers are skipped */
typedef struct vdp_cvi_session vdp_cvi_session;
int vdp_cvi_session_create(vdp_cvi_session** session, const vdp_cvi_problem* problem_host_arrays, int64_t chunk,
                           int streams, int device);
int vdp_cvi_session_site_update(vdp_cvi_session* session, const double* obs_y_host, double lr_data,
                                double lr_girsanov, int flags, double* elbo_host, double* ve_host, double* kl_host,
                                double* fx_mus_host, double* fx_covs_host, int64_t* status_host2);
int vdp_cvi_session_download_state(vdp_cvi_session* session, const vdp_cvi_problem* problem_host_arrays);
int vdp_cvi_session_destroy(vdp_cvi_session* session);

/* ------------------------------------------------------------------------------------------------
 * VDP: VariationalMarkovGP (markovflow/models/vi_sde.py:63-482), batched over B sequences.
 *   vdp_vgp_forward_pass               forward_pass              vi_sde.py:171-204  -> m, S
 *   vdp_vgp_update_lagrange            update_lagrange           vi_sde.py:289-347  (m, S) -> lam, psi
 *                                      (+ _grad_E_sde :206-240, _jump_conditions :262-287)
 *   vdp_vgp_update_param               update_param              vi_sde.py:377-414  (m, S, lam, psi) -> A, b
 *   vdp_vgp_update_initial_statistics  update_initial_statistics vi_sde.py:242-260  -> q_mean, q_chol
 *   vdp_vgp_elbo                       elbo                      vi_sde.py:436-455  (m, S current) -> elbo
 *   vdp_vgp_sweep                      one iteration of VIMarkovGPTrainer.perform_inference
 *                                      (docs/diffusion_processes/vi_markov_gp_trainer.py:62-73)
 * ---------------------------------------------------------------------------------------------- */
typedef struct vdp_vgp_problem {
    int64_t B, N, n_obs;
    int32_t D, drift_id;
    double drift_params[16];
    double q[9];               /* spectral density, row-major D x D */
    double dt;                 /* float(grid[1] - grid[0]) */
    double x0_mu[3], x0_cov[9];/* prior on the initial state */
    double lik_chol[9];        /* Cholesky factor of the Gaussian likelihood covariance */
    double lr, x0_lr;          /* damping of update_param / update_initial_statistics */
    const double* time_grid;   /* [N+1] */
    double *A, *b;             /* variational drift  [B,N,D,D] [B,N,D]   (in/out) */
    double *lam, *psi;         /* Lagrange multipliers [B,N,D] [B,N,D,D] (in/out) */
    double *q_mean, *q_chol;   /* q(x0): [B,D] [B,D,D]                   (in/out) */
    const int64_t* obs_idx;    /* [B,n_obs] */
    const double* obs_y;       /* [B,n_obs,D] */
    double *m, *S;             /* marginals [B,N+1,D] [B,N+1,D,D] */
    double* elbo;              /* [B] */
    void* workspace;
    size_t workspace_bytes;    /* >= vdp_vgp_workspace(B,N,D,n_obs) */
    int64_t* status;
    int32_t flags;             /* VDP_VGP_* */
    int32_t reserved;
} vdp_vgp_problem;

#define VDP_VGP_STABILIZE 1    /* stabilize_system=True (clips of vi_sde.py:184-197, 314-323, 393-397) */
#define VDP_VGP_UPDATE_X0 2    /* vdp_vgp_sweep also runs update_initial_statistics */
#define VDP_VGP_GH_QUADRATURE 4 /* as VDP_FLAG_GH_QUADRATURE */
#define VDP_VGP_REFACTOR_PRECISION 8 /* forward_pass: marginal covariances through the rebuilt precision (banded
                                  Cholesky + inverse band, state_space_model.py:262) as the reference, instead of the
                                  forward recursion S_{k+1} = A S A^T + Q riding on the mean sweep (equal to rounding) */
#define VDP_VGP_STATE_CARRIED 16 /* vdp_vgp_sweep only.  The caller's promise: the previous call on this problem was a
                                  vdp_vgp_sweep and nothing it read or wrote (A, b, the initial statistics, m, S, the
                                  observations, the workspace) has changed since.  The sweep ends with the forward pass
                                  and the E_sde quadrature of the parameters it has just written -- for its ELBO -- and
                                  the next sweep would begin by computing exactly those again (vi_markov_gp_trainer.py:
                                  62-73 calls update_param, which starts from forward_pass, then elbo, which runs it
                                  again): with the flag it starts from the marginals and dE_sde/dm, dE_sde/dS the
                                  workspace still holds.  Same results bit for bit. */

/* LinearDrift.to_ssm (markovflow/sde/drift.py:64-116): A_ssm = I + A delta_k, b_ssm = b delta_k,
 * cholQ = chol(q delta_k).  A [B,N,D,D], b [B,N,D] -> same shapes. */
int vdp_linear_drift_to_ssm(const double* A, const double* b, const double* q_host, const double* time_grid,
                            double* ssm_A, double* ssm_b, double* ssm_cholQ, int64_t B, int64_t N, int D,
                            int64_t* status, void* stream);

/* squared_drift_difference_along_Gaussian_path (sde_utils.py:182-249), H = 20, linear drift (-A, b):
 *   value[b] = 0.5 dt sum_k E_{N(m_k,S_k)} || -A_k x + b_k - f(x) ||^2_{q^-1},  k = 0..N-1,
 * m [B,N+1,D], S [B,N+1,D,D] (the last grid point is not used); optional dEdm [B,N,D], dEdS [B,N,D,D]
 * = gradients wrt (m_k, S_k) divided by dt (vi_sde.py:206-240). */
size_t vdp_sq_drift_difference_workspace(int64_t B, int64_t N, int D);
int vdp_sq_drift_difference(int drift_id, const double* params_host, const double* q_host, double dt,
                            const double* A, const double* b, const double* m, const double* S, double* value,
                            double* dEdm, double* dEdS, void* workspace, size_t workspace_bytes, int64_t B,
                            int64_t N, int D, int64_t* status, void* stream);

size_t vdp_vgp_workspace(int64_t B, int64_t N, int D, int64_t n_obs);
int vdp_vgp_forward_pass(const vdp_vgp_problem* problem_host, void* stream);
int vdp_vgp_update_lagrange(const vdp_vgp_problem* problem_host, void* stream);
int vdp_vgp_update_param(const vdp_vgp_problem* problem_host, void* stream);
int vdp_vgp_update_initial_statistics(const vdp_vgp_problem* problem_host, void* stream);
int vdp_vgp_elbo(const vdp_vgp_problem* problem_host, void* stream);
int vdp_vgp_sweep(const vdp_vgp_problem* problem_host, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Time-segment sharding of one long sequence over ranks (BASELINE config 5).  Rank r owns the contiguous
 * transitions [s_r, e_r) of the grid (N_local = e_r - s_r) and runs the tiled sweeps on them; ranks exchange
 * one associative element per sweep ("total") and the natural initial state of the first / last rank
 * ("init"), e.g. with an NCCL all-gather of vdp_dist_blob_bytes() bytes each.  Sequence per sweep:
 *     vdp_dist_sweep_summary(p, total, init)            -> this rank's blobs (device memory)
 *     all-gather -> all_totals [nranks x total_bytes], all_inits [nranks x init_bytes] (device memory)
 *     vdp_dist_sweep_finish(p, all_totals, all_inits)   -> outputs of the local range
 * VDP_SWEEP_NAT2SSM has two dependent recursions: summary(stage 0) -> gather -> summary(stage 1, with
 * prev_totals / prev_inits = the gathered stage-0 blobs) -> gather -> finish (stage-1 blobs).
 *
 * Local arrays (one sequence, B = 1; D x D blocks row-major) and their halo slots:
 *   VDP_SWEEP_CHOL     in[0] Kd [N_local+1]  (last block = first node of the next rank; on the last rank the
 *                      final node), in[1] Ks [N_local]; scaleD / scaleS multiply them (-2 / -1 for naturals);
 *                      out[0] Ld [N_local+1], out[1] Ls [N_local]
 *   VDP_SWEEP_NAT2SSM  in[0] Ld, in[1] Ls, in[2] th1 [N_local(+1 on the last rank), D], in[3] thd, in[4] ths;
 *                      out[0] A [N_local], out[1] b, out[2] cholQ: pointers to the slot of transition s_r, the
 *                      slot BEFORE it must exist on ranks > 0 (it receives the output of node s_r);
 *                      out[3] mu0, out[4] cholP0 (rank 0 only)
 *   VDP_SWEEP_MARG     in[0] A [N_local+1] (last = first transition of the next rank, unused on the last
 *                      rank), in[1] b [N_local], in[2] cholQ (slot -1 = factor of node s_r on ranks > 0, slot
 *                      N_local = halo as for A), in[3] mu0, in[4] cholP0 (rank 0 only);
 *                      out[0] m [N_local+1, D], out[1] Ld [N_local+1], out[2] Ls [N_local]
 *   VDP_SWEEP_INVBAND  in[0] Ld, in[1] Ls; out[0] Sd [N_local (+1 on the last rank)], out[1] Ss or NULL
 *   VDP_SWEEP_MEAN     in[0] A [N_local], in[1] b [N_local], in[2] mu0 (rank 0 only); out[0] m [N_local+1, D]
 *                      (marginal means alone, state_space_model.py:251)
 * VDP_SWEEP_NAT2SSM with out[5] = Sd [N_local (+1 on the last rank), D, D] != NULL also writes the marginal
 * covariances of the Gaussian the naturals describe (the diagonal of the inverse band the sweep walks through);
 * together with VDP_SWEEP_MEAN that is the rebuild's marginals without VDP_SWEEP_MARG + VDP_SWEEP_INVBAND (a second
 * factorisation of the same precision).
 *
 * The rebuild in two sweeps instead of three (no factor stored; equal to the route above in exact arithmetic):
 *   VDP_SWEEP_INFO     naturals_to_ssm_params as ONE backward sweep in information form (the Schur complements of
 *                      the precision taken from the bottom, the linear term riding along):
 *                      in[0] th1 [N_local (+1 on the last rank), D], in[1] thd, in[2] ths [N_local];
 *                      out[0] A, out[1] b, out[2] cholQ [N_local] (each at the index of its transition, no halo
 *                      slots), out[3] mu0, out[4] cholP0 (rank 0 only)
 *   VDP_SWEEP_MOMENTS  marginal means and covariances by the moment recursion m' = A m + b, S' = A S A^T + Q:
 *                      in[0] A, in[1] b, in[2] cholQ [N_local], in[3] mu0, in[4] cholP0 (rank 0 only);
 *                      out[0] m [N_local+1, D], out[1] Sd [N_local+1, D, D] (last row = the first grid point of the
 *                      next rank)
 * ---------------------------------------------------------------------------------------------- */
enum { VDP_SWEEP_CHOL = 0, VDP_SWEEP_NAT2SSM = 1, VDP_SWEEP_MARG = 2, VDP_SWEEP_INVBAND = 3, VDP_SWEEP_MEAN = 4,
       VDP_SWEEP_INFO = 5, VDP_SWEEP_MOMENTS = 6 };

typedef struct vdp_dist_sweep {
    int32_t kind, stage;
    int32_t rank, nranks;
    int64_t N_local;
    int32_t D, reserved;
    double scaleD, scaleS;
    const double* in[8];
    double* out[8];
    const void* prev_totals;   /* NAT2SSM stage 1: gathered stage-0 blobs */
    const void* prev_inits;
    void* workspace;           /* vdp_dist_workspace(D) bytes, the SAME buffer for all calls of one sweep */
    size_t workspace_bytes;
    int64_t* status;
} vdp_dist_sweep;

/* Exchange between ranks issued from C on the compute stream: a communicator over NCCL (bound at run time with
 * dlopen("libnccl.so.2"); the library has no link-time dependency on NCCL).  One process per GPU:
 *     rank 0: vdp_comm_unique_id(id)  ->  the 128 bytes travel to every rank by any out-of-band channel
 *     every rank, its GPU current:  vdp_comm_create(&comm, id, rank, nranks)
 *     vdp_comm_all_gather(comm, send, recv, bytes_per_rank, stream): recv [nranks x bytes_per_rank], device memory,
 *     asynchronous on `stream` (ncclAllGather); this is the all-gather of the sequence above. */
typedef struct vdp_comm vdp_comm;
int vdp_comm_unique_id(void* id128_host);
int vdp_comm_create(vdp_comm** comm, const void* id128_host, int rank, int nranks);
int vdp_comm_all_gather(vdp_comm* comm, const void* send, void* recv, size_t bytes_per_rank, void* stream);
int vdp_comm_destroy(vdp_comm* comm);

size_t vdp_dist_workspace(int D);
int vdp_dist_blob_bytes(int kind, int stage, int D, size_t* total_bytes, size_t* init_bytes);
int vdp_dist_sweep_summary(const vdp_dist_sweep* sweep_host, void* total_blob, void* init_blob, void* stream);
int vdp_dist_sweep_finish(const vdp_dist_sweep* sweep_host, const void* all_totals, const void* all_inits,
                          void* stream);

/* ------------------------------------------------------------------------------------------------
 * diagnostics: per-phase device time of the fused pipelines (CUDA events on the launching stream).
 * vdp_profile_enable(1) starts recording (and clears earlier marks); vdp_profile_read synchronises on
 * the last mark, returns accumulated milliseconds and number of spans per phase, and clears the marks.
 * Not thread safe; meant for bench.py.
 * ---------------------------------------------------------------------------------------------- */
enum {
    VDP_PHASE_DATA_SITES = 0,   /* k_data_site_update */
    VDP_PHASE_SITES_TO_NAT = 1, /* k_ssm_to_naturals (+ k_scatter_data) */
    VDP_PHASE_CHOL_FWD = 2,     /* k_chol_fwd */
    VDP_PHASE_NAT2SSM_BWD = 3,  /* k_nat2ssm_bwd */
    VDP_PHASE_MARG_FWD = 4,     /* k_marg_fwd */
    VDP_PHASE_INVBAND_BWD = 5,  /* k_invband_bwd */
    VDP_PHASE_KL_GRAD = 6,      /* memsets + k_scatter_data + k_kl_grad */
    VDP_PHASE_ELBO = 7,         /* k_variational_expectation + k_kl_value + k_reduce_terms + k_elbo_combine */
    VDP_PHASE_COUNT = 8,
    VDP_PHASE_END = -1
};
/* Diagnostic: the library's float64 reciprocal / square root / division primitives (csrc/fastmath.cuh:
 * hardware seed + Newton steps, no IEEE slow path), element-wise on device arrays of n positive normal numbers:
 * rcp = 1/x, sq = sqrt(x), rsq = 1/sqrt(x), quo = y/x.  Used by the tests to bound their error in ulps. */
int vdp_test_fastmath(const double* x, const double* y, double* rcp, double* sq, double* rsq, double* quo,
                      int64_t n, void* stream);

/* Diagnostic: float64 FMA throughput (the roofline denominator of the quadrature kernels, which are bound by the
 * FP64 pipe): blocks x threads threads run `iters` rounds of 16 independent fused multiply-adds each;
 * flops = 2 * 16 * iters * blocks * threads.  out [blocks * threads] receives the chains' sums.  Time it with events. */
int vdp_test_fp64_fma(double* out, int64_t blocks, int threads, int64_t iters, void* stream);

int vdp_profile_enable(int on);
int vdp_profile_read(double* ms_by_phase, int64_t* spans_by_phase, int n_phases);

#ifdef __cplusplus
}
#endif
#endif /* VDP_B200_H */
