"""
Oracle, layer L3: the CVI-DP and VDP models (one site update / one fixed-point sweep) and the
Gaussian likelihood they touch, in NumPy float64.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Nothing here is shipped.

Follows (paths relative to /root/reference):
  markovflow/models/variational_cvi_sde.py   CVISitesSSM, CVISitesSDE
  markovflow/models/vi_sde.py                VariationalMarkovGP
  markovflow/likelihoods/multivariate_gaussian.py:82-116   variational_expectations
  markovflow/models/variational_cvi.py:448-462             gradient_transformation_mean_var_to_expectation
  markovflow/ssm_natgrad.py:121-218                        SSMNaturalGradient (plain step, no momentum)
  docs/diffusion_processes/cvi_dp_trainer.py:63-136        loop structure of one trainer iteration

Reference quirks reproduced on purpose (each is what "parity" means here):
  * ``dist_q`` is rebuilt from the sites on every access (variational_cvi_sde.py:176-192);
  * KL_q_p squares the Cholesky factor ELEMENT-WISE (variational_cvi_sde.py:452) -- equal to Q only for D = 1;
  * the "trace" term multiplies element-wise (sde_utils.py:327);
  * VDP backward sweep uses psi A + psi A (vi_sde.py:338), not psi A + A^T psi.
"""
from __future__ import annotations

import numpy as np

from .gauss_markov_np import (BTDGaussian, StateSpaceModel, btd_cholesky, btd_solve, chol, chol_solve, expectations_to_ssm_params, mT,
                              naturals_to_ssm_params, ssm_to_expectations, ssm_to_naturals, tri_solve)
from .sde_np import (SDE, Gaussian, LinearDrift, SDE_SSM_KL_with_grads_wrt_exp_params, SSM_KL_along_Gaussian_path,
                     gaussian_kl, grad_squared_drift_difference, linearize_sde, mvnquad, remove_batch_from_ssm,
                     squared_drift_difference_along_Gaussian_path, ssm_to_btd_nat, tranform_girsanov_sites)


class MultivariateGaussian:
    """likelihoods/multivariate_gaussian.py: Gaussian likelihood with Cholesky factor ``chol_covariance``."""

    def __init__(self, chol_covariance):
        self.chol_covariance = np.atleast_2d(np.asarray(chol_covariance, dtype=np.float64))
        self.obs_dim = self.chol_covariance.shape[-1]

    def log_probability_density(self, fs, observations):
        D = self.obs_dim
        z = tri_solve(self.chol_covariance, (observations - fs)[..., None])[..., 0]
        logdet = np.sum(np.log(np.abs(np.diag(self.chol_covariance))))
        return -0.5 * np.sum(z * z, axis=-1) - logdet - 0.5 * D * np.log(2.0 * np.pi)

    def variational_expectations(self, f_means, f_covariances, observations):
        """multivariate_gaussian.py:82-116: -1/2 Tr(Sigma^-1 S_i) + log N(y_i; mu_i, Sigma)."""
        inv_cov = chol_solve(self.chol_covariance, np.eye(self.obs_dim))
        return -0.5 * np.sum(inv_cov * f_covariances, axis=(-1, -2)) + self.log_probability_density(
            f_means, observations)

    def variational_expectations_grads(self, f_means, f_covariances, observations):
        """Closed-form tape result of variational_cvi_sde.py:211-216: d sum VE / d(mu, S)."""
        inv_cov = chol_solve(self.chol_covariance, np.eye(self.obs_dim))
        g_mu = (0.5 * (inv_cov + inv_cov.T) @ (observations - f_means)[..., None])[..., 0]
        g_S = np.broadcast_to(-0.5 * inv_cov, f_covariances.shape).copy()
        return g_mu, g_S

    def predict_mean_and_var(self, f_means, f_covariances):
        return f_means, self.chol_covariance @ self.chol_covariance.T + f_covariances


def gradient_transformation_mean_var_to_expectation(inputs, grads):
    """models/variational_cvi.py:448-462."""
    if grads[1].ndim == 2:
        return grads[0] - 2.0 * grads[1] * inputs[0], grads[1]
    return grads[0] - 2.0 * (grads[1] @ inputs[0][..., None])[:, :, 0], grads[1]


class GaussianSitesNat:
    """kalman_filter.py:382-436 container (nat1, nat2, log_norm)."""

    def __init__(self, nat1, nat2, log_norm=None):
        self.nat1 = np.array(nat1, dtype=np.float64)
        self.nat2 = np.array(nat2, dtype=np.float64)
        self.log_norm = log_norm


def obs_indices(time_grid, obs_times):
    """variational_cvi_sde.py:117: tf.where(tf.equal(grid[:, None], obs_times))[:, 0] -- sorted grid indices."""
    return np.where(np.equal(np.asarray(time_grid)[..., None], np.asarray(obs_times)))[0]


class CVISitesSDE:
    """
    variational_cvi_sde.py:49-518 (CVISitesSSM + the CVISitesSDE overrides), non-batched.
    """

    def __init__(self, prior_sde: SDE, time_grid, input_data, likelihood: MultivariateGaussian,
                 prior_initial_state: Gaussian = None, initial_posterior_path: Gaussian = None,
                 stabilize_ssm: bool = True, clip_state_transitions=(-1.0, 1.0)):
        self.state_dim = input_data[-1].shape[-1]
        self.prior_sde = prior_sde
        if prior_initial_state is None:  # :434-444
            prior_initial_state = Gaussian(np.zeros(self.state_dim),
                                           prior_sde.q * np.ones((self.state_dim, self.state_dim)))
        self.prior_initial_state = prior_initial_state
        self.likelihood = likelihood
        self.time_grid = np.asarray(time_grid, dtype=np.float64)
        self.dt = float(self.time_grid[1] - self.time_grid[0])
        self.dist_p = None
        self._observations_time_points = np.asarray(input_data[0], dtype=np.float64)
        self._observations = np.asarray(input_data[1], dtype=np.float64)
        self.output_dim = self._observations.shape[-1]
        n1, D, n_obs = self.time_grid.shape[0], self.state_dim, self._observations_time_points.shape[0]
        # :135-151
        self.girsanov_sites = BTDGaussian(np.zeros((n1, D)), np.ones((n1, D, D)) * -1e-10,
                                          np.ones((n1 - 1, D, D)) * -1e-10)
        # :97-104
        self.data_sites = GaussianSitesNat(np.zeros((n_obs, D)), np.broadcast_to(np.eye(D), (n_obs, D, D)) * 1e-10)
        if initial_posterior_path is None:  # :125-133
            initial_posterior_path = Gaussian(np.zeros((n1, D)), np.broadcast_to(np.eye(D), (n1, D, D)).copy())
        self.fx_mus = initial_posterior_path.mu.reshape(n1, D).astype(np.float64)
        self.fx_covs = initial_posterior_path.cov.reshape(n1, D, D).astype(np.float64)
        self.obs_sites_indices = obs_indices(self.time_grid, self._observations_time_points)
        assert self.obs_sites_indices.shape[0] == self._observations.shape[0]
        self.dist_p_linearized = None
        self.stabilize_ssm = stabilize_ssm
        self.clip_state_transitions = clip_state_transitions
        self.set_linearized_prior()

    # ---- :408-432
    def set_linearized_prior(self):
        posterior_path = Gaussian(self.fx_mus[1:], self.fx_covs[1:])
        lin = linearize_sde(self.prior_sde, self.time_grid, posterior_path, self.prior_initial_state)
        self.dist_p_linearized = remove_batch_from_ssm(lin)
        if self.stabilize_ssm:
            lo, hi = self.clip_state_transitions
            self.dist_p = StateSpaceModel(
                self.dist_p_linearized.initial_mean, self.dist_p_linearized.cholesky_initial_covariance,
                np.clip(self.dist_p_linearized.state_transitions, lo, hi),
                np.clip(self.dist_p_linearized.state_offsets, lo, hi),
                self.dist_p_linearized.cholesky_process_covariances)
        else:
            self.dist_p = self.dist_p_linearized

    def _scatter_data_sites(self):
        n1, D = self.girsanov_sites.nat1.shape
        nat1 = np.zeros((n1, D))
        nat2 = np.zeros((n1, D, D))
        np.add.at(nat1, self.obs_sites_indices, self.data_sites.nat1)  # tf.scatter_nd sums duplicates
        np.add.at(nat2, self.obs_sites_indices, self.data_sites.nat2)
        return nat1, nat2

    # ---- :161-174
    def full_sites(self):
        lin = ssm_to_btd_nat(self.dist_p)
        d1, d2 = self._scatter_data_sites()
        q_nat1 = lin.nat1 + self.girsanov_sites.nat1 + d1
        q_nat2 = lin.nat2 + self.girsanov_sites.nat2 + d2
        return q_nat1, q_nat2

    # ---- :176-192
    @property
    def dist_q(self) -> StateSpaceModel:
        q_nat1, q_nat2 = self.full_sites()
        As, offsets, chol_P0, chol_Q, mu0 = naturals_to_ssm_params(q_nat1, q_nat2.block_diagonal,
                                                                   q_nat2.block_sub_diagonal)
        return StateSpaceModel(mu0, chol_P0, As, offsets, chol_Q)

    # ---- :204-220
    def local_objective_and_gradients(self, Fmu, Fvar):
        local_obj = np.sum(self.likelihood.variational_expectations(Fmu, Fvar, self._observations))
        grads = self.likelihood.variational_expectations_grads(Fmu, Fvar, self._observations)
        grads = gradient_transformation_mean_var_to_expectation((Fmu, Fvar), grads)
        return local_obj, grads

    # ---- :488-493
    def grad_kl_wrt_exp_param(self):
        return SDE_SSM_KL_with_grads_wrt_exp_params(self.dist_q, self.prior_sde, self.dt, self.prior_initial_state,
                                                    self.time_grid)

    # ---- :279-299
    def update_girsanov_sites(self, lr: float):
        _, grad_kl = self.grad_kl_wrt_exp_param()
        d1, d2 = self._scatter_data_sites()
        g = self.girsanov_sites
        new1 = g.nat1 + lr * (d1 - grad_kl[0])
        new2 = g.nat2.block_diagonal + lr * (d2 - grad_kl[1])
        new3 = g.nat2.block_sub_diagonal - lr * grad_kl[2]
        self.girsanov_sites = BTDGaussian(new1, new2, new3)
        self.fx_mus, self.fx_covs = self.dist_q.marginals

    # ---- :301-317
    def update_data_sites(self, lr: float):
        fx_mus = self.fx_mus.reshape(-1, self.state_dim)[self.obs_sites_indices]
        fx_covs = self.fx_covs.reshape(-1, self.state_dim, self.state_dim)[self.obs_sites_indices]
        _, grads_ve = self.local_objective_and_gradients(fx_mus, fx_covs)
        self.data_sites.nat1 = (1 - lr) * self.data_sites.nat1 + lr * grads_ve[0]
        self.data_sites.nat2 = (1 - lr) * self.data_sites.nat2 + lr * grads_ve[1]
        self.fx_mus, self.fx_covs = self.dist_q.marginals

    # ---- :319-337
    def variational_expectation(self, fx_mus=None, fx_covs=None):
        if fx_mus is None or fx_covs is None:
            fx_mus, fx_covs = self.dist_q.marginals
        mu_obs = fx_mus.reshape(-1, self.state_dim)[self.obs_sites_indices]
        cov_obs = fx_covs.reshape(-1, self.state_dim, self.state_dim)[self.obs_sites_indices]
        return np.sum(self.likelihood.variational_expectations(mu_obs, cov_obs, self._observations))

    # ---- :446-486
    def KL_q_p(self):
        dist_q = self.dist_q  # every access in the reference rebuilds the same values
        m = dist_q.marginal_means
        S = dist_q.marginal_covariances
        Q_q = np.square(dist_q.cholesky_process_covariances)            # element-wise square (:452)
        Q_p = self.dt * np.repeat(self.prior_sde.q[None], Q_q.shape[0], axis=0)
        A, b = dist_q.state_transitions, dist_q.state_offsets

        def f_q(x):
            return (A[None] @ x[..., None])[..., 0] + b[None]

        def f_p(x):
            return x + self.dt * self.prior_sde.drift(x)

        kl_q_p = SSM_KL_along_Gaussian_path(f_q, f_p, Q_q, Q_p, m, S)
        cq = dist_q.cholesky_initial_covariance
        cp = self.dist_p.cholesky_initial_covariance
        kl_0 = gaussian_kl(dist_q.initial_mean, cq @ cq.T, self.dist_p.initial_mean, cp @ cp.T)
        return kl_q_p + kl_0

    # ---- :339-352
    def classic_elbo(self):
        fx_mus, fx_covs = self.dist_q.marginals
        return self.variational_expectation(fx_mus, fx_covs) - self.KL_q_p()

    # ---- :495-504
    def grad_KL_wrt_prior_params(self):
        """
        d KL_q_p / d(prior_sde parameters), every entry of ``prior_sde.params``.  The tape of the reference watches
        the drift's trainable variables only; dist_q, dist_p and f_q are constants on it (f_q is wrapped in
        stop_gradient, :466), so the only dependence is through ``prior_sde.drift`` inside the H = 20 quadrature:
            d/dp 1/2 sum_t E[d^T Q_p^-1 d] = sum_t E[d^T sym(Q_p^-1) dt df/dp],   d = x + dt f(x) - (A x + b).
        """
        dist_q = self.dist_q
        m, S = dist_q.marginal_means[:-1], dist_q.marginal_covariances[:-1]
        A, b = dist_q.state_transitions, dist_q.state_offsets
        D, n = self.state_dim, A.shape[0]
        Qp_inv = chol_solve(chol(self.dt * self.prior_sde.q), np.eye(D))
        W = 0.5 * (Qp_inv + Qp_inv.T)
        n_par = len(np.asarray(self.prior_sde.params))

        def func(x):
            x = x.reshape(-1, n, D)
            d = x + self.dt * self.prior_sde.drift(x) - ((A[None] @ x[..., None])[..., 0] + b[None])
            df, _ = self.prior_sde.param_grads(x)
            return (self.dt * np.einsum("pni,ij,kpnj->pnk", d, W, df)).reshape(-1, n_par)

        return np.sum(mvnquad(func, m, S, H=20, Din=D, Dout=(n_par,)), axis=0)

    # ---- :506-518
    def grad_VE_wrt_prior_params(self):
        """
        d(-VE)/d(prior_sde parameters) with the tape running through set_linearized_prior -> dist_q.marginals ->
        variational_expectation.  As in the reference the call RE-LINEARISES the prior around the stored
        (fx_mus, fx_covs) as a side effect and leaves fx_mus / fx_covs untouched.

        Route of the restatement (forward mode in the parameters, reverse mode in the chain):
          * dA_k/dp, db_k/dp of the linearised SSM: H = 10 expectations of dJ/dp, df/dp (sde_utils.py:156-166,
            drift.py:102-108), zero where tf.clip_by_value saturates (:417-430);
          * q has precision P and linear natural parameter h (prior naturals + sites); with L = -VE,
            gbar_j = dL/dm_j, G_j = dL/dSigma_jj at the observed points:  dL = u^T dh - u^T dP m - tr(V dP),
            u = P^-1 gbar,  V = P^-1 Gbar P^-1 (its block-tridiagonal band), both from block solves with the
            Cholesky factor of P (one solve per observed point and state dimension);
          * dP, dh from the differential of ssm_to_naturals (ssm_gaussian_transformations.py:182-253).
        Checked against torch.autograd on a dense restatement and central differences
        (tests/test_oracle_prior_grads.py).
        """
        self.set_linearized_prior()
        sde, D = self.prior_sde, self.state_dim
        n_par = len(np.asarray(sde.params))
        path_m, path_S = self.fx_mus[1:], self.fx_covs[1:]
        N = path_m.shape[0]
        deltas = (self.time_grid[1:] - self.time_grid[:-1])[:, None, None]
        E_dJ = mvnquad(lambda x: np.moveaxis(sde.param_grads(x)[1], 0, -1), path_m, path_S, H=10, Din=D,
                       Dout=(D, D, n_par))
        E_df = mvnquad(lambda x: np.moveaxis(sde.param_grads(x)[0], 0, -1), path_m, path_S, H=10, Din=D,
                       Dout=(D, n_par))
        dA = deltas[..., None] * E_dJ                                               # [N, D, D, P]
        db = deltas[..., 0, :, None] * (E_df - np.einsum("nijp,nj->nip", E_dJ, path_m))  # [N, D, P]
        if self.stabilize_ssm:
            lo, hi = self.clip_state_transitions
            A_lin, b_lin = self.dist_p_linearized.state_transitions, self.dist_p_linearized.state_offsets
            dA = dA * ((A_lin >= lo) & (A_lin <= hi))[..., None]
            db = db * ((b_lin >= lo) & (b_lin <= hi))[..., None]
        # posterior q and the cotangents of L = -VE at the observed grid points
        q_nat1, q_nat2 = self.full_sites()
        Pd, Ps = -2.0 * q_nat2.block_diagonal, -q_nat2.block_sub_diagonal
        Ld, Ls = btd_cholesky(Pd, Ps)
        fx_mus, fx_covs = self.dist_q.marginals
        idx = self.obs_sites_indices
        g_mu, g_S = self.likelihood.variational_expectations_grads(fx_mus[idx], fx_covs[idx], self._observations)
        g_mu, g_S = -g_mu, -0.5 * (g_S + mT(g_S))
        n1 = N + 1
        gbar = np.zeros((n1, D))
        np.add.at(gbar, idx, g_mu)
        solve = lambda rhs: btd_solve(Ld, Ls, btd_solve(Ld, Ls, rhs, transpose=False), transpose=True)
        u = solve(gbar)
        Vd, Vs = np.zeros((n1, D, D)), np.zeros((N, D, D))
        for j, Gj in zip(idx, g_S):
            cols = []
            for c in range(D):
                rhs = np.zeros((n1, D))
                rhs[j, c] = 1.0
                cols.append(solve(rhs))
            Z = np.stack(cols, axis=-1)                                   # P^-1 E_j, [n1, D, D]
            Vd += Z @ Gj @ mT(Z)
            Vs += Z[1:] @ Gj @ mT(Z[:-1])
        # differential of the prior's naturals in (A_k, b_k)
        A, b = self.dist_p.state_transitions, self.dist_p.state_offsets
        cQ = self.dist_p.cholesky_process_covariances
        W = chol_solve(cQ, np.broadcast_to(np.eye(D), cQ.shape))
        grads = np.zeros(n_par)
        m = fx_mus
        for p in range(n_par):
            dAp, dbp = dA[..., p], db[..., p]
            dPd = np.zeros((n1, D, D))
            dPd[:-1] = mT(dAp) @ W @ A + mT(A) @ W @ dAp
            dPs = -W @ dAp
            dh = np.zeros((n1, D))
            dh[:-1] += -(mT(dAp) @ W @ b[..., None] + mT(A) @ W @ dbp[..., None])[..., 0]
            dh[1:] += (W @ dbp[..., None])[..., 0]
            dPm = (dPd @ m[..., None])[..., 0]
            dPm[1:] += (dPs @ m[:-1, :, None])[..., 0]
            dPm[:-1] += (mT(dPs) @ m[1:, :, None])[..., 0]
            grads[p] = (np.sum(u * dh) - np.sum(u * dPm) - np.sum(Vd * dPd) - 2.0 * np.sum(Vs * dPs))
        return grads

    # ---- cvi_dp_trainer.py:71-77 -- the unit the benchmark counts
    def site_update(self, data_lr: float, girsanov_lr: float):
        self.update_data_sites(data_lr)
        self.update_girsanov_sites(girsanov_lr)
        return self.classic_elbo()

    # ---- cvi_dp_trainer.py:110-134
    def relinearize(self):
        """Move the sites onto the un-clipped linearisation, re-linearise, move them again."""
        self.girsanov_sites = tranform_girsanov_sites(self.girsanov_sites, self.dist_p, self.dist_p_linearized)
        self.dist_p = self.dist_p_linearized
        dist_p_last = self.dist_p
        self.set_linearized_prior()
        self.girsanov_sites = tranform_girsanov_sites(self.girsanov_sites, dist_p_last, self.dist_p)


class VariationalMarkovGP:
    """vi_sde.py:63-482 (Archambeau et al. 2007), non-batched."""

    CLIP_MIN, CLIP_MAX = -5000.0, 5000.0

    def __init__(self, input_data, prior_sde: SDE, grid, likelihood: MultivariateGaussian,
                 prior_initial_state: Gaussian = None, stabilize_system: bool = False):
        self._time_points, self.observations = (np.asarray(input_data[0], dtype=np.float64),
                                                np.asarray(input_data[1], dtype=np.float64))
        self.prior_sde = prior_sde
        self.likelihood = likelihood
        self.state_dim = prior_sde.state_dim
        self.grid = np.asarray(grid, dtype=np.float64)
        self.num_states = self.grid.shape[0]
        self.num_transitions = self.num_states - 1
        self.dt = float(self.grid[1] - self.grid[0])
        D, N = self.state_dim, self.num_transitions
        self.A = np.zeros((N, D, D))
        self.b = np.zeros((N, D))
        if prior_initial_state is None:  # :88-94
            prior_initial_state = Gaussian(np.zeros(D), prior_sde.q * np.eye(D))
        self.prior_initial_state = prior_initial_state
        self.q_initial_mean = prior_initial_state.mu.copy()
        self.q_initial_chol = chol(prior_initial_state.cov)
        self.lambda_lagrange = np.zeros((N, D))
        self.psi_lagrange = np.broadcast_to(np.eye(D), (N, D, D)) * 1e-10
        self.dist_q_ssm = None
        self.stabilize_system = stabilize_system
        self.obs_indices = obs_indices(self.grid, self._time_points)

    # ---- :171-204
    @property
    def forward_pass(self):
        q = np.repeat(self.prior_sde.q.reshape(1, self.state_dim, self.state_dim), self.A.shape[0], axis=0)
        self.dist_q_ssm = LinearDrift(A=-1 * self.A, b=self.b).to_ssm(
            q=q, initial_mean=self.q_initial_mean, transition_times=self.grid,
            initial_chol_covariance=self.q_initial_chol)
        if self.stabilize_system:
            st = np.where(np.isnan(self.dist_q_ssm.state_transitions), 1e-8, self.dist_q_ssm.state_transitions)
            so = np.where(np.isnan(self.dist_q_ssm.state_offsets), 1e-8, self.dist_q_ssm.state_offsets)
            self.dist_q_ssm = StateSpaceModel(self.dist_q_ssm.initial_mean,
                                              self.dist_q_ssm.cholesky_initial_covariance,
                                              np.clip(st, -1.0, 1.0), np.clip(so, -1.0, 1.0),
                                              self.dist_q_ssm.cholesky_process_covariances)
        mm, mc = self.dist_q_ssm.marginals
        return mm[0], mc[0]

    # ---- :422-434
    def E_sde(self, m=None, S=None):
        if m is None or S is None:
            m, S = self.forward_pass
            m, S = m[:-1], S[:-1]
        return squared_drift_difference_along_Gaussian_path(self.prior_sde, LinearDrift(-1 * self.A, self.b),
                                                            Gaussian(m, S), self.dt)

    # ---- :206-240
    def _grad_E_sde(self, m, S):
        _, dm, dS = grad_squared_drift_difference(self.prior_sde, -1 * self.A, self.b, m[:-1], S[:-1], self.dt)
        return dm / self.dt, dS / self.dt

    # ---- :262-287
    def _jump_conditions(self, m, S):
        m_obs, S_obs = m[self.obs_indices], S[self.obs_indices]
        g_mu, g_S = self.likelihood.variational_expectations_grads(m_obs, S_obs, self.observations)
        d_obs_m = np.zeros_like(m)
        d_obs_S = np.zeros_like(S)
        np.add.at(d_obs_m, self.obs_indices, g_mu)
        np.add.at(d_obs_S, self.obs_indices, g_S)
        return d_obs_m, d_obs_S

    # ---- :289-347
    def update_lagrange(self, m, S):
        dEdm, dEdS = self._grad_E_sde(m, S)
        d_obs_m, d_obs_S = self._jump_conditions(m, S)
        if self.stabilize_system:
            fix = lambda x: np.clip(np.where(np.isnan(x), 1e-8, x), self.CLIP_MIN, self.CLIP_MAX)
            dEdm, dEdS, d_obs_m, d_obs_S = fix(dEdm), fix(dEdS), fix(d_obs_m), fix(d_obs_S)
        D, N = self.state_dim, self.num_transitions
        lam = np.zeros((N, D))
        psi = np.broadcast_to(np.eye(D), (N, D, D)) * 1e-10
        for t in range(N - 1, 0, -1):
            d_psi = psi[t] @ self.A[t] + psi[t] @ self.A[t] - dEdS[t]
            d_lambda = (self.A[t] @ lam[t][..., None])[..., 0] - dEdm[t]
            psi[t - 1] = psi[t] - self.dt * d_psi - d_obs_S[t]
            lam[t - 1] = lam[t] - self.dt * d_lambda - d_obs_m[t]
        self.lambda_lagrange, self.psi_lagrange = lam, psi

    # ---- :377-414
    def update_param(self, m, S, lr: float):
        m, S = m[:-1], S[:-1]
        if self.stabilize_system:
            fix = lambda x: np.clip(np.where(np.isnan(x), 1e-8, x), self.CLIP_MIN, self.CLIP_MAX)
            self.psi_lagrange, self.lambda_lagrange = fix(self.psi_lagrange), fix(self.lambda_lagrange)
        var = self.prior_sde.q
        egd = -self.prior_sde.expected_gradient_drift(m[None], S[None])[0]
        ed = self.prior_sde.expected_drift(m[None], S[None])[0]
        if self.state_dim == 1 and egd.ndim == 2:
            egd = egd[..., None]
        A_tilde = egd + 2.0 * var[None] @ self.psi_lagrange
        b_tilde = ed + (A_tilde @ m[..., None])[..., 0] - (var[None] @ self.lambda_lagrange[..., None])[..., 0]
        self.A = (1 - lr) * self.A + lr * A_tilde
        self.b = (1 - lr) * self.b + lr * b_tilde

    # ---- :242-260
    def update_initial_statistics(self, lr: float):
        Lp = chol(self.prior_initial_state.cov)
        q_mean = self.prior_initial_state.mu - ((Lp @ Lp.T) @ self.lambda_lagrange[0][..., None])[..., 0]
        prior_chol_cov_inv = chol_solve(Lp, np.eye(self.state_dim))
        prior_cov_inv = prior_chol_cov_inv @ prior_chol_cov_inv.T   # as written in the reference (:250-251)
        q_cov = np.linalg.inv(prior_cov_inv + 2 * self.psi_lagrange[0])
        self.q_initial_mean = (1 - lr) * self.q_initial_mean + lr * q_mean
        q_cov = (1 - lr) * (self.q_initial_chol @ self.q_initial_chol.T) + lr * q_cov
        self.q_initial_chol = chol(q_cov)

    def KL_initial_state(self):
        return gaussian_kl(self.q_initial_mean, self.q_initial_chol @ self.q_initial_chol.T,
                           self.prior_initial_state.mu, self.prior_initial_state.cov)

    # ---- :436-455
    def elbo(self, m=None, S=None):
        if m is None:
            m, S = self.forward_pass
        E_sde = self.E_sde()
        E_obs = np.sum(self.likelihood.variational_expectations(m[self.obs_indices], S[self.obs_indices],
                                                                self.observations))
        return float(E_obs - E_sde - self.KL_initial_state())

    # ---- :457-470
    def grad_prior_sde_params(self):
        """
        d E_sde / d(prior_sde parameters), every entry of ``prior_sde.params``, with (m, S) = forward_pass[1:]
        held constant -- the reference pairs the marginals at grid points 1..N with A[0..N-1] here (:461-465),
        unlike E_sde() itself, which uses grid points 0..N-1 (:428-430).
            d/dp 1/2 dt sum_t E||r||^2_{q^-1} = -dt sum_t E[r^T sym(q^-1) df/dp],   r = -A x + b - f(x).
        """
        m, S = self.forward_pass
        m, S = m[1:], S[1:]
        D, n = self.state_dim, self.A.shape[0]
        A, b = -1 * self.A, self.b
        sig_inv = chol_solve(chol(self.prior_sde.q), np.eye(D))
        W = 0.5 * (sig_inv + sig_inv.T)
        n_par = len(np.asarray(self.prior_sde.params))

        def func(x):
            x = x.reshape(-1, n, D)
            r = ((A[None] @ x[..., None])[..., 0] + b[None]) - self.prior_sde.drift(x)
            df, _ = self.prior_sde.param_grads(x)
            return (-np.einsum("pni,ij,kpnj->pnk", r, W, df)).reshape(-1, n_par)

        return self.dt * np.sum(mvnquad(func, m, S, H=20, Din=D, Dout=(n_par,)), axis=0)

    # ---- :472-482
    def grad_initial_state(self):
        """d KL[q(x0) || p(x0)] / d(prior loc, prior scale_tril), closed form of the tfp tape result."""
        mu_p, cov_p = self.prior_initial_state.mu, self.prior_initial_state.cov
        Lp = chol(cov_p)
        P_inv = chol_solve(Lp, np.eye(self.state_dim))
        Sq = self.q_initial_chol @ self.q_initial_chol.T
        dm = self.q_initial_mean - mu_p
        g_loc = -(P_inv @ dm)
        g_cov = 0.5 * (P_inv - P_inv @ (Sq + np.outer(dm, dm)) @ P_inv)
        g_tril = np.tril(2.0 * g_cov @ Lp)
        return g_loc, g_tril

    # ---- vi_markov_gp_trainer.py:62-70 -- one fixed-point sweep
    def sweep(self, lr: float, update_x0: bool = False, x0_lr: float = 0.1):
        m, S = self.forward_pass
        self.update_lagrange(m, S)
        self.update_param(m, S, lr)
        if update_x0:
            self.update_initial_statistics(x0_lr)
        m, S = self.forward_pass
        return self.elbo(m, S)


def ssm_natgrad_step(ssm: StateSpaceModel, dL_dssm, gamma: float):
    """
    ssm_natgrad.py:121-218, plain step (momentum=False): theta <- theta - gamma dL/d eta with
    dL/d eta = J^T dL/d(ssm params), J the Jacobian of expectations_to_ssm_params at the current
    expectation parameters (sde_np.expectations_to_ssm_params_vjp, checked against torch.autograd in
    tests/test_oracle_ssm.py).  ``dL_dssm`` = gradients with respect to (A, b, chol_P0, chol_Q, mu0).
    """
    from .sde_np import expectations_to_ssm_params_vjp
    etas = ssm_to_expectations(ssm)
    thetas = ssm_to_naturals(ssm)
    dL_dA, dL_db, dL_dP0, dL_dQ, dL_dmu0 = dL_dssm
    grads = expectations_to_ssm_params_vjp(*etas, dL_dA, dL_db, dL_dP0, dL_dQ, dL_dmu0)
    new_thetas = [t - gamma * g for t, g in zip(thetas, grads)]
    As, offsets, chol_P0, chol_Q, mu0 = naturals_to_ssm_params(*new_thetas)
    return StateSpaceModel(mu0, chol_P0, As, offsets, chol_Q)
