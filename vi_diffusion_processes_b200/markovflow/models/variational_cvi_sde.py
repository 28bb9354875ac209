"""
Mirror of ``markovflow/models/variational_cvi_sde.py``: CVI-DP with a non-linear SDE prior (``CVISitesSDE``,
variational_cvi_sde.py:368-518, with the ``CVISitesSSM`` machinery it inherits, :49-366).  Same constructor
arguments, attributes and method names; the work is done by :class:`vi_diffusion_processes_b200.cvi.CviEngine`
(``vdp_cvi_site_update`` / ``vdp_cvi_elbo`` / ``vdp_cvi_kl_and_grad`` of the C ABI).
"""
from __future__ import annotations

from typing import Tuple

import numpy as np
import torch

from ... import _lib, ops
from ...cvi import CviEngine
from .._util import as_np
from ..gauss_markov import BTDGaussian
from ..likelihoods import MultivariateGaussian
from ..sde.sde import SDE
from ..sde.sde_utils import Gaussian
from ..state_space_model import StateSpaceModel


class GaussianSitesNat:
    """kalman_filter.py:382-436: container of the data sites' natural parameters."""

    def __init__(self, nat1, nat2):
        self.nat1, self.nat2 = nat1, nat2


class CVISitesSDE:
    """variational_cvi_sde.py:368-518.  Batching is not supported by the reference either (:395); use
    :class:`CviEngine` directly for B independent sequences."""

    def __init__(self, prior_sde: SDE, time_grid, input_data: Tuple, likelihood: MultivariateGaussian,
                 prior_initial_state: Gaussian = None, initial_posterior_path: Gaussian = None,
                 stabilize_ssm: bool = True, clip_state_transitions: Tuple = (-1.0, 1.0)) -> None:
        obs_t, obs_y = as_np(input_data[0]), as_np(input_data[1])
        self.state_dim = D = obs_y.shape[-1]
        self.prior_sde = prior_sde
        if prior_initial_state is None:   # :434-444
            prior_initial_state = Gaussian(np.zeros(D), prior_sde.q * np.ones((D, D)))
        self._prior_initial_state = prior_initial_state
        self.likelihood = likelihood
        tg = as_np(time_grid).reshape(-1)
        self.time_grid = tg
        self._observations_time_points, self._observations = obs_t, obs_y
        # :117  grid index of every observation
        self.obs_sites_indices = np.where(np.equal(tg[..., None], obs_t.reshape(-1)))[0]
        assert self.obs_sites_indices.shape[0] == obs_y.shape[0], "every observation must sit on a grid point"
        init_m = init_S = None
        if initial_posterior_path is not None:   # :125-133
            init_m, init_S = as_np(initial_posterior_path.mu), as_np(initial_posterior_path.cov)
        self.stabilize_ssm = stabilize_ssm
        self.clip_state_transitions = clip_state_transitions
        self._engine = CviEngine(prior_sde.drift_id, prior_sde.params, prior_sde.q, tg, self.obs_sites_indices, obs_y,
                              likelihood.chol_covariance, as_np(prior_initial_state.mu), as_np(prior_initial_state.cov),
                              stabilize_ssm=stabilize_ssm, clip_state_transitions=clip_state_transitions,
                              initial_mean=None if init_m is None else init_m[None],
                              initial_cov=None if init_S is None else init_S[None])

    @property
    def prior_initial_state(self) -> Gaussian:
        return self._prior_initial_state

    @prior_initial_state.setter
    def prior_initial_state(self, value: Gaussian):
        """Assignable as in the reference (cvi_dp_trainer.py:231-235)."""
        self._prior_initial_state = value
        self._engine.set_prior_initial_state(as_np(value.mu), as_np(value.cov))

    # ---- state, in the reference's names
    @property
    def time_points(self):
        return self._observations_time_points

    @property
    def observations(self):
        return self._observations

    @property
    def fx_mus(self):
        return self._eng.fx_mus[0]

    @property
    def fx_covs(self):
        return self._eng.fx_covs[0]

    @property
    def girsanov_sites(self) -> BTDGaussian:
        return BTDGaussian(self._eng.g1[0], self._eng.gd[0], self._eng.gs[0])

    @girsanov_sites.setter
    def girsanov_sites(self, sites: BTDGaussian):
        self._eng.g1 = sites.nat1.reshape(self._eng.g1.shape).contiguous()
        self._eng.gd = sites.nat2.block_diagonal.reshape(self._eng.gd.shape).contiguous()
        self._eng.gs = sites.nat2.block_sub_diagonal.reshape(self._eng.gs.shape).contiguous()

    @property
    def data_sites(self) -> GaussianSitesNat:
        return GaussianSitesNat(self._eng.d1[0], self._eng.d2[0])

    @property
    def dist_p(self) -> StateSpaceModel:
        e = self._eng
        return StateSpaceModel(e.pmu0[0], e.pcholP0[0], e.pA[0], e.pb[0], e.pcholQ[0])

    @property
    def dist_p_linearized(self) -> StateSpaceModel:
        e = self._eng
        return StateSpaceModel(e.pmu0[0], e.pcholP0[0], e.pA_lin[0], e.pb_lin[0], e.pcholQ[0])

    @property
    def _eng(self) -> CviEngine:
        """The engine with the drift parameters of ``prior_sde`` (prior learning changes them between calls)."""
        self._engine.drift_params = [float(v) for v in self.prior_sde.params]
        return self._engine

    def set_linearized_prior(self):
        """variational_cvi_sde.py:408-432"""
        self._eng.set_linearized_prior()

    @property
    def dist_q(self) -> StateSpaceModel:
        """variational_cvi_sde.py:176-192: rebuilt from the sites on every access, as in the reference."""
        A, b, cQ, mu0, cP0 = self._eng.posterior_ssm()
        return StateSpaceModel(mu0[0].clone(), cP0[0].clone(), A[0].clone(), b[0].clone(), cQ[0].clone())

    # ---- the three steps of the trainer's hot loop (cvi_dp_trainer.py:71-77)
    def update_data_sites(self, lr: float):
        """variational_cvi_sde.py:301-317"""
        self._eng.site_update(lr, 0.0, flags=_lib.FLAG_SKIP_GIRSANOV | _lib.FLAG_SKIP_ELBO)

    def update_girsanov_sites(self, lr: float):
        """variational_cvi_sde.py:279-299"""
        self._eng.site_update(0.0, lr, flags=_lib.FLAG_SKIP_DATA_UPDATE | _lib.FLAG_SKIP_ELBO)

    def classic_elbo(self):
        """variational_cvi_sde.py:339-352.  dist_q is rebuilt from the sites; the stored ``fx_mus`` / ``fx_covs`` keep
        the values of the last site update, as in the reference."""
        return self._eng.classic_elbo_keep_marginals()[0]

    # ---- prior-parameter learning (cvi_dp_trainer.py:207-250)
    def grad_KL_wrt_prior_params(self):
        """variational_cvi_sde.py:495-504 -> one gradient per ``prior_sde.trainable_variables`` entry."""
        g = self._eng.grad_kl_prior_params()[0]
        return [g[i] for i in self.prior_sde.trainable_indices]

    def grad_VE_wrt_prior_params(self):
        """variational_cvi_sde.py:506-518 (re-linearises the prior as a side effect, :512)."""
        g = self._eng.grad_ve_prior_params()[0]
        return [g[i] for i in self.prior_sde.trainable_indices]

    def site_update(self, data_lr: float, girsanov_lr: float):
        """The three calls above fused in one launch sequence; returns the ELBO."""
        return self._eng.site_update(data_lr, girsanov_lr)[0]

    def optimize_sites(self, max_itr: int, optim_tol: float, data_lr: float, girsanov_lr: float, test_indices=None):
        """``CVISitesTrainer._optimize_sites_under_stable_prior`` (cvi_dp_trainer.py:63-101) as one device-side
        loop (``vdp_cvi_optimize_sites``).  Returns (elbo list, test means [n, n_test, D], test covariances, data_lr,
        girsanov_lr) with n = the updates the loop took."""
        ti = None if test_indices is None else np.asarray(test_indices, dtype=np.int64)[None]
        r = self._eng.optimize_sites(max_itr, optim_tol, data_lr, girsanov_lr, test_idx=ti)
        n = int(r["n_itr"][0])
        elbo = [float(v) for v in r["elbo_trace"][1:n + 1, 0]]
        tm = r["test_mus"][:n, 0] if "test_mus" in r else None
        tc = r["test_covs"][:n, 0] if "test_covs" in r else None
        return elbo, tm, tc, float(r["lr_data"][0]), float(r["lr_girsanov"][0])

    def variational_expectation(self):
        """variational_cvi_sde.py:319-337 (of the current posterior)"""
        self.classic_elbo()
        return self._eng.ve[0]

    def KL_q_p(self):
        """variational_cvi_sde.py:446-486"""
        self.classic_elbo()
        return self._eng.kl[0]

    def grad_kl_wrt_exp_param(self):
        """variational_cvi_sde.py:488-493 -> [dKL/d eta1, dKL/d eta_diag, dKL/d eta_subdiag]"""
        e = self._eng
        q = self.dist_q
        m, S = q.marginals
        _, g = ops.cvi_kl_and_grad(e.drift_id, e.drift_params, e.q, e.dt, e.time_grid, e.x0_mu, e.x0_cov,
                                   q.state_transitions[None].contiguous(), m[None].contiguous(), S[None].contiguous())
        return [t[0] for t in g]

    def relinearize(self):
        """cvi_dp_trainer.py:110-134: tranform_girsanov_sites around set_linearized_prior."""
        self._eng.relinearize()

    def transform_sites_to_linearized_prior(self):
        """cvi_dp_trainer.py:110-113: ``tranform_girsanov_sites(sites, dist_p, dist_p_linearized)``; dist_p <- p_lin."""
        self._eng.to_linearized_prior()

    def relinearize_and_transform_sites(self):
        """cvi_dp_trainer.py:128-134: ``set_linearized_prior`` + ``tranform_girsanov_sites(sites, p_last, dist_p)``."""
        self._eng.relinearize_new()

    def loss(self):
        return -self.classic_elbo()

    def check_status(self):
        """Raise if a kernel met a non positive-definite block since the model was built (the analogue of the
        InvalidArgumentError TensorFlow's Cholesky raises); synchronises the stream."""
        self._engine.check_status()
