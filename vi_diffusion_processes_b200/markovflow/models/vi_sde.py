"""
Mirror of ``markovflow/models/vi_sde.py``: the VDP model of Archambeau et al. (2007), ``VariationalMarkovGP``
(vi_sde.py:63-482).  Same constructor arguments, attributes and method names; the work is done by
:class:`vi_diffusion_processes_b200.vgp.VgpEngine` (``vdp_vgp_*`` entries of the C ABI).
"""
from __future__ import annotations

import numpy as np

from ...vgp import VgpEngine
from .._util import as_dev, as_np
from ..likelihoods import MultivariateGaussian
from ..sde.sde import SDE
from ..sde.sde_utils import Gaussian
from ..state_space_model import StateSpaceModel


class VariationalMarkovGP:
    """vi_sde.py:63-482 (non-batched, like the reference)."""

    def __init__(self, input_data, prior_sde: SDE, grid, likelihood: MultivariateGaussian,
                 prior_initial_state: Gaussian = None, stabilize_system: bool = False):
        self._time_points, self.observations = as_np(input_data[0]).reshape(-1), as_np(input_data[1])
        self.prior_sde = prior_sde
        self.likelihood = likelihood
        self.state_dim = D = prior_sde.state_dim
        self.grid = tg = as_np(grid).reshape(-1)
        self.num_states = tg.shape[0]
        self.num_transitions = self.num_states - 1
        self.dt = float(tg[1] - tg[0])
        if prior_initial_state is None:   # vi_sde.py:88-94
            prior_initial_state = Gaussian(np.zeros(D), prior_sde.q * np.eye(D))
        self.prior_initial_state = prior_initial_state
        self.stabilize_system = stabilize_system
        self.obs_indices = np.where(np.equal(tg[..., None], self._time_points))[0]
        assert self.obs_indices.shape[0] == self.observations.shape[0], "every observation must sit on a grid point"
        self._engine = VgpEngine(prior_sde.drift_id, prior_sde.params, prior_sde.q, tg, self.obs_indices,
                              self.observations.reshape(-1, D), likelihood.chol_covariance,
                              as_np(prior_initial_state.mu), as_np(prior_initial_state.cov),
                              stabilize_system=stabilize_system)

    # ---- variational parameters, in the reference's names (assignable, as the reference's tf.Variables)
    def _get(self, name):
        return getattr(self._eng, name)[0]

    def _set(self, name, value):
        cur = getattr(self._eng, name)
        setattr(self._eng, name, as_dev(value).reshape(cur.shape).contiguous())

    A = property(lambda s: s._get("A"), lambda s, v: s._set("A", v))
    b = property(lambda s: s._get("b"), lambda s, v: s._set("b", v))
    lambda_lagrange = property(lambda s: s._get("lambda_lagrange"), lambda s, v: s._set("lambda_lagrange", v))
    psi_lagrange = property(lambda s: s._get("psi_lagrange"), lambda s, v: s._set("psi_lagrange", v))
    q_initial_mean = property(lambda s: s._get("q_initial_mean"), lambda s, v: s._set("q_initial_mean", v))
    q_initial_chol = property(lambda s: s._get("q_initial_chol"), lambda s, v: s._set("q_initial_chol", v))

    m = property(lambda s: s._eng.m[0])     # marginals of the last forward pass
    S = property(lambda s: s._eng.S[0])

    @property
    def forward_pass(self):
        """vi_sde.py:171-204 -> (m [N+1, D], S [N+1, D, D])"""
        m, S = self._eng.forward_pass()
        return m[0], S[0]

    def update_lagrange(self, m=None, S=None):
        """vi_sde.py:289-347.  ``m, S`` are the marginals of the last forward pass (kept on the device)."""
        self._eng.update_lagrange()

    def update_param(self, m=None, S=None, lr: float = 0.1):
        """vi_sde.py:377-414"""
        self._eng.update_param(lr)

    def update_initial_statistics(self, lr: float):
        """vi_sde.py:242-260"""
        self._eng.update_initial_statistics(lr)

    def elbo(self, m=None, S=None):
        """vi_sde.py:436-455"""
        if m is None:
            self._eng.forward_pass()
        return self._eng.elbo()[0]

    # ---- prior-parameter learning (vi_markov_gp_trainer.py:163-235)
    @property
    def _eng(self) -> VgpEngine:
        """The engine with the drift parameters of ``prior_sde`` (prior learning changes them between calls)."""
        self._engine.drift_params = [float(v) for v in self.prior_sde.params]
        return self._engine

    def grad_prior_sde_params(self):
        """vi_sde.py:457-470 -> one gradient of E_sde per ``prior_sde.trainable_variables`` entry."""
        g = self._eng.grad_prior_sde_params()[0]
        return [g[i] for i in self.prior_sde.trainable_indices]

    def KL_initial_state(self):
        """vi_sde.py:416-420: KL[q(x0) || p(x0)] (D x D algebra on the host side of the device tensors)."""
        import torch
        q0 = torch.distributions.MultivariateNormal(self.q_initial_mean, scale_tril=self.q_initial_chol)
        p0 = torch.distributions.MultivariateNormal(as_dev(self.prior_initial_state.mu),
                                                    covariance_matrix=as_dev(self.prior_initial_state.cov))
        return torch.distributions.kl_divergence(q0, p0)

    def grad_initial_state(self):
        """vi_sde.py:472-482: d KL[q(x0) || p(x0)] / d(prior loc, prior scale_tril)."""
        import torch
        loc = as_dev(self.prior_initial_state.mu).clone().requires_grad_(True)
        tril = torch.linalg.cholesky(as_dev(self.prior_initial_state.cov)).clone().requires_grad_(True)
        q0 = torch.distributions.MultivariateNormal(self.q_initial_mean, scale_tril=self.q_initial_chol)
        p0 = torch.distributions.MultivariateNormal(loc, scale_tril=torch.tril(tril))
        g = torch.autograd.grad(torch.distributions.kl_divergence(q0, p0), [loc, tril])
        return [g[0], torch.tril(g[1])]

    def set_prior_initial_state(self, mu, cov):
        """Assign p(x0) (vi_markov_gp_trainer.py:217-218)."""
        self.prior_initial_state = Gaussian(as_np(mu), as_np(cov))
        self._eng.x0_mu = as_np(mu).reshape(-1).astype(np.float64)
        self._eng.x0_cov = np.atleast_2d(as_np(cov)).astype(np.float64)

    def sweep(self, lr: float, update_x0: bool = False, x0_lr: float = 0.1):
        """One iteration of VIMarkovGPTrainer.perform_inference (vi_markov_gp_trainer.py:62-73), fused."""
        return self._eng.sweep(lr, update_x0=update_x0, x0_lr=x0_lr)[0]

    def check_status(self):
        """Raise if a kernel met a non positive-definite block (synchronises the stream)."""
        self._engine.check_status()

    @property
    def dist_q_ssm(self) -> StateSpaceModel:
        """The SSM of the last forward pass (vi_sde.py:180-197)."""
        from ..sde.drift import LinearDrift
        return LinearDrift(-1.0 * self.A, self.b).to_ssm(self.prior_sde.q, self.grid, self.q_initial_mean,
                                                         self.q_initial_chol)
