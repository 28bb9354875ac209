"""
Mirror of ``markovflow/state_space_model.py``: the linear-Gaussian state-space model
    x_0 ~ N(mu_0, P_0),   x_{k+1} = A_k x_k + b_k + N(0, Q_k)
parameterised by Cholesky factors.  Marginals come from ``vdp_ssm_marginals`` (forward affine recursion,
rebuilt precision, banded Cholesky, inverse band -- the reference's route, state_space_model.py:231-262).
"""
from __future__ import annotations

import torch

from .. import ops
from ._util import as_dev, batched, unbatched
from .block_tri_diag import LowerTriangularBlockTriDiagonal, SymmetricBlockTriDiagonal


class StateSpaceModel:
    """state_space_model.py:35-664 (the part the diffusion-process models use)."""

    def __init__(self, initial_mean, chol_initial_covariance, state_transitions, state_offsets,
                 chol_process_covariances):
        self._mu_0 = as_dev(initial_mean)
        self._chol_P_0 = as_dev(chol_initial_covariance)
        self._A_s = as_dev(state_transitions)
        self._b_s = as_dev(state_offsets)
        self._chol_Q_s = as_dev(chol_process_covariances)

    # ---- shapes and parameters (state_space_model.py:126-229)
    @property
    def batch_shape(self):
        return tuple(self._A_s.shape[:-3])

    @property
    def state_dim(self) -> int:
        return self._A_s.shape[-1]

    @property
    def num_transitions(self) -> int:
        return self._A_s.shape[-3]

    @property
    def event_shape(self):
        return (self.num_transitions + 1, self.state_dim)

    @property
    def cholesky_process_covariances(self):
        return self._chol_Q_s

    @property
    def cholesky_initial_covariance(self):
        return self._chol_P_0

    @property
    def initial_covariance(self):
        return self._chol_P_0 @ self._chol_P_0.transpose(-1, -2)

    @property
    def concatenated_cholesky_process_covariance(self):
        return torch.cat([self._chol_P_0.unsqueeze(-3), self._chol_Q_s], dim=-3)

    @property
    def state_offsets(self):
        return self._b_s

    @property
    def initial_mean(self):
        return self._mu_0

    @property
    def concatenated_state_offsets(self):
        return torch.cat([self._mu_0.unsqueeze(-2), self._b_s], dim=-2)

    @property
    def state_transitions(self):
        return self._A_s

    def _b(self):
        A, had = batched(self._A_s, 3)
        return (A, batched(self._b_s, 2)[0], batched(self._chol_Q_s, 3)[0], batched(self._mu_0, 1)[0],
                batched(self._chol_P_0, 2)[0], had)

    # ---- marginals (state_space_model.py:231-262, gauss_markov.py:107-117)
    @property
    def marginals(self):
        A, b, cQ, mu0, cP0, had = self._b()
        status = ops.new_status(A.device)
        m, Sd = ops.ssm_marginals(A, b, cQ, mu0, cP0, status=status)
        code, idx = ops.read_status(status)
        if code != 0:
            raise ValueError(f"StateSpaceModel.marginals: precision block {idx} is not positive definite")
        return unbatched(m, had), unbatched(Sd, had)

    @property
    def marginal_means(self):
        return self.marginals[0]

    @property
    def marginal_covariances(self):
        return self.marginals[1]

    def subsequent_covariances(self, marginal_covariances):
        """Cov(x_{k+1}, x_k) = A_k Sigma_kk   (state_space_model.py:326-341)."""
        return self._A_s @ marginal_covariances[..., :-1, :, :]

    def covariance_blocks(self):
        """state_space_model.py:264-275"""
        S = self.marginal_covariances
        return S, self.subsequent_covariances(S)

    # ---- precision and the A^-1 block (state_space_model.py:277-296, 431-483)
    @property
    def precision(self) -> SymmetricBlockTriDiagonal:
        return self._build_precision()

    def _build_precision(self) -> SymmetricBlockTriDiagonal:
        """K = -2 theta_diag, sub-diagonal -theta_subdiag of ssm_to_naturals: the same blocks as :431-483."""
        A, b, cQ, mu0, cP0, had = self._b()
        _, thd, ths = ops.ssm_to_naturals(A, b, cQ, mu0, cP0)
        return SymmetricBlockTriDiagonal(unbatched(-2.0 * thd, had), unbatched(-1.0 * ths, had))

    @property
    def a_inv_block(self) -> LowerTriangularBlockTriDiagonal:
        """[[I], [-A_0, I], ...]   (state_space_model.py:277-296)"""
        D, n = self.state_dim, self.num_transitions + 1
        eye = torch.eye(D, dtype=self._A_s.dtype, device=self._A_s.device)
        diag = eye.expand(self.batch_shape + (n, D, D)).contiguous()
        return LowerTriangularBlockTriDiagonal(diag, (-1.0 * self._A_s).contiguous())

    def log_det_precision(self):
        """-2 sum log diag of the Cholesky factors   (state_space_model.py:343-373)"""
        c = self.concatenated_cholesky_process_covariance
        return -2.0 * torch.log(torch.abs(torch.diagonal(c, dim1=-2, dim2=-1))).sum(dim=(-1, -2))

    # ---- densities (state_space_model.py:298-324, 485-593)
    def _ssm(self):
        A, b, cQ, mu0, cP0, had = self._b()
        return (A.contiguous(), b.contiguous(), cQ.contiguous(), mu0.contiguous(), cP0.contiguous()), had

    def kl_divergence(self, dist: "StateSpaceModel"):
        """KL(self || dist) (state_space_model.py:528-593) -> tensor of shape batch_shape."""
        assert self.event_shape == dist.event_shape and self.batch_shape == dist.batch_shape   # check_compatible
        s1, had = self._ssm()
        s2, _ = dist._ssm()
        status = ops.new_status(s1[0].device)
        kl = ops.ssm_kl_divergence(s1, s2, status=status)
        code, idx = ops.read_status(status)
        if code != 0:
            raise ValueError(f"StateSpaceModel.kl_divergence: precision block {idx} is not positive definite")
        return unbatched(kl, had)

    def log_pdf(self, states):
        """log p(states) (state_space_model.py:485-526); states: sample_shape + batch_shape + event_shape."""
        s, had = self._ssm()
        x = as_dev(states)
        sample_shape = tuple(x.shape[:x.dim() - 2 - len(self.batch_shape)])
        B = s[0].shape[0]
        x = x.reshape((-1, B) + tuple(self.event_shape)).contiguous()
        out = ops.ssm_log_pdf(*s, x)
        return out.reshape(sample_shape + tuple(self.batch_shape))

    def sample(self, sample_shape, seed: int = 0, eps=None):
        """Sample trajectories (state_space_model.py:298-324) -> sample_shape + batch_shape + event_shape.
        ``eps`` (same shape) replaces the device generator's standard normals."""
        s, had = self._ssm()
        shape = (sample_shape,) if isinstance(sample_shape, int) else tuple(sample_shape)
        S = 1
        for v in shape:
            S *= int(v)
        B = s[0].shape[0]
        full = shape + tuple(self.batch_shape) + tuple(self.event_shape)
        if S == 0:
            return torch.empty(full, dtype=s[0].dtype, device=s[0].device)
        e = None if eps is None else as_dev(eps).reshape((S, B) + tuple(self.event_shape)).contiguous()
        return ops.ssm_sample(*s, S, eps=e, seed=seed).reshape(full)

    def create_non_trainable_copy(self) -> "StateSpaceModel":
        return StateSpaceModel(self._mu_0.clone(), self._chol_P_0.clone(), self._A_s.clone(), self._b_s.clone(),
                               self._chol_Q_s.clone())

    create_trainable_copy = create_non_trainable_copy
