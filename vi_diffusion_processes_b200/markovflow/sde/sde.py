"""
Mirror of ``markovflow/sde/sde.py``: SDE base class and the drifts of the paper.  A drift is a (drift_id,
parameter block) pair understood by the CUDA kernels (csrc/drifts.cuh); ``expected_drift`` and
``expected_gradient_drift`` are the Gauss-Hermite expectations of sde.py:92-131 (H = 10) computed by
``vdp_gh_expect``.  ``drift`` / ``gradient_drift`` on explicit points are the closed forms, evaluated with
torch element-wise ops on the caller's tensor (they are not on the hot path; the reference uses them for
simulation and plotting only).
User-defined Python drifts are not supported (there is no CPU path to call back into).
"""
from __future__ import annotations

import numpy as np
import torch

from ... import ops
from .._util import as_dev, as_np

DRIFT_OU, DRIFT_DW, DRIFT_BENES, DRIFT_SINE, DRIFT_SQRT, DRIFT_VDP, DRIFT_MLP131, DRIFT_LORENZ63, DRIFT_OU_ND = range(9)


class SDE:
    """sde.py:24-131"""

    drift_id = None
    full_jacobian = False   # True when gradient_drift returns the [.., D, D] Jacobian (batch_jacobian subclasses)

    param_names = ()        # attributes behind ``params``, in the order of the kernels' parameter block

    def __init__(self, state_dim=1):
        self._state_dim = state_dim
        if not hasattr(self, "_trainable"):
            self._trainable = [False] * len(self.param_names)

    @property
    def state_dim(self) -> int:
        return self._state_dim

    # ---- the reference's tf.Variable(trainable=...) parameters (sde.py:150, 197-198, 242, 286, 330, 451-452)
    @property
    def trainable_indices(self):
        """Positions in ``params`` of the parameters the reference would list in ``trainable_variables``."""
        return [i for i, t in enumerate(self._trainable) if t]

    @property
    def trainable_variables(self):
        p = list(self.params)
        return [p[i] for i in self.trainable_indices]

    def assign_trainable(self, values):
        """Write new values of the trainable parameters (what ``optimizer.apply_gradients`` does to the variables)."""
        for i, v in zip(self.trainable_indices, values):
            setattr(self, self.param_names[i], float(v))

    @property
    def params(self):
        raise NotImplementedError

    def drift(self, x, t=None):
        raise NotImplementedError

    def diffusion(self, x, t=None):
        """chol(q), constant in x and t for every SDE of the reference (sde.py:165-176)."""
        q = as_dev(self.q)
        return torch.linalg.cholesky(q)

    def gradient_drift(self, x, t=None):
        """sde.py:75-90: d(sum_i f_i)/dx -> same shape as x (full Jacobian for batch_jacobian subclasses)."""
        raise NotImplementedError

    def _expect(self, q_mean, q_covar):
        m = as_dev(q_mean)
        S = as_dev(q_covar)
        D = m.shape[-1]
        lead = m.shape[:-1]
        Ef, EJ = ops.gh_expect(self.drift_id, self.params, m.reshape(-1, D).contiguous(),
                               S.reshape(-1, D, D).contiguous(), H=10, full_jacobian=self.full_jacobian)
        return Ef.reshape(lead + (D,)), EJ.reshape(lead + ((D, D) if self.full_jacobian else (D,)))

    def expected_drift(self, q_mean, q_covar):
        """E_{N(q_mean, q_covar)}[f(x)], q_mean [n_batch, N, D], q_covar [n_batch, N, D, D]   (sde.py:92-112)."""
        return self._expect(q_mean, q_covar)[0]

    def expected_gradient_drift(self, q_mean, q_covar):
        """sde.py:114-131 (VanderPol: :501-518)"""
        return self._expect(q_mean, q_covar)[1]


def _q(q, D=None):
    q = np.atleast_2d(as_np(q)).astype(np.float64)
    return q


class OrnsteinUhlenbeckSDE(SDE):
    """f(x) = -decay x   (sde.py:134-176)"""
    drift_id = DRIFT_OU

    def __init__(self, decay: float = 1.0, q=np.ones((1, 1)), trainable: bool = False):
        self.q = _q(q)
        self._trainable = [bool(trainable)]
        super().__init__(self.q.shape[0])
        self.decay = float(decay)

    param_names = ("decay",)

    @property
    def params(self):
        return [self.decay]

    def drift(self, x, t=None):
        return -self.decay * x

    def gradient_drift(self, x, t=None):
        return torch.full_like(x, -self.decay)


class DoubleWellSDE(SDE):
    """f(x) = scale x (c - x^2)   (sde.py:179-224)"""
    drift_id = DRIFT_DW

    def __init__(self, q=np.ones((1, 1)), scale_trainable: bool = False, c_trainable: bool = False, scale=4.0, c=1.0):
        self.q = _q(q)
        self._trainable = [bool(scale_trainable), bool(c_trainable)]
        super().__init__(self.q.shape[0])
        self.scale, self.c = float(scale), float(c)

    param_names = ("scale", "c")

    @property
    def params(self):
        return [self.scale, self.c]

    def drift(self, x, t=None):
        return self.scale * x * (self.c - torch.square(x))

    def gradient_drift(self, x, t=None):
        return self.scale * (self.c - 3.0 * torch.square(x))


class BenesSDE(SDE):
    """f(x) = theta tanh(x)   (sde.py:227-268)"""
    drift_id = DRIFT_BENES

    def __init__(self, theta: float = 1, q=np.ones((1, 1)), trainable=False):
        self.q = _q(q)
        self._trainable = [bool(trainable)]
        super().__init__(self.q.shape[0])
        self.theta = float(theta)

    param_names = ("theta",)

    @property
    def params(self):
        return [self.theta]

    def drift(self, x, t=None):
        return self.theta * torch.tanh(x)

    def gradient_drift(self, x, t=None):
        return self.theta * (1.0 - torch.square(torch.tanh(x)))


class SineDiffusionSDE(SDE):
    """f(x) = sin(x - theta)   (sde.py:271-312)"""
    drift_id = DRIFT_SINE

    def __init__(self, theta: float = 0, q=np.ones((1, 1)), trainable=False):
        self.q = _q(q)
        self._trainable = [bool(trainable)]
        super().__init__(self.q.shape[0])
        self.theta = float(theta)

    param_names = ("theta",)

    @property
    def params(self):
        return [self.theta]

    def drift(self, x, t=None):
        return torch.sin(x - self.theta)

    def gradient_drift(self, x, t=None):
        return torch.cos(x - self.theta)


class SqrtDiffusionSDE(SDE):
    """f(x) = sqrt(theta |x|)   (sde.py:315-356)"""
    drift_id = DRIFT_SQRT

    def __init__(self, theta: float = 1, q=np.ones((1, 1)), trainable=False):
        self.q = _q(q)
        self._trainable = [bool(trainable)]
        super().__init__(self.q.shape[0])
        self.theta = float(theta)

    param_names = ("theta",)

    @property
    def params(self):
        return [self.theta]

    def drift(self, x, t=None):
        return torch.sqrt(self.theta * torch.abs(x))

    def gradient_drift(self, x, t=None):
        return 0.5 / torch.sqrt(self.theta * torch.abs(x)) * self.theta * torch.sign(x)


class MLPDrift(SDE):
    """1-3-1 relu MLP applied per coordinate (sde.py:359-430); evaluated in float64."""
    drift_id = DRIFT_MLP131

    def __init__(self, q=np.ones((1, 1)), weights=None, seed=0):
        self.q = _q(q)
        super().__init__(self.q.shape[0])
        if weights is None:
            weights = np.random.default_rng(seed).normal(size=10) * 0.5
        self.weights = [float(v) for v in np.asarray(weights).reshape(-1)]
        assert len(self.weights) == 10, "w1[3], b1[3], w2[3], b2"
        self._trainable = [True] * 10     # Keras layers: every weight is a trainable variable (sde.py:373-377)

    def assign_trainable(self, values):
        self.weights = [float(v) for v in values]

    @property
    def params(self):
        return self.weights

    def drift(self, x, t=None):
        w = torch.as_tensor(self.weights, dtype=x.dtype, device=x.device)
        pre = x.unsqueeze(-1) * w[0:3] + w[3:6]
        return (torch.relu(pre) * w[6:9]).sum(-1) + w[9]

    def gradient_drift(self, x, t=None):
        w = torch.as_tensor(self.weights, dtype=x.dtype, device=x.device)
        pre = x.unsqueeze(-1) * w[0:3] + w[3:6]
        return ((pre > 0) * (w[0:3] * w[6:9])).sum(-1)


class VanderPolOscillatorSDE(SDE):
    """f(x) = tau [a (x1 - x1^3/3 - x2), x1 / a], D = 2   (sde.py:432-518)"""
    drift_id = DRIFT_VDP
    full_jacobian = True

    def __init__(self, a: float = 1.0, tau: float = 1.0, q=np.eye(2), trainable=False):
        self.q = _q(q)
        self._trainable = [bool(trainable), bool(trainable)]
        super().__init__(2)
        self.a, self.tau = float(a), float(tau)

    param_names = ("a", "tau")

    @property
    def params(self):
        return [self.a, self.tau]

    def drift(self, x, t=None):
        x1, x2 = x[..., 0], x[..., 1]
        return torch.stack([self.tau * (self.a * (x1 - (1.0 / 3.0) * x1 ** 3 - x2)), self.tau * ((1.0 / self.a) * x1)], -1)

    def gradient_drift(self, x, t=None):
        x1 = x[..., 0]
        z, o = torch.zeros_like(x1), torch.ones_like(x1)
        return torch.stack([torch.stack([self.tau * self.a * (1.0 - x1 * x1), -self.tau * self.a * o], -1),
                            torch.stack([self.tau / self.a * o, z], -1)], -2)


class Lorenz63SDE(SDE):
    """Extension (no 3-D SDE exists in the reference): Lorenz-63 drift, full Jacobian."""
    drift_id = DRIFT_LORENZ63
    full_jacobian = True

    param_names = ("sigma", "rho", "beta")

    def __init__(self, sigma=10.0, rho=28.0, beta=8.0 / 3.0, q=np.eye(3), trainable=False):
        self.q = _q(q)
        self._trainable = [bool(trainable)] * 3
        super().__init__(3)
        self.sigma, self.rho, self.beta = float(sigma), float(rho), float(beta)

    @property
    def params(self):
        return [self.sigma, self.rho, self.beta]

    def drift(self, x, t=None):
        X, Y, Z = x[..., 0], x[..., 1], x[..., 2]
        return torch.stack([self.sigma * (Y - X), X * (self.rho - Z) - Y, X * Y - self.beta * Z], -1)


class OUNdSDE(SDE):
    """Extension: D-dimensional OU with the full (diagonal) Jacobian returned, usable for D > 1."""
    drift_id = DRIFT_OU_ND
    full_jacobian = True

    param_names = ("decay",)

    def __init__(self, decay=0.5, q=np.eye(3), trainable=False):
        self.q = _q(q)
        self._trainable = [bool(trainable)]
        super().__init__(self.q.shape[0])
        self.decay = float(decay)

    @property
    def params(self):
        return [self.decay]

    def drift(self, x, t=None):
        return -self.decay * x
