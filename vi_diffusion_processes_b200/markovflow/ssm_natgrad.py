"""
Mirror of ``markovflow/ssm_natgrad.py``: natural-gradient optimiser over a state-space model,
theta <- theta - gamma dL/d eta (Salimbeni et al. 2018, eq. 10), with dL/d eta obtained by the chain rule
through ``expectations_to_ssm_params`` (ssm_natgrad.py:142-172) -- here one CUDA kernel
(``k_expectations_to_ssm_vjp``) between the marginals sweep and ``naturals_to_ssm_params``.

The Euclidean gradient dL/d(ssm parameters) comes from torch autograd on the caller's ``loss_fn`` (the
reference uses a tf.GradientTape the same way); the plain step (``momentum=False``) is implemented.
"""
from __future__ import annotations

from typing import Callable

import torch

from .. import ops
from ._util import batched
from .state_space_model import StateSpaceModel


class SSMNaturalGradient:
    """ssm_natgrad.py:33-244"""

    def __init__(self, gamma: float = 0.1, momentum: bool = False, beta1: float = 0.9, beta2: float = 0.99,
                 epsilon: float = 1e-08, name=None) -> None:
        if momentum:
            raise NotImplementedError("the momentum variant needs the adjoint of naturals_to_ssm_params "
                                      "(ssm_natgrad.py:163-172); only momentum=False is built")
        self.gamma = gamma
        self._momentum = momentum

    def minimize(self, loss_fn: Callable, ssm: StateSpaceModel) -> None:
        """``loss_fn()`` must compute the loss from ``ssm``'s parameter tensors with torch ops (so that autograd
        can give dL/d(A, b, chol_P0, chol_Q, mu0)); the parameters of ``ssm`` are then updated in place."""
        params = [ssm._A_s, ssm._b_s, ssm._chol_P_0, ssm._chol_Q_s, ssm._mu_0]
        leaves = [p.detach().requires_grad_(True) for p in params]
        ssm._A_s, ssm._b_s, ssm._chol_P_0, ssm._chol_Q_s, ssm._mu_0 = leaves
        try:
            loss = loss_fn()
            grads = torch.autograd.grad(loss, leaves, allow_unused=True)
        finally:
            ssm._A_s, ssm._b_s, ssm._chol_P_0, ssm._chol_Q_s, ssm._mu_0 = params
        grads = [torch.zeros_like(p) if g is None else g for g, p in zip(grads, params)]
        self.apply_gradients(grads, ssm)

    def apply_gradients(self, dL_dssm, ssm: StateSpaceModel) -> None:
        """The step itself, given dL/d(A, b, chol_P0, chol_Q, mu0) (ssm_natgrad.py:157-218)."""
        gA, gb, gP0, gQ, gmu0 = [g.detach().contiguous() for g in dL_dssm]
        A, had = batched(ssm._A_s.detach(), 3)
        b, cP0 = batched(ssm._b_s.detach(), 2)[0], batched(ssm._chol_P_0.detach(), 2)[0]
        cQ, mu0 = batched(ssm._chol_Q_s.detach(), 3)[0], batched(ssm._mu_0.detach(), 1)[0]
        A, b, cP0, cQ, mu0 = [t.clone() for t in (A, b, cP0, cQ, mu0)]
        status = ops.new_status(A.device)
        ops.ssm_natgrad_step(A, b, cP0, cQ, mu0, batched(gA, 3)[0], batched(gb, 2)[0], batched(gP0, 2)[0],
                             batched(gQ, 3)[0], batched(gmu0, 1)[0], self.gamma, status=status)
        code, idx = ops.read_status(status)
        if code != 0:
            raise ValueError(f"natural-gradient step left the positive-definite cone at block {idx}; reduce gamma")
        shape = lambda t, ref: t.reshape(ref.shape)
        ssm._A_s, ssm._b_s = shape(A, ssm._A_s), shape(b, ssm._b_s)
        ssm._chol_P_0, ssm._chol_Q_s, ssm._mu_0 = shape(cP0, ssm._chol_P_0), shape(cQ, ssm._chol_Q_s), shape(mu0, ssm._mu_0)
