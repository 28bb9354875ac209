"""
Mirror of ``markovflow/ssm_natgrad.py``: natural-gradient optimiser over a state-space model,
theta <- theta - gamma dL/d eta (Salimbeni et al. 2018, eq. 10), with dL/d eta obtained by the chain rule
through ``expectations_to_ssm_params`` (ssm_natgrad.py:142-172) -- here one CUDA kernel
(``k_expectations_to_ssm_vjp``) between the marginals sweep and ``naturals_to_ssm_params``.

The Euclidean gradient dL/d(ssm parameters) comes from torch autograd on the caller's ``loss_fn`` (the
reference uses a tf.GradientTape the same way).  ``momentum=False`` is one fused C call (``vdp_ssm_natgrad_step``);
``momentum=True`` (ssm_natgrad.py:163-208) is composed from the block-tridiagonal ops: the moving averages live in
expectation-parameter space and the Adam-like normaliser is the Fisher norm of the natural gradient,
sum_c w_c <dL/d eta_c, dL/d theta_c> with w = (1, 1, 2) and dL/d theta = tape.gradient(naturals_to_ssm_params(theta),
theta, output_gradients=dL_dssm) = (d eta / d theta)^T dL/d eta.  That scalar is <g, J (W g)> with J = d eta / d theta,
i.e. ONE directional derivative of theta -> eta along W g (``naturals_to_expectations_jvp``: tangent of the Cholesky
factor and of its inverse band, ``vdp_btd_inverse_band_jvp``, plus two triangular solves for the mean), so no
reverse-mode pass through naturals_to_ssm_params is needed.
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
        self.gamma = gamma
        self._momentum = momentum
        self._beta1, self._beta2, self._epsilon = beta1, beta2, epsilon
        self._ms = None            # moving averages of dL/d eta (ssm_natgrad.py:89, :104-108)
        self._v = 0.0              # moving average of the Fisher norm (:90)
        self._step_counter = 1.0   # (:91)
        self._effective_lr = None

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
        if self._momentum:
            A, b, cP0, cQ, mu0 = self._momentum_step(A, b, cP0, cQ, mu0, batched(gA, 3)[0], batched(gb, 2)[0],
                                                     batched(gP0, 2)[0], batched(gQ, 3)[0], batched(gmu0, 1)[0], status)
        else:
            ops.ssm_natgrad_step(A, b, cP0, cQ, mu0, batched(gA, 3)[0], batched(gb, 2)[0], batched(gP0, 2)[0],
                                 batched(gQ, 3)[0], batched(gmu0, 1)[0], self.gamma, status=status)
        code, idx = ops.read_status(status)
        if code != 0:
            raise ValueError(f"natural-gradient step left the positive-definite cone at block {idx}; reduce gamma")
        shape = lambda t, ref: t.reshape(ref.shape)
        ssm._A_s, ssm._b_s = shape(A, ssm._A_s), shape(b, ssm._b_s)
        ssm._chol_P_0, ssm._chol_Q_s, ssm._mu_0 = shape(cP0, ssm._chol_P_0), shape(cQ, ssm._chol_Q_s), shape(mu0, ssm._mu_0)

    # ---- momentum=True ------------------------------------------------------------------------------------
    def _momentum_step(self, A, b, cP0, cQ, mu0, gA, gb, gP0, gQ, gmu0, status):
        """ssm_natgrad.py:147-208 on batched tensors [B, ...]; the moving averages are shared by the batch entries'
        own slots (one optimiser per model, as in the reference) and the Fisher norm is summed over the batch, which
        is what ``tf.reduce_sum`` over the leading dimensions does."""
        import math
        th1, thd, ths = ops.ssm_to_naturals(A, b, cQ, mu0, cP0)
        m, Sd = ops.ssm_marginals(A, b, cQ, mu0, cP0, status=status)
        e1, ed, es = ops.ssm_to_expectations(A, m, Sd)
        g = ops.expectations_to_ssm_vjp(e1, ed, es, gA, gb, gP0, gQ, gmu0, status=status)      # dL/d eta  (:168)
        if self._ms is None:
            self._ms = [torch.zeros_like(x) for x in g]
        t = self._step_counter
        lr = self.gamma * math.sqrt(1.0 - self._beta2 ** t) / (1.0 - self._beta1 ** t)        # (:180-184)
        ms_new = [mm_ * self._beta1 + (1.0 - self._beta1) * gi for mm_, gi in zip(self._ms, g)]
        # Fisher norm (:190-192): sum_c w_c <g_c, (J^T g)_c> = <g, J (W g)>, W = diag(1, 1, 2)
        d1, dd, ds = naturals_to_expectations_jvp(th1, thd, ths, g[0], g[1], 2.0 * g[2], status=status)
        norm = float((g[0] * d1).sum() + (g[1] * dd).sum() + (g[2] * ds).sum())
        v_new = self._v * self._beta2 + (1.0 - self._beta2) * norm
        eff = lr / (math.sqrt(v_new) + self._epsilon)
        new = [th - eff * mm_ for th, mm_ in zip((th1, thd, ths), ms_new)]
        self._ms, self._v = ms_new, v_new
        self._step_counter += 1.0
        self._effective_lr = eff
        A2, b2, cP02, cQ2, mu02 = ops.naturals_to_ssm(*[x.contiguous() for x in new], status=status)
        return A2, b2, cP02, cQ2, mu02


def _mirror_lower(X):
    """the symmetric matrix a banded representation stores: lower triangle, mirrored"""
    L = torch.tril(X)
    return L + torch.tril(X, -1).transpose(-1, -2)


def naturals_to_expectations_jvp(th1, thd, ths, w1, wd, ws, status=None):
    """d/d eps of (eta_1, eta_diag, eta_subdiag)(theta + eps w) at eps = 0, theta = (th1 [B,n,D], thd [B,n,D,D],
    ths [B,n-1,D,D]).  Precision K = (-2 thd, -ths) (its lower band is what the reference factorises,
    ssm_gaussian_transformations.py:443), mean K m = th1, eta = (m, Sigma_kk + m m^T, Sigma_{k+1,k} + m_{k+1} m_k^T)."""
    Kd, Ks = (-2.0 * thd).contiguous(), (-1.0 * ths).contiguous()
    Ld, Ls = ops.btd_cholesky(Kd, Ks, status=status)
    m = ops.btd_solve(Ld, Ls, ops.btd_solve(Ld, Ls, th1.contiguous(), transpose=False), transpose=True)
    dKd, dKs = (-2.0 * _mirror_lower(wd)).contiguous(), (-1.0 * ws).contiguous()
    dSd, dSs = ops.btd_inverse_band_jvp(Ld, Ls, dKd, dKs)
    rhs = (w1 - ops.btd_matvec(dKd, dKs, m, symmetric=True)).contiguous()                # K dm = w1 - dK m
    dm = ops.btd_solve(Ld, Ls, ops.btd_solve(Ld, Ls, rhs, transpose=False), transpose=True)
    outer = lambda a, c: a.unsqueeze(-1) * c.unsqueeze(-2)
    d1 = dm
    dd = dSd + outer(dm, m) + outer(m, dm)
    ds = dSs + outer(dm[:, 1:], m[:, :-1]) + outer(m[:, 1:], dm[:, :-1])
    return d1, dd, ds
