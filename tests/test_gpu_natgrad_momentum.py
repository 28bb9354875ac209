"""
SSMNaturalGradient(momentum=True)  (ssm_natgrad.py:147-208).

The Adam-like normaliser of the reference is the Fisher norm of the natural gradient,
    sum_c w_c <dL/d eta_c, dL/d theta_c>,   w = (1, 1, 2),
    dL/d theta = tape.gradient(naturals_to_ssm_params(theta), theta, output_gradients=dL_dssm)   (:172).
The CUDA path gets that scalar from ONE directional derivative of theta -> eta (vdp_btd_inverse_band_jvp + triangular
solves, markovflow/ssm_natgrad.py::naturals_to_expectations_jvp).  Checked here against torch.autograd on a dense
float64 restatement of the reference functions (CPU; the checker, as tests/torch_tape.py is for the KL gradient):

  * the directional derivative itself against torch.func.jvp of the dense theta -> eta map;
  * the Fisher norm against the reference's expression with both gradients from reverse mode;
  * three optimiser steps against the reference's update rule written out in torch.

The banded precision is read through its lower triangle (ssm_gaussian_transformations.py:443), which the dense
restatement does too; whether TensorFlow's registered gradients use the same convention for D > 1 cannot be checked
here (no TensorFlow): D = 1 has no such freedom.
"""
import math

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
F64 = torch.float64


def _mirror_lower(X):
    return torch.tril(X) + torch.tril(X, -1).transpose(-1, -2)


def _nat_to_exp_dense(th1, thd, ths):
    """theta -> eta through the dense precision (one sequence): K = (-2 thd, -ths), m = K^-1 th1."""
    n, D = th1.shape
    K = torch.zeros((n * D, n * D), dtype=F64)
    Kd = -2.0 * _mirror_lower(thd)
    for k in range(n):
        K[k * D:(k + 1) * D, k * D:(k + 1) * D] = Kd[k]
    for k in range(n - 1):
        K[(k + 1) * D:(k + 2) * D, k * D:(k + 1) * D] = -ths[k]
        K[k * D:(k + 1) * D, (k + 1) * D:(k + 2) * D] = -ths[k].T
    S = torch.linalg.inv(K)
    m = (S @ th1.reshape(-1)).reshape(n, D)
    ed = torch.stack([S[k * D:(k + 1) * D, k * D:(k + 1) * D] for k in range(n)]) + m[:, :, None] * m[:, None, :]
    es = torch.stack([S[(k + 1) * D:(k + 2) * D, k * D:(k + 1) * D] for k in range(n - 1)]) \
        + m[1:, :, None] * m[:-1, None, :]
    return m, ed, es


def _exp_to_ssm_torch(e1, ed, es):
    """expectations_to_ssm_params (ssm_gaussian_transformations.py:93-178) op by op; es[k] = E x_{k+1} x_k^T."""
    el = e1[..., None]
    mcov = ed - el @ el.transpose(-1, -2)
    csub = es.transpose(-1, -2) - el[:-1] @ el[1:].transpose(-1, -2)
    mch = torch.linalg.cholesky(mcov)
    As = torch.cholesky_solve(csub, mch[:-1]).transpose(-1, -2)
    b = (el[1:] - As @ el[:-1])[..., 0]
    ccov = mcov[1:] - As @ (mcov[:-1] @ As.transpose(-1, -2))
    return As, b, mch[0], torch.linalg.cholesky(ccov), el[0, :, 0]


def _rand_theta(rng, n, D):
    """naturals of a random stable chain (from the CUDA op, so both sides start from the same theta)"""
    from vi_diffusion_processes_b200 import ops
    A = 0.6 * np.eye(D) + 0.1 * rng.normal(size=(1, n - 1, D, D))
    b = 0.3 * rng.normal(size=(1, n - 1, D))
    cQ = np.tril(0.1 * rng.normal(size=(1, n - 1, D, D))) + 0.5 * np.eye(D)
    mu0 = rng.normal(size=(1, D))
    cP0 = (np.tril(0.1 * rng.normal(size=(1, D, D))) + 0.7 * np.eye(D))
    dev = lambda x: torch.as_tensor(x, dtype=F64, device="cuda").contiguous()
    ssm = [dev(x) for x in (A, b, cQ, mu0, cP0)]
    th = ops.ssm_to_naturals(*ssm)
    return ssm, th


@pytest.mark.parametrize("D", [1, 2, 3])
def test_directional_derivative_of_naturals_to_expectations(D):
    from vi_diffusion_processes_b200.markovflow.ssm_natgrad import naturals_to_expectations_jvp
    rng = np.random.default_rng(70 + D)
    n = 7
    _, th = _rand_theta(rng, n, D)
    w = [torch.as_tensor(rng.normal(size=tuple(t.shape)), dtype=F64, device="cuda") for t in th]
    got = naturals_to_expectations_jvp(*th, *w)
    tc = [t[0].cpu() for t in th]
    wc = [t[0].cpu() for t in w]
    _, want = torch.func.jvp(_nat_to_exp_dense, tuple(tc), tuple(wc))
    for g, r in zip(got, want):
        assert float((g[0].cpu() - r).abs().max()) <= 1e-10 * float(r.abs().max())


@pytest.mark.parametrize("D", [1, 2, 3])
def test_fisher_norm_equals_the_reference_expression(D):
    """sum_c w_c <dL/d eta_c, dL/d theta_c> with both factors from reverse mode (the reference's tape) against
    <g, J (W g)> with J (W g) from the CUDA directional derivative."""
    from vi_diffusion_processes_b200.markovflow.ssm_natgrad import naturals_to_expectations_jvp
    rng = np.random.default_rng(80 + D)
    n = 6
    _, th = _rand_theta(rng, n, D)
    tc = [t[0].cpu().clone().requires_grad_(True) for t in th]
    outs = _exp_to_ssm_torch(*_nat_to_exp_dense(*tc))                 # naturals_to_ssm_params
    cot = [torch.as_tensor(rng.normal(size=tuple(o.shape)), dtype=F64) for o in outs]
    cot[2], cot[3] = torch.tril(cot[2]), torch.tril(cot[3])
    gt = torch.autograd.grad(outs, tc, grad_outputs=cot)              # dL/d theta   (:172)
    eta = [e.detach().clone().requires_grad_(True) for e in _nat_to_exp_dense(*tc)]
    g = torch.autograd.grad(_exp_to_ssm_torch(*eta), eta, grad_outputs=cot)   # dL/d eta  (:168)
    want = float((g[0] * gt[0]).sum() + (g[1] * gt[1]).sum() + 2.0 * (g[2] * gt[2]).sum())
    gd = [x[None].cuda().contiguous() for x in g]
    d = naturals_to_expectations_jvp(*th, gd[0], gd[1], 2.0 * gd[2])
    got = float((gd[0] * d[0]).sum() + (gd[1] * d[1]).sum() + (gd[2] * d[2]).sum())
    assert abs(got - want) <= 1e-9 * abs(want), (got, want)
    assert want > 0.0                                                 # a Fisher norm


def _reference_momentum_steps(ssm, grads_fn, gamma, beta1, beta2, eps, steps):
    """ssm_natgrad.py:147-218 in torch float64 on the CPU, dense restatements, one sequence."""
    A, b, cQ, mu0, cP0 = [x[0].cpu() for x in ssm]
    ms, v, t = None, 0.0, 1.0
    for _ in range(steps):
        cot = grads_fn(A, b, cP0, cQ, mu0)
        # ssm -> theta through the dense joint: theta = (K m, -K_kk / 2, -K_{k+1,k})
        from vi_diffusion_processes_b200 import ops
        th = [x[0].cpu() for x in ops.ssm_to_naturals(*[x[None].cuda().contiguous() for x in (A, b, cQ, mu0, cP0)])]
        tc = [x.clone().requires_grad_(True) for x in th]
        outs = _exp_to_ssm_torch(*_nat_to_exp_dense(*tc))
        gt = torch.autograd.grad(outs, tc, grad_outputs=cot)
        eta = [e.detach().clone().requires_grad_(True) for e in _nat_to_exp_dense(*tc)]
        g = torch.autograd.grad(_exp_to_ssm_torch(*eta), eta, grad_outputs=cot)
        if ms is None:
            ms = [torch.zeros_like(x) for x in g]
        lr = gamma * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
        ms = [m * beta1 + (1.0 - beta1) * gi for m, gi in zip(ms, g)]
        norm = float((g[0] * gt[0]).sum() + (g[1] * gt[1]).sum() + 2.0 * (g[2] * gt[2]).sum())
        v = v * beta2 + (1.0 - beta2) * norm
        new = [x - lr * m / (math.sqrt(v) + eps) for x, m in zip(th, ms)]
        t += 1.0
        with torch.no_grad():
            A, b, cP0, cQ, mu0 = _exp_to_ssm_torch(*_nat_to_exp_dense(*new))
    return A, b, cP0, cQ, mu0


@pytest.mark.parametrize("D", [1, 2])
def test_three_momentum_steps_follow_the_reference_rule(D):
    from vi_diffusion_processes_b200.markovflow.ssm_natgrad import SSMNaturalGradient
    from vi_diffusion_processes_b200.markovflow.state_space_model import StateSpaceModel
    rng = np.random.default_rng(90 + D)
    n = 6
    ssm_t, _ = _rand_theta(rng, n, D)
    A, b, cQ, mu0, cP0 = ssm_t
    tA = torch.as_tensor(rng.normal(size=(n - 1, D, D)), dtype=F64)
    tb = torch.as_tensor(rng.normal(size=(n - 1, D)), dtype=F64)

    def grads_fn(A_, b_, cP0_, cQ_, mu0_):      # dL/d(A, b, chol_P0, chol_Q, mu0) of a fixed quadratic loss
        return [0.3 * (A_ - tA), 2.0 * (b_ - tb), 0.5 * torch.tril(cP0_), 0.2 * torch.tril(cQ_), 2.0 * mu0_]

    want = _reference_momentum_steps(ssm_t, grads_fn, 0.05, 0.9, 0.99, 1e-8, 3)
    model = StateSpaceModel(mu0[0].clone(), cP0[0].clone(), A[0].clone(), b[0].clone(), cQ[0].clone())
    opt = SSMNaturalGradient(gamma=0.05, momentum=True, beta1=0.9, beta2=0.99, epsilon=1e-8)
    for _ in range(3):
        cur = [model.state_transitions, model.state_offsets, model.cholesky_initial_covariance,
               model.cholesky_process_covariances, model.initial_mean]
        g = grads_fn(*[x.detach().cpu() for x in cur])
        opt.apply_gradients([x.cuda() for x in g], model)
    got = [model.state_transitions, model.state_offsets, model.cholesky_initial_covariance,
           model.cholesky_process_covariances, model.initial_mean]
    tol = 1e-9 if D == 1 else 1e-8
    for x, r in zip(got, want):
        assert float((x.detach().cpu() - r).abs().max()) <= tol * max(float(r.abs().max()), 1.0)
    assert opt._effective_lr is not None and opt._step_counter == 4.0
