"""
Test helper: the taped forward pass of SDE_SSM_KL_with_grads_wrt_exp_params
(/root/reference/markovflow/sde/sde_utils.py:496-545) restated op by op in torch float64 so that
torch.autograd plays the role of tf.GradientTape.  Used only to check the oracle's hand-written adjoint.
"""
import itertools
import math

import numpy as np
import torch


def _drift_torch(sde, x):
    from oracle import sde_np as S
    if sde.drift_id == S.DRIFT_OU or sde.drift_id == S.DRIFT_OU_ND:
        return -sde.decay * x
    if sde.drift_id == S.DRIFT_DW:
        return sde.scale * x * (sde.c - x * x)
    if sde.drift_id == S.DRIFT_BENES:
        return sde.theta * torch.tanh(x)
    if sde.drift_id == S.DRIFT_SINE:
        return torch.sin(x - sde.theta)
    if sde.drift_id == S.DRIFT_SQRT:
        return torch.sqrt(sde.theta * torch.abs(x))
    if sde.drift_id == S.DRIFT_VDP:
        dx1 = sde.a * (x[..., 0] - (1 / 3) * x[..., 0] ** 3 - x[..., 1])
        dx2 = (1 / sde.a) * x[..., 0]
        return sde.tau * torch.stack([dx1, dx2], -1)
    if sde.drift_id == S.DRIFT_LORENZ63:
        X, Y, Z = x[..., 0], x[..., 1], x[..., 2]
        return torch.stack([sde.sigma * (Y - X), X * (sde.rho - Z) - Y, X * Y - sde.beta * Z], -1)
    if sde.drift_id == S.DRIFT_MLP131:
        w = torch.tensor(sde.weights)
        h = torch.relu(x[..., None] * w[0:3] + w[3:6])
        return h @ w[6:9] + w[9]
    raise NotImplementedError


def kl_and_grads_torch(exp1, exp_diag, exp_sub, sde, dt, prior_mu, prior_cov, time_grid, H=20):
    e1 = torch.tensor(exp1, dtype=torch.float64, requires_grad=True)
    ed = torch.tensor(exp_diag, dtype=torch.float64, requires_grad=True)
    es = torch.tensor(exp_sub, dtype=torch.float64, requires_grad=True)
    D = e1.shape[-1]
    el = e1[..., None]
    mcov = ed - el @ el.transpose(-1, -2)
    csub = es.transpose(-1, -2) - el[:-1] @ el[1:].transpose(-1, -2)
    mch = torch.linalg.cholesky(mcov)
    As = torch.cholesky_solve(csub, mch[:-1]).transpose(-1, -2)
    mu0 = el[0, :, 0]
    cholP0 = mch[0]
    b = (el[1:] - As @ el[:-1])[..., 0]
    ccov = mcov[1:] - As @ (mcov[:-1] @ As.transpose(-1, -2))
    cholQ = torch.linalg.cholesky(ccov)
    covar = ed - el @ el.transpose(-1, -2)
    Qq = cholQ @ cholQ.transpose(-1, -2)
    tg = torch.tensor(time_grid, dtype=torch.float64)
    deltas = (tg[1:] - tg[:-1])[:, None, None]
    Qp = (deltas * torch.tensor(sde.q)[None]).detach()
    I = torch.eye(D, dtype=torch.float64).expand(Qp.shape)
    Qpinv = torch.cholesky_solve(I, torch.linalg.cholesky(Qp))
    C = -(torch.logdet(Qq) - torch.logdet(Qp)) - D + torch.diagonal(Qpinv * Qq, dim1=-2, dim2=-1).sum(-1)
    gx, gw = np.polynomial.hermite.hermgauss(H)
    xn = torch.tensor(np.array(list(itertools.product(*(gx,) * D))))
    wn = torch.tensor(np.prod(np.array(list(itertools.product(*(gw,) * D))), 1)) * math.pi ** (-D * 0.5)
    m = e1[:-1]
    L = torch.linalg.cholesky(covar[:-1])
    X = 2.0 ** 0.5 * torch.einsum("nij,pj->pni", L, xn) + m[None]
    fq = torch.einsum("nij,pnj->pni", As, X) + b[None]
    fp = X + dt * _drift_torch(sde, X)
    diff = fp - fq
    val = torch.einsum("pni,nij,pnj->pn", diff, Qpinv, diff)
    quad = (val * wn[:, None]).sum(0)
    kl_path = 0.5 * (quad + C).sum()
    P0 = cholP0 @ cholP0.T
    pm = torch.tensor(prior_mu, dtype=torch.float64)
    pc = torch.tensor(prior_cov, dtype=torch.float64)
    q0 = torch.distributions.MultivariateNormal(mu0, covariance_matrix=P0)
    p0 = torch.distributions.MultivariateNormal(pm, covariance_matrix=pc)
    kl = torch.distributions.kl_divergence(q0, p0) + kl_path
    g = torch.autograd.grad(kl, [e1, ed, es])
    return kl.item(), [t.numpy() for t in g]
