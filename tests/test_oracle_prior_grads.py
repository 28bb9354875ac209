"""
Prior-parameter learning: the oracle's closed-form restatements of
  CVISitesSDE.grad_KL_wrt_prior_params / grad_VE_wrt_prior_params   (variational_cvi_sde.py:495-518)
  VariationalMarkovGP.grad_prior_sde_params / grad_initial_state    (vi_sde.py:457-482)
pinned by (i) central differences of the oracle's own forward functions and (ii) torch.autograd on a dense
op-by-op restatement of the taped forward pass (torch plays the role of tf.GradientTape).
"""
import itertools
import math
import types

import numpy as np
import pytest
import torch

from oracle import models_np as M
from oracle import sde_np as S
from torch_tape import _drift_torch

torch.set_default_dtype(torch.float64)


def all_sdes(rng):
    return [S.OrnsteinUhlenbeckSDE(0.7), S.DoubleWellSDE(scale=3.0, c=1.2), S.BenesSDE(0.8), S.SineDiffusionSDE(0.3),
            S.SqrtDiffusionSDE(1.3), S.MLPDrift(weights=rng.standard_normal(10)), S.VanderPolOscillatorSDE(1.5, 2.0),
            S.Lorenz63SDE(), S.OUNdSDE(0.4)]


def test_param_grads_match_central_differences():
    rng = np.random.default_rng(0)
    for sde in all_sdes(rng):
        x = rng.standard_normal((7, 5, sde.state_dim))
        df, dJ = sde.param_grads(x)
        p = np.asarray(sde.params, dtype=np.float64)
        assert df.shape == (len(p),) + x.shape and dJ.shape == (len(p),) + x.shape + (sde.state_dim,)
        for i in range(len(p)):
            h = 1e-6
            pp, pm = p.copy(), p.copy()
            pp[i] += h
            pm[i] -= h
            a, b = sde.with_params(pp), sde.with_params(pm)
            np.testing.assert_allclose((a.drift(x) - b.drift(x)) / (2 * h), df[i], atol=2e-8)
            np.testing.assert_allclose((a.jacobian(x) - b.jacobian(x)) / (2 * h), dJ[i], atol=2e-8)


def make_cvi(sde, N=30, n_obs=6, seed=1, dt=0.01, updates=2, lik=0.3, clip=(-1.0, 1.0)):
    rng = np.random.default_rng(seed)
    D = sde.state_dim
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(rng.choice(np.arange(1, N + 1), n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.4 + 0.5
    model = M.CVISitesSDE(sde, grid, (grid[idx], y), M.MultivariateGaussian(lik * np.eye(D)),
                          prior_initial_state=S.Gaussian(0.3 * np.ones(D), 0.5 * np.eye(D)),
                          clip_state_transitions=clip)
    for _ in range(updates):
        model.site_update(0.5, 0.5)
    return model


def clone_with_params(model, params):
    import copy
    m2 = copy.deepcopy(model)
    m2.prior_sde = model.prior_sde.with_params(params)
    return m2


CVI_CASES = [(S.DoubleWellSDE(scale=2.0, c=0.8), (-1.0, 1.0)), (S.OrnsteinUhlenbeckSDE(0.9), (-1.0, 1.0)),
             (S.BenesSDE(0.7), (-1.0, 1.0)), (S.VanderPolOscillatorSDE(1.3, 1.7, q=np.eye(2)), (-1.0, 1.0)),
             (S.MLPDrift(weights=np.random.default_rng(5).standard_normal(10)), (-1.0, 1.0)),
             # a clip window that saturates part of the linearised transitions (zero gradient there)
             (S.DoubleWellSDE(scale=2.0, c=0.8), (-0.994, 0.994))]


@pytest.mark.parametrize("case", range(len(CVI_CASES)))
def test_grad_kl_wrt_prior_params_fd(case):
    sde, clip = CVI_CASES[case]
    model = make_cvi(sde, clip=clip)
    g = model.grad_KL_wrt_prior_params()
    p = np.asarray(sde.params, dtype=np.float64)
    for i in range(len(p)):
        h = 1e-5
        pp, pm = p.copy(), p.copy()
        pp[i] += h
        pm[i] -= h
        fd = (clone_with_params(model, pp).KL_q_p() - clone_with_params(model, pm).KL_q_p()) / (2 * h)
        assert abs(fd - g[i]) <= 1e-7 * max(1.0, abs(g[i])), (i, fd, g[i])


def neg_ve_after_relinearisation(model, params):
    m2 = clone_with_params(model, params)
    m2.set_linearized_prior()
    return -m2.variational_expectation(*m2.dist_q.marginals)


@pytest.mark.parametrize("case", range(len(CVI_CASES)))
def test_grad_ve_wrt_prior_params_fd(case):
    sde, clip = CVI_CASES[case]
    model = make_cvi(sde, clip=clip)
    fx_before = model.fx_mus.copy()
    g = model.grad_VE_wrt_prior_params()
    np.testing.assert_array_equal(fx_before, model.fx_mus)   # the stored marginals are not refreshed (:513)
    if clip != (-1.0, 1.0):
        A_lin = model.dist_p_linearized.state_transitions
        assert ((A_lin > clip[1]) | (A_lin < clip[0])).any() and ((A_lin <= clip[1]) & (A_lin >= clip[0])).any()
    p = np.asarray(sde.params, dtype=np.float64)
    for i in range(len(p)):
        h = 1e-5
        pp, pm = p.copy(), p.copy()
        pp[i] += h
        pm[i] -= h
        fd = (neg_ve_after_relinearisation(model, pp) - neg_ve_after_relinearisation(model, pm)) / (2 * h)
        assert abs(fd - g[i]) <= 2e-6 * max(1.0, abs(g[i])), (i, fd, g[i])


# ---- torch.autograd on a dense restatement of the taped forward pass of grad_VE_wrt_prior_params


def _torch_sde(sde, params):
    names = {S.DRIFT_OU: ["decay"], S.DRIFT_OU_ND: ["decay"], S.DRIFT_DW: ["scale", "c"], S.DRIFT_BENES: ["theta"],
             S.DRIFT_SINE: ["theta"], S.DRIFT_SQRT: ["theta"], S.DRIFT_VDP: ["a", "tau"],
             S.DRIFT_LORENZ63: ["sigma", "rho", "beta"]}[sde.drift_id]
    shim = types.SimpleNamespace(drift_id=sde.drift_id, q=sde.q)
    for k, name in enumerate(names):
        setattr(shim, name, params[k])
    return shim


def _mvnquad_torch(fn, m, Sg, H, D):
    gx, gw = np.polynomial.hermite.hermgauss(H)
    xn = torch.tensor(np.array(list(itertools.product(*(gx,) * D))))
    wn = torch.tensor(np.prod(np.array(list(itertools.product(*(gw,) * D))), 1)) * math.pi ** (-D * 0.5)
    L = torch.linalg.cholesky(Sg)
    X = 2.0 ** 0.5 * torch.einsum("nij,pj->pni", L, xn) + m[None]
    vals = fn(X)
    return (vals * wn.reshape((-1,) + (1,) * (vals.dim() - 1))).sum(0)


def neg_ve_torch(model, params):
    """set_linearized_prior -> dist_q.marginals -> -VE with every op in torch; the chain is inverted densely."""
    sde = _torch_sde(model.prior_sde, params)
    D = model.state_dim
    path_m, path_S = torch.tensor(model.fx_mus[1:]), torch.tensor(model.fx_covs[1:])
    N = path_m.shape[0]
    n1 = N + 1

    def drift_and_jac(X):
        Xr = X.detach().clone().requires_grad_(True)
        f = _drift_torch(sde, Xr)
        rows = [torch.autograd.grad(f[..., i].sum(), Xr, create_graph=True)[0] for i in range(D)]
        return f, torch.stack(rows, dim=-2)

    E_f = _mvnquad_torch(lambda X: drift_and_jac(X)[0], path_m, path_S, 10, D)
    E_J = _mvnquad_torch(lambda X: drift_and_jac(X)[1], path_m, path_S, 10, D)
    tg = torch.tensor(model.time_grid)
    deltas = (tg[1:] - tg[:-1])
    b_lin = E_f - torch.einsum("nij,nj->ni", E_J, path_m)
    A = torch.eye(D)[None] + E_J * deltas[:, None, None]
    b = b_lin * deltas[:, None]
    if model.stabilize_ssm:
        lo, hi = model.clip_state_transitions
        A, b = torch.clamp(A, lo, hi), torch.clamp(b, lo, hi)
    cQ = torch.tensor(model.dist_p.cholesky_process_covariances)
    W = torch.cholesky_solve(torch.eye(D).expand(N, D, D), cQ)
    cP0 = torch.tensor(model.dist_p.cholesky_initial_covariance)
    P0inv = torch.cholesky_solve(torch.eye(D), cP0)
    mu0 = torch.tensor(model.dist_p.initial_mean)
    # dense precision and linear natural parameter of the prior chain x_{k+1} = A_k x_k + b_k + N(0, Q_k)
    P = torch.zeros(n1 * D, n1 * D)
    h = torch.zeros(n1 * D)
    P[:D, :D] = P[:D, :D] + P0inv
    h[:D] = h[:D] + P0inv @ mu0
    for k in range(N):
        a, c = slice(k * D, (k + 1) * D), slice((k + 1) * D, (k + 2) * D)
        WA = W[k] @ A[k]
        P[a, a] = P[a, a] + A[k].T @ WA
        P[c, c] = P[c, c] + W[k]
        P[c, a] = P[c, a] - WA
        P[a, c] = P[a, c] - WA.T
        h[a] = h[a] - A[k].T @ (W[k] @ b[k])
        h[c] = h[c] + W[k] @ b[k]
    # sites (constants): theta_d = -1/2 P_kk, theta_s = -P_{k+1,k}
    g = model.girsanov_sites
    d1, d2 = model._scatter_data_sites()
    s1 = torch.tensor(g.nat1 + d1)
    sd = torch.tensor(g.nat2.block_diagonal + d2)
    ss = torch.tensor(g.nat2.block_sub_diagonal)
    h = h + s1.reshape(-1)
    for k in range(n1):
        a = slice(k * D, (k + 1) * D)
        P[a, a] = P[a, a] - 2.0 * 0.5 * (sd[k] + sd[k].T)
    for k in range(N):
        a, c = slice(k * D, (k + 1) * D), slice((k + 1) * D, (k + 2) * D)
        P[c, a] = P[c, a] - ss[k]
        P[a, c] = P[a, c] - ss[k].T
    Sigma = torch.linalg.inv(P)
    mean = (Sigma @ h).reshape(n1, D)
    idx = model.obs_sites_indices
    Linv = torch.tensor(np.linalg.inv(model.likelihood.chol_covariance))
    Rinv = Linv.T @ Linv
    y = torch.tensor(model._observations)
    ve = 0.0
    for j, yj in zip(idx, y):
        Sj = Sigma[j * D:(j + 1) * D, j * D:(j + 1) * D]
        r = yj - mean[j]
        ve = ve - 0.5 * torch.sum(Rinv * Sj) - 0.5 * r @ Rinv @ r   # + constants
    return -ve


TORCH_CASES = [0, 1, 2, 3, 5]   # no MLP (its torch drift detaches the weights)


@pytest.mark.parametrize("case", TORCH_CASES)
def test_grad_ve_wrt_prior_params_torch_autograd(case):
    sde, clip = CVI_CASES[case]
    model = make_cvi(sde, clip=clip)
    g = model.grad_VE_wrt_prior_params()
    params = [torch.tensor(float(v), requires_grad=True) for v in np.asarray(sde.params)]
    loss = neg_ve_torch(model, params)
    # the dense restatement reproduces the oracle's forward value (up to the dropped constants) ...
    ref = torch.autograd.grad(loss, params)
    for i in range(len(params)):
        assert abs(ref[i].item() - g[i]) <= 1e-9 * max(1.0, abs(g[i])), (i, ref[i].item(), g[i])


def test_grad_kl_wrt_prior_params_torch_autograd():
    from torch_tape import _drift_torch as drift
    for case in (0, 3):
        sde, clip = CVI_CASES[case]
        model = make_cvi(sde, clip=clip)
        g = model.grad_KL_wrt_prior_params()
        params = [torch.tensor(float(v), requires_grad=True) for v in np.asarray(sde.params)]
        tsde = _torch_sde(sde, params)
        q = model.dist_q
        D = model.state_dim
        m, Sg = torch.tensor(q.marginal_means[:-1]), torch.tensor(q.marginal_covariances[:-1])
        A, b = torch.tensor(q.state_transitions), torch.tensor(q.state_offsets)
        Qpinv = torch.tensor(np.linalg.inv(model.dt * sde.q))

        def fn(X):
            d = X + model.dt * drift(tsde, X) - (torch.einsum("nij,pnj->pni", A, X) + b[None])
            return torch.einsum("pni,ij,pnj->pn", d, Qpinv, d)

        kl = 0.5 * _mvnquad_torch(fn, m, Sg, 20, D).sum()
        ref = torch.autograd.grad(kl, params)
        for i in range(len(params)):
            assert abs(ref[i].item() - g[i]) <= 1e-10 * max(1.0, abs(g[i])), (i, ref[i].item(), g[i])


def make_vgp(sde, N=30, n_obs=6, seed=2, dt=0.01, sweeps=3):
    rng = np.random.default_rng(seed)
    D = sde.state_dim
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(rng.choice(np.arange(1, N + 1), n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.4 + 0.5
    model = M.VariationalMarkovGP((grid[idx], y), sde, grid, M.MultivariateGaussian(0.3 * np.eye(D)),
                                  prior_initial_state=S.Gaussian(0.3 * np.ones(D), 0.5 * np.eye(D)))
    for _ in range(sweeps):
        model.sweep(0.3)
    return model


@pytest.mark.parametrize("sde", [S.DoubleWellSDE(scale=2.0, c=0.8), S.VanderPolOscillatorSDE(1.3, 1.7, q=np.eye(2)),
                                 S.SineDiffusionSDE(0.4)])
def test_vgp_grad_prior_sde_params_fd(sde):
    model = make_vgp(sde)
    g = model.grad_prior_sde_params()
    m, Sg = model.forward_pass
    p = np.asarray(sde.params, dtype=np.float64)
    import copy
    for i in range(len(p)):
        h = 1e-5
        vals = []
        for s in (+1, -1):
            pp = p.copy()
            pp[i] += s * h
            m2 = copy.deepcopy(model)
            m2.prior_sde = sde.with_params(pp)
            vals.append(m2.E_sde(m[1:], Sg[1:]))
        fd = (vals[0] - vals[1]) / (2 * h)
        assert abs(fd - g[i]) <= 1e-7 * max(1.0, abs(g[i])), (i, fd, g[i])


def test_vgp_grad_initial_state_torch_autograd():
    model = make_vgp(S.VanderPolOscillatorSDE(1.3, 1.7, q=np.eye(2)))
    model.update_initial_statistics(0.5)
    g_loc, g_tril = model.grad_initial_state()
    loc = torch.tensor(model.prior_initial_state.mu, requires_grad=True)
    tril = torch.tensor(np.linalg.cholesky(model.prior_initial_state.cov), requires_grad=True)
    q0 = torch.distributions.MultivariateNormal(torch.tensor(model.q_initial_mean),
                                                scale_tril=torch.tensor(model.q_initial_chol))
    p0 = torch.distributions.MultivariateNormal(loc, scale_tril=torch.tril(tril))
    kl = torch.distributions.kl_divergence(q0, p0)
    r = torch.autograd.grad(kl, [loc, tril])
    np.testing.assert_allclose(r[0].numpy(), g_loc, rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(np.tril(r[1].numpy()), g_tril, rtol=1e-10, atol=1e-12)
