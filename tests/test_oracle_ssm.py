"""
Ports of the reference tests that pin StateSpaceModel and the parameter transforms, on the oracle.
Reference: /root/reference/tests/unit/test_state_space_model.py (:40-101 marginals/precision/log-det,
:145-184 kl_divergence) and tests/unit/test_ssm_gaussian_transformations.py (:65-106 round trips).
"""
import numpy as np
import pytest

from oracle.gauss_markov_np import (StateSpaceModel, expectations_to_ssm_params, naturals_to_ssm_params,
                                    naturals_to_ssm_params_no_smoothing, ssm_to_expectations, ssm_to_naturals,
                                    ssm_to_naturals_no_smoothing, state_space_model_from_covariances)

BATCH = [(3,), (), (2, 1)]


def _pd(rng, shape, D):
    a = rng.normal(size=shape + (D, D))
    return a @ np.swapaxes(a, -1, -2) + D * np.eye(D)


def _ssm(rng, batch, D, N, stable=False):
    A = rng.normal(size=batch + (N, D, D))
    if stable:
        A = 0.5 * A / np.sqrt(D)
    return state_space_model_from_covariances(rng.normal(size=batch + (D,)), _pd(rng, batch, D), A,
                                              rng.normal(size=batch + (N, D)), _pd(rng, batch + (N,), D))


def _dense_joint(ssm):
    """Explicit forward recursion for means, covariances and the dense precision."""
    A, b, mu0 = ssm.state_transitions, ssm.state_offsets, ssm.initial_mean
    P0 = ssm.initial_covariance
    Q = ssm.cholesky_process_covariances @ np.swapaxes(ssm.cholesky_process_covariances, -1, -2)
    N = A.shape[-3]
    means, covs = [mu0], [P0]
    for k in range(N):
        means.append((A[..., k, :, :] @ means[-1][..., None])[..., 0] + b[..., k, :])
        covs.append(A[..., k, :, :] @ covs[-1] @ np.swapaxes(A[..., k, :, :], -1, -2) + Q[..., k, :, :])
    return np.stack(means, -2), np.stack(covs, -3)


@pytest.mark.parametrize("batch", BATCH)
@pytest.mark.parametrize("D", [1, 3, 5])
@pytest.mark.parametrize("N", [1, 3, 5])
def test_marginals(batch, D, N):
    rng = np.random.default_rng(71892305)
    ssm = _ssm(rng, batch, D, N)
    means, covs = _dense_joint(ssm)
    np.testing.assert_allclose(ssm.marginal_means, means, rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(ssm.marginal_covariances, covs, rtol=1e-7, atol=1e-7)
    np.testing.assert_allclose(ssm.subsequent_covariances(covs), ssm.state_transitions @ covs[..., :-1, :, :])


@pytest.mark.parametrize("batch", BATCH)
@pytest.mark.parametrize("D", [1, 3])
def test_log_det_precision(batch, D):
    rng = np.random.default_rng(7)
    ssm = _ssm(rng, batch, D, 4)
    _, ld = np.linalg.slogdet(ssm.precision.to_dense())
    np.testing.assert_allclose(ssm.log_det_precision(), ld, rtol=1e-9)


def _dense_kl(s1, s2):
    m1, m2 = s1.marginal_means.reshape(-1), s2.marginal_means.reshape(-1)
    P1, P2 = s1.precision.to_dense(), s2.precision.to_dense()
    S1 = np.linalg.inv(P1)
    d = m2 - m1
    return 0.5 * (np.trace(P2 @ S1) + d @ P2 @ d - m1.size - np.linalg.slogdet(P2)[1] + np.linalg.slogdet(P1)[1])


@pytest.mark.parametrize("D", [1, 3])
def test_kl_divergence(D):
    rng = np.random.default_rng(8)
    s1, s2 = _ssm(rng, (), D, 4, stable=True), _ssm(rng, (), D, 4, stable=True)
    np.testing.assert_allclose(s1.kl_divergence(s2), _dense_kl(s1, s2), rtol=1e-8)
    np.testing.assert_allclose(s1.kl_divergence(s1), 0.0, atol=1e-8)


@pytest.mark.parametrize("batch", BATCH)
@pytest.mark.parametrize("D", [1, 2, 3])
def test_round_trips(batch, D):
    rng = np.random.default_rng(9)
    ssm = _ssm(rng, batch, D, 30, stable=True)
    ref = (ssm.state_transitions, ssm.state_offsets, ssm.cholesky_initial_covariance,
           ssm.cholesky_process_covariances, ssm.initial_mean)
    for fwd, bwd in [(ssm_to_expectations, expectations_to_ssm_params), (ssm_to_naturals, naturals_to_ssm_params),
                     (ssm_to_naturals_no_smoothing, naturals_to_ssm_params_no_smoothing)]:
        out = bwd(*fwd(ssm))
        for o, r in zip(out, ref):
            np.testing.assert_allclose(o, r, rtol=1e-7, atol=1e-6)


def test_naturals_are_the_dense_precision():
    rng = np.random.default_rng(10)
    ssm = _ssm(rng, (), 2, 5, stable=True)
    th1, thd, ths = ssm_to_naturals(ssm)
    P = ssm.precision
    np.testing.assert_allclose(-2 * thd, P.block_diagonal, rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(-ths, P.block_sub_diagonal, rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(th1.reshape(-1), P.to_dense() @ ssm.marginal_means.reshape(-1), rtol=1e-8, atol=1e-8)


def test_log_pdf_matches_dense():
    rng = np.random.default_rng(11)
    ssm = _ssm(rng, (), 2, 4, stable=True)
    x = rng.normal(size=(5, 2))
    m = ssm.marginal_means.reshape(-1)
    P = ssm.precision.to_dense()
    d = x.reshape(-1) - m
    expect = -0.5 * d @ P @ d + 0.5 * np.linalg.slogdet(P)[1] - 0.5 * d.size * np.log(2 * np.pi)
    np.testing.assert_allclose(ssm.log_pdf(x), expect, rtol=1e-9)


def _exp_to_ssm_torch(e1, ed, es):
    """expectations_to_ssm_params (ssm_gaussian_transformations.py:93-178) op by op in torch float64."""
    import torch
    el = e1[..., None]
    mcov = ed - el @ el.transpose(-1, -2)
    csub = es.transpose(-1, -2) - el[:-1] @ el[1:].transpose(-1, -2)
    mch = torch.linalg.cholesky(mcov)
    As = torch.cholesky_solve(csub, mch[:-1]).transpose(-1, -2)
    b = (el[1:] - As @ el[:-1])[..., 0]
    ccov = mcov[1:] - As @ (mcov[:-1] @ As.transpose(-1, -2))
    return As, b, mch[0], torch.linalg.cholesky(ccov), el[0, :, 0]


@pytest.mark.parametrize("D", [1, 2, 3])
def test_expectations_to_ssm_vjp_matches_autograd(D):
    """The chain-rule step of SSMNaturalGradient (ssm_natgrad.py:168): hand adjoint vs torch.autograd (torch's
    Cholesky backward is the same symmetrised formula TensorFlow registers)."""
    torch = pytest.importorskip("torch")
    from oracle.sde_np import expectations_to_ssm_params_vjp
    rng = np.random.default_rng(12 + D)
    ssm = _ssm(rng, (), D, 6, stable=True)
    e = [np.ascontiguousarray(x) for x in ssm_to_expectations(ssm)]
    te = [torch.tensor(x, dtype=torch.float64, requires_grad=True) for x in e]
    outs = _exp_to_ssm_torch(*te)
    cot = [rng.normal(size=tuple(o.shape)) for o in outs]
    cot[2], cot[3] = np.tril(cot[2]), np.tril(cot[3])
    want = torch.autograd.grad(outs, te, grad_outputs=[torch.tensor(c) for c in cot])
    got = expectations_to_ssm_params_vjp(*e, cot[0], cot[1], cot[2], cot[3], cot[4])
    sym = lambda X: 0.5 * (X + np.swapaxes(X, -1, -2))
    np.testing.assert_allclose(got[0], want[0].numpy(), rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(sym(got[1]), sym(want[1].numpy()), rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(got[2], want[2].numpy(), rtol=1e-9, atol=1e-10)


def test_natgrad_step_with_unit_gamma_solves_a_conjugate_problem():
    """Natural gradient with gamma = 1 on L = KL[q || p] jumps to q = p when the loss is expressed in the
    expectation parameters linearly (conjugacy): here L = <eta(q), theta(q0) - theta(p)> has dL/d eta constant,
    so one step from q0 with that gradient lands exactly on theta(p)."""
    from oracle.models_np import ssm_natgrad_step
    from oracle.sde_np import expectations_to_ssm_params_vjp
    rng = np.random.default_rng(13)
    q0 = _ssm(rng, (), 2, 5, stable=True)
    p = _ssm(rng, (), 2, 5, stable=True)
    tq, tp = ssm_to_naturals(q0), ssm_to_naturals(p)
    g_eta = [a - b for a, b in zip(tq, tp)]          # desired dL/d eta
    # find dL/d(ssm params) with J^T dL_dssm = g_eta by solving the small linear system column by column
    e = ssm_to_expectations(q0)
    shapes = [(5, 2, 2), (5, 2), (2, 2), (5, 2, 2), (2,)]
    sizes = [int(np.prod(s)) for s in shapes]
    cols = []
    for i in range(sum(sizes)):
        v = np.zeros(sum(sizes))
        v[i] = 1.0
        parts = [x.reshape(s) for x, s in zip(np.split(v, np.cumsum(sizes)[:-1]), shapes)]
        if np.any(np.triu(parts[2], 1)) or np.any(np.triu(parts[3], 1)):
            cols.append(None)
            continue
        g = expectations_to_ssm_params_vjp(*e, parts[0], parts[1], parts[2], parts[3], parts[4])
        gd = np.tril(g[1]) + np.tril(g[1], -1)      # count the symmetric pair once (lower storage)
        cols.append(np.concatenate([g[0].ravel(), gd.ravel(), g[2].ravel()]))
    keep = [i for i, c in enumerate(cols) if c is not None]
    J = np.stack([cols[i] for i in keep], axis=1)
    gdl = np.tril(g_eta[1]) + np.tril(g_eta[1], -1)
    rhs = np.concatenate([g_eta[0].ravel(), gdl.ravel(), g_eta[2].ravel()])
    sol, *_ = np.linalg.lstsq(J, rhs, rcond=None)
    np.testing.assert_allclose(J @ sol, rhs, atol=1e-8)
    v = np.zeros(sum(sizes))
    v[keep] = sol
    dL = [x.reshape(s) for x, s in zip(np.split(v, np.cumsum(sizes)[:-1]), shapes)]
    new = ssm_natgrad_step(q0, dL, 1.0)
    np.testing.assert_allclose(new.state_transitions, p.state_transitions, atol=1e-6)
    np.testing.assert_allclose(new.state_offsets, p.state_offsets, atol=1e-6)
    np.testing.assert_allclose(new.cholesky_process_covariances, p.cholesky_process_covariances, atol=1e-6)
