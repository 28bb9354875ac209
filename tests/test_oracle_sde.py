"""
Oracle checks for the SDE layer.
Ports: /root/reference/tests/unit/test_sde.py :66-106 (linearize_sde on OU), :149-197 (drift-difference KL
vs StateSpaceModel.kl_divergence), :200-226 (LinearDrift.set_from_ssm), :109-146 (Euler-Maruyama).
Own pins (the reference has none for these): Gauss-Hermite closed forms, hand adjoint vs torch.autograd
(tests/torch_tape.py plays tf.GradientTape) and vs central differences.
"""
import numpy as np
import pytest

from oracle import sde_np as S
from oracle.gauss_markov_np import StateSpaceModel, ssm_to_expectations, state_space_model_from_covariances

from torch_tape import kl_and_grads_torch


def test_mvnquad_polynomial_is_exact():
    rng = np.random.default_rng(0)
    m = rng.normal(size=(7, 1))
    s = rng.uniform(0.1, 2.0, size=(7, 1, 1))
    sde = S.DoubleWellSDE(scale=4.0, c=1.0)
    # E[4x(1-x^2)] = 4(m - m^3 - 3 m s);  E[4(1-3x^2)] = 4(1 - 3 m^2 - 3 s)
    ef = sde.expected_drift(m[None], s[None])[0]
    eg = sde.expected_gradient_drift(m[None], s[None])[0]
    np.testing.assert_allclose(ef, 4 * (m - m ** 3 - 3 * m * s[:, 0]), rtol=1e-12)
    np.testing.assert_allclose(eg, 4 * (1 - 3 * m ** 2 - 3 * s[:, 0]), rtol=1e-12)


def test_mvnquad_2d_moments():
    rng = np.random.default_rng(1)
    a = rng.normal(size=(5, 2, 2))
    cov = a @ np.swapaxes(a, -1, -2) + 0.5 * np.eye(2)
    m = rng.normal(size=(5, 2))
    val = S.mvnquad(lambda x: x[:, :, None] * x[:, None, :], m, cov, H=6, Din=2, Dout=(2, 2))
    np.testing.assert_allclose(val, cov + m[:, :, None] * m[:, None, :], rtol=1e-11, atol=1e-12)


@pytest.mark.parametrize("nbatch", [3, 1])
def test_linearize_sde_ou_statedim_1(nbatch):
    rng = np.random.default_rng(33)
    N = 1000
    dt = 1.0 / N
    grid = np.linspace(dt, 1.0, N)
    decay = float(rng.normal())
    sde = S.OrnsteinUhlenbeckSDE(decay, np.eye(1))
    x0 = rng.normal(size=(nbatch, 1))
    q_mean = S.euler_maruyama(sde, x0, grid, rng)[:, :-1, :]
    q_cov = 1e-4 * np.ones(q_mean.shape + (1,))
    ssm = S.linearize_sde(sde, grid, S.Gaussian(q_mean, q_cov), S.Gaussian(x0, q_cov[:, 0]))
    dtg = grid[1] - grid[0]
    np.testing.assert_allclose(ssm.state_transitions, np.full(q_cov.shape, 1 - decay * dtg), atol=1e-3)
    np.testing.assert_allclose(ssm.state_offsets, 0.0, atol=1e-3)
    np.testing.assert_allclose(ssm.cholesky_process_covariances, np.full(q_cov.shape, np.sqrt(dtg)), atol=1e-3)


def test_euler_maruyama_shapes_and_zero_noise():
    rng = np.random.default_rng(2)
    grid = np.linspace(0.01, 1.0, 100)
    sde = S.OrnsteinUhlenbeckSDE(0.5, 1e-20 * np.eye(1))
    x0 = np.ones((4, 1))
    vals = S.euler_maruyama(sde, x0, grid, rng)
    assert vals.shape == (4, 100, 1)
    expect = np.cumprod(np.concatenate([[1.0], 1 - 0.5 * np.diff(np.concatenate([[0.0], grid]))[:-1]]))
    np.testing.assert_allclose(vals[0, :, 0], expect, atol=1e-5)


def test_KL_sde_matches_ssm_kl():
    """test_sde.py:149-197: drift-difference functional ~ KL of the two discretised OU SSMs (2 decimals)."""
    N, dt = 400, 0.0025
    grid = np.arange(N + 1) * dt
    q = np.eye(1)
    ssm_q = S.LinearDrift(A=-0.5 * np.ones((N, 1, 1)), b=np.zeros((N, 1))).to_ssm(
        q=np.repeat(q[None], N, 0), transition_times=grid, initial_mean=np.zeros(1), initial_chol_covariance=np.eye(1))
    ssm_p = S.LinearDrift(A=-1.5 * np.ones((N, 1, 1)), b=np.zeros((N, 1))).to_ssm(
        q=np.repeat(q[None], N, 0), transition_times=grid, initial_mean=np.zeros(1), initial_chol_covariance=np.eye(1))
    m, Sg = ssm_q.marginals
    val = S.squared_drift_difference_along_Gaussian_path(S.OrnsteinUhlenbeckSDE(1.5, q),
                                                         S.LinearDrift(-0.5 * np.ones((N, 1, 1)), np.zeros((N, 1))),
                                                         S.Gaussian(m[0, :-1], Sg[0, :-1]), dt)
    np.testing.assert_almost_equal(val, ssm_q.kl_divergence(ssm_p)[0], decimal=2)


def test_ssm_to_linear_drift():
    rng = np.random.default_rng(3)
    N, D, dt = 10, 2, 0.01
    A = rng.normal(size=(N, D, D))
    b = rng.normal(size=(N, D))
    grid = np.arange(N + 1) * dt
    ssm = S.LinearDrift(A, b).to_ssm(np.repeat(np.eye(D)[None], N, 0), grid, np.zeros(D), np.eye(D))
    ld = S.LinearDrift()
    ld.set_from_ssm(ssm, dt)
    np.testing.assert_allclose(ld.A, A, rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(ld.b, b, rtol=1e-7, atol=1e-9)


def _rand_q(rng, N, D):
    A = np.eye(D) * 0.95 + 0.05 * rng.standard_normal((N, D, D))
    b = 0.1 * rng.standard_normal((N, D))
    R = rng.standard_normal((N, D, D))
    Lq = np.linalg.cholesky(0.01 * np.eye(D) + 0.002 * R @ np.swapaxes(R, -1, -2))
    return StateSpaceModel(rng.standard_normal(D), np.linalg.cholesky(0.5 * np.eye(D)), A, b, Lq)


SDES = [S.DoubleWellSDE(), S.BenesSDE(0.7), S.SineDiffusionSDE(0.3), S.OrnsteinUhlenbeckSDE(1.2), S.MLPDrift(),
        S.SqrtDiffusionSDE(1.1), S.VanderPolOscillatorSDE(1.3, 0.8, np.eye(2)), S.Lorenz63SDE(q=np.eye(3))]


@pytest.mark.parametrize("sde", SDES, ids=lambda s: type(s).__name__)
def test_kl_gradient_matches_autograd(sde):
    rng = np.random.default_rng(4)
    D, N = sde.state_dim, 12
    ssm = _rand_q(rng, N, D)
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    eta = ssm_to_expectations(ssm)
    p0 = S.Gaussian(np.zeros(D) + 0.2, np.eye(D) * 0.7)
    H = 20 if D < 3 else 6
    kl, g = S.sde_ssm_kl_and_grads(*eta, sde, 0.01, p0, grid, H=H)
    assert kl == S.sde_ssm_kl_forward(*eta, sde, 0.01, p0, grid, H=H)
    klt, gt = kl_and_grads_torch(*eta, sde, 0.01, p0.mu, p0.cov, grid, H=H)
    np.testing.assert_allclose(kl, klt, rtol=1e-13)
    for a, b in zip(g, gt):
        assert np.abs(a - b).max() <= 1e-12 * np.abs(b).max()


def test_kl_gradient_matches_finite_differences():
    rng = np.random.default_rng(5)
    sde = S.DoubleWellSDE()
    N = 5
    ssm = _rand_q(rng, N, 1)
    grid = np.arange(N + 1) * 0.01
    eta = [e.copy() for e in ssm_to_expectations(ssm)]
    p0 = S.Gaussian(np.zeros(1), np.eye(1))
    _, g = S.sde_ssm_kl_and_grads(*eta, sde, 0.01, p0, grid)
    for i in range(3):
        for idx in np.ndindex(eta[i].shape):
            h = 1e-6
            ep = [e.copy() for e in eta]
            em = [e.copy() for e in eta]
            ep[i][idx] += h
            em[i][idx] -= h
            fd = (S.sde_ssm_kl_forward(*ep, sde, 0.01, p0, grid) - S.sde_ssm_kl_forward(*em, sde, 0.01, p0, grid)) / (2 * h)
            np.testing.assert_allclose(g[i][idx], fd, rtol=2e-5, atol=1e-6)


def test_drift_difference_gradient_matches_finite_differences():
    rng = np.random.default_rng(6)
    sde = S.VanderPolOscillatorSDE(1.2, 0.9, np.eye(2))
    n, D = 4, 2
    A = rng.normal(size=(n, D, D))
    b = rng.normal(size=(n, D))
    m = rng.normal(size=(n, D))
    R = rng.normal(size=(n, D, D))
    Sg = R @ np.swapaxes(R, -1, -2) + np.eye(D)
    val, dm, dS = S.grad_squared_drift_difference(sde, A, b, m, Sg, 0.01, H=8)
    np.testing.assert_allclose(val, S.squared_drift_difference_along_Gaussian_path(
        sde, S.LinearDrift(A, b), S.Gaussian(m, Sg), 0.01, 8), rtol=1e-12)
    f = lambda mm, SS: S.squared_drift_difference_along_Gaussian_path(sde, S.LinearDrift(A, b), S.Gaussian(mm, SS), 0.01, 8)
    h = 1e-6
    for idx in np.ndindex(m.shape):
        mp, mn = m.copy(), m.copy()
        mp[idx] += h
        mn[idx] -= h
        np.testing.assert_allclose(dm[idx], (f(mp, Sg) - f(mn, Sg)) / (2 * h), rtol=1e-5, atol=1e-7)
    for k in range(n):
        for (i, j) in [(0, 0), (1, 1), (1, 0)]:
            Sp, Sn = Sg.copy(), Sg.copy()
            for (a, c) in {(i, j), (j, i)}:
                Sp[k, a, c] += h
                Sn[k, a, c] -= h
            fd = (f(mp * 0 + m, Sp) - f(m, Sn)) / (2 * h)
            expect = dS[k, i, j] if i == j else dS[k, i, j] + dS[k, j, i]
            np.testing.assert_allclose(expect, fd, rtol=1e-5, atol=1e-7)
