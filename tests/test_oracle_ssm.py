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
