"""
StateSpaceModel.kl_divergence / log_pdf / sample on the GPU (vdp_ssm_kl_divergence, vdp_ssm_log_pdf, vdp_ssm_sample)
against the oracle (oracle/gauss_markov_np.py:418-446) and against dense Gaussian algebra, the way the reference's
tests/unit/test_state_space_model.py:145-184 checks them (there against tfp's dense multivariate normal).
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import gauss_markov_np as G

pytestmark = pytest.mark.gpu

TOL = 1e-9


def _rand_ssm(rng, B, N, D, shift=0.0):
    A = np.eye(D) * 0.9 + 0.1 * rng.standard_normal((B, N, D, D))
    b = 0.1 * rng.standard_normal((B, N, D)) + shift
    R = rng.standard_normal((B, N, D, D))
    cQ = np.linalg.cholesky(0.01 * np.eye(D) + 0.002 * R @ G.mT(R))
    R0 = rng.standard_normal((B, D, D))
    cP0 = np.linalg.cholesky(0.5 * np.eye(D) + 0.1 * R0 @ G.mT(R0))
    mu0 = rng.standard_normal((B, D))
    return G.StateSpaceModel(mu0, cP0, A, b, cQ)


def _dev(ssm):
    from vi_diffusion_processes_b200.markovflow.state_space_model import StateSpaceModel
    return StateSpaceModel(ssm.initial_mean, ssm.cholesky_initial_covariance, ssm.state_transitions,
                           ssm.state_offsets, ssm.cholesky_process_covariances)


def _dense_kl(s1, s2):
    """KL of the equivalent dense multivariate normals (what tfp computes in the reference's test)."""
    out = []
    m1, m2 = s1.marginal_means, s2.marginal_means
    P1, P2 = s1.precision.to_dense(), s2.precision.to_dense()
    for b in range(m1.shape[0]):
        S1 = np.linalg.inv(P1[b])
        d = (m2[b] - m1[b]).reshape(-1)
        out.append(0.5 * (np.trace(P2[b] @ S1) + d @ P2[b] @ d - d.size - np.linalg.slogdet(P2[b])[1]
                          + np.linalg.slogdet(P1[b])[1]))
    return np.asarray(out)


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("B,N", [(3, 1), (2, 40), (2, 700), (1, 4001)])
def test_kl_divergence(D, B, N):
    rng = np.random.default_rng(100 * D + N)
    s1, s2 = _rand_ssm(rng, B, N, D), _rand_ssm(rng, B, N, D, shift=0.05)
    want = s1.kl_divergence(s2)
    got = _dev(s1).kl_divergence(_dev(s2)).cpu().numpy()
    assert np.abs(got - want).max() <= TOL * np.abs(want).max()
    back = _dev(s2).kl_divergence(_dev(s1)).cpu().numpy()
    assert np.abs(back - s2.kl_divergence(s1)).max() <= TOL * np.abs(back).max()
    if N <= 40:   # the reference's own check: equal to the KL of the dense Gaussians (rtol 1e-3 there)
        np.testing.assert_allclose(got, _dense_kl(s1, s2), rtol=1e-8)
    self_kl = _dev(s1).kl_divergence(_dev(s1)).cpu().numpy()   # test_self_kl_divergence: atol 1e-6
    assert np.abs(self_kl).max() < 1e-8 * max(1.0, np.abs(want).max())


def test_kl_divergence_unbatched_and_multi_batch():
    rng = np.random.default_rng(5)
    s1, s2 = _rand_ssm(rng, 2, 5, 2), _rand_ssm(rng, 2, 5, 2)
    want = s1.kl_divergence(s2)
    from vi_diffusion_processes_b200.markovflow.state_space_model import StateSpaceModel

    def pick(s, f):
        return StateSpaceModel(f(s.initial_mean), f(s.cholesky_initial_covariance), f(s.state_transitions),
                               f(s.state_offsets), f(s.cholesky_process_covariances))
    one = pick(s1, lambda t: t[0]).kl_divergence(pick(s2, lambda t: t[0]))
    assert one.dim() == 0 and abs(float(one) - want[0]) <= TOL * abs(want[0])
    two = pick(s1, lambda t: t.reshape((2, 1) + t.shape[1:])).kl_divergence(
        pick(s2, lambda t: t.reshape((2, 1) + t.shape[1:])))
    assert tuple(two.shape) == (2, 1)
    np.testing.assert_allclose(two.cpu().numpy().reshape(-1), want, rtol=TOL)


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("S,B,N", [(1, 1, 1), (4, 3, 37), (2, 2, 3000)])
def test_log_pdf_and_sample(D, S, B, N):
    rng = np.random.default_rng(7 * D + N)
    ssm = _rand_ssm(rng, B, N, D)
    dev = _dev(ssm)
    eps = rng.standard_normal((S, B, N + 1, D))
    # sample with the caller's normals == a_inv_block.solve(offsets + chol eps)   (state_space_model.py:311-322)
    z = (ssm.concatenated_cholesky_process_covariance @ eps[..., None])[..., 0]
    want_x = ssm.a_inv_block.solve(ssm.concatenated_state_offsets + z)
    got_x = dev.sample((S,), eps=eps).cpu().numpy()
    assert got_x.shape == (S, B, N + 1, D)
    assert np.abs(got_x - want_x).max() <= TOL * np.abs(want_x).max()
    want_lp = ssm.log_pdf(want_x)
    got_lp = dev.log_pdf(torch.as_tensor(want_x).cuda()).cpu().numpy()
    assert got_lp.shape == (S, B)
    assert np.abs(got_lp - want_lp).max() <= TOL * np.abs(want_lp).max()


def test_sample_statistics_with_the_device_generator():
    """Philox draws: sample mean / variance of many trajectories approach the marginals (4 sigma bands)."""
    rng = np.random.default_rng(11)
    ssm = _rand_ssm(rng, 1, 30, 2)
    dev = _dev(ssm)
    S = 20000
    x = dev.sample(S, seed=1234)
    assert tuple(x.shape) == (S, 1, 31, 2)
    x2 = dev.sample(S, seed=1234)
    assert torch.equal(x, x2)                      # reproducible per seed
    assert not torch.equal(x, dev.sample(S, seed=1235))
    m, Sd = ssm.marginals
    xs = x.cpu().numpy()[:, 0]
    var = np.diagonal(Sd[0], axis1=-2, axis2=-1)
    assert (np.abs(xs.mean(0) - m[0]) < 4.0 * np.sqrt(var / S) + 1e-12).all()
    assert (np.abs(xs.var(0) - var) < 4.0 * var * np.sqrt(2.0 / S) + 1e-12).all()
    assert tuple(dev.sample((0,)).shape) == (0, 1, 31, 2)
