"""
GPU tests of the reference-shaped Python interface (vi_diffusion_processes_b200.markovflow): the same calls
a user of markovflow makes, checked against dense NumPy linear algebra (as the reference's own unit tests
tests/unit/test_block_tri_diag.py, test_state_space_model.py, test_ssm_gaussian_transformations.py do) and
against the oracle for the models.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import gauss_markov_np as G
from oracle import models_np as M
from oracle import sde_np as S

pytestmark = pytest.mark.gpu
TOL = 1e-9


def rel(a, b, scale=None):
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    b = np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    den = np.abs(b).max() if scale is None else scale
    return float(np.abs(a - b).max() / (den if den > 0 else 1.0))


def dev(x):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float64).cuda()


@pytest.fixture(scope="module")
def mf():
    import vi_diffusion_processes_b200.markovflow as m
    from vi_diffusion_processes_b200.markovflow import (block_tri_diag, gauss_markov, likelihoods,  # noqa: F401
                                                        ssm_gaussian_transformations, ssm_natgrad, state_space_model)
    from vi_diffusion_processes_b200.markovflow import models, sde  # noqa: F401
    from vi_diffusion_processes_b200.markovflow.sde import sde_utils  # noqa: F401
    return m


def _dense_sym(Kd, Ks):
    n, D = Kd.shape[0], Kd.shape[1]
    K = np.zeros((n * D, n * D))
    for k in range(n):
        K[k * D:(k + 1) * D, k * D:(k + 1) * D] = Kd[k]
        if k + 1 < n:
            K[(k + 1) * D:(k + 2) * D, k * D:(k + 1) * D] = Ks[k]
            K[k * D:(k + 1) * D, (k + 1) * D:(k + 2) * D] = Ks[k].T
    return K


@pytest.mark.parametrize("batch", [(), (3,)])
@pytest.mark.parametrize("D,n", [(1, 6), (2, 5), (3, 40)])
def test_block_tri_diag_against_dense(mf, batch, D, n):
    """tests/unit/test_block_tri_diag.py: cholesky, solve, block_diagonal_of_inverse, abs_log_det, dense_mult."""
    rng = np.random.default_rng(31 + D + n)
    Ld0 = np.tril(rng.normal(size=batch + (n, D, D))) * 0.3 + 2.0 * np.eye(D)
    Ls0 = rng.normal(size=batch + (n - 1, D, D)) * 0.5
    Kd = Ld0 @ G.mT(Ld0)
    Kd[..., 1:, :, :] += Ls0 @ G.mT(Ls0)
    Ks = Ls0 @ G.mT(Ld0[..., :-1, :, :])
    sym = mf.block_tri_diag.SymmetricBlockTriDiagonal(dev(Kd), dev(Ks))
    chol = sym.cholesky
    rhs = rng.normal(size=batch + (n, D))
    flat = lambda x: x.reshape(-1, *x.shape[len(batch):])
    got_dense = flat(sym.to_dense().cpu().numpy())
    got_Ld, got_Ls = flat(chol.block_diagonal.cpu().numpy()), flat(chol.block_sub_diagonal.cpu().numpy())
    got_sol = flat(chol.solve(dev(rhs)).cpu().numpy())
    got_solT = flat(chol.solve(dev(rhs), transpose_left=True).cpu().numpy())
    got_inv = flat(chol.block_diagonal_of_inverse().cpu().numpy())
    got_ld = chol.abs_log_det().cpu().numpy().reshape(-1)
    got_mul = flat(sym.dense_mult(dev(rhs)).cpu().numpy())
    for i, (kd, ks, r) in enumerate(zip(flat(Kd), flat(Ks), flat(rhs))):
        K = _dense_sym(kd, ks)
        Lfull = np.linalg.cholesky(K)
        assert rel(got_dense[i], K) < TOL
        assert rel(got_Ld[i], np.stack([Lfull[k * D:(k + 1) * D, k * D:(k + 1) * D] for k in range(n)])) < TOL
        assert rel(got_Ls[i], np.stack([Lfull[(k + 1) * D:(k + 2) * D, k * D:(k + 1) * D] for k in range(n - 1)])) < TOL
        assert rel(got_sol[i].reshape(-1), np.linalg.solve(Lfull, r.reshape(-1))) < TOL
        assert rel(got_solT[i].reshape(-1), np.linalg.solve(Lfull.T, r.reshape(-1))) < TOL
        Kinv = np.linalg.inv(K)
        assert rel(got_inv[i], np.stack([Kinv[k * D:(k + 1) * D, k * D:(k + 1) * D] for k in range(n)])) < TOL
        assert abs(got_ld[i] - 0.5 * np.linalg.slogdet(K)[1]) < TOL * n
        assert rel(got_mul[i].reshape(-1), K @ r.reshape(-1)) < TOL


def test_cholesky_of_indefinite_matrix_raises(mf):
    Kd = np.tile(np.eye(2), (5, 1, 1))
    Kd[3] = -np.eye(2)
    with pytest.raises(ValueError):
        mf.block_tri_diag.SymmetricBlockTriDiagonal(dev(Kd), dev(np.zeros((4, 2, 2)))).cholesky


def _rand_ssm_np(rng, N, D):
    A = np.eye(D) * 0.9 + 0.1 * rng.standard_normal((N, D, D))
    b = 0.1 * rng.standard_normal((N, D))
    R = rng.standard_normal((N, D, D))
    cQ = np.linalg.cholesky(0.01 * np.eye(D) + 0.002 * R @ G.mT(R))
    R0 = rng.standard_normal((D, D))
    cP0 = np.linalg.cholesky(0.5 * np.eye(D) + 0.1 * R0 @ R0.T)
    mu0 = rng.standard_normal(D)
    return mu0, cP0, A, b, cQ


@pytest.mark.parametrize("D,N", [(1, 30), (2, 12), (3, 700)])
def test_state_space_model_and_transformations(mf, D, N):
    """test_state_space_model.py (marginals vs the dense joint) and test_ssm_gaussian_transformations.py (round trips)."""
    rng = np.random.default_rng(41 + D)
    mu0, cP0, A, b, cQ = _rand_ssm_np(rng, N, D)
    ssm = mf.state_space_model.StateSpaceModel(mu0, cP0, A, b, cQ)
    ref = G.StateSpaceModel(mu0, cP0, A, b, cQ)
    m, Sd = ssm.marginals
    # dense truth: x = Ainv_block^-1 (offsets + noise)
    if N <= 30:
        n = N + 1
        Ab = np.eye(n * D)
        for k in range(N):
            Ab[(k + 1) * D:(k + 2) * D, k * D:(k + 1) * D] = -A[k]
        Abi = np.linalg.inv(Ab)
        off = np.concatenate([mu0[None], b]).reshape(-1)
        chols = np.concatenate([cP0[None], cQ])
        Qb = np.zeros((n * D, n * D))
        for k in range(n):
            Qb[k * D:(k + 1) * D, k * D:(k + 1) * D] = chols[k] @ chols[k].T
        cov = Abi @ Qb @ Abi.T
        assert rel(m.reshape(-1), Abi @ off) < TOL
        assert rel(Sd, np.stack([cov[k * D:(k + 1) * D, k * D:(k + 1) * D] for k in range(n)])) < TOL
        assert rel(ssm.precision.to_dense(), np.linalg.inv(cov)) < 1e-7
    om, oS = ref.marginals
    assert rel(m, om) < TOL and rel(Sd, oS) < TOL
    T = mf.ssm_gaussian_transformations
    th = T.ssm_to_naturals(ssm)
    back = T.naturals_to_ssm_params(*th)
    for got, want in zip(back, (A, b, cP0, cQ, mu0)):
        assert rel(got, want) < 1e-8
    et = T.ssm_to_expectations(ssm)
    back = T.expectations_to_ssm_params(*et)
    for got, want in zip(back, (A, b, cP0, cQ, mu0)):
        assert rel(got, want) < 1e-7
    assert abs(ssm.log_det_precision().item() - ref.log_det_precision()) < TOL * N


def test_sde_expectations_and_linearize(mf):
    sde = mf.sde.DoubleWellSDE(q=np.ones((1, 1)))
    osde = S.DoubleWellSDE()
    rng = np.random.default_rng(43)
    n = 60
    m = rng.normal(size=(1, n, 1))
    Sg = 0.05 + 0.1 * rng.random(size=(1, n, 1, 1))
    assert rel(sde.expected_drift(m, Sg), osde.expected_drift(m, Sg)) < TOL
    assert rel(sde.expected_gradient_drift(m, Sg), osde.expected_gradient_drift(m, Sg)) < TOL
    # closed form: E[4x(1-x^2)] = 4m - 4(m^3 + 3 m S)
    want = 4 * m - 4 * (m ** 3 + 3 * m * Sg[..., 0])
    assert rel(sde.expected_drift(m, Sg), want) < 1e-12
    grid = np.round(np.arange(n + 1) * 0.01, 2)
    U = mf.sde.sde_utils
    lin = U.linearize_sde(sde, grid, U.Gaussian(m[0], Sg[0]), U.Gaussian(np.zeros(1), np.eye(1)))
    o = S.linearize_sde(osde, grid, S.Gaussian(m[0], Sg[0]), S.Gaussian(np.zeros(1), np.eye(1)))
    assert rel(lin.state_transitions, o.state_transitions) < TOL
    assert rel(lin.cholesky_process_covariances, o.cholesky_process_covariances) < TOL
    x = dev(rng.normal(size=(5, 1)))
    assert rel(sde.drift(x), osde.drift(x.cpu().numpy())) < 1e-14
    # E_sde of a linear drift against the oracle
    A = rng.normal(size=(n, 1, 1)) * 0.2 - 1.0
    b = rng.normal(size=(n, 1)) * 0.2
    ld = mf.sde.LinearDrift(A, b)
    val = U.squared_drift_difference_along_Gaussian_path(sde, ld, U.Gaussian(m[0], Sg[0]), 0.01)
    want = S.squared_drift_difference_along_Gaussian_path(osde, S.LinearDrift(A, b), S.Gaussian(m[0], Sg[0]), 0.01)
    assert abs(val.item() - want) < TOL * abs(want)
    ssm = ld.to_ssm(np.ones((1, 1)), grid, np.zeros(1), np.eye(1))
    ossm = S.LinearDrift(A, b).to_ssm(q=np.ones((n, 1, 1)), transition_times=grid, initial_mean=np.zeros(1),
                                      initial_chol_covariance=np.eye(1))
    assert rel(ssm.state_transitions, ossm.state_transitions) < TOL
    assert rel(ssm.state_offsets, ossm.state_offsets) < TOL
    assert rel(ssm.cholesky_process_covariances, ossm.cholesky_process_covariances) < TOL


def _cvi_case(D=1):
    rng = np.random.default_rng(47)
    N, n_obs = 300, 14
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.3 + 0.8
    return grid, idx, y


def test_cvi_sites_sde_reference_calls(mf):
    """The reference's hot loop, call by call (cvi_dp_trainer.py:71-77), against the oracle model."""
    grid, idx, y = _cvi_case()
    U = mf.sde.sde_utils
    model = mf.models.CVISitesSDE(mf.sde.DoubleWellSDE(), grid, (grid[idx], y),
                                  mf.likelihoods.MultivariateGaussian(0.1 * np.eye(1)),
                                  prior_initial_state=U.Gaussian(np.ones(1), 1e-2 * np.eye(1)))
    ref = M.CVISitesSDE(S.DoubleWellSDE(), grid, (grid[idx], y), M.MultivariateGaussian(0.1 * np.eye(1)),
                        prior_initial_state=S.Gaussian(np.ones(1), 1e-2 * np.eye(1)))
    assert np.array_equal(model.obs_sites_indices, ref.obs_sites_indices)
    for it in range(3):
        model.update_data_sites(0.5)
        ref.update_data_sites(0.5)
        assert rel(model.data_sites.nat1, ref.data_sites.nat1) < TOL and rel(model.fx_mus, ref.fx_mus) < TOL
        model.update_girsanov_sites(0.5)
        ref.update_girsanov_sites(0.5)
        assert rel(model.girsanov_sites.nat1, ref.girsanov_sites.nat1) < TOL
        assert rel(model.girsanov_sites.nat2.block_diagonal, ref.girsanov_sites.nat2.block_diagonal) < TOL
        assert rel(model.girsanov_sites.nat2.block_sub_diagonal, ref.girsanov_sites.nat2.block_sub_diagonal) < TOL
        e, r = model.classic_elbo().item(), ref.classic_elbo()
        assert abs(e - r) < TOL * abs(r)
        q, oq = model.dist_q, ref.dist_q
        assert rel(q.state_transitions, oq.state_transitions) < TOL and rel(q.state_offsets, oq.state_offsets) < TOL
        if it == 1:
            model.relinearize()
            ref.relinearize()
            assert rel(model.dist_p.state_transitions, ref.dist_p.state_transitions) < TOL
    g = model.grad_kl_wrt_exp_param()
    _, og = ref.grad_kl_wrt_exp_param()
    for a, b in zip(g, og):
        assert rel(a, b) < 1e-8


def test_variational_markov_gp_reference_calls(mf):
    grid, idx, y = _cvi_case()
    idx = idx[idx > 0]
    y = y[-idx.shape[0]:] * 0.2 + 0.9
    U = mf.sde.sde_utils
    model = mf.models.VariationalMarkovGP((grid[idx], y), mf.sde.DoubleWellSDE(), grid,
                                          mf.likelihoods.MultivariateGaussian(0.1 * np.eye(1)),
                                          prior_initial_state=U.Gaussian(np.ones(1), 1e-2 * np.eye(1)))
    ref = M.VariationalMarkovGP((grid[idx], y), S.DoubleWellSDE(), grid, M.MultivariateGaussian(0.1 * np.eye(1)),
                                prior_initial_state=S.Gaussian(np.ones(1), 1e-2 * np.eye(1)))
    model.A = model.A + 2.0
    model.b = model.b + 2.0
    ref.A, ref.b = ref.A + 2.0, ref.b + 2.0
    for it in range(3):
        m, Sg = model.forward_pass
        rm, rS = ref.forward_pass
        model.update_lagrange(m, Sg)
        ref.update_lagrange(rm, rS)
        model.update_param(m, Sg, lr=0.1)
        ref.update_param(rm, rS, 0.1)
        if it > 0:
            model.update_initial_statistics(0.1)
            ref.update_initial_statistics(0.1)
        m, Sg = model.forward_pass
        rm, rS = ref.forward_pass
        assert rel(m, rm) < TOL and rel(Sg, rS) < TOL
        e, r = model.elbo(m, Sg).item(), ref.elbo(rm, rS)
        assert abs(e - r) < TOL * abs(r)
        assert rel(model.A, ref.A) < TOL and rel(model.psi_lagrange, ref.psi_lagrange) < TOL


def _exp_to_ssm_torch(e1, ed, es):
    """expectations_to_ssm_params (ssm_gaussian_transformations.py:93-178) op by op in torch (CPU, float64)."""
    el = e1[..., None]
    mcov = ed - el @ el.transpose(-1, -2)
    csub = es.transpose(-1, -2) - el[:-1] @ el[1:].transpose(-1, -2)
    mch = torch.linalg.cholesky(mcov)
    As = torch.cholesky_solve(csub, mch[:-1]).transpose(-1, -2)
    b = (el[1:] - As @ el[:-1])[..., 0]
    ccov = mcov[1:] - As @ (mcov[:-1] @ As.transpose(-1, -2))
    return As, b, mch[0], torch.linalg.cholesky(ccov), el[0, :, 0]


@pytest.mark.parametrize("D,N", [(1, 9), (2, 7), (3, 50)])
def test_expectations_to_ssm_vjp_against_autograd(D, N):
    from vi_diffusion_processes_b200 import ops
    rng = np.random.default_rng(53 + D)
    ssm = G.StateSpaceModel(*_rand_ssm_np(rng, N, D))
    e = [np.ascontiguousarray(x) for x in G.ssm_to_expectations(ssm)]
    te = [torch.tensor(x, dtype=torch.float64, requires_grad=True) for x in e]
    outs = _exp_to_ssm_torch(*te)
    cot = [torch.tensor(rng.normal(size=tuple(o.shape))) for o in outs]
    cot[2], cot[3] = torch.tril(cot[2]), torch.tril(cot[3])
    want = torch.autograd.grad(outs, te, grad_outputs=cot)
    got = ops.expectations_to_ssm_vjp(dev(e[0][None]), dev(e[1][None]), dev(e[2][None]), dev(cot[0].numpy()[None]),
                                      dev(cot[1].numpy()[None]), dev(cot[2].numpy()[None]), dev(cot[3].numpy()[None]),
                                      dev(cot[4].numpy()[None]))
    for g, w in zip(got, want):
        w = w.numpy()
        if w.ndim == 3 and w.shape[0] == N + 1:
            w = 0.5 * (w + np.swapaxes(w, -1, -2))   # eta_diag is symmetric: compare the symmetric part
            g = 0.5 * (g[0] + g[0].transpose(-1, -2))
        else:
            g = g[0]
        assert rel(g, w) < 1e-9


def test_ssm_natural_gradient_step(mf):
    """SSMNaturalGradient.minimize (momentum=False) against the oracle's finite-difference step."""
    rng = np.random.default_rng(59)
    D, N = 2, 4
    mu0, cP0, A, b, cQ = _rand_ssm_np(rng, N, D)
    ssm = mf.state_space_model.StateSpaceModel(mu0, cP0, A, b, cQ)
    target = dev(rng.normal(size=(N + 1, D)))
    W = dev(rng.normal(size=(N, D, D)))

    def loss_fn():
        # any differentiable function of the SSM parameters
        return ((ssm.state_offsets - target[1:]) ** 2).sum() + (ssm.state_transitions * W).sum() \
            + (ssm.cholesky_process_covariances ** 2).sum() + ((ssm.initial_mean - target[0]) ** 2).sum() \
            + (torch.tril(ssm.cholesky_initial_covariance) ** 2).sum()

    tn, Wn = target.cpu().numpy(), W.cpu().numpy()
    dL = (Wn, 2 * (b - tn[1:]), 2 * np.tril(cP0), 2 * np.tril(cQ), 2 * (mu0 - tn[0]))
    want = M.ssm_natgrad_step(G.StateSpaceModel(mu0, cP0, A, b, cQ), dL, 0.01)
    mf.ssm_natgrad.SSMNaturalGradient(gamma=0.01, momentum=False).minimize(loss_fn, ssm)
    assert rel(ssm.state_transitions, want.state_transitions) < 1e-5
    assert rel(ssm.state_offsets, want.state_offsets, scale=1.0) < 1e-5
    assert rel(ssm.cholesky_process_covariances, want.cholesky_process_covariances) < 1e-5
    assert rel(ssm.initial_mean, want.initial_mean) < 1e-5
    assert rel(ssm.cholesky_initial_covariance, want.cholesky_initial_covariance) < 1e-5
