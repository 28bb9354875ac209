"""
GPU parity tests: the CUDA path, called through the C ABI, against the NumPy oracle on the same inputs.
Tolerance (north_star): 1e-9 relative on ELBO, site natural parameters and marginals; relative error
here = max |cuda - oracle| / max |oracle| per tensor.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import gauss_markov_np as G
from oracle import models_np as M
from oracle import scalar_d1
from oracle import sde_np as S

pytestmark = pytest.mark.gpu

TOL = 1e-9


def rel(a, b, scale=None):
    """max |a - b| / max |b|  (or / scale when the tensor is a cancellation result near zero)."""
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    b = np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    den = np.abs(b).max() if scale is None else scale
    return float(np.abs(a - b).max() / (den if den > 0 else 1.0))


def dev(x):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float64).cuda()


@pytest.fixture(scope="module")
def ops():
    from vi_diffusion_processes_b200 import ops as o
    return o


@pytest.fixture(params=["seq", "tiled", "tiled-seg3"])
def sweeps(request, vdp_cfg):
    """Run the test once per sweep mapping: thread-per-sequence, parallel-in-time tiles with one segment
    per sequence, and tiles with every sequence split in three time segments (carry exchange path)."""
    mode = request.param
    vdp_cfg(SWEEPS="seq" if mode == "seq" else "tiled")
    if mode == "tiled-seg3":
        vdp_cfg(NSEG="3")
    else:
        vdp_cfg(NSEG=None)
    return mode


def _spd(rng, B, n, D):
    Ld = np.tril(rng.normal(size=(B, n, D, D))) * 0.3 + 2.0 * np.eye(D)
    Ls = rng.normal(size=(B, n - 1, D, D)) * 0.5
    Kd = Ld @ G.mT(Ld)
    Kd[:, 1:] += Ls @ G.mT(Ls)
    Ks = Ls @ G.mT(Ld[:, :-1])
    return Kd, Ks


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("B,n", [(1, 1), (3, 2), (5, 64), (2, 1001), (2, 5000)])
def test_btd_ops(ops, sweeps, D, B, n):
    rng = np.random.default_rng(D * 100 + n)
    Kd, Ks = _spd(rng, B, max(n, 2), D)
    Kd, Ks = Kd[:, :n], Ks[:, :n - 1]
    st = ops.new_status("cuda")
    if n == 1:
        Ld, _ = ops.btd_cholesky(dev(Kd), None, st)
        assert rel(Ld, G.btd_cholesky(Kd, None)[0]) < TOL
        return
    Ld, Ls = ops.btd_cholesky(dev(Kd), dev(Ks), st)
    oLd, oLs = G.btd_cholesky(Kd, Ks)
    assert ops.read_status(st)[0] == 0
    assert rel(Ld, oLd) < TOL and rel(Ls, oLs) < TOL
    Sd, Ss = ops.btd_inverse_band(Ld, Ls)
    oSd, oSs = G.btd_inverse_band(oLd, oLs)
    assert rel(Sd, oSd) < TOL and rel(Ss, oSs) < TOL
    rhs = rng.normal(size=(B, n, D))
    for tr in (False, True):
        assert rel(ops.btd_solve(Ld, Ls, dev(rhs), tr), G.btd_solve(oLd, oLs, rhs, tr)) < TOL
        assert rel(ops.btd_matvec(Ld, Ls, dev(rhs), False, tr), G.btd_matvec(np.tril(oLd), oLs, rhs, False, tr)) < TOL
    assert rel(ops.btd_matvec(dev(Kd), dev(Ks), dev(rhs), True), G.btd_matvec(Kd, Ks, rhs, True)) < TOL
    assert rel(ops.btd_logdet(Ld), G.LowerTriangularBlockTriDiagonal(oLd, oLs).abs_log_det()) < TOL


def test_cholesky_reports_non_pd(ops):
    Kd = np.tile(np.eye(2), (1, 5, 1, 1))
    Kd[0, 3] = -np.eye(2)
    st = ops.new_status("cuda")
    ops.btd_cholesky(dev(Kd), dev(np.zeros((1, 4, 2, 2))), st)
    code, idx = ops.read_status(st)
    assert code == 1 and idx == 3


def _rand_ssm(rng, B, N, D):
    A = np.eye(D) * 0.9 + 0.1 * rng.standard_normal((B, N, D, D))
    b = 0.1 * rng.standard_normal((B, N, D))
    R = rng.standard_normal((B, N, D, D))
    cQ = np.linalg.cholesky(0.01 * np.eye(D) + 0.002 * R @ G.mT(R))
    R0 = rng.standard_normal((B, D, D))
    cP0 = np.linalg.cholesky(0.5 * np.eye(D) + 0.1 * R0 @ G.mT(R0))
    mu0 = rng.standard_normal((B, D))
    return G.StateSpaceModel(mu0, cP0, A, b, cQ)


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("B,N", [(1, 1), (4, 37), (2, 500), (3, 4001)])
def test_ssm_transforms(ops, sweeps, D, B, N):
    rng = np.random.default_rng(7 + D + N)
    ssm = _rand_ssm(rng, B, N, D)
    args = [dev(x) for x in (ssm.state_transitions, ssm.state_offsets, ssm.cholesky_process_covariances,
                             ssm.initial_mean, ssm.cholesky_initial_covariance)]
    th = ops.ssm_to_naturals(*args)
    oth = G.ssm_to_naturals(ssm)
    for a, b in zip(th, oth):
        assert rel(a, b) < TOL
    st = ops.new_status("cuda")
    back = ops.naturals_to_ssm(*th, status=st)
    oback = G.naturals_to_ssm_params(*oth)
    assert ops.read_status(st)[0] == 0
    for a, b in zip(back, oback):
        assert rel(a, b) < TOL
    m, Sd = ops.ssm_marginals(*args, status=st)
    om, oS = ssm.marginals
    assert rel(m, om) < TOL and rel(Sd, oS) < TOL
    e = ops.ssm_to_expectations(args[0], m, Sd)
    oe = G.ssm_to_expectations(ssm)
    for a, b in zip(e, oe):
        assert rel(a, b) < TOL
    back = ops.expectations_to_ssm(*e, status=st)
    oback = G.expectations_to_ssm_params(*oe)
    for a, b in zip(back, oback):
        assert rel(a, b) < 1e-8   # chol(S1 - A S0 A^T) amplifies the rounding of the marginals


SDES = [S.OrnsteinUhlenbeckSDE(1.2), S.DoubleWellSDE(), S.BenesSDE(0.7), S.SineDiffusionSDE(0.3),
        S.SqrtDiffusionSDE(1.1), S.MLPDrift(), S.VanderPolOscillatorSDE(1.3, 0.8, np.eye(2)),
        S.Lorenz63SDE(q=np.eye(3)), S.OUNdSDE(0.5, np.eye(3))]
IDS = [type(s).__name__ for s in SDES]


@pytest.mark.parametrize("sde", SDES, ids=IDS)
def test_gh_expect_and_linearize(ops, sde):
    rng = np.random.default_rng(11)
    D, n = sde.state_dim, 50
    m = rng.normal(size=(n, D))
    R = rng.normal(size=(n, D, D))
    Sg = 0.05 * R @ G.mT(R) + 0.02 * np.eye(D)
    Ef, EJ = ops.gh_expect(sde.drift_id, sde.params, dev(m), dev(Sg), H=10, full_jacobian=sde.full_jacobian)
    assert rel(Ef, sde.expected_drift(m[None], Sg[None])[0]) < TOL
    assert rel(EJ, sde.expected_gradient_drift(m[None], Sg[None])[0]) < TOL
    grid = np.round(np.arange(n + 1) * 0.01, 2)
    A, b, cQ = ops.linearize_to_ssm(sde.drift_id, sde.params, sde.q, dev(grid), dev(m[None]), dev(Sg[None]))
    o = S.linearize_sde(sde, grid, S.Gaussian(m, Sg), S.Gaussian(np.zeros(D), np.eye(D)))
    # b = (E[f] - A m) dt is a cancellation (exactly 0 for OU): compare at the scale of its operands
    assert rel(A, o.state_transitions) < TOL
    assert rel(b, o.state_offsets, scale=0.01 * np.abs(sde.expected_drift(m[None], Sg[None])).max()) < TOL
    assert rel(cQ, o.cholesky_process_covariances) < TOL


@pytest.mark.parametrize("sde", SDES, ids=IDS)
def test_kl_and_grad(ops, sde):
    rng = np.random.default_rng(13)
    D = sde.state_dim
    N = 40 if D < 3 else 6
    ssm = _rand_ssm(rng, 1, N, D)
    ssm1 = S.remove_batch_from_ssm(ssm)
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    p0 = S.Gaussian(np.zeros(D) + 0.2, np.eye(D) * 0.7)
    m, Sg = ssm.marginals
    kl, g = ops.cvi_kl_and_grad(sde.drift_id, sde.params, sde.q, 0.01, dev(grid), p0.mu, p0.cov,
                                dev(ssm.state_transitions), dev(m), dev(Sg))
    okl, og = S.SDE_SSM_KL_with_grads_wrt_exp_params(ssm1, sde, 0.01, p0, grid)
    assert abs(kl.item() - okl) < TOL * abs(okl)
    for a, b in zip(g, og):
        assert rel(a[0], b) < TOL


def _d1_problem(rng, B, N, n_obs, dt):
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs)) * 0.3 + np.sign(rng.normal(size=(B, 1)))
    return grid, idx, y


@pytest.mark.parametrize("drift,params,B,N", [(S.DRIFT_DW, [4.0, 1.0], 3, 800), (S.DRIFT_OU, [1.2], 2, 1000),
                                              (S.DRIFT_BENES, [1.0], 2, 300), (S.DRIFT_DW, [4.0, 1.0], 64, 2000)])
def test_site_update_d1(sweeps, drift, params, B, N):
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(17)
    grid, idx, y = _d1_problem(rng, B, N, 20, 0.005)
    eng = CviEngine(drift, params, np.ones((1, 1)), grid, idx, y[..., None], [[0.1]], [1.0], [[1e-2]])
    ref = scalar_d1.BatchedCVISitesSDE1D(drift, params, 1.0, grid, idx, y, 0.1, 1.0, 1e-2)
    for it in range(3):
        e = eng.site_update(0.5, 0.5)
        r = ref.site_update(0.5, 0.5)
        eng.check_status()
        assert rel(e, r) < TOL
        assert rel(eng.fx_mus[..., 0], ref.fx_mus) < TOL and rel(eng.fx_covs[..., 0, 0], ref.fx_covs) < TOL
        # the sites are additive parts of q's natural parameters theta_q: a site tensor whose entries are
        # tiny next to theta_q (OU prior: the exact answer is 0) is compared at 1e-3 of theta_q's scale
        th = scalar_d1.ssm_to_naturals(ref.p["A"], ref.p["b"], ref.p["cQ"], ref.p["mu0"], ref.p["cP0"])
        for got, want, t in ((eng.g1[..., 0], ref.g1, th[0]), (eng.gd[..., 0, 0], ref.gd, th[1]),
                             (eng.gs[..., 0, 0], ref.gs, th[2])):
            scale = max(np.abs(want).max(), 1e-3 * np.abs(t + want).max())
            assert rel(got, want, scale=scale) < TOL
        assert rel(eng.d1[..., 0], ref.d1) < TOL and rel(eng.d2[..., 0, 0], ref.d2) < TOL
        if it == 1:
            eng.relinearize()
            ref.relinearize()


def test_site_update_vdp_2d(sweeps):
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(19)
    N, n_obs = 200, 16
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, 2))
    sde = S.VanderPolOscillatorSDE(1.0, 1.0, np.eye(2))
    x0 = S.Gaussian(np.ones(2), 1e-2 * np.eye(2))
    lik = M.MultivariateGaussian(0.1 * np.eye(2))
    ref = M.CVISitesSDE(sde, grid, (grid[idx], y), lik, prior_initial_state=x0)
    eng = CviEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, 0.1 * np.eye(2), x0.mu, x0.cov)
    for it in range(3):
        e = eng.site_update(0.5, 0.5)
        r = ref.site_update(0.5, 0.5)
        eng.check_status()
        assert abs(e.item() - r) < TOL * abs(r)
        assert rel(eng.fx_mus[0], ref.fx_mus) < TOL and rel(eng.fx_covs[0], ref.fx_covs) < TOL
        assert rel(eng.g1[0], ref.girsanov_sites.nat1) < TOL
        assert rel(eng.gd[0], ref.girsanov_sites.nat2.block_diagonal) < TOL
        assert rel(eng.gs[0], ref.girsanov_sites.nat2.block_sub_diagonal) < TOL
        if it == 1:
            eng.relinearize()
            ref.relinearize()


# ---------------------------------------------------------------------------------------------------
# VDP (VariationalMarkovGP, vi_sde.py) against the oracle
def _vgp_case(sde, N, n_obs, dt, seed, noise=0.1):
    rng = np.random.default_rng(seed)
    D = sde.state_dim
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(rng.choice(np.arange(1, N + 1), n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.2 + 1.0
    lik = M.MultivariateGaussian(noise * np.eye(D))
    x0 = S.Gaussian(np.full(D, 1.0), 0.01 * np.eye(D))
    return grid, idx, y, lik, x0


@pytest.mark.parametrize("sde,N,stab", [(S.DoubleWellSDE(), 600, False), (S.OrnsteinUhlenbeckSDE(0.8), 2500, False),
                                        (S.BenesSDE(0.7), 300, True),
                                        (S.VanderPolOscillatorSDE(1.0, 1.0, np.eye(2)), 250, False),
                                        (S.Lorenz63SDE(q=np.eye(3)), 120, False)],
                         ids=["dw", "ou", "benes-stab", "vdp2d", "lorenz3d"])
def test_vgp_sweeps(sweeps, sde, N, stab):
    from vi_diffusion_processes_b200.vgp import VgpEngine
    D = sde.state_dim
    dt = 0.01 if D < 3 else 0.002
    grid, idx, y, lik, x0 = _vgp_case(sde, N, 12, dt, seed=23 + N)
    ref = M.VariationalMarkovGP((grid[idx], y), sde, grid, lik, prior_initial_state=x0, stabilize_system=stab)
    eng = VgpEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, lik.chol_covariance, x0.mu, x0.cov,
                    stabilize_system=stab)
    # mean-reverting start, as the reference's driver does (docs/diffusion_processes/vi_markov_gp.py:83,114-115)
    ref.A = ref.A + 2.0 * np.eye(D)
    ref.b = ref.b + 2.0
    eng.A += 2.0 * torch.eye(D, dtype=torch.float64, device="cuda")
    eng.b += 2.0
    for it in range(3):
        # piecewise on the first iteration, fused afterwards
        upd = it >= 1
        if it == 0:
            m, Sg = eng.forward_pass()
            rm, rS = ref.forward_pass
            assert rel(m[0], rm) < TOL and rel(Sg[0], rS) < TOL
            eng.update_lagrange()
            ref.update_lagrange(rm, rS)
            assert rel(eng.lambda_lagrange[0], ref.lambda_lagrange) < TOL
            assert rel(eng.psi_lagrange[0], ref.psi_lagrange) < TOL
            eng.update_param(0.1)
            ref.update_param(rm, rS, 0.1)
            eng.forward_pass()
            e = eng.elbo()
            r = ref.elbo(*ref.forward_pass)
        else:
            e = eng.sweep(0.1, update_x0=upd, x0_lr=0.1)
            r = ref.sweep(0.1, update_x0=upd, x0_lr=0.1)
        eng.check_status()
        assert rel(eng.A[0], ref.A) < TOL and rel(eng.b[0], ref.b) < TOL
        assert rel(eng.q_initial_mean[0], ref.q_initial_mean) < TOL
        assert rel(eng.q_initial_chol[0], ref.q_initial_chol) < TOL
        assert abs(e.item() - r) < TOL * max(1.0, abs(r))


def test_vgp_batched_matches_single():
    """B sequences in one call == the same sequences one at a time (bitwise: no cross-sequence coupling)."""
    from vi_diffusion_processes_b200.vgp import VgpEngine
    sde = S.DoubleWellSDE()
    rng = np.random.default_rng(5)
    N, n_obs, B = 700, 10, 3
    grid = np.round(np.arange(N + 1) * 0.01, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 1)) * 0.2 + 1.0
    eng = VgpEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, [[0.1]], [1.0], [[0.01]])
    eng.A += 2.0
    eng.b += 2.0
    for _ in range(2):
        eng.sweep(0.1)
    for bi in range(B):
        one = VgpEngine(sde.drift_id, sde.params, sde.q, grid, idx[bi], y[bi], [[0.1]], [1.0], [[0.01]])
        one.A += 2.0
        one.b += 2.0
        for _ in range(2):
            one.sweep(0.1)
        assert torch.equal(one.A[0], eng.A[bi]) and torch.equal(one.m[0], eng.m[bi])
        assert torch.equal(one.elbo_value[0], eng.elbo_value[bi])


# ---------------------------------------------------------------------------------------------------
# band storage of banded_matrices and the triangular band solve of naturals_to_ssm_params
@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("with_sub", [True, False])
def test_band_pack_roundtrip_and_dense(ops, D, with_sub):
    rng = np.random.default_rng(D)
    B, n = 3, 7
    Kd, Ks = _spd(rng, B, n, D)
    band = ops.btd_block_to_band(dev(Kd), dev(Ks) if with_sub else None).cpu().numpy()
    dense = G.btd_to_dense(Kd, Ks if with_sub else None, symmetric=True)
    R = 2 * D if with_sub else D
    assert band.shape == (B, R, n * D)
    for r in range(R):        # band[r, j] = M[j + r, j]  (lower band, row 0 = main diagonal)
        for j in range(n * D):
            want = dense[:, j + r, j] if j + r < n * D else 0.0
            if not with_sub and (j % D) + r >= D:
                want = 0.0
            assert np.all(band[:, r, j] == want)
    Md, Ms = ops.btd_band_to_block(torch.as_tensor(band).cuda(), D, symmetric=True, with_sub=with_sub)
    assert np.array_equal(Md.cpu().numpy(), G.symmetrise_from_lower(Kd))
    if with_sub:
        assert np.array_equal(Ms.cpu().numpy(), Ks)
    Ml, _ = ops.btd_band_to_block(torch.as_tensor(band).cuda(), D, symmetric=False, with_sub=with_sub)
    assert np.array_equal(Ml.cpu().numpy(), np.tril(Kd))


@pytest.mark.parametrize("D", [1, 2, 3])
def test_conditional_precisions_match_dense_solve(ops, D):
    rng = np.random.default_rng(10 + D)
    B, n = 2, 6
    Kd, Ks = _spd(rng, B, n, D)
    As = rng.normal(size=(B, n - 1, D, D)) * 0.4
    got = ops.btd_conditional_precisions(dev(As), dev(Kd), dev(Ks)).cpu().numpy()
    P = G.btd_to_dense(Kd, Ks, symmetric=True)
    eye = np.broadcast_to(np.eye(D), (B, n, D, D))
    Ainv = G.btd_to_dense(eye, -As, symmetric=False)
    full = np.linalg.solve(np.swapaxes(Ainv, -1, -2), P)
    for k in range(n):
        blk = full[:, k * D:(k + 1) * D, k * D:(k + 1) * D]
        assert rel(got[:, k], blk) < 1e-12


@pytest.mark.parametrize("which", ["ou3", "lorenz"])
def test_site_update_3d(sweeps, which):
    """BASELINE config 5 shape (D = 3) at a size the oracle finishes in seconds (8000 nodes per transition)."""
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(23)
    N, n_obs = 40, 8
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    if which == "ou3":
        sde, y, mu0 = S.OUNdSDE(0.5, np.eye(3)), rng.normal(size=(n_obs, 3)) * 0.5, np.zeros(3)
    else:
        mu0 = np.array([1.0, 1.0, 20.0])
        sde, y = S.Lorenz63SDE(q=np.eye(3)), rng.normal(size=(n_obs, 3)) * 0.3 + mu0
    x0 = S.Gaussian(mu0, 0.5 * np.eye(3))
    lik = M.MultivariateGaussian(0.1 * np.eye(3))
    ref = M.CVISitesSDE(sde, grid, (grid[idx], y), lik, prior_initial_state=x0)
    eng = CviEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, 0.1 * np.eye(3), x0.mu, x0.cov)
    for it in range(2):
        e = eng.site_update(0.3, 0.3)
        r = ref.site_update(0.3, 0.3)
        eng.check_status()
        assert abs(e.item() - r) < TOL * abs(r)
        assert rel(eng.fx_mus[0], ref.fx_mus) < TOL and rel(eng.fx_covs[0], ref.fx_covs) < TOL
        th = ref.dist_p_nat() if hasattr(ref, "dist_p_nat") else None
        for got, want in ((eng.g1[0], ref.girsanov_sites.nat1), (eng.gd[0], ref.girsanov_sites.nat2.block_diagonal),
                          (eng.gs[0], ref.girsanov_sites.nat2.block_sub_diagonal)):
            scale = max(np.abs(want).max(), 1e-3 * np.abs(eng.fx_covs[0].cpu().numpy()).max() ** -1)
            assert rel(got, want, scale=scale) < TOL


# ---------------------------------------------------------------------------------------------------
# euler_maruyama (sde_utils.py:36-96)
class _RecordingRng:
    """np.random.Generator stand-in that records every standard_normal draw (to hand the same noise to CUDA)."""

    def __init__(self, seed):
        self.rng, self.draws = np.random.default_rng(seed), []

    def standard_normal(self, shape):
        z = self.rng.standard_normal(shape)
        self.draws.append(z)
        return z


@pytest.mark.parametrize("sde", [S.DoubleWellSDE(q=0.7 * np.ones((1, 1))), S.OrnsteinUhlenbeckSDE(0.8, np.ones((1, 1))),
                                 S.VanderPolOscillatorSDE(1.0, 2.0, np.array([[1.0, 0.2], [0.2, 0.5]])),
                                 S.Lorenz63SDE(q=np.eye(3))])
def test_euler_maruyama_given_noise(ops, sde):
    D = sde.q.shape[0]
    B, n = 5, 300
    rec = _RecordingRng(9)
    x0 = np.random.default_rng(1).normal(size=(B, D))
    grid = np.arange(n) * 0.004
    want = S.euler_maruyama(sde, x0, grid, rec)
    noise = np.stack(rec.draws[:n - 1], axis=1)                       # [B, n-1, D]; the reference discards the last draw
    got = ops.euler_maruyama(sde.drift_id, sde.params, sde.q, dev(x0), dev(grid), noise=dev(noise))
    assert rel(got, want) < 1e-12


def test_euler_maruyama_philox_statistics(ops):
    """OU started in its stationary law stays in it: mean 0, variance q / (2 decay); paths reproducible by seed."""
    B, n, decay = 8192, 400, 1.5
    x0 = torch.randn((B, 1), dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(3))
    x0 = x0 * (1.0 / (2 * decay)) ** 0.5
    grid = dev(np.arange(n) * 0.01)
    a = ops.euler_maruyama(S.DRIFT_OU, [decay], np.ones((1, 1)), x0, grid, seed=11)
    b = ops.euler_maruyama(S.DRIFT_OU, [decay], np.ones((1, 1)), x0, grid, seed=11)
    c = ops.euler_maruyama(S.DRIFT_OU, [decay], np.ones((1, 1)), x0, grid, seed=12)
    assert torch.equal(a, b) and not torch.equal(a, c)
    assert torch.equal(a[:, 0], x0)
    last = a[:, -1, 0]
    assert abs(last.mean().item()) < 0.03 and abs(last.var().item() - 1.0 / (2 * decay)) < 0.03
    # a sequence's path does not depend on the batch it is simulated in
    sub = ops.euler_maruyama(S.DRIFT_OU, [decay], np.ones((1, 1)), x0[:100].contiguous(), grid, seed=11)
    assert torch.equal(sub, a[:100])


def test_fused_naturals_equal_separate_pass(vdp_cfg):
    """The girsanov-update kernel emits the naturals of the posterior that follows; with VDP_NO_FUSED_NATURALS the
    rebuild makes its own pass over the sites (k_ssm_to_naturals + scatter).  Same statements, same state."""
    from oracle.sde_np import DRIFT_DW
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(5)
    N, B, n_obs = 700, 3, 9
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    idx[0, 0], idx[1, -1] = 0, N                      # first and last grid point observed
    y = rng.normal(size=(B, n_obs, 1)) * 0.3 + 1.0
    mk = lambda: CviEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]],
                           lazy_sites=False)
    a, c = mk(), mk()
    for _ in range(3):
        ea = a.site_update(0.5, 0.5).clone()
    vdp_cfg(NO_FUSED_NATURALS="1")
    for _ in range(3):
        ec = c.site_update(0.5, 0.5).clone()
    for name in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
        ref = getattr(c, name)
        assert (getattr(a, name) - ref).abs().max() <= 1e-12 * ref.abs().max(), name
    assert (ea - ec).abs().max() <= 1e-12 * ec.abs().max()


def test_carried_naturals_equal_full_pass():
    """Consecutive site updates patch the naturals left in the workspace at the observed grid points
    (VDP_FLAG_NATURALS_CARRIED); an engine that recomputes them over the whole grid every time ends in bitwise the
    same state -- across a re-linearisation, an in-place edit of the sites through torch, an ELBO call in between
    and data-sites-only updates (carried, and not carried: that one leaves no dense data-site copy to carry on)."""
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.markovflow.sde.sde import DRIFT_DW
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(21)
    N, B, n_obs = 900, 3, 11
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    idx[0, 0], idx[1, -1] = 0, N
    y = rng.normal(size=(B, n_obs, 1)) * 0.3 + 1.0
    mk = lambda carry, lazy=False: CviEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0],
                                             [[1e-2]], carry_naturals=carry, lazy_sites=lazy)
    # a: carried naturals, sites written every update; c: nothing carried; z: carried + sites held in theta between
    # updates (VDP_FLAG_SITES_IN_NATURALS, the engine's default): same state up to the rounding of theta + delta
    a, c, z = mk(True), mk(False), mk(True, True)
    carried, lazy = [], []
    for it in range(12):
        for e in (a, c, z):
            if it == 3:
                e.relinearize()
            if it == 5:
                e.g1.mul_(0.999)
            if it in (6, 9):
                e.classic_elbo()
        carried.append(a._nat_sig is not None and a._nat_sig == a._nat_signature())
        flags = _lib.FLAG_SKIP_GIRSANOV if it in (7, 9) else 0
        ea = a.site_update(0.4, 0.4, flags).clone()
        ec = c.site_update(0.4, 0.4, flags).clone()
        ez = z.site_update(0.4, 0.4, flags).clone()
        lazy.append(z._sites_stale)
        assert torch.equal(ea, ec), it
        assert rel(ez, ec.cpu().numpy()) < 1e-11, it
    a.check_status()
    z.check_status()
    assert carried == [False, True, True, False, True, False, False, True, True, False, False, True]
    assert lazy == [False, True, True, False, True, False, False, False, True, False, False, True]
    for name in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
        assert torch.equal(getattr(a, name), getattr(c, name)), name
        scale = None
        if name in ("g1", "gd", "gs"):   # a site is a summand of theta: right to 1e-11 at that scale
            th = dict(zip(("g1", "gd", "gs"), c._prior_naturals(c.pA, c.pb)))[name]
            scale = float((th + getattr(c, name)).abs().max())
        assert rel(getattr(z, name), getattr(c, name).cpu().numpy(), scale=scale) < 1e-11, name
    assert not z._sites_stale            # reading the sites brought them up to date


def test_carried_naturals_repeated_observation_points():
    """Two observations on one grid point add up (tf.scatter_nd); the carried path sums them in entry order, the
    scatter in arrival order: equal up to that rounding (amplified by the cancellation in the sites over 4 updates)."""
    from vi_diffusion_processes_b200.markovflow.sde.sde import DRIFT_DW
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(22)
    N, B, n_obs = 400, 2, 8
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    idx[0, 3] = idx[0, 4]
    idx[1, 0] = idx[1, 1] = 0
    y = rng.normal(size=(B, n_obs, 1)) * 0.3 + 1.0
    mk = lambda carry: CviEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]],
                                 carry_naturals=carry)
    a, c = mk(True), mk(False)
    for _ in range(4):
        ea, ec = a.site_update(0.5, 0.5), c.site_update(0.5, 0.5)
    a.check_status()
    for name in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
        ref = getattr(c, name)
        assert (getattr(a, name) - ref).abs().max() <= 1e-10 * ref.abs().max(), name
    assert (ea - ec).abs().max() <= 1e-10 * ec.abs().max()


@pytest.mark.parametrize("drift,params,noise", [("dw", [4.0, 1.0], 0.1), ("dw", [2.5, 0.7], 1e-3), ("ou", [0.8], 0.05)])
def test_closed_form_moments_equal_gauss_hermite(vdp_cfg, drift, params, noise):
    """For polynomial scalar drifts the 20-point Gauss-Hermite sums of the KL term and of its gradient are taken in
    closed form (central Gaussian moments; the rule is exact for these integrands).  VDP_GH_QUADRATURE=1 runs the
    quadrature loops: same sites, marginals and ELBO to rounding, also with a small observation noise (posterior
    variances of 1e-6 next to the observations, where an expansion about zero instead of the mean would cancel).
    With that noise the girsanov sites next to an observation are themselves ill-conditioned (the gradient with
    respect to the expectation parameters carries S^-1 ~ 1e6; the NumPy oracle and either variant differ by 1e-4
    there, tests/closed_diag.py, with the closed form the closer one), so the sites are compared on the scale of
    the total naturals, data sites included -- what the posterior is built from."""
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    from vi_diffusion_processes_b200.cvi import CviEngine
    did = {"dw": sde_mod.DRIFT_DW, "ou": sde_mod.DRIFT_OU}[drift]
    rng = np.random.default_rng(33)
    N, B, n_obs = 800, 4, 10
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 1)) * 0.3 + np.array([1.0, -1.0, 0.2, 1.4])[:, None, None]
    mk = lambda: CviEngine(did, params, np.ones((1, 1)), grid, idx, y, [[noise]], [1.0], [[1e-2]])
    a, c = mk(), mk()
    ea, ec = [], []
    for it in range(4):
        if it == 2:
            a.relinearize()
        ea.append(a.site_update(0.5, 0.5).clone())
    vdp_cfg(GH_QUADRATURE="1")
    for it in range(4):
        if it == 2:
            c.relinearize()
        ec.append(c.site_update(0.5, 0.5).clone())
    a.check_status()
    c.check_status()
    th = c._prior_naturals(c.pA, c.pb)
    for name, prior, data in (("g1", th[0], c.d1), ("gd", th[1], c.d2), ("gs", th[2], c.d2), ("fx_mus", None, None),
                              ("fx_covs", None, None)):
        ref = getattr(c, name)
        scale = ref.abs().max() if prior is None else max((ref + prior).abs().max(), data.abs().max())
        assert (getattr(a, name) - ref).abs().max() <= (1e-11 if prior is None else 1e-9) * scale, name
    for x, r in zip(ea, ec):
        assert (x - r).abs().max() <= 1e-11 * r.abs().max()


@pytest.mark.parametrize("drift,params", [("dw", [4.0, 1.0]), ("ou", [0.8])])
def test_vdp_closed_form_moments_equal_gauss_hermite(vdp_cfg, drift, params):
    """The same closed forms in the VDP sweep (E_sde and its gradients, 20 points; expected drift and slope of
    update_param, 10 points) against the quadrature loops (VDP_GH_QUADRATURE=1)."""
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    from vi_diffusion_processes_b200.vgp import VgpEngine
    did = {"dw": sde_mod.DRIFT_DW, "ou": sde_mod.DRIFT_OU}[drift]
    rng = np.random.default_rng(41)
    N, n_obs, B = 700, 10, 3
    grid = np.round(np.arange(N + 1) * 0.01, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 1)) * 0.2 + 1.0

    def run():
        e = VgpEngine(did, params, np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[0.01]])
        e.A += 2.0
        e.b += 2.0
        return e, [e.sweep(0.1, update_x0=it > 1).clone() for it in range(4)]

    a, ea = run()
    vdp_cfg(GH_QUADRATURE="1")
    c, ec = run()
    for name in ("A", "b", "m", "S", "lambda_lagrange", "psi_lagrange"):
        ref = getattr(c, name)
        assert (getattr(a, name) - ref).abs().max() <= 1e-11 * ref.abs().max(), name
    for x, r in zip(ea, ec):
        assert (x - r).abs().max() <= 1e-11 * r.abs().max()


@pytest.mark.parametrize("drift,params,D,N,nseg", [("dw", [4.0, 1.0], 1, 3000, 0), ("dw", [4.0, 1.0], 1, 6000, 3),
                                                  ("vdp", [1.0, 1.0], 2, 1500, 0), ("ou", [0.7], 1, 40, 0)])
def test_marginals_from_first_factor(vdp_cfg, drift, params, D, N, nseg):
    """Posterior rebuild: the marginal covariances come from the Cholesky factor naturals_to_ssm started from (the
    factor of the matrix the reference rebuilds from the state-space model and factorises a second time), the means
    from the forward recursion alone.  VDP_REFACTOR_PRECISION=1 runs the second factorisation: same state to
    rounding, also with the chain cut into segments."""
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    from vi_diffusion_processes_b200.cvi import CviEngine
    did = {"dw": sde_mod.DRIFT_DW, "ou": sde_mod.DRIFT_OU, "vdp": sde_mod.DRIFT_VDP}[drift]
    if nseg:
        vdp_cfg(NSEG=str(nseg))
    rng = np.random.default_rng(52)
    B, n_obs = 3, 12
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, D)) * 0.3 + 1.0
    mk = lambda: CviEngine(did, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.ones(D), 1e-2 * np.eye(D))
    a, c = mk(), mk()
    ea = [a.site_update(0.5, 0.5).clone() for _ in range(3)]
    vdp_cfg(REFACTOR_PRECISION="1")
    ec = [c.site_update(0.5, 0.5).clone() for _ in range(3)]
    a.check_status()
    c.check_status()
    th = c._prior_naturals(c.pA, c.pb)
    for name, prior in (("g1", th[0]), ("gd", th[1]), ("gs", th[2]), ("fx_mus", None), ("fx_covs", None)):
        ref = getattr(c, name)
        scale = ref.abs().max() if prior is None else (ref + prior).abs().max()
        assert (getattr(a, name) - ref).abs().max() <= 1e-11 * scale, name
    for x, r in zip(ea, ec):
        assert (x - r).abs().max() <= 1e-11 * r.abs().max()


@pytest.mark.parametrize("drift,params,N,nseg", [("dw", [4.0, 1.0], 3000, 0), ("dw", [4.0, 1.0], 7000, 3),
                                                ("ou", [0.7], 2500, 0)])
def test_kl_terms_summed_in_the_mean_sweep(vdp_cfg, drift, params, N, nseg):
    """Second rebuild of a site update, D = 1, polynomial drift: the mean sweep adds up the KL terms of classic_elbo on
    its way (per thread, per CTA, per tile; k_reduce_terms adds the tiles).  VDP_SEPARATE_KL_VALUE=1 runs k_kl_value
    over the grid instead: same ELBO and state up to the summation order."""
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    from vi_diffusion_processes_b200.cvi import CviEngine
    did = {"dw": sde_mod.DRIFT_DW, "ou": sde_mod.DRIFT_OU}[drift]
    if nseg:
        vdp_cfg(NSEG=str(nseg))
    rng = np.random.default_rng(61)
    B, n_obs = 3, 12
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 1)) * 0.3 + 1.0
    mk = lambda: CviEngine(did, params, np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]])
    a, c = mk(), mk()
    ea = [a.site_update(0.5, 0.5).clone() for _ in range(3)]
    vdp_cfg(SEPARATE_KL_VALUE="1")
    ec = [c.site_update(0.5, 0.5).clone() for _ in range(3)]
    a.check_status()
    th = c._prior_naturals(c.pA, c.pb)
    for name, prior in (("g1", th[0]), ("gd", th[1]), ("gs", th[2]), ("fx_mus", None), ("fx_covs", None)):
        ref = getattr(c, name)
        scale = ref.abs().max() if prior is None else (ref + prior).abs().max()
        assert (getattr(a, name) - ref).abs().max() <= 1e-11 * scale, name
    for x, r in zip(ea, ec):
        assert (x - r).abs().max() <= 1e-12 * r.abs().max()
    assert (a.kl - c.kl).abs().max() <= 1e-12 * c.kl.abs().max()


def test_cuda_graph_replay_equals_direct_calls():
    """site_update_graphed / sweep_graphed (CUDA-graph replay for launch-bound single sequences) leave bitwise the
    state of the direct calls, across a re-linearisation (buffers replaced -> re-capture) and a step-size change."""
    from oracle.sde_np import DRIFT_DW
    from vi_diffusion_processes_b200.cvi import CviEngine
    from vi_diffusion_processes_b200.vgp import VgpEngine
    rng = np.random.default_rng(8)
    N, n_obs = 600, 12
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))[None]
    y = rng.normal(size=(1, n_obs, 1)) * 0.3 + 1.0
    mk = lambda: CviEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]],
                           lazy_sites=False)   # (a replay writes the sites every update: bitwise comparison)
    a, b = mk(), mk()
    for it in range(6):
        lr = 0.5 if it < 4 else 0.25
        ea = a.site_update(lr, lr).clone()
        eb = b.site_update_graphed(lr, lr).clone()
        assert torch.equal(ea, eb), it
        if it == 2:
            a.relinearize()
            b.relinearize()
    for name in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    b.check_status()
    mv = lambda: VgpEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]])
    c, d = mv(), mv()
    for e in (c, d):
        e.A += 2.0
        e.b += 2.0
    for it in range(5):
        ec = c.sweep(0.1, update_x0=it > 1).clone()
        ed = d.sweep_graphed(0.1, update_x0=it > 1).clone()
        assert torch.equal(ec, ed), it
    for name in ("A", "b", "lambda_lagrange", "psi_lagrange", "q_initial_mean", "q_initial_chol", "m", "S"):
        assert torch.equal(getattr(c, name), getattr(d, name)), name


@pytest.mark.parametrize("D,N,nseg", [(1, 3000, 0), (1, 6000, 3), (2, 1500, 0), (3, 700, 2)])
def test_vdp_forward_recursion_equals_rebuilt_precision(vdp_cfg, D, N, nseg):
    """VariationalMarkovGP.forward_pass on the tiled route takes the marginal covariances from the forward recursion
    S_{k+1} = A S A^T + Q riding on the mean sweep (one fused kernel); VGP_REFACTOR_PRECISION keeps the reference's
    route (rebuilt precision, banded Cholesky, inverse band; state_space_model.py:262).  Same marginals to rounding,
    over sweeps, with and without stabilize_system, with several time segments per sequence."""
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    from vi_diffusion_processes_b200.vgp import VgpEngine
    vdp_cfg(SWEEPS="tiled")
    if nseg:
        vdp_cfg(NSEG=nseg)
    did, params = {1: (sde_mod.DRIFT_DW, [4.0, 1.0]), 2: (sde_mod.DRIFT_VDP, [1.0, 1.0]),
                   3: (sde_mod.DRIFT_OU_ND, [0.5])}[D]
    rng = np.random.default_rng(51 + D)
    n_obs, B = 10, 2
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, D)) * 0.2 + 1.0
    for stab in (False, True):
        def run(flags):
            e = VgpEngine(did, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.ones(D), 0.01 * np.eye(D),
                          stabilize_system=stab)
            e.route_flags = flags
            e.A += 2.0 * torch.eye(D, dtype=torch.float64, device="cuda")
            e.b += 2.0
            return e, [e.sweep(0.05, update_x0=it > 1).clone() for it in range(3)]
        a, ea = run(0)
        c, ec = run(_lib.VGP_REFACTOR_PRECISION)
        a.check_status()
        for name in ("m", "S", "A", "b", "lambda_lagrange", "psi_lagrange"):
            ref = getattr(c, name)
            assert (getattr(a, name) - ref).abs().max() <= 1e-11 * ref.abs().max(), (stab, name)
        for x, r in zip(ea, ec):
            assert (x - r).abs().max() <= 1e-11 * r.abs().max()
