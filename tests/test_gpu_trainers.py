"""The trainer loops (cvi_dp_trainer.py / vi_markov_gp_trainer.py mirrors) on the CUDA models against the same
loops driven over the NumPy oracle models: same number of iterations, ELBO / NLPD / RMSE traces to 1e-9."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import models_np as M
from oracle import sde_np as S

pytestmark = pytest.mark.gpu


class _OracleCvi(M.CVISitesSDE):
    """The oracle model with the two halves of relinearize() under the names the trainer calls."""

    def transform_sites_to_linearized_prior(self):
        self.girsanov_sites = M.tranform_girsanov_sites(self.girsanov_sites, self.dist_p, self.dist_p_linearized)
        self.dist_p = self.dist_p_linearized

    def relinearize_and_transform_sites(self):
        last = self.dist_p
        self.set_linearized_prior()
        self.girsanov_sites = M.tranform_girsanov_sites(self.girsanov_sites, last, self.dist_p)


class _OracleVgp(M.VariationalMarkovGP):
    m = property(lambda s: s._last[0])
    S = property(lambda s: s._last[1])

    def sweep(self, lr, update_x0=False, x0_lr=0.1):
        e = super().sweep(lr, update_x0=update_x0, x0_lr=x0_lr)
        self._last = self.forward_pass
        return e


def _data(rng, N, n_obs, dt, D=1):
    grid = np.round(np.arange(N + 1) * dt, 6)
    pick = np.sort(rng.choice(N + 1, 2 * n_obs, replace=False))
    tr, te = pick[::2], pick[1::2]
    y = rng.normal(size=(2 * n_obs, D)) * 0.2 + 1.0
    return grid, (grid[tr], y[::2]), (grid[te], y[1::2])


def test_cvi_sites_trainer_matches_oracle_loop():
    from vi_diffusion_processes_b200.markovflow.likelihoods import MultivariateGaussian
    from vi_diffusion_processes_b200.markovflow.models import CVISitesSDE
    from vi_diffusion_processes_b200.markovflow.sde import DoubleWellSDE
    from vi_diffusion_processes_b200.markovflow.sde.sde_utils import Gaussian
    from vi_diffusion_processes_b200.trainers import CVISitesTrainer
    rng = np.random.default_rng(41)
    grid, train, test = _data(rng, 300, 12, 0.01)
    kw = dict(max_itr=3, optim_tol=1e-3, max_itr_sites_optim=4, girsanov_sites_lr=0.5, data_sites_lr=0.5)
    model = CVISitesSDE(DoubleWellSDE(q=np.ones((1, 1))), grid, train, MultivariateGaussian(0.1 * np.eye(1)),
                        prior_initial_state=Gaussian(np.ones(1), 0.01 * np.eye(1)))
    ref = _OracleCvi(S.DoubleWellSDE(q=np.ones((1, 1))), grid, train, M.MultivariateGaussian(0.1 * np.eye(1)),
                     prior_initial_state=S.Gaussian(np.ones(1), 0.01 * np.eye(1)))
    got = CVISitesTrainer(model, test, **kw).optimize()
    want = CVISitesTrainer(ref, test, **kw).optimize()
    for g, w, name in zip(got[:3], want[:3], ("elbo", "nlpd", "rmse")):
        assert len(g) == len(w) and len(g) >= 4, name
        np.testing.assert_allclose(g, w, rtol=1e-9, err_msg=name)


def test_vi_markov_gp_trainer_matches_oracle_loop():
    from vi_diffusion_processes_b200.markovflow.likelihoods import MultivariateGaussian
    from vi_diffusion_processes_b200.markovflow.models import VariationalMarkovGP
    from vi_diffusion_processes_b200.markovflow.sde import OrnsteinUhlenbeckSDE
    from vi_diffusion_processes_b200.markovflow.sde.sde_utils import Gaussian
    from vi_diffusion_processes_b200.trainers import VIMarkovGPTrainer
    rng = np.random.default_rng(42)
    grid, train, test = _data(rng, 200, 10, 0.01)
    x0 = (np.ones(1), 0.01 * np.eye(1))
    model = VariationalMarkovGP(train, OrnsteinUhlenbeckSDE(decay=1.0, q=np.ones((1, 1))), grid,
                                MultivariateGaussian(0.1 * np.eye(1)), prior_initial_state=Gaussian(*x0))
    ref = _OracleVgp(train, S.OrnsteinUhlenbeckSDE(decay=1.0, q=np.ones((1, 1))), grid,
                     M.MultivariateGaussian(0.1 * np.eye(1)), prior_initial_state=S.Gaussian(*x0))
    for mdl in (model, ref):           # mean-reverting start, as the reference driver (vi_markov_gp.py:83,114)
        mdl.A = mdl.A + 2.0
        mdl.b = mdl.b + 2.0
    kw = dict(q_lr=0.1, x0_lr=0.1, max_itr=6, lr_tol=1e-4, optim_tol=1e-6, warmup_x0_itr=2)
    got = VIMarkovGPTrainer(model, test, **kw).perform_inference()
    want = VIMarkovGPTrainer(ref, test, **kw).perform_inference()
    for g, w, name in zip(got, want, ("elbo", "nlpd", "rmse")):
        assert len(g) == len(w) == 6, name
        np.testing.assert_allclose(g, w, rtol=1e-9, err_msg=name)


def test_device_side_site_loop_equals_host_loop_per_sequence():
    """vdp_cvi_optimize_sites (CUDA-graph WHILE node, convergence / step-size decay decided by a kernel, per sequence)
    against the host loop of cvi_dp_trainer.py:63-101 run on one-sequence engines: same number of updates, ELBO
    traces, decayed step sizes, test-point marginals and final state for every sequence of a batch whose members
    stop at different iterations."""
    from oracle.sde_np import DRIFT_DW
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(77)
    B, N, n_obs, n_test = 3, 500, 9, 5
    grid = np.round(np.arange(N + 1) * 0.01, 6)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 1)) * np.array([0.05, 0.3, 0.6])[:, None, None] + 1.0
    tidx = np.sort(np.stack([rng.choice(N + 1, n_test, replace=False) for _ in range(B)]), axis=1)
    mk = lambda sl: CviEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx[sl], y[sl], [[0.1]], [1.0], [[1e-2]])
    max_itr, tol, lr0 = 12, 0.5, 0.7
    batch = mk(slice(None))
    res = batch.optimize_sites(max_itr, tol, lr0, lr0, test_idx=tidx)
    batch.check_status()
    n_seen = set()
    for b in range(B):
        one = mk(slice(b, b + 1))
        lr = lr0
        elbo = [one.classic_elbo_keep_marginals().item()]
        tm = []
        while len(elbo) - 1 < max_itr:
            elbo.append(one.site_update(lr, lr).item())
            tm.append(one.fx_mus[0, tidx[b], 0].cpu().numpy().copy())
            if len(elbo) > 2 and elbo[-2] > elbo[-1]:
                lr /= 10
            if len(elbo) > 2 and abs(elbo[-2] - elbo[-1]) < tol:
                break
        n = len(elbo) - 1
        n_seen.add(n)
        assert int(res["n_itr"][b]) == n, (b, int(res["n_itr"][b]), n)
        np.testing.assert_allclose(res["elbo_trace"][:n + 1, b].numpy(), elbo, rtol=1e-10)
        np.testing.assert_allclose(float(res["lr_data"][b]), lr, rtol=1e-15)
        np.testing.assert_allclose(res["test_mus"][:n, b, :, 0].numpy(), np.stack(tm), rtol=1e-9, atol=1e-12)
        for name in ("fx_mus", "fx_covs", "d1", "d2"):
            a, w = getattr(batch, name)[b], getattr(one, name)[0]
            assert float((a - w).abs().max()) <= 1e-10 * float(w.abs().max()), (b, name)
        th = one._prior_naturals(one.pA, one.pb)
        for name, t in (("g1", th[0]), ("gd", th[1]), ("gs", th[2])):
            a, w = getattr(batch, name)[b], getattr(one, name)[0]
            assert float((a - w).abs().max()) <= 1e-10 * float((t[0] + w).abs().max()), (b, name)
    assert len(n_seen) > 1, "the sequences were meant to stop at different iterations"
    # a second call on the same engine reuses the captured loop and starts from the decayed step sizes
    res2 = batch.optimize_sites(max_itr, tol, res["lr_data"].numpy(), res["lr_girsanov"].numpy(), test_idx=tidx)
    batch.check_status()
    assert (res2["n_itr"] >= 2).all()
