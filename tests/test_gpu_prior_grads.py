"""
Prior-parameter learning on the GPU against the oracle:
  vdp_cvi_grad_kl_prior_params   <- CVISitesSDE.grad_KL_wrt_prior_params   (variational_cvi_sde.py:495-504)
  vdp_cvi_grad_ve_prior_params   <- CVISitesSDE.grad_VE_wrt_prior_params   (variational_cvi_sde.py:506-518)
  vdp_drift_param_grad           <- VariationalMarkovGP.grad_prior_sde_params (vi_sde.py:457-470)
and the trainers' learning loops (cvi_dp_trainer.py:207-250, vi_markov_gp_trainer.py:163-201).
Tolerance: 1e-9 relative to the largest gradient entry of the sequence (float64 throughout).
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import models_np as M
from oracle import sde_np as S

pytestmark = pytest.mark.gpu

TOL = 1e-9


def _rel(got, want):
    want = np.asarray(want, dtype=np.float64)
    return np.abs(np.asarray(got) - want).max() / max(np.abs(want).max(), 1e-300)


def _cvi_pair(sde, N, n_obs, B, seed, dt=0.01, clip=(-1.0, 1.0), lik=0.3, updates=2, x0=(0.3, 0.5)):
    from vi_diffusion_processes_b200.cvi import CviEngine
    rng = np.random.default_rng(seed)
    D = sde.state_dim
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, D)) * 0.4 + 0.5
    x0_mu, x0_cov = x0[0] * np.ones(D), x0[1] * np.eye(D)
    eng = CviEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, lik * np.eye(D), x0_mu, x0_cov,
                    clip_state_transitions=clip)
    refs = [M.CVISitesSDE(sde, grid, (grid[idx[b]], y[b]), M.MultivariateGaussian(lik * np.eye(D)),
                          prior_initial_state=S.Gaussian(x0_mu, x0_cov), clip_state_transitions=clip)
            for b in range(B)]
    for _ in range(updates):
        eng.site_update(0.5, 0.5)
        for r in refs:
            r.site_update(0.5, 0.5)
    return eng, refs


CASES = {
    "dw": (S.DoubleWellSDE(scale=2.0, c=0.8), (-1.0, 1.0)),
    "ou": (S.OrnsteinUhlenbeckSDE(0.9), (-1.0, 1.0)),
    "benes": (S.BenesSDE(0.7), (-1.0, 1.0)),
    "sine": (S.SineDiffusionSDE(0.3), (-1.0, 1.0)),
    "sqrt": (S.SqrtDiffusionSDE(1.3), (-1.0, 1.0)),
    "mlp": (S.MLPDrift(weights=np.random.default_rng(5).standard_normal(10)), (-1.0, 1.0)),
    "vdp": (S.VanderPolOscillatorSDE(1.3, 1.7, q=np.eye(2)), (-1.0, 1.0)),
    "lorenz": (S.Lorenz63SDE(2.0, 3.0, 1.0, q=np.eye(3)), (-1.0, 1.0)),
    "ou3": (S.OUNdSDE(0.6, q=np.eye(3)), (-1.0, 1.0)),
    "dw_clipped": (S.DoubleWellSDE(scale=2.0, c=0.8), (-1.001, 1.001)),
}


@pytest.mark.parametrize("name", list(CASES))
def test_cvi_prior_param_grads_match_oracle(name):
    sde, clip = CASES[name]
    N = 40 if sde.state_dim < 3 else 24
    eng, refs = _cvi_pair(sde, N=N, n_obs=6, B=3, seed=11, clip=clip)
    fx_before = eng.fx_mus.clone()
    g_kl = eng.grad_kl_prior_params().cpu().numpy()
    assert torch.equal(fx_before, eng.fx_mus)           # the stored marginals are not refreshed
    g_ve = eng.grad_ve_prior_params().cpu().numpy()
    assert torch.equal(fx_before, eng.fx_mus)
    eng.check_status()
    for b, r in enumerate(refs):
        want_kl = r.grad_KL_wrt_prior_params()
        want_ve = r.grad_VE_wrt_prior_params()
        assert g_kl.shape[1] == want_kl.shape[0]
        assert _rel(g_kl[b], want_kl) < TOL, (name, b, g_kl[b], want_kl)
        assert _rel(g_ve[b], want_ve) < TOL, (name, b, g_ve[b], want_ve)
        # side effect of grad_VE: the prior is re-linearised around the stored path (:512)
        assert _rel(eng.pA[b].cpu().numpy(), r.dist_p.state_transitions) < TOL
    if name == "dw_clipped":
        A_lin = refs[0].dist_p_linearized.state_transitions
        assert ((A_lin > clip[1]).any() and (A_lin <= clip[1]).any())   # part of the chain is saturated


def test_cvi_prior_param_grads_long_chain_many_chunks():
    """N large enough for several chunks per sequence in the sensitivity sweeps and several blocks per sequence in
    the parameter reductions."""
    sde = S.DoubleWellSDE(scale=2.0, c=0.8)
    eng, refs = _cvi_pair(sde, N=3000, n_obs=25, B=2, seed=12, dt=0.005, updates=2)
    # re-base the sites on the linearisation around the current posterior (cvi_dp_trainer.py:110-134) so that the
    # re-linearisation inside grad_VE keeps the posterior precision positive definite
    eng.relinearize()
    eng.site_update(0.3, 0.3)
    for r in refs:
        r.relinearize()
        r.site_update(0.3, 0.3)
    g_kl = eng.grad_kl_prior_params().cpu().numpy()
    g_ve = eng.grad_ve_prior_params().cpu().numpy()
    eng.check_status()
    for b, r in enumerate(refs):
        assert _rel(g_kl[b], r.grad_KL_wrt_prior_params()) < TOL
        assert _rel(g_ve[b], r.grad_VE_wrt_prior_params()) < TOL


def test_cvi_prior_param_grads_vdp_d2_long():
    sde = S.VanderPolOscillatorSDE(1.2, 1.5, q=np.eye(2))
    eng, refs = _cvi_pair(sde, N=500, n_obs=40, B=1, seed=13, dt=0.01, lik=0.1)
    g_kl = eng.grad_kl_prior_params().cpu().numpy()
    g_ve = eng.grad_ve_prior_params().cpu().numpy()
    eng.check_status()
    assert _rel(g_kl[0], refs[0].grad_KL_wrt_prior_params()) < TOL
    assert _rel(g_ve[0], refs[0].grad_VE_wrt_prior_params()) < TOL


@pytest.mark.parametrize("name", ["dw", "sine", "vdp", "mlp"])
def test_vgp_grad_prior_sde_params_match_oracle(name):
    from vi_diffusion_processes_b200.vgp import VgpEngine
    sde, _ = CASES[name]
    rng = np.random.default_rng(21)
    D, N, n_obs, B = sde.state_dim, 60, 7, 2
    grid = np.round(np.arange(N + 1) * 0.01, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, D)) * 0.4 + 0.5
    x0_mu, x0_cov = 0.3 * np.ones(D), 0.5 * np.eye(D)
    eng = VgpEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, 0.3 * np.eye(D), x0_mu, x0_cov)
    refs = [M.VariationalMarkovGP((grid[idx[b]], y[b]), sde, grid, M.MultivariateGaussian(0.3 * np.eye(D)),
                                  prior_initial_state=S.Gaussian(x0_mu, x0_cov)) for b in range(B)]
    for _ in range(3):
        eng.sweep(0.3)
        for r in refs:
            r.sweep(0.3)
    g = eng.grad_prior_sde_params().cpu().numpy()
    eng.check_status()
    for b, r in enumerate(refs):
        assert _rel(g[b], r.grad_prior_sde_params()) < TOL, (name, b)


# ---- trainer loops with learning


class _OracleCvi(M.CVISitesSDE):
    def transform_sites_to_linearized_prior(self):
        self.girsanov_sites = M.tranform_girsanov_sites(self.girsanov_sites, self.dist_p, self.dist_p_linearized)
        self.dist_p = self.dist_p_linearized

    def relinearize_and_transform_sites(self):
        last = self.dist_p
        self.set_linearized_prior()
        self.girsanov_sites = M.tranform_girsanov_sites(self.girsanov_sites, last, self.dist_p)

    def grad_KL_wrt_prior_params(self):
        g = super().grad_KL_wrt_prior_params()
        return [g[i] for i in self.prior_sde.trainable_indices]

    def grad_VE_wrt_prior_params(self):
        g = super().grad_VE_wrt_prior_params()
        return [g[i] for i in self.prior_sde.trainable_indices]


class _OracleVgp(M.VariationalMarkovGP):
    m = property(lambda s: s._last[0])
    S = property(lambda s: s._last[1])

    def sweep(self, lr, update_x0=False, x0_lr=0.1):
        e = super().sweep(lr, update_x0=update_x0, x0_lr=x0_lr)
        self._last = self.forward_pass
        return e

    def grad_prior_sde_params(self):
        g = super().grad_prior_sde_params()
        return [g[i] for i in self.prior_sde.trainable_indices]


def _data(rng, N, n_obs, dt, D=1):
    grid = np.round(np.arange(N + 1) * dt, 6)
    pick = np.sort(rng.choice(np.arange(1, N + 1), 2 * n_obs, replace=False))
    tr, te = pick[::2], pick[1::2]
    y = rng.normal(size=(2 * n_obs, D)) * 0.2 + 1.0
    return grid, (grid[tr], y[::2]), (grid[te], y[1::2])


@pytest.mark.parametrize("which", ["dw_scale", "ou", "vdp"])
def test_cvi_sites_trainer_with_prior_learning_matches_oracle_loop(which):
    from vi_diffusion_processes_b200.markovflow import sde as msde
    from vi_diffusion_processes_b200.markovflow.likelihoods import MultivariateGaussian
    from vi_diffusion_processes_b200.markovflow.models import CVISitesSDE
    from vi_diffusion_processes_b200.markovflow.sde.sde_utils import Gaussian
    from vi_diffusion_processes_b200.trainers import CVISitesTrainer
    rng = np.random.default_rng(43)
    if which == "dw_scale":     # configs/prior_sde/dw.yaml style: only the scale is learnt
        D = 1
        mine = msde.DoubleWellSDE(q=np.ones((1, 1)), scale_trainable=True, scale=2.0, c=1.0)
        ref_sde = S.DoubleWellSDE(q=np.ones((1, 1)), scale=2.0, c=1.0)
        ref_sde.trainable = [0]
    elif which == "ou":
        D = 1
        mine = msde.OrnsteinUhlenbeckSDE(decay=1.2, q=np.ones((1, 1)), trainable=True)
        ref_sde = S.OrnsteinUhlenbeckSDE(decay=1.2, q=np.ones((1, 1)))
    else:                       # BASELINE config 3: Van der Pol with drift-parameter learning
        D = 2
        mine = msde.VanderPolOscillatorSDE(a=1.0, tau=1.0, q=np.eye(2), trainable=True)
        ref_sde = S.VanderPolOscillatorSDE(a=1.0, tau=1.0, q=np.eye(2))
    grid, train, test = _data(rng, 200 if D == 1 else 120, 12, 0.01, D)
    x0 = (np.ones(D), 0.01 * np.eye(D))
    kw = dict(max_itr=2, optim_tol=1e-3, max_itr_sites_optim=3, girsanov_sites_lr=0.5, data_sites_lr=0.5,
              learn_prior_sde=True, prior_sde_lr=1e-2, learning_max_itr=3, learning_tol=1e-6)
    model = CVISitesSDE(mine, grid, train, MultivariateGaussian(0.1 * np.eye(D)), prior_initial_state=Gaussian(*x0))
    ref = _OracleCvi(ref_sde, grid, train, M.MultivariateGaussian(0.1 * np.eye(D)),
                     prior_initial_state=S.Gaussian(*x0))
    got = CVISitesTrainer(model, test, **kw).optimize()
    want = CVISitesTrainer(ref, test, **kw).optimize()
    for g, w, name in zip(got[:3], want[:3], ("elbo", "nlpd", "rmse")):
        assert len(g) == len(w) and len(g) >= 8, (name, len(g), len(w))
        np.testing.assert_allclose(g, w, rtol=1e-9, err_msg=name)
    assert got[3].keys() == want[3].keys() and len(got[3]) >= 1
    for k in got[3]:
        np.testing.assert_allclose(got[3][k], want[3][k], rtol=1e-9)
        assert len(got[3][k]) >= 4 and got[3][k][0] != got[3][k][-1]       # the parameters did move
    np.testing.assert_allclose(list(mine.params), np.asarray(ref_sde.params), rtol=1e-9)


def test_vi_markov_gp_trainer_with_prior_learning_matches_oracle_loop():
    from vi_diffusion_processes_b200.markovflow import sde as msde
    from vi_diffusion_processes_b200.markovflow.likelihoods import MultivariateGaussian
    from vi_diffusion_processes_b200.markovflow.models import VariationalMarkovGP
    from vi_diffusion_processes_b200.markovflow.sde.sde_utils import Gaussian
    from vi_diffusion_processes_b200.trainers import VIMarkovGPTrainer
    rng = np.random.default_rng(44)
    grid, train, test = _data(rng, 150, 10, 0.01)
    x0 = (np.ones(1), 0.01 * np.eye(1))
    mine = msde.DoubleWellSDE(q=np.ones((1, 1)), scale_trainable=True, c_trainable=True, scale=2.0, c=1.0)
    ref_sde = S.DoubleWellSDE(q=np.ones((1, 1)), scale=2.0, c=1.0)
    model = VariationalMarkovGP(train, mine, grid, MultivariateGaussian(0.1 * np.eye(1)),
                                prior_initial_state=Gaussian(*x0))
    ref = _OracleVgp(train, ref_sde, grid, M.MultivariateGaussian(0.1 * np.eye(1)),
                     prior_initial_state=S.Gaussian(*x0))
    for mdl in (model, ref):
        mdl.A = mdl.A + 2.0
        mdl.b = mdl.b + 2.0
    kw = dict(q_lr=0.1, x0_lr=0.1, max_itr=2, lr_tol=1e-4, optim_tol=1e-9, warmup_x0_itr=2, learn_prior_sde=True,
              prior_sde_lr=1e-2, learning_max_itr=3, learning_tol=1e-9)
    tr_a, tr_b = VIMarkovGPTrainer(model, test, **kw), VIMarkovGPTrainer(ref, test, **kw)
    tr_a.max_itr = tr_b.max_itr = 3      # inner sweeps per inference round (the reference shares max_itr)
    got, want = tr_a.optimize(), tr_b.optimize()
    for g, w, name in zip(got[:3], want[:3], ("elbo", "nlpd", "rmse")):
        assert len(g) == len(w) and len(g) >= 6, (name, len(g), len(w))
        np.testing.assert_allclose(g, w, rtol=1e-9, err_msg=name)
    for k in got[3]:
        np.testing.assert_allclose(got[3][k], want[3][k], rtol=1e-9)
    np.testing.assert_allclose(list(mine.params), np.asarray(ref_sde.params), rtol=1e-9)
