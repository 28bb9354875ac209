"""Checkpoint round trips on the GPU models (cvi_dp.py:140-149, vi_markov_gp.py:171-174) and the data-set generator
(generate_data.py:25-50 scheme over the Euler-Maruyama kernel)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _models(D=1):
    from vi_diffusion_processes_b200.markovflow.likelihoods import MultivariateGaussian
    from vi_diffusion_processes_b200.markovflow.models import CVISitesSDE, VariationalMarkovGP
    from vi_diffusion_processes_b200.markovflow.sde import DoubleWellSDE
    from vi_diffusion_processes_b200.markovflow.sde.sde_utils import Gaussian
    rng = np.random.default_rng(3)
    grid = np.round(np.arange(201) * 0.01, 2)
    idx = np.sort(rng.choice(np.arange(1, 201), 10, replace=False))
    data = (grid[idx], rng.normal(size=(10, 1)) * 0.3 + 1.0)
    mk_cvi = lambda: CVISitesSDE(DoubleWellSDE(q=np.ones((1, 1))), grid, data, MultivariateGaussian(0.1 * np.eye(1)),
                                 prior_initial_state=Gaussian(np.ones(1), 0.01 * np.eye(1)))
    mk_vgp = lambda: VariationalMarkovGP(data, DoubleWellSDE(q=np.ones((1, 1))), grid,
                                         MultivariateGaussian(0.1 * np.eye(1)),
                                         prior_initial_state=Gaussian(np.ones(1), 0.01 * np.eye(1)))
    return mk_cvi, mk_vgp


def test_cvi_checkpoint_round_trip(tmp_path):
    from vi_diffusion_processes_b200 import io as vio
    mk_cvi, _ = _models()
    a = mk_cvi()
    for _ in range(3):
        a.site_update(0.5, 0.5)
    vio.save_cvi_model(str(tmp_path), a)
    d = np.load(tmp_path / "cvi_model.npz")
    assert set(d.keys()) == set(vio.CVI_MODEL_KEYS)
    assert d["girsanov_sites_nat2_subdiag"].shape == (200, 1, 1) and d["data_sites_nat1"].shape == (10, 1)
    b = vio.load_cvi_model(str(tmp_path / "cvi_model.npz"), mk_cvi())
    assert float(a.classic_elbo()) == float(b.classic_elbo())       # both rebuilt from the same site tensors
    # a's stored marginals are those of its last site update, which carried the natural parameters theta along and
    # held the girsanov sites inside them; b's come from theta rebuilt as nat(p) + g + data sites out of the saved
    # sites: the same numbers up to the rounding of that sum
    for x, r in ((a.fx_mus, b.fx_mus), (a.fx_covs, b.fx_covs)):
        assert float((x - r).abs().max()) <= 1e-12 * float(r.abs().max())
    # training continues from the restored sites as it does from the live model
    ea, eb = float(a.site_update(0.5, 0.5)), float(b.site_update(0.5, 0.5))
    assert abs(ea - eb) <= 1e-11 * abs(eb)


def test_vi_gp_checkpoint_round_trip(tmp_path):
    from vi_diffusion_processes_b200 import io as vio
    _, mk_vgp = _models()
    a = mk_vgp()
    a.A = a.A + 2.0
    a.b = a.b + 2.0
    for i in range(4):
        a.sweep(0.1, update_x0=i > 1)
    vio.save_vi_gp_model(str(tmp_path), a)
    d = np.load(tmp_path / "vi_gp_model.npz")
    assert set(d.keys()) == set(vio.VI_GP_MODEL_KEYS) and d["A"].shape == (200, 1, 1) and d["x0_S"].shape == (1, 1)
    b = vio.load_vi_gp_model(str(tmp_path / "vi_gp_model.npz"), mk_vgp())
    np.testing.assert_allclose(float(a.elbo()), float(b.elbo()), rtol=1e-13)
    np.testing.assert_allclose(float(a.sweep(0.1)), float(b.sweep(0.1)), rtol=1e-12)


def test_generate_dataset_and_run_outputs(tmp_path):
    from vi_diffusion_processes_b200 import io as vio
    from vi_diffusion_processes_b200.markovflow.sde import VanderPolOscillatorSDE
    sde = VanderPolOscillatorSDE(a=2.0, tau=5.0, q=np.eye(2))
    latent, og, ov, tg, tgrid, tv = vio.generate_dataset(sde, 0.0, 5.0, np.zeros(2), 0.01, 0.1, 40, seed=33)
    assert tg.shape == (501,) and latent.shape == (501, 2) and ov.shape == (40, 2) and tv.shape == (8, 2)
    assert np.isin(og, tg).all() and np.isin(tgrid, tg).all() and np.all(np.diff(og) > 0)
    assert np.abs(ov - latent[np.where(np.equal(tg[:, None], og))[0]]).max() < 0.6      # noise_std = 0.1
    vio.save_exp_data(tmp_path / "33.npz", "vanderpol", np.eye(2), np.zeros((1, 2)), 0.1, latent, ov, og, tg, tv, tgrid)
    Q, x0, noise, lat2, obs, tg2, test = vio.load_exp_data(str(tmp_path / "33.npz"))
    assert np.array_equal(lat2, latent) and np.array_equal(obs[1], ov)
    vio.save_run_outputs(str(tmp_path), "cvi", latent, np.ones((501, 2, 2)), tg, [1.0, 2.0], [0.1], [0.2], {"0": [1.0, 1.1]})
    assert set(np.load(tmp_path / "posteriors.npz").keys()) == {"cvi_m", "cvi_S", "time_grid"}
    assert set(np.load(tmp_path / "training_statistics.npz").keys()) == {"elbo", "nlpd", "rmse"}
    assert np.load(tmp_path / "learnt_prior_params.npz")["0"].shape == (2,)
