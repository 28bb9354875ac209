"""
On-disk formats of the reference's experiment drivers, so that data sets and checkpoints move
between the reference and this build unchanged:

* data set ``.npz``      -- ``docs/diffusion_processes/generate_data.py:128-140`` (writer) and
  ``exp_dp_utils.py:108-124`` ``load_exp_data`` (reader): keys ``sde, decay, Q, x0, sigma, latent_process,
  observations, observation_grid, time_grid, test_observations, test_grid``;
* CVI-DP checkpoint       -- ``cvi_dp.py:140-149`` ``cvi_model.npz``: ``data_sites_nat1, data_sites_nat2,
  girsanov_sites_nat1, girsanov_sites_nat2_diag, girsanov_sites_nat2_subdiag``;
* VDP checkpoint          -- ``vi_markov_gp.py:171-174`` ``vi_gp_model.npz``: ``A, b, lambda_lagrange, psi_lagrange,
  x0_m, x0_S``;
* run outputs             -- ``posteriors.npz`` (``cvi_m / cvi_S`` or ``vi_m / vi_S``, ``time_grid``),
  ``training_statistics.npz`` (``elbo, nlpd, rmse``), ``learnt_prior_params.npz`` (one array per trainable
  parameter index) -- ``cvi_dp.py:136-155``, ``vi_markov_gp.py:166-180``.

``generate_dataset`` is ``get_observations`` of ``generate_data.py:25-50`` with the path simulated by the batched
Euler-Maruyama kernel (``vdp_euler_maruyama``); the draws come from NumPy's generator, not TensorFlow's, so a seed
does not reproduce the reference's files -- the schema and the sampling scheme are what is mirrored.
"""
from __future__ import annotations

import os

import numpy as np

DATASET_KEYS = ("sde", "decay", "Q", "x0", "sigma", "latent_process", "observations", "observation_grid", "time_grid",
                "test_observations", "test_grid")
CVI_MODEL_KEYS = ("data_sites_nat1", "data_sites_nat2", "girsanov_sites_nat1", "girsanov_sites_nat2_diag",
                  "girsanov_sites_nat2_subdiag")
VI_GP_MODEL_KEYS = ("A", "b", "lambda_lagrange", "psi_lagrange", "x0_m", "x0_S")


def _np(x):
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


# ---------------------------------------------------------------------------------------------- data sets


def save_exp_data(path, sde: str, Q, x0, sigma, latent_process, observations, observation_grid, time_grid,
                  test_observations, test_grid, decay: float = 0.5):
    """generate_data.py:136-138"""
    np.savez(path, sde=sde, decay=decay, Q=_np(Q), x0=_np(x0), sigma=sigma, latent_process=_np(latent_process),
             observations=_np(observations), observation_grid=_np(observation_grid), time_grid=_np(time_grid),
             test_observations=_np(test_observations), test_grid=_np(test_grid))


def load_exp_data(data_path: str):
    """exp_dp_utils.py:108-124 -> (Q, x0, noise_stddev [1,1], latent_process, (obs_grid, obs), time_grid,
    (test_grid, test_obs)) as float64 NumPy arrays."""
    data = np.load(data_path)
    Q = data["Q"]
    x0 = data["x0"]
    noise_stddev = data["sigma"].reshape((1, 1))
    latent_process = data["latent_process"]
    observations = (data["observation_grid"], data["observations"])
    test_obs = (data["test_grid"], data["test_observations"])
    return Q, x0, noise_stddev, latent_process, observations, data["time_grid"], test_obs


def generate_dataset(sde, t0: float, t1: float, x0, dt: float, noise_std: float, num_observations: int, seed: int = 0,
                     device="cuda"):
    """
    get_observations (generate_data.py:25-50): simulate the SDE on the dense grid, draw ``num_observations`` training
    and ``0.2 num_observations`` test points on grid points without replacement, add Gaussian noise.
    ``sde`` is a markovflow-mirror SDE (drift_id, params, q).  Returns the tuple of the reference:
    (simulated_vals, observation_grid, observations, time_grid, test_observation_grid, test_observations).
    """
    import torch

    from . import ops
    rng = np.random.default_rng(seed)
    n_decimals = str(dt)[::-1].find('.')
    time_grid = np.round(np.arange(t0, t1 + dt, dt, dtype=np.float64), decimals=n_decimals)
    D = sde.state_dim
    x0 = np.asarray(x0, dtype=np.float64).reshape(1, D)
    noise = torch.as_tensor(rng.standard_normal((1, time_grid.shape[0] - 1, D)), device=device)
    path = ops.euler_maruyama(sde.drift_id, sde.params, sde.q, torch.as_tensor(x0, device=device),
                              torch.as_tensor(time_grid, device=device), noise=noise)[0].cpu().numpy()

    def pick(count):
        grid = np.sort(rng.choice(time_grid, count, replace=False)).reshape((-1,))
        idx = np.where(np.equal(time_grid[..., None], grid))[0]
        return grid, path[idx] + noise_std * rng.standard_normal((idx.shape[0], D))

    obs_grid, obs = pick(num_observations)
    test_grid, test_obs = pick(int(.2 * num_observations))
    return path, obs_grid, obs, time_grid, test_grid, test_obs


# ---------------------------------------------------------------------------------------------- checkpoints


def save_cvi_model(output_dir: str, model):
    """cvi_dp.py:140-146: the sites of a CVISitesSDE."""
    g = model.girsanov_sites
    np.savez(os.path.join(output_dir, "cvi_model.npz"), data_sites_nat1=_np(model.data_sites.nat1),
             data_sites_nat2=_np(model.data_sites.nat2), girsanov_sites_nat1=_np(g.nat1),
             girsanov_sites_nat2_diag=_np(g.nat2.block_diagonal),
             girsanov_sites_nat2_subdiag=_np(g.nat2.block_sub_diagonal))


def load_cvi_model(path: str, model):
    """Restore the sites written by ``save_cvi_model`` (or by the reference's cvi_dp.py) into a CVISitesSDE and
    refresh its stored marginals from them (``dist_q.marginals``)."""
    import torch
    d = np.load(path)
    eng = model._engine
    dev, f64 = eng.device, torch.float64
    put = lambda name, arr: setattr(eng, name, torch.as_tensor(np.asarray(arr), dtype=f64).to(dev).reshape(
        getattr(eng, name).shape).contiguous())
    put("d1", d["data_sites_nat1"])
    put("d2", d["data_sites_nat2"])
    put("g1", d["girsanov_sites_nat1"])
    put("gd", d["girsanov_sites_nat2_diag"])
    put("gs", d["girsanov_sites_nat2_subdiag"])
    eng.classic_elbo()          # rebuilds dist_q and writes its marginals into fx_mus / fx_covs
    return model


def save_vi_gp_model(output_dir: str, model):
    """vi_markov_gp.py:171-174: the variational parameters of a VariationalMarkovGP."""
    L = _np(model.q_initial_chol)
    np.savez(os.path.join(output_dir, "vi_gp_model.npz"), A=_np(model.A), b=_np(model.b),
             lambda_lagrange=_np(model.lambda_lagrange), psi_lagrange=_np(model.psi_lagrange),
             x0_m=_np(model.q_initial_mean), x0_S=L @ L.T)


def load_vi_gp_model(path: str, model):
    d = np.load(path)
    model.A, model.b = d["A"], d["b"]
    model.lambda_lagrange, model.psi_lagrange = d["lambda_lagrange"], d["psi_lagrange"]
    model.q_initial_mean = d["x0_m"]
    model.q_initial_chol = np.linalg.cholesky(np.atleast_2d(d["x0_S"]))
    return model


def save_run_outputs(output_dir: str, kind: str, m, S, time_grid, elbo, nlpd, rmse, prior_params=None):
    """posteriors.npz / training_statistics.npz / learnt_prior_params.npz (cvi_dp.py:136-155,
    vi_markov_gp.py:166-180); ``kind`` = "cvi" or "vi" selects the key prefix of the posterior."""
    assert kind in ("cvi", "vi")
    np.savez(os.path.join(output_dir, "posteriors.npz"), **{f"{kind}_m": _np(m), f"{kind}_S": _np(S),
                                                            "time_grid": _np(time_grid)})
    np.savez(os.path.join(output_dir, "training_statistics.npz"), elbo=np.asarray(elbo), nlpd=np.asarray(nlpd),
             rmse=np.asarray(rmse))
    if prior_params:
        np.savez(os.path.join(output_dir, "learnt_prior_params.npz"),
                 **{str(k): np.asarray(v) for k, v in prior_params.items()})
