#!/usr/bin/env python
"""
Generate tests/golden/*.npz: frozen outputs of the oracle (oracle/models_np.py) on the data files the reference
ships (docs/diffusion_processes/data.zip) for BASELINE.json configs C1-C3.

The reference itself cannot be imported here (TensorFlow / gpflow / banded_matrices are not installable in this
image), so these vectors pin the ORACLE, not TensorFlow: they guard the oracle against regressions (CPU test)
and give the CUDA path fixed targets that do not depend on re-running the oracle (GPU test).  Run in the build
container only (it reads /root/reference); the GPU box uses the committed files.

    python tests/golden/make_golden.py
"""
import io
import os
import sys
import zipfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import models_np as M  # noqa: E402
from oracle import sde_np as S  # noqa: E402

DATA_ZIP = "/root/reference/docs/diffusion_processes/data.zip"
OUT = os.path.join(ROOT, "tests", "golden")

CASES = {
    # name: (file in data.zip, prior SDE of the reference's config, x0 prior (mean, cov), lr, iterations)
    "c1_ou": ("data/linear/15/0.npz", lambda d: S.OrnsteinUhlenbeckSDE(1.2, np.ones((1, 1)) * float(d["Q"])),
              None, 1.0),                                    # configs/cvi_linear_process.yaml, prior_sde/ou.yaml
    "c2_dw": ("data/dw/3/0.npz", lambda d: S.DoubleWellSDE(np.ones((1, 1)) * float(d["Q"])),
              (1.0, 1e-4), 0.5),                             # configs/cvi_non_linear_process.yaml
    "c3_vdp": ("data/vanderpol/33/0.npz", lambda d: S.VanderPolOscillatorSDE(1.0, 1.0, np.asarray(d["Q"])),
               None, 0.5),                                   # configs/prior_sde/vanderpol.yaml (a = tau = 1)
}


def load(name):
    with zipfile.ZipFile(DATA_ZIP) as z:
        return np.load(io.BytesIO(z.read(name)), allow_pickle=True)


def main():
    os.makedirs(OUT, exist_ok=True)
    for case, (fname, make_sde, x0, lr) in CASES.items():
        d = load(fname)
        grid = np.asarray(d["time_grid"], dtype=np.float64)
        obs_t = np.asarray(d["observation_grid"], dtype=np.float64)
        y = np.asarray(d["observations"], dtype=np.float64)
        D = y.shape[-1]
        sde = make_sde(d)
        sigma = float(d["sigma"])
        lik = M.MultivariateGaussian(sigma * np.eye(D))
        if x0 is None:
            x0m = np.asarray(d["x0"], dtype=np.float64).reshape(D)
            prior = S.Gaussian(x0m, 1e-2 * np.eye(D))
        else:
            prior = S.Gaussian(np.full(D, x0[0]), x0[1] * np.eye(D))
        out = dict(time_grid=grid, obs_t=obs_t, obs_y=y, sigma=sigma, q=sde.q, drift_id=sde.drift_id,
                   drift_params=np.asarray(sde.params, dtype=np.float64), x0_mu=prior.mu, x0_cov=prior.cov, lr=lr)
        # CVI-DP: three site updates with a re-linearisation after the second (cvi_dp_trainer.py:71-77, 110-134)
        cvi = M.CVISitesSDE(sde, grid, (obs_t, y), lik, prior_initial_state=prior)
        elbos = []
        for it in range(3):
            elbos.append(cvi.site_update(lr, lr))
            if it == 1:
                cvi.relinearize()
        out.update(cvi_elbo=np.asarray(elbos), cvi_g1=cvi.girsanov_sites.nat1,
                   cvi_gd=cvi.girsanov_sites.nat2.block_diagonal, cvi_gs=cvi.girsanov_sites.nat2.block_sub_diagonal,
                   cvi_d1=cvi.data_sites.nat1, cvi_d2=cvi.data_sites.nat2, cvi_m=cvi.fx_mus, cvi_S=cvi.fx_covs,
                   obs_idx=cvi.obs_sites_indices)
        # rounding floor of the float64 reference route itself: the same three updates re-evaluated in 80-bit
        # long double (oracle/scalar_d1.py, D = 1 only).  The site natural parameters are ill-conditioned
        # functions of the marginals (1 / (S_{k+1} - A S_k A^T)): two float64 evaluations that round
        # differently cannot agree better than this, so the GPU test allows max(1e-9, 3 x floor).
        floor = np.zeros(3)
        if D == 1:
            from oracle import scalar_d1
            runs = []
            for dtype in (np.float64, np.longdouble):
                m1 = scalar_d1.BatchedCVISitesSDE1D(sde.drift_id, np.asarray(sde.params), float(sde.q[0, 0]), grid,
                                                    cvi.obs_sites_indices[None], y[None, :, 0], sigma,
                                                    float(prior.mu[0]), float(prior.cov[0, 0]), dtype=dtype)
                for it in range(3):
                    m1.site_update(lr, lr)
                    if it == 1:
                        m1.relinearize()
                runs.append(m1)
            f64, f80 = runs
            floor = np.array([float(np.abs(getattr(f64, k) - getattr(f80, k)).max() / np.abs(getattr(f80, k)).max())
                              for k in ("g1", "gd", "gs")])
        out.update(site_floor_f64_vs_f80=floor)
        # VDP: three fixed-point sweeps from a mean-reverting start (vi_markov_gp.py:83, trainer :62-73)
        vgp_stab = case != "c1_ou"   # sigma = 0.01 makes psi jump by 5000 at every observation: the reference's stabilize_system
        vgp = M.VariationalMarkovGP((obs_t, y), sde, grid, lik, prior_initial_state=prior, stabilize_system=vgp_stab)
        vgp.A = vgp.A + 2.0 * np.eye(D)
        vgp.b = vgp.b + 2.0 * prior.mu
        vgp_lr = 0.1 if case == "c1_ou" else 0.01   # the noisier the jumps, the smaller the stable step
        velbo = [vgp.sweep(vgp_lr, update_x0=it >= 1, x0_lr=vgp_lr) for it in range(3)]
        m, Sg = vgp.forward_pass
        out.update(vgp_elbo=np.asarray(velbo), vgp_A=vgp.A, vgp_b=vgp.b, vgp_lam=vgp.lambda_lagrange,
                   vgp_psi=vgp.psi_lagrange, vgp_m=m, vgp_S=Sg, vgp_q_mean=vgp.q_initial_mean,
                   vgp_q_chol=vgp.q_initial_chol, vgp_lr=vgp_lr, vgp_stabilize=vgp_stab)
        path = os.path.join(OUT, case + ".npz")
        np.savez_compressed(path, **out)
        print(case, "grid", grid.shape[0], "obs", y.shape[0], "cvi elbo", elbos, "vgp elbo", velbo,
              "->", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
