#!/usr/bin/env python
"""
Dump the REAL reference's outputs for the golden cases.

NOT RUNNABLE IN THE BUILD IMAGE: it imports the reference itself (AaltoML/vi-diffusion-processes `markovflow`,
TensorFlow 2.4, gpflow 2.2, tensorflow-probability 0.11, banded-matrices 0.0.6, Python 3.7), none of which can be
installed here.  It is the one artefact that turns the parity status from "pinned to the oracle" into "pinned to
the reference": run it once on a box that has the reference's environment,

    cd <checkout of AaltoML/vi-diffusion-processes> && poetry install
    poetry run python /path/to/this/repo/baseline/run_tf_reference.py --out /path/to/this/repo/tests/golden_tf

and then `python -m pytest tests/test_golden_tf.py` compares the oracle (CPU) and the CUDA path (GPU) with those
files at 1e-9.  It performs exactly the steps of tests/golden/make_golden.py -- same members of the reference's
docs/diffusion_processes/data.zip, same models, same three CVI-DP site updates with a re-linearisation after the
second (docs/diffusion_processes/cvi_dp_trainer.py:71-77, :110-134) and the same three VDP sweeps
(vi_markov_gp_trainer.py:62-73) -- through the reference's own classes, and writes the same keys with a `tf_` prefix.
Nothing in this repository imports this file.
"""
import argparse
import io
import os
import zipfile

import numpy as np

CASES = {
    # name: (member of data.zip, prior SDE factory, (x0 mean, x0 cov) or None = data x0 with cov 1e-2, site lr)
    "c1_ou": ("data/linear/15/0.npz", "ou", None, 1.0),
    "c2_dw": ("data/dw/3/0.npz", "dw", (1.0, 1e-4), 0.5),
    "c3_vdp": ("data/vanderpol/33/0.npz", "vdp", None, 0.5),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default=".", help="checkout of AaltoML/vi-diffusion-processes")
    ap.add_argument("--out", default="golden_tf")
    args = ap.parse_args()

    import tensorflow as tf
    from gpflow.probability_distributions import Gaussian
    from markovflow.likelihoods import MultivariateGaussian
    from markovflow.models.variational_cvi_sde import CVISitesSDE
    from markovflow.models.vi_sde import VariationalMarkovGP
    from markovflow.sde.sde import DoubleWellSDE, OrnsteinUhlenbeckSDE, VanderPolOscillatorSDE
    from markovflow.sde.sde_utils import tranform_girsanov_sites
    from tensorflow_probability import distributions

    f64 = tf.float64
    os.makedirs(args.out, exist_ok=True)
    data_zip = os.path.join(args.reference, "docs", "diffusion_processes", "data.zip")

    def load(name):
        with zipfile.ZipFile(data_zip) as z:
            return np.load(io.BytesIO(z.read(name)), allow_pickle=True)

    def make_sde(kind, d):
        D = np.asarray(d["observations"]).shape[-1]
        if kind == "ou":
            return OrnsteinUhlenbeckSDE(decay=1.2, q=tf.ones((1, 1), dtype=f64) * float(d["Q"]))
        if kind == "dw":
            return DoubleWellSDE(q=tf.ones((1, 1), dtype=f64) * float(d["Q"]))
        return VanderPolOscillatorSDE(a=1.0, tau=1.0, q=tf.convert_to_tensor(np.asarray(d["Q"]).reshape(D, D), dtype=f64))

    def relinearize(model):
        """cvi_dp_trainer.py:110-113 and :128-134."""
        model.girsanov_sites = tranform_girsanov_sites(model.girsanov_sites, model.dist_p, model.dist_p_linearized)
        model.dist_p = model.dist_p_linearized
        dist_p_last = model.dist_p
        model.set_linearized_prior()
        model.girsanov_sites = tranform_girsanov_sites(model.girsanov_sites, dist_p_last, model.dist_p)

    for case, (member, kind, x0, lr) in CASES.items():
        d = load(member)
        grid = tf.convert_to_tensor(np.asarray(d["time_grid"], dtype=np.float64))
        obs_t = tf.convert_to_tensor(np.asarray(d["observation_grid"], dtype=np.float64))
        y = tf.convert_to_tensor(np.asarray(d["observations"], dtype=np.float64))
        D = y.shape[-1]
        sigma = float(d["sigma"])
        if x0 is None:
            mu0, cov0 = np.asarray(d["x0"], dtype=np.float64).reshape(D), 1e-2 * np.eye(D)
        else:
            mu0, cov0 = np.full(D, x0[0]), x0[1] * np.eye(D)
        out = {}

        # ---- CVI-DP: update_data_sites, update_girsanov_sites, classic_elbo  x 3
        sde = make_sde(kind, d)
        cvi = CVISitesSDE(sde, grid, (obs_t, y), MultivariateGaussian(sigma * tf.eye(D, dtype=f64)),
                          prior_initial_state=Gaussian(mu=tf.convert_to_tensor(mu0), cov=tf.convert_to_tensor(cov0)))
        elbos = []
        for it in range(3):
            cvi.update_data_sites(lr)
            cvi.update_girsanov_sites(lr)
            elbos.append(float(cvi.classic_elbo()))
            if it == 1:
                relinearize(cvi)
        out.update(tf_cvi_elbo=np.asarray(elbos), tf_cvi_g1=cvi.girsanov_sites.nat1.numpy(),
                   tf_cvi_gd=cvi.girsanov_sites.nat2.block_diagonal.numpy(),
                   tf_cvi_gs=cvi.girsanov_sites.nat2.block_sub_diagonal.numpy(),
                   tf_cvi_d1=cvi.data_sites.nat1.numpy(), tf_cvi_d2=cvi.data_sites.nat2.numpy(),
                   tf_cvi_m=cvi.fx_mus.numpy(), tf_cvi_S=cvi.fx_covs.numpy())
        # intermediate quantities whose third-party semantics the oracle restates from memory (gpflow mvnquad node
        # order, banded_matrices band conventions, TensorFlow's Cholesky gradient): one value of each
        q = cvi.dist_q
        out.update(tf_q_A=q.state_transitions.numpy(), tf_q_b=q.state_offsets.numpy(),
                   tf_q_cholQ=q.cholesky_process_covariances.numpy())
        kl_grads = cvi.grad_kl_wrt_exp_param()
        out.update(tf_klgrad_e1=kl_grads[0].numpy(), tf_klgrad_ed=kl_grads[1].numpy(), tf_klgrad_es=kl_grads[2].numpy(),
                   tf_kl_q_p=float(cvi.KL_q_p()))

        # ---- VDP: three fixed-point sweeps from a mean-reverting start (vi_markov_gp.py:83; trainer :62-73)
        stab = case != "c1_ou"
        vlr = 0.1 if case == "c1_ou" else 0.01
        prior_x0 = distributions.MultivariateNormalFullCovariance(loc=tf.convert_to_tensor(mu0),
                                                                   covariance_matrix=tf.convert_to_tensor(cov0))
        vgp = VariationalMarkovGP((obs_t, y), make_sde(kind, d), grid, MultivariateGaussian(sigma * tf.eye(D, dtype=f64)),
                                  prior_initial_state=prior_x0, stabilize_system=stab)
        vgp.A = vgp.A + 2.0 * tf.eye(D, dtype=f64)
        vgp.b = vgp.b + 2.0 * tf.convert_to_tensor(mu0)
        velbo = []
        for it in range(3):
            m, S = vgp.forward_pass
            vgp.update_lagrange(m, S)
            vgp.update_param(m, S, lr=vlr)
            if it >= 1:
                vgp.update_initial_statistics(lr=vlr)
            m, S = vgp.forward_pass
            velbo.append(float(vgp.elbo(m, S)))
        out.update(tf_vgp_elbo=np.asarray(velbo), tf_vgp_A=vgp.A.numpy(), tf_vgp_b=vgp.b.numpy(),
                   tf_vgp_lam=vgp.lambda_lagrange.numpy(), tf_vgp_psi=vgp.psi_lagrange.numpy(), tf_vgp_m=m.numpy(),
                   tf_vgp_S=S.numpy(), tf_vgp_q_mean=vgp.q_initial_mean.numpy(), tf_vgp_q_chol=vgp.q_initial_chol.numpy())
        path = os.path.join(args.out, case + "_tf.npz")
        np.savez_compressed(path, **out)
        print(case, "cvi elbo", elbos, "vgp elbo", velbo, "->", path)


if __name__ == "__main__":
    main()
