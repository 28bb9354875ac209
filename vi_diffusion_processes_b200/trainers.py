"""
Mirrors of the reference's trainer loops around the hot path:

* :class:`CVISitesTrainer`  -- ``docs/diffusion_processes/cvi_dp_trainer.py:18-160``: site updates until the ELBO
  converges under a fixed linearised prior (learning-rate decay when it drops), then re-linearisation with the
  girsanov sites re-based by ``tranform_girsanov_sites`` (``sde_utils.py:550-568``), until the outer ELBO converges;
* :class:`VIMarkovGPTrainer` -- ``docs/diffusion_processes/vi_markov_gp_trainer.py:17-92``: the VDP fixed-point
  iterations with learning-rate decay;
* :func:`calculate_nlpd`, :func:`calculate_rmse` -- ``docs/diffusion_processes/exp_dp_utils.py:189-224``.

Same constructor arguments, loop structure, break conditions and return values as the reference, including
prior-parameter learning (``learn_prior_sde=True``: ``optimize_prior_sde`` of ``cvi_dp_trainer.py:207-250`` and
``vi_markov_gp_trainer.py:163-235`` with a host-side Adam on the few drift parameters, :class:`Adam` below); logging
and wandb are not part of this build.  The models are the ``markovflow`` mirrors of this package, i.e. every step
runs through the C ABI on the GPU; the loop itself reads one ELBO scalar (and, when learning, one gradient per
drift parameter) per iteration, as the reference does.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .markovflow._util import as_np


def _f(x) -> float:
    return float(x.item()) if hasattr(x, "item") else float(x)


def _test_indices(time_grid, test_t):
    tg = as_np(time_grid).reshape(-1)
    tt = as_np(test_t).reshape(-1)
    return np.where(np.equal(tg[..., None], tt))[0]


def calculate_nlpd(m, S, time_grid, test_data) -> float:
    """exp_dp_utils.py:189-206: minus the mean Gaussian log predictive density at the test points.
    m [N, D], S [N, D, D] (already passed through ``likelihood.predict_mean_and_var``)."""
    m, S = torch.as_tensor(m), torch.as_tensor(S)
    idx = torch.as_tensor(_test_indices(time_grid, test_data[0]), device=m.device)
    y = torch.as_tensor(as_np(test_data[1]), dtype=m.dtype, device=m.device).reshape(idx.shape[0], -1)
    mt, St = m.index_select(0, idx), S.index_select(0, idx)
    L = torch.linalg.cholesky(St)
    z = torch.linalg.solve_triangular(L, (y - mt)[..., None], upper=False)[..., 0]
    D = mt.shape[-1]
    logp = -0.5 * (z * z).sum(-1) - torch.log(torch.diagonal(L, dim1=-2, dim2=-1)).sum(-1) - 0.5 * D * math.log(2 * math.pi)
    return float(-logp.mean().item())


def calculate_rmse(m, time_grid, test_data) -> float:
    """exp_dp_utils.py:209-224"""
    m = torch.as_tensor(m)
    idx = torch.as_tensor(_test_indices(time_grid, test_data[0]), device=m.device)
    y = torch.as_tensor(as_np(test_data[1]), dtype=m.dtype, device=m.device).reshape(idx.shape[0], -1)
    return float(torch.sqrt(torch.mean((m.index_select(0, idx) - y) ** 2)).item())


class Adam:
    """tf.optimizers.Adam on a short list of float64 scalars (the drift parameters): beta1 = 0.9, beta2 = 0.999,
    epsilon = 1e-7, lr_t = lr sqrt(1 - beta2^t) / (1 - beta1^t), theta -= lr_t m / (sqrt(v) + epsilon)."""

    def __init__(self, learning_rate: float = 1e-2, beta_1: float = 0.9, beta_2: float = 0.999, epsilon: float = 1e-7):
        self.lr, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
        self.t, self.m, self.v = 0, None, None

    def apply_gradients(self, grads, values):
        g = np.asarray([_f(x) for x in grads], dtype=np.float64)
        x = np.asarray([_f(v) for v in values], dtype=np.float64)
        if self.m is None:
            self.m, self.v = np.zeros_like(g), np.zeros_like(g)
        self.t += 1
        self.m = self.beta_1 * self.m + (1.0 - self.beta_1) * g
        self.v = self.beta_2 * self.v + (1.0 - self.beta_2) * g * g
        lr_t = self.lr * math.sqrt(1.0 - self.beta_2 ** self.t) / (1.0 - self.beta_1 ** self.t)
        return x - lr_t * self.m / (np.sqrt(self.v) + self.epsilon)


def _is_ou(sde) -> bool:
    return type(sde).__name__ == "OrnsteinUhlenbeckSDE"


def _is_mlp(sde) -> bool:
    return type(sde).__name__ == "MLPDrift"


class CVISitesTrainer:
    """cvi_dp_trainer.py:18-250."""

    def __init__(self, model, test_data, prior_sde=None, max_itr: int = 100, optim_tol: float = 1e-2,
                 max_itr_sites_optim: int = 20, girsanov_sites_lr: float = 0.1, data_sites_lr: float = 0.1,
                 learn_prior_sde: bool = False, prior_sde_lr: float = 1e-2, learning_max_itr: int = 100,
                 learning_tol: float = 1e-2, device_loop: bool = True, **unused):
        self.model, self.test_data, self.prior_sde = model, test_data, prior_sde
        # the inner loop runs on the device (vdp_cvi_optimize_sites: CUDA-graph WHILE node, convergence and step-size
        # decay decided by a kernel) when the model offers it; device_loop=False keeps the host loop below
        self.device_loop = bool(device_loop) and hasattr(model, "optimize_sites")
        self.max_itr, self.optim_tol, self.max_itr_sites_optim = max_itr, optim_tol, max_itr_sites_optim
        self.girsanov_sites_lr, self.data_sites_lr = girsanov_sites_lr, data_sites_lr
        self.learn_prior_sde, self.prior_sde_lr = learn_prior_sde, prior_sde_lr
        self.learning_max_itr, self.learning_tol = learning_max_itr, learning_tol
        self.prior_params = {}
        self.n_prior_step = 0
        if self.learn_prior_sde:
            self.prior_sde_optim = Adam(learning_rate=self.prior_sde_lr)
            self.store_prior_param_vals()

    def store_prior_param_vals(self):
        """cvi_dp_trainer.py:52-61"""
        if _is_mlp(self.model.prior_sde):
            return
        for i, v in enumerate(self.model.prior_sde.trainable_variables):
            self.prior_params.setdefault(str(i), []).append(float(v))

    def _elbo(self) -> float:
        e = _f(self.model.classic_elbo())
        self._check()
        return e

    def _check(self):
        """Raise on a non positive-definite block (the reference fails inside tf.linalg.cholesky / cholesky_band;
        here the kernels report through the device status word, read where the ELBO has just been synchronised)."""
        check = getattr(self.model, "check_status", None)
        if check is not None:
            check()

    def _marginals(self):
        q = self.model.dist_q                     # cvi_dp_trainer.py:192-207: dist_q.marginals, rebuilt from the sites
        q = q() if callable(q) else q
        ms = q.marginals
        return ms() if callable(ms) else ms

    def model_nlpd(self) -> float:
        m, S = self.model.likelihood.predict_mean_and_var(*self._marginals())
        return calculate_nlpd(m, S, self.model.time_grid, self.test_data)

    def model_rmse(self) -> float:
        m, S = self.model.likelihood.predict_mean_and_var(*self._marginals())
        return calculate_rmse(m, self.model.time_grid, self.test_data)

    def _optimize_sites_on_device(self):
        """cvi_dp_trainer.py:63-101 with the loop on the device: one call, one read-back."""
        idx = _test_indices(self.model.time_grid, self.test_data[0])
        elbo, tm, tc, self.data_sites_lr, self.girsanov_sites_lr = self.model.optimize_sites(
            self.max_itr_sites_optim, self.optim_tol, self.data_sites_lr, self.girsanov_sites_lr, test_indices=idx)
        self._check()
        y = torch.as_tensor(as_np(self.test_data[1]), dtype=tm.dtype).reshape(idx.shape[0], -1)
        nlpd, rmse = [], []
        for m, S in zip(tm, tc):       # NLPD / RMSE of every iteration from the recorded test marginals
            m, S = self.model.likelihood.predict_mean_and_var(m, S)
            L = torch.linalg.cholesky(S)
            z = torch.linalg.solve_triangular(L, (y - m)[..., None], upper=False)[..., 0]
            D = m.shape[-1]
            logp = (-0.5 * (z * z).sum(-1) - torch.log(torch.diagonal(L, dim1=-2, dim2=-1)).sum(-1)
                    - 0.5 * D * math.log(2 * math.pi))
            nlpd.append(float(-logp.mean().item()))
            rmse.append(float(torch.sqrt(torch.mean((m - y) ** 2)).item()))
        return elbo, nlpd, rmse

    def _optimize_sites_under_stable_prior(self):
        """cvi_dp_trainer.py:63-101"""
        if self.device_loop:
            return self._optimize_sites_on_device()
        elbo, nlpd, rmse = [self._elbo()], [], []
        while (len(elbo) - 1) < self.max_itr_sites_optim:
            # update_data_sites; update_girsanov_sites; classic_elbo  == one fused vdp_cvi_site_update
            elbo.append(_f(self.model.site_update(self.data_sites_lr, self.girsanov_sites_lr)))
            self._check()
            nlpd.append(self.model_nlpd())
            rmse.append(self.model_rmse())
            if len(elbo) > 2 and elbo[-2] > elbo[-1]:
                self.girsanov_sites_lr /= 10
                self.data_sites_lr /= 10
            if len(elbo) > 2 and abs(elbo[-2] - elbo[-1]) < self.optim_tol:
                break
        return elbo[1:], nlpd, rmse

    def perform_inference(self):
        """cvi_dp_trainer.py:103-136"""
        elbo_vals, nlpd_vals, rmse_vals = [self._elbo()], [], []
        for i in range(self.max_itr):
            before = self._elbo()
            e, n, r = self._optimize_sites_under_stable_prior()
            self.model.transform_sites_to_linearized_prior()      # :110-113
            after = self._elbo()
            elbo_vals, nlpd_vals, rmse_vals = elbo_vals + e, nlpd_vals + n, rmse_vals + r
            if abs(before - after) < self.optim_tol:
                return elbo_vals[1:], nlpd_vals, rmse_vals
            if i != self.max_itr - 1:                             # :128-134
                self.model.relinearize_and_transform_sites()
        return elbo_vals[1:], nlpd_vals, rmse_vals

    def optimize_prior_sde(self):
        """cvi_dp_trainer.py:207-250: Adam steps on the drift parameters along d(KL - VE)/d params."""
        from .markovflow.sde.sde_utils import Gaussian
        model, sde = self.model, self.model.prior_sde
        elbo, nlpd, rmse = [self._elbo()], [], []
        for _ in range(self.learning_max_itr):
            grads_kl = model.grad_KL_wrt_prior_params()
            grads_ve = model.grad_VE_wrt_prior_params()
            grads = [_f(a) + _f(b) for a, b in zip(grads_kl, grads_ve)]
            sde.assign_trainable(self.prior_sde_optim.apply_gradients(grads, sde.trainable_variables))
            elbo.append(self._elbo())
            nlpd.append(self.model_nlpd())
            rmse.append(self.model_rmse())
            self.n_prior_step += 1
            self.store_prior_param_vals()
            if _is_ou(sde):   # :231-235: p(x0) follows the stationary variance
                steady_cov = np.asarray(sde.q, dtype=np.float64) / (2.0 * sde.decay)
                model.prior_initial_state = Gaussian(np.zeros((1,)), steady_cov * np.ones((1, 1)))
            if elbo[-1] < elbo[-2]:
                self.prior_sde_optim.lr = self.prior_sde_optim.lr / 2
            if abs(elbo[-2] - elbo[-1]) < self.learning_tol:
                break
        return elbo[1:], nlpd, rmse

    def optimize(self):
        """cvi_dp_trainer.py:138-190"""
        elbo, nlpd, rmse = [self._elbo()], [self.model_nlpd()], [self.model_rmse()]
        previous = []
        for _ in range(self.max_itr):
            e, n, r = self.perform_inference()
            elbo, nlpd, rmse = elbo + e, nlpd + n, rmse + r
            if not self.learn_prior_sde:
                break
            pe, pn, pr = self.optimize_prior_sde()
            done = abs(elbo[-1] - pe[-1]) < self.optim_tol
            elbo, nlpd, rmse = elbo + pe, nlpd + pn, rmse + pr
            if done:
                break
            if len(previous) > 4 and (abs(previous[-1] - previous[-3]) < 1e-4 or abs(previous[-2] - previous[-4]) < 1e-4):
                break   # the objective jumps between two values (:176-183)
            previous.append(elbo[-1])
        return elbo, nlpd, rmse, self.prior_params


class VIMarkovGPTrainer:
    """vi_markov_gp_trainer.py:17-235."""

    def __init__(self, model, test_data, q_lr: float = 0.1, x0_lr: float = 0.1, max_itr: int = 100,
                 lr_tol: float = 1e-2, optim_tol: float = 1e-2, warmup_x0_itr: int = 10,
                 learn_prior_sde: bool = False, prior_sde_lr: float = 1e-2, learning_max_itr: int = 100,
                 learning_tol: float = 1e-2, optimize_prior_initial_state: bool = False,
                 prior_initial_state_lr: float = 0.1, **unused):
        self.model, self.test_data = model, test_data
        self.q_lr, self.x0_lr, self.max_itr = q_lr, x0_lr, max_itr
        self.lr_tol, self.optim_tol, self.warmup_x0_itr = lr_tol, optim_tol, warmup_x0_itr
        self.learn_prior_sde, self.prior_sde_lr = learn_prior_sde, prior_sde_lr
        self.learning_max_itr, self.learning_tol = learning_max_itr, learning_tol
        self.optimize_prior_initial_state = optimize_prior_initial_state
        self.prior_initial_state_lr = prior_initial_state_lr
        self.prior_params = {}
        self.n_prior_step = 0
        if self.learn_prior_sde:
            self.prior_sde_optim = Adam(learning_rate=self.prior_sde_lr)
            self.store_prior_param_vals()

    def store_prior_param_vals(self):
        """vi_markov_gp_trainer.py:152-161"""
        if _is_mlp(self.model.prior_sde):
            return
        for i, v in enumerate(self.model.prior_sde.trainable_variables):
            self.prior_params.setdefault(i, []).append(float(v))

    def _check(self):
        check = getattr(self.model, "check_status", None)
        if check is not None:
            check()

    def model_nlpd(self, m, S) -> float:
        m, S = self.model.likelihood.predict_mean_and_var(m, S)
        return calculate_nlpd(m, S, self.model.grid, self.test_data)

    def model_rmse(self, m, S) -> float:
        m, S = self.model.likelihood.predict_mean_and_var(m, S)
        return calculate_rmse(m, self.model.grid, self.test_data)

    def perform_inference(self):
        """vi_markov_gp_trainer.py:50-92"""
        elbo, nlpd, rmse = [_f(self.model.elbo())], [], []
        q_lr, x0_lr = self.q_lr, self.x0_lr
        for i in range(self.max_itr):
            # forward_pass; update_lagrange; update_param; [update_initial_statistics]; forward_pass; elbo
            e = self.model.sweep(q_lr, update_x0=i > self.warmup_x0_itr, x0_lr=x0_lr)
            m, S = self.model.m, self.model.S
            elbo.append(_f(e))
            self._check()
            nlpd.append(self.model_nlpd(m, S))
            rmse.append(self.model_rmse(m, S))
            if elbo[-2] > elbo[-1] or abs(elbo[-2] - elbo[-1]) < self.lr_tol:
                q_lr, x0_lr = q_lr / 10, x0_lr / 10
            if abs(elbo[-2] - elbo[-1]) < self.optim_tol:
                break
        return elbo[1:], nlpd, rmse

    def optimize_prior_x0(self):
        """vi_markov_gp_trainer.py:203-218 (the reference multiplies the new factor with itself, not its transpose)."""
        model, sde = self.model, self.model.prior_sde
        mu = torch.as_tensor(as_np(model.prior_initial_state.mu))
        if _is_ou(sde):
            tril = torch.as_tensor(np.asarray(sde.q, dtype=np.float64) / (2.0 * sde.decay))
        else:
            g = model.grad_initial_state()
            tril0 = torch.linalg.cholesky(torch.as_tensor(as_np(model.prior_initial_state.cov)))
            mu = mu - self.prior_initial_state_lr * g[0].cpu()
            tril = tril0 - self.prior_initial_state_lr * g[1].cpu()
        model.set_prior_initial_state(mu, tril @ tril)

    def optimize_prior_sde(self):
        """vi_markov_gp_trainer.py:163-201"""
        model, sde = self.model, self.model.prior_sde
        elbo, nlpd, rmse = [_f(model.elbo())], [], []
        for _ in range(self.learning_max_itr):
            grads = model.grad_prior_sde_params()
            sde.assign_trainable(self.prior_sde_optim.apply_gradients(grads, sde.trainable_variables))
            if self.optimize_prior_initial_state:
                self.optimize_prior_x0()
            m, S = model.forward_pass
            elbo.append(_f(model.elbo(m, S)))
            self._check()
            nlpd.append(self.model_nlpd(m, S))
            rmse.append(self.model_rmse(m, S))
            self.store_prior_param_vals()
            self.n_prior_step += 1
            if abs(elbo[-2] - elbo[-1]) < self.learning_tol:
                break
        return elbo[1:], nlpd, rmse

    def optimize(self):
        """vi_markov_gp_trainer.py:94-135: 20 warm-up sweeps at lr 1e-6, then inference (and learning) rounds."""
        for _ in range(20):
            self.model.sweep(1e-6)
        m, S = self.model.forward_pass
        elbo, nlpd, rmse = [_f(self.model.elbo(m, S))], [self.model_nlpd(m, S)], [self.model_rmse(m, S)]
        for _ in range(self.max_itr):
            e, n, r = self.perform_inference()
            elbo, nlpd, rmse = elbo + e, nlpd + n, rmse + r
            if not self.learn_prior_sde:
                break
            pe, pn, pr = self.optimize_prior_sde()
            elbo, nlpd, rmse = elbo + pe, nlpd + pn, rmse + pr
            # :131 as written: (len > 2 and |diff|) < optim_tol, i.e. a short list (False < tol) also stops
            if (abs(pe[-2] - pe[-1]) if len(pe) > 2 else 0.0) < self.optim_tol:
                break
        return elbo, nlpd, rmse, self.prior_params
