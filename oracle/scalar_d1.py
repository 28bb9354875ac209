"""
Oracle, state dimension 1, batched and dtype-generic.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Nothing here is shipped.

The same operations, in the same order, as gauss_markov_np / sde_np / models_np specialised to
D = 1, written with element-wise NumPy so that
  * a batch of B independent sequences runs vectorised (BASELINE config 4 sized checks), and
  * the whole site update can be re-evaluated in 80-bit ``np.longdouble`` to measure how far the
    float64 reference route itself is from the exactly-rounded answer (the floor for any parity claim).

Arrays are ``[B, N+1]`` / ``[B, N]``.  Reference lines are cited on each function.
"""
from __future__ import annotations

import numpy as np

from .sde_np import (DRIFT_BENES, DRIFT_DW, DRIFT_MLP131, DRIFT_OU, DRIFT_SINE, DRIFT_SQRT)


def drift_and_grad(drift_id, params, x):
    """sde.py drifts (D = 1) and their derivative."""
    p = params
    if drift_id == DRIFT_OU:
        return -p[0] * x, -p[0] * np.ones_like(x)
    if drift_id == DRIFT_DW:
        return p[0] * x * (p[1] - x * x), p[0] * (p[1] - 3.0 * (x * x))
    if drift_id == DRIFT_BENES:
        t = np.tanh(x)
        return p[0] * t, p[0] * (1.0 - t * t)
    if drift_id == DRIFT_SINE:
        return np.sin(x - p[0]), np.cos(x - p[0])
    if drift_id == DRIFT_SQRT:
        r = np.sqrt(p[0] * np.abs(x))
        return r, 0.5 / r * p[0] * np.sign(x)
    if drift_id == DRIFT_MLP131:
        pre = x[..., None] * p[0:3] + p[3:6]
        h = np.maximum(pre, 0)
        return h @ p[6:9] + p[9], ((pre > 0) * (p[0:3] * p[6:9])).sum(-1)
    raise NotImplementedError(drift_id)


def ssm_to_naturals(A, b, cQ, mu0, cP0):
    """ssm_gaussian_transformations.py:182-253 for D = 1."""
    c = np.concatenate([cP0[:, None], cQ], axis=1)
    o = np.concatenate([mu0[:, None], b], axis=1)
    Linv_A = A / cQ
    ths = Linv_A / cQ
    tmp = (o / c) / c
    th1 = np.concatenate([tmp[:, :-1] - A * tmp[:, 1:], tmp[:, -1:]], axis=1)
    tmp2 = np.concatenate([Linv_A * Linv_A, np.zeros_like(A[:, :1])], axis=1)
    prec = (1.0 / c) / c
    thd = -0.5 * (prec + tmp2)
    return th1, thd, ths


def chol_fwd(Kd, Ks):
    """cholesky_band (block_tri_diag.py:440)."""
    n = Kd.shape[1]
    l = np.empty_like(Kd)
    ls = np.empty_like(Ks)
    l[:, 0] = np.sqrt(Kd[:, 0])
    for k in range(n - 1):
        ls[:, k] = Ks[:, k] / l[:, k]
        l[:, k + 1] = np.sqrt(Kd[:, k + 1] - ls[:, k] * ls[:, k])
    return l, ls


def inv_band_bwd(l, ls):
    """inverse_from_cholesky_band (ssm_gaussian_transformations.py:443-445)."""
    n = l.shape[1]
    Sd = np.empty_like(l)
    Ss = np.empty_like(ls)
    U = 1.0 / l
    Sd[:, n - 1] = U[:, n - 1] * U[:, n - 1]
    for k in range(n - 2, -1, -1):
        G = ls[:, k] * U[:, k]
        Ss[:, k] = -Sd[:, k + 1] * G
        Sd[:, k] = U[:, k] * U[:, k] - Ss[:, k] * G
    return Sd, Ss


def naturals_to_ssm_params(th1, thd, ths):
    """ssm_gaussian_transformations.py:333-511 for D = 1."""
    Kd = -2.0 * thd
    Ks = -ths
    l, ls = chol_fwd(Kd, Ks)
    Sd, Ss = inv_band_bwd(l, ls)
    As = Ss / Sd[:, :-1]
    cp = Kd.copy()
    cp[:, :-1] = Kd[:, :-1] + As * Ks
    ccp = np.sqrt(cp)
    cov = (1.0 / ccp) / ccp
    chols = np.sqrt(cov)
    n = th1.shape[1]
    y = np.empty_like(th1)
    y[:, n - 1] = th1[:, n - 1]
    for k in range(n - 2, -1, -1):
        y[:, k] = th1[:, k] + As[:, k] * y[:, k + 1]
    off = cov * y
    return As, off[:, 1:], chols[:, 0], chols[:, 1:], off[:, 0]


def ssm_marginals(A, b, cQ, mu0, cP0):
    """state_space_model.py:231-262, 431-483 for D = 1."""
    n = A.shape[1] + 1
    m = np.empty((A.shape[0], n), dtype=A.dtype)
    m[:, 0] = mu0
    for k in range(n - 1):
        m[:, k + 1] = b[:, k] + A[:, k] * m[:, k]
    inv_q_a = (A / cQ) / cQ
    aqa = A * inv_q_a
    c = np.concatenate([cP0[:, None], cQ], axis=1)
    cinv = (1.0 / c) / c
    diag = cinv + np.concatenate([aqa, np.zeros_like(aqa[:, :1])], axis=1)
    l, ls = chol_fwd(diag, -inv_q_a)
    Sd, _ = inv_band_bwd(l, ls)
    return m, Sd


def kl_and_grads(A_q, m, S, drift_id, params, q, dt, time_grid, p0_mu, p0_cov, gh_x, gh_w):
    """
    sde_utils.py:473-547 for D = 1 (see sde_np.sde_ssm_kl_and_grads for the derivation).
    ``m, S`` are the marginals of q, ``A_q`` its transitions; returns KL and d KL / d(eta1, eta_d, eta_s).
    """
    dtype = m.dtype
    e1 = m
    ed = S + m * m
    es = A_q * S[:, :-1] + m[:, 1:] * m[:, :-1]
    # expectations_to_ssm_params (:93-178)
    Sc = ed - e1 * e1
    Cs = es - e1[:, :-1] * e1[:, 1:]
    LS = np.sqrt(Sc)
    A = (Cs / LS[:, :-1]) / LS[:, :-1]
    b = e1[:, 1:] - A * e1[:, :-1]
    Qc = Sc[:, 1:] - A * (Sc[:, :-1] * A)
    LQ = np.sqrt(Qc)
    Qq = LQ * LQ
    deltas = (time_grid[1:] - time_grid[:-1]).astype(dtype)
    Qp = deltas[None, :] * q
    cQp = np.sqrt(Qp)
    Qpinv = (1.0 / cQp) / cQp
    ldq = 2.0 * np.log(LQ)          # logdet = 2 log diag chol(Qq); chol(Qq) = sqrt(LQ*LQ)
    ldq = 2.0 * np.log(np.sqrt(Qq))
    ldp = 2.0 * np.log(cQp)
    Cterm = -(ldq - ldp) - 1.0 + Qpinv * Qq
    wt = (gh_w * np.pi ** (-0.5)).astype(dtype)
    xn = gh_x.astype(dtype)
    sqrt2 = dtype.type(2.0) ** dtype.type(0.5)
    X = sqrt2 * (LS[:, :-1][None] * xn[:, None, None]) + e1[:, :-1][None]          # [P, B, N]
    f, fp = drift_and_grad(drift_id, params, X)
    d = (X + dt * f) - (A[None] * X + b[None])
    r = Qpinv[None] * d
    quad = np.sum(wt[:, None, None] * (d * r), axis=0)
    kl_path = 0.5 * np.sum(quad + Cterm, axis=1)
    P0 = LS[:, 0] * LS[:, 0]
    kl0 = 0.5 * (P0 / p0_cov + (p0_mu - e1[:, 0]) ** 2 / p0_cov - 1.0 + np.log(p0_cov) - np.log(P0))
    kl = kl0 + kl_path
    # reverse
    g_m = np.zeros_like(m)
    g_S = np.zeros_like(S)
    rb = wt[:, None, None] * r
    g_b = -np.sum(rb, axis=0)
    g_A = -np.sum(rb * X, axis=0)
    Xb = ((1.0 + dt * fp) - A[None]) * rb
    g_m[:, :-1] += np.sum(Xb, axis=0)
    LSb = sqrt2 * np.sum(Xb * xn[:, None, None], axis=0)
    g_S[:, :-1] += LSb / (2.0 * LS[:, :-1])
    g_Qc = 0.5 * (-1.0 / Qq + Qpinv)
    g_m[:, 1:] += g_b
    g_A += -g_b * e1[:, :-1]
    g_m[:, :-1] += -A * g_b
    g_S[:, 1:] += g_Qc
    g_A += -2.0 * g_Qc * A * Sc[:, :-1]
    g_S[:, :-1] += -A * g_Qc * A
    g_C = (g_A / LS[:, :-1]) / LS[:, :-1]
    g_S[:, :-1] += -g_C * A
    g_m[:, 0] += (e1[:, 0] - p0_mu) / p0_cov
    g_S[:, 0] += 0.5 * (1.0 / p0_cov - 1.0 / P0)
    g_es = g_C
    g_m[:, :-1] += -g_C * e1[:, 1:]
    g_m[:, 1:] += -g_C * e1[:, :-1]
    g_ed = g_S
    g_m += -2.0 * g_S * e1
    return kl, (g_m, g_ed, g_es)


def kl_q_p(A, b, cQ, mu0, cP0, m, S, drift_id, params, q, dt, p_mu0, p_cP0, gh_x, gh_w):
    """variational_cvi_sde.py:446-486 for D = 1."""
    dtype = m.dtype
    Qq = cQ * cQ
    Qp = dt * q * np.ones_like(Qq)
    cQp = np.sqrt(Qp)
    Qpinv = (1.0 / cQp) / cQp
    Cterm = -(2.0 * np.log(np.sqrt(Qq)) - 2.0 * np.log(cQp)) - 1.0 + Qpinv * Qq
    wt = (gh_w * np.pi ** (-0.5)).astype(dtype)
    xn = gh_x.astype(dtype)
    sqrt2 = dtype.type(2.0) ** dtype.type(0.5)
    X = sqrt2 * (np.sqrt(S[:, :-1])[None] * xn[:, None, None]) + m[:, :-1][None]
    f, _ = drift_and_grad(drift_id, params, X)
    d = (X + dt * f) - (A[None] * X + b[None])
    quad = np.sum(wt[:, None, None] * (d * Qpinv[None] * d), axis=0)
    kl_path = 0.5 * np.sum(quad + Cterm, axis=1)
    P0 = cP0 * cP0
    Pp = p_cP0 * p_cP0
    kl0 = 0.5 * (P0 / Pp + (p_mu0 - mu0) ** 2 / Pp - 1.0 + np.log(Pp) - np.log(P0))
    return kl_path + kl0


class BatchedCVISitesSDE1D:
    """
    B independent copies of models_np.CVISitesSDE with D = 1 (same grid, per-sequence observations at
    per-sequence grid indices).  ``site_update`` = cvi_dp_trainer.py:71-77 once.
    """

    def __init__(self, drift_id, params, q, time_grid, obs_idx, obs_y, noise_std, p0_mu, p0_cov,
                 dtype=np.float64, clip=(-1.0, 1.0), stabilize=True):
        self.dtype = np.dtype(dtype)
        T = self.dtype.type
        self.drift_id, self.params = drift_id, np.asarray(params, dtype=self.dtype)
        self.q = T(q)
        self.time_grid = np.asarray(time_grid, dtype=np.float64)
        self.dt = T(float(self.time_grid[1] - self.time_grid[0]))
        self.obs_idx = np.asarray(obs_idx)                 # [B, n_obs] sorted grid indices
        self.obs_y = np.asarray(obs_y, dtype=self.dtype)   # [B, n_obs]
        self.noise_std = T(noise_std)
        self.p0_mu, self.p0_cov = T(p0_mu), T(p0_cov)
        B, n_obs = self.obs_idx.shape
        n1 = self.time_grid.shape[0]
        self.B, self.n1 = B, n1
        self.g1 = np.zeros((B, n1), dtype=self.dtype)
        self.gd = np.full((B, n1), -1e-10, dtype=self.dtype)
        self.gs = np.full((B, n1 - 1), -1e-10, dtype=self.dtype)
        self.d1 = np.zeros((B, n_obs), dtype=self.dtype)
        self.d2 = np.full((B, n_obs), 1e-10, dtype=self.dtype)
        self.fx_mus = np.zeros((B, n1), dtype=self.dtype)
        self.fx_covs = np.ones((B, n1), dtype=self.dtype)
        self.clip, self.stabilize = clip, stabilize
        self.gh10 = np.polynomial.hermite.hermgauss(10)
        self.gh20 = np.polynomial.hermite.hermgauss(20)
        self.set_linearized_prior()

    def set_linearized_prior(self):
        """variational_cvi_sde.py:408-432 + sde_utils.py:119-179 + drift.py:64-116."""
        T = self.dtype.type
        m, S = self.fx_mus[:, 1:], self.fx_covs[:, 1:]
        xn, wn = self.gh10
        xn = xn.astype(self.dtype)
        wt = (wn * np.pi ** (-0.5)).astype(self.dtype)
        X = T(2.0) ** T(0.5) * (np.sqrt(S)[None] * xn[:, None, None]) + m[None]
        f, fp = drift_and_grad(self.drift_id, self.params, X)
        E_f = np.sum(f * wt[:, None, None], axis=0)
        A = np.sum(fp * wt[:, None, None], axis=0)
        b = E_f - A * m
        deltas = (self.time_grid[1:] - self.time_grid[:-1]).astype(self.dtype)[None]
        cq = np.sqrt(self.q)
        self.lin = dict(A=A * deltas + T(1.0), b=b * deltas,
                        cQ=np.sqrt((cq * cq) * deltas) * np.ones_like(A),
                        mu0=np.full(self.B, self.p0_mu, dtype=self.dtype),
                        cP0=np.full(self.B, np.sqrt(self.p0_cov), dtype=self.dtype))
        self.p = dict(self.lin)
        if self.stabilize:
            self.p["A"] = np.clip(self.lin["A"], T(self.clip[0]), T(self.clip[1]))
            self.p["b"] = np.clip(self.lin["b"], T(self.clip[0]), T(self.clip[1]))

    def _prior_nat(self, p):
        return ssm_to_naturals(p["A"], p["b"], p["cQ"], p["mu0"], p["cP0"])

    def _scatter(self):
        s1 = np.zeros((self.B, self.n1), dtype=self.dtype)
        s2 = np.zeros((self.B, self.n1), dtype=self.dtype)
        rows = np.arange(self.B)[:, None]
        np.add.at(s1, (rows, self.obs_idx), self.d1)
        np.add.at(s2, (rows, self.obs_idx), self.d2)
        return s1, s2

    def dist_q(self):
        p1, pd, ps = self._prior_nat(self.p)
        s1, s2 = self._scatter()
        return naturals_to_ssm_params(p1 + self.g1 + s1, pd + self.gd + s2, ps + self.gs)

    def marginals(self):
        A, b, cP0, cQ, mu0 = self.dist_q()
        return ssm_marginals(A, b, cQ, mu0, cP0)

    def update_data_sites(self, lr):
        rows = np.arange(self.B)[:, None]
        mu = self.fx_mus[rows, self.obs_idx]
        inv = (1.0 / self.noise_std) / self.noise_std
        g_mu = inv * (self.obs_y - mu)
        g_S = -0.5 * inv * np.ones_like(mu)
        g1 = g_mu - 2.0 * g_S * mu
        self.d1 = (1 - lr) * self.d1 + lr * g1
        self.d2 = (1 - lr) * self.d2 + lr * g_S
        self.fx_mus, self.fx_covs = self.marginals()

    def update_girsanov_sites(self, lr):
        A, b, cP0, cQ, mu0 = self.dist_q()
        m, S = ssm_marginals(A, b, cQ, mu0, cP0)
        _, (g1, g2, g3) = kl_and_grads(A, m, S, self.drift_id, self.params, self.q, self.dt, self.time_grid,
                                       self.p0_mu, self.p0_cov, *self.gh20)
        s1, s2 = self._scatter()
        self.g1 = self.g1 + lr * (s1 - g1)
        self.gd = self.gd + lr * (s2 - g2)
        self.gs = self.gs - lr * g3
        self.fx_mus, self.fx_covs = self.marginals()

    def classic_elbo(self):
        A, b, cP0, cQ, mu0 = self.dist_q()
        m, S = ssm_marginals(A, b, cQ, mu0, cP0)
        rows = np.arange(self.B)[:, None]
        mu, cov = m[rows, self.obs_idx], S[rows, self.obs_idx]
        inv = (1.0 / self.noise_std) / self.noise_std
        z = (self.obs_y - mu) / self.noise_std
        ve = np.sum(-0.5 * inv * cov + (-0.5 * z * z - np.log(np.abs(self.noise_std)) - 0.5 * np.log(2.0 * np.pi)),
                    axis=1)
        kl = kl_q_p(A, b, cQ, mu0, cP0, m, S, self.drift_id, self.params, self.q, self.dt, self.p["mu0"],
                    self.p["cP0"], *self.gh20)
        return ve - kl

    def site_update(self, data_lr, girsanov_lr):
        self.update_data_sites(data_lr)
        self.update_girsanov_sites(girsanov_lr)
        return self.classic_elbo()

    def relinearize(self):
        """cvi_dp_trainer.py:110-134 (sde_utils.py:550-568 twice, around set_linearized_prior)."""
        def move(cur, new):
            c, n = self._prior_nat(cur), self._prior_nat(new)
            self.g1 = self.g1 + c[0] - n[0]
            self.gd = self.gd + c[1] - n[1]
            self.gs = self.gs + c[2] - n[2]
        move(self.p, self.lin)
        self.p = dict(self.lin)
        last = self.p
        self.set_linearized_prior()
        move(last, self.p)
