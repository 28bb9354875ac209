"""
Oracle, layer L1b: ``naturals_to_ssm_params`` and ``StateSpaceModel.marginals`` by the TWO-SWEEP route the CUDA path
uses on a time-sharded chain (csrc/scan_elems.cuh InfoOps, csrc/tile_sweeps.cuh k_info_tiled / k_moments_tiled),
in NumPy float64.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Nothing here is shipped.

This is NOT how the reference computes these quantities -- ssm_gaussian_transformations.py:333-511 factorises the
precision from the top, takes the band of its inverse and forms A from the covariances (oracle/gauss_markov_np.py
``naturals_to_ssm_params`` restates that) -- but a re-ordering of the same elimination:

    K = SymmetricBlockTriDiagonal(-2 Theta_d, -Theta_s), linear term theta_1           (:440)
    J_N = K_NN, h_N = theta_1[N];   downwards, k = N-1 .. 0:
        Q_k = J_{k+1}^-1,  A_k = Q_k Theta_s[k],  b_k = Q_k h_{k+1}
        J_k = K_kk - Theta_s[k]^T A_k,   h_k = theta_1[k] + Theta_s[k]^T b_k
    P_0 = J_0^-1,  mu_0 = P_0 h_0

J_k is the precision of x_k given x_{k-1} with everything after k integrated out, i.e. the conditional precision the
reference obtains at :470-481, and A_k, b_k, Q_k are the forward transitions of the same Gaussian.  tests/
test_oracle_info_form.py holds the two restatements together on random chains (the pin of the CUDA route is then the
pin of the reference-order oracle).  ``moments`` is state_space_model.py:231-262 by the textbook recursion.
"""
from __future__ import annotations

import numpy as np

from .gauss_markov_np import chol, chol_solve, eye_like, mT, symmetrise_from_lower


def naturals_to_ssm_params_info(theta_linear, theta_diag, theta_subdiag):
    """Same signature and return order as gauss_markov_np.naturals_to_ssm_params; leading batch dimensions allowed."""
    t1 = np.asarray(theta_linear, dtype=np.float64)
    Kd = symmetrise_from_lower(-2.0 * np.asarray(theta_diag, dtype=np.float64))
    Ts = np.asarray(theta_subdiag, dtype=np.float64)
    N = Ts.shape[-3]
    As, bs, cQ = np.empty_like(Ts), np.empty_like(t1[..., 1:, :]), np.empty_like(Ts)
    J, h = Kd[..., N, :, :], t1[..., N, :]
    for k in range(N - 1, -1, -1):
        cJ = chol(J)
        Q = chol_solve(cJ, eye_like(cJ))                       # chol -> cholesky_solve(I) -> chol, as :493-495
        cQ[..., k, :, :] = chol(Q)
        As[..., k, :, :] = Q @ Ts[..., k, :, :]
        bs[..., k, :] = (Q @ h[..., None])[..., 0]
        J = symmetrise_from_lower(Kd[..., k, :, :] - mT(Ts[..., k, :, :]) @ As[..., k, :, :])
        h = t1[..., k, :] + (mT(Ts[..., k, :, :]) @ bs[..., k, :, None])[..., 0]
    cJ = chol(J)
    P0 = chol_solve(cJ, eye_like(cJ))
    return As, bs, chol(P0), cQ, (P0 @ h[..., None])[..., 0]


def moments(As, bs, chol_P0, chol_Q, mu0):
    """marginal means [.., N+1, D] and covariances [.., N+1, D, D] by m' = A m + b, S' = A S A^T + Q"""
    As, bs = np.asarray(As, dtype=np.float64), np.asarray(bs, dtype=np.float64)
    N, D = As.shape[-3], As.shape[-1]
    m = np.empty(bs.shape[:-2] + (N + 1, D))
    S = np.empty(As.shape[:-3] + (N + 1, D, D))
    m[..., 0, :] = mu0
    S[..., 0, :, :] = chol_P0 @ mT(chol_P0)
    for k in range(N):
        A, c = As[..., k, :, :], chol_Q[..., k, :, :]
        m[..., k + 1, :] = (A @ m[..., k, :, None])[..., 0] + bs[..., k, :]
        S[..., k + 1, :, :] = A @ S[..., k, :, :] @ mT(A) + c @ mT(c)
    return m, S


# ---- the associative elements of the scan (scan_elems.cuh InfoOps), for the test of their algebra
def info_step(Ts, Kd, t1):
    D = Ts.shape[-1]
    return dict(La=np.zeros((D, D)), Lab=Ts, Lb=Kd, va=np.zeros(D), vb=t1)


def info_combine(e1, e2):
    """e1 (the upper stretch) first, then e2"""
    Mi = np.linalg.inv(e1["Lb"] + e2["La"])
    U, g = e1["Lab"], e1["vb"] + e2["va"]
    return dict(La=e1["La"] - U @ Mi @ U.T, Lab=U @ Mi @ e2["Lab"], Lb=e2["Lb"] - e2["Lab"].T @ Mi @ e2["Lab"],
                va=e1["va"] + U @ Mi @ g, vb=e2["vb"] + e2["Lab"].T @ Mi @ g)


def info_apply(e, J, h):
    Xi = np.linalg.inv(J + e["La"])
    return e["Lb"] - e["Lab"].T @ Xi @ e["Lab"], e["vb"] + e["Lab"].T @ Xi @ (h + e["va"])
