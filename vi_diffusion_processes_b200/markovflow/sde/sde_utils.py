"""
Mirror of ``markovflow/sde/sde_utils.py`` for the functions on the VDP / CVI-DP path.  Functions that take
arbitrary Python callables in the reference (``SSM_KL_along_Gaussian_path``) are offered in the closed form
the models use them in (a registered SDE drift against a state-space model).
"""
from __future__ import annotations

from collections import namedtuple

import numpy as np
import torch

from ... import ops
from .._util import as_dev, as_np, batched, unbatched
from ..gauss_markov import BTDGaussian
from ..ssm_gaussian_transformations import ssm_to_naturals
from ..state_space_model import StateSpaceModel
from .drift import LinearDrift
from .sde import SDE

Gaussian = namedtuple("Gaussian", ["mu", "cov"])   # stands in for gpflow.probability_distributions.Gaussian


def handle_tensor_shape(tensor, desired_dimensions=2):
    """sde_utils.py:99-116: squeeze / expand leading axes to get ``desired_dimensions`` axes."""
    t = tensor
    while t.dim() > desired_dimensions:
        t = t.squeeze(0)
    while t.dim() < desired_dimensions:
        t = t.unsqueeze(0)
    return t


def euler_maruyama(sde: SDE, x0, time_grid, seed: int = 0, noise=None):
    """sde_utils.py:36-96 on the GPU (``vdp_euler_maruyama``): x0 [n_batch, D], time_grid [n] -> [n_batch, n, D],
    first state = x0, the reference's clock convention (dt_0 = time_grid[0] - 0).  Randomness: Philox with ``seed``
    (tf.random.normal in the reference), or the standard normals ``noise`` [n_batch, n-1, D]."""
    x = as_dev(x0)
    x = x.reshape(-1, x.shape[-1]).contiguous()
    tg = as_dev(time_grid).reshape(-1).contiguous()
    nz = None if noise is None else as_dev(noise).contiguous()
    return ops.euler_maruyama(sde.drift_id, sde.params, sde.q, x, tg, noise=nz, seed=seed)


def linearize_sde(sde: SDE, transition_times, linearization_path: Gaussian, initial_state: Gaussian) -> StateSpaceModel:
    """sde_utils.py:119-179: statistical linearisation A = E[grad f], b = E[f] - A m on the path (N points for the
    N transitions), then LinearDrift.to_ssm.  Returns the SSM with a leading batch axis of 1, as the reference."""
    m, _ = batched(as_dev(linearization_path.mu), 2)
    S, _ = batched(as_dev(linearization_path.cov), 3)
    tg = as_dev(transition_times).reshape(-1)
    A, b, cQ = ops.linearize_to_ssm(sde.drift_id, sde.params, sde.q, tg, m, S)
    mu0 = batched(as_dev(initial_state.mu), 1)[0]
    P0 = as_np(initial_state.cov)
    cP0 = batched(as_dev(np.linalg.cholesky(P0)), 2)[0]
    return StateSpaceModel(mu0, cP0, A, b, cQ)


def squared_drift_difference_along_Gaussian_path(sde_p: SDE, linear_drift: LinearDrift, q: Gaussian, dt: float,
                                                 quadrature_pnts: int = 20):
    """sde_utils.py:182-249: 0.5 dt sum_t E_q ||A_t x + b_t - f(x)||^2_{q^-1} (H = 20 only)."""
    if quadrature_pnts != 20:
        raise ValueError("the CUDA path implements the reference's H = 20 rule")
    A, had = batched(linear_drift.A, 3)
    b, _ = batched(linear_drift.b, 2)
    m, _ = batched(as_dev(q.mu), 2)
    S, _ = batched(as_dev(q.cov), 3)
    pad_m = torch.cat([m, m[:, -1:]], dim=1).contiguous()       # the kernel indexes a grid of N + 1 points
    pad_S = torch.cat([S, S[:, -1:]], dim=1).contiguous()
    val = ops.sq_drift_difference(sde_p.drift_id, sde_p.params, sde_p.q, dt, (-1.0 * A).contiguous(), b, pad_m, pad_S)
    return val[0] if had is False else val


def remove_batch_from_ssm(ssm: StateSpaceModel) -> StateSpaceModel:
    """sde_utils.py:464-470"""
    sq = lambda t, core: t.reshape(t.shape[t.dim() - core:])
    return StateSpaceModel(sq(ssm.initial_mean, 1), sq(ssm.cholesky_initial_covariance, 2), sq(ssm.state_transitions, 3),
                           sq(ssm.state_offsets, 2), sq(ssm.cholesky_process_covariances, 3))


def ssm_to_btd_nat(ssm: StateSpaceModel) -> BTDGaussian:
    """sde_utils.py:362-373"""
    return BTDGaussian(*ssm_to_naturals(ssm))


def tranform_girsanov_sites(girsanov_sites: BTDGaussian, current_prior: StateSpaceModel,
                            new_prior: StateSpaceModel) -> BTDGaussian:
    """sites + nat(current prior) - nat(new prior)   (sde_utils.py:550-568; the reference's spelling is kept)."""
    c, n = ssm_to_btd_nat(current_prior), ssm_to_btd_nat(new_prior)
    return BTDGaussian(girsanov_sites.nat1 + c.nat1 - n.nat1,
                       girsanov_sites.nat2.block_diagonal + c.nat2.block_diagonal - n.nat2.block_diagonal,
                       girsanov_sites.nat2.block_sub_diagonal + c.nat2.block_sub_diagonal
                       - n.nat2.block_sub_diagonal)


def SDE_SSM_KL_with_grads_wrt_exp_params(ssm_q: StateSpaceModel, sde_p: SDE, dt: float, prior_initial_state: Gaussian,
                                         time_grid):
    """sde_utils.py:473-547: KL[q || p_SDE] (H = 20) and its gradient wrt q's expectation parameters."""
    A, had = batched(ssm_q.state_transitions, 3)
    m, Sd = ssm_q.marginals
    m, _ = batched(m, 2)
    Sd, _ = batched(Sd, 3)
    kl, g = ops.cvi_kl_and_grad(sde_p.drift_id, sde_p.params, sde_p.q, dt, as_dev(time_grid).reshape(-1),
                                as_np(prior_initial_state.mu), as_np(prior_initial_state.cov), A, m, Sd)
    kl = kl[0] if had is False else kl
    return kl, [unbatched(t, had) for t in g]
