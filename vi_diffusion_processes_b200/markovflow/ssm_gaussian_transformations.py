"""
Mirror of ``markovflow/ssm_gaussian_transformations.py``: conversions between the state-space, natural and
expectation parameterisations of a Gauss-Markov chain.  Each function is one call of the C ABI.
"""
from __future__ import annotations

from .. import ops
from ._util import batched, unbatched


def _ssm_tensors(ssm):
    A, had = batched(ssm.state_transitions, 3)
    return (A, batched(ssm.state_offsets, 2)[0], batched(ssm.cholesky_process_covariances, 3)[0],
            batched(ssm.initial_mean, 1)[0], batched(ssm.cholesky_initial_covariance, 2)[0], had)


def ssm_to_expectations(ssm):
    """ssm_gaussian_transformations.py:32-89 -> (eta1 [..,N+1,D], eta_diag [..,N+1,D,D], eta_subdiag [..,N,D,D])."""
    A, b, cQ, mu0, cP0, had = _ssm_tensors(ssm)
    m, Sd = ops.ssm_marginals(A, b, cQ, mu0, cP0)
    return tuple(unbatched(t, had) for t in ops.ssm_to_expectations(A, m, Sd))


def expectations_to_ssm_params(exp_1, exp_2_diag, exp_2_subdiag):
    """ssm_gaussian_transformations.py:93-178 -> (As, offsets, chol_initial_covariance, chol_process_covariances, mu0)."""
    e1, had = batched(exp_1, 2)
    out = ops.expectations_to_ssm(e1, batched(exp_2_diag, 3)[0], batched(exp_2_subdiag, 3)[0])   # reference order
    return tuple(unbatched(t, had) for t in out)


def ssm_to_naturals(ssm):
    """ssm_gaussian_transformations.py:182-253 -> (theta_linear, theta_diag, theta_subdiag)."""
    A, b, cQ, mu0, cP0, had = _ssm_tensors(ssm)
    return tuple(unbatched(t, had) for t in ops.ssm_to_naturals(A, b, cQ, mu0, cP0))


def naturals_to_ssm_params(theta_linear, theta_diag, theta_subdiag):
    """ssm_gaussian_transformations.py:333-511 -> (As, offsets, chol_initial_covariance, chol_process_covariances, mu0)."""
    t1, had = batched(theta_linear, 2)
    status = ops.new_status(t1.device)
    A, b, cP0, cQ, mu0 = ops.naturals_to_ssm(t1, batched(theta_diag, 3)[0], batched(theta_subdiag, 3)[0], status=status)
    code, idx = ops.read_status(status)
    if code != 0:
        raise ValueError(f"naturals_to_ssm_params: precision block {idx} is not positive definite")
    return tuple(unbatched(t, had) for t in (A, b, cP0, cQ, mu0))
