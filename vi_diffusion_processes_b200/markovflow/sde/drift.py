"""Mirror of ``markovflow/sde/drift.py``: the time-varying linear drift f_L(x, t) = A_t x + b_t."""
from __future__ import annotations

import numpy as np
import torch

from ... import ops
from .._util import as_dev, as_np, batched
from ..state_space_model import StateSpaceModel


class LinearDrift:
    """drift.py:24-116"""

    def __init__(self, A=None, b=None):
        self.A = None if A is None else as_dev(A)
        self.b = None if b is None else as_dev(b)

    def set_from_ssm(self, ssm: StateSpaceModel, dt: float):
        """A = (A_ssm - I) / dt, b = b_ssm / dt   (drift.py:39-62)"""
        st = ssm.state_transitions
        eye = torch.eye(ssm.state_dim, dtype=st.dtype, device=st.device)
        self.A = (st - eye) / dt
        self.b = ssm.state_offsets / dt

    def to_ssm(self, q, transition_times, initial_mean, initial_chol_covariance) -> StateSpaceModel:
        """A_ssm = I + A delta, b_ssm = b delta, cholQ = chol(q delta)   (drift.py:64-116).
        ``q`` is the [D, D] spectral density (or [.., N, D, D] of identical blocks, as the reference passes)."""
        if self.A is None or self.b is None:
            raise Exception("Linear drift is empty so SSM can't be created!")
        A, _ = batched(self.A, 3)
        b, _ = batched(self.b, 2)
        qn = as_np(q)
        D = A.shape[-1]
        qn = qn.reshape(-1, D, D)[0]
        sA, sb, scQ = ops.linear_drift_to_ssm(A, b, np.ascontiguousarray(qn), as_dev(transition_times).reshape(-1))
        mu0 = batched(as_dev(initial_mean), 1)[0]
        cP0 = batched(as_dev(initial_chol_covariance), 2)[0]
        return StateSpaceModel(mu0, cP0, sA, sb, scQ)
