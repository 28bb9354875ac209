"""Mirror of ``markovflow/likelihoods/multivariate_gaussian.py`` (the likelihood the diffusion models use)."""
from __future__ import annotations

import numpy as np

from ._util import as_np


class MultivariateGaussian:
    """Gaussian likelihood given by the Cholesky factor of its covariance (multivariate_gaussian.py:28-177).
    The expectations at the observed grid points are evaluated inside the fused CUDA entries
    (k_data_site_update, k_variational_expectation); this object carries the parameters."""

    def __init__(self, chol_covariance):
        self.chol_covariance = np.atleast_2d(as_np(chol_covariance)).astype(np.float64)
        self.obs_dim = self.chol_covariance.shape[-1]

    @property
    def covariance(self):
        return self.chol_covariance @ self.chol_covariance.T

    def predict_mean_and_var(self, f_means, f_covariances):
        """multivariate_gaussian.py:118-140"""
        import torch
        cov = torch.as_tensor(self.covariance, dtype=f_covariances.dtype, device=f_covariances.device)
        return f_means, f_covariances + cov
