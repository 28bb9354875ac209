"""
Oracle, layer L1: block-tridiagonal algebra, the state-space model and the
SSM <-> natural / expectation parameter transforms, in NumPy float64.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Nothing here is shipped.

Follows, function by function (paths relative to /root/reference):
  markovflow/block_tri_diag.py            (SymmetricBlockTriDiagonal, LowerTriangularBlockTriDiagonal)
  markovflow/state_space_model.py         (StateSpaceModel)
  markovflow/ssm_gaussian_transformations.py
  markovflow/gauss_markov.py:220-242      (BTDGaussian)

The reference reaches its sequential arithmetic through the un-vendored wheel
``banded-matrices==0.0.6`` (pyproject.toml:17).  Its ops are restated here in
*block* form from their published contracts:
  cholesky_band                -> the unique lower Cholesky factor of the banded SPD matrix
  inverse_from_cholesky_band   -> the lower band (same bandwidth) of (L L^T)^-1
  solve_triang_mat             -> L^-1 b or L^-T b
  solve_triang_band            -> the requested band of L^-1 B / L^-T B
  product_band_mat             -> banded matrix times dense
Both factorisations are unique, so the block recursions below return the same
numbers as the scalar band kernels up to floating-point rounding.

All arrays are row-major with layout ``batch_shape + [n, D, D]`` (block
diagonal), ``batch_shape + [n-1, D, D]`` (block sub-diagonal) and
``batch_shape + [n, D]`` (vectors), exactly as the reference tensors.
"""
from __future__ import annotations

import numpy as np

# ----------------------------------------------------------------------------------------------
# small batched dense helpers on [..., D, D] (stand-ins for tf.linalg.* on D x D blocks)
# ----------------------------------------------------------------------------------------------


def mT(x):
    """tf.linalg.matrix_transpose."""
    return np.swapaxes(x, -1, -2)


def chol(x):
    """tf.linalg.cholesky: lower factor, reads the lower triangle only."""
    return np.linalg.cholesky(x)


def tri_solve(L, B, adjoint=False):
    """
    tf.linalg.triangular_solve(L, B, lower=True, adjoint=adjoint) for ``[..., D, D]`` lower
    triangular ``L`` and ``[..., D, M]`` right-hand sides, by explicit substitution
    (batched; D is tiny so the loop is over rows).
    """
    L = np.asarray(L)
    B = np.asarray(B)
    D = L.shape[-1]
    X = np.zeros(np.broadcast_shapes(L.shape[:-2], B.shape[:-2]) + B.shape[-2:], dtype=np.result_type(L, B))
    if not adjoint:
        for i in range(D):
            acc = B[..., i, :]
            for j in range(i):
                acc = acc - L[..., i, j, None] * X[..., j, :]
            X[..., i, :] = acc / L[..., i, i, None]
    else:
        for i in range(D - 1, -1, -1):
            acc = B[..., i, :]
            for j in range(i + 1, D):
                acc = acc - L[..., j, i, None] * X[..., j, :]
            X[..., i, :] = acc / L[..., i, i, None]
    return X


def chol_solve(L, B):
    """tf.linalg.cholesky_solve(L, B) = (L L^T)^-1 B."""
    return tri_solve(L, tri_solve(L, B), adjoint=True)


def eye_like(x):
    D = x.shape[-1]
    return np.broadcast_to(np.eye(D, dtype=x.dtype), x.shape).copy()


def symmetrise_from_lower(X):
    """band_to_block(..., symmetric=True) on a diagonal block: mirror the lower triangle."""
    lo = np.tril(X)
    return lo + mT(np.tril(X, -1))


# ----------------------------------------------------------------------------------------------
# banded_matrices ops, block form
# ----------------------------------------------------------------------------------------------


def btd_cholesky(Kd, Ks):
    """
    cholesky_band on a symmetric block-tridiagonal matrix (block_tri_diag.py:428-440).

    :param Kd: ``[..., n, D, D]`` symmetric diagonal blocks K_kk.
    :param Ks: ``[..., n-1, D, D]`` sub-diagonal blocks K_{k+1,k} (or None).
    :return: (Ld, Ls) lower block-bidiagonal Cholesky factor.
    """
    Kd = np.asarray(Kd, dtype=np.float64)
    n = Kd.shape[-3]
    Ld = np.empty_like(Kd)
    if Ks is None:
        return chol(Kd), None
    Ks = np.asarray(Ks, dtype=np.float64)
    Ls = np.empty_like(Ks)
    Ld[..., 0, :, :] = chol(Kd[..., 0, :, :])
    for k in range(n - 1):
        # L_{k+1,k} = K_{k+1,k} L_kk^-T
        Ls[..., k, :, :] = mT(tri_solve(Ld[..., k, :, :], mT(Ks[..., k, :, :])))
        S = Kd[..., k + 1, :, :] - Ls[..., k, :, :] @ mT(Ls[..., k, :, :])
        Ld[..., k + 1, :, :] = chol(S)
    return Ld, Ls


def btd_inverse_band(Ld, Ls):
    """
    inverse_from_cholesky_band restricted to the block band (block_tri_diag.py:331,
    ssm_gaussian_transformations.py:443-445): block diagonal Sigma_kk and block sub-diagonal
    Sigma_{k+1,k} of (L L^T)^-1, by the backward (Takahashi) recursion

        Sigma_nn     = L_nn^-T L_nn^-1
        Sigma_{k+1,k} = - Sigma_{k+1,k+1} G_k,           G_k = L_{k+1,k} L_kk^-1
        Sigma_kk     = L_kk^-T L_kk^-1 + G_k^T Sigma_{k+1,k+1} G_k
    """
    Ld = np.asarray(Ld, dtype=np.float64)
    n = Ld.shape[-3]
    I = eye_like(Ld[..., 0, :, :])
    Sd = np.empty_like(Ld)
    U = tri_solve(Ld[..., n - 1, :, :], I)
    Sd[..., n - 1, :, :] = mT(U) @ U
    if Ls is None:
        for k in range(n - 2, -1, -1):
            U = tri_solve(Ld[..., k, :, :], I)
            Sd[..., k, :, :] = mT(U) @ U
        return Sd, None
    Ls = np.asarray(Ls, dtype=np.float64)
    Ss = np.empty_like(Ls)
    for k in range(n - 2, -1, -1):
        U = tri_solve(Ld[..., k, :, :], I)
        G = Ls[..., k, :, :] @ U
        Ss[..., k, :, :] = -Sd[..., k + 1, :, :] @ G
        Sk = mT(U) @ U - mT(Ss[..., k, :, :]) @ G
        Sd[..., k, :, :] = symmetrise_from_lower(Sk)  # only the lower band is formed, then mirrored
    return Sd, Ss


def btd_solve(Ld, Ls, rhs, transpose=False):
    """
    solve_triang_mat (block_tri_diag.py:339-351): x = L^-1 rhs, or L^-T rhs when ``transpose``.
    ``rhs`` has shape ``[..., n, D]``.
    """
    Ld = np.asarray(Ld, dtype=np.float64)
    rhs = np.asarray(rhs, dtype=np.float64)
    n = Ld.shape[-3]
    x = np.empty(np.broadcast_shapes(Ld.shape[:-3], rhs.shape[:-2]) + rhs.shape[-2:], dtype=np.float64)
    if not transpose:
        for k in range(n):
            r = rhs[..., k, :]
            if Ls is not None and k > 0:
                r = r - (Ls[..., k - 1, :, :] @ x[..., k - 1, :, None])[..., 0]
            x[..., k, :] = tri_solve(Ld[..., k, :, :], r[..., None])[..., 0]
    else:
        for k in range(n - 1, -1, -1):
            r = rhs[..., k, :]
            if Ls is not None and k < n - 1:
                r = r - (mT(Ls[..., k, :, :]) @ x[..., k + 1, :, None])[..., 0]
            x[..., k, :] = tri_solve(Ld[..., k, :, :], r[..., None], adjoint=True)[..., 0]
    return x


def btd_matvec(Md, Ms, x, symmetric, transpose=False):
    """product_band_mat (block_tri_diag.py:175-199): dense vector times a block-(tri/bi)diagonal matrix."""
    Md = np.asarray(Md, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    if symmetric:
        y = (Md @ x[..., None])[..., 0]
        if Ms is not None:
            y[..., 1:, :] += (Ms @ x[..., :-1, :, None])[..., 0]
            y[..., :-1, :] += (mT(Ms) @ x[..., 1:, :, None])[..., 0]
        return y
    if not transpose:
        y = (Md @ x[..., None])[..., 0]
        if Ms is not None:
            y[..., 1:, :] += (Ms @ x[..., :-1, :, None])[..., 0]
    else:
        y = (mT(Md) @ x[..., None])[..., 0]
        if Ms is not None:
            y[..., :-1, :] += (mT(Ms) @ x[..., 1:, :, None])[..., 0]
    return y


def btd_to_dense(Md, Ms, symmetric):
    """to_dense (block_tri_diag.py:150-173), for tests."""
    Md = np.asarray(Md)
    n, D = Md.shape[-3], Md.shape[-1]
    out = np.zeros(Md.shape[:-3] + (n * D, n * D), dtype=Md.dtype)
    for k in range(n):
        blk = Md[..., k, :, :]
        if not symmetric:
            blk = np.tril(blk)
        out[..., k * D:(k + 1) * D, k * D:(k + 1) * D] = blk
        if Ms is not None and k < n - 1:
            out[..., (k + 1) * D:(k + 2) * D, k * D:(k + 1) * D] = Ms[..., k, :, :]
            if symmetric:
                out[..., k * D:(k + 1) * D, (k + 1) * D:(k + 2) * D] = mT(Ms[..., k, :, :])
    return out


class LowerTriangularBlockTriDiagonal:
    """block_tri_diag.py:291-381."""

    def __init__(self, diagonal, sub_diagonal=None):
        self.block_diagonal = np.asarray(diagonal, dtype=np.float64)
        self.block_sub_diagonal = None if sub_diagonal is None else np.asarray(sub_diagonal, dtype=np.float64)

    @property
    def inner_dim(self):
        return self.block_diagonal.shape[-1]

    @property
    def outer_dim(self):
        return self.block_diagonal.shape[-3]

    def to_dense(self):
        return btd_to_dense(self.block_diagonal, self.block_sub_diagonal, symmetric=False)

    def solve(self, right, transpose_left=False):
        return btd_solve(self.block_diagonal, self.block_sub_diagonal, right, transpose=transpose_left)

    def dense_mult(self, right, transpose_left=False):
        return btd_matvec(np.tril(self.block_diagonal), self.block_sub_diagonal, right, symmetric=False,
                          transpose=transpose_left)

    def block_diagonal_of_inverse(self):
        return btd_inverse_band(self.block_diagonal, self.block_sub_diagonal)[0]

    def abs_log_det(self):
        d = np.diagonal(self.block_diagonal, axis1=-2, axis2=-1)
        return 0.5 * np.sum(np.log(np.square(d)), axis=(-1, -2))

    def __add__(self, other):
        sub = self.block_sub_diagonal
        if sub is not None:
            if other.block_sub_diagonal is not None:
                sub = sub + other.block_sub_diagonal
        else:
            sub = other.block_sub_diagonal
        return LowerTriangularBlockTriDiagonal(self.block_diagonal + other.block_diagonal, sub)


class SymmetricBlockTriDiagonal:
    """block_tri_diag.py:384-549."""

    def __init__(self, diagonal, sub_diagonal=None):
        self.block_diagonal = np.asarray(diagonal, dtype=np.float64)
        self.block_sub_diagonal = None if sub_diagonal is None else np.asarray(sub_diagonal, dtype=np.float64)

    @property
    def inner_dim(self):
        return self.block_diagonal.shape[-1]

    @property
    def outer_dim(self):
        return self.block_diagonal.shape[-3]

    def to_dense(self):
        return btd_to_dense(self.block_diagonal, self.block_sub_diagonal, symmetric=True)

    def dense_mult(self, right, transpose_left=False):
        return btd_matvec(self.block_diagonal, self.block_sub_diagonal, right, symmetric=True)

    def __add__(self, other):
        if not isinstance(other, (SymmetricBlockTriDiagonal, LowerTriangularBlockTriDiagonal)):
            return SymmetricBlockTriDiagonal(self.block_diagonal + other, self.block_sub_diagonal)
        sub = self.block_sub_diagonal
        if sub is not None:
            if other.block_sub_diagonal is not None:
                sub = sub + other.block_sub_diagonal
        else:
            sub = other.block_sub_diagonal
        return SymmetricBlockTriDiagonal(self.block_diagonal + other.block_diagonal, sub)

    @property
    def cholesky(self):
        Ld, Ls = btd_cholesky(self.block_diagonal, self.block_sub_diagonal)
        return LowerTriangularBlockTriDiagonal(Ld, Ls)

    def upper_diagonal_lower(self):
        """block_tri_diag.py:442-549: backward Schur recursion D_k = K_kk - K_{k+1,k}^T D_{k+1}^-1 K_{k+1,k}."""
        Kd, Ks = self.block_diagonal, self.block_sub_diagonal
        assert Ks is not None
        n = self.outer_dim
        cholD = np.empty_like(Kd)
        cholD[..., n - 1, :, :] = chol(Kd[..., n - 1, :, :])
        for k in range(n - 2, -1, -1):
            d_inv_k = chol_solve(cholD[..., k + 1, :, :], Ks[..., k, :, :])
            cholD[..., k, :, :] = chol(Kd[..., k, :, :] - mT(Ks[..., k, :, :]) @ d_inv_k)
        u_s = chol_solve(cholD[..., 1:, :, :], Ks)
        return (LowerTriangularBlockTriDiagonal(eye_like(Kd), u_s), LowerTriangularBlockTriDiagonal(cholD))


# ----------------------------------------------------------------------------------------------
# state_space_model.py
# ----------------------------------------------------------------------------------------------


class StateSpaceModel:
    """
    state_space_model.py:35-593.  x_{k+1} = A_k x_k + b_k + q_k, q_k ~ N(0, Q_k), x_0 ~ N(mu_0, P_0).
    Argument order as the reference constructor (state_space_model.py:70-77).
    """

    def __init__(self, initial_mean, chol_initial_covariance, state_transitions, state_offsets,
                 chol_process_covariances):
        self._mu_0 = np.asarray(initial_mean, dtype=np.float64)
        self._chol_P_0 = np.asarray(chol_initial_covariance, dtype=np.float64)
        self._A_s = np.asarray(state_transitions, dtype=np.float64)
        self._b_s = np.asarray(state_offsets, dtype=np.float64)
        self._chol_Q_s = np.asarray(chol_process_covariances, dtype=np.float64)

    # -- accessors (state_space_model.py:118-229)
    @property
    def batch_shape(self):
        return self._A_s.shape[:-3]

    @property
    def state_dim(self):
        return self._A_s.shape[-2]

    @property
    def num_transitions(self):
        return self._A_s.shape[-3]

    @property
    def cholesky_process_covariances(self):
        return self._chol_Q_s

    @property
    def cholesky_initial_covariance(self):
        return self._chol_P_0

    @property
    def initial_covariance(self):
        return self._chol_P_0 @ mT(self._chol_P_0)

    @property
    def concatenated_cholesky_process_covariance(self):
        return np.concatenate([self._chol_P_0[..., None, :, :], self._chol_Q_s], axis=-3)

    @property
    def state_offsets(self):
        return self._b_s

    @property
    def initial_mean(self):
        return self._mu_0

    @property
    def concatenated_state_offsets(self):
        return np.concatenate([self._mu_0[..., None, :], self._b_s], axis=-2)

    @property
    def state_transitions(self):
        return self._A_s

    # -- marginals (state_space_model.py:231-262)
    @property
    def a_inv_block(self):
        """state_space_model.py:277-296."""
        n = self.num_transitions + 1
        D = self.state_dim
        I = np.broadcast_to(np.eye(D), self.batch_shape + (n, D, D)).copy()
        return LowerTriangularBlockTriDiagonal(I, -self._A_s)

    @property
    def marginal_means(self):
        return self.a_inv_block.solve(self.concatenated_state_offsets)

    @property
    def marginal_covariances(self):
        return self.precision.cholesky.block_diagonal_of_inverse()

    @property
    def marginals(self):
        """gauss_markov.py:107-117."""
        return self.marginal_means, self.marginal_covariances

    def covariance_blocks(self):
        mc = self.marginal_covariances
        return mc, self.subsequent_covariances(mc)

    def subsequent_covariances(self, marginal_covariances):
        """state_space_model.py:326-341."""
        return self._A_s @ marginal_covariances[..., :-1, :, :]

    def log_det_precision(self):
        """state_space_model.py:343-373."""
        d0 = np.diagonal(self._chol_P_0, axis1=-2, axis2=-1)
        dq = np.diagonal(self._chol_Q_s, axis1=-2, axis2=-1)
        return -(np.sum(np.log(np.square(d0)), axis=-1) + np.sum(np.log(np.square(dq)), axis=(-1, -2)))

    def _build_precision(self):
        """state_space_model.py:431-483."""
        inv_q_a = chol_solve(self._chol_Q_s, self._A_s)
        aqa = mT(self._A_s) @ inv_q_a
        chols = self.concatenated_cholesky_process_covariance
        concatted_inv_q_s = chol_solve(chols, eye_like(chols))
        pad = np.zeros_like(self._chol_P_0)[..., None, :, :]
        diag = concatted_inv_q_s + np.concatenate([aqa, pad], axis=-3)
        return SymmetricBlockTriDiagonal(diag, -inv_q_a)

    @property
    def precision(self):
        return self._build_precision()

    def log_pdf(self, states):
        """state_space_model.py:485-526 (tfp MultivariateNormalTriL log_prob restated)."""
        states = np.asarray(states, dtype=np.float64)
        D = self.state_dim

        def mvn_tril_logpdf(loc, tril, x):
            z = tri_solve(tril, (x - loc)[..., None])[..., 0]
            logdet = np.sum(np.log(np.abs(np.diagonal(tril, axis1=-2, axis2=-1))), axis=-1)
            return -0.5 * np.sum(z * z, axis=-1) - logdet - 0.5 * D * np.log(2.0 * np.pi)

        initial = mvn_tril_logpdf(self._mu_0, self._chol_P_0, states[..., 0, :])
        cond_means = (self._A_s @ states[..., :-1, :, None])[..., 0] + self._b_s
        rest = mvn_tril_logpdf(cond_means, self._chol_Q_s, states[..., 1:, :])
        return initial + np.sum(rest, axis=-1)

    def kl_divergence(self, dist):
        """state_space_model.py:528-593."""
        marginal_covs_1 = self.marginal_covariances
        precision_2 = dist.precision
        subsequent_covs_1 = self.subsequent_covariances(marginal_covs_1)
        trace = np.sum(precision_2.block_diagonal * marginal_covs_1, axis=(-3, -2, -1)) + 2.0 * np.sum(
            precision_2.block_sub_diagonal * subsequent_covs_1, axis=(-3, -2, -1))
        mean_diff = dist.marginal_means - self.marginal_means
        l_mean_diff = precision_2.cholesky.dense_mult(mean_diff, transpose_left=True)
        mahalanobis = np.sum(l_mean_diff * l_mean_diff, axis=(-2, -1))
        dim = float((self.num_transitions + 1) * self.state_dim)
        return 0.5 * (trace + mahalanobis - dim - dist.log_det_precision() + self.log_det_precision())

    def sample(self, sample_shape, rng):
        """state_space_model.py:298-324 with a NumPy generator in place of tf.random."""
        sample_shape = tuple(np.atleast_1d(sample_shape)) if np.ndim(sample_shape) else (int(sample_shape),)
        b = self.concatenated_state_offsets
        eps = rng.standard_normal(sample_shape + b.shape + (1,))
        z = (self.concatenated_cholesky_process_covariance @ eps)[..., 0]
        return self.a_inv_block.solve(b + z)


def state_space_model_from_covariances(initial_mean, initial_covariance, state_transitions, state_offsets,
                                       process_covariances):
    """state_space_model.py:612-664 (Cholesky, or zero where the covariance is exactly zero)."""

    def cholesky_or_zero(cov):
        cov = np.asarray(cov, dtype=np.float64)
        mask = np.all(cov == 0.0, axis=(-2, -1))
        fix = np.where(mask[..., None, None], np.eye(cov.shape[-1]), 0.0)
        return np.where(mask[..., None, None], 0.0, chol(cov + fix))

    return StateSpaceModel(initial_mean, cholesky_or_zero(initial_covariance), state_transitions, state_offsets,
                           cholesky_or_zero(process_covariances))


# ----------------------------------------------------------------------------------------------
# ssm_gaussian_transformations.py
# ----------------------------------------------------------------------------------------------


def ssm_to_expectations(ssm: StateSpaceModel):
    """ssm_gaussian_transformations.py:32-89."""
    mm = ssm.marginal_means[..., None]
    mc = ssm.marginal_covariances
    As = ssm.state_transitions
    eta_linear = mm[..., 0]
    eta_diag = mc + mm @ mT(mm)
    eta_subdiag = As @ mc[..., :-1, :, :] + mm[..., 1:, :, :] @ mT(mm[..., :-1, :, :])
    return eta_linear, eta_diag, eta_subdiag


def expectations_to_ssm_params(eta_linear, eta_diag, eta_subdiag):
    """ssm_gaussian_transformations.py:93-178."""
    e = np.asarray(eta_linear, dtype=np.float64)[..., None]
    eta_diag = np.asarray(eta_diag, dtype=np.float64)
    eta_subdiag = np.asarray(eta_subdiag, dtype=np.float64)
    marginal_covs = eta_diag - e @ mT(e)
    covs_sub_diag = mT(eta_subdiag) - e[..., :-1, :, :] @ mT(e[..., 1:, :, :])
    marginal_chols = chol(marginal_covs)
    As = mT(chol_solve(marginal_chols[..., :-1, :, :], covs_sub_diag))
    initial_mean = e[..., 0, :, 0]
    chol_initial_covariance = marginal_chols[..., 0, :, :]
    offsets = (e[..., 1:, :, :] - As @ e[..., :-1, :, :])[..., 0]
    conditional_covs = marginal_covs[..., 1:, :, :] - As @ (marginal_covs[..., :-1, :, :] @ mT(As))
    chol_process_covariances = chol(conditional_covs)
    return As, offsets, chol_initial_covariance, chol_process_covariances, initial_mean


def ssm_to_naturals(ssm: StateSpaceModel):
    """ssm_gaussian_transformations.py:182-253."""
    As = ssm.state_transitions
    offsets = ssm.concatenated_state_offsets[..., None]
    chols = ssm.concatenated_cholesky_process_covariance
    Linv_As = tri_solve(chols[..., 1:, :, :], As)
    theta_subdiag = tri_solve(chols[..., 1:, :, :], Linv_As, adjoint=True)
    tmp = chol_solve(chols, offsets)
    theta_linear = np.concatenate(
        [tmp[..., :-1, :, :] - mT(As) @ tmp[..., 1:, :, :], tmp[..., -1:, :, :]], axis=-3)[..., 0]
    tmp2 = mT(Linv_As) @ Linv_As
    tmp2 = np.concatenate([tmp2, np.zeros_like(tmp2[..., :1, :, :])], axis=-3)
    precisions = chol_solve(chols, eye_like(chols))
    theta_diag = -0.5 * (precisions + tmp2)
    return theta_linear, theta_diag, theta_subdiag


def ssm_to_naturals_no_smoothing(ssm: StateSpaceModel):
    """ssm_gaussian_transformations.py:257-329."""
    As = ssm.state_transitions
    offsets = ssm.concatenated_state_offsets[..., None]
    chols = ssm.concatenated_cholesky_process_covariance
    theta_subdiag = chol_solve(chols[..., 1:, :, :], As)
    theta_linear = chol_solve(chols, offsets)[..., 0]
    theta_diag = -0.5 * chol_solve(chols, eye_like(chols))
    return theta_linear, theta_diag, theta_subdiag


def naturals_to_ssm_params(theta_linear, theta_diag, theta_subdiag):
    """
    ssm_gaussian_transformations.py:333-511.

    Steps, with the reference line each one restates:
      :440      P = SymmetricBlockTriDiagonal(-2 Theta_d, -Theta_s)
      :443-445  band of P^-1 from the banded Cholesky -> Sigma_kk and Sigma_{k,k+1}
      :462      A_k = (Sigma_kk^-1 Sigma_{k,k+1})^T            (tf.linalg.solve = LU, np.linalg.solve here)
      :470-481  block diagonal of (A^-1)^-T P  = conditional precisions; only the lower band of the
                result is formed and band_to_block mirrors it, hence symmetrise_from_lower
      :493-495  chol -> cholesky_solve(I) -> chol
      :502-504  offsets = Q (A^-T)-solve(theta_linear)
    """
    theta_linear = np.asarray(theta_linear, dtype=np.float64)
    theta_diag = np.asarray(theta_diag, dtype=np.float64)
    theta_subdiag = np.asarray(theta_subdiag, dtype=np.float64)
    Kd = -2.0 * theta_diag
    Ks = -theta_subdiag
    precision = SymmetricBlockTriDiagonal(Kd, Ks)
    L = precision.cholesky
    marginal_covs, cov_sub = btd_inverse_band(L.block_diagonal, L.block_sub_diagonal)  # Sigma_kk, Sigma_{k+1,k}
    sub_diag = mT(cov_sub)  # Sigma_{k,k+1}
    As = mT(np.linalg.solve(marginal_covs[..., :-1, :, :], sub_diag))

    # block diagonal of the back-substitution (A^-1)^T X = P restricted to the lower band:
    #   X_nn = K_nn ;  X_{k+1,k} = K_{k+1,k} ;  X_kk = K_kk + A_{k+1}^T X_{k+1,k}
    cond_prec = Kd.copy()
    cond_prec[..., :-1, :, :] = Kd[..., :-1, :, :] + mT(As) @ Ks
    cond_prec = symmetrise_from_lower(cond_prec)

    chol_cond_prec = chol(cond_prec)
    covariances = chol_solve(chol_cond_prec, eye_like(chol_cond_prec))
    chols = chol(covariances)
    chol_initial_covariance = chols[..., 0, :, :]
    chol_process_covariances = chols[..., 1:, :, :]

    a_inv_block = LowerTriangularBlockTriDiagonal(eye_like(Kd), -As)
    precision_times_offsets = a_inv_block.solve(theta_linear, transpose_left=True)
    offsets = (covariances @ precision_times_offsets[..., None])
    initial_mean = offsets[..., 0, :, 0]
    offsets = offsets[..., 1:, :, 0]
    return As, offsets, chol_initial_covariance, chol_process_covariances, initial_mean


def naturals_to_ssm_params_no_smoothing(theta_linear, theta_diag, theta_subdiag):
    """ssm_gaussian_transformations.py:515-593."""
    theta_linear = np.asarray(theta_linear, dtype=np.float64)
    chol_conditional_prec = chol(-2.0 * np.asarray(theta_diag, dtype=np.float64))
    As = chol_solve(chol_conditional_prec[..., 1:, :, :], np.asarray(theta_subdiag, dtype=np.float64))
    offsets = chol_solve(chol_conditional_prec, theta_linear[..., None])
    initial_mean = offsets[..., 0, :, 0]
    offsets = offsets[..., 1:, :, 0]
    conditional_covs = chol_solve(chol_conditional_prec, eye_like(chol_conditional_prec))
    chols = chol(conditional_covs)
    return As, offsets, chols[..., 0, :, :], chols[..., 1:, :, :], initial_mean


# ----------------------------------------------------------------------------------------------
# gauss_markov.py:220-242
# ----------------------------------------------------------------------------------------------


class BTDGaussian:
    """Gaussian in natural parameters with a symmetric block-tridiagonal nat2 (gauss_markov.py:220-242)."""

    def __init__(self, nat1, nat2_diag, nat2_subdiag):
        self.nat1 = np.array(nat1, dtype=np.float64)
        self.nat2 = SymmetricBlockTriDiagonal(np.array(nat2_diag, dtype=np.float64),
                                              np.array(nat2_subdiag, dtype=np.float64))
