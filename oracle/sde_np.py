"""
Oracle, layer L2: SDE drifts, Gauss-Hermite expectations, linearisation, and the
KL / drift-difference functionals with their gradients, in NumPy float64.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Nothing here is shipped.

Follows (paths relative to /root/reference):
  markovflow/sde/sde.py        SDE, the seven drifts, expected_drift / expected_gradient_drift
  markovflow/sde/drift.py      LinearDrift.to_ssm / set_from_ssm
  markovflow/sde/sde_utils.py  euler_maruyama, linearize_sde, squared_drift_difference_along_Gaussian_path,
                               SSM_KL_along_Gaussian_path, SDE_SSM_KL_with_grads_wrt_exp_params,
                               ssm_to_btd_nat, tranform_girsanov_sites

Third-party arithmetic restated from its published algorithm (package absent here):
  gpflow 2.2.1  gpflow.quadrature.mvnquad / mvhermgauss  (poetry.lock:418-419) -> mvnquad() below.
  tensorflow    reverse-mode gradients.  The reference obtains d KL / d eta from a GradientTape
                (sde_utils.py:496-545); the tape differentiates the *discretised* quadrature sum and
                the D x D Choleskys, so the hand-written adjoint below differentiates exactly those
                operations (Cholesky adjoint after I. Murray 2016, the formula TensorFlow registers).
"""
from __future__ import annotations

import itertools

import numpy as np

from .gauss_markov_np import (BTDGaussian, StateSpaceModel, chol, chol_solve, expectations_to_ssm_params, mT,
                              ssm_to_expectations, ssm_to_naturals, tri_solve)

# ----------------------------------------------------------------------------------------------
# gpflow.quadrature.mvnquad
# ----------------------------------------------------------------------------------------------


def mvhermgauss(H: int, D: int):
    """gpflow.quadrature.mvhermgauss: Cartesian-product Gauss-Hermite nodes ``[H^D, D]`` and weights ``[H^D]``."""
    gh_x, gh_w = np.polynomial.hermite.hermgauss(H)
    x = np.array(list(itertools.product(*(gh_x,) * D)))
    w = np.prod(np.array(list(itertools.product(*(gh_w,) * D))), 1)
    return x, w


def mvnquad(func, means, covs, H: int, Din: int = None, Dout=None):
    """
    gpflow.quadrature.mvnquad (call sites sde.py:109,128,515; sde_utils.py:247,354).

    X = sqrt(2) chol(cov) xi + mean, rows ordered node-major then N (``[H^D * N, D]``), result
    ``sum_i w_i pi^(-D/2) f(X_i)`` over the node axis.
    """
    means = np.asarray(means, dtype=np.float64)
    covs = np.asarray(covs, dtype=np.float64)
    if Din is None:
        Din = means.shape[1]
    xn, wn = mvhermgauss(H, Din)
    N = means.shape[0]
    cholXcov = chol(covs)
    Xt = cholXcov @ np.broadcast_to(xn.T, (N, Din, xn.shape[0]))
    X = 2.0 ** 0.5 * Xt + means[:, :, None]
    Xr = np.reshape(np.transpose(X, (2, 0, 1)), (-1, Din))
    fevals = np.asarray(func(Xr))
    if Dout is None:
        Dout = tuple(fevals.shape[1:])
    fX = np.reshape(fevals, (H ** Din, N) + tuple(Dout))
    wr = np.reshape(wn * np.pi ** (-Din * 0.5), (-1,) + (1,) * (1 + len(Dout)))
    return np.sum(fX * wr, axis=0)


# ----------------------------------------------------------------------------------------------
# sde.py
# ----------------------------------------------------------------------------------------------

DRIFT_OU, DRIFT_DW, DRIFT_BENES, DRIFT_SINE, DRIFT_SQRT, DRIFT_VDP, DRIFT_MLP131, DRIFT_LORENZ63, DRIFT_OU_ND = range(9)


class SDE:
    """
    sde.py:24-131.  ``drift`` maps ``[..., D] -> [..., D]``.  ``gradient_drift`` reproduces the tape
    semantics of the reference: d(sum_j f_j)/dx, shape ``[n, D]`` (sde.py:75-90), unless the subclass
    overrides it with a batch Jacobian ``[n, D, D]`` (sde.py:484-499).  ``jacobian`` is always the full
    ``[..., D, D]`` Jacobian df_i/dx_j, used by the hand-written adjoints.
    """
    drift_id = -1
    full_jacobian = False  # True => expected_gradient_drift returns [.., D, D] (VdP-style)

    def __init__(self, q):
        self.q = np.atleast_2d(np.asarray(q, dtype=np.float64))
        self.state_dim = self.q.shape[0]

    @property
    def params(self):
        raise NotImplementedError

    def drift(self, x, t=None):
        raise NotImplementedError

    def jacobian(self, x):
        raise NotImplementedError

    def with_params(self, params):
        """A copy of this SDE with the parameter block replaced (used by the prior-learning restatement)."""
        raise NotImplementedError

    # the reference's tf.Variable(trainable=...) flags: ``trainable`` = positions in ``params`` (None = all)
    trainable = None

    @property
    def trainable_indices(self):
        n = len(np.asarray(self.params))
        return list(range(n)) if self.trainable is None else list(self.trainable)

    @property
    def trainable_variables(self):
        p = np.asarray(self.params, dtype=np.float64)
        return [p[i] for i in self.trainable_indices]

    def assign_trainable(self, values):
        p = np.asarray(self.params, dtype=np.float64).copy()
        for i, v in zip(self.trainable_indices, values):
            p[i] = float(v)
        keep = self.trainable
        self.__dict__.update(self.with_params(p).__dict__)
        self.trainable = keep

    def param_grads(self, x):
        """
        (df/dp ``[P, ..., D]``, dJ/dp ``[P, ..., D, D]``) for every entry p of ``params``: what a tape watching
        ``prior_sde.trainable_variables`` propagates through ``drift`` / ``gradient_drift``
        (variational_cvi_sde.py:495-518, vi_sde.py:457-470).  Closed forms, checked against central
        differences in tests/test_oracle_prior_grads.py.
        """
        raise NotImplementedError

    def diffusion(self, x, t=None):
        return np.ones_like(x[..., None]) * chol(self.q)

    def gradient_drift(self, x, t=None):
        J = self.jacobian(x)
        if self.full_jacobian:
            return J
        return np.sum(J, axis=-2)  # d(sum_i f_i)/dx_j

    def expected_drift(self, q_mean, q_covar):
        """sde.py:92-112 (H=10)."""
        n_batch, n_states, D = q_mean.shape
        val = mvnquad(lambda x: self.drift(x), q_mean.reshape(-1, D), q_covar.reshape(-1, D, D), H=10, Din=D,
                      Dout=(D,))
        return val.reshape(n_batch, n_states, D)

    def expected_gradient_drift(self, q_mean, q_covar):
        """sde.py:114-131 and the Van der Pol override sde.py:501-518 (H=10)."""
        n_batch, n_states, D = q_mean.shape
        Dout = (D, D) if self.full_jacobian else (D,)
        val = mvnquad(self.gradient_drift, q_mean.reshape(-1, D), q_covar.reshape(-1, D, D), H=10, Din=D, Dout=Dout)
        return val.reshape((n_batch, n_states) + Dout)


def _diag_jac(g):
    """[..., D] elementwise derivative -> [..., D, D] diagonal Jacobian."""
    D = g.shape[-1]
    return g[..., :, None] * np.eye(D)


class OrnsteinUhlenbeckSDE(SDE):
    """sde.py:134-176: f(x) = -decay x."""
    drift_id = DRIFT_OU

    def __init__(self, decay=1.0, q=np.ones((1, 1))):
        super().__init__(q)
        self.decay = float(decay)

    @property
    def params(self):
        return np.array([self.decay])

    def drift(self, x, t=None):
        return -self.decay * x

    def jacobian(self, x):
        return _diag_jac(-self.decay * np.ones_like(x))

    def with_params(self, params):
        return type(self)(decay=params[0], q=self.q)

    def param_grads(self, x):
        return np.stack([-x]), np.stack([_diag_jac(-np.ones_like(x))])


class DoubleWellSDE(SDE):
    """sde.py:179-224: f(x) = scale x (c - x^2)."""
    drift_id = DRIFT_DW

    def __init__(self, q=np.ones((1, 1)), scale=4.0, c=1.0):
        super().__init__(q)
        self.scale = float(scale)
        self.c = float(c)

    @property
    def params(self):
        return np.array([self.scale, self.c])

    def drift(self, x, t=None):
        return self.scale * x * (self.c - np.square(x))

    def jacobian(self, x):
        return _diag_jac(self.scale * (self.c - 3.0 * np.square(x)))

    def with_params(self, params):
        return type(self)(q=self.q, scale=params[0], c=params[1])

    def param_grads(self, x):
        x2 = np.square(x)
        return (np.stack([x * (self.c - x2), self.scale * x]),
                np.stack([_diag_jac(self.c - 3.0 * x2), _diag_jac(self.scale * np.ones_like(x))]))


class BenesSDE(SDE):
    """sde.py:227-268: f(x) = theta tanh(x)."""
    drift_id = DRIFT_BENES

    def __init__(self, theta=1.0, q=np.ones((1, 1))):
        super().__init__(q)
        self.theta = float(theta)

    @property
    def params(self):
        return np.array([self.theta])

    def drift(self, x, t=None):
        return self.theta * np.tanh(x)

    def jacobian(self, x):
        return _diag_jac(self.theta * (1.0 - np.square(np.tanh(x))))

    def with_params(self, params):
        return type(self)(theta=params[0], q=self.q)

    def param_grads(self, x):
        t = np.tanh(x)
        return np.stack([t]), np.stack([_diag_jac(1.0 - np.square(t))])


class SineDiffusionSDE(SDE):
    """sde.py:271-312: f(x) = sin(x - theta)."""
    drift_id = DRIFT_SINE

    def __init__(self, theta=0.0, q=np.ones((1, 1))):
        super().__init__(q)
        self.theta = float(theta)

    @property
    def params(self):
        return np.array([self.theta])

    def drift(self, x, t=None):
        return np.sin(x - self.theta)

    def jacobian(self, x):
        return _diag_jac(np.cos(x - self.theta))

    def with_params(self, params):
        return type(self)(theta=params[0], q=self.q)

    def param_grads(self, x):
        return np.stack([-np.cos(x - self.theta)]), np.stack([_diag_jac(np.sin(x - self.theta))])


class SqrtDiffusionSDE(SDE):
    """sde.py:315-356: f(x) = sqrt(theta |x|)."""
    drift_id = DRIFT_SQRT

    def __init__(self, theta=1.0, q=np.ones((1, 1))):
        super().__init__(q)
        self.theta = float(theta)

    @property
    def params(self):
        return np.array([self.theta])

    def drift(self, x, t=None):
        return np.sqrt(self.theta * np.abs(x))

    def jacobian(self, x):
        # d/dx sqrt(theta |x|) = theta sign(x) / (2 sqrt(theta |x|))   (tape: sqrt' * theta * sign)
        return _diag_jac(0.5 / np.sqrt(self.theta * np.abs(x)) * self.theta * np.sign(x))

    def with_params(self, params):
        return type(self)(theta=params[0], q=self.q)

    def param_grads(self, x):
        r = np.sqrt(self.theta * np.abs(x))
        return np.stack([0.5 * np.abs(x) / r]), np.stack([_diag_jac(0.25 * np.sign(x) / r)])


class MLPDrift(SDE):
    """
    sde.py:359-430: 1-3-1 relu MLP on each coordinate.  The reference's Keras layers run in float32
    and cast back (sde.py:411); this restatement (and the CUDA path) evaluates the same network in
    float64 -- a documented deviation, 1e-9 parity is not defined for a float32 network.
    ``weights`` = (w1[3], b1[3], w2[3], b2).
    """
    drift_id = DRIFT_MLP131

    def __init__(self, q=np.ones((1, 1)), weights=None, rng=None):
        super().__init__(q)
        if weights is None:
            rng = np.random.default_rng(0) if rng is None else rng
            weights = np.concatenate([rng.standard_normal(3), np.zeros(3), rng.standard_normal(3), np.zeros(1)])
        self.weights = np.asarray(weights, dtype=np.float64)

    @property
    def params(self):
        return self.weights

    def drift(self, x, t=None):
        w1, b1, w2, b2 = self.weights[0:3], self.weights[3:6], self.weights[6:9], self.weights[9]
        h = np.maximum(x[..., None] * w1 + b1, 0.0)
        return h @ w2 + b2

    def jacobian(self, x):
        w1, b1, w2 = self.weights[0:3], self.weights[3:6], self.weights[6:9]
        act = (x[..., None] * w1 + b1) > 0.0
        return _diag_jac((act * (w1 * w2)).sum(-1))

    def with_params(self, params):
        return type(self)(q=self.q, weights=np.asarray(params, dtype=np.float64))

    def param_grads(self, x):
        w1, b1, w2 = self.weights[0:3], self.weights[3:6], self.weights[6:9]
        pre = x[..., None] * w1 + b1
        on = (pre > 0.0).astype(np.float64)          # relu'(0) = 0, as TensorFlow
        df, dJ = [], []
        for h in range(3):                           # d/dw1[h]
            df.append(on[..., h] * x * w2[h]); dJ.append(on[..., h] * w2[h])
        for h in range(3):                           # d/db1[h]
            df.append(on[..., h] * w2[h]); dJ.append(np.zeros_like(x))
        for h in range(3):                           # d/dw2[h]
            df.append(np.maximum(pre[..., h], 0.0)); dJ.append(on[..., h] * w1[h])
        df.append(np.ones_like(x)); dJ.append(np.zeros_like(x))   # d/db2
        return np.stack(df), np.stack([_diag_jac(g) for g in dJ])


class VanderPolOscillatorSDE(SDE):
    """sde.py:433-518: f = tau [a (x1 - x1^3/3 - x2), x1 / a]."""
    drift_id = DRIFT_VDP
    full_jacobian = True

    def __init__(self, a=1.0, tau=1.0, q=np.eye(2)):
        super().__init__(q)
        self.a = float(a)
        self.tau = float(tau)

    @property
    def params(self):
        return np.array([self.a, self.tau])

    def drift(self, x, t=None):
        dx1 = self.a * (x[..., 0] - (1 / 3) * np.power(x[..., 0], 3) - x[..., 1])
        dx2 = (1 / self.a) * x[..., 0]
        return self.tau * np.stack([dx1, dx2], axis=-1)

    def jacobian(self, x):
        J = np.zeros(x.shape + (2,))
        J[..., 0, 0] = self.tau * self.a * (1.0 - np.square(x[..., 0]))
        J[..., 0, 1] = -self.tau * self.a
        J[..., 1, 0] = self.tau / self.a
        return J

    def with_params(self, params):
        return type(self)(a=params[0], tau=params[1], q=self.q)

    def param_grads(self, x):
        a, tau = self.a, self.tau
        x1, x2 = x[..., 0], x[..., 1]
        g = x1 - (1 / 3) * np.power(x1, 3) - x2
        df = np.stack([np.stack([tau * g, -tau * x1 / a ** 2], -1), np.stack([a * g, x1 / a], -1)])
        dJ = np.zeros((2,) + x.shape + (2,))
        dJ[0, ..., 0, 0] = tau * (1.0 - np.square(x1)); dJ[0, ..., 0, 1] = -tau; dJ[0, ..., 1, 0] = -tau / a ** 2
        dJ[1, ..., 0, 0] = a * (1.0 - np.square(x1)); dJ[1, ..., 0, 1] = -a; dJ[1, ..., 1, 0] = 1.0 / a
        return df, dJ


class Lorenz63SDE(SDE):
    """
    EXTENSION (no 3-D SDE exists in the reference; BASELINE config 5).  Written the way the
    reference's Van der Pol class is (batch Jacobian): f = [s (y - x), x (r - z) - y, x y - b z].
    """
    drift_id = DRIFT_LORENZ63
    full_jacobian = True

    def __init__(self, sigma=10.0, rho=28.0, beta=8.0 / 3.0, q=np.eye(3)):
        super().__init__(q)
        self.sigma, self.rho, self.beta = float(sigma), float(rho), float(beta)

    @property
    def params(self):
        return np.array([self.sigma, self.rho, self.beta])

    def drift(self, x, t=None):
        X, Y, Z = x[..., 0], x[..., 1], x[..., 2]
        return np.stack([self.sigma * (Y - X), X * (self.rho - Z) - Y, X * Y - self.beta * Z], axis=-1)

    def jacobian(self, x):
        X, Y, Z = x[..., 0], x[..., 1], x[..., 2]
        J = np.zeros(x.shape + (3,))
        J[..., 0, 0] = -self.sigma
        J[..., 0, 1] = self.sigma
        J[..., 1, 0] = self.rho - Z
        J[..., 1, 1] = -1.0
        J[..., 1, 2] = -X
        J[..., 2, 0] = Y
        J[..., 2, 1] = X
        J[..., 2, 2] = -self.beta
        return J

    def with_params(self, params):
        return type(self)(sigma=params[0], rho=params[1], beta=params[2], q=self.q)

    def param_grads(self, x):
        X, Y, Z = x[..., 0], x[..., 1], x[..., 2]
        zero = np.zeros_like(X)
        df = np.stack([np.stack([Y - X, zero, zero], -1), np.stack([zero, X, zero], -1),
                       np.stack([zero, zero, -Z], -1)])
        dJ = np.zeros((3,) + x.shape + (3,))
        dJ[0, ..., 0, 0] = -1.0; dJ[0, ..., 0, 1] = 1.0
        dJ[1, ..., 1, 0] = 1.0
        dJ[2, ..., 2, 2] = -1.0
        return df, dJ


class OUNdSDE(SDE):
    """EXTENSION: D-dimensional OU, f(x) = -decay x, with a full (diagonal) batch Jacobian."""
    drift_id = DRIFT_OU_ND
    full_jacobian = True

    def __init__(self, decay=0.5, q=np.eye(3)):
        super().__init__(q)
        self.decay = float(decay)

    @property
    def params(self):
        return np.array([self.decay])

    def drift(self, x, t=None):
        return -self.decay * x

    def jacobian(self, x):
        return _diag_jac(-self.decay * np.ones_like(x))

    def with_params(self, params):
        return type(self)(decay=params[0], q=self.q)

    def param_grads(self, x):
        return np.stack([-x]), np.stack([_diag_jac(-np.ones_like(x))])


# ----------------------------------------------------------------------------------------------
# drift.py
# ----------------------------------------------------------------------------------------------


class Gaussian:
    """gpflow.probability_distributions.Gaussian: a (mu, cov) pair."""

    def __init__(self, mu, cov):
        self.mu = np.asarray(mu, dtype=np.float64)
        self.cov = np.asarray(cov, dtype=np.float64)


def handle_tensor_shape(tensor, desired_dimensions=2):
    """sde_utils.py:99-116."""
    tensor = np.asarray(tensor)
    if tensor.ndim == desired_dimensions:
        return tensor
    if tensor.ndim == desired_dimensions + 1 and tensor.shape[0] == 1:
        return tensor[0]
    if tensor.ndim == desired_dimensions - 1:
        return tensor[None]
    raise Exception("Batch present!")


class LinearDrift:
    """drift.py:24-116: f_L(x, t) = A_t x + b_t."""

    def __init__(self, A=None, b=None):
        self.A = A
        self.b = b

    def set_from_ssm(self, ssm: StateSpaceModel, dt: float):
        st = handle_tensor_shape(ssm.state_transitions, 3)
        so = handle_tensor_shape(ssm.state_offsets, 2)
        self.A = (st - np.eye(ssm.state_dim)) / dt
        self.b = so / dt

    def to_ssm(self, q, transition_times, initial_mean, initial_chol_covariance):
        if self.A is None or self.b is None:
            raise Exception("Linear drift is empty so SSM can't be created!")
        self.A = handle_tensor_shape(self.A, 4)
        self.b = handle_tensor_shape(self.b, 3)
        transition_times = handle_tensor_shape(transition_times, 1)
        q = handle_tensor_shape(q, 4)
        initial_mean = handle_tensor_shape(initial_mean, 2)
        initial_chol_covariance = handle_tensor_shape(initial_chol_covariance, 3)
        deltas = np.reshape(transition_times[1:] - transition_times[:-1], (1, -1, 1))
        state_transitions = self.A * deltas[..., None] + np.eye(self.A.shape[-1])
        state_offsets = self.b * deltas
        chol_process_covariances = chol(q * deltas[..., None])
        return StateSpaceModel(initial_mean, initial_chol_covariance, state_transitions, state_offsets,
                               chol_process_covariances)


# ----------------------------------------------------------------------------------------------
# sde_utils.py
# ----------------------------------------------------------------------------------------------


def euler_maruyama(sde: SDE, x0, time_grid, rng):
    """sde_utils.py:36-96 with a NumPy generator in place of tf.random.normal (one draw per step)."""
    x0 = np.asarray(x0, dtype=np.float64)
    time_grid = np.asarray(time_grid, dtype=np.float64)
    n = time_grid.shape[0]
    vals = np.empty((n,) + x0.shape)
    x = x0
    t = 0.0
    for k in range(n):
        dt = time_grid[k] - t
        diffusion_term = sde.diffusion(x, None) * np.sqrt(dt)
        x = x + sde.drift(x, None) * dt + (diffusion_term @ rng.standard_normal(x.shape)[..., None])[..., 0]
        t = time_grid[k]
        vals[k] = x
    vals = np.transpose(vals, (1, 0, 2))
    return np.concatenate([x0[:, None, :], vals[:, :-1, :]], axis=1)


def remove_batch_from_ssm(ssm: StateSpaceModel):
    """sde_utils.py:464-470."""
    return StateSpaceModel(ssm.initial_mean[0], ssm.cholesky_initial_covariance[0], ssm.state_transitions[0],
                           ssm.state_offsets[0], ssm.cholesky_process_covariances[0])


def linearize_sde(sde: SDE, transition_times, linearization_path: Gaussian, initial_state: Gaussian):
    """sde_utils.py:119-179."""
    q_mean = handle_tensor_shape(linearization_path.mu, 3)
    q_covar = handle_tensor_shape(linearization_path.cov, 4)
    initial_mean = handle_tensor_shape(initial_state.mu, 2)
    initial_chol_covariance = handle_tensor_shape(chol(initial_state.cov), 3)
    B, N, D = q_mean.shape
    E_f = sde.expected_drift(q_mean, q_covar)
    A = sde.expected_gradient_drift(q_mean, q_covar)
    if A.shape == (B, N, D) and D == 1:
        A = A[..., None]
    b = E_f - (A @ q_mean[..., None])[..., 0]
    chol_q = sde.diffusion(q_mean, None)
    return LinearDrift(A=A, b=b).to_ssm(q=chol_q @ chol_q, transition_times=transition_times,
                                        initial_mean=initial_mean, initial_chol_covariance=initial_chol_covariance)


def squared_drift_difference_along_Gaussian_path(sde_p: SDE, linear_drift: LinearDrift, q: Gaussian, dt: float,
                                                 quadrature_pnts: int = 20):
    """sde_utils.py:182-249: 0.5 dt sum_t E_q || A_t x + b_t - f(x) ||^2_{q^-1}."""
    m = handle_tensor_shape(q.mu, 2)
    S = handle_tensor_shape(q.cov, 3)
    D = S.shape[-1]
    n = S.shape[0]
    A, b = linear_drift.A, linear_drift.b
    sigma_inv = chol_solve(chol(sde_p.q), np.eye(D))

    def func(x):
        x = x.reshape(-1, n, D)
        tmp = ((A[None] @ x[..., None])[..., 0] + b[None]) - sde_p.drift(x)
        return np.einsum("pni,ij,pnj->pn", tmp, sigma_inv, tmp).reshape(-1)

    val = mvnquad(func, m, S, H=quadrature_pnts, Din=D)
    return 0.5 * np.sum(val) * dt


def grad_squared_drift_difference(sde_p: SDE, A, b, m, S, dt: float, H: int = 20):
    """
    Value and (m, S)-gradient of squared_drift_difference_along_Gaussian_path, i.e. what the tape
    of vi_sde.py:206-225 returns *before* the division by dt.  Hand adjoint through the
    discretised quadrature and chol(S).
    """
    m = np.asarray(m, dtype=np.float64)
    S = np.asarray(S, dtype=np.float64)
    n, D = m.shape
    xn, wn = mvhermgauss(H, D)
    wt = wn * np.pi ** (-D * 0.5)
    sigma_inv = chol_solve(chol(sde_p.q), np.eye(D))
    L = chol(S)
    X = 2.0 ** 0.5 * np.einsum("nij,pj->pni", L, xn) + m[None]
    d = (np.einsum("nij,pnj->pni", A, X) + b[None]) - sde_p.drift(X)
    r = np.einsum("ij,pnj->pni", sigma_inv, d)
    val = 0.5 * dt * np.sum(wt[:, None] * np.einsum("pni,pni->pn", d, r))
    # d val / d d_p = dt * w_p * sigma_inv d_p
    rb = dt * wt[:, None, None] * r
    Jd = A[None] - sde_p.jacobian(X)  # d d / d X
    Xb = np.einsum("pnji,pnj->pni", Jd, rb)
    dm = np.sum(Xb, axis=0)
    Lb = np.tril(2.0 ** 0.5 * np.einsum("pni,pj->nij", Xb, xn))
    dS = cholesky_adjoint(L, Lb)
    return val, dm, dS


def cholesky_adjoint(L, Lbar):
    """
    Reverse-mode of L = chol(S) as TensorFlow registers it (I. Murray 2016):
    Sbar = 0.5 (P + P^T),  P = L^-T Phi(L^T Lbar) L^-1,  Phi = lower triangle with halved diagonal.
    """
    middle = mT(L) @ Lbar
    D = L.shape[-1]
    middle = np.tril(middle)
    idx = np.arange(D)
    middle[..., idx, idx] *= 0.5
    # L^-T middle L^-1
    P = tri_solve(L, middle, adjoint=True)          # L^-T middle
    P = mT(tri_solve(L, mT(P), adjoint=True))       # (.) L^-1 = (L^-T (.)^T)^T
    return 0.5 * (P + mT(P))


def expectations_to_ssm_params_vjp(exp1, exp_diag, exp_sub_diag, g_A, g_b, g_cholP0, g_cholQ, g_mu0):
    """
    Vector-Jacobian product of expectations_to_ssm_params (ssm_gaussian_transformations.py:93-178) -- what
    ``tape.gradient(ssm_params, etas, output_gradients=dL_dssm)`` returns at ssm_natgrad.py:168.  Same
    statement-by-statement adjoint as in :func:`sde_ssm_kl_and_grads`; the gradient through a Cholesky is the
    symmetrised one TensorFlow registers.  Cotangents of the Cholesky factors are lower triangular.
    """
    m = np.asarray(exp1, dtype=np.float64)
    sym = lambda X: 0.5 * (X + mT(X))
    mcol = m[..., None]
    S = exp_diag - mcol @ mT(mcol)
    Csub = mT(exp_sub_diag) - mcol[:-1] @ mT(mcol[1:])
    LS = chol(S)
    A = mT(chol_solve(LS[:-1], Csub))
    Qc = S[1:] - A @ (S[:-1] @ mT(A))
    LQ = chol(Qc)
    g_m = np.zeros_like(m)
    g_S = np.zeros_like(S)
    g_A = np.array(g_A, dtype=np.float64)
    g_b = np.asarray(g_b, dtype=np.float64)
    g_Qc = cholesky_adjoint(LQ, np.tril(g_cholQ))
    g_m[1:] += g_b
    g_A += -g_b[..., None] @ mT(mcol[:-1])
    g_m[:-1] += -(mT(A) @ g_b[..., None])[..., 0]
    g_S[1:] += g_Qc
    g_A += -2.0 * g_Qc @ A @ S[:-1]
    g_S[:-1] += -mT(A) @ g_Qc @ A
    g_C = chol_solve(LS[:-1], mT(g_A))
    g_S[:-1] += sym(-g_C @ A)
    g_m[0] += g_mu0
    g_S[0] += cholesky_adjoint(LS[0], np.tril(g_cholP0))
    g_eta_s = mT(g_C)
    g_m[:-1] += -(g_C @ mcol[1:])[..., 0]
    g_m[1:] += -(mT(g_C) @ mcol[:-1])[..., 0]
    g_m += -2.0 * (g_S @ mcol)[..., 0]
    return g_m, g_S, g_eta_s


def gaussian_kl(mu_q, cov_q, mu_p, cov_p):
    """tfp kl_divergence(MultivariateNormalFullCovariance q, p) in closed form."""
    D = mu_q.shape[-1]
    Lp = chol(cov_p)
    Lq = chol(cov_q)
    M = tri_solve(Lp, Lq)
    diff = tri_solve(Lp, (mu_p - mu_q)[..., None])[..., 0]
    logdet_p = 2.0 * np.sum(np.log(np.diagonal(Lp, axis1=-2, axis2=-1)), axis=-1)
    logdet_q = 2.0 * np.sum(np.log(np.diagonal(Lq, axis1=-2, axis2=-1)), axis=-1)
    return 0.5 * (np.sum(M * M, axis=(-1, -2)) + np.sum(diff * diff, axis=-1) - D + logdet_p - logdet_q)


def ssm_kl_constant_terms(Q_q, Q_p):
    """
    sde_utils.py:317-330.  C = -(logdet Q_q - logdet Q_p) - D + trace(Q_p^-1 * Q_q) where ``*`` is the
    reference's ELEMENT-WISE product (so the "trace" is sum_i (Q_p^-1)_ii (Q_q)_ii).  tf.linalg.logdet
    is 2 sum log diag chol(.) and reads the lower triangle only.
    """
    D = Q_q.shape[-1]
    Qp_inv = chol_solve(chol(Q_p), np.broadcast_to(np.eye(D), Q_p.shape))
    ld_q = 2.0 * np.sum(np.log(np.diagonal(chol(Q_q), axis1=-2, axis2=-1)), axis=-1)
    ld_p = 2.0 * np.sum(np.log(np.diagonal(chol(Q_p), axis1=-2, axis2=-1)), axis=-1)
    tr = np.trace(Qp_inv * Q_q, axis1=-2, axis2=-1)
    return -(ld_q - ld_p) - D + tr, Qp_inv


def SSM_KL_along_Gaussian_path(func_q, func_p, ssm_q_process_covar, ssm_p_process_covar, ssm_q_marginals_mean,
                               ssm_q_marginals_covar, quadrature_pnts: int = 20):
    """sde_utils.py:262-359."""
    n = ssm_p_process_covar.shape[0]
    D = ssm_p_process_covar.shape[-1]
    C, Qp_inv = ssm_kl_constant_terms(ssm_q_process_covar, ssm_p_process_covar)

    def func(x):
        x = x.reshape(-1, n, D)
        diff = func_p(x) - func_q(x)
        return np.einsum("pni,nij,pnj->pn", diff, Qp_inv, diff).reshape(-1)

    fn_difference = mvnquad(func, ssm_q_marginals_mean[:-1], ssm_q_marginals_covar[:-1], H=quadrature_pnts, Din=D)
    return 0.5 * np.sum(fn_difference + C)


def ssm_to_btd_nat(ssm: StateSpaceModel) -> BTDGaussian:
    """sde_utils.py:362-373."""
    nat1, nat_diag, nat_subdiag = ssm_to_naturals(ssm)
    return BTDGaussian(nat1, nat_diag, nat_subdiag)


def tranform_girsanov_sites(girsanov_sites: BTDGaussian, current_prior: StateSpaceModel,
                            new_prior: StateSpaceModel) -> BTDGaussian:
    """sde_utils.py:550-568."""
    new = ssm_to_btd_nat(new_prior)
    cur = ssm_to_btd_nat(current_prior)
    return BTDGaussian(
        girsanov_sites.nat1 + cur.nat1 - new.nat1,
        girsanov_sites.nat2.block_diagonal + cur.nat2.block_diagonal - new.nat2.block_diagonal,
        girsanov_sites.nat2.block_sub_diagonal + cur.nat2.block_sub_diagonal - new.nat2.block_sub_diagonal)


def sde_ssm_kl_forward(exp1, exp_diag, exp_sub_diag, sde_p: SDE, dt: float, prior_initial_state: Gaussian,
                       transition_times, H: int = 20):
    """
    The taped forward pass of SDE_SSM_KL_with_grads_wrt_exp_params (sde_utils.py:496-543) as a
    plain function of the expectation parameters.  Returns the KL value.
    """
    A, b, chol_P0, chol_Q, mu0 = expectations_to_ssm_params(exp1, exp_diag, exp_sub_diag)
    covar = exp_diag - exp1[..., None] @ mT(exp1[..., None])
    Q_q = chol_Q @ mT(chol_Q)
    deltas = (transition_times[1:] - transition_times[:-1])[..., None, None]
    Q_p = deltas * np.repeat(sde_p.q[None], deltas.shape[0], axis=0)

    def f_p(x):
        return x + dt * sde_p.drift(x)

    def f_q(x):
        return (A[None] @ x[..., None])[..., 0] + b[None]

    kl_q_p = SSM_KL_along_Gaussian_path(f_q, f_p, Q_q, Q_p, exp1, covar, quadrature_pnts=H)
    kl_0 = gaussian_kl(mu0, chol_P0 @ mT(chol_P0), prior_initial_state.mu, prior_initial_state.cov)
    return kl_0 + kl_q_p


def sde_ssm_kl_and_grads(exp1, exp_diag, exp_sub_diag, sde_p: SDE, dt: float, prior_initial_state: Gaussian,
                         transition_times, H: int = 20):
    """
    Value and reverse-mode gradient of :func:`sde_ssm_kl_forward` with respect to the three
    expectation parameters -- what ``g.gradient(kl_val, [exp1, exp_diag, exp_sub_diag])`` returns at
    sde_utils.py:545.  Hand-written adjoint, transition by transition:

      forward (k = 0..N-1):  S_k = eta_d,k - m_k m_k^T,  C_k = eta_s,k^T - m_k m_{k+1}^T,
                             A_k = (S_k^-1 C_k)^T,  b_k = m_{k+1} - A_k m_k,
                             Qc_k = S_{k+1} - A_k S_k A_k^T,  L_Q = chol(Qc_k),  Q_q = L_Q L_Q^T,
                             X = m_k + sqrt2 chol(S_k) xi,  d = X + dt f(X) - A_k X - b_k,
                             term_k = 0.5 ( sum_p w_p d^T Qp^-1 d  - logdet Q_q + logdet Q_p - D
                                            + sum_i (Qp^-1)_ii (Q_q)_ii )
      the gradient with respect to a symmetric matrix that enters through a Cholesky is the
      symmetrised one (that is what the registered Cholesky gradient returns).
    """
    m = np.asarray(exp1, dtype=np.float64)
    eta_d = np.asarray(exp_diag, dtype=np.float64)
    eta_s = np.asarray(exp_sub_diag, dtype=np.float64)
    n1, D = m.shape
    N = n1 - 1
    transition_times = np.asarray(transition_times, dtype=np.float64)

    sym = lambda M: 0.5 * (M + mT(M))
    I = np.eye(D)

    # ---- forward
    mcol = m[..., None]
    S = eta_d - mcol @ mT(mcol)                       # [N+1, D, D]
    Csub = mT(eta_s) - mcol[:-1] @ mT(mcol[1:])       # Sigma_{k,k+1}
    LS = chol(S)
    A = mT(chol_solve(LS[:-1], Csub))
    b = (mcol[1:] - A @ mcol[:-1])[..., 0]
    Qc = S[1:] - A @ (S[:-1] @ mT(A))
    LQ = chol(Qc)
    Q_q = LQ @ mT(LQ)
    deltas = (transition_times[1:] - transition_times[:-1])[..., None, None]
    Q_p = deltas * np.repeat(sde_p.q[None], N, axis=0)
    Cterm, Qp_inv = ssm_kl_constant_terms(Q_q, Q_p)

    xn, wn = mvhermgauss(H, D)
    wt = wn * np.pi ** (-D * 0.5)
    X = 2.0 ** 0.5 * np.einsum("nij,pj->pni", LS[:-1], xn) + m[None, :-1]      # [P, N, D]
    fX = sde_p.drift(X)
    d = (X + dt * fX) - (np.einsum("nij,pnj->pni", A, X) + b[None])
    r = np.einsum("nij,pnj->pni", Qp_inv, d)
    quad = np.sum(wt[:, None] * np.einsum("pni,pni->pn", d, r), axis=0)
    kl_path = 0.5 * np.sum(quad + Cterm)
    P0 = LS[0] @ LS[0].T
    kl0 = gaussian_kl(m[0], P0, prior_initial_state.mu, prior_initial_state.cov)
    kl = kl0 + kl_path

    # ---- reverse
    g_m = np.zeros_like(m)
    g_S = np.zeros_like(S)          # adjoint of S_k (symmetric matrices)
    # quadrature: d term / d d_p = w_p (Qp^-1 + Qp^-T)/2 * 2 * 0.5 d = w_p sym(Qp^-1) d
    Qp_inv_s = sym(Qp_inv)
    rb = wt[:, None, None] * np.einsum("nij,pnj->pni", Qp_inv_s, d)            # [P, N, D]
    g_b = -np.sum(rb, axis=0)                                                   # [N, D]
    g_A = -np.einsum("pni,pnj->nij", rb, X)                                     # [N, D, D]
    Jd = (I + dt * sde_p.jacobian(X)) - A[None]                                 # d d / d X
    Xb = np.einsum("pnji,pnj->pni", Jd, rb)
    g_m[:-1] += np.sum(Xb, axis=0)
    LSb = np.tril(2.0 ** 0.5 * np.einsum("pni,pj->nij", Xb, xn))
    g_S[:-1] += cholesky_adjoint(LS[:-1], LSb)
    # constant terms: d/dQ_q [0.5(-logdet Q_q + sum_i Qpinv_ii Qq_ii)], symmetrised through chol(Qc)
    Qq_inv = chol_solve(LQ, np.broadcast_to(I, LQ.shape))
    g_Qc = sym(0.5 * (-Qq_inv + Qp_inv * I))
    # b = m_{k+1} - A m_k
    g_m[1:] += g_b
    g_A += -g_b[..., None] @ mT(mcol[:-1])
    g_m[:-1] += -(mT(A) @ g_b[..., None])[..., 0]
    # Qc = S_{k+1} - A S_k A^T
    g_S[1:] += g_Qc
    g_A += -2.0 * g_Qc @ A @ S[:-1]
    g_S[:-1] += -mT(A) @ g_Qc @ A
    # A = C^T S_k^-1  (through cholesky_solve(chol(S_k), C)): Y = S^-1 C, A = Y^T
    g_Y = mT(g_A)
    g_C = chol_solve(LS[:-1], g_Y)
    g_S[:-1] += sym(-g_C @ A)            # -S^-1 Ybar Y^T, Y^T = A
    # KL_0 (mu0 = m_0, P0 = S_0)
    Pp_inv = chol_solve(chol(prior_initial_state.cov), I)
    g_m[0] += Pp_inv @ (m[0] - prior_initial_state.mu)
    g_S[0] += 0.5 * (Pp_inv - chol_solve(LS[0], I))
    # C = eta_s^T - m_k m_{k+1}^T
    g_eta_s = mT(g_C)
    g_m[:-1] += -(g_C @ mcol[1:])[..., 0]
    g_m[1:] += -(mT(g_C) @ mcol[:-1])[..., 0]
    # S = eta_d - m m^T
    g_eta_d = g_S
    g_m += -2.0 * (g_S @ mcol)[..., 0]
    return kl, (g_m, g_eta_d, g_eta_s)


def SDE_SSM_KL_with_grads_wrt_exp_params(ssm_q: StateSpaceModel, sde_p: SDE, dt: float,
                                         prior_initial_state: Gaussian, transition_times):
    """sde_utils.py:473-547."""
    exp1, exp_diag, exp_sub_diag = ssm_to_expectations(ssm_q)
    return sde_ssm_kl_and_grads(exp1, exp_diag, exp_sub_diag, sde_p, dt, prior_initial_state, transition_times)
