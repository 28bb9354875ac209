"""
Batched VDP engine (Archambeau et al. 2007): device-resident state of B independent sequences and the
fixed-point sweep of ``markovflow/models/vi_sde.py`` (VariationalMarkovGP), executed by the
``vdp_vgp_*`` entries of the C ABI.  One :meth:`VgpEngine.sweep` = one iteration of
``VIMarkovGPTrainer.perform_inference`` (docs/diffusion_processes/vi_markov_gp_trainer.py:62-73).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, ops
from ._lib import VgpProblem, check

F64 = torch.float64


class VgpEngine:
    """
    State (float64 CUDA tensors, reference layouts with a leading batch axis):
      A [B,N,D,D], b [B,N,D]                 variational drift            (vi_sde.py:84-85)
      lambda_lagrange [B,N,D], psi_lagrange [B,N,D,D]                      (vi_sde.py:100-103)
      q_initial_mean [B,D], q_initial_chol [B,D,D]                         (vi_sde.py:96-98)
      m [B,N+1,D], S [B,N+1,D,D]             marginals of the last forward pass
    """

    def __init__(self, drift_id, drift_params, q, time_grid, obs_idx, obs_y, lik_chol, x0_mu=None, x0_cov=None,
                 device="cuda", stabilize_system=False):
        self.device = dev = torch.device(device)
        self.drift_id = int(drift_id)
        self.drift_params = [float(v) for v in np.asarray(drift_params).reshape(-1)]
        self.q = np.atleast_2d(np.asarray(q, dtype=np.float64))
        self.D = D = self.q.shape[0]
        tg = np.asarray(time_grid, dtype=np.float64)
        self.dt = float(tg[1] - tg[0])
        self.time_grid = torch.as_tensor(tg, dtype=F64).to(dev)
        self.N = N = tg.shape[0] - 1
        obs_idx = np.asarray(obs_idx, dtype=np.int64)
        obs_y = np.asarray(obs_y, dtype=np.float64)
        if obs_idx.ndim == 1:
            obs_idx, obs_y = obs_idx[None], obs_y[None]
        self.B = B = obs_idx.shape[0]
        self.n_obs = n_obs = obs_idx.shape[1]
        if obs_idx.size and (obs_idx.min() < 0 or obs_idx.max() > N):
            raise _lib.VdpError(f"observation grid indices must lie in [0, {N}] (got {obs_idx.min()}..{obs_idx.max()})")
        self.obs_idx = torch.as_tensor(obs_idx).to(dev).contiguous()
        self.obs_y = torch.as_tensor(obs_y.reshape(B, n_obs, D), dtype=F64).to(dev).contiguous()
        self.lik_chol = np.atleast_2d(np.asarray(lik_chol, dtype=np.float64))
        # prior on x0 (vi_sde.py:88-94): N(0, q I) when not given
        self.x0_mu = np.zeros(D) if x0_mu is None else np.asarray(x0_mu, dtype=np.float64).reshape(D)
        self.x0_cov = self.q * np.eye(D) if x0_cov is None else np.atleast_2d(np.asarray(x0_cov, dtype=np.float64))
        self.stabilize_system = bool(stabilize_system)
        self.route_flags = 0          # VGP_GH_QUADRATURE: OR-ed into every call
        self.carry_state = True       # consecutive sweeps start from the previous sweep's forward pass (see sweep)
        self._carry_sig = None
        z = lambda *s: torch.zeros(s, dtype=F64, device=dev)
        eye = torch.eye(D, dtype=F64, device=dev)
        self.A, self.b = z(B, N, D, D), z(B, N, D)
        self.lambda_lagrange = z(B, N, D)
        self.psi_lagrange = (eye * 1e-10).expand(B, N, D, D).contiguous()
        self.q_initial_mean = torch.as_tensor(self.x0_mu, dtype=F64).to(dev).expand(B, D).contiguous()
        self.q_initial_chol = torch.as_tensor(np.linalg.cholesky(self.x0_cov), dtype=F64).to(dev).expand(
            B, D, D).contiguous()
        self.m, self.S = z(B, N + 1, D), z(B, N + 1, D, D)
        self.elbo_value = z(B)
        self.status = ops.new_status(dev)
        lib = _lib.load()
        self._ws_bytes = lib.vdp_vgp_workspace(B, N, D, n_obs)
        self._ws = torch.empty(self._ws_bytes, dtype=torch.uint8, device=dev)

    def _problem(self, lr=0.0, x0_lr=0.0, update_x0=False):
        p = VgpProblem()
        p.B, p.N, p.n_obs, p.D, p.drift_id = self.B, self.N, self.n_obs, self.D, self.drift_id
        for i, v in enumerate(self.drift_params[:16]):
            p.drift_params[i] = v
        for i, v in enumerate(self.q.reshape(-1)):
            p.q[i] = v
        p.dt = self.dt
        for i, v in enumerate(self.x0_mu):
            p.x0_mu[i] = v
        for i, v in enumerate(self.x0_cov.reshape(-1)):
            p.x0_cov[i] = v
        for i, v in enumerate(self.lik_chol.reshape(-1)):
            p.lik_chol[i] = v
        p.lr, p.x0_lr = float(lr), float(x0_lr)
        for field, attr in (("time_grid", "time_grid"), ("A", "A"), ("b", "b"), ("lam", "lambda_lagrange"),
                            ("psi", "psi_lagrange"), ("q_mean", "q_initial_mean"), ("q_chol", "q_initial_chol"),
                            ("obs_idx", "obs_idx"), ("obs_y", "obs_y"), ("m", "m"), ("S", "S"),
                            ("elbo", "elbo_value"), ("status", "status")):
            t = getattr(self, attr)
            assert t.is_contiguous(), attr
            setattr(p, field, t.data_ptr())
        p.workspace = self._ws.data_ptr()
        p.workspace_bytes = self._ws_bytes
        p.flags = ((_lib.VGP_STABILIZE if self.stabilize_system else 0) | (_lib.VGP_UPDATE_X0 if update_x0 else 0) |
                   self.route_flags)
        return p

    def _call(self, name, p):
        self._carry_sig = None                 # any call may touch the workspace / the state a sweep would carry
        with torch.cuda.device(self.device):   # the launches go to the current device
            stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            check(getattr(_lib.load(), name)(C.byref(p), stream), name)

    # vi_sde.py:171-204
    def forward_pass(self):
        self._call("vdp_vgp_forward_pass", self._problem())
        return self.m, self.S

    # vi_sde.py:289-347 (uses the marginals of the last forward pass)
    def update_lagrange(self):
        self._call("vdp_vgp_update_lagrange", self._problem())

    # vi_sde.py:377-414
    def update_param(self, lr):
        self._call("vdp_vgp_update_param", self._problem(lr=lr))

    # vi_sde.py:242-260
    def update_initial_statistics(self, lr):
        self._call("vdp_vgp_update_initial_statistics", self._problem(x0_lr=lr))

    # vi_sde.py:436-455 (marginals of the last forward pass)
    def elbo(self):
        self._call("vdp_vgp_elbo", self._problem())
        return self.elbo_value

    # vi_markov_gp_trainer.py:62-73
    _CARRY_FIELDS = ("A", "b", "q_initial_mean", "q_initial_chol", "m", "S", "obs_idx", "obs_y", "time_grid", "_ws")

    def _carry_signature(self):
        """everything the head of a sweep (forward pass + E_sde quadrature) is a function of: storage and torch version
        counter of the tensors, the scalar parameters, the route"""
        return (tuple((t.data_ptr(), t._version) for t in (getattr(self, n) for n in self._CARRY_FIELDS)),
                tuple(self.drift_params), tuple(self.q.reshape(-1).tolist()), self.dt, self.stabilize_system,
                self.route_flags, tuple(_lib.config_get(k) for k in _lib.CONFIG_KEYS))

    def sweep(self, lr, update_x0=False, x0_lr=0.1, carry=None):
        """One iteration of VIMarkovGPTrainer.perform_inference.  A sweep ends with the forward pass and the E_sde
        quadrature of the parameters it has just written (for its ELBO); the next sweep would begin with the same two
        computations.  If nothing they depend on has changed in between -- checked here: no other call on this engine,
        no tensor replaced or modified through torch -- ``VDP_VGP_STATE_CARRIED`` lets it start from what the
        workspace still holds (bitwise the same results)."""
        carry = self.carry_state if carry is None else carry
        p = self._problem(lr=lr, x0_lr=x0_lr, update_x0=update_x0)
        if carry and self._carry_sig is not None and self._carry_sig == self._carry_signature():
            p.flags |= _lib.VGP_STATE_CARRIED
        self._call("vdp_vgp_sweep", p)
        self._carry_sig = self._carry_signature() if carry else None
        return self.elbo_value

    # vi_sde.py:457-470
    def grad_prior_sde_params(self):
        """d E_sde / d(drift parameters) at (m, S) = forward_pass[1:] held constant -> [B, n_params]."""
        m, S = self.forward_pass()
        return ops.drift_param_grad(self.drift_id, self.drift_params, self.q, 1.0, 0.0, 1.0, -1.0, self.dt, self.A,
                                    self.b, m, S, point_offset=1, status=self.status)

    def sweep_graphed(self, lr, update_x0=False, x0_lr=0.1):
        """:meth:`sweep` replayed from a CUDA graph (launch-bound for single short sequences, see
        ``CviEngine.site_update_graphed``); re-captured when a buffer, a step size or a drift parameter changes."""
        names = ("time_grid", "A", "b", "lambda_lagrange", "psi_lagrange", "q_initial_mean", "q_initial_chol",
                 "obs_idx", "obs_y", "m", "S", "elbo_value", "status")
        key = (float(lr), bool(update_x0), float(x0_lr), self.route_flags, tuple(self.drift_params),
               tuple(_lib.config_get(k) for k in _lib.CONFIG_KEYS), tuple(self.x0_mu.tolist()),
               tuple(self.x0_cov.reshape(-1).tolist()), tuple(self.q.reshape(-1).tolist()),
               tuple(self.lik_chol.reshape(-1).tolist()), self.dt, self.stabilize_system, tuple(getattr(self, n).data_ptr() for n in names))
        cache = getattr(self, "_graphs", None)
        if cache is None:
            cache = self._graphs = {}
        g = cache.get(key)
        if g is None:
            if len(cache) >= 8:
                cache.clear()
            keep = ("A", "b", "lambda_lagrange", "psi_lagrange", "q_initial_mean", "q_initial_chol", "m", "S")
            state = [getattr(self, n).clone() for n in keep]
            self.sweep(lr, update_x0=update_x0, x0_lr=x0_lr, carry=False)   # warm-up outside the capture, undone below
            g = torch.cuda.CUDAGraph()
            with torch.cuda.device(self.device), torch.cuda.graph(g):
                self.sweep(lr, update_x0=update_x0, x0_lr=x0_lr, carry=False)
            for n, t in zip(keep, state):
                getattr(self, n).copy_(t)
            cache[key] = g
        self._carry_sig = None
        g.replay()
        return self.elbo_value

    def check_status(self):
        code, idx = ops.read_status(self.status)
        if code != 0:
            raise _lib.VdpError(f"non positive-definite block at flat grid index {idx} (code {code})")
