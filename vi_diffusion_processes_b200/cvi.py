"""
Batched CVI-DP engine: device-resident state of B independent sequences and the fused site update.

This is the object the reference-shaped model classes (``markovflow/models/variational_cvi_sde.py``
mirror in :mod:`vi_diffusion_processes_b200.markovflow.models`) and ``bench.py`` drive.  One call of
:meth:`CviEngine.site_update` = one iteration of the reference's hot loop
(``docs/diffusion_processes/cvi_dp_trainer.py:71-77``): update_data_sites, update_girsanov_sites,
classic_elbo -- executed by ``vdp_cvi_site_update`` on the GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, ops
from ._lib import CviProblem, check

F64 = torch.float64


def _cfg_on(name):
    return _lib.config_get(name) != 0


class CviEngine:
    """
    State (all float64 CUDA tensors, reference layouts with a leading batch axis):
      girsanov sites  g1 [B,N+1,D], gd [B,N+1,D,D], gs [B,N,D,D]     (variational_cvi_sde.py:135-151)
      data sites      d1 [B,n_obs,D], d2 [B,n_obs,D,D]               (:97-104)
      posterior       fx_mus [B,N+1,D], fx_covs [B,N+1,D,D]          (:110-115)
      linearised prior p (A, b, cholQ, mu0, cholP0), and its un-clipped version p_lin (:408-432)
    """

    def __init__(self, drift_id, drift_params, q, time_grid, obs_idx, obs_y, lik_chol, x0_mu, x0_cov,
                 device="cuda", stabilize_ssm=True, clip_state_transitions=(-1.0, 1.0),
                 initial_mean=None, initial_cov=None, carry_naturals=True, lazy_sites=True):
        self.device = torch.device(device)
        self.carry_naturals = bool(carry_naturals)
        # consecutive carried updates keep the girsanov sites implicitly in theta (VDP_FLAG_SITES_IN_NATURALS) and
        # form g1 / gd / gs when somebody reads them (the g1 / gd / gs properties below)
        self.lazy_sites = bool(lazy_sites) and self.carry_naturals
        self.route_flags = 0          # FLAG_GH_QUADRATURE | FLAG_REFACTOR_PRECISION | ...: OR-ed into every call
        self._sites_stale = False
        self._nat_sig = None
        self.drift_id = int(drift_id)
        self.drift_params = [float(v) for v in np.asarray(drift_params).reshape(-1)]
        self.q = np.atleast_2d(np.asarray(q, dtype=np.float64))
        self.D = D = self.q.shape[0]
        tg = np.asarray(time_grid, dtype=np.float64)
        self.dt = float(tg[1] - tg[0])
        self.time_grid = torch.as_tensor(tg, dtype=F64).to(self.device)
        self.N = N = tg.shape[0] - 1
        obs_idx = np.asarray(obs_idx, dtype=np.int64)
        obs_y = np.asarray(obs_y, dtype=np.float64)
        if obs_idx.ndim == 1:
            obs_idx, obs_y = obs_idx[None], obs_y[None]
        self.B = B = obs_idx.shape[0]
        self.n_obs = n_obs = obs_idx.shape[1]
        if obs_idx.size and (obs_idx.min() < 0 or obs_idx.max() > N):
            raise _lib.VdpError(f"observation grid indices must lie in [0, {N}] (got {obs_idx.min()}..{obs_idx.max()})")
        self.obs_idx = torch.as_tensor(obs_idx).to(self.device).contiguous()
        self.obs_y = torch.as_tensor(obs_y.reshape(B, n_obs, D), dtype=F64).to(self.device).contiguous()
        self.lik_chol = np.atleast_2d(np.asarray(lik_chol, dtype=np.float64))
        self.x0_mu = np.asarray(x0_mu, dtype=np.float64).reshape(D)
        self.x0_cov = np.atleast_2d(np.asarray(x0_cov, dtype=np.float64))
        self.stabilize_ssm = stabilize_ssm
        self.clip = tuple(float(c) for c in clip_state_transitions)
        dev = self.device
        z = lambda *s: torch.zeros(s, dtype=F64, device=dev)
        self._g1 = z(B, N + 1, D)
        self._gd = torch.full((B, N + 1, D, D), -1e-10, dtype=F64, device=dev)
        self._gs = torch.full((B, N, D, D), -1e-10, dtype=F64, device=dev)
        self.d1 = z(B, n_obs, D)
        self.d2 = (torch.eye(D, dtype=F64, device=dev) * 1e-10).expand(B, n_obs, D, D).contiguous()
        self.fx_mus = z(B, N + 1, D) if initial_mean is None else torch.as_tensor(
            initial_mean, dtype=F64).to(dev).reshape(B, N + 1, D).contiguous()
        self.fx_covs = (torch.eye(D, dtype=F64, device=dev).expand(B, N + 1, D, D).contiguous()
                        if initial_cov is None else
                        torch.as_tensor(initial_cov, dtype=F64).to(dev).reshape(B, N + 1, D, D).contiguous())
        self.qA, self.qb, self.qcholQ = z(B, N, D, D), z(B, N, D), z(B, N, D, D)
        self.qmu0, self.qcholP0 = z(B, D), z(B, D, D)
        self.elbo, self.ve, self.kl = z(B), z(B), z(B)
        self.status = ops.new_status(dev)
        lib = _lib.load()
        self._ws_bytes = lib.vdp_cvi_workspace(B, N, D, n_obs)
        self._ws = torch.empty(self._ws_bytes, dtype=torch.uint8, device=dev)
        # prior p: initial state from the prior on x0 (variational_cvi_sde.py:408-412, sde_utils.py:145-147)
        self.set_linearized_prior()

    # ---- girsanov sites: plain tensors that are brought up to date on access when the last updates kept them in theta
    def _sync_sites(self):
        if self._sites_stale:
            self._sites_stale = False
            p = self._problem(0.0, 0.0, keep_carry=True)
            self._run("vdp_cvi_materialize_sites", C.byref(p))

    def _get_site(self, name):
        self._sync_sites()
        return getattr(self, name)

    def _set_site(self, name, value):
        self._sync_sites()
        setattr(self, name, value)
        self._nat_sig = None

    g1 = property(lambda s: s._get_site("_g1"), lambda s, v: s._set_site("_g1", v))
    gd = property(lambda s: s._get_site("_gd"), lambda s, v: s._set_site("_gd", v))
    gs = property(lambda s: s._get_site("_gs"), lambda s, v: s._set_site("_gs", v))

    # variational_cvi_sde.py:408-432
    def set_prior_initial_state(self, mu, cov):
        """Assign the prior on x0 (cvi_dp_trainer.py:231-235); dist_p picks it up at the next linearisation."""
        self.x0_mu = np.asarray(mu, dtype=np.float64).reshape(self.D)
        self.x0_cov = np.atleast_2d(np.asarray(cov, dtype=np.float64))
        self._graphs = {}             # captured graphs hold the old x0 prior by value

    def set_linearized_prior(self):
        dev, B, D = self.device, self.B, self.D
        self._sync_sites()            # implicit sites are relative to the prior that is about to be replaced
        self._graphs = {}             # the caching allocator may hand the new prior the old addresses
        self.pmu0 = torch.as_tensor(self.x0_mu, dtype=F64).to(dev).expand(B, D).contiguous()
        self.pcholP0 = torch.as_tensor(np.linalg.cholesky(self.x0_cov), dtype=F64).to(dev).expand(B, D, D).contiguous()
        m = self.fx_mus[:, 1:].contiguous()
        S = self.fx_covs[:, 1:].contiguous()
        clip = self.clip if self.stabilize_ssm else None
        A, b, cholQ, A_lin, b_lin = ops.linearize_to_ssm(self.drift_id, self.drift_params, self.q, self.time_grid, m,
                                                         S, clip=clip, status=self.status, want_unclipped=True)
        self.pA, self.pb, self.pcholQ = A, b, cholQ
        self.pA_lin, self.pb_lin = A_lin, b_lin

    _RAW = {"g1": "_g1", "gd": "_gd", "gs": "_gs"}

    def _problem(self, lr_data, lr_girsanov, flags=0, keep_carry=False):
        if not keep_carry:
            self._sync_sites()        # every entry but the lazy site update / the materialisation reads g1, gd, gs
        p = CviProblem()
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
        p.lr_data, p.lr_girsanov = float(lr_data), float(lr_girsanov)
        for name in ("time_grid", "pA", "pb", "pcholQ", "pmu0", "pcholP0", "g1", "gd", "gs", "d1", "d2", "obs_idx",
                     "obs_y", "fx_mus", "fx_covs", "qA", "qb", "qcholQ", "qmu0", "qcholP0", "elbo", "ve", "kl",
                     "status"):
            t = getattr(self, self._RAW.get(name, name))
            assert t.is_contiguous(), name
            setattr(p, name, t.data_ptr())
        p.workspace = self._ws.data_ptr()
        p.workspace_bytes = self._ws_bytes
        p.flags = flags | self.route_flags
        if not keep_carry:
            self._nat_sig = None      # whatever is called with this problem may rewrite the workspace
        return p

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _run(self, name, *args):
        """Call a library entry on this engine's device (the launches go to the CURRENT device) and stream."""
        with torch.cuda.device(self.device):
            check(getattr(_lib.load(), name)(*args, self._stream()), name)

    # what the natural parameters left in the workspace by a site update depend on (besides the update itself)
    _NAT_FIELDS = ("pA", "pb", "pcholQ", "pmu0", "pcholP0", "_g1", "_gd", "_gs", "d1", "d2", "obs_idx", "_ws")

    def _nat_signature(self):
        return tuple((t.data_ptr(), t._version) for t in (getattr(self, n) for n in self._NAT_FIELDS))

    def site_update(self, lr_data, lr_girsanov, flags=0, carry=None):
        """
        cvi_dp_trainer.py:71-77 once.  Returns the ELBO tensor [B] (device).

        Between two consecutive updates only the data sites change before the first posterior rebuild, so the
        naturals the previous update left in the workspace are patched at the observed grid points instead of being
        recomputed over the whole grid (``VDP_FLAG_NATURALS_CARRIED``).  The promise behind that flag is checked
        here: no other library call on this engine in between, and none of the tensors the naturals are made of
        replaced or modified in place through torch (storage pointer + version counter).
        """
        carry = self.carry_naturals if carry is None else carry
        flags |= self.route_flags
        lazy = False
        if carry and self._nat_sig is not None and self._nat_sig == self._nat_signature():
            flags |= _lib.FLAG_NATURALS_CARRIED
            # the library honours the lazy sites under these conditions (vdp_cvi.inl run_site_update)
            lazy = (self.lazy_sites and self.n_obs <= 128 and not (flags & _lib.FLAG_SKIP_GIRSANOV)
                    and not _cfg_on("NO_CARRIED_NATURALS") and not _cfg_on("NO_FUSED_NATURALS")
                    and not (flags & _lib.FLAG_NO_FUSED_NATURALS))
            if lazy:
                flags |= _lib.FLAG_SITES_IN_NATURALS
        p = self._problem(lr_data, lr_girsanov, flags, keep_carry=lazy)
        self._run("vdp_cvi_site_update", C.byref(p))
        if lazy:
            self._sites_stale = True
        if carry and (not (flags & _lib.FLAG_SKIP_GIRSANOV) or (flags & _lib.FLAG_NATURALS_CARRIED)):
            self._nat_sig = self._nat_signature()      # (a data-sites-only update leaves no dense data-site copy)
        else:
            self._nat_sig = None
        return self.elbo

    def optimize_sites(self, max_itr, optim_tol, lr_data, lr_girsanov, test_idx=None):
        """
        ``CVISitesTrainer._optimize_sites_under_stable_prior`` (cvi_dp_trainer.py:63-101) on the device
        (``vdp_cvi_optimize_sites``): up to ``max_itr`` site updates with the ELBO-drop => step sizes / 10 rule and the
        convergence test evaluated per sequence by a kernel, the loop a CUDA-graph WHILE node -- no host
        synchronisation between updates.  ``lr_data`` / ``lr_girsanov``: scalars or [B]; ``test_idx`` [B, n_test] grid
        indices whose marginals are recorded after every update (for NLPD / RMSE).

        Returns a dict of CPU tensors: ``n_itr`` [B], ``elbo_trace`` [max_itr + 1, B] (row 0 = ELBO before the loop),
        ``lr_data`` / ``lr_girsanov`` [B] (decayed), ``test_mus`` [max_itr, B, n_test, D], ``test_covs``.
        """
        B, D, dev = self.B, self.D, self.device
        mk = lambda v: torch.as_tensor(np.broadcast_to(np.asarray(v, dtype=np.float64), (B,)).copy()).to(dev)
        lrd, lrg = mk(lr_data), mk(lr_girsanov)
        trace = torch.zeros((max_itr + 1, B), dtype=F64, device=dev)
        trace[0] = self.classic_elbo_keep_marginals()      # tmp_elbo_vals = [classic_elbo()]   (:67)
        active = torch.zeros(B, dtype=torch.int32, device=dev)
        n_itr = torch.zeros(B, dtype=torch.int32, device=dev)
        counters = torch.zeros(2, dtype=torch.int32, device=dev)
        lp = _lib.CviLoop()
        lp.max_itr, lp.optim_tol = int(max_itr), float(optim_tol)
        n_test = 0
        tm = tc = ti = None
        if test_idx is not None:
            ti = torch.as_tensor(np.asarray(test_idx, dtype=np.int64).reshape(B, -1)).to(dev).contiguous()
            n_test = ti.shape[1]
            tm = torch.zeros((max_itr, B, n_test, D), dtype=F64, device=dev)
            tc = torch.zeros((max_itr, B, n_test, D, D), dtype=F64, device=dev)
            lp.test_idx, lp.test_mus, lp.test_covs = ti.data_ptr(), tm.data_ptr(), tc.data_ptr()
        lp.n_test = n_test
        lp.lr_data, lp.lr_girsanov = lrd.data_ptr(), lrg.data_ptr()
        lp.active, lp.n_itr, lp.elbo_trace, lp.counters = (active.data_ptr(), n_itr.data_ptr(), trace.data_ptr(),
                                                           counters.data_ptr())
        lazy = (self.carry_naturals and self.lazy_sites and self.n_obs <= 128 and max_itr >= 2
                and not _cfg_on("NO_CARRIED_NATURALS") and not _cfg_on("NO_FUSED_NATURALS")
                and not (self.route_flags & _lib.FLAG_NO_FUSED_NATURALS))
        later = self.route_flags
        if self.carry_naturals:
            later |= _lib.FLAG_NATURALS_CARRIED | (_lib.FLAG_SITES_IN_NATURALS if lazy else 0)
        lp.later_flags = later
        p = self._problem(0.0, 0.0, self.route_flags)      # first update: naturals rebuilt from the sites
        self._run("vdp_cvi_optimize_sites", C.byref(p), C.byref(lp))
        self._sites_stale = lazy
        self._nat_sig = self._nat_signature() if self.carry_naturals else None
        out = {"n_itr": n_itr.cpu(), "elbo_trace": trace.cpu(), "lr_data": lrd.cpu(), "lr_girsanov": lrg.cpu()}
        if tm is not None:
            out["test_mus"], out["test_covs"] = tm.cpu(), tc.cpu()
        return out

    _PTR_FIELDS = ("time_grid", "pA", "pb", "pcholQ", "pmu0", "pcholP0", "_g1", "_gd", "_gs", "d1", "d2", "obs_idx", "obs_y",
                   "fx_mus", "fx_covs", "qA", "qb", "qcholQ", "qmu0", "qcholP0", "elbo", "ve", "kl", "status")

    def site_update_graphed(self, lr_data, lr_girsanov, flags=0):
        """
        The same update replayed from a CUDA graph: for one or a few short sequences (the reference's own configs:
        N = 500..4000, B = 1) an update is ~20 launches of a few microseconds each, i.e. launch-bound; the graph
        removes the per-launch host cost.  The graph is captured on first use and re-captured whenever a buffer of
        the problem has been replaced (re-linearisation, site assignment) or the step sizes / drift parameters change.
        """
        # everything a captured launch sequence holds by value or by address
        key = (float(lr_data), float(lr_girsanov), int(flags | self.route_flags), tuple(self.drift_params),
               tuple(_lib.config_get(k) for k in _lib.CONFIG_KEYS), tuple(self.x0_mu.tolist()),
               tuple(self.x0_cov.reshape(-1).tolist()), tuple(self.q.reshape(-1).tolist()),
               tuple(self.lik_chol.reshape(-1).tolist()), self.dt, self.clip, self.stabilize_ssm,
               tuple(getattr(self, n).data_ptr() for n in self._PTR_FIELDS))
        cache = getattr(self, "_graphs", None)
        if cache is None:
            cache = self._graphs = {}
        g = cache.get(key)
        if g is None:
            if len(cache) >= 8:
                cache.clear()
            state = [getattr(self, n).clone() for n in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs")]
            self.site_update(lr_data, lr_girsanov, flags, carry=False)   # warm-up outside the capture,
            g = torch.cuda.CUDAGraph()
            with torch.cuda.device(self.device), torch.cuda.graph(g):
                self.site_update(lr_data, lr_girsanov, flags, carry=False)
            for n, t in zip(("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"), state):
                getattr(self, n).copy_(t)                          # ... whose effect is undone: the replay below is
            cache[key] = g                                         # the one update of this call
        g.replay()
        self._nat_sig = None
        return self.elbo

    def site_update_sharded(self, phase, node_begin, node_end, lr_data=0.0, lr_girsanov=0.0):
        """One phase of ``vdp_cvi_site_update_sharded`` on the flat node range [node_begin, node_end)."""
        p = self._problem(lr_data, lr_girsanov)
        self._run("vdp_cvi_site_update_sharded", C.byref(p), int(phase), int(node_begin), int(node_end))

    def classic_elbo(self):
        p = self._problem(0.0, 0.0)
        self._run("vdp_cvi_elbo", C.byref(p))
        return self.elbo

    def posterior_ssm(self):
        """dist_q (variational_cvi_sde.py:176-192): rebuild q as an SSM from the sites; marginals untouched."""
        p = self._problem(0.0, 0.0, flags=_lib.FLAG_POSTERIOR_SSM_ONLY)
        self._run("vdp_cvi_elbo", C.byref(p))
        return self.qA, self.qb, self.qcholQ, self.qmu0, self.qcholP0

    # ---- prior-parameter learning (cvi_dp_trainer.py:207-250)
    def _scratch_marginals(self):
        """Buffers for dist_q.marginals that must not overwrite the stored fx_mus / fx_covs (the reference keeps
        self.fx_mus as of the last site update while classic_elbo / grad_* rebuild dist_q, :339-352, :495-518)."""
        if getattr(self, "_tmp_mus", None) is None:
            self._tmp_mus, self._tmp_covs = torch.empty_like(self.fx_mus), torch.empty_like(self.fx_covs)
        return self._tmp_mus, self._tmp_covs

    def _problem_with_scratch_marginals(self):
        p = self._problem(0.0, 0.0)
        tm, tS = self._scratch_marginals()
        p.fx_mus, p.fx_covs = tm.data_ptr(), tS.data_ptr()
        return p

    def classic_elbo_keep_marginals(self):
        """classic_elbo (variational_cvi_sde.py:339-352) with the stored fx_mus / fx_covs left as they are."""
        p = self._problem_with_scratch_marginals()
        self._run("vdp_cvi_elbo", C.byref(p))
        return self.elbo

    def n_drift_params(self):
        return ops.drift_num_params(self.drift_id)

    def grad_kl_prior_params(self):
        """CVISitesSDE.grad_KL_wrt_prior_params (variational_cvi_sde.py:495-504) -> [B, n_params]."""
        p = self._problem_with_scratch_marginals()
        g = torch.empty((self.B, 16), dtype=F64, device=self.device)
        self._run("vdp_cvi_grad_kl_prior_params", C.byref(p), C.c_void_p(g.data_ptr()))
        return g[:, :self.n_drift_params()]

    def grad_ve_prior_params(self):
        """CVISitesSDE.grad_VE_wrt_prior_params (variational_cvi_sde.py:506-518) -> [B, n_params] = d(-VE)/dp.
        Re-linearises the prior around the stored posterior path as a side effect, as the reference does (:512)."""
        self.set_linearized_prior()
        p = self._problem_with_scratch_marginals()
        g = torch.empty((self.B, 16), dtype=F64, device=self.device)
        clip = 1 if self.stabilize_ssm else 0
        self._run("vdp_cvi_grad_ve_prior_params", C.byref(p), C.c_void_p(self.fx_mus.data_ptr()),
                  C.c_void_p(self.fx_covs.data_ptr()), clip, self.clip[0], self.clip[1], C.c_void_p(g.data_ptr()))
        return g[:, :self.n_drift_params()]

    def check_status(self):
        code, idx = ops.read_status(self.status)
        if code != 0:
            raise _lib.VdpError(f"non positive-definite block at flat grid index {idx} (code {code})")

    # sde_utils.py:550-568 applied as in cvi_dp_trainer.py:110-134
    def _prior_naturals(self, A, b):
        return ops.ssm_to_naturals(A, b, self.pcholQ, self.pmu0, self.pcholP0)

    def _rebase_sites(self, cur, new):
        self.g1 = (self.g1 + cur[0] - new[0]).contiguous()
        self.gd = (self.gd + cur[1] - new[1]).contiguous()
        self.gs = (self.gs + cur[2] - new[2]).contiguous()

    def to_linearized_prior(self):
        """cvi_dp_trainer.py:110-113: re-base the sites from the (clipped) prior onto p_lin and adopt p_lin."""
        self._rebase_sites(self._prior_naturals(self.pA, self.pb), self._prior_naturals(self.pA_lin, self.pb_lin))
        self.pA, self.pb = self.pA_lin, self.pb_lin

    def relinearize_new(self):
        """cvi_dp_trainer.py:128-134: linearise around the current posterior and re-base the sites onto it."""
        last = self._prior_naturals(self.pA, self.pb)
        self.set_linearized_prior()
        self._rebase_sites(last, self._prior_naturals(self.pA, self.pb))

    def relinearize(self):
        self.to_linearized_prior()
        self.relinearize_new()


class CviHostEngine:
    """
    The same site update with the whole state in pinned HOST memory (``vdp_cvi_host_site_update``): what a host
    program holding NumPy arrays calls.  Every call uploads the state (and the prior unless ``resident_prior``),
    runs the update and downloads sites, marginals, posterior SSM and ELBO, chunked over sequences so that
    copies and kernels overlap.  ``host`` maps the field names of ``vdp_cvi_problem`` to pinned CPU tensors.
    """

    STATE = ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs")
    PRIOR = ("pA", "pb", "pcholQ", "pmu0", "pcholP0")
    CONST = ("time_grid", "obs_idx", "obs_y")
    OUT = ("qA", "qb", "qcholQ", "qmu0", "qcholP0", "elbo", "ve", "kl")

    def __init__(self, eng: CviEngine, chunk=256, slots=3, resident_prior=False, want_posterior_ssm=True):
        self._eng = eng
        self.B, self.N, self.D, self.n_obs = eng.B, eng.N, eng.D, eng.n_obs
        self.resident_prior = resident_prior
        names = self.STATE + self.PRIOR + self.CONST + self.OUT
        self.host = {k: getattr(eng, k).detach().to("cpu").contiguous().pin_memory() for k in names}
        self.skip = set() if want_posterior_ssm else {"qA", "qb", "qcholQ", "qmu0", "qcholP0"}
        self._ctx = C.c_void_p()
        dev = eng.device.index if eng.device.index is not None else torch.cuda.current_device()
        check(_lib.load().vdp_cvi_host_create(C.byref(self._ctx), min(chunk, self.B), self.N, self.D, self.n_obs,
                                              slots, dev), "vdp_cvi_host_create")
        if resident_prior:
            self.set_prior()

    def set_prior(self):
        h = self.host
        check(_lib.load().vdp_cvi_host_set_prior(self._ctx, self.B, *(h[k].data_ptr() for k in self.PRIOR)),
              "vdp_cvi_host_set_prior")

    def bytes_per_update(self):
        """(host->device, device->host) bytes one site_update moves."""
        h = self.host
        nb = lambda names: sum(h[k].numel() * h[k].element_size() for k in names if k not in self.skip)
        mu_obs = self.B * self.n_obs * self.D * 8          # of fx_mus only the observed grid points are uploaded
        up = (nb([k for k in self.STATE if k not in ("fx_covs", "fx_mus")]) + mu_obs + nb(self.CONST) +
              (0 if self.resident_prior else nb(self.PRIOR)))
        return up, nb(self.STATE) + nb(self.OUT)

    def site_update(self, lr_data, lr_girsanov, flags=0):
        p = self._eng._problem(lr_data, lr_girsanov, flags)
        for k, t in self.host.items():
            setattr(p, k, None if k in self.skip or (self.resident_prior and k in self.PRIOR) else t.data_ptr())
        p.workspace, p.workspace_bytes, p.status = None, 0, None
        st = (C.c_int64 * 2)()
        check(_lib.load().vdp_cvi_host_site_update(self._ctx, C.byref(p), st), "vdp_cvi_host_site_update")
        if st[0] != 0:
            raise _lib.VdpError(f"non positive-definite block (code {st[0]}, chunk-local grid index {st[1]})")
        return self.host["elbo"]

    def close(self):
        if self._ctx:
            _lib.load().vdp_cvi_host_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CviSession:
    """
    Resident-state session of the host-buffer interface (``vdp_cvi_session_*``): the state is uploaded once from
    pinned host tensors and stays on the device; every :meth:`site_update` takes that update's observation values from
    host memory and writes the ELBO terms and the posterior marginals back to host memory.  ``host`` maps the field
    names of ``vdp_cvi_problem`` to the pinned CPU tensors (``host["fx_mus"]`` etc. receive the results).
    """

    FIELDS = CviHostEngine.STATE + CviHostEngine.PRIOR + CviHostEngine.CONST + CviHostEngine.OUT

    def __init__(self, eng: CviEngine, chunk=1024, streams=3):
        self._eng = eng
        self.B, self.N, self.D, self.n_obs = eng.B, eng.N, eng.D, eng.n_obs
        self.host = {k: getattr(eng, k).detach().to("cpu").contiguous().pin_memory() for k in self.FIELDS}
        p = self._host_problem()
        self._s = C.c_void_p()
        dev = eng.device.index if eng.device.index is not None else torch.cuda.current_device()
        check(_lib.load().vdp_cvi_session_create(C.byref(self._s), C.byref(p), int(chunk), int(streams), dev),
              "vdp_cvi_session_create")

    def _host_problem(self):
        p = self._eng._problem(0.0, 0.0)
        for k, t in self.host.items():
            setattr(p, k, t.data_ptr())
        p.workspace, p.workspace_bytes, p.status = None, 0, None
        return p

    def bytes_per_update(self, marginals=True):
        """(host->device, device->host) bytes one site_update moves: observation values up; ELBO terms and (with
        ``marginals``) the posterior marginals down."""
        h = self.host
        nb = lambda names: sum(h[k].numel() * h[k].element_size() for k in names)
        return nb(["obs_y"]), nb(["elbo", "ve", "kl"] + (["fx_mus", "fx_covs"] if marginals else []))

    def site_update(self, lr_data, lr_girsanov, flags=0, obs_y=None, marginals=True):
        """``marginals=False``: only the ELBO terms come down -- what the reference's trainer reads every iteration
        (cvi_dp_trainer.py:75, the marginals stay with the model); ``download_state`` fetches them when wanted."""
        h = self.host
        y = h["obs_y"] if obs_y is None else obs_y
        st = (C.c_int64 * 2)()
        check(_lib.load().vdp_cvi_session_site_update(
            self._s, C.c_void_p(y.data_ptr()), float(lr_data), float(lr_girsanov), int(flags),
            C.c_void_p(h["elbo"].data_ptr()), C.c_void_p(h["ve"].data_ptr()), C.c_void_p(h["kl"].data_ptr()),
            C.c_void_p(h["fx_mus"].data_ptr()) if marginals else None,
            C.c_void_p(h["fx_covs"].data_ptr()) if marginals else None, st), "vdp_cvi_session_site_update")
        if st[0] != 0:
            raise _lib.VdpError(f"non positive-definite block (code {st[0]}, grid index {st[1]})")
        return h["elbo"]

    def download_state(self):
        p = self._host_problem()
        check(_lib.load().vdp_cvi_session_download_state(self._s, C.byref(p)), "vdp_cvi_session_download_state")
        return self.host

    def close(self):
        if self._s:
            _lib.load().vdp_cvi_session_destroy(self._s)
            self._s = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
