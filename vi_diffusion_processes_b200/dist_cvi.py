"""
CVI-DP site update of ONE long sequence with the chain sharded in TIME over ranks (BASELINE config 5: 10^7 grid steps,
D = 3, "parallel-in-time scan sharded over 2/4/8 GPUs ... with a small NCCL-over-NVLink exchange of scan carries").

No rank holds the whole chain.  Rank r owns the grid points [n0, n1) and the transitions [s, e) of
``time_partition`` and keeps, for every array of the model, its own stretch extended by ONE halo grid point at either
end (the "extended local chain"); on it the pointwise stages of the update run through the same C entry points a single
GPU uses, restricted to the owned node range:

    update_data_sites     vdp_cvi_update_data_sites             variational_cvi_sde.py:301-314   (own observations)
    full_sites            vdp_cvi_full_sites                    :161-174
    dist_q + marginals    TimeShardedChain (vdp_dist_sweep_*)   :176-192, gauss_markov.py:107-117   <- scan carries
    update_girsanov_sites vdp_cvi_update_girsanov_sites_range   :279-296  (H = 20 quadrature, own grid points)
    classic_elbo          vdp_cvi_elbo_terms_range              :339-352  (partial sums, added in rank order)

What crosses ranks per update: the scan elements / initial states of the sweeps (two rebuilds x 5 all-gathers of a
few hundred bytes), one halo blob per rebuild (marginal mean + covariance of a rank's first and last grid point and
its last transition matrix: (2 D + 3 D^2) doubles) and 2 doubles of ELBO terms.  All of it goes through
``Comm.all_gather`` -- ``vdp_comm_all_gather`` (NCCL issued from C on the compute stream) in a multi-process run,
plain lists when the ranks are emulated in one process (tests).  The linearised prior needs no exchange: transition
k is linearised around grid point k + 1 alone (variational_cvi_sde.py:412), which the extended chain holds.

The state after an update equals the single-GPU state up to the rounding of the scan carries (1e-12 relative,
tests/test_gpu_dist_cvi.py); the ELBO terms are added in rank order, so a run is reproducible bit for bit.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import numpy as np
import torch

from . import _lib
from ._lib import check
from .cvi import CviEngine
from .dist_time import Comm, TimeShardedChain, time_partition

F64 = torch.float64


class _RankState:
    def __init__(self, rank, nranks, N, n0, n1, s, e):
        self.rank, self.nranks = rank, nranks
        self.first, self.last = rank == 0, rank == nranks - 1
        self.n0, self.n1, self.s, self.e = n0, n1, s, e
        self.x0, self.x1 = max(n0 - 1, 0), min(n1 + 1, N + 1)        # extended chain: grid points [x0, x1)
        self.lo = n0 - self.x0                                        # owned range in extended-local indices
        self.hi = self.lo + (n1 - n0)
        self.nt = e - s                                               # owned transitions, local index lo .. lo + nt
        self.eng: CviEngine = None


class TimeShardedCvi:
    """
    ``drift_id, drift_params, q, lik_chol, x0_mu, x0_cov`` as :class:`CviEngine`; ``time_grid`` [N+1],
    ``obs_idx`` [n_obs] (grid indices, sorted or not) and ``obs_y`` [n_obs, D] describe the WHOLE sequence: every rank
    slices its own part (synthetic inputs are cheap to regenerate per rank; a loader would read the rank's stretch).
    """

    def __init__(self, drift_id, drift_params, q, time_grid, obs_idx, obs_y, lik_chol, x0_mu, x0_cov, comm: Comm,
                 device="cuda", stabilize_ssm=True, clip_state_transitions=(-1.0, 1.0), refactor_precision=False,
                 carry_naturals=True, info_rebuild=True):
        tg = np.asarray(time_grid, dtype=np.float64)
        self.N = N = tg.shape[0] - 1
        self.D = D = np.atleast_2d(np.asarray(q)).shape[0]
        self.comm = comm
        self.refactor_precision = bool(refactor_precision)
        self.carry_naturals, self._th_sig = bool(carry_naturals), None
        self.info_rebuild = bool(info_rebuild) and not self.refactor_precision   # (the reference's route wins)
        self.device = torch.device(device)
        self.chain = TimeShardedChain(N, D, comm, device=device)
        obs_idx = np.asarray(obs_idx, dtype=np.int64).reshape(-1)
        obs_y = np.asarray(obs_y, dtype=np.float64).reshape(obs_idx.shape[0], D)
        self.ranks: List[_RankState] = []
        for r in comm.local_ranks:
            s, e = self.chain.parts[r]
            n0, n1 = self.chain.owned_nodes(r)
            st = _RankState(r, comm.nranks, N, n0, n1, s, e)
            mine = (obs_idx >= n0) & (obs_idx < n1)
            st.eng = CviEngine(drift_id, drift_params, q, tg[st.x0:st.x1], (obs_idx[mine] - st.x0)[None],
                               obs_y[mine][None], lik_chol, x0_mu, x0_cov, device=device, stabilize_ssm=stabilize_ssm,
                               clip_state_transitions=clip_state_transitions, carry_naturals=False)
            n_ext = st.x1 - st.x0
            z = lambda *shape: torch.zeros(shape, dtype=F64, device=self.device)
            st.th1, st.thd, st.ths = z(1, n_ext, D), z(1, n_ext, D, D), z(1, n_ext - 1, D, D)
            self.ranks.append(st)
        self.elbo = None

    # ---- C calls on one rank's extended chain
    @staticmethod
    def _call(st: _RankState, name, lr_data=0.0, lr_girsanov=0.0, *args):
        p = st.eng._problem(lr_data, lr_girsanov)
        st.eng._run(name, C.byref(p), *args)

    def _own(self, st: _RankState):
        """views of the owned part of theta on the extended chain"""
        return (st.th1[0, st.lo:st.hi], st.thd[0, st.lo:st.hi], st.ths[0, st.lo:st.lo + st.nt])

    _TH_FIELDS = ("pA", "pb", "pcholQ", "pmu0", "pcholP0", "_g1", "_gd", "_gs", "d1", "d2", "obs_idx")

    def _th_signature(self):
        """what the full naturals held in ``st.th1 / thd / ths`` were computed from (torch-side writes bump _version;
        the C calls of this class that write them are followed by a rebuild or reset ``_th_sig``).  Re-linearise through
        :meth:`relinearize`, not through the engines directly: a replaced prior tensor can land on the address of the
        old one with the same version count, which this signature cannot tell apart."""
        return tuple((t.data_ptr(), t._version) for st in self.ranks
                     for t in [getattr(st.eng, n) for n in self._TH_FIELDS] + [st.th1, st.thd, st.ths])

    def _rebuild(self, data_sites_before=None):
        """dist_q + marginals on the time-sharded chain, results into the engines' extended arrays + halos.
        ``data_sites_before``: per local rank the (d1, d2) the naturals in ``st.th*`` were built with, when nothing
        but the data sites has changed since -- theta is then corrected at the observed grid points only instead of
        being rebuilt from the prior at every grid point (0.5 GB read + written per 10^6 steps in D = 3)."""
        D = self.D
        for i, st in enumerate(self.ranks):
            if data_sites_before is not None:
                e = st.eng
                if e.obs_idx.numel():
                    old1, old2 = data_sites_before[i]
                    st.th1[0].index_add_(0, e.obs_idx[0], (e.d1 - old1)[0])
                    st.thd[0].index_add_(0, e.obs_idx[0], (e.d2 - old2)[0])
                continue
            self._call(st, "vdp_cvi_full_sites", 0.0, 0.0, C.c_void_p(st.th1.data_ptr()), C.c_void_p(st.thd.data_ptr()),
                       C.c_void_p(st.ths.data_ptr()))
        own = [self._own(st) for st in self.ranks]
        if self.info_rebuild:
            # two sweeps: information form from the bottom (transitions straight from the naturals, no factor, no
            # inverse band), moments from the top; straight into the engines' arrays
            outs, mouts = [], []
            for st in self.ranks:
                e, lo, nt = st.eng, st.lo, st.nt
                outs.append(dict(A=e.qA[0, lo:lo + nt], b=e.qb[0, lo:lo + nt], cholQ=e.qcholQ[0, lo:lo + nt]))
                mouts.append((e.fx_mus[0, lo:lo + nt + 1], e.fx_covs[0, lo:lo + nt + 1]))
            ssm = self.chain.naturals_to_ssm_info([o[0] for o in own], [o[1] for o in own], [o[2] for o in own],
                                                  out=outs)
            marg = self.chain.ssm_moments([o["A"] for o in ssm], [o["b"] for o in ssm], [o["cholQ"] for o in ssm],
                                          [o["mu0"] for o in ssm], [o["cholP0"] for o in ssm], out=mouts)
        elif self.refactor_precision:
            # the reference's route (gauss_markov.py:107-117): precision rebuilt from the SSM, factorised again,
            # diagonal of its inverse band
            ssm = self.chain.naturals_to_ssm([o[0] for o in own], [o[1] for o in own], [o[2] for o in own])
            marg = self.chain.ssm_marginals([o["A"] for o in ssm], [o["b"] for o in ssm], [o["cholQ"] for o in ssm],
                                            [o["mu0"] for o in ssm], [o["cholP0"] for o in ssm])
        else:
            # the same matrix is the one naturals_to_ssm started from: its backward sweep writes the marginal
            # covariances on the way, a mean-only forward sweep follows (as the single-GPU rebuild does).  The sweeps
            # write straight into the engines' extended arrays (no staging copies: 0.7 GB each at 10^7 steps)
            outs, mouts = [], []
            for st in self.ranks:
                e, lo, nt = st.eng, st.lo, st.nt
                o = dict(A=e.qA[0, lo:lo + nt], b=e.qb[0, lo:lo + nt], cholQ=e.qcholQ[0, lo:lo + nt],
                         S=e.fx_covs[0, lo:st.hi])
                if not st.first:
                    o["b_prev"], o["cholQ_prev"] = e.qb[0, lo - 1:lo], e.qcholQ[0, lo - 1:lo]
                if not st.last:
                    o["thd_ext"] = st.thd[0, lo:st.hi + 1]
                outs.append(o)
                mouts.append(e.fx_mus[0, lo:lo + nt + 1])
            ssm = self.chain.naturals_to_ssm([o[0] for o in own], [o[1] for o in own], [o[2] for o in own],
                                             want_cov=True, out=outs)
            means = self.chain.ssm_means([o["A"] for o in ssm], [o["b"] for o in ssm], [o["mu0"] for o in ssm],
                                         out=mouts)
            marg = [(m, o["S"]) for m, o in zip(means, ssm)]
        blobs = []
        for st, o, (m, S) in zip(self.ranks, ssm, marg):
            e = st.eng
            if self.refactor_precision:
                e.qA[0, st.lo:st.lo + st.nt] = o["A"]
                e.qb[0, st.lo:st.lo + st.nt] = o["b"]
                e.qcholQ[0, st.lo:st.lo + st.nt] = o["cholQ"]
                e.fx_mus[0, st.lo:st.hi] = m
                e.fx_covs[0, st.lo:st.hi] = S
            if st.first:
                e.qmu0[0] = o["mu0"]
                e.qcholP0[0] = o["cholP0"]
            blobs.append(torch.cat([m[0], S[0].reshape(-1), m[-1], S[-1].reshape(-1), o["A"][-1].reshape(-1)]))
        allb = self.comm.all_gather(blobs)
        DD = D * D
        for st in self.ranks:
            e = st.eng
            if not st.first:      # front halo: last grid point and last transition of the rank before
                bl = allb[st.rank - 1]
                e.fx_mus[0, 0] = bl[D + DD:2 * D + DD]
                e.fx_covs[0, 0] = bl[2 * D + DD:2 * D + 2 * DD].reshape(D, D)
                e.qA[0, 0] = bl[2 * D + 2 * DD:].reshape(D, D)
            if not st.last:       # back halo: first grid point of the rank after
                bl = allb[st.rank + 1]
                e.fx_mus[0, -1] = bl[:D]
                e.fx_covs[0, -1] = bl[D:D + DD].reshape(D, D)

    def _site_update_device(self, lr_data, lr_girsanov):
        """one update as a stream-ordered sequence of launches and exchanges (no host read): capturable"""
        carried = self.carry_naturals and self._th_sig is not None and self._th_sig == self._th_signature()
        before = [(st.eng.d1.clone(), st.eng.d2.clone()) for st in self.ranks] if carried else None
        self._th_sig = None
        for st in self.ranks:
            self._call(st, "vdp_cvi_update_data_sites", lr_data, lr_girsanov)
        self._rebuild(before)
        for st in self.ranks:
            self._call(st, "vdp_cvi_update_girsanov_sites_range", lr_data, lr_girsanov, st.lo, st.hi)
        self._rebuild()
        self._th_sig = self._th_signature()       # theta of (prior, both kinds of sites) as they stand now
        self._elbo_device()

    def site_update(self, lr_data, lr_girsanov):
        """cvi_dp_trainer.py:71-77 once; returns the ELBO (float), the same on every rank."""
        self._site_update_device(lr_data, lr_girsanov)
        return float(self.elbo.item())

    def _graph_key(self, lr_data, lr_girsanov):
        """everything the captured launch sequence holds by value or by address"""
        ptrs = []
        for st in self.ranks:
            e = st.eng
            ptrs += [t.data_ptr() for t in (st.th1, st.thd, st.ths)]
            ptrs += [getattr(e, n).data_ptr() for n in e._PTR_FIELDS]
            # ... and by value (as CviEngine.site_update_graphed keys them)
            ptrs += [int(e.route_flags), e.dt, e.clip, e.stabilize_ssm] + [float(x) for x in e.drift_params]
            ptrs += e.x0_mu.tolist() + e.x0_cov.reshape(-1).tolist() + e.q.reshape(-1).tolist() \
                + e.lik_chol.reshape(-1).tolist()
        return (float(lr_data), float(lr_girsanov), self.refactor_precision, self.info_rebuild,
                tuple(_lib.config_get(k) for k in _lib.CONFIG_KEYS), tuple(ptrs))

    def site_update_graphed(self, lr_data, lr_girsanov):
        """
        :meth:`site_update` replayed from a CUDA graph: the update is ~90 launches and 15 exchanges with no host
        read in between, i.e. one capturable sequence.

        The first call with a given (step sizes, buffers, configuration) runs the update eagerly -- which also sizes
        every workspace -- and then captures the sequence (a capture executes nothing); later calls replay it.

        Measured (B200, D = 3): the replay buys 2 % on the 1.25e6-step stretch a rank of 8 owns (3.21 -> 3.14 ms:
        the update is not host-bound there, its fixed part is short kernels), and costs 9 % on the whole 10^7-step
        chain.  With ranks in separate processes the one attempt to capture the ``ncclAllGather`` calls hung at
        replay (2 GPUs), so that combination is refused here; the eager sequence is what bench.py times.
        """
        if self.comm.distributed:
            raise NotImplementedError("graph replay of the time-sharded update across processes is not enabled "
                                      "(see the docstring); use site_update")
        key = self._graph_key(lr_data, lr_girsanov)
        cache = getattr(self, "_graphs", None)
        if cache is None:
            cache = self._graphs = {}
        # the captured sequence is the one of a carried update (theta corrected at the observations only)
        fresh = self._th_sig is not None and self._th_sig == self._th_signature()
        ent = cache.get(key) if (fresh or not self.carry_naturals) else None
        if ent is None:
            out = self.site_update(lr_data, lr_girsanov)
            cache.clear()                      # a graph keeps its temporaries: hold one
            g = torch.cuda.CUDAGraph()
            torch.cuda.synchronize(self.device)
            with torch.cuda.device(self.device), torch.cuda.graph(g, capture_error_mode="thread_local"):
                self._site_update_device(lr_data, lr_girsanov)
            cache[key] = (g, self.elbo, self.ve, self.kl)
            return out
        g, self.elbo, self.ve, self.kl = ent
        g.replay()
        return float(self.elbo.item())

    def _elbo_device(self):
        parts = []
        for st in self.ranks:
            self._call(st, "vdp_cvi_elbo_terms_range", 0.0, 0.0, st.lo, st.hi)
            parts.append(torch.cat([st.eng.ve, st.eng.kl]))
        allp = self.comm.all_gather(parts)
        tot = allp[0].clone()
        for t in allp[1:]:                 # fixed rank order: reproducible bit for bit
            tot = tot + t.to(tot.device)
        self.ve, self.kl = tot[0], tot[1]
        self.elbo = tot[0] - tot[1]

    def classic_elbo(self, rebuilt=False):
        if not rebuilt:
            self._rebuild()
        self._elbo_device()
        return float(self.elbo.item())

    def relinearize(self):
        """cvi_dp_trainer.py:110-134 on every rank's extended chain (pointwise: no exchange, the halos are there)."""
        self._th_sig = None
        for st in self.ranks:
            st.eng.relinearize()

    def check_status(self):
        self.chain.check_status()
        for st in self.ranks:
            st.eng.check_status()

    # ---- owned slices, for gathering / comparing
    def owned(self, name):
        """list over local ranks of the owned stretch of ``name`` (g1, gd, fx_mus, fx_covs: grid points;
        gs, qA, qb, qcholQ: transitions)"""
        out = []
        for st in self.ranks:
            t = getattr(st.eng, name)[0]
            if name in ("gs", "qA", "qb", "qcholQ", "pA", "pb", "pcholQ"):
                out.append(t[st.lo:st.lo + st.nt])
            else:
                out.append(t[st.lo:st.hi])
        return out
