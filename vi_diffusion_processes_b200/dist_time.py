"""
Time-segment sharding of ONE long sequence over ranks (BASELINE config 5): the posterior rebuild
(naturals -> state-space model -> marginals; variational_cvi_sde.py:176-192 + gauss_markov.py:107-117) with
every rank holding a contiguous range of the grid.

What crosses ranks per sweep is one associative element and one initial state (``vdp_dist_blob_bytes`` bytes,
a few hundred) plus one-block halos of the inputs; the exchange is an all-gather over ``torch.distributed``
(NCCL on GPUs; gloo in the CPU tests of the plumbing).  The sweeps themselves are the tiled CUDA kernels
(``vdp_dist_sweep_summary`` / ``vdp_dist_sweep_finish``).

The algorithm is written in lockstep over a list of rank states so that the same code runs
  * distributed: the list holds this process's rank, ``Comm`` gathers over the process group;
  * emulated: the list holds all R ranks of one process / one GPU (tests, no second GPU needed).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from . import _lib
from ._lib import DistSweep, check

F64 = torch.float64


def time_partition(N: int, nranks: int):
    """Contiguous transition ranges [s_r, e_r) covering [0, N); the last rank also owns grid point N."""
    base, extra = divmod(N, nranks)
    out, s = [], 0
    for r in range(nranks):
        e = s + base + (1 if r < extra else 0)
        out.append((s, e))
        s = e
    if any(e - s < 1 for s, e in out):
        raise ValueError(f"cannot split {N} transitions over {nranks} ranks")
    return out


class Comm:
    """
    All-gather of one small tensor per rank.  ``local_ranks`` = the ranks driven by this process.

    Transports: a plain list when all ranks live in this process (``emulated``: tests); ``torch.distributed`` (gloo
    in the CPU tests of the plumbing); or -- the GPU data path -- ``vdp_comm_all_gather``: NCCL issued from C on the
    compute stream (``from_process_group(backend="nccl_c")``; torch.distributed then only carries the 128-byte NCCL
    id to the ranks once).
    """

    def __init__(self, nranks: int, local_ranks: List[int], group=None, c_comm=None):
        self.nranks, self.local_ranks, self.group = nranks, list(local_ranks), group
        self.distributed = len(self.local_ranks) != nranks
        self._c_comm = c_comm
        self.n_gathers = 0            # exchanges issued (reported by the bench)

    @classmethod
    def emulated(cls, nranks):
        return cls(nranks, list(range(nranks)))

    @classmethod
    def from_process_group(cls, group=None, backend="torch"):
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if backend == "torch" or world == 1:
            return cls(world, [rank], group)
        assert backend == "nccl_c", backend
        lib = _lib.load()
        ident = (C.c_char * 128)()
        if rank == 0:
            check(lib.vdp_comm_unique_id(ident), "vdp_comm_unique_id")
        box = [bytes(ident)]
        dist.broadcast_object_list(box, src=0, group=group)
        handle = C.c_void_p()
        check(lib.vdp_comm_create(C.byref(handle), box[0], rank, world), "vdp_comm_create")   # current CUDA device
        return cls(world, [rank], group, c_comm=handle)

    def close(self):
        if self._c_comm is not None:
            _lib.load().vdp_comm_destroy(self._c_comm)
            self._c_comm = None

    def all_gather(self, per_local: List[torch.Tensor]) -> List[torch.Tensor]:
        """per_local[i] = the tensor of local rank i -> list over ALL ranks (same for every local rank)."""
        self.n_gathers += 1
        if not self.distributed:
            return list(per_local)
        t = per_local[0].contiguous()
        if self._c_comm is not None:
            out = torch.empty((self.nranks,) + tuple(t.shape), dtype=t.dtype, device=t.device)
            stream = C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)
            check(_lib.load().vdp_comm_all_gather(self._c_comm, C.c_void_p(t.data_ptr()), C.c_void_p(out.data_ptr()),
                                                  t.numel() * t.element_size(), stream), "vdp_comm_all_gather")
            return list(out.unbind(0))
        import torch.distributed as dist
        out = [torch.empty_like(t) for _ in range(self.nranks)]
        dist.all_gather(out, t, group=self.group)
        return out


def halo_from_next(comm: Comm, blocks: List[torch.Tensor]):
    """blocks[i] = what local rank i offers to the rank before it -> what local rank i receives from the rank
    after it (None on the last rank)."""
    allv = comm.all_gather(blocks)
    return [allv[r + 1] if r + 1 < comm.nranks else None for r in comm.local_ranks]


def halo_from_prev(comm: Comm, blocks: List[torch.Tensor]):
    """blocks[i] = what local rank i offers to the rank after it -> what it receives from the rank before it."""
    allv = comm.all_gather(blocks)
    return [allv[r - 1] if r > 0 else None for r in comm.local_ranks]


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class _Rank:
    """Per-rank scratch for the C calls."""

    def __init__(self, rank, nranks, s, e, D, device):
        self.rank, self.nranks, self.s, self.e, self.D = rank, nranks, s, e, D
        self.N_local = e - s
        self.first, self.last = rank == 0, rank == nranks - 1
        self.device = device
        lib = _lib.load()
        self.ws = torch.empty(lib.vdp_dist_workspace(D), dtype=torch.uint8, device=device)
        self.status = torch.zeros(2, dtype=torch.int64, device=device)

    def sweep(self, kind, stage, ins, outs, scaleD=1.0, scaleS=1.0, prev=None):
        p = DistSweep()
        p.kind, p.stage, p.rank, p.nranks, p.N_local, p.D = kind, stage, self.rank, self.nranks, self.N_local, self.D
        p.scaleD, p.scaleS = scaleD, scaleS
        for i, t in enumerate(ins):
            p.__getattribute__("in")[i] = _ptr(t)
        for i, t in enumerate(outs):
            p.out[i] = _ptr(t)
        if prev is not None:
            p.prev_totals, p.prev_inits = prev[0].data_ptr(), prev[1].data_ptr()
        p.workspace, p.workspace_bytes = self.ws.data_ptr(), self.ws.numel()
        p.status = self.status.data_ptr()
        return p


def _blob_bytes(kind, stage, D):
    tb, ib = C.c_size_t(), C.c_size_t()
    check(_lib.load().vdp_dist_blob_bytes(kind, stage, D, C.byref(tb), C.byref(ib)), "vdp_dist_blob_bytes")
    return tb.value, ib.value


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)   # callers hold torch.cuda.device(self.device)


def _run_sweep(comm: Comm, ranks: List[_Rank], kind, problems_stage, D):
    """summary -> all-gather -> [summary stage 1 -> all-gather] -> finish, in lockstep over the local ranks.
    ``problems_stage(rank_state, stage, prev)`` builds the DistSweep of one rank."""
    lib = _lib.load()
    stages = 2 if kind == _lib.SWEEP_NAT2SSM else 1
    prev = [None] * len(ranks)
    gathered = None
    for stage in range(stages):
        tb, ib = _blob_bytes(kind, stage, D)
        # one blob per rank = [scan element of the rank's stretch | its natural initial state]: one all-gather per stage
        tbp = (tb + 15) // 16 * 16
        blobs = []
        for i, rk in enumerate(ranks):
            blob = torch.zeros(tbp + ib, dtype=torch.uint8, device=rk.device)
            p = problems_stage(rk, stage, prev[i])
            check(lib.vdp_dist_sweep_summary(C.byref(p), blob.data_ptr(), blob.data_ptr() + tbp, _stream()),
                  "vdp_dist_sweep_summary")
            blobs.append(blob)
        allb = comm.all_gather(blobs)
        all_t = torch.cat([x[:tb] for x in allb]).contiguous()
        all_i = torch.cat([x[tbp:] for x in allb]).contiguous()
        gathered = (all_t, all_i)
        prev = [gathered] * len(ranks)
    for i, rk in enumerate(ranks):
        p = problems_stage(rk, stages - 1, prev[i] if stages == 2 else None)
        check(lib.vdp_dist_sweep_finish(C.byref(p), gathered[0].data_ptr(), gathered[1].data_ptr(), _stream()),
              "vdp_dist_sweep_finish")


class TimeShardedChain:
    """
    Posterior rebuild of one sequence of N transitions, state dimension D, sharded in time.

    Per local rank r with range [s, e): ``th1`` [n_own, D], ``thd`` [n_own, D, D] for the owned grid points
    (n_own = e - s, plus the final point N on the last rank), ``ths`` [e - s, D, D].
    """

    def __init__(self, N: int, D: int, comm: Comm, device="cuda"):
        self.N, self.D, self.comm = N, D, comm
        self.parts = time_partition(N, comm.nranks)
        self.ranks = [_Rank(r, comm.nranks, *self.parts[r], D, torch.device(device)) for r in comm.local_ranks]

    def owned_nodes(self, r):
        s, e = self.parts[r]
        return s, e + (1 if r == self.comm.nranks - 1 else 0)

    # ---- halo helpers: element j of the result = what local rank j needs from its neighbour
    def _from_next(self, firsts: List[torch.Tensor]):
        return halo_from_next(self.comm, firsts)

    def _from_prev(self, lasts: List[torch.Tensor]):
        return halo_from_prev(self.comm, lasts)

    def naturals_to_ssm(self, th1: List[torch.Tensor], thd: List[torch.Tensor], ths: List[torch.Tensor],
                        want_cov: bool = False, out: Optional[List[dict]] = None):
        """naturals_to_ssm_params (ssm_gaussian_transformations.py:333-511), time-sharded.
        Returns per local rank a dict: A [n,D,D], b [n,D], cholQ [n,D,D] for its transitions (n = e - s), the
        node-s factor ``cholQ_prev`` (cholQ of transition s-1; None on rank 0) and mu0 / cholP0 on rank 0.
        ``want_cov``: also ``S`` [n_own, D, D], the marginal covariances of the owned grid points (the diagonal of
        the inverse band the backward sweep walks through -- no second factorisation needed for them).

        ``out`` (optional, per local rank): the caller's own arrays for the sweeps to write into instead of fresh
        ones (a chain of 10^7 steps moves 0.7 GB per copy): ``A`` [n,D,D], ``b`` [n,D], ``cholQ`` [n,D,D], ``S``
        [n_own,D,D]; ``b_prev`` [1,D] / ``cholQ_prev`` [1,D,D] = the slots right BEFORE ``b`` / ``cholQ`` in memory
        on every rank but the first (a rank's first node produces offset and factor of the transition into it);
        ``thd_ext`` = ``thd`` with one more row behind it, which receives the next rank's first block (unused on
        the last rank)."""
        D, dev = self.D, self.ranks[0].device
        halo = self._from_next([t[:1].contiguous() for t in thd])
        Kd, Ld, Ls = [], [], []
        for i, (rk, d_own, h) in enumerate(zip(self.ranks, thd, halo)):
            if rk.last:
                kd = d_own
            elif out is not None and out[i].get("thd_ext") is not None:
                kd = out[i]["thd_ext"]
                assert kd.data_ptr() == d_own.data_ptr() and kd.is_contiguous()
                kd[-1] = h[0]
            else:
                kd = torch.cat([d_own, h])
            assert kd.shape[0] == rk.N_local + 1
            Kd.append(kd.contiguous())
            Ld.append(torch.empty_like(Kd[-1]))
            Ls.append(torch.empty((rk.N_local, D, D), dtype=F64, device=dev))
        idx = {id(rk): i for i, rk in enumerate(self.ranks)}
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_CHOL,
                   lambda rk, st, prev: rk.sweep(_lib.SWEEP_CHOL, 0, [Kd[idx[id(rk)]], ths[idx[id(rk)]]],
                                                 [Ld[idx[id(rk)]], Ls[idx[id(rk)]]], -2.0, -1.0), D)
        A, bo, qo, bprev, qprev, mu0, cP0, Sd = [], [], [], [], [], [], [], []
        for i, rk in enumerate(self.ranks):
            n, n_own = rk.N_local, rk.N_local + (1 if rk.last else 0)
            o = out[i] if out is not None else {}
            A.append(o["A"] if "A" in o else torch.empty((n, D, D), dtype=F64, device=dev))
            if "b" in o:
                bo.append(o["b"])
                qo.append(o["cholQ"])
                bprev.append(None if rk.first else o["b_prev"])
                qprev.append(None if rk.first else o["cholQ_prev"])
                if not rk.first:       # slot j - 1 <-> transition s - 1 + j: one slot before the owned ones
                    assert bprev[-1].data_ptr() + D * 8 == bo[-1].data_ptr()
                    assert qprev[-1].data_ptr() + D * D * 8 == qo[-1].data_ptr()
            else:
                bb = torch.empty((n + 1, D), dtype=F64, device=dev)
                qb = torch.empty((n + 1, D, D), dtype=F64, device=dev)
                bo.append(bb[1:])
                qo.append(qb[1:])
                bprev.append(None if rk.first else bb[:1])
                qprev.append(None if rk.first else qb[:1])
            mu0.append(torch.zeros(D, dtype=F64, device=dev))
            cP0.append(torch.zeros((D, D), dtype=F64, device=dev))
            Sd.append((o["S"] if "S" in o else torch.empty((n_own, D, D), dtype=F64, device=dev)) if want_cov else None)
            for t, shape in ((A[-1], (n, D, D)), (bo[-1], (n, D)), (qo[-1], (n, D, D))):
                assert tuple(t.shape) == shape and t.is_contiguous(), (tuple(t.shape), shape)
            assert Sd[-1] is None or (tuple(Sd[-1].shape) == (n_own, D, D) and Sd[-1].is_contiguous())

        def n2s(rk, stage, prev):
            i = idx[id(rk)]
            return rk.sweep(_lib.SWEEP_NAT2SSM, stage, [Ld[i], Ls[i], th1[i], thd[i], ths[i]],
                            [A[i], bo[i], qo[i], mu0[i], cP0[i], Sd[i]], prev=prev)
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_NAT2SSM, n2s, D)
        # transition e-1 (output of node e) lives in the slot before the next rank's own (offset and factor in one
        # exchange; the first rank has no such slot and offers zeros nobody reads)
        zero = torch.zeros(D + D * D, dtype=F64, device=dev)
        nbq = self._from_next([zero if bp is None else torch.cat([bp.reshape(-1), qp.reshape(-1)])
                               for bp, qp in zip(bprev, qprev)])
        res = []
        for i, rk in enumerate(self.ranks):
            if not rk.last:
                bo[i][-1] = nbq[i][:D]
                qo[i][-1] = nbq[i][D:].reshape(D, D)
            res.append(dict(A=A[i], b=bo[i], cholQ=qo[i], cholQ_prev=qprev[i],
                            mu0=mu0[i] if rk.first else None, cholP0=cP0[i] if rk.first else None, S=Sd[i]))
        return res

    def ssm_means(self, A: List[torch.Tensor], b: List[torch.Tensor], mu0: List[Optional[torch.Tensor]],
                  out: Optional[List[torch.Tensor]] = None):
        """StateSpaceModel.marginal_means (state_space_model.py:251), time-sharded: per local rank m [n_own, D].
        ``out`` (optional): per local rank an array [n + 1, D] to write into -- the sweep of a rank also produces the
        mean of the next rank's first grid point, in the last row."""
        D, dev = self.D, self.ranks[0].device
        m = [torch.empty((rk.N_local + 1, D), dtype=F64, device=dev) for rk in self.ranks] if out is None else out
        for rk, t in zip(self.ranks, m):
            assert tuple(t.shape) == (rk.N_local + 1, D) and t.is_contiguous()
        idx = {id(rk): i for i, rk in enumerate(self.ranks)}
        dummy_v = torch.zeros(D, dtype=F64, device=dev)

        def mean(rk, stage, prev):
            i = idx[id(rk)]
            return rk.sweep(_lib.SWEEP_MEAN, 0, [A[i].contiguous(), b[i].contiguous(),
                                                 mu0[i] if rk.first else dummy_v], [m[i]])
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_MEAN, mean, D)
        return [m[i][:rk.N_local + (1 if rk.last else 0)] for i, rk in enumerate(self.ranks)]

    # ---- the rebuild in two sweeps: information form from the bottom, moments from the top
    def naturals_to_ssm_info(self, th1: List[torch.Tensor], thd: List[torch.Tensor], ths: List[torch.Tensor],
                             out: Optional[List[dict]] = None):
        """naturals_to_ssm_params as ONE backward sweep in information form (``VDP_SWEEP_INFO``: the Schur complements
        of the precision taken from the bottom with the linear term riding along; csrc/scan_elems.cuh InfoOps).  No
        factor is stored, no halo of the inputs is exchanged, every output lands at the index of its transition.
        Returns per local rank a dict A [n,D,D], b [n,D], cholQ [n,D,D], mu0 / cholP0 (rank 0, else None).
        ``out`` (optional, per local rank): ``A``, ``b``, ``cholQ`` arrays to write into."""
        D, dev = self.D, self.ranks[0].device
        A, bo, qo, mu0, cP0 = [], [], [], [], []
        for i, rk in enumerate(self.ranks):
            n = rk.N_local
            o = out[i] if out is not None else {}
            A.append(o["A"] if "A" in o else torch.empty((n, D, D), dtype=F64, device=dev))
            bo.append(o["b"] if "b" in o else torch.empty((n, D), dtype=F64, device=dev))
            qo.append(o["cholQ"] if "cholQ" in o else torch.empty((n, D, D), dtype=F64, device=dev))
            mu0.append(torch.zeros(D, dtype=F64, device=dev))
            cP0.append(torch.zeros((D, D), dtype=F64, device=dev))
            for t, shape in ((A[-1], (n, D, D)), (bo[-1], (n, D)), (qo[-1], (n, D, D))):
                assert tuple(t.shape) == shape and t.is_contiguous(), (tuple(t.shape), shape)
            n_own = n + (1 if rk.last else 0)
            assert th1[i].shape[0] == n_own and thd[i].shape[0] == n_own and ths[i].shape[0] == n
        idx = {id(rk): i for i, rk in enumerate(self.ranks)}

        def info(rk, stage, prev):
            i = idx[id(rk)]
            return rk.sweep(_lib.SWEEP_INFO, 0, [th1[i].contiguous(), thd[i].contiguous(), ths[i].contiguous()],
                            [A[i], bo[i], qo[i], mu0[i], cP0[i]])
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_INFO, info, D)
        return [dict(A=A[i], b=bo[i], cholQ=qo[i], mu0=mu0[i] if rk.first else None,
                     cholP0=cP0[i] if rk.first else None) for i, rk in enumerate(self.ranks)]

    def ssm_moments(self, A: List[torch.Tensor], b: List[torch.Tensor], cholQ: List[torch.Tensor],
                    mu0: List[Optional[torch.Tensor]], cholP0: List[Optional[torch.Tensor]],
                    out: Optional[List[tuple]] = None):
        """StateSpaceModel.marginals (state_space_model.py:231-262) by the moment recursion m' = A m + b,
        S' = A S A^T + Q (``VDP_SWEEP_MOMENTS``), time-sharded.  Returns per local rank (m [n_own, D], S [n_own, D, D]).
        ``out`` (optional): per local rank (m [n + 1, D], S [n + 1, D, D]) to write into -- the sweep of a rank also
        produces the moments of the next rank's first grid point, in the last row."""
        D, dev = self.D, self.ranks[0].device
        if out is None:
            out = [(torch.empty((rk.N_local + 1, D), dtype=F64, device=dev),
                    torch.empty((rk.N_local + 1, D, D), dtype=F64, device=dev)) for rk in self.ranks]
        for rk, (m, S) in zip(self.ranks, out):
            assert tuple(m.shape) == (rk.N_local + 1, D) and m.is_contiguous()
            assert tuple(S.shape) == (rk.N_local + 1, D, D) and S.is_contiguous()
        idx = {id(rk): i for i, rk in enumerate(self.ranks)}
        dummy_v = torch.zeros(D, dtype=F64, device=dev)
        dummy_m = torch.zeros((D, D), dtype=F64, device=dev)

        def mom(rk, stage, prev):
            i = idx[id(rk)]
            return rk.sweep(_lib.SWEEP_MOMENTS, 0,
                            [A[i].contiguous(), b[i].contiguous(), cholQ[i].contiguous(),
                             mu0[i] if rk.first else dummy_v, cholP0[i] if rk.first else dummy_m],
                            [out[i][0], out[i][1]])
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_MOMENTS, mom, D)
        res = []
        for rk, (m, S) in zip(self.ranks, out):
            k = rk.N_local + (1 if rk.last else 0)
            res.append((m[:k], S[:k]))
        return res

    def ssm_marginals(self, A: List[torch.Tensor], b: List[torch.Tensor], cholQ: List[torch.Tensor],
                      mu0: List[Optional[torch.Tensor]], cholP0: List[Optional[torch.Tensor]]):
        """StateSpaceModel.marginals (state_space_model.py:231-262), time-sharded.  Returns per local rank
        (m [n_own, D], Sd [n_own, D, D]) for its owned grid points."""
        D, dev = self.D, self.ranks[0].device
        qprev = self._from_prev([t[-1:].contiguous() for t in cholQ])
        anext = self._from_next([t[:1].contiguous() for t in A])
        qnext = self._from_next([t[:1].contiguous() for t in cholQ])
        Abuf, qbuf, m, Ld, Ls, Sd = [], [], [], [], [], []
        zero = torch.zeros((1, D, D), dtype=F64, device=dev)
        for i, rk in enumerate(self.ranks):
            n = rk.N_local
            Abuf.append(torch.cat([A[i], zero if rk.last else anext[i]]).contiguous())
            qbuf.append(torch.cat([zero if rk.first else qprev[i], cholQ[i], zero if rk.last else qnext[i]]).contiguous())
            m.append(torch.empty((n + 1, D), dtype=F64, device=dev))
            Ld.append(torch.empty((n + 1, D, D), dtype=F64, device=dev))
            Ls.append(torch.empty((n, D, D), dtype=F64, device=dev))
            Sd.append(torch.empty((n + 1, D, D), dtype=F64, device=dev))
        idx = {id(rk): i for i, rk in enumerate(self.ranks)}
        dummy_v = torch.zeros(D, dtype=F64, device=dev)
        dummy_m = torch.zeros((D, D), dtype=F64, device=dev)

        def marg(rk, stage, prev):
            i = idx[id(rk)]
            return rk.sweep(_lib.SWEEP_MARG, 0,
                            [Abuf[i], b[i].contiguous(), qbuf[i][1:], mu0[i] if rk.first else dummy_v,
                             cholP0[i] if rk.first else dummy_m], [m[i], Ld[i], Ls[i]])
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_MARG, marg, D)
        _run_sweep(self.comm, self.ranks, _lib.SWEEP_INVBAND,
                   lambda rk, st, prev: rk.sweep(_lib.SWEEP_INVBAND, 0, [Ld[idx[id(rk)]], Ls[idx[id(rk)]]],
                                                 [Sd[idx[id(rk)]], None]), D)
        out = []
        for i, rk in enumerate(self.ranks):
            n_own = rk.N_local + (1 if rk.last else 0)
            out.append((m[i][:n_own], Sd[i][:n_own]))
        return out

    def check_status(self):
        for rk in self.ranks:
            code = int(rk.status[0].item())
            if code != 0:
                raise _lib.VdpError(f"rank {rk.rank}: non positive-definite block at local index {int(rk.status[1])}")
