"""
Sharded quadrature for ONE long sequence over ranks (BASELINE config 5, D = 3: the 8000-node Gauss-Hermite
quadrature is > 95 % of a site update, the block-tridiagonal sweeps ~2 %).

Every rank holds the whole chain (11 GB at N = 10^7, D = 3) and repeats the cheap posterior rebuilds; the KL
gradient + girsanov site update and the KL value run on the rank's own range of grid points
(``vdp_cvi_site_update_sharded``).  Per update the ranks exchange their updated site ranges (one broadcast per rank
and site tensor over NCCL / NVLink, (2 D^2 + D) * 8 bytes per grid point in total) and one partial KL each.
The result is the single-GPU result: the same kernels evaluate the same terms, a rank's first grid point gets the
contribution of the transition before it from the kernel's halo group exactly as an interior block boundary does.

For chains that do not fit one GPU, ``dist_time.TimeShardedChain`` shards the rebuild itself.

Written in lockstep over a list of engines so that the same code runs distributed (one engine per process) and
emulated (all ranks in one process; tests).
"""
from __future__ import annotations

from typing import List

import torch

from .cvi import CviEngine
from .dist_time import Comm, time_partition


class ShardedQuadratureCvi:
    def __init__(self, engines: List[CviEngine], comm: Comm):
        assert len(engines) == len(comm.local_ranks)
        e0 = engines[0]
        assert e0.B == 1, "one long sequence (use sequence sharding for batches)"
        self.engines, self.comm = engines, comm
        self.N, self.D = e0.N, e0.D
        # rank r owns grid points [s_r, e_r) (+ the final point N on the last rank) and transitions [s_r, e_r)
        self.parts = time_partition(self.N, comm.nranks)

    def node_range(self, r):
        s, e = self.parts[r]
        return s, e + (1 if r == self.comm.nranks - 1 else 0)

    def _exchange_sites(self):
        comm = self.comm
        for r in range(comm.nranks):
            n0, n1 = self.node_range(r)
            s, e = self.parts[r]
            views = lambda eng: (eng.g1[0, n0:n1], eng.gd[0, n0:n1], eng.gs[0, s:e])
            if comm.distributed:
                import torch.distributed as dist
                for v in views(self.engines[0]):
                    dist.broadcast(v, src=r, group=comm.group)   # contiguous slab of a [1, n, ...] tensor
            else:
                src = views(self.engines[r])
                for j, eng in enumerate(self.engines):
                    if j != r:
                        for dst, sv in zip(views(eng), src):
                            dst.copy_(sv)

    def site_update(self, lr_data, lr_girsanov):
        """cvi_dp_trainer.py:71-77 once; returns the ELBO (float) -- the same on every rank."""
        for eng, r in zip(self.engines, self.comm.local_ranks):
            eng.site_update_sharded(0, *self.node_range(r), lr_data=lr_data, lr_girsanov=lr_girsanov)
        self._exchange_sites()
        for eng, r in zip(self.engines, self.comm.local_ranks):
            eng.site_update_sharded(1, *self.node_range(r))
        parts = self.comm.all_gather([eng.kl.clone() for eng in self.engines])
        kl = parts[0].clone()
        for t in parts[1:]:                       # fixed rank order: bitwise reproducible
            kl = kl + t.to(kl.device)
        for eng in self.engines:
            eng.kl.copy_(kl.to(eng.kl.device))
            eng.elbo.copy_(eng.ve - eng.kl)
        return float(self.engines[0].elbo[0].item())
