"""
CPU tests (gloo, world_size 2) of the multi-rank host logic: the all-gather plumbing and halo exchange of the
time-segment sharding (vi_diffusion_processes_b200.dist_time) and the sequence sharding bench.py uses.
No CUDA kernel runs here; the kernels of the same path are covered by tests/test_gpu_dist_time.py.
"""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vi_diffusion_processes_b200.dist_time import Comm, halo_from_next, halo_from_prev, time_partition
        comm = Comm.from_process_group()
        assert comm.nranks == world and comm.local_ranks == [rank] and comm.distributed
        N = 11
        parts = time_partition(N, world)
        s, e = parts[rank]
        grid = torch.arange(N + 1, dtype=torch.float64)
        own = grid[s:e + (1 if rank == world - 1 else 0)]
        # blobs: every rank sees all ranks' blobs, in rank order
        blob = torch.full((5,), rank, dtype=torch.uint8)
        allb = torch.cat(comm.all_gather([blob]))
        assert allb.tolist() == sum(([r] * 5 for r in range(world)), [])
        # halos: first owned point of the next rank / last owned point of the previous rank
        nxt = halo_from_next(comm, [own[:1].clone()])[0]
        prv = halo_from_prev(comm, [own[-1:].clone()])[0]
        if rank < world - 1:
            assert nxt.item() == float(parts[rank + 1][0])
        else:
            assert nxt is None
        if rank > 0:
            last_prev = parts[rank - 1][1] - 1
            assert prv.item() == float(last_prev)
        else:
            assert prv is None
        # a forward carry scan done the way the kernels do it: every rank combines the totals of the ranks
        # before it, in rank order, onto rank 0's initial state (affine maps x -> a x + c on doubles)
        rng = np.random.default_rng(100 + rank)
        a, c = rng.uniform(0.5, 1.0), rng.normal()
        tot = comm.all_gather([torch.tensor([a, c], dtype=torch.float64)])
        init = comm.all_gather([torch.tensor([3.0 if rank == 0 else float("nan")], dtype=torch.float64)])
        x = init[0].item()
        for r in range(rank):
            x = tot[r][0].item() * x + tot[r][1].item()
        q.put((rank, x, [t.tolist() for t in tot]))
    finally:
        dist.destroy_process_group()


def test_time_sharding_plumbing_world_size_2():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tot = res[0][2]
    assert res[0][1] == 3.0
    assert res[1][1] == pytest.approx(tot[0][0] * 3.0 + tot[0][1])
    assert res[0][2] == res[1][2]


def _seq_worker(rank, world, port, q):
    """bench.py's sharding: independent sequences per rank, the only collective is the MAX over ranks of the time."""
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        grid, idx, y = bench.make_problem(4, 50, seed=4096 + rank)
        ms = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        q.put((rank, float(ms.item()), float(y.sum()), idx.shape))
    finally:
        dist.destroy_process_group()


def test_sequence_sharding_world_size_2():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_seq_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == res[1][1] == 11.0          # max over ranks
    assert res[0][2] != res[1][2]                  # different sequences per rank
    assert res[0][3] == res[1][3] == (4, 50)


class _FakeEngine:
    """Stands in for CviEngine in the exchange logic of dist_quad (CPU tensors, no kernels)."""

    def __init__(self, N, D, rank):
        self.B, self.N, self.D = 1, N, D
        self.g1 = torch.full((1, N + 1, D), float(rank), dtype=torch.float64)
        self.gd = torch.full((1, N + 1, D, D), float(rank), dtype=torch.float64)
        self.gs = torch.full((1, N, D, D), float(rank), dtype=torch.float64)
        self.kl = torch.tensor([1.0 + rank], dtype=torch.float64)
        self.ve = torch.tensor([10.0], dtype=torch.float64)
        self.elbo = torch.zeros(1, dtype=torch.float64)
        self.calls = []

    def site_update_sharded(self, phase, n0, n1, lr_data=0.0, lr_girsanov=0.0):
        self.calls.append((phase, n0, n1))


def _quad_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vi_diffusion_processes_b200.dist_quad import ShardedQuadratureCvi
        from vi_diffusion_processes_b200.dist_time import Comm
        N, D = 13, 2
        eng = _FakeEngine(N, D, rank)
        sh = ShardedQuadratureCvi([eng], Comm.from_process_group())
        elbo = sh.site_update(0.5, 0.5)
        q.put((rank, elbo, eng.calls, eng.g1[0, :, 0].tolist(), eng.gs[0, :, 0, 0].tolist(), sh.parts))
    finally:
        dist.destroy_process_group()


def test_sharded_quadrature_exchange_world_size_2():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_quad_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    parts = res[0][5]
    (s0, e0), (s1, e1) = parts
    for rank, elbo, calls, g1, gs, _ in res:
        assert elbo == 10.0 - (1.0 + 2.0)                     # ve - sum of the ranks' partial KL
        n0, n1 = (s0, e0) if rank == 0 else (s1, e1 + 1)      # the last rank owns the final grid point
        assert calls == [(0, n0, n1), (1, n0, n1)]
        assert g1 == [0.0] * e0 + [1.0] * (e1 + 1 - s1)       # every rank holds every rank's site range
        assert gs == [0.0] * e0 + [1.0] * (e1 - s1)
