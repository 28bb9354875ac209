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


def _pack(e):
    return torch.as_tensor(np.concatenate([e[k].reshape(-1) for k in ("La", "Lab", "Lb", "va", "vb")]))


def _unpack(t, D):
    a = t.numpy()
    DD = D * D
    return dict(La=a[:DD].reshape(D, D), Lab=a[DD:2 * DD].reshape(D, D), Lb=a[2 * DD:3 * DD].reshape(D, D),
                va=a[3 * DD:3 * DD + D], vb=a[3 * DD + D:])


def _worker_info(rank, world, port, q):
    """The two-sweep rebuild of a time-sharded chain with the ranks in separate processes, the kernels replaced by the
    oracle's NumPy elements (oracle/info_form_np.py): SUMMARY (the element of the rank's stretch) -> all-gather ->
    incoming carry from the ranks AFTER this one (backward sweep) / BEFORE it (forward sweep) -> recursion over the
    own stretch.  What the CUDA path does per rank in vdp_dist_sweep_summary / finish (VDP_SWEEP_INFO, VDP_SWEEP_MOMENTS)."""
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import gauss_markov_np as G
        from oracle import info_form_np as I
        from vi_diffusion_processes_b200.dist_time import Comm, time_partition
        comm = Comm.from_process_group()
        D, N = 2, 13
        rng = np.random.default_rng(5)                       # the same chain on every rank; each uses its stretch
        A = 0.6 * np.eye(D) + 0.1 * rng.normal(size=(N, D, D))
        b = rng.normal(size=(N, D))
        R = rng.normal(size=(N, D, D))
        cQ = np.linalg.cholesky(0.05 * np.eye(D) + 0.02 * R @ G.mT(R))
        ssm = G.StateSpaceModel(rng.normal(size=D), np.linalg.cholesky(0.5 * np.eye(D)), A, b, cQ)
        t1, td, ts = G.ssm_to_naturals(ssm)
        Kd = G.symmetrise_from_lower(-2.0 * td)
        s, e = time_partition(N, world)[rank]
        last = rank == world - 1
        # ---- backward sweep in information form
        el = None
        for k in range(e - 1, s - 1, -1):
            st = I.info_step(ts[k], Kd[k], t1[k])
            el = st if el is None else I.info_combine(el, st)
        tot = comm.all_gather([_pack(el)])
        init = comm.all_gather([torch.as_tensor(np.concatenate([Kd[N].reshape(-1), t1[N]]) if last
                                                else np.full(D * D + D, np.nan))])
        J, h = init[world - 1][:D * D].numpy().reshape(D, D), init[world - 1][D * D:].numpy()
        for r in range(world - 1, rank, -1):
            J, h = I.info_apply(_unpack(tot[r], D), J, h)
        As, bs, Qs = np.empty((e - s, D, D)), np.empty((e - s, D)), np.empty((e - s, D, D))
        for k in range(e - 1, s - 1, -1):
            Q = np.linalg.inv(J)
            As[k - s], bs[k - s], Qs[k - s] = Q @ ts[k], Q @ h, Q
            J, h = Kd[k] - ts[k].T @ As[k - s], t1[k] + ts[k].T @ bs[k - s]
        # ---- forward sweep of the moments: element (M, v, C) of m' = M m + v, S' = M S M^T + C
        M, v, Cc = np.eye(D), np.zeros(D), np.zeros((D, D))
        for k in range(e - s):
            M, v, Cc = As[k] @ M, As[k] @ v + bs[k], As[k] @ Cc @ As[k].T + Qs[k]
        tot = comm.all_gather([torch.as_tensor(np.concatenate([M.reshape(-1), v, Cc.reshape(-1)]))])
        P0 = np.linalg.inv(J) if rank == 0 else np.full((D, D), np.nan)
        mu0 = P0 @ h if rank == 0 else np.full(D, np.nan)
        init = comm.all_gather([torch.as_tensor(np.concatenate([mu0, P0.reshape(-1)]))])
        m, S = init[0][:D].numpy(), init[0][D:].numpy().reshape(D, D)
        for r in range(rank):
            t = tot[r].numpy()
            Mr, vr, Cr = t[:D * D].reshape(D, D), t[D * D:D * D + D], t[D * D + D:].reshape(D, D)
            m, S = Mr @ m + vr, Mr @ S @ Mr.T + Cr
        ms, Ss = [m], [S]
        for k in range(e - s):
            ms.append(As[k] @ ms[-1] + bs[k])
            Ss.append(As[k] @ Ss[-1] @ As[k].T + Qs[k])
        n_own = e - s + (1 if last else 0)
        wm, wS = ssm.marginals
        n0 = s
        err = max(np.abs(As - A[s:e]).max(), np.abs(bs - b[s:e]).max(),
                  np.abs(Qs - cQ[s:e] @ G.mT(cQ[s:e])).max(),
                  np.abs(np.stack(ms[:n_own]) - wm[n0:n0 + n_own]).max(),
                  np.abs(np.stack(Ss[:n_own]) - wS[n0:n0 + n_own]).max())
        q.put((rank, float(err)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_sweep_rebuild_across_processes(world):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_info, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r for r, _ in res] == list(range(world))
    assert max(err for _, err in res) <= 1e-11, res
