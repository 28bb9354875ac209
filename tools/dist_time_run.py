#!/usr/bin/env python
"""
Time-segment sharding on real ranks (launch under torchrun, one process per GPU; also runs with 1 process):
  1. correctness: the posterior rebuild of one sequence split over the ranks == the same rebuild on one GPU;
  2. timing of BASELINE config 5 (one sequence, N = 10^7 grid steps, D = 3): naturals -> SSM -> marginals.
Prints one JSON line on rank 0.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
      tools/dist_time_run.py --grid 10000000
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vi_diffusion_processes_b200 import ops  # noqa: E402
from vi_diffusion_processes_b200.dist_time import Comm, TimeShardedChain  # noqa: E402

F64 = torch.float64


def make_ssm(N, D, dev, seed=7):
    """A stable random state-space model of N transitions (the same on every rank)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    eye = torch.eye(D, dtype=F64, device=dev)
    A = 0.97 * eye + 0.02 * torch.randn((1, N, D, D), dtype=F64, device=dev, generator=g) / D ** 0.5
    b = 0.05 * torch.randn((1, N, D), dtype=F64, device=dev, generator=g)
    cQ = 0.1 * eye + torch.tril(0.01 * torch.randn((1, N, D, D), dtype=F64, device=dev, generator=g))
    cQ = cQ.contiguous()
    mu0 = torch.zeros((1, D), dtype=F64, device=dev)
    cP0 = (0.7 * eye)[None].contiguous()
    return A.contiguous(), b.contiguous(), cQ, mu0, cP0


def run(N, D, comm, rank, world, dev, steps, check):
    A, b, cQ, mu0, cP0 = make_ssm(N, D, dev)
    th1, thd, ths = ops.ssm_to_naturals(A, b, cQ, mu0, cP0)
    chain = TimeShardedChain(N, D, comm, device=dev)
    s, e = chain.parts[rank]
    n0, n1 = chain.owned_nodes(rank)
    l1, ld, ls = [th1[0, n0:n1].contiguous()], [thd[0, n0:n1].contiguous()], [ths[0, s:e].contiguous()]
    ref = None
    if check:
        rA, rb, rcP0, rcQ, rmu0 = ops.naturals_to_ssm(th1, thd, ths)
        rm, rS = ops.ssm_marginals(rA, rb, rcQ, rmu0, rcP0)
        ref = (rA[0, s:e], rb[0, s:e], rcQ[0, s:e], rm[0, n0:n1], rS[0, n0:n1])
    del A, b, cQ, th1, thd, ths
    torch.cuda.empty_cache()

    def step():
        out = chain.naturals_to_ssm(l1, ld, ls)[0]
        marg = chain.ssm_marginals([out["A"]], [out["b"]], [out["cholQ"]], [out["mu0"]], [out["cholP0"]])[0]
        return out, marg

    out, marg = step()
    chain.check_status()
    err = 0.0
    if check:
        for got, want in zip((out["A"], out["b"], out["cholQ"], marg[0], marg[1]), ref):
            err = max(err, float((got - want).abs().max() / want.abs().max()))
    for _ in range(2):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=F64, device=dev)
    errt = torch.tensor([err], dtype=F64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(errt, op=dist.ReduceOp.MAX)
    return float(ms.item()), float(errt.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=10 ** 7)
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--check-grid", type=int, default=200001)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        comm = Comm.from_process_group()
    else:
        comm = Comm.emulated(1)
    _, err = run(args.check_grid, args.dim, comm, rank, world, dev, 1, check=True)
    ms, _ = run(args.grid, args.dim, comm, rank, world, dev, args.steps, check=False)
    if rank == 0:
        print(json.dumps({"workload": "C5: one sequence, time-segment sharded posterior rebuild (naturals -> SSM -> marginals)",
                          "grid_steps": args.grid, "state_dim": args.dim, "n_gpus": world, "ms_per_rebuild": ms,
                          "grid_steps_per_s": args.grid / (ms * 1e-3),
                          "max_rel_err_vs_single_gpu": err, "check_grid": args.check_grid}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
