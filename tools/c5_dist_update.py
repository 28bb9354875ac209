#!/usr/bin/env python
"""
BASELINE config 5 over several GPUs (launch under torchrun, one process per GPU): ONE sequence, N grid steps
(default 10^7), D = 3, full CVI-DP site update with the quadrature sharded over ranks
(vi_diffusion_processes_b200.dist_quad.ShardedQuadratureCvi).  Rank 0 prints one JSON line.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
      tools/c5_dist_update.py
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vi_diffusion_processes_b200.cvi import CviEngine  # noqa: E402
from vi_diffusion_processes_b200.dist_quad import ShardedQuadratureCvi  # noqa: E402
from vi_diffusion_processes_b200.dist_time import Comm  # noqa: E402

DRIFT_OU_ND = 8


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=10_000_000)
    ap.add_argument("--obs", type=int, default=10_000)
    ap.add_argument("--steps", type=int, default=2)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    comm = Comm.from_process_group() if world > 1 else Comm.emulated(1)
    N, D = a.grid, 3
    rng = np.random.default_rng(7)                       # the same problem on every rank
    grid = np.arange(N + 1) * 1e-3
    idx = np.sort(rng.choice(N + 1, a.obs, replace=False))
    y = rng.normal(size=(a.obs, D)) * 0.5
    eng = CviEngine(DRIFT_OU_ND, [0.5], np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.zeros(D), np.eye(D), device=dev)
    sh = ShardedQuadratureCvi([eng], comm)
    sh.site_update(0.5, 0.5)                              # warm
    eng.check_status()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        elbo = sh.site_update(0.5, 0.5)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    eng.check_status()
    if rank == 0:
        print(json.dumps({"workload": f"C5: 1 sequence x {N} grid steps, D=3, OU prior, {a.obs} observations, one "
                                      "CVI-DP site update, quadrature sharded over ranks (rebuilds replicated)",
                          "n_gpus": world, "ms_per_update": float(ms.item()),
                          "grid_steps_per_s": N / (float(ms.item()) * 1e-3), "elbo": elbo,
                          "site_bytes_exchanged_per_update": (N + 1) * (D + 2 * D * D) * 8}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
