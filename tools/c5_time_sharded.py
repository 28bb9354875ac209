#!/usr/bin/env python
"""
BASELINE config 5 with the chain sharded in TIME over the ranks (dist_cvi.TimeShardedCvi): ONE sequence of N grid
steps (default 10^7), D = 3 (3-D OU prior), 10^4 observations, full CVI-DP site update per step.  Strong scaling:

    python tools/c5_time_sharded.py                                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        tools/c5_time_sharded.py

Rank 0 prints one JSON line; `run()` is what bench.py --workload c5 calls.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

DRIFT_OU_ND = 8


def make_problem(N, n_obs, seed=7, dt=1e-3, D=3):
    rng = np.random.default_rng(seed)
    grid = np.arange(N + 1) * dt
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.5
    return grid, idx, y


def run(N=10_000_000, n_obs=10_000, steps=2, warmup=1, lr=0.5, keep=False, gh20=False, graphed=False, carry=True, info=None):
    """Returns (result dict on rank 0 else None[, the TimeShardedCvi]).  The process group, if any, must exist."""
    import torch.distributed as dist
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    device = torch.device("cuda", torch.cuda.current_device())
    D = 3
    grid, idx, y = make_problem(N, n_obs)
    comm = Comm.from_process_group(backend="nccl_c") if world > 1 else Comm.emulated(1)
    t_build = time.perf_counter()
    sh = TimeShardedCvi(DRIFT_OU_ND, [0.5], np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.zeros(D), np.eye(D), comm=comm,
                        device=device, carry_naturals=carry,
                        **({} if info is None else {"info_rebuild": bool(info)}))
    del grid
    if gh20:      # the reference's 20 points per dimension instead of the exact 2-point rule of the linear drift
        from vi_diffusion_processes_b200 import _lib
        for st in sh.ranks:
            st.eng.route_flags |= _lib.FLAG_GH_QUADRATURE
    update = sh.site_update_graphed if graphed else sh.site_update      # graphed: first call eager + capture
    # two updates at least: the second is the first one on the carried route (theta corrected at the observations)
    warmup = max(2, warmup)
    for _ in range(warmup + (1 if graphed else 0)):
        elbo = update(lr, lr)
    sh.check_status()
    t_build = time.perf_counter() - t_build

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    barrier()
    g0 = comm.n_gathers
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        elbo = update(lr, lr)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    sh.check_status()
    res = None
    if rank == 0:
        ms_step = float(ms.item()) / steps
        res = {"workload": f"C5: 1 sequence x {N} grid steps, D=3 OU prior, {n_obs} observations, one CVI-DP site "
                           f"update per step, chain sharded in time over {world} rank(s)",
               "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_step,
               "value": N / (ms_step * 1e-3), "unit": "grid-steps/s", "scaling": "strong", "elbo": elbo,
               "quadrature_nodes_per_transition": 8000 if gh20 else 8,
               "rebuild": "information form + moments (2 sweeps)" if sh.info_rebuild else "Cholesky + inverse band + means (3 sweeps)",
               "launch": "CUDA graph replay (launches and ncclAllGather captured once)" if graphed else "eager",
               "collective": "vdp_comm_all_gather (ncclAllGather issued from C on the compute stream)" if world > 1
               else "none (1 rank)",
               "all_gathers_per_step": (comm.n_gathers - g0) // steps,
               "grid_steps_per_rank": sh.ranks[0].nt,
               "peak_mem_gb_rank0": torch.cuda.max_memory_allocated() / 1e9, "setup_s": t_build}
    if keep:
        return res, sh
    comm.close()
    return res


def main():
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=10_000_000)
    ap.add_argument("--obs", type=int, default=10_000)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--graphed", action="store_true", help="replay the update from a CUDA graph")
    ap.add_argument("--no-carry", action="store_true", help="rebuild theta from the prior in both rebuilds")
    ap.add_argument("--info", type=int, default=None, help="1 / 0: the two-sweep rebuild in information form on / off")
    ap.add_argument("--gh20", action="store_true", help="the reference's 20 points per dimension")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    res = run(a.grid, a.obs, a.steps, a.warmup, gh20=a.gh20, graphed=a.graphed, carry=not a.no_carry, info=a.info)
    if res is not None:
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
