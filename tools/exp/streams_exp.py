"""Experiment: does splitting the batch over CUDA streams overlap the FP64-bound and the memory-bound kernels?"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from bench import make_problem, DW_PARAMS, NOISE_STD, X0_MU, X0_COV, LR
DRIFT_DW = 1   # VDP_DRIFT_DW (include/vdp_b200.h)
from vi_diffusion_processes_b200.cvi import CviEngine

B, N = 4096, 10000
for k in (1, 2, 4, 8):
    engs, streams = [], []
    for j in range(k):
        grid, idx, y = make_problem(B // k, N, seed=4096 + j)
        engs.append(CviEngine(DRIFT_DW, DW_PARAMS, np.ones((1, 1)), grid, idx, y[..., None], [[NOISE_STD]], [X0_MU],
                              [[X0_COV]]))
        streams.append(torch.cuda.Stream())
    def step():
        for e, s in zip(engs, streams):
            with torch.cuda.stream(s):
                e.site_update(LR, LR)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for _ in range(5):
        step()
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    print(k, "streams:", e0.elapsed_time(e1) / 5, "ms/update", flush=True)
    del engs
    torch.cuda.empty_cache()
