"""Tuning run: end-to-end (host buffers) time of one C4 site update for several chunk sizes / slot counts."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bench
DRIFT_DW = 1   # VDP_DRIFT_DW (include/vdp_b200.h)
from vi_diffusion_processes_b200.cvi import CviEngine, CviHostEngine

B, N = 4096, 10000
grid, idx, y = bench.make_problem(B, N, seed=4096)
eng = CviEngine(DRIFT_DW, bench.DW_PARAMS, np.ones((1, 1)), grid, idx, y[..., None], [[bench.NOISE_STD]],
                [bench.X0_MU], [[bench.X0_COV]], device="cuda:0")
for _ in range(3):
    eng.site_update(bench.LR, bench.LR)
for chunk, slots in [(256, 3), (128, 3), (512, 3), (256, 4), (128, 4), (64, 4), (512, 2), (1024, 3)]:
    h = CviHostEngine(eng, chunk=chunk, slots=slots, want_posterior_ssm=False)
    h.site_update(bench.LR, bench.LR)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        h.site_update(bench.LR, bench.LR)
    torch.cuda.synchronize()
    print(chunk, slots, round((time.perf_counter() - t0) / 4 * 1e3, 2), "ms", flush=True)
    h.close()
