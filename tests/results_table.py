#!/usr/bin/env python
"""
(test infrastructure, not collected by pytest: it runs the oracle as the checker and as the CPU baseline)
BASELINE.md section 5 rows for configs C1-C3 (the golden fixtures made from the reference's shipped data files):
per config, on cuda:0 -- max relative error of ELBO / marginals / sites against the fixture, device time of one
CVI-DP site update and of one VDP sweep (CUDA events, sites after the 3 fixture updates), and the NumPy oracle's time
for the same update on one host core.  Prints one JSON object (kept as profiles/r02_c123_results.json).
"""
import glob
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import models_np as M          # noqa: E402  (checker / CPU baseline leg only)
from oracle import sde_np as S             # noqa: E402
from test_golden import _sde, relerr       # noqa: E402
from vi_diffusion_processes_b200.cvi import CviEngine   # noqa: E402
from vi_diffusion_processes_b200.vgp import VgpEngine   # noqa: E402


def device_ms(fn, reps=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


out = {}
for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz"))):
    g = np.load(path)
    name = os.path.basename(path)[:-4]
    D = g["obs_y"].shape[-1]
    N = g["time_grid"].shape[0] - 1
    lik = float(g["sigma"]) * np.eye(D)
    eng = CviEngine(int(g["drift_id"]), g["drift_params"], g["q"], g["time_grid"], g["obs_idx"], g["obs_y"], lik,
                    g["x0_mu"], g["x0_cov"])
    lr = float(g["lr"])
    elbos = []
    for it in range(3):
        elbos.append(eng.site_update(lr, lr).item())
        if it == 1:
            eng.relinearize()
    eng.check_status()
    th = [t[0].cpu().numpy() for t in eng._prior_naturals(eng.pA, eng.pb)]
    site_err = max(relerr(got[0].cpu().numpy(), g[k], scale=np.abs(t + g[k]).max())
                   for got, k, t in ((eng.g1, "cvi_g1", th[0]), (eng.gd, "cvi_gd", th[1]), (eng.gs, "cvi_gs", th[2])))
    row = {"D": int(D), "grid_steps": int(N), "n_obs": int(g["obs_idx"].shape[-1]),
           "err_elbo": relerr(elbos, g["cvi_elbo"]),
           "err_marginals": max(relerr(eng.fx_mus[0].cpu().numpy(), g["cvi_m"]),
                                relerr(eng.fx_covs[0].cpu().numpy(), g["cvi_S"])),
           "err_sites_vs_posterior_naturals": site_err,
           # UNSCALED: max |cuda - fixture| / max |fixture| per site tensor, beside the float64 floor of the route
           # itself on this data (float64 vs 80-bit evaluation of the oracle, make_golden.py)
           "err_sites_unscaled": {k: relerr(t[0].cpu().numpy(), g["cvi_" + k])
                                  for k, t in (("g1", eng.g1), ("gd", eng.gd), ("gs", eng.gs))},
           "site_floor_f64_vs_f80": {k: float(v) for k, v in zip(("g1", "gd", "gs"), g["site_floor_f64_vs_f80"])}}
    ms = device_ms(lambda: eng.site_update(0.1 * lr, 0.1 * lr))
    eng.check_status()
    row["cvi_ms_per_update"] = ms
    row["cvi_ms_per_update_cuda_graph"] = device_ms(lambda: eng.site_update_graphed(0.1 * lr, 0.1 * lr))
    eng.check_status()
    row["cvi_grid_steps_per_s"] = N / (ms * 1e-3)
    # the trainer's inner loop on the device (vdp_cvi_optimize_sites, CUDA-graph WHILE node): 20 updates per call
    # (tolerance 0: never stops early), wall time per update including the one read-back per call
    eng.optimize_sites(20, 0.0, 0.1 * lr, 0.1 * lr)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        r = eng.optimize_sites(20, 0.0, 0.1 * lr, 0.1 * lr)
    row["cvi_ms_per_update_device_loop_wall"] = (time.perf_counter() - t0) / (5 * 20) * 1e3
    assert int(r["n_itr"][0]) == 20
    eng.check_status()
    vg = VgpEngine(int(g["drift_id"]), g["drift_params"], g["q"], g["time_grid"], g["obs_idx"], g["obs_y"], lik,
                   g["x0_mu"], g["x0_cov"], stabilize_system=bool(g["vgp_stabilize"]))
    vg.A += 2.0 * torch.eye(D, dtype=torch.float64, device="cuda")
    vg.b += 2.0 * torch.as_tensor(g["x0_mu"], device="cuda")
    vlr = float(g["vgp_lr"])
    velbo = [vg.sweep(vlr, update_x0=it >= 1, x0_lr=vlr).item() for it in range(3)]
    row["err_vdp_elbo"] = relerr(velbo, g["vgp_elbo"])
    row["err_vdp_marginals"] = max(relerr(vg.m[0].cpu().numpy(), g["vgp_m"]), relerr(vg.S[0].cpu().numpy(), g["vgp_S"]))
    vms = device_ms(lambda: vg.sweep(0.01 * vlr))
    row["vdp_ms_per_sweep"] = vms
    row["vdp_ms_per_sweep_cuda_graph"] = device_ms(lambda: vg.sweep_graphed(0.01 * vlr))
    row["vdp_grid_steps_per_s"] = N / (vms * 1e-3)
    # CPU: the NumPy oracle, one process, the same update (all 14 posterior rebuilds of the reference's iteration)
    ref = M.CVISitesSDE(_sde(g), g["time_grid"], (g["obs_t"], g["obs_y"]), M.MultivariateGaussian(lik),
                        prior_initial_state=S.Gaussian(g["x0_mu"], g["x0_cov"]))
    ref.site_update(lr, lr)
    t0 = time.perf_counter()
    reps = 3 if D == 1 else 1
    for _ in range(reps):
        ref.site_update(lr, lr)
    cpu_s = (time.perf_counter() - t0) / reps
    row["numpy_oracle_s_per_update_1core"] = cpu_s
    row["numpy_oracle_grid_steps_per_s"] = N / cpu_s
    out[name] = row
print(json.dumps(out, indent=1))
