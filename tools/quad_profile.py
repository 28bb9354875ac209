#!/usr/bin/env python
"""
Evidence for the D >= 2 quadrature: measured FP64 FMA peak of the device (vdp_test_fp64_fma) and
one CVI-DP site update each on a D = 2 batch (Van der Pol, C3-like: 4096 sequences x 500 grid steps, 400 nodes per
transition) and a D = 3 chain (OU, C5 slice: 2 x 10^5 grid steps, 8000 nodes per transition), timed with CUDA events
and per phase (vdp_profile_*).  Under ncu --set full the same script gives the pipe / stall counters of
k_kl_grad<.., D=2> and <.., D=3>.  Prints one JSON object.
"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vi_diffusion_processes_b200 import _lib  # noqa: E402
from vi_diffusion_processes_b200.cvi import CviEngine  # noqa: E402

DRIFT_VDP, DRIFT_OU_ND = 5, 8
# float64 operations per quadrature node (counted from kl_kernels.cuh)
# D = 3 nested form (kl_quad_nested3, inner loop over the third coordinate): point 2, OU drift 3, residual 12,
# Q_p^-1-weighted residual 18, quadratic form 6 = 41 (value); + adjoint: (I + dt J - A)^T u 36, accumulators 14,
# per-(i0, i1)-pair work / 20 ~ 3 = 94 (gradient).  D = 2 flat loop (Van der Pol): point 8, drift + Jacobian 12,
# A x 6, residual 8, quadratic form 11 = 45 (value); + weighted residual 8, adjoint 16, accumulators 21 = 90 (gradient).
FLOPS = {2: {"grad_node": 90, "value_node": 45, "nodes": 400, "algebra": 2000},
         3: {"grad_node": 94, "value_node": 42, "nodes": 8000, "algebra": 6000}}


def fp64_peak(lib):
    blocks, threads, iters = 148 * 8, 256, 20000
    out = torch.empty(blocks * threads, dtype=torch.float64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    lib.vdp_test_fp64_fma(C.c_void_p(out.data_ptr()), blocks, threads, 100, st)
    torch.cuda.synchronize()
    best = 0.0
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        lib.vdp_test_fp64_fma(C.c_void_p(out.data_ptr()), blocks, threads, iters, st)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, 2.0 * 16 * iters * blocks * threads / (e0.elapsed_time(e1) * 1e-3))
    return best


def timed_update(eng, lib, steps=2):
    eng.site_update(0.5, 0.5)
    eng.check_status()
    torch.cuda.synchronize()
    lib.vdp_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.site_update(0.5, 0.5)
    e1.record()
    torch.cuda.synchronize()
    nph = len(_lib.PHASES)
    ph_ms, ph_n = (C.c_double * nph)(), (C.c_int64 * nph)()
    lib.vdp_profile_read(ph_ms, ph_n, nph)
    lib.vdp_profile_enable(0)
    eng.check_status()
    return e0.elapsed_time(e1) / steps, {nm: ph_ms[i] / max(1, ph_n[i]) for i, nm in enumerate(_lib.PHASES)}


def main():
    lib = _lib.load()
    rng = np.random.default_rng(5)
    peak = fp64_peak(lib)
    out = {"fp64_fma_peak_tflops_measured": peak / 1e12}
    # D = 2: Van der Pol batch
    B, N, n_obs = 4096, 500, 40
    grid = np.round(np.arange(N + 1) * 0.01, 6)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 2))
    e2 = CviEngine(DRIFT_VDP, [1.0, 1.0], np.eye(2), grid, idx, y, 0.1 * np.eye(2), np.zeros(2), 1e-2 * np.eye(2))
    ms, ph = timed_update(e2, lib)
    f = FLOPS[2]
    steps = B * N
    out["d2_vdp_4096x500"] = {"ms_per_update": ms, "grid_steps_per_s": steps / (ms * 1e-3), "phases_ms": ph,
                               "kl_grad_tflops": steps * f["nodes"] * f["grad_node"] / (ph["kl_grad"] * 1e-3) / 1e12,
                               "kl_grad_frac_of_measured_fp64_peak":
                                   steps * f["nodes"] * f["grad_node"] / (ph["kl_grad"] * 1e-3) / peak}
    del e2
    torch.cuda.empty_cache()
    # D = 3: one chain
    N, n_obs = 200_000, 200
    grid = np.arange(N + 1) * 1e-3
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, 3)) * 0.5
    e3 = CviEngine(DRIFT_OU_ND, [0.5], np.eye(3), grid, idx, y, 0.1 * np.eye(3), np.zeros(3), np.eye(3))
    ms, ph = timed_update(e3, lib)
    f = FLOPS[3]
    out["d3_ou_1x200000"] = {"ms_per_update": ms, "grid_steps_per_s": N / (ms * 1e-3), "phases_ms": ph,
                             "kl_grad_tflops": N * f["nodes"] * f["grad_node"] / (ph["kl_grad"] * 1e-3) / 1e12,
                             "kl_grad_frac_of_measured_fp64_peak":
                                 N * f["nodes"] * f["grad_node"] / (ph["kl_grad"] * 1e-3) / peak}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
