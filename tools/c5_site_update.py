#!/usr/bin/env python
"""
BASELINE config 5 on one GPU: ONE sequence, N grid steps (default 10^7), D = 3 (3-D OU prior by default,
--drift lorenz for Lorenz-63), full CVI-DP site update (update_data_sites, update_girsanov_sites with the
H = 20 quadrature = 8000 nodes per transition, classic_elbo).  Prints one JSON line.
"""
import argparse
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

DRIFT_LORENZ63, DRIFT_OU_ND = 7, 8


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=10_000_000)
    ap.add_argument("--obs", type=int, default=10_000)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--drift", default="ou", choices=["ou", "lorenz"])
    a = ap.parse_args()
    N, D, dt = a.grid, 3, 1e-3
    rng = np.random.default_rng(7)
    grid = np.arange(N + 1) * dt
    idx = np.sort(rng.choice(N + 1, a.obs, replace=False))
    if a.drift == "ou":
        drift, params = DRIFT_OU_ND, [0.5]
        y = rng.normal(size=(a.obs, D)) * 0.5
    else:
        drift, params = DRIFT_LORENZ63, [10.0, 28.0, 8.0 / 3.0]
        y = rng.normal(size=(a.obs, D)) * 2.0 + np.array([1.0, 1.0, 20.0])
    eng = CviEngine(drift, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), y[0] if a.drift == "lorenz" else
                    np.zeros(D), np.eye(D))
    lib = _lib.load()
    eng.site_update(0.5, 0.5)          # warm (moves the sites off their initial values)
    eng.check_status()
    torch.cuda.synchronize()
    lib.vdp_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        elbo = eng.site_update(0.5, 0.5)
    e1.record()
    torch.cuda.synchronize()
    nph = len(_lib.PHASES)
    ph_ms, ph_n = (C.c_double * nph)(), (C.c_int64 * nph)()
    lib.vdp_profile_read(ph_ms, ph_n, nph)
    lib.vdp_profile_enable(0)
    eng.check_status()
    ms = e0.elapsed_time(e1) / a.steps
    print(json.dumps({"workload": f"C5: 1 sequence x {N} grid steps, D=3, {a.drift} prior, {a.obs} observations, "
                                  "one CVI-DP site update", "n_gpus": 1, "ms_per_update": ms,
                      "grid_steps_per_s": N / (ms * 1e-3), "elbo": float(elbo[0]),
                      "quadrature_nodes_per_transition": 8000,
                      "phases_ms_per_launch": {nm: ph_ms[i] / max(1, ph_n[i]) for i, nm in enumerate(_lib.PHASES)},
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}))


if __name__ == "__main__":
    main()
