#!/usr/bin/env python
"""
bench.py -- grid-steps/s of one CVI-DP site update (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--seqs B] [--grid N]

Workload (config.workload "C4"): B independent double-well sequences (drift 4x(1-x^2), q = 1) on a dense
grid of N = 10^4 steps, dt = 0.005, 50 noisy observations each, float64.  One "step" = one site update
of every sequence = docs/diffusion_processes/cvi_dp_trainer.py:71-77 once (update_data_sites,
update_girsanov_sites, classic_elbo).  Sequences are sharded over ranks with no collective on the data
path; per-GPU work is fixed (B = 4096 per GPU) so scaling is "weak".  Timing state: sites after 3 warm
updates.

value  = grid steps of all ranks / max-over-ranks device time, state resident in HBM.
e2e    = same metric through the host-buffer C ABI with the copies inside the timed region: resident-state session
         (vdp_cvi_session_*): per step the observation values go host->device and the step's result, the ELBO
         terms per sequence, device->host.  e2e_marginals_down = the same call also bringing the posterior marginals
         of every grid point down every step (round 2a's e2e); e2e_full_state_roundtrip = every array of the problem
         both ways every step (vdp_cvi_host_site_update, round 1's e2e); these two come with the box's measured raw
         copy ceiling.
c5     = BASELINE config 5 (one sequence, 10^7 grid steps, D = 3) with the chain sharded in time over the ranks and
         the scan carries / halos exchanged over NCCL (strong scaling); `--workload c5` prints it as its own line.
parity_check = first 4 sequences of the timed state against the CPU oracle; the run exits non-zero above 1e-9.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DT = 0.005
N_OBS = 50
NOISE_STD = 0.1
X0_MU, X0_COV = 1.0, 1e-2
DW_PARAMS = [4.0, 1.0]
LR = 0.5
ALGO_BYTES_PER_STEP_D1 = 80  # 8 * (4D + 6D^2) for D = 1
# float64 operations per grid step of one D = 3 site update (counted from kl_kernels.cuh kl_quad_nested3, see
# tools/quad_profile.py / BASELINE.md section 4): the KL-gradient quadrature (8000 nodes x 94 flops: point, OU drift,
# residual, Q_p^-1-weighted residual, quadratic form, adjoint accumulators) + the KL value quadrature of classic_elbo
# (8000 x 42) + ~6 kflop of block algebra
FP64_PEAK_MEASURED = 33.67e12   # flop/s: vdp_test_fp64_fma on this pool's B200 (profiles/r02_quadrature_d2_d3.json)
C5_FLOPS_PER_STEP = 8000 * 94 + 8000 * 42 + 6000
# D = 3 update, every input and output once: prior A, b, chol Q (21) + sites g1, gd, gs in and out (2 x 21) + q as an
# SSM A, b, chol Q (21) + marginals m, S (12) + time grid (1) = 97 doubles; the 10^4 observations add < 0.1 %
C5_BYTES_PER_STEP = 97 * 8


def make_problem(B, N, seed):
    """Synthetic C4 inputs: Euler-Maruyama double-well paths (sde_utils.py:36-96 restated), 50 obs each."""
    rng = np.random.default_rng(seed)
    grid = np.round(np.arange(N + 1) * DT, 6)
    x = np.empty((B, N + 1))
    x[:, 0] = np.where(rng.random(B) < 0.5, -1.0, 1.0)
    sq = np.sqrt(DT)
    for k in range(N):
        xk = x[:, k]
        x[:, k + 1] = xk + DW_PARAMS[0] * xk * (DW_PARAMS[1] - xk * xk) * DT + sq * rng.standard_normal(B)
    idx = np.sort(np.stack([rng.choice(N + 1, N_OBS, replace=False) for _ in range(B)]), axis=1)
    y = np.take_along_axis(x, idx, axis=1) + NOISE_STD * rng.standard_normal((B, N_OBS))
    return grid, idx, y


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(mx)) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return json.load(open(path)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _oracle_worker(args):
    """One host process: the NumPy oracle on its own sub-batch of sequences (fork-safe, no GPU touched)."""
    B, N, seed, updates = args
    from oracle import scalar_d1
    from oracle.sde_np import DRIFT_DW
    grid, idx, y = make_problem(B, N, seed)
    ref = scalar_d1.BatchedCVISitesSDE1D(DRIFT_DW, DW_PARAMS, 1.0, grid, idx, y, NOISE_STD, X0_MU, X0_COV)
    ref.site_update(LR, LR)  # warm (also moves the sites off their initial values)
    t0 = time.perf_counter()
    for _ in range(updates):
        ref.site_update(LR, LR)
    return time.perf_counter() - t0


def cpu_oracle_steps_per_s(B_per_core, N, seed, updates=1, cores=None):
    """The oracle (NumPy port, oracle/scalar_d1.py) timed on the host cores on a bounded sample of the workload:
    one process per core, B_per_core sequences each (sequences are independent, like the GPU sharding).
    Returns (grid-steps/s over all cores, wall seconds of the slowest worker, cores)."""
    import multiprocessing as mp
    cores = cores or len(os.sched_getaffinity(0))
    jobs = [(B_per_core, N, seed + 1000 * c, updates) for c in range(cores)]
    if cores == 1:
        times = [_oracle_worker(jobs[0])]
    else:
        with mp.get_context("fork").Pool(cores) as pool:
            times = pool.map(_oracle_worker, jobs)
    dt = max(times)
    return B_per_core * cores * N * updates / dt, dt, cores


def cpu_c_steps_per_s(B, N, seed, updates, threads):
    """The C / OpenMP restatement of the oracle (oracle/cvi_d1.c) on `threads` host cores: B sequences x
    `updates` site updates after one warm update.  Returns (grid-steps/s, seconds)."""
    from oracle import c_port
    from oracle.sde_np import DRIFT_DW
    grid, idx, y = make_problem(B, N, seed)
    ref = c_port.CBatchedCVISitesSDE1D(DRIFT_DW, DW_PARAMS, 1.0, grid, idx, y, NOISE_STD, X0_MU, X0_COV,
                                       threads=threads)
    ref.site_update(LR, LR)
    t0 = time.perf_counter()
    for _ in range(updates):
        ref.site_update(LR, LR)
    dt = time.perf_counter() - t0
    return B * N * updates / dt, dt


# algorithmic (compulsory) float64 values moved per grid step by each phase's kernels: (vectors, matrices)
# read + written once; bytes = 8 * (v * D + M * D^2).  Stated per kernel in DESIGN.md section 4.
PHASE_TRAFFIC = {
    "data_sites": (0, 0),
    # consecutive updates carry the naturals of the previous update in the workspace and rewrite them at the observed
    # grid points only (VDP_FLAG_NATURALS_CARRIED): O(n_obs) per sequence instead of a pass over the grid
    "sites_to_nat": (0, 0),
    # read thd, ths ; write Ld, Ls.  Once per update: the factor of the second rebuild comes out of the fused forward
    # pass (phase kl_grad)
    "chol_fwd": (0, 2 + 2),
    # read th1 | Ld, Ls ; write b | A, cholQ, Sd (K blocks rebuilt from L in the tile; the marginal covariances are
    # the diagonal of the inverse band the sweep walks through, so no separate inverse-band sweep)
    # (the rebuild between the two site updates does not write chol Q: 48 and 56 B, 52 on average)
    "nat2ssm_bwd": (1 + 1, 2 + 2.5),
    # mean sweep of the second rebuild: read b | A, S, cholQ ; write m -- it adds up the KL terms of classic_elbo on
    # its way.  (The means of the first rebuild never leave the fused forward pass.)
    "marg_fwd": (1 + 1, 1 + 1 + 1),
    "invband_bwd": (0, 2 + 1),        # (only with VDP_FLAG_REFACTOR_PRECISION) read Ld, Ls ; write Sd
    # fused forward pass (csrc/fwd_fused.cuh: means of the first rebuild -> KL gradient -> girsanov update of theta ->
    # Cholesky factor of the second rebuild): read b, th1 | A, S, thd, ths ; write th1 | thd, ths, Ld, Ls.  The sites
    # are held in theta between carried updates (VDP_FLAG_SITES_IN_NATURALS); the data-site part of the step is added
    # at the observed grid points by a per-observation kernel
    "kl_grad": (1 + 1 + 1, 1 + 1 + 2 + 2 + 2),
    "elbo": (0, 0),                   # variational expectations at the observations + sum of the KL partials
}
LAUNCHES_PER_UPDATE = 3 + 2 + 2 + 3       # kernels of ours in one vdp_cvi_site_update in steady state (data sites +
#                                           patch, Cholesky, naturals_to_ssm; data step + fused forward pass;
#                                           naturals_to_ssm + means; VE + reduce + combine)


def phase_bytes_per_step(name, D):
    v, m = PHASE_TRAFFIC[name]
    return 8 * (v * D + m * D * D)


def parity_check(eng, grid, idx, y, n_updates, n_seq=4):
    """First `n_seq` sequences of the resident engine against the C port of the oracle after the same `n_updates`
    site updates (checker only: never inside a timed region)."""
    from oracle import c_port
    from oracle.sde_np import DRIFT_DW
    n_seq = min(n_seq, idx.shape[0])
    ref = c_port.CBatchedCVISitesSDE1D(DRIFT_DW, DW_PARAMS, 1.0, grid, idx[:n_seq], y[:n_seq], NOISE_STD, X0_MU, X0_COV,
                                       threads=min(n_seq, len(os.sched_getaffinity(0))))
    for _ in range(n_updates):
        elbo = ref.site_update(LR, LR)

    def rel(a, b):
        a = a.detach().cpu().numpy()
        return float(np.abs(a - b).max() / np.abs(b).max())
    out = {"sequences": n_seq, "updates": n_updates, "checker": "oracle/cvi_d1.c",
           "max_rel_elbo": rel(eng.elbo[:n_seq], elbo),
           "max_rel_marginal_mean": rel(eng.fx_mus[:n_seq, :, 0], ref.fx_mus),
           "max_rel_marginal_cov": rel(eng.fx_covs[:n_seq, :, 0, 0], ref.fx_covs),
           "max_rel_sites_unscaled": {"g1": rel(eng.g1[:n_seq, :, 0], ref.g1), "gd": rel(eng.gd[:n_seq, :, 0, 0], ref.gd),
                                      "gs": rel(eng.gs[:n_seq, :, 0, 0], ref.gs)},
           "tolerance": 1e-9}
    out["ok"] = bool(max(out["max_rel_elbo"], out["max_rel_marginal_mean"], out["max_rel_marginal_cov"]) < 1e-9)
    return out


def copy_ceiling_ms(h2d_bytes, d2h_bytes, device, barrier, world, dist):
    import torch
    up_h = torch.empty(h2d_bytes // 8, dtype=torch.float64).pin_memory()
    up_d = torch.empty(h2d_bytes // 8, dtype=torch.float64, device=device)
    dn_d = torch.zeros(d2h_bytes // 8, dtype=torch.float64, device=device)
    dn_h = torch.empty(d2h_bytes // 8, dtype=torch.float64).pin_memory()
    s1, s2 = torch.cuda.Stream(device), torch.cuda.Stream(device)
    best = float("inf")
    for _ in range(4):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_stream(torch.cuda.current_stream())
        s2.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s1):
            up_d.copy_(up_h, non_blocking=True)
        with torch.cuda.stream(s2):
            dn_h.copy_(dn_d, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        best = min(best, float(ms.item()))
    return best


def run_ours(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from vi_diffusion_processes_b200.markovflow.sde.sde import DRIFT_DW
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.cvi import CviEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    B, N, D = args.seqs, args.grid, 1
    grid, idx, y = make_problem(B, N, seed=4096 + rank)
    eng = CviEngine(DRIFT_DW, DW_PARAMS, np.ones((1, 1)), grid, idx, y[..., None], [[NOISE_STD]], [X0_MU], [[X0_COV]],
                    device=device)
    lib = _lib.load()
    for _ in range(3):                      # timing state: sites after 3 warm updates
        eng.site_update(LR, LR)
    eng.check_status()

    # end-to-end leg: the whole state in pinned HOST memory, through vdp_cvi_host_site_update (C ABI with host
    # buffers): per step, upload state + prior + observations, update, download sites + marginals + ELBO
    from vi_diffusion_processes_b200.cvi import CviHostEngine
    if not args.quick:
        host_eng = CviHostEngine(eng, chunk=args.chunk, slots=3, want_posterior_ssm=False)
        h2d, d2h = host_eng.bytes_per_update()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def step_resident():
        eng.site_update(LR, LR)

    def step_e2e():
        host_eng.site_update(LR, LR)        # returns when the results are in the host buffers

    def step_session():
        session.site_update(LR, LR, marginals=False)   # obs values up, update, ELBO terms down, returns when done

    def step_session_marginals():
        session.site_update(LR, LR)         # ... and the posterior marginals of every grid point down as well

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t_load = time.perf_counter()
    ms_total = timed(step_resident, args.steps, args.warmup)
    # self-check of the state the timed updates produced (outside the timed region): the first 4 sequences after
    # the same number of updates on the CPU checker (oracle/cvi_d1.c).  Sequences are independent, so a 4-sequence
    # CPU run reproduces them.  Bar: 1e-9 relative on ELBO and marginals; a miss makes the run exit non-zero.
    parity = None
    if rank == 0 and not args.no_check:
        parity = parity_check(eng, grid, idx, y, 3 + args.warmup + args.steps)
    # per-phase device times from a second pass with the library's phase timer on (it serialises the update on one
    # stream, so it is kept out of the timed region above)
    lib.vdp_profile_enable(1)
    timed(step_resident, args.steps, 1)
    nph = len(_lib.PHASES)
    ph_ms = (C.c_double * nph)()
    ph_n = (C.c_int64 * nph)()
    lib.vdp_profile_read(ph_ms, ph_n, nph)    # covers warm-up + timed steps: averages per launch below
    lib.vdp_profile_enable(0)
    # the same update with the 20-point Gauss-Hermite loops instead of their closed form (polynomial drift), and with
    # the naturals recomputed over the whole grid every update: reported beside the headline, not part of it
    eng.route_flags = _lib.FLAG_GH_QUADRATURE          # per-call route flag of vdp_cvi_problem (no environment)
    ms_gh = timed(step_resident, args.steps, 1)
    eng.carry_naturals = False
    ms_gh_full = timed(step_resident, args.steps, 1)
    eng.route_flags, eng.carry_naturals = 0, True
    # a drift that is not a polynomial (Benes, f = theta tanh(x)): no closed form and no fused forward pass, the
    # 20-point loops and the three-kernel route are all there is -- what 4 of the reference's 7 drifts get
    ms_benes = None
    if not args.quick:
        from vi_diffusion_processes_b200.markovflow.sde.sde import DRIFT_BENES
        eng_b = CviEngine(DRIFT_BENES, [1.0], np.ones((1, 1)), grid, idx, y[..., None], [[NOISE_STD]], [X0_MU],
                          [[X0_COV]], device=device)
        ms_benes = timed(lambda: eng_b.site_update(0.1, 0.1), args.steps, 2)   # (same kernels at any step size; the
        try:                                                                   # bench data come from a double well)
            eng_b.check_status()
            benes_ok = True
        except _lib.VdpError:
            benes_ok = False
        del eng_b
        torch.cuda.empty_cache()
    # nvidia-smi samples every 100 ms: keep the same load running (untimed) until at least ~0.6 s have been sampled
    while time.perf_counter() - t_load < 0.6:
        step_resident()
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    eng.check_status()
    if args.quick:
        if rank == 0:
            print(json.dumps({"parity_check": parity, "ms_per_step": ms_total / args.steps, "ms_per_step_gh_loops": ms_gh / args.steps,
                              "ms_per_step_gh_loops_no_carry": ms_gh_full / args.steps, "phases": {
                nm: {"ms_per_launch": ph_ms[i] / max(1, ph_n[i])} for i, nm in enumerate(_lib.PHASES)}}))
        if world > 1:
            dist.destroy_process_group()
        return
    e2e_steps = max(2, args.steps // 2)
    # (1) resident-state session (vdp_cvi_session_*): the state was uploaded once and lives on the device, as the
    # reference's model object keeps its variables between trainer iterations; every step moves that step's inputs
    # (the observation values, pinned host memory) up and its result (the ELBO terms: ve, kl, elbo per sequence -- the
    # loss the reference's trainer reads every iteration, cvi_dp_trainer.py:75; NLPD / RMSE there are two more scalars
    # from marginals that stay with the model) down.  This is the `e2e` of the line.  `e2e_marginals_down` is the same
    # call that also brings the posterior marginals of every grid point to the host every step (the `e2e` of round 2a),
    # with the raw copy ceiling of its bytes.
    from vi_diffusion_processes_b200.cvi import CviSession
    # (with only the ELBO terms coming down there is no download to overlap with the next chunk's kernels: one chunk
    # of all sequences keeps the launches full; the marginal download is pipelined over chunks of --session-chunk)
    session = CviSession(eng, chunk=B, streams=1)
    s_h2d, s_d2h = session.bytes_per_update(marginals=False)
    ms_session = timed(step_session, args.steps, 2)
    session.close()
    del session
    session = CviSession(eng, chunk=args.session_chunk, streams=3)
    sm_h2d, sm_d2h = session.bytes_per_update(marginals=True)
    ms_session_m = timed(step_session_marginals, args.steps, 2)
    ms_session_ceiling = copy_ceiling_ms(sm_h2d, sm_d2h, device, barrier, world, dist)
    session.close()
    del session
    # (2) the whole state in host memory, every array both ways every step (vdp_cvi_host_site_update): the round-1
    # end-to-end figure, kept for comparison under "e2e_full_state_roundtrip"
    ms_e2e = timed(step_e2e, e2e_steps, 1)
    # what the box allows: the same bytes as raw pinned cudaMemcpyAsync, upload and download on two streams at once,
    # all ranks at the same time (max over ranks) -- the ceiling of any implementation of the host-buffer call
    ms_ceiling = copy_ceiling_ms(h2d, d2h, device, barrier, world, dist)
    host_eng.resident_prior = True          # variant: linearised prior uploaded once (it only changes at a
    host_eng.set_prior()                    # re-linearisation), state still travels every step
    h2d_rp, _ = host_eng.bytes_per_update()
    ms_e2e_rp = timed(step_e2e, e2e_steps, 1)
    host_eng.close()

    # second workload of BASELINE.json (config 5): ONE sequence of 10^7 grid steps, D = 3, the chain sharded in TIME
    # over the ranks with the scan carries / halos exchanged over NCCL (strong scaling).  Reported under "c5" of the
    # same line so that the scaling run records a workload with a collective on its data path.
    c5 = None
    if not args.no_c5:
        del host_eng
        torch.cuda.empty_cache()
        c5 = run_c5_leg(args)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    steps_total = B * N * world
    value = steps_total * args.steps / (ms_total * 1e-3)
    e2e_value = steps_total * e2e_steps / (ms_e2e * 1e-3)
    session_value = steps_total * args.steps / (ms_session * 1e-3)
    phases = {nm: {"ms_per_launch": ph_ms[i] / max(1, ph_n[i]), "launches": int(ph_n[i]), "ms_total": ph_ms[i]}
              for i, nm in enumerate(_lib.PHASES)}
    dom = max(phases, key=lambda k: phases[k]["ms_total"])
    peak, peak_src = measured_peaks()
    dom_bytes = phase_bytes_per_step(dom, D) * B * N
    achieved = dom_bytes / (phases[dom]["ms_per_launch"] * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
    if os.path.exists(tpath):
        t = json.load(open(tpath))
        if t.get("B") == B and t.get("N") == N:
            traffic = t.get("dram_bytes_per_launch_by_phase", {}).get(dom)
    cpu_cores = len(os.sched_getaffinity(0))
    cpu_B = 64 * cpu_cores
    cpu_val, cpu_s = cpu_c_steps_per_s(cpu_B, N, seed=4096, updates=10, threads=cpu_cores)
    cpu_1, cpu_1s = cpu_c_steps_per_s(16, N, seed=4096, updates=10, threads=1)
    np_val, np_s, _ = cpu_oracle_steps_per_s(16, N, seed=4096, updates=1, cores=1)
    line = {
        "metric": "grid-steps/sec per CVI-DP site update", "value": value, "unit": "grid-steps/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C4: independent double-well sequences x dense grid, one CVI-DP site update per step",
                   "sequences_per_gpu": B, "grid_steps": N, "state_dim": D, "dt": DT, "n_obs": N_OBS, "quad_H": 20,
                   "quadrature": "double-well drift is a cubic: the 20-point Gauss-Hermite sums are taken in closed "
                                 "form (central Gaussian moments, equal to rounding); variants has the loops",
                   "parallelism": f"sequence-sharded x{world}, no data-path collective",
                   "l2": "working set per step (~%.1f GB) exceeds the 126 MB L2" % (B * N * 8 * 30 / 1e9)},
        "clocks": clocks,
        "variants": {"gh20_loops": {"ms_per_step": ms_gh / args.steps,
                                    "value": steps_total * args.steps / (ms_gh * 1e-3)},
                     "benes_drift_gh20_loops": {"ms_per_step": ms_benes / args.steps,
                                                "value": steps_total * args.steps / (ms_benes * 1e-3),
                                                "status_ok": benes_ok, "step_sizes": 0.1,
                                                "note": "f = tanh(x): not a polynomial, so neither the closed-form "
                                                        "moments nor the fused forward pass apply"},
                     "gh20_loops_naturals_recomputed_every_update": {
                         "ms_per_step": ms_gh_full / args.steps,
                         "value": steps_total * args.steps / (ms_gh_full * 1e-3)}},
        "e2e": {"value": session_value, "unit": "grid-steps/s", "h2d_bytes_per_step": s_h2d,
                "d2h_bytes_per_step": s_d2h, "ms_per_step": ms_session / args.steps,
                "api": "vdp_cvi_session_site_update (C ABI, host buffers): state resident on the device between calls "
                       "(uploaded once by vdp_cvi_session_create, like the reference's model variables); per step the "
                       "observation values [B, n_obs] go up from pinned host memory and the step's result, the ELBO "
                       "terms ve, kl, elbo [B], comes down to pinned host memory (what cvi_dp_trainer.py:75 reads "
                       "every iteration); the call returns when they are there; one chunk of all %d sequences" % B},
        "e2e_marginals_down": {
            "value": steps_total * args.steps / (ms_session_m * 1e-3), "unit": "grid-steps/s",
            "h2d_bytes_per_step": sm_h2d, "d2h_bytes_per_step": sm_d2h, "ms_per_step": ms_session_m / args.steps,
            "api": "the same call with fx_mus_host / fx_covs_host given: the posterior marginals of every grid point "
                   "[B, N+1] also come down every step (this was `e2e` in round 2a); %d-sequence chunks on 3 streams"
                   % min(args.session_chunk, B),
            "copy_ceiling_ms": ms_session_ceiling, "frac_of_copy_ceiling": ms_session_ceiling / (ms_session_m / args.steps),
            "copy_ceiling": "raw pinned cudaMemcpyAsync of the same h2d + d2h bytes on two streams at once, all "
                            "ranks concurrently, max over ranks, best of 4"},
        "e2e_full_state_roundtrip": {
            "value": e2e_value, "unit": "grid-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "ms_per_step": ms_e2e / e2e_steps,
            "api": "vdp_cvi_host_site_update (C ABI, pinned host buffers, %d-sequence chunks on 3 streams): EVERY array "
                   "of the problem (sites, prior, data sites, marginals) travels host -> device -> host every step; "
                   "this was `e2e` in round 1" % min(args.chunk, B),
            "copy_ceiling_ms": ms_ceiling, "frac_of_copy_ceiling": ms_ceiling / (ms_e2e / e2e_steps)},
        "e2e_prior_resident": {"value": steps_total * e2e_steps / (ms_e2e_rp * 1e-3), "unit": "grid-steps/s",
                               "h2d_bytes_per_step": h2d_rp, "d2h_bytes_per_step": d2h,
                               "ms_per_step": ms_e2e_rp / e2e_steps},
        "gpu_launches": LAUNCHES_PER_UPDATE * args.steps,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": dom_bytes,
                     "kernel_share_of_step": phases[dom]["ms_total"] / sum(p["ms_total"] for p in phases.values())},
        "roofline_update": {"algorithmic_bytes_per_grid_step": ALGO_BYTES_PER_STEP_D1,
                            "achieved": value * ALGO_BYTES_PER_STEP_D1 / world / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": value * ALGO_BYTES_PER_STEP_D1 / world / 1e9 / peak},
        "phases": phases,
        "parity_check": parity,
        "c5": c5,
        "cpu_baseline": {"value": cpu_val, "unit": "grid-steps/s", "cores": cpu_cores, "kind": "port",
                         "sample": f"C/OpenMP restatement of the oracle (oracle/cvi_d1.c, gcc -O2), {cpu_cores} "
                                   f"threads x 64 sequences x {N} grid steps x 10 site updates, {cpu_s:.1f} s",
                         "single_thread": {"value": cpu_1, "sample": f"16 sequences x 10 updates, {cpu_1s:.1f} s"},
                         "numpy_oracle_1_process": {"value": np_val,
                                                    "sample": f"oracle/scalar_d1.py, 16 sequences x 1 update, "
                                                              f"{np_s:.1f} s"}},
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit("bench.py: parity check against the CPU oracle failed: %s" % json.dumps(parity))


def run_c5_leg(args):
    """BASELINE config 5 through tools/c5_time_sharded.run (time-sharded chain, dist_cvi.TimeShardedCvi)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import c5_time_sharded
    # default route: the drift is linear, so the 2-point rule per dimension is exact (csrc/drifts.cuh DriftPoly) and the
    # update is bound by the sweeps' HBM traffic; then the same update with the reference's 20 points per dimension
    # (8000 nodes per transition), which is bound by the float64 pipe
    res = c5_time_sharded.run(N=args.c5_grid, n_obs=args.c5_obs, steps=args.c5_steps, warmup=1)
    gh = c5_time_sharded.run(N=args.c5_grid, n_obs=args.c5_obs, steps=1, warmup=1, gh20=True)
    if res is not None:
        peak_hbm, peak_src = measured_peaks()
        bytes_step = C5_BYTES_PER_STEP
        gbs = bytes_step * res["value"] / 1e9 / res["n_gpus"]
        res["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": peak_hbm, "unit": "GB/s", "frac": gbs / peak_hbm,
                           "algorithmic_bytes_per_grid_step": bytes_step, "peak_source": peak_src,
                           "note": "inputs and outputs of one update once: prior SSM, sites and data in; sites, q as an "
                                   "SSM and marginals out (D = 3: 97 doubles per grid step)"}
        peak_fp64 = FP64_PEAK_MEASURED   # vdp_test_fp64_fma on this pool's B200 (tools/quad_profile.py)
        flops = C5_FLOPS_PER_STEP * gh["value"]
        res["gh20_reference_order"] = {
            "ms_per_step": gh["ms_per_step"], "value": gh["value"], "unit": "grid-steps/s", "elbo": gh["elbo"],
            "quadrature_nodes_per_transition": 8000,
            "roofline": {"bound": "fp64", "achieved": flops / 1e12 / gh["n_gpus"], "peak": peak_fp64 / 1e12,
                         "unit": "TFLOP/s", "frac": flops / gh["n_gpus"] / peak_fp64,
                         "flops_per_grid_step": C5_FLOPS_PER_STEP,
                         "peak_source": "measured float64 FMA throughput of this pool's B200 (vdp_test_fp64_fma, "
                                        "profiles/r02_quadrature_d2_d3.json)"}}
    return res


def run_c5(args):
    """`--workload c5`: the C5 leg alone, as a bench line of its own."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    sampler = ClockSampler(local_rank)
    if int(os.environ.get("RANK", "0")) == 0:
        sampler.start()
    res = run_c5_leg(args)
    if res is not None:
        line = {"metric": "grid-steps/sec per CVI-DP site update", "value": res["value"], "unit": "grid-steps/s",
                "n_gpus": world, "steps": res["steps"], "warmup": res["warmup"], "ms_per_step": res["ms_per_step"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": {"workload": res["workload"], "grid_steps": args.c5_grid,
                                                "state_dim": 3, "n_obs": args.c5_obs,
                                                "quad_H": "2 per dimension (exact for the linear drift); "
                                                          "gh20_reference_order: 20",
                                                "parallelism": f"time-sharded x{world}: {res['collective']}"},
                "clocks": sampler.stop(), "roofline": res["roofline"], "c5": res}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_vdp(args):
    """Second line of SURVEY 8(d): one VDP fixed-point sweep (vi_markov_gp_trainer.py:62-73) on the C4 shape."""
    import torch
    from vi_diffusion_processes_b200.markovflow.sde.sde import DRIFT_DW
    from vi_diffusion_processes_b200.vgp import VgpEngine
    torch.cuda.set_device(0)
    B, N = args.seqs, args.grid
    grid, idx, y = make_problem(B, N, seed=4096)
    idx = np.maximum(idx, 1)
    eng = VgpEngine(DRIFT_DW, DW_PARAMS, np.ones((1, 1)), grid, idx, y[..., None], [[NOISE_STD]], [X0_MU], [[X0_COV]],
                    stabilize_system=True)
    eng.A += 2.0
    eng.b += 2.0 * torch.as_tensor(np.sign(y[:, :1]), device="cuda").reshape(-1, 1, 1)
    for _ in range(max(3, args.warmup)):
        eng.sweep(0.01)
    eng.check_status()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        eng.sweep(0.01)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    peak, peak_src = measured_peaks()
    bytes_per_step = 8 * (2 + 2 + 2 + 2)   # read A, b, write A, b, lam, psi, m, S (D = 1)
    print(json.dumps({"metric": "grid-steps/sec per VDP fixed-point sweep", "value": B * N / (ms * 1e-3),
                      "unit": "grid-steps/s", "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
                      "ms_per_step": ms, "higher_is_better": True, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": "C4 shape, VDP: forward pass, Lagrange sweep, parameter update, forward "
                                             "pass, ELBO", "sequences_per_gpu": B, "grid_steps": N, "state_dim": 1},
                      "roofline_update": {"algorithmic_bytes_per_grid_step": bytes_per_step,
                                          "achieved": B * N * bytes_per_step / (ms * 1e-3) / 1e9, "peak": peak,
                                          "unit": "GB/s", "frac": B * N * bytes_per_step / (ms * 1e-3) / 1e9 / peak,
                                          "peak_source": peak_src}}))


def run_reference(args):
    """The reference's CPU path = the C/OpenMP port of the oracle on all host cores (TensorFlow is not installable
    here)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps)
    cores = len(os.sched_getaffinity(0))
    B = args.ref_seqs * cores
    value, dt = cpu_c_steps_per_s(B, args.grid, seed=4096, updates=steps, threads=cores)
    line = {
        "impl": "reference", "metric": "grid-steps/sec per CVI-DP site update", "value": value,
        "unit": "grid-steps/s", "n_gpus": args.gpus, "steps": steps, "warmup": 1, "ms_per_step": dt / steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C4: independent double-well sequences x dense grid, one CVI-DP site update per step",
                   "sequences": B, "grid_steps": args.grid, "state_dim": 1, "dt": DT, "n_obs": N_OBS,
                   "quad_H": 20},
        "cpu_baseline": {"value": value, "unit": "grid-steps/s", "cores": cores, "kind": "port",
                         "sample": f"C/OpenMP restatement of the oracle (oracle/cvi_d1.c), {cores} threads, {B} sequences "
                                   f"x {args.grid} grid steps x {steps} site updates (bounded sample of the "
                                   f"4096-sequence workload), {dt:.1f} s"},
        "e2e": {"value": value, "unit": "grid-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cvi", choices=["cvi", "vdp", "c5"],
                    help="cvi = the BASELINE.json metric on config 4 with config 5 under the key c5 (default); vdp = the "
                         "VDP sweep on the config-4 shape (1 GPU); c5 = config 5 alone (time-sharded over the ranks)")
    ap.add_argument("--c5-grid", type=int, default=10_000_000, help="grid steps of the config-5 sequence")
    ap.add_argument("--c5-obs", type=int, default=10_000)
    ap.add_argument("--c5-steps", type=int, default=2)
    ap.add_argument("--no-c5", action="store_true", help="skip the config-5 leg of the default run")
    ap.add_argument("--seqs", type=int, default=4096, help="sequences per GPU")
    ap.add_argument("--grid", type=int, default=10000, help="grid steps per sequence")
    ap.add_argument("--quick", action="store_true", help="tuning runs: resident timing and phases only")
    ap.add_argument("--no-check", action="store_true", help="skip the parity self-check against the CPU oracle")
    ap.add_argument("--chunk", type=int, default=256, help="sequences per chunk of the host-buffer pipeline")
    ap.add_argument("--session-chunk", type=int, default=1024, help="sequences per chunk of the resident-state session")
    ap.add_argument("--ref-seqs", type=int, default=64, help="sequences per host core in the bounded CPU sample")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "vdp":
        run_vdp(args)
    elif args.workload == "c5":
        run_c5(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
