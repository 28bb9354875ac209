"""
Golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py from the data files the reference ships):
  * CPU: the oracle still reproduces them (regression guard for the checker itself);
  * GPU: the CUDA path, through the C ABI, reproduces them to 1e-9 without running the oracle.
"""
import glob
import os

import numpy as np
import pytest

from oracle import models_np as M
from oracle import sde_np as S

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))
TOL = 1e-9


def _sde(g):
    did, p, q = int(g["drift_id"]), g["drift_params"], g["q"]
    if did == S.DRIFT_OU:
        return S.OrnsteinUhlenbeckSDE(p[0], q)
    if did == S.DRIFT_DW:
        return S.DoubleWellSDE(q, p[0], p[1])
    if did == S.DRIFT_VDP:
        return S.VanderPolOscillatorSDE(p[0], p[1], q)
    raise NotImplementedError(did)


def relerr(a, b, scale=None):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    den = np.abs(b).max() if scale is None else scale
    return float(np.abs(a - b).max() / (den if den > 0 else 1.0))


def test_fixtures_exist():
    assert len(GOLDEN) >= 3, "run tests/golden/make_golden.py"


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_reproduces_golden_cvi(path):
    g = np.load(path)
    D = g["obs_y"].shape[-1]
    cvi = M.CVISitesSDE(_sde(g), g["time_grid"], (g["obs_t"], g["obs_y"]), M.MultivariateGaussian(float(g["sigma"]) * np.eye(D)),
                        prior_initial_state=S.Gaussian(g["x0_mu"], g["x0_cov"]))
    lr = float(g["lr"])
    elbos = []
    for it in range(3):
        elbos.append(cvi.site_update(lr, lr))
        if it == 1:
            cvi.relinearize()
    np.testing.assert_allclose(elbos, g["cvi_elbo"], rtol=1e-10)
    np.testing.assert_allclose(cvi.fx_mus, g["cvi_m"], rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(cvi.girsanov_sites.nat2.block_diagonal, g["cvi_gd"], rtol=1e-7, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_reproduces_golden_cvi(path):
    torch = pytest.importorskip("torch")
    from vi_diffusion_processes_b200.cvi import CviEngine
    g = np.load(path)
    D = g["obs_y"].shape[-1]
    eng = CviEngine(int(g["drift_id"]), g["drift_params"], g["q"], g["time_grid"], g["obs_idx"], g["obs_y"],
                    float(g["sigma"]) * np.eye(D), g["x0_mu"], g["x0_cov"])
    lr = float(g["lr"])
    elbos = []
    for it in range(3):
        elbos.append(eng.site_update(lr, lr).item())
        if it == 1:
            eng.relinearize()
    eng.check_status()
    assert relerr(elbos, g["cvi_elbo"]) < TOL
    assert relerr(eng.fx_mus[0].cpu().numpy(), g["cvi_m"]) < TOL
    assert relerr(eng.fx_covs[0].cpu().numpy(), g["cvi_S"]) < TOL
    assert relerr(eng.d1[0].cpu().numpy(), g["cvi_d1"]) < TOL
    # site tensors: at the scale of the posterior natural parameters they are added to (OU prior: exact sites are 0)
    # the fixture records how far the float64 reference route itself is from its 80-bit evaluation on this data
    # (site_floor_f64_vs_f80, e.g. 1.4e-9 for g1 on c2): no float64 implementation can be pinned tighter than that
    th = [t[0].cpu().numpy() for t in eng._prior_naturals(eng.pA, eng.pb)]
    for got, want, t, floor in ((eng.g1, g["cvi_g1"], th[0], g["site_floor_f64_vs_f80"][0]),
                                (eng.gd, g["cvi_gd"], th[1], g["site_floor_f64_vs_f80"][1]),
                                (eng.gs, g["cvi_gs"], th[2], g["site_floor_f64_vs_f80"][2])):
        scale = max(np.abs(want).max(), 1e-3 * np.abs(t + want).max())
        floor = floor if floor < 1e-6 else 0.0   # OU prior: the exact sites are 0, a relative floor means nothing
        assert relerr(got[0].cpu().numpy(), want, scale=scale) < max(TOL, 3.0 * floor)
        # as a summand of q's natural parameters the site is right to 1e-11
        assert relerr(got[0].cpu().numpy(), want, scale=np.abs(t + want).max()) < 1e-11


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_reproduces_golden_vgp(path):
    torch = pytest.importorskip("torch")
    from vi_diffusion_processes_b200.vgp import VgpEngine
    g = np.load(path)
    D = g["obs_y"].shape[-1]
    eng = VgpEngine(int(g["drift_id"]), g["drift_params"], g["q"], g["time_grid"], g["obs_idx"], g["obs_y"],
                    float(g["sigma"]) * np.eye(D), g["x0_mu"], g["x0_cov"], stabilize_system=bool(g["vgp_stabilize"]))
    eng.A += 2.0 * torch.eye(D, dtype=torch.float64, device="cuda")
    eng.b += 2.0 * torch.as_tensor(g["x0_mu"], device="cuda")
    lr = float(g["vgp_lr"])
    elbos = [eng.sweep(lr, update_x0=it >= 1, x0_lr=lr).item() for it in range(3)]
    eng.check_status()
    assert relerr(elbos, g["vgp_elbo"]) < TOL
    for got, want in ((eng.A, g["vgp_A"]), (eng.b, g["vgp_b"]), (eng.lambda_lagrange, g["vgp_lam"]),
                      (eng.psi_lagrange, g["vgp_psi"]), (eng.m, g["vgp_m"]), (eng.S, g["vgp_S"]),
                      (eng.q_initial_mean, g["vgp_q_mean"]), (eng.q_initial_chol, g["vgp_q_chol"])):
        assert relerr(got[0].cpu().numpy(), want) < TOL
