"""
Polynomial drifts in D >= 2: the KL term along the Gaussian path and its gradient through the exact low-order
Gauss-Hermite rule (degree p drift -> p + 1 points per dimension: csrc/drifts.cuh DriftPoly / ExactOrder) against the
reference's 20 points per dimension (sde_utils.py:332-354; VDP_FLAG_GH_QUADRATURE keeps them).  Both rules integrate
every integrand of the term and of its adjoint exactly, so site updates, marginals and ELBO agree to rounding:

    Van der Pol   D = 2, cubic      16 nodes instead of 400
    double well   D = 2, cubic      16 / 400    (the scalar drift applied per coordinate)
    Lorenz-63     D = 3, quadratic  27 / 8000
    OU (matrix)   D = 3, linear      8 / 8000

The oracle comparison of these drifts (tests/test_gpu_parity.py::test_site_update_vdp_2d, the D = 3 tests of
tests/test_gpu_dist_cvi.py, the golden fixtures) runs on the default route, i.e. on the low-order rule.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _case(name):
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    if name == "vdp":
        return sde_mod.DRIFT_VDP, [1.0, 1.0], 2, 400, 0.01
    if name == "dw2":
        return sde_mod.DRIFT_DW, [2.0, 1.0], 2, 300, 0.01
    if name == "lorenz":
        return sde_mod.DRIFT_LORENZ63, [10.0, 28.0, 8.0 / 3.0], 3, 150, 0.002
    if name == "ou3":
        return sde_mod.DRIFT_OU_ND, [0.5], 3, 250, 0.01
    raise KeyError(name)


@pytest.mark.parametrize("name", ["vdp", "dw2", "lorenz", "ou3"])
def test_exact_low_order_rule_equals_twenty_points(name):
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.cvi import CviEngine
    did, params, D, N, dt = _case(name)
    rng = np.random.default_rng(5 + D)
    B, n_obs = 3, 9
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = 1.0 + 0.3 * rng.normal(size=(B, n_obs, D))
    x0 = np.ones(D)

    def make(flags):
        eng = CviEngine(did, params, np.eye(D), grid, idx, y, 0.2 * np.eye(D), x0, 0.05 * np.eye(D))
        eng.route_flags = flags
        return eng

    a, c = make(0), make(_lib.FLAG_GH_QUADRATURE)
    for it in range(3):
        ea = a.site_update(0.4, 0.4)
        ec = c.site_update(0.4, 0.4)
        a.check_status()
        c.check_status()
        assert (ea - ec).abs().max() <= 1e-11 * ec.abs().max(), it
        for nm in ("fx_mus", "fx_covs", "qA", "qb"):
            ref = getattr(c, nm)
            assert (getattr(a, nm) - ref).abs().max() <= 1e-11 * ref.abs().max(), (it, nm)
        if it == 0:
            a.relinearize()
            c.relinearize()
    th = c._prior_naturals(c.pA, c.pb)
    for nm, prior in (("g1", th[0]), ("gd", th[1]), ("gs", th[2])):
        ref = getattr(c, nm)
        scale = max(ref.abs().max(), 1e-3 * (ref + prior).abs().max())
        assert (getattr(a, nm) - ref).abs().max() <= 1e-9 * scale, nm
        assert (getattr(a, nm) - ref).abs().max() <= 1e-11 * (ref + prior).abs().max(), nm


def test_kl_gradient_entry_point_low_order_equals_twenty_points(vdp_cfg):
    """vdp_cvi_kl_and_grad (value and gradient of the KL alone) for Lorenz-63: process-wide switch instead of the flag"""
    from vi_diffusion_processes_b200 import ops
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    rng = np.random.default_rng(2)
    B, N, D = 2, 60, 3
    dev = "cuda"
    A_np = np.eye(D) * 0.97 + 0.01 * rng.normal(size=(B, N, D, D))
    S_np = np.empty((B, N + 1, D, D))
    L0 = np.tril(0.05 * rng.normal(size=(B, D, D))) + 0.3 * np.eye(D)
    S_np[:, 0] = L0 @ np.swapaxes(L0, -1, -2)
    for k in range(N):                                   # marginal covariances of a valid chain: S' = A S A^T + Q
        S_np[:, k + 1] = A_np[:, k] @ S_np[:, k] @ np.swapaxes(A_np[:, k], -1, -2) + 0.004 * np.eye(D)
    A = torch.as_tensor(A_np, device=dev)
    m = torch.as_tensor(1.0 + 0.2 * rng.normal(size=(B, N + 1, D)), device=dev)
    Sd = torch.as_tensor(S_np, device=dev)
    grid = torch.as_tensor(np.arange(N + 1) * 0.002, device=dev)
    args = (sde_mod.DRIFT_LORENZ63, [10.0, 28.0, 8.0 / 3.0], np.eye(D), 0.002, grid, np.ones(D), 0.1 * np.eye(D), A, m, Sd)
    kl_a, g_a = ops.cvi_kl_and_grad(*args)
    vdp_cfg(GH_QUADRATURE=1)
    kl_c, g_c = ops.cvi_kl_and_grad(*args)
    assert bool(torch.isfinite(kl_c).all())
    assert (kl_a - kl_c).abs().max() <= 1e-11 * kl_c.abs().max()
    for x, r in zip(g_a, g_c):
        assert (x - r).abs().max() <= 1e-10 * r.abs().max()


@pytest.mark.parametrize("name", ["vdp", "lorenz", "ou3"])
def test_linearisation_low_order_equals_ten_points(name, vdp_cfg):
    """E f and E J of the linearised prior (sde.py:109-128, H = 10 in the reference) through the low-order rule"""
    from vi_diffusion_processes_b200 import ops
    did, params, D, N, dt = _case(name)
    rng = np.random.default_rng(11 + D)
    B = 2
    grid = torch.as_tensor(np.arange(N + 1) * dt, device="cuda")
    m = torch.as_tensor(1.0 + 0.3 * rng.normal(size=(B, N, D)), device="cuda")
    L = np.tril(0.1 * rng.normal(size=(B, N, D, D))) + 0.4 * np.eye(D)
    S = torch.as_tensor(L @ np.swapaxes(L, -1, -2), device="cuda")
    a = ops.linearize_to_ssm(did, params, np.eye(D), grid, m, S, clip=(-1.0, 1.0), want_unclipped=True)
    vdp_cfg(GH_QUADRATURE=1)
    c = ops.linearize_to_ssm(did, params, np.eye(D), grid, m, S, clip=(-1.0, 1.0), want_unclipped=True)
    for x, r in zip(a, c):
        assert float((x - r).abs().max()) <= 1e-12 * max(float(r.abs().max()), 1.0)


@pytest.mark.parametrize("name", ["vdp", "lorenz"])
def test_vdp_sweeps_low_order_equal_reference_order(name):
    """VariationalMarkovGP sweeps (vi_sde.py:171-414): E_sde with its gradients (H = 20) and the parameter update
    (H = 10) through the low-order rule against VDP_VGP_GH_QUADRATURE"""
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.vgp import VgpEngine
    did, params, D, N, dt = _case(name)
    N = 200 if D == 2 else 100
    rng = np.random.default_rng(21 + D)
    B, n_obs = 2, 8
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = 1.0 + 0.2 * rng.normal(size=(B, n_obs, D))

    def make(flags):
        eng = VgpEngine(did, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.ones(D), 0.01 * np.eye(D))
        eng.route_flags = flags
        return eng

    a, c = make(0), make(_lib.VGP_GH_QUADRATURE)
    for it in range(3):
        a.sweep(0.05)
        c.sweep(0.05)
        for nm in ("A", "b", "m", "S"):
            ref = getattr(c, nm)
            assert float((getattr(a, nm) - ref).abs().max()) <= 1e-11 * max(float(ref.abs().max()), 1.0), (it, nm)
    ea, ec = a.elbo(), c.elbo()
    assert float((ea - ec).abs().max()) <= 1e-11 * float(ec.abs().max())
