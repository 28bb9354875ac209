"""
The one-kernel forward pass of the site update (csrc/fwd_fused.cuh: marginal means -> KL gradient / girsanov site
update -> Cholesky factor of the next rebuild; D = 1, polynomial drift, sites held in theta, one segment per sequence)
against

  * the three-kernel route it replaces (VDP_FLAG_NO_FUSED_FORWARD), to 1e-11 of the posterior natural parameters, at
    tile-boundary chain lengths (one partial tile, exactly one tile, one step more, several tiles), with duplicate
    observation times and with observations on the first and last grid point;
  * the C restatement of the oracle (oracle/cvi_d1.c) over several updates with a re-linearisation, at 1e-9;
  * the device-side trainer loop (per-sequence step sizes read from device memory) against the host loop.

Reference: variational_cvi_sde.py:279-299 (update_girsanov_sites), :176-192 (dist_q), sde_utils.py:473-547.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import c_port, scalar_d1
from oracle import sde_np as S

pytestmark = pytest.mark.gpu

TILE = 128 * 9          # grid steps per tile of the D = 1 sweeps (TileCfg<1>)


def _problem(B, N, n_obs, seed, dup=False, ends=False):
    rng = np.random.default_rng(seed)
    dt = 0.01
    grid = np.arange(N + 1) * dt
    idx = np.sort(rng.choice(np.arange(1, N), size=(B, n_obs), replace=True), axis=1)
    if dup:
        idx[:, 1] = idx[:, 0]
        idx[:, -1] = idx[:, -2]
    if ends:
        idx[:, 0] = 0
        idx[:, -1] = N
    y = 0.7 * np.sin(grid[idx]) + 0.1 * rng.standard_normal((B, n_obs))
    return grid, idx.astype(np.int64), y


def _engine(drift, params, grid, idx, y, flags=0):
    from vi_diffusion_processes_b200.cvi import CviEngine
    eng = CviEngine(drift, params, np.ones((1, 1)), grid, idx, y[..., None], [[0.2]], [0.3], [[0.5]])
    eng.route_flags = flags
    return eng


def _theta_scale(eng):
    p = {k: getattr(eng, k).cpu().numpy() for k in ("pA", "pb", "pcholQ", "pmu0", "pcholP0")}
    th = scalar_d1.ssm_to_naturals(p["pA"][..., 0, 0], p["pb"][..., 0], p["pcholQ"][..., 0, 0], p["pmu0"][..., 0],
                                   p["pcholP0"][..., 0, 0])
    return th


@pytest.mark.parametrize("drift,params", [(S.DRIFT_DW, [1.0, 1.0]), (S.DRIFT_OU, [0.8])])
@pytest.mark.parametrize("N", [2, 3, 10, 100, TILE - 1, TILE, TILE + 1, 2 * TILE, 2 * TILE + 1, 5000])
def test_fused_forward_equals_three_kernel_route(drift, params, N, vdp_cfg):
    from vi_diffusion_processes_b200 import _lib
    vdp_cfg(SWEEPS="tiled", NSEG=1)
    B, n_obs = 5, (12 if N >= 100 else 2)
    grid, idx, y = _problem(B, N, n_obs, seed=N, dup=True, ends=(N % 2 == 0))
    fused = _engine(drift, params, grid, idx, y)
    plain = _engine(drift, params, grid, idx, y, flags=_lib.FLAG_NO_FUSED_FORWARD)
    for it in range(4):
        ef = fused.site_update(0.4, 0.5)
        ep = plain.site_update(0.4, 0.5)
        fused.check_status()
        plain.check_status()
        assert torch.allclose(ef, ep, rtol=1e-11, atol=0), it
        for name in ("fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
            a, b = getattr(fused, name), getattr(plain, name)
            assert float((a - b).abs().max() / b.abs().max()) < 1e-11, (it, name)
        if it == 1:
            fused.relinearize()
            plain.relinearize()
    th = _theta_scale(plain)
    for name, t in (("g1", th[0]), ("gd", th[1]), ("gs", th[2])):
        a = getattr(fused, name).cpu().numpy().reshape(t.shape)
        b = getattr(plain, name).cpu().numpy().reshape(t.shape)
        assert np.abs(a - b).max() / np.abs(t + b).max() < 1e-11, name


@pytest.mark.parametrize("N", [700, 3000])
def test_fused_forward_against_the_oracle(N, vdp_cfg):
    vdp_cfg(SWEEPS="tiled", NSEG=1)
    B, n_obs = 6, 15
    grid, idx, y = _problem(B, N, n_obs, seed=7 + N)
    eng = _engine(S.DRIFT_DW, [1.0, 1.0], grid, idx, y)
    ref = c_port.CBatchedCVISitesSDE1D(S.DRIFT_DW, [1.0, 1.0], 1.0, grid, idx, y, 0.2, 0.3, 0.5,
                                       threads=min(8, c_port.max_threads()))
    for it in range(5):
        e = eng.site_update(0.5, 0.5).cpu().numpy()
        r = ref.site_update(0.5, 0.5)
        eng.check_status()
        assert np.abs(e - r).max() / np.abs(r).max() < 1e-9, it
        assert np.abs(eng.fx_mus[..., 0].cpu().numpy() - ref.fx_mus).max() / np.abs(ref.fx_mus).max() < 1e-9, it
        assert np.abs(eng.fx_covs[..., 0, 0].cpu().numpy() - ref.fx_covs).max() / np.abs(ref.fx_covs).max() < 1e-9, it
        if it == 2:
            eng.relinearize()
            ref.relinearize()
    th = scalar_d1.ssm_to_naturals(ref.p["A"], ref.p["b"], ref.p["cQ"], ref.p["mu0"], ref.p["cP0"])
    for name, got, want, t in (("g1", eng.g1[..., 0], ref.g1, th[0]), ("gd", eng.gd[..., 0, 0], ref.gd, th[1]),
                               ("gs", eng.gs[..., 0, 0], ref.gs, th[2])):
        got = got.cpu().numpy()
        scale = max(np.abs(want).max(), 1e-3 * np.abs(t + want).max())
        assert np.abs(got - want).max() / scale < 1e-9, name


def test_fused_forward_reports_a_lost_positive_definiteness(vdp_cfg):
    """a step size far too large drives the precision out of the cone: the status word must say so, on the same
    update as on the three-kernel route"""
    from vi_diffusion_processes_b200 import _lib
    vdp_cfg(SWEEPS="tiled", NSEG=1)
    grid, idx, y = _problem(3, 1500, 10, seed=3)

    def first_failure(flags):
        eng = _engine(S.DRIFT_DW, [1.0, 1.0], grid, idx, 5.0 * y, flags=flags)
        eng.site_update(0.5, 0.5)
        for it in range(6):
            eng.site_update(1.0, 60.0)
            try:
                eng.check_status()
            except Exception:
                return it
        return None

    fused, plain = first_failure(0), first_failure(_lib.FLAG_NO_FUSED_FORWARD)
    assert fused is not None and fused == plain


def test_fused_forward_in_the_device_loop(vdp_cfg):
    """per-sequence step sizes and stop flags in device memory (vdp_cvi_optimize_sites) on the fused route"""
    from vi_diffusion_processes_b200 import _lib
    vdp_cfg(SWEEPS="tiled", NSEG=1)
    B, N = 4, 1300
    grid, idx, y = _problem(B, N, 10, seed=11)
    y = y * np.array([0.05, 0.3, 0.6, 1.0])[:, None]
    a = _engine(S.DRIFT_DW, [1.0, 1.0], grid, idx, y)
    b = _engine(S.DRIFT_DW, [1.0, 1.0], grid, idx, y, flags=_lib.FLAG_NO_FUSED_FORWARD)
    ra = a.optimize_sites(8, 0.5, 0.7, 0.7)
    rb = b.optimize_sites(8, 0.5, 0.7, 0.7)
    a.check_status()
    b.check_status()
    assert torch.equal(ra["n_itr"], rb["n_itr"])
    assert int(ra["n_itr"].max()) >= 3 and int(ra["n_itr"].min()) < 8      # some converge early, some do not
    assert torch.allclose(ra["elbo_trace"], rb["elbo_trace"], rtol=1e-10, atol=0)
    assert float((a.fx_mus - b.fx_mus).abs().max()) < 1e-10


@pytest.mark.parametrize("D,stab", [(1, False), (1, True), (2, False)])
def test_vdp_sweep_carried_state_is_bitwise_the_plain_sweep(D, stab):
    """VDP_VGP_STATE_CARRIED: a sweep starts from the forward pass and the E_sde quadrature the previous sweep ended
    with (vi_markov_gp_trainer.py:62-73 computes both twice).  Bitwise the same parameters, multipliers, marginals and
    ELBO; an in-place change of A between sweeps (torch version counter) or another call on the engine drops the
    carry."""
    from vi_diffusion_processes_b200.vgp import VgpEngine
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    did, params = (sde_mod.DRIFT_DW, [4.0, 1.0]) if D == 1 else (sde_mod.DRIFT_VDP, [1.0, 1.0])
    rng = np.random.default_rng(60 + D)
    B, N, n_obs = 3, 1500 if D == 1 else 300, 9
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(np.stack([rng.choice(np.arange(1, N + 1), n_obs, replace=False) for _ in range(B)]), axis=1)
    y = 1.0 + 0.2 * rng.normal(size=(B, n_obs, D))
    mk = lambda: VgpEngine(did, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.ones(D), 0.01 * np.eye(D),
                           stabilize_system=stab)
    a, c = mk(), mk()
    for e in (a, c):
        e.A += 1.5
        e.b += 1.5
    carried = []
    for it in range(6):
        if it == 3:                       # in-place edit: the carry must notice
            a.A.mul_(0.99)
            c.A.mul_(0.99)
        if it == 4:                       # another entry point in between
            a.elbo()
        sig_ok = a._carry_sig is not None and a._carry_sig == a._carry_signature()
        carried.append(sig_ok)
        ea = a.sweep(0.1, update_x0=(it % 2 == 1)).clone()
        ec = c.sweep(0.1, update_x0=(it % 2 == 1), carry=False).clone()
        assert torch.equal(ea, ec), it
        for nm in ("A", "b", "m", "S", "lambda_lagrange", "psi_lagrange", "q_initial_mean", "q_initial_chol"):
            assert torch.equal(getattr(a, nm), getattr(c, nm)), (it, nm)
    assert carried == [False, True, True, False, False, True]
    a.check_status()
