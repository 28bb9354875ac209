"""
Time-sharded CVI-DP site update (dist_cvi.TimeShardedCvi: no rank holds the whole chain; BASELINE config 5 path):
emulated ranks on one GPU -- same kernels, same blobs, the all-gather replaced by a list -- against the single-GPU
engine on the same problem and against the oracle.  The multi-process run over NCCL is tools/c5_time_sharded.py /
bench.py --workload c5.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import models_np as M
from oracle import scalar_d1
from oracle import sde_np as S

pytestmark = pytest.mark.gpu


def _gather(sh, name):
    return torch.cat(sh.owned(name))


def _close(a, b, tol, scale=None):
    den = float(b.abs().max()) if scale is None else scale
    return float((a - b).abs().max()) <= tol * max(den, 1e-300)


@pytest.mark.parametrize("R", [1, 2, 3, 5])
def test_time_sharded_update_d1_equals_single_gpu_and_oracle(R, vdp_cfg):
    from vi_diffusion_processes_b200.cvi import CviEngine
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    vdp_cfg(SWEEPS="tiled")
    rng = np.random.default_rng(31)
    N, n_obs = 2300, 21
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    idx[0], idx[-1] = 0, N                                   # first and last grid point observed
    y = rng.normal(size=(n_obs, 1)) * 0.3 + 1.0
    args = (S.DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]])
    single = CviEngine(*args[:4], idx[None], y[None], *args[6:], lazy_sites=False)
    ref = scalar_d1.BatchedCVISitesSDE1D(S.DRIFT_DW, [4.0, 1.0], 1.0, grid, idx[None], y[None, :, 0], 0.1, 1.0, 1e-2)
    sh = TimeShardedCvi(*args, comm=Comm.emulated(R))
    for it in range(3):
        e = sh.site_update(0.5, 0.5)
        s = single.site_update(0.5, 0.5).item()
        r = ref.site_update(0.5, 0.5)[0]
        sh.check_status()
        assert abs(e - s) <= 1e-11 * abs(s), (it, e, s)
        assert abs(e - r) <= 1e-9 * abs(r), (it, e, r)
        th = single._prior_naturals(single.pA, single.pb)
        for k, t in (("g1", th[0]), ("gd", th[1]), ("gs", th[2])):
            want = getattr(single, k)[0]
            assert _close(_gather(sh, k), want, 1e-11, scale=float((t[0] + want).abs().max())), (it, k)
        for k in ("fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
            assert _close(_gather(sh, k), getattr(single, k)[0], 1e-10), (it, k)
        if it == 1:
            sh.relinearize()
            single.relinearize()
            ref.relinearize()


def test_time_sharded_update_3d():
    from vi_diffusion_processes_b200.cvi import CviEngine
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    rng = np.random.default_rng(32)
    N, n_obs = 41, 7
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, 3)) * 0.5
    sde = S.OUNdSDE(0.5, np.eye(3))
    x0 = S.Gaussian(np.zeros(3), 0.5 * np.eye(3))
    ref = M.CVISitesSDE(sde, grid, (grid[idx], y), M.MultivariateGaussian(0.1 * np.eye(3)), prior_initial_state=x0)
    args = (sde.drift_id, sde.params, sde.q, grid, idx, y, 0.1 * np.eye(3), x0.mu, x0.cov)
    single = CviEngine(*args, lazy_sites=False)
    sh = TimeShardedCvi(*args, comm=Comm.emulated(4))
    for it in range(2):
        e = sh.site_update(0.3, 0.3)
        s = single.site_update(0.3, 0.3).item()
        r = ref.site_update(0.3, 0.3)
        sh.check_status()
        assert abs(e - s) <= 1e-11 * abs(s) and abs(e - r) <= 1e-9 * abs(r)
        for k in ("fx_mus", "fx_covs", "qA"):
            assert _close(_gather(sh, k), getattr(single, k)[0], 1e-10), k
        th = single._prior_naturals(single.pA, single.pb)
        for k, t in (("g1", th[0]), ("gd", th[1]), ("gs", th[2])):
            want = getattr(single, k)[0]
            # (OU prior with zero offsets: nat_1(p) and the exact g1 are 0, theta_1 is made of the data sites)
            scale = max(float((t[0] + want).abs().max()), float(single.d1.abs().max()) if k == "g1" else 0.0)
            assert _close(_gather(sh, k), want, 1e-11, scale=scale), k


@pytest.mark.parametrize("D,R,N", [(1, 3, 3100), (2, 3, 1500), (3, 4, 900)])
def test_time_sharded_marginals_from_first_factor(D, R, N, vdp_cfg):
    """The time-sharded rebuild takes the marginal covariances from the backward sweep of naturals_to_ssm and the means
    from a mean-only forward sweep (VDP_SWEEP_MEAN, VDP_SWEEP_NAT2SSM out[5]); ``refactor_precision=True`` keeps the
    reference's route (precision rebuilt from the SSM, second factorisation, inverse band: gauss_markov.py:107-117).
    Same update to rounding, several segments per rank."""
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    vdp_cfg(SWEEPS="tiled", NSEG=3)
    did, params = {1: (sde_mod.DRIFT_DW, [4.0, 1.0]), 2: (sde_mod.DRIFT_VDP, [1.0, 1.0]),
                   3: (sde_mod.DRIFT_OU_ND, [0.5])}[D]
    rng = np.random.default_rng(40 + D)
    n_obs = 17
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.3 + 1.0
    args = (did, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.ones(D), 0.05 * np.eye(D))
    a = TimeShardedCvi(*args, comm=Comm.emulated(R), info_rebuild=False)
    c = TimeShardedCvi(*args, comm=Comm.emulated(R), info_rebuild=False, refactor_precision=True)
    for it in range(3):
        ea, ec = a.site_update(0.5, 0.5), c.site_update(0.5, 0.5)
        a.check_status()
        c.check_status()
        assert abs(ea - ec) <= 1e-11 * abs(ec), (it, ea, ec)
        for k in ("fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
            assert _close(_gather(a, k), _gather(c, k), 1e-10), (it, k)
        if it == 0:
            a.relinearize()
            c.relinearize()


@pytest.mark.parametrize("D,R,N", [(1, 3, 2300), (3, 2, 700)])
def test_time_sharded_update_replayed_from_a_graph(D, R, N):
    """TimeShardedCvi.site_update_graphed: the first call runs eagerly and captures, the others replay -- the same
    launches on the same buffers, so ELBO and state equal the plain updates bit for bit"""
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    rng = np.random.default_rng(33 + D)
    n_obs = 17
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, D)) * 0.3 + 1.0
    if D == 1:
        args = (S.DRIFT_DW, [4.0, 1.0], np.eye(1), grid, idx, y, 0.1 * np.eye(1), np.ones(1), 1e-2 * np.eye(1))
    else:
        args = (S.DRIFT_OU_ND, [0.5], np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.zeros(D), 0.5 * np.eye(D))
    plain = TimeShardedCvi(*args, comm=Comm.emulated(R))
    graphed = TimeShardedCvi(*args, comm=Comm.emulated(R))
    for it in range(4):
        a = plain.site_update(0.4, 0.4)
        b = graphed.site_update_graphed(0.4, 0.4)
        plain.check_status()
        graphed.check_status()
        assert a == b, (it, a, b)
        for k in ("g1", "gd", "gs", "fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
            assert torch.equal(_gather(plain, k), _gather(graphed, k)), (it, k)
    assert len(graphed._graphs) == 1
    graphed.relinearize()                   # new prior buffers: the next call re-captures
    plain.relinearize()
    for it in range(2):
        assert plain.site_update(0.4, 0.4) == graphed.site_update_graphed(0.4, 0.4)
    assert len(graphed._graphs) == 1


@pytest.mark.parametrize("D,R", [(1, 3), (3, 2)])
def test_time_sharded_naturals_carried_between_updates(D, R):
    """carry_naturals (default): the first rebuild of an update corrects theta at the observed grid points instead of
    rebuilding it from the prior everywhere -- same state as the full rebuild to rounding; a torch-side write to the
    sites or a re-linearisation in between falls back to the full rebuild"""
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    rng = np.random.default_rng(43 + D)
    N, n_obs = 900, 15
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    idx[0], idx[-1] = 0, N
    y = rng.normal(size=(n_obs, D)) * 0.3 + 1.0
    if D == 1:
        args = (S.DRIFT_DW, [4.0, 1.0], np.eye(1), grid, idx, y, 0.1 * np.eye(1), np.ones(1), 1e-2 * np.eye(1))
    else:
        args = (S.DRIFT_OU_ND, [0.5], np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.zeros(D), 0.5 * np.eye(D))
    full = TimeShardedCvi(*args, comm=Comm.emulated(R), carry_naturals=False)
    lazy = TimeShardedCvi(*args, comm=Comm.emulated(R))
    for it in range(6):
        a, b = full.site_update(0.4, 0.4), lazy.site_update(0.4, 0.4)
        lazy.check_status()
        assert abs(a - b) <= 1e-11 * abs(a), (it, a, b)
        for k in ("g1", "gd", "gs", "fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
            # the sites are small differences of naturals of size 1 / (q dt) (zero up to rounding for a linear
            # drift): measured against theta, as the other tests of this file do
            scale = float(max(st.thd.abs().max() for st in full.ranks)) if k[0] == "g" else None
            assert _close(_gather(lazy, k), _gather(full, k), 1e-11, scale=scale), (it, k)
        assert lazy._th_sig == lazy._th_signature()
        if it == 1:
            full.relinearize()
            lazy.relinearize()
            assert lazy._th_sig is None
        if it == 3:                         # data sites rewritten behind the engine's back
            for eng in (full, lazy):
                for st in eng.ranks:
                    st.eng.d1.mul_(0.5)
            assert lazy._th_sig != lazy._th_signature()


@pytest.mark.parametrize("D,R,N", [(1, 1, 700), (1, 3, 3100), (2, 3, 1500), (3, 1, 500), (3, 4, 900), (3, 5, 2100)])
def test_time_sharded_rebuild_in_information_form(D, R, N, vdp_cfg):
    """``info_rebuild=True``: dist_q as ONE backward sweep in information form (VDP_SWEEP_INFO: the Schur complements of
    the precision from the bottom, transitions straight from them) and the marginals by the moment recursion
    (VDP_SWEEP_MOMENTS) -- against the default rebuild (Cholesky sweep, inverse band, means): the same Gaussian, so the
    same update to rounding; several segments per rank, first / last grid point observed."""
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    from vi_diffusion_processes_b200.markovflow.sde import sde as sde_mod
    vdp_cfg(SWEEPS="tiled", NSEG=3)
    did, params = {1: (sde_mod.DRIFT_DW, [4.0, 1.0]), 2: (sde_mod.DRIFT_VDP, [1.0, 1.0]),
                   3: (sde_mod.DRIFT_OU_ND, [0.5])}[D]
    rng = np.random.default_rng(50 + D)
    n_obs = 17
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    idx[0], idx[-1] = 0, N
    y = rng.normal(size=(n_obs, D)) * 0.3 + 1.0
    args = (did, params, np.eye(D), grid, idx, y, 0.1 * np.eye(D), np.ones(D), 0.05 * np.eye(D))
    a = TimeShardedCvi(*args, comm=Comm.emulated(R), info_rebuild=True)
    c = TimeShardedCvi(*args, comm=Comm.emulated(R), info_rebuild=False)
    for it in range(4):
        ea, ec = a.site_update(0.5, 0.5), c.site_update(0.5, 0.5)
        a.check_status()
        c.check_status()
        assert abs(ea - ec) <= 1e-11 * abs(ec), (it, ea, ec)
        for k in ("fx_mus", "fx_covs", "qA", "qb", "qcholQ"):
            assert _close(_gather(a, k), _gather(c, k), 1e-10), (it, k)
        scale = float(max(st.thd.abs().max() for st in c.ranks))
        for k in ("g1", "gd", "gs"):
            assert _close(_gather(a, k), _gather(c, k), 1e-11, scale=scale), (it, k)
        for x, r in zip((a.ranks[0].eng.qmu0, a.ranks[0].eng.qcholP0), (c.ranks[0].eng.qmu0, c.ranks[0].eng.qcholP0)):
            assert _close(x, r, 1e-10)
        if it == 1:
            a.relinearize()
            c.relinearize()


def test_information_form_rebuild_against_the_oracle_3d():
    """the two-sweep rebuild against the NumPy restatement of the reference (D = 3, N small enough for the oracle)"""
    from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    rng = np.random.default_rng(61)
    N, n_obs = 41, 7
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, 3)) * 0.5
    sde = S.OUNdSDE(0.5, np.eye(3))
    x0 = S.Gaussian(np.zeros(3), 0.5 * np.eye(3))
    ref = M.CVISitesSDE(sde, grid, (grid[idx], y), M.MultivariateGaussian(0.1 * np.eye(3)), prior_initial_state=x0)
    args = (sde.drift_id, sde.params, sde.q, grid, idx, y, 0.1 * np.eye(3), x0.mu, x0.cov)
    sh = TimeShardedCvi(*args, comm=Comm.emulated(3), info_rebuild=True)
    for it in range(2):
        e = sh.site_update(0.3, 0.3)
        r = ref.site_update(0.3, 0.3)
        sh.check_status()
        assert abs(e - r) <= 1e-9 * abs(r), (it, e, r)
        assert _close(_gather(sh, "fx_mus").cpu(), torch.as_tensor(np.asarray(ref.fx_mus).reshape(N + 1, 3)), 1e-9)
        assert _close(_gather(sh, "fx_covs").cpu(), torch.as_tensor(np.asarray(ref.fx_covs).reshape(N + 1, 3, 3)), 1e-9)
