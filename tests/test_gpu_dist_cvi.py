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
