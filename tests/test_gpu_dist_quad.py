"""Sharded quadrature (vdp_cvi_site_update_sharded, dist_quad.ShardedQuadratureCvi), emulated ranks on one GPU,
against the single-GPU site update and the oracle."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import models_np as M
from oracle import scalar_d1
from oracle import sde_np as S

pytestmark = pytest.mark.gpu


def _engines(R, mk):
    return [mk() for _ in range(R)]


@pytest.mark.parametrize("R", [1, 2, 3, 5])
def test_sharded_quadrature_d1_equals_single_gpu_and_oracle(R):
    from vi_diffusion_processes_b200.cvi import CviEngine
    from vi_diffusion_processes_b200.dist_quad import ShardedQuadratureCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    rng = np.random.default_rng(11)
    N, n_obs = 900, 14
    grid = np.round(np.arange(N + 1) * 0.005, 6)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))[None]
    y = rng.normal(size=(1, n_obs)) * 0.3 + 1.0
    mk = lambda: CviEngine(S.DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y[..., None], [[0.1]], [1.0], [[1e-2]],
                           lazy_sites=False)   # bitwise comparisons below: sites written by every update
    single = mk()
    ref = scalar_d1.BatchedCVISitesSDE1D(S.DRIFT_DW, [4.0, 1.0], 1.0, grid, idx, y, 0.1, 1.0, 1e-2)
    sh = ShardedQuadratureCvi(_engines(R, mk), Comm.emulated(R))
    for it in range(3):
        e = sh.site_update(0.5, 0.5)
        s = single.site_update(0.5, 0.5).item()
        r = ref.site_update(0.5, 0.5)[0]
        assert abs(e - s) <= 1e-13 * abs(s)
        assert abs(e - r) <= 1e-9 * abs(r)
        for eng in sh.engines:          # every rank ends with the same, complete state
            for k in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
                assert torch.equal(getattr(eng, k), getattr(single, k)), k


def test_sharded_quadrature_3d():
    from vi_diffusion_processes_b200.cvi import CviEngine
    from vi_diffusion_processes_b200.dist_quad import ShardedQuadratureCvi
    from vi_diffusion_processes_b200.dist_time import Comm
    rng = np.random.default_rng(12)
    N, n_obs = 37, 6
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))
    y = rng.normal(size=(n_obs, 3)) * 0.5
    sde = S.OUNdSDE(0.5, np.eye(3))
    x0 = S.Gaussian(np.zeros(3), 0.5 * np.eye(3))
    ref = M.CVISitesSDE(sde, grid, (grid[idx], y), M.MultivariateGaussian(0.1 * np.eye(3)), prior_initial_state=x0)
    mk = lambda: CviEngine(sde.drift_id, sde.params, sde.q, grid, idx, y, 0.1 * np.eye(3), x0.mu, x0.cov, lazy_sites=False)
    single = mk()
    sh = ShardedQuadratureCvi(_engines(4, mk), Comm.emulated(4))
    for it in range(2):
        e = sh.site_update(0.3, 0.3)
        s = single.site_update(0.3, 0.3).item()
        r = ref.site_update(0.3, 0.3)
        assert abs(e - s) <= 1e-13 * abs(s) and abs(e - r) <= 1e-9 * abs(r)
        for k in ("g1", "gd", "gs", "fx_mus", "fx_covs"):
            assert torch.equal(getattr(sh.engines[2], k), getattr(single, k)), k
