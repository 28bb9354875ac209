"""The C / OpenMP restatement of the D = 1 oracle (oracle/cvi_d1.c) against the NumPy oracle it restates."""
import shutil

import numpy as np
import pytest

from oracle import scalar_d1
from oracle import sde_np as S

pytestmark = pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")


def _problem(rng, B, N, n_obs, dt):
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs)) * 0.3 + np.sign(rng.normal(size=(B, 1)))
    return grid, idx, y


@pytest.mark.parametrize("drift,params,threads", [(S.DRIFT_DW, [4.0, 1.0], 1), (S.DRIFT_DW, [4.0, 1.0], 3),
                                                  (S.DRIFT_OU, [1.2], 2), (S.DRIFT_BENES, [1.0], 1),
                                                  (S.DRIFT_SINE, [0.3], 2)])
def test_c_port_matches_numpy_oracle(drift, params, threads):
    from oracle import c_port
    rng = np.random.default_rng(3)
    grid, idx, y = _problem(rng, 5, 400, 15, 0.005)
    a = scalar_d1.BatchedCVISitesSDE1D(drift, params, 1.0, grid, idx, y, 0.1, 1.0, 1e-2)
    c = c_port.CBatchedCVISitesSDE1D(drift, params, 1.0, grid, idx, y, 0.1, 1.0, 1e-2, threads=threads)
    for it in range(3):
        ea, ec = a.site_update(0.5, 0.5), c.site_update(0.5, 0.5)
        np.testing.assert_allclose(ec, ea, rtol=1e-11)
        for k in ("fx_mus", "fx_covs", "d1", "d2"):
            np.testing.assert_allclose(getattr(c, k), getattr(a, k), rtol=1e-10, atol=1e-13, err_msg=k)
        th = scalar_d1.ssm_to_naturals(a.p["A"], a.p["b"], a.p["cQ"], a.p["mu0"], a.p["cP0"])
        for k, t in (("g1", th[0]), ("gd", th[1]), ("gs", th[2])):
            scale = max(np.abs(getattr(a, k)).max(), 1e-3 * np.abs(t + getattr(a, k)).max())
            assert np.abs(getattr(c, k) - getattr(a, k)).max() / scale < 1e-10, k
        if it == 1:
            a.relinearize()
            c.relinearize()
