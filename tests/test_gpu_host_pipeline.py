"""
GPU tests of the host-buffer entry (vdp_cvi_host_site_update): state in pinned host memory, chunked over
sequences with copies and kernels overlapped.  Checked against the NumPy oracle (1e-9) and against the
device-resident entry on the same inputs.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import scalar_d1
from oracle import sde_np as S

pytestmark = pytest.mark.gpu

TOL = 1e-9


def rel(a, b, scale=None):
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    b = b.detach().cpu().numpy() if hasattr(b, "detach") else np.asarray(b)
    den = np.abs(b).max() if scale is None else scale
    return float(np.abs(a - b).max() / (den if den > 0 else 1.0))


def _problem(rng, B, N, n_obs, dt):
    grid = np.round(np.arange(N + 1) * dt, 6)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs)) * 0.3 + np.sign(rng.normal(size=(B, 1)))
    return grid, idx, y


@pytest.mark.parametrize("B,chunk,slots,resident", [(10, 4, 3, False), (10, 4, 3, True), (7, 7, 1, False),
                                                    (5, 2, 2, False), (3, 8, 4, True)])
def test_host_site_update_matches_oracle_and_resident(B, chunk, slots, resident):
    from vi_diffusion_processes_b200.cvi import CviEngine, CviHostEngine
    rng = np.random.default_rng(31)
    N = 700
    grid, idx, y = _problem(rng, B, N, 12, 0.005)
    mk = lambda: CviEngine(S.DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y[..., None], [[0.1]], [1.0], [[1e-2]])
    eng, dev_eng = mk(), mk()
    ref = scalar_d1.BatchedCVISitesSDE1D(S.DRIFT_DW, [4.0, 1.0], 1.0, grid, idx, y, 0.1, 1.0, 1e-2)
    host = CviHostEngine(eng, chunk=chunk, slots=slots, resident_prior=resident)
    up, down = host.bytes_per_update()
    assert up > 0 and down > 0
    for it in range(3):
        e = host.site_update(0.5, 0.5)
        d = dev_eng.site_update(0.5, 0.5)
        r = ref.site_update(0.5, 0.5)
        h = host.host
        assert rel(e, r) < TOL and rel(e, d) < 1e-12
        assert rel(h["fx_mus"][..., 0], ref.fx_mus) < TOL and rel(h["fx_covs"][..., 0, 0], ref.fx_covs) < TOL
        th = scalar_d1.ssm_to_naturals(ref.p["A"], ref.p["b"], ref.p["cQ"], ref.p["mu0"], ref.p["cP0"])
        for got, want, t in ((h["g1"][..., 0], ref.g1, th[0]), (h["gd"][..., 0, 0], ref.gd, th[1]),
                             (h["gs"][..., 0, 0], ref.gs, th[2])):
            scale = max(np.abs(want).max(), 1e-3 * np.abs(t + want).max())
            assert rel(got, want, scale=scale) < TOL
        assert rel(h["d1"][..., 0], ref.d1) < TOL and rel(h["d2"][..., 0, 0], ref.d2) < TOL
        # same numbers as the device-resident entry (chunking only changes the launch geometry)
        for k in ("g1", "gd", "gs", "fx_mus", "fx_covs", "qA", "qb", "qcholQ", "qmu0", "qcholP0", "ve", "kl"):
            assert rel(h[k], getattr(dev_eng, k)) < 1e-10, k
    host.close()


def test_host_site_update_2d_and_optional_outputs():
    from vi_diffusion_processes_b200.cvi import CviEngine, CviHostEngine
    rng = np.random.default_rng(5)
    N, n_obs, B = 150, 9, 5
    grid = np.round(np.arange(N + 1) * 0.01, 2)
    idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
    y = rng.normal(size=(B, n_obs, 2))
    mk = lambda: CviEngine(S.DRIFT_VDP, [1.0, 1.0], np.eye(2), grid, idx, y, 0.1 * np.eye(2), np.ones(2),
                           1e-2 * np.eye(2))
    eng, dev_eng = mk(), mk()
    host = CviHostEngine(eng, chunk=2, slots=3, want_posterior_ssm=False)
    before = host.host["qA"].clone()
    for _ in range(2):
        e = host.site_update(0.3, 0.3)
        d = dev_eng.site_update(0.3, 0.3)
        assert rel(e, d) < 1e-12
        for k in ("g1", "gd", "gs", "fx_mus", "fx_covs", "d1", "d2"):
            assert rel(host.host[k], getattr(dev_eng, k)) < 1e-10, k
    assert torch.equal(host.host["qA"], before)      # NULL output pointers are left alone
    host.close()


def test_host_errors():
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.cvi import CviEngine, CviHostEngine
    rng = np.random.default_rng(2)
    grid, idx, y = _problem(rng, 3, 64, 4, 0.01)
    eng = CviEngine(S.DRIFT_OU, [1.0], np.ones((1, 1)), grid, idx, y[..., None], [[0.1]], [0.0], [[1.0]])
    host = CviHostEngine(eng, chunk=2, slots=2)
    host.resident_prior = True            # pA = NULL without vdp_cvi_host_set_prior
    with pytest.raises(_lib.VdpError):
        host.site_update(0.1, 0.1)
    host.resident_prior = False
    host.site_update(0.1, 0.1)
    host.close()


@pytest.mark.parametrize("B,chunk,streams", [(10, 4, 3), (7, 16, 1), (5, 2, 2)])
def test_resident_session_matches_oracle_and_engine(B, chunk, streams):
    """vdp_cvi_session_*: state uploaded once, per update observation values up and ELBO terms + marginals down.
    Against the oracle (1e-9) and the device engine; the observation values change between updates; the state read
    back at the end (girsanov sites materialised from theta) matches the oracle's."""
    from vi_diffusion_processes_b200.cvi import CviEngine, CviSession
    rng = np.random.default_rng(37)
    N = 700
    grid, idx, y = _problem(rng, B, N, 12, 0.005)
    mk = lambda: CviEngine(S.DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y[..., None], [[0.1]], [1.0], [[1e-2]])
    ses = CviSession(mk(), chunk=chunk, streams=streams)
    dev_eng = mk()
    ref = scalar_d1.BatchedCVISitesSDE1D(S.DRIFT_DW, [4.0, 1.0], 1.0, grid, idx, y, 0.1, 1.0, 1e-2)
    up, down = ses.bytes_per_update()
    assert up == B * 12 * 8 and down == (3 * B + 2 * B * (N + 1)) * 8
    for it in range(4):
        if it == 2:          # new observation values for this update
            y = y + 0.05 * rng.normal(size=y.shape)
            ref.obs_y = y.copy()
            dev_eng.obs_y.copy_(torch.as_tensor(y[..., None]))
            ses.host["obs_y"].copy_(torch.as_tensor(y[..., None]))
        e = ses.site_update(0.5, 0.5)
        d = dev_eng.site_update(0.5, 0.5)
        r = ref.site_update(0.5, 0.5)
        assert rel(e, r) < TOL and rel(e, d) < 1e-11
        assert rel(ses.host["fx_mus"][..., 0], ref.fx_mus) < TOL and rel(ses.host["fx_covs"][..., 0, 0], ref.fx_covs) < TOL
    h = ses.download_state()
    th = scalar_d1.ssm_to_naturals(ref.p["A"], ref.p["b"], ref.p["cQ"], ref.p["mu0"], ref.p["cP0"])
    for got, want, t in ((h["g1"][..., 0], ref.g1, th[0]), (h["gd"][..., 0, 0], ref.gd, th[1]),
                         (h["gs"][..., 0, 0], ref.gs, th[2])):
        scale = max(np.abs(want).max(), 1e-3 * np.abs(t + want).max())
        assert rel(got, want, scale=scale) < TOL
    assert rel(h["d1"][..., 0], ref.d1) < TOL and rel(h["d2"][..., 0, 0], ref.d2) < TOL
    for k in ("qA", "qb", "qcholQ"):
        assert rel(h[k], getattr(dev_eng, k)) < 1e-10, k
    ses.close()


def test_resident_session_without_marginal_download():
    """``site_update(marginals=False)``: only the ELBO terms travel device -> host per update (what the reference's
    trainer reads every iteration); the host marginal buffers stay as they are until ``download_state`` fills them."""
    from vi_diffusion_processes_b200.cvi import CviEngine, CviSession
    rng = np.random.default_rng(41)
    B, N = 6, 900
    grid, idx, y = _problem(rng, B, N, 9, 0.005)
    mk = lambda: CviEngine(S.DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y[..., None], [[0.1]], [1.0], [[1e-2]])
    ses, full = CviSession(mk(), chunk=4, streams=2), CviSession(mk(), chunk=4, streams=2)
    up, down = ses.bytes_per_update(marginals=False)
    assert up == B * 9 * 8 and down == 3 * B * 8
    ses.host["fx_mus"].fill_(-7.0)
    ses.host["fx_covs"].fill_(-7.0)
    for it in range(3):
        e = ses.site_update(0.5, 0.5, marginals=False)
        f = full.site_update(0.5, 0.5)
        assert torch.equal(e, f)
        for k in ("ve", "kl"):
            assert torch.equal(ses.host[k], full.host[k])
    assert bool((ses.host["fx_mus"] == -7.0).all()) and bool((ses.host["fx_covs"] == -7.0).all())
    h = ses.download_state()
    assert torch.equal(h["fx_mus"], full.host["fx_mus"]) and torch.equal(h["fx_covs"], full.host["fx_covs"])
    ses.close()
    full.close()
