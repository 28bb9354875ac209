"""
Error behaviour of the host layer on the GPU path: a non positive-definite update must raise (the reference fails
inside tf.linalg.cholesky / cholesky_band), CUDA-graph replays must follow a changed prior on x0, observation indices
outside the grid are refused, and ops / engines run on the device of their tensors.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _engine(rng, N=400, n_obs=10, **kw):
    from oracle.sde_np import DRIFT_DW
    from vi_diffusion_processes_b200.cvi import CviEngine
    grid = np.round(np.arange(N + 1) * 0.005, 3)
    idx = np.sort(rng.choice(N + 1, n_obs, replace=False))[None]
    y = rng.normal(size=(1, n_obs, 1)) * 0.3 + 1.0
    return CviEngine(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, idx, y, [[0.1]], [1.0], [[1e-2]], **kw)


def test_trainer_raises_on_a_non_pd_update():
    """A girsanov step that leaves the positive-definite cone gives NaN everywhere; the trainer loop must stop with
    an exception instead of running to max_itr on NaN comparisons."""
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.markovflow.likelihoods import MultivariateGaussian
    from vi_diffusion_processes_b200.markovflow.models import CVISitesSDE
    from vi_diffusion_processes_b200.markovflow.sde import DoubleWellSDE
    from vi_diffusion_processes_b200.markovflow.sde.sde_utils import Gaussian
    from vi_diffusion_processes_b200.trainers import CVISitesTrainer
    rng = np.random.default_rng(5)
    N = 300
    grid = np.round(np.arange(N + 1) * 0.01, 6)
    pick = np.sort(rng.choice(N + 1, 20, replace=False))
    y = rng.normal(size=(20, 1)) * 0.2 + 1.0
    model = CVISitesSDE(DoubleWellSDE(q=np.ones((1, 1))), grid, (grid[pick[::2]], y[::2]),
                        MultivariateGaussian(0.1 * np.eye(1)),
                        prior_initial_state=Gaussian(np.ones(1), 0.01 * np.eye(1)))
    # sites that make the posterior precision indefinite: + 1e6 on the diagonal natural parameter (precision -2 e6)
    model._engine.gd += 1e6
    trainer = CVISitesTrainer(model, (grid[pick[1::2]], y[1::2]), max_itr=2, max_itr_sites_optim=3,
                              girsanov_sites_lr=0.5, data_sites_lr=0.5)
    with pytest.raises(_lib.VdpError, match="positive-definite"):
        trainer.optimize()


def test_graph_replay_follows_a_changed_prior_initial_state():
    """site_update_graphed after set_prior_initial_state must use the new p(x0) in the KL_0 gradient (the captured
    launch holds x0_mu / x0_cov by value): replay == direct call, bitwise."""
    rng = np.random.default_rng(9)
    a, b = _engine(np.random.default_rng(9), lazy_sites=False), _engine(np.random.default_rng(9), lazy_sites=False)
    for it in range(5):
        if it == 2:
            for e in (a, b):
                e.set_prior_initial_state([0.7], [[0.05]])
        ea = a.site_update(0.5, 0.5).clone()
        eb = b.site_update_graphed(0.5, 0.5).clone()
        assert torch.equal(ea, eb), it
    for name in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    # and the change is not a no-op: an engine that keeps the old prior differs
    c = _engine(np.random.default_rng(9), lazy_sites=False)
    for it in range(5):
        ec = c.site_update(0.5, 0.5)
    assert not torch.equal(ec, ea)
    del rng


def test_observation_indices_outside_the_grid_are_refused():
    from oracle.sde_np import DRIFT_DW
    from vi_diffusion_processes_b200 import _lib
    from vi_diffusion_processes_b200.cvi import CviEngine
    from vi_diffusion_processes_b200.vgp import VgpEngine
    grid = np.arange(11) * 0.01
    y = np.zeros((1, 2, 1))
    for bad in ([[-1, 3]], [[2, 11]]):
        for cls in (CviEngine, VgpEngine):
            with pytest.raises(_lib.VdpError, match="observation grid indices"):
                cls(DRIFT_DW, [4.0, 1.0], np.ones((1, 1)), grid, np.asarray(bad), y, [[0.1]], [1.0], [[1e-2]])


def test_non_pd_spectral_density_is_reported_by_linearize():
    """k_linearize folds the Cholesky of q (and of q dt) into the status word."""
    from oracle.sde_np import DRIFT_OU
    from vi_diffusion_processes_b200 import ops
    B, N = 1, 8
    grid = torch.arange(N + 1, dtype=torch.float64, device="cuda") * 0.01
    m = torch.zeros(B, N, 1, dtype=torch.float64, device="cuda")
    S = torch.ones(B, N, 1, 1, dtype=torch.float64, device="cuda")
    st = ops.new_status("cuda")
    ops.linearize_to_ssm(DRIFT_OU, [1.0], -np.ones((1, 1)), grid, m, S, status=st)
    assert ops.read_status(st)[0] == 1


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_ops_and_engines_run_on_the_device_of_their_tensors():
    from vi_diffusion_processes_b200 import ops
    rng = np.random.default_rng(3)
    torch.cuda.set_device(0)
    Kd = torch.as_tensor(np.tile(2.0 * np.eye(2), (2, 50, 1, 1))).to("cuda:1")
    Ks = torch.as_tensor(0.1 * rng.normal(size=(2, 49, 2, 2))).to("cuda:1")
    Ld, Ls = ops.btd_cholesky(Kd, Ks)
    assert Ld.device.index == 1 and torch.isfinite(Ld).all()
    e1 = _engine(np.random.default_rng(3), device="cuda:1")
    e0 = _engine(np.random.default_rng(3), device="cuda:0")
    for _ in range(2):
        a, b = e1.site_update(0.5, 0.5), e0.site_update(0.5, 0.5)
    assert torch.equal(a.cpu(), b.cpu())
    assert torch.cuda.current_device() == 0
