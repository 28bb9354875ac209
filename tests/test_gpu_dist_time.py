"""
Time-segment sharding (BASELINE config 5 path): the posterior rebuild of one long sequence split over R ranks
must match the oracle's sequential result.  Ranks are emulated in one process on one GPU (same kernels, same
blobs, all-gather replaced by a list) -- the real multi-process runs are tools/dist_time_run.py and tools/c5_time_sharded.py under torchrun.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import gauss_markov_np as G

pytestmark = pytest.mark.gpu
TOL = 1e-9


def rel(a, b):
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def dev(x):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float64).cuda()


def rand_ssm(rng, N, D):
    A = np.eye(D) * 0.9 + 0.1 * rng.standard_normal((N, D, D)) / np.sqrt(D)
    b = 0.1 * rng.standard_normal((N, D))
    R = rng.standard_normal((N, D, D))
    cQ = np.linalg.cholesky(0.01 * np.eye(D) + 0.002 * R @ G.mT(R))
    R0 = rng.standard_normal((D, D))
    cP0 = np.linalg.cholesky(0.5 * np.eye(D) + 0.1 * R0 @ R0.T)
    return G.StateSpaceModel(rng.standard_normal(D), cP0, A, b, cQ)


@pytest.mark.parametrize("nseg", [None, "2"])
@pytest.mark.parametrize("R", [1, 2, 3, 5])
@pytest.mark.parametrize("D,N", [(1, 6000), (2, 1500), (3, 901)])
def test_time_sharded_posterior_rebuild(vdp_cfg, D, N, R, nseg):
    from vi_diffusion_processes_b200.dist_time import Comm, TimeShardedChain
    vdp_cfg(SWEEPS="tiled")
    if nseg:
        vdp_cfg(NSEG=nseg)
    else:
        vdp_cfg(NSEG=None)
    rng = np.random.default_rng(61 + D + R)
    ssm = rand_ssm(rng, N, D)
    th1, thd, ths = G.ssm_to_naturals(ssm)
    want = G.naturals_to_ssm_params(th1, thd, ths)           # As, offsets, chol_P0, chol_Q, mu0
    wm, wS = ssm.marginals
    chain = TimeShardedChain(N, D, Comm.emulated(R))
    l1, ld, ls = [], [], []
    for r in range(R):
        s, e = chain.parts[r]
        n0, n1 = chain.owned_nodes(r)
        l1.append(dev(th1[n0:n1]))
        ld.append(dev(thd[n0:n1]))
        ls.append(dev(ths[s:e]))
    out = chain.naturals_to_ssm(l1, ld, ls)
    chain.check_status()
    for r in range(R):
        s, e = chain.parts[r]
        assert rel(out[r]["A"], want[0][s:e]) < TOL
        assert rel(out[r]["b"], want[1][s:e]) < 1e-8
        assert rel(out[r]["cholQ"], want[3][s:e]) < TOL
    assert rel(out[0]["mu0"], want[4]) < 1e-8 and rel(out[0]["cholP0"], want[2]) < TOL
    # marginals of the SSM we started from, sharded the same way
    A = [dev(ssm.state_transitions[s:e]) for s, e in chain.parts]
    b = [dev(ssm.state_offsets[s:e]) for s, e in chain.parts]
    cQ = [dev(ssm.cholesky_process_covariances[s:e]) for s, e in chain.parts]
    mu0 = [dev(ssm.initial_mean)] + [None] * (R - 1)
    cP0 = [dev(ssm.cholesky_initial_covariance)] + [None] * (R - 1)
    marg = chain.ssm_marginals(A, b, cQ, mu0, cP0)
    chain.check_status()
    for r in range(R):
        n0, n1 = chain.owned_nodes(r)
        assert rel(marg[r][0], wm[n0:n1]) < TOL and rel(marg[r][1], wS[n0:n1]) < TOL


def test_time_partition_covers_the_grid():
    from vi_diffusion_processes_b200.dist_time import time_partition
    for N, R in [(10, 3), (7, 7), (10 ** 7, 8), (1001, 2)]:
        parts = time_partition(N, R)
        assert parts[0][0] == 0 and parts[-1][1] == N
        assert all(parts[i][1] == parts[i + 1][0] for i in range(R - 1))
        assert max(e - s for s, e in parts) - min(e - s for s, e in parts) <= 1
    with pytest.raises(ValueError):
        time_partition(3, 5)


def test_time_sharded_rebuild_long_chain_8_ranks_equals_single_rank():
    """N = 10^6 transitions, D = 3, 8 emulated ranks against the single-rank CUDA result (the NumPy oracle walks the
    chain block by block in Python and is too slow at this length; the single-rank kernels are pinned to it above)."""
    from vi_diffusion_processes_b200.dist_time import Comm, TimeShardedChain
    N, D, R = 1_000_000, 3, 8
    rng = np.random.default_rng(7)
    ssm = rand_ssm(rng, N, D)
    th1, thd, ths = G.ssm_to_naturals(ssm)
    res = {}
    for nr in (1, R):
        chain = TimeShardedChain(N, D, Comm.emulated(nr))
        l1, ld, ls = [], [], []
        for r in range(nr):
            s, e = chain.parts[r]
            n0, n1 = chain.owned_nodes(r)
            l1.append(dev(th1[n0:n1]))
            ld.append(dev(thd[n0:n1]))
            ls.append(dev(ths[s:e]))
        out = chain.naturals_to_ssm(l1, ld, ls)
        marg = chain.ssm_marginals([o["A"] for o in out], [o["b"] for o in out], [o["cholQ"] for o in out],
                                   [o["mu0"] for o in out], [o["cholP0"] for o in out])
        chain.check_status()
        res[nr] = {k: torch.cat([o[k] for o in out]) for k in ("A", "b", "cholQ")}
        res[nr]["m"] = torch.cat([m for m, _ in marg])
        res[nr]["S"] = torch.cat([S for _, S in marg])
    for k in ("A", "b", "cholQ", "m", "S"):
        a, b = res[R][k], res[1][k]
        assert a.shape == b.shape
        assert float((a - b).abs().max()) <= 1e-11 * float(b.abs().max()), k
    # and the marginal means satisfy the forward recursion of the model they were built from
    m = res[R]["m"].cpu().numpy()
    assert np.abs(m[1:] - (np.einsum("nij,nj->ni", ssm.state_transitions, m[:-1]) + ssm.state_offsets)).max() < 1e-9


@pytest.mark.parametrize("N,D,R", [(1_000_000, 3, 8), (30_000, 2, 3), (50_000, 1, 5), (7, 3, 2)])
def test_information_form_rebuild_recovers_the_model_the_naturals_came_from(N, D, R):
    """naturals of a random state-space model -> VDP_SWEEP_INFO -> the same A, b, chol Q, mu0, chol P0 back (the
    information-form sweep inverts ssm_to_naturals exactly), and VDP_SWEEP_MOMENTS on them -> the marginals of the
    three-sweep route; 10^6 transitions over 8 emulated ranks down to stretches shorter than a tile."""
    from vi_diffusion_processes_b200.dist_time import Comm, TimeShardedChain
    rng = np.random.default_rng(11 + D)
    ssm = rand_ssm(rng, N, D)
    th1, thd, ths = G.ssm_to_naturals(ssm)
    chain = TimeShardedChain(N, D, Comm.emulated(R))
    l1, ld, ls = [], [], []
    for r in range(R):
        s, e = chain.parts[r]
        n0, n1 = chain.owned_nodes(r)
        l1.append(dev(th1[n0:n1]))
        ld.append(dev(thd[n0:n1]))
        ls.append(dev(ths[s:e]))
    out = chain.naturals_to_ssm_info(l1, ld, ls)
    marg = chain.ssm_moments([o["A"] for o in out], [o["b"] for o in out], [o["cholQ"] for o in out],
                             [o["mu0"] for o in out], [o["cholP0"] for o in out])
    ref = chain.naturals_to_ssm(l1, ld, ls, want_cov=True)
    ref_m = chain.ssm_means([o["A"] for o in ref], [o["b"] for o in ref], [o["mu0"] for o in ref])
    chain.check_status()
    got = {k: torch.cat([o[k] for o in out]).cpu().numpy() for k in ("A", "b", "cholQ")}
    want = {"A": ssm.state_transitions, "b": ssm.state_offsets, "cholQ": ssm.cholesky_process_covariances}
    for k in got:
        assert np.abs(got[k] - want[k]).max() <= 1e-10 * np.abs(want[k]).max(), k
    assert np.abs(out[0]["mu0"].cpu().numpy() - ssm.initial_mean).max() <= 1e-11
    assert np.abs(out[0]["cholP0"].cpu().numpy() - ssm.cholesky_initial_covariance).max() <= 1e-11
    m, S = torch.cat([x for x, _ in marg]), torch.cat([x for _, x in marg])
    rm, rS = torch.cat(ref_m), torch.cat([o["S"] for o in ref])
    assert m.shape == rm.shape and S.shape == rS.shape
    assert float((m - rm).abs().max()) <= 1e-10 * max(float(rm.abs().max()), 1.0)
    assert float((S - rS).abs().max()) <= 1e-10 * float(rS.abs().max())
