"""
The two-sweep rebuild (information form from the bottom, moments from the top) restated in NumPy
(oracle/info_form_np.py) against the reference-order restatement (oracle/gauss_markov_np.py: Cholesky from the top,
inverse band, A from the covariances -- ssm_gaussian_transformations.py:333-511, state_space_model.py:231-262), and
the algebra of its scan elements -- in NumPy and, compiled for the host, in the C++ the kernels use
(csrc/scan_elems.cuh InfoOps).  CPU only.
"""
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import gauss_markov_np as G
from oracle import info_form_np as I

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rand_ssm(rng, batch, N, D):
    A = 0.6 * np.eye(D) + 0.2 * rng.normal(size=batch + (N, D, D)) / np.sqrt(D)
    b = rng.normal(size=batch + (N, D))
    R = rng.normal(size=batch + (N, D, D))
    cQ = np.linalg.cholesky(0.05 * np.eye(D) + 0.02 * R @ G.mT(R))
    R0 = rng.normal(size=batch + (D, D))
    cP0 = np.linalg.cholesky(0.5 * np.eye(D) + 0.1 * R0 @ G.mT(R0))
    return G.StateSpaceModel(rng.normal(size=batch + (D,)), cP0, A, b, cQ)


@pytest.mark.parametrize("batch", [(), (3,)])
@pytest.mark.parametrize("D,N", [(1, 1), (1, 40), (2, 25), (3, 1), (3, 60)])
def test_information_form_equals_the_reference_order(batch, D, N):
    rng = np.random.default_rng(100 * D + N)
    ssm = _rand_ssm(rng, batch, N, D)
    th = G.ssm_to_naturals(ssm)
    want = G.naturals_to_ssm_params(*th)
    got = I.naturals_to_ssm_params_info(*th)
    for name, g, w in zip(("A", "b", "chol_P0", "chol_Q", "mu0"), got, want):
        assert g.shape == w.shape, name
        assert np.abs(g - w).max() <= 1e-11 * max(np.abs(w).max(), 1.0), name
    # ... and both give back the model the naturals were made from
    for g, w in zip(got, (ssm.state_transitions, ssm.state_offsets, ssm.cholesky_initial_covariance,
                          ssm.cholesky_process_covariances, ssm.initial_mean)):
        assert np.abs(g - w).max() <= 1e-11 * max(np.abs(w).max(), 1.0)


@pytest.mark.parametrize("D,N", [(1, 30), (2, 20), (3, 45)])
def test_moment_recursion_equals_marginals_through_the_precision(D, N):
    rng = np.random.default_rng(7 * D + N)
    ssm = _rand_ssm(rng, (2,), N, D)
    m, S = I.moments(ssm.state_transitions, ssm.state_offsets, ssm.cholesky_initial_covariance,
                     ssm.cholesky_process_covariances, ssm.initial_mean)
    wm, wS = ssm.marginals                                    # state_space_model.py:231-262 (precision, inverse band)
    assert np.abs(m - wm).max() <= 1e-11 * np.abs(wm).max()
    assert np.abs(S - wS).max() <= 1e-11 * np.abs(wS).max()


@pytest.mark.parametrize("D", [1, 2, 3])
def test_scan_elements_compose(D):
    """any bracketing of the per-step maps gives the state the step-by-step recursion reaches"""
    rng = np.random.default_rng(5 + D)
    N = 23
    ssm = _rand_ssm(rng, (), N, D)
    t1, td, ts = G.ssm_to_naturals(ssm)
    Kd = G.symmetrise_from_lower(-2.0 * td)
    J, h = Kd[N], t1[N]
    for k in range(N - 1, -1, -1):
        J, h = I.info_apply(I.info_step(ts[k], Kd[k], t1[k]), J, h)
    es = [I.info_step(ts[k], Kd[k], t1[k]) for k in range(N - 1, -1, -1)]
    while len(es) > 1:                                        # balanced tree
        es = [I.info_combine(es[i], es[i + 1]) if i + 1 < len(es) else es[i] for i in range(0, len(es), 2)]
    J2, h2 = I.info_apply(es[0], Kd[N], t1[N])
    e = I.info_step(ts[N - 1], Kd[N - 1], t1[N - 1])          # left fold
    for k in range(N - 2, -1, -1):
        e = I.info_combine(e, I.info_step(ts[k], Kd[k], t1[k]))
    J3, h3 = I.info_apply(e, Kd[N], t1[N])
    for a, b in ((J2, J), (h2, h), (J3, J), (h3, h)):
        assert np.abs(a - b).max() <= 1e-11 * max(np.abs(b).max(), 1.0)
    P0 = np.linalg.inv(J)
    assert np.abs(P0 @ h - ssm.initial_mean).max() <= 1e-11


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not on PATH")
def test_cxx_scan_elements_on_the_host(tmp_path):
    """csrc/scan_elems.cuh InfoOps (the functions the kernels call are __host__ __device__) compiled for the host:
    tree-ordered composition with `push` against the step-by-step recursion, and the transitions it yields against
    the state-space model the naturals were built from, D = 1, 2, 3"""
    exe = tmp_path / "info_ops_host"
    subprocess.run(["nvcc", "-std=c++17", "-O1", "-x", "cu", "-o", str(exe),
                    os.path.join(ROOT, "tools", "exp", "info_ops_host.cu")], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    assert "PASS" in out, out
