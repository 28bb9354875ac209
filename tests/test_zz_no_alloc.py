"""
vdp_ssm_marginals_static: the allocation-free variant of vdp_ssm_marginals (scratch in a static __device__ array of
the library; no counterpart in the reference).  Called through the C
ABI (ctypes) and compared with the existing entry point on the same inputs, two calls in a row on different inputs
(the array keeps the previous call's contents).  The file sorts last on purpose.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

TOL = 1e-9   # the parity tolerance of tests/test_gpu_parity.py (both entries run the same kernels: expect 0)


def _rand_ssm(rng, B, N, D):
    A = np.eye(D) * 0.9 + 0.1 * rng.standard_normal((B, N, D, D))
    b = 0.1 * rng.standard_normal((B, N, D))
    R = rng.standard_normal((B, N, D, D))
    cQ = np.linalg.cholesky(0.01 * np.eye(D) + 0.002 * R @ np.swapaxes(R, -1, -2))
    R0 = rng.standard_normal((B, D, D))
    cP0 = np.linalg.cholesky(0.5 * np.eye(D) + 0.1 * R0 @ np.swapaxes(R0, -1, -2))
    mu0 = rng.standard_normal((B, D))
    return [torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float64).cuda() for x in (A, b, cQ, mu0, cP0)]


def _p(t):
    return C.c_void_p(t.data_ptr())


def _both(lib, args, B, N, D):
    """(m, Sd) from the workspace entry point and from the static-scratch one, raw C-ABI calls."""
    from vi_diffusion_processes_b200._lib import check
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for static in (False, True):
        m = torch.full((B, N + 1, D), float("nan"), dtype=torch.float64, device="cuda")
        Sd = torch.full((B, N + 1, D, D), float("nan"), dtype=torch.float64, device="cuda")
        status = torch.zeros(2, dtype=torch.int64, device="cuda")
        ptrs = [_p(t) for t in args] + [_p(m), _p(Sd)]
        if static:
            rc = lib.vdp_ssm_marginals_static(*ptrs, B, N, D, _p(status), stream)
            check(rc, "vdp_ssm_marginals_static")
        else:
            nbytes = lib.vdp_ssm_marginals_workspace(B, N, D)
            ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
            rc = lib.vdp_ssm_marginals(*ptrs, _p(ws), nbytes, B, N, D, _p(status), stream)
            check(rc, "vdp_ssm_marginals")
        torch.cuda.synchronize()
        assert int(status[0]) == 0
        outs.append((m.cpu().numpy(), Sd.cpu().numpy()))
    return outs


@pytest.mark.parametrize("D,B,N", [(1, 4, 4001), (2, 3, 1500), (3, 2, 700), (1, 300, 40), (2, 1, 1)])
def test_static_scratch_variant_equals_workspace_variant(D, B, N):
    from vi_diffusion_processes_b200 import _lib
    lib = _lib.load()
    assert lib.vdp_static_scratch_bytes() == 16 << 20
    rng = np.random.default_rng(1000 * D + N)
    for call in range(2):   # two calls in a row, different inputs: the second sees the first one's scratch contents
        args = _rand_ssm(rng, B, N, D)
        (m0, S0), (m1, S1) = _both(lib, args, B, N, D)
        assert np.isfinite(m1).all() and np.isfinite(S1).all()
        assert np.abs(m1 - m0).max() <= TOL * np.abs(m0).max(), (call, np.abs(m1 - m0).max())
        assert np.abs(S1 - S0).max() <= TOL * np.abs(S0).max(), (call, np.abs(S1 - S0).max())


def test_static_scratch_variant_rejects_sizes_beyond_the_array():
    from vi_diffusion_processes_b200 import _lib
    lib = _lib.load()
    D, B, N = 1, 1, 600_000          # 8 (2 N + 1) B D^2 = 9.6 MB > the 8 MiB the factor may take
    args = _rand_ssm(np.random.default_rng(0), 1, 4, D)   # never read: the call must refuse before launching
    m = torch.zeros(8, dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.vdp_ssm_marginals_static(*[_p(t) for t in args], _p(m), _p(m), B, N, D, None, stream)
    assert rc == -3   # VDP_E_WORKSPACE
    msg = lib.vdp_last_error().decode()
    assert "vdp_ssm_marginals_static" in msg and "limit" in msg
    torch.cuda.synchronize()
