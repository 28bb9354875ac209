"""
Thin torch-tensor front end of the C ABI: every function takes float64 CUDA tensors in the
reference's layouts (leading batch axis), allocates the outputs with torch, and launches the CUDA
kernels on torch's current stream.  No arithmetic happens in Python and nothing falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import CviProblem, check, dvec

F64 = torch.float64


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _on_tensor_device(fn):
    """Run `fn` with the device of its first CUDA tensor argument current: the library launches on the CURRENT
    device, so an op on tensors of cuda:1 while cuda:0 is current would otherwise run on the wrong GPU."""
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, torch.Tensor) and a.is_cuda:
                if a.device.index == torch.cuda.current_device():
                    break
                with torch.cuda.device(a.device):
                    return fn(*args, **kwargs)
        return fn(*args, **kwargs)
    return wrapped


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _req(*tensors):
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.VdpError("vdp_b200 ops need CUDA tensors (there is no CPU path)")
        if t.dtype not in (F64, torch.int64):
            raise _lib.VdpError(f"expected float64/int64, got {t.dtype}")
        if not t.is_contiguous():
            raise _lib.VdpError("tensors must be contiguous")


def new_status(device):
    return torch.zeros(2, dtype=torch.int64, device=device)


def read_status(status):
    """(code, flat index) of the first non-positive-definite block seen since the word was zeroed."""
    out = (C.c_int64 * 2)()
    _lib.load().vdp_read_status(_p(status), out, _stream())
    return int(out[0]), int(out[1])


def _bnd(diag):
    """[B, n, D, D] -> B, n, D (a missing batch axis is an error: callers add it)."""
    assert diag.dim() == 4 and diag.shape[-1] == diag.shape[-2], diag.shape
    return diag.shape[0], diag.shape[1], diag.shape[2]


# ------------------------------------------------------------------ op level


def btd_cholesky(Kd, Ks=None, status=None):
    _req(Kd, Ks, status)
    B, n, D = _bnd(Kd)
    Ld = torch.empty_like(Kd)
    Ls = None if Ks is None else torch.empty_like(Ks)
    check(_lib.load().vdp_btd_cholesky(_p(Kd), _p(Ks), _p(Ld), _p(Ls), B, n, D, _p(status), _stream()),
          "vdp_btd_cholesky")
    return Ld, Ls


def btd_inverse_band(Ld, Ls=None, want_sub=True):
    _req(Ld, Ls)
    B, n, D = _bnd(Ld)
    Sd = torch.empty_like(Ld)
    Ss = torch.empty_like(Ls) if (Ls is not None and want_sub) else None
    check(_lib.load().vdp_btd_inverse_band(_p(Ld), _p(Ls), _p(Sd), _p(Ss), B, n, D, _stream()),
          "vdp_btd_inverse_band")
    return Sd, Ss


def btd_inverse_band_jvp(Ld, Ls, dKd, dKs):
    """d(Sigma_kk), d(Sigma_{k+1,k}) of d(K^-1) = -K^-1 dK K^-1 for K = L L^T and a symmetric block-tridiagonal
    direction dK (lower triangle of the diagonal blocks read)."""
    _req(Ld, Ls, dKd, dKs)
    B, n, D = _bnd(Ld)
    dSd, dSs = torch.empty_like(Ld), torch.empty_like(Ls)
    check(_lib.load().vdp_btd_inverse_band_jvp(_p(Ld), _p(Ls), _p(dKd), _p(dKs), _p(dSd), _p(dSs), B, n, D, _stream()),
          "vdp_btd_inverse_band_jvp")
    return dSd, dSs


def btd_solve(Ld, Ls, rhs, transpose=False):
    _req(Ld, Ls, rhs)
    B, n, D = _bnd(Ld)
    x = torch.empty_like(rhs)
    check(_lib.load().vdp_btd_solve(_p(Ld), _p(Ls), _p(rhs), _p(x), int(transpose), B, n, D, _stream()),
          "vdp_btd_solve")
    return x


def btd_matvec(Md, Ms, x, symmetric, transpose=False):
    _req(Md, Ms, x)
    B, n, D = _bnd(Md)
    y = torch.empty_like(x)
    check(_lib.load().vdp_btd_matvec(_p(Md), _p(Ms), _p(x), _p(y), int(symmetric), int(transpose), B, n, D,
                                     _stream()), "vdp_btd_matvec")
    return y


def btd_logdet(Ld):
    _req(Ld)
    B, n, D = _bnd(Ld)
    out = torch.empty(B, dtype=F64, device=Ld.device)
    check(_lib.load().vdp_btd_logdet(_p(Ld), _p(out), B, n, D, _stream()), "vdp_btd_logdet")
    return out


def btd_block_to_band(Md, Ms=None):
    """block_tri_diag.py:206-237 (`as_band`): [B,n,D,D] (+ [B,n-1,D,D]) -> [B, R, n*D], R = 2D or D."""
    _req(Md, Ms)
    B, n, D = _bnd(Md)
    band = torch.empty((B, (2 * D if Ms is not None else D), n * D), dtype=F64, device=Md.device)
    check(_lib.load().vdp_btd_block_to_band(_p(Md), _p(Ms), _p(band), B, n, D, _stream()), "vdp_btd_block_to_band")
    return band


def btd_band_to_block(band, D, symmetric, with_sub=True):
    _req(band)
    B, R, cols = band.shape
    n = cols // D
    assert R == (2 * D if with_sub else D) and cols == n * D, band.shape
    Md = torch.empty((B, n, D, D), dtype=F64, device=band.device)
    Ms = torch.empty((B, n - 1, D, D), dtype=F64, device=band.device) if with_sub else None
    check(_lib.load().vdp_btd_band_to_block(_p(band), _p(Md), _p(Ms), int(symmetric), B, n, D, _stream()),
          "vdp_btd_band_to_block")
    return Md, Ms


def btd_conditional_precisions(As, Kd, Ks):
    """ssm_gaussian_transformations.py:470-483: block diagonal of A_inv^-T P."""
    _req(As, Kd, Ks)
    B, n, D = _bnd(Kd)
    out = torch.empty_like(Kd)
    check(_lib.load().vdp_btd_conditional_precisions(_p(As), _p(Kd), _p(Ks), _p(out), B, n, D, _stream()),
          "vdp_btd_conditional_precisions")
    return out


# ------------------------------------------------------------------ fused level


def gh_expect(drift_id, params, m, S, H=10, full_jacobian=False, status=None):
    """m [count, D], S [count, D, D] -> E[f] [count, D], E[grad f] ([count, D, D] or [count, D])."""
    _req(m, S, status)
    count, D = m.shape
    Ef = torch.empty_like(m)
    EJ = torch.empty((count, D, D) if full_jacobian else (count, D), dtype=F64, device=m.device)
    check(_lib.load().vdp_gh_expect(int(drift_id), dvec(params, 16), _p(m), _p(S), int(H), _p(Ef), _p(EJ), count, D,
                                    _p(status), _stream()), "vdp_gh_expect")
    return Ef, EJ


def linearize_to_ssm(drift_id, params, q, time_grid, m, S, clip=None, status=None, want_unclipped=False):
    """m [B,N,D], S [B,N,D,D] on grid points 1..N -> A [B,N,D,D], b [B,N,D], cholQ [B,N,D,D]."""
    _req(time_grid, m, S, status)
    B, N, D = m.shape
    A = torch.empty((B, N, D, D), dtype=F64, device=m.device)
    b = torch.empty((B, N, D), dtype=F64, device=m.device)
    cholQ = torch.empty_like(A)
    A_lin = torch.empty_like(A) if want_unclipped else None
    b_lin = torch.empty_like(b) if want_unclipped else None
    lo, hi = (clip if clip is not None else (0.0, 0.0))
    check(_lib.load().vdp_linearize_to_ssm(int(drift_id), dvec(params, 16), dvec(q.reshape(-1), D * D),
                                           _p(time_grid), _p(m), _p(S), _p(A), _p(b), _p(cholQ), _p(A_lin), _p(b_lin),
                                           int(clip is not None), float(lo), float(hi), B, N, D, _p(status),
                                           _stream()), "vdp_linearize_to_ssm")
    return (A, b, cholQ, A_lin, b_lin) if want_unclipped else (A, b, cholQ)


def ssm_to_naturals(A, b, cholQ, mu0, cholP0):
    _req(A, b, cholQ, mu0, cholP0)
    B, N, D = b.shape
    th1 = torch.empty((B, N + 1, D), dtype=F64, device=A.device)
    thd = torch.empty((B, N + 1, D, D), dtype=F64, device=A.device)
    ths = torch.empty((B, N, D, D), dtype=F64, device=A.device)
    check(_lib.load().vdp_ssm_to_naturals(_p(A), _p(b), _p(cholQ), _p(mu0), _p(cholP0), _p(th1), _p(thd), _p(ths), B,
                                          N, D, _stream()), "vdp_ssm_to_naturals")
    return th1, thd, ths


def _workspace(nbytes, device):
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def naturals_to_ssm(th1, thd, ths, status=None):
    _req(th1, thd, ths, status)
    B, n, D = th1.shape
    N = n - 1
    dev = th1.device
    A = torch.empty((B, N, D, D), dtype=F64, device=dev)
    b = torch.empty((B, N, D), dtype=F64, device=dev)
    cholQ = torch.empty_like(A)
    mu0 = torch.empty((B, D), dtype=F64, device=dev)
    cholP0 = torch.empty((B, D, D), dtype=F64, device=dev)
    lib = _lib.load()
    nbytes = lib.vdp_naturals_to_ssm_workspace(B, N, D)
    ws = _workspace(nbytes, dev)
    check(lib.vdp_naturals_to_ssm(_p(th1), _p(thd), _p(ths), _p(A), _p(b), _p(cholQ), _p(mu0), _p(cholP0), _p(ws),
                                  nbytes, B, N, D, _p(status), _stream()), "vdp_naturals_to_ssm")
    return A, b, cholP0, cholQ, mu0   # reference return order (ssm_gaussian_transformations.py:511)


def ssm_marginals(A, b, cholQ, mu0, cholP0, status=None):
    _req(A, b, cholQ, mu0, cholP0, status)
    B, N, D = b.shape
    dev = A.device
    m = torch.empty((B, N + 1, D), dtype=F64, device=dev)
    Sd = torch.empty((B, N + 1, D, D), dtype=F64, device=dev)
    lib = _lib.load()
    nbytes = lib.vdp_ssm_marginals_workspace(B, N, D)
    ws = _workspace(nbytes, dev)
    check(lib.vdp_ssm_marginals(_p(A), _p(b), _p(cholQ), _p(mu0), _p(cholP0), _p(m), _p(Sd), _p(ws), nbytes, B, N, D,
                                _p(status), _stream()), "vdp_ssm_marginals")
    return m, Sd


def ssm_kl_divergence(ssm1, ssm2, status=None):
    """StateSpaceModel.kl_divergence (state_space_model.py:528-593); ssm = (A, b, cholQ, mu0, cholP0) -> [B]."""
    _req(*ssm1, *ssm2, status)
    B, N, D = ssm1[0].shape[0], ssm1[0].shape[1], ssm1[0].shape[-1]
    lib = _lib.load()
    nbytes = lib.vdp_ssm_kl_divergence_workspace(B, N, D)
    ws = _workspace(nbytes, ssm1[0].device)
    kl = torch.empty(B, dtype=F64, device=ssm1[0].device)
    check(lib.vdp_ssm_kl_divergence(*[_p(t) for t in ssm1], *[_p(t) for t in ssm2], _p(kl), _p(ws), nbytes, B, N, D,
                                    _p(status), _stream()), "vdp_ssm_kl_divergence")
    return kl


def ssm_log_pdf(A, b, cholQ, mu0, cholP0, states):
    """StateSpaceModel.log_pdf (state_space_model.py:485-526); states [S, B, N+1, D] -> [S, B]."""
    _req(A, b, cholQ, mu0, cholP0, states)
    B, N, D = A.shape[0], A.shape[1], A.shape[-1]
    S = states.shape[0]
    assert states.shape == (S, B, N + 1, D), states.shape
    lib = _lib.load()
    nbytes = lib.vdp_ssm_log_pdf_workspace(S, B, N)
    ws = _workspace(nbytes, A.device)
    out = torch.empty((S, B), dtype=F64, device=A.device)
    check(lib.vdp_ssm_log_pdf(_p(A), _p(b), _p(cholQ), _p(mu0), _p(cholP0), _p(states), _p(out), _p(ws), nbytes, S, B,
                              N, D, _stream()), "vdp_ssm_log_pdf")
    return out


def ssm_sample(A, b, cholQ, mu0, cholP0, num_samples, eps=None, seed=0):
    """StateSpaceModel.sample (state_space_model.py:298-324) -> [S, B, N+1, D]; eps [S, B, N+1, D] standard normals
    to use instead of the device generator (Philox, `seed`)."""
    _req(A, b, cholQ, mu0, cholP0, eps)
    B, N, D = A.shape[0], A.shape[1], A.shape[-1]
    S = int(num_samples)
    if eps is not None:
        assert eps.shape == (S, B, N + 1, D), eps.shape
    out = torch.empty((S, B, N + 1, D), dtype=F64, device=A.device)
    check(_lib.load().vdp_ssm_sample(_p(A), _p(b), _p(cholQ), _p(mu0), _p(cholP0), _p(eps), int(seed), _p(out), S, B,
                                     N, D, _stream()), "vdp_ssm_sample")
    return out


def ssm_to_expectations(A, m, Sd):
    _req(A, m, Sd)
    B, n, D = m.shape
    e1 = torch.empty_like(m)
    ed = torch.empty_like(Sd)
    es = torch.empty_like(A)
    check(_lib.load().vdp_ssm_to_expectations(_p(A), _p(m), _p(Sd), _p(e1), _p(ed), _p(es), B, n - 1, D, _stream()),
          "vdp_ssm_to_expectations")
    return e1, ed, es


def expectations_to_ssm(e1, ed, es, status=None):
    _req(e1, ed, es, status)
    B, n, D = e1.shape
    N = n - 1
    dev = e1.device
    A = torch.empty((B, N, D, D), dtype=F64, device=dev)
    b = torch.empty((B, N, D), dtype=F64, device=dev)
    cholQ = torch.empty_like(A)
    mu0 = torch.empty((B, D), dtype=F64, device=dev)
    cholP0 = torch.empty((B, D, D), dtype=F64, device=dev)
    check(_lib.load().vdp_expectations_to_ssm(_p(e1), _p(ed), _p(es), _p(A), _p(b), _p(cholQ), _p(mu0), _p(cholP0), B,
                                              N, D, _p(status), _stream()), "vdp_expectations_to_ssm")
    return A, b, cholP0, cholQ, mu0


def cvi_kl_and_grad(drift_id, params, q, dt, time_grid, x0_mu, x0_cov, A, m, Sd, status=None):
    _req(time_grid, A, m, Sd, status)
    B, n, D = m.shape
    N = n - 1
    dev = m.device
    kl = torch.empty(B, dtype=F64, device=dev)
    g1 = torch.empty_like(m)
    gd = torch.empty_like(Sd)
    gs = torch.empty_like(A)
    lib = _lib.load()
    nbytes = lib.vdp_cvi_kl_workspace(B, N, D)
    ws = _workspace(nbytes, dev)
    check(lib.vdp_cvi_kl_and_grad(int(drift_id), dvec(params, 16), dvec(q.reshape(-1), D * D), float(dt),
                                  _p(time_grid), dvec(x0_mu.reshape(-1), D), dvec(x0_cov.reshape(-1), D * D), _p(A),
                                  _p(m), _p(Sd), _p(kl), _p(g1), _p(gd), _p(gs), _p(ws), nbytes, B, N, D, _p(status),
                                  _stream()), "vdp_cvi_kl_and_grad")
    return kl, (g1, gd, gs)


def linear_drift_to_ssm(A, b, q, time_grid, status=None):
    """LinearDrift.to_ssm (drift.py:64-116): A [B,N,D,D], b [B,N,D] -> SSM transitions, offsets, chol Q."""
    _req(A, b, time_grid, status)
    B, N, D = b.shape
    sA, sb, scQ = torch.empty_like(A), torch.empty_like(b), torch.empty_like(A)
    check(_lib.load().vdp_linear_drift_to_ssm(_p(A), _p(b), dvec(q.reshape(-1), D * D), _p(time_grid), _p(sA), _p(sb),
                                              _p(scQ), B, N, D, _p(status), _stream()), "vdp_linear_drift_to_ssm")
    return sA, sb, scQ


def sq_drift_difference(drift_id, params, q, dt, A, b, m, S, want_grads=False, status=None):
    """sde_utils.py:182-249 with the linear drift (-A, b); m [B,N+1,D], S [B,N+1,D,D]."""
    _req(A, b, m, S, status)
    B, N, D = b.shape
    lib = _lib.load()
    value = torch.empty(B, dtype=F64, device=A.device)
    dEdm = torch.empty_like(b) if want_grads else None
    dEdS = torch.empty_like(A) if want_grads else None
    nbytes = lib.vdp_sq_drift_difference_workspace(B, N, D)
    ws = _workspace(nbytes, A.device)
    check(lib.vdp_sq_drift_difference(int(drift_id), dvec(params, 16), dvec(q.reshape(-1), D * D), float(dt), _p(A),
                                      _p(b), _p(m), _p(S), _p(value), _p(dEdm), _p(dEdS), _p(ws), nbytes, B, N, D,
                                      _p(status), _stream()), "vdp_sq_drift_difference")
    return (value, dEdm, dEdS) if want_grads else value


def drift_num_params(drift_id):
    return int(_lib.load().vdp_drift_num_params(int(drift_id)))


def drift_param_grad(drift_id, params, q, q_scale, c_x, c_f, a_sign, out_scale, A, b, m, S, point_offset=0,
                     status=None):
    """
    grads[b, p] = out_scale sum_t E_{N(m_t, S_t)}[r^T sym((q_scale q)^-1) df/dp],  r = c_x x + c_f f(x) - (a_sign A_t x
    + b_t), H = 20 (``vdp_drift_param_grad``).  A [B,N,D,D], b [B,N,D]; m [B,P,D], S [B,P,D,D] with P >= N +
    point_offset grid points per sequence, transition t reads grid point t + point_offset.  -> [B, n_params].
    """
    _req(A, b, m, S, status)
    B, N, D = b.shape
    pts = m.shape[1]
    assert pts >= N + point_offset and S.shape[1] == pts
    lib = _lib.load()
    import numpy as np
    qh = np.atleast_2d(np.asarray(q, dtype=np.float64))
    grads = torch.empty((B, 16), dtype=F64, device=A.device)
    nbytes = lib.vdp_drift_param_grad_workspace(B, N, D)
    ws = _workspace(nbytes, A.device)
    mp = C.c_void_p(m.data_ptr() + 8 * D * point_offset)
    Sp = C.c_void_p(S.data_ptr() + 8 * D * D * point_offset)
    check(lib.vdp_drift_param_grad(int(drift_id), dvec(params, 16), dvec(qh.reshape(-1), D * D), float(q_scale),
                                   float(c_x), float(c_f), float(a_sign), float(out_scale), _p(A), _p(b), mp, Sp, pts,
                                   _p(grads), _p(ws), nbytes, B, N, D, _p(status), _stream()), "vdp_drift_param_grad")
    return grads[:, :drift_num_params(drift_id)]


def expectations_to_ssm_vjp(e1, ed, es, gA, gb, gcholP0, gcholQ, gmu0, status=None):
    """VJP of expectations_to_ssm_params (ssm_natgrad.py:168) -> dL/d(eta1, eta_diag, eta_subdiag)."""
    _req(e1, ed, es, gA, gb, gcholP0, gcholQ, gmu0, status)
    B, n, D = e1.shape
    g1, gd, gs = torch.empty_like(e1), torch.empty_like(ed), torch.empty_like(es)
    check(_lib.load().vdp_expectations_to_ssm_vjp(_p(e1), _p(ed), _p(es), _p(gA), _p(gb), _p(gcholP0), _p(gcholQ),
                                                  _p(gmu0), _p(g1), _p(gd), _p(gs), B, n - 1, D, _p(status),
                                                  _stream()), "vdp_expectations_to_ssm_vjp")
    return g1, gd, gs


def ssm_natgrad_step(A, b, cholP0, cholQ, mu0, dL_dA, dL_db, dL_dcholP0, dL_dcholQ, dL_dmu0, gamma, status=None):
    """In-place natural-gradient step on the SSM parameters (ssm_natgrad.py:121-218, momentum=False)."""
    _req(A, b, cholP0, cholQ, mu0, dL_dA, dL_db, dL_dcholP0, dL_dcholQ, dL_dmu0, status)
    B, N, D = b.shape
    lib = _lib.load()
    nbytes = lib.vdp_ssm_natgrad_workspace(B, N, D)
    ws = _workspace(nbytes, A.device)
    check(lib.vdp_ssm_natgrad_step(_p(A), _p(b), _p(cholP0), _p(cholQ), _p(mu0), _p(dL_dA), _p(dL_db), _p(dL_dcholP0),
                                   _p(dL_dcholQ), _p(dL_dmu0), float(gamma), _p(ws), nbytes, B, N, D, _p(status),
                                   _stream()), "vdp_ssm_natgrad_step")


def euler_maruyama(drift_id, params, q, x0, time_grid, noise=None, seed=0):
    """sde_utils.py:36-96, batched: x0 [B, D], time_grid [n] -> paths [B, n, D] (paths[:, 0] = x0).
    ``noise`` [B, n-1, D] standard normals replaces the Philox generator (exact reproduction of a given draw)."""
    _req(x0, time_grid, noise)
    B, D = x0.shape
    n = time_grid.shape[0]
    out = torch.empty((B, n, D), dtype=F64, device=x0.device)
    import numpy as np
    qh = np.atleast_2d(np.asarray(q, dtype=np.float64))
    check(_lib.load().vdp_euler_maruyama(int(drift_id), dvec(params, 16), dvec(qh.reshape(-1), D * D),
                                         _p(x0), _p(time_grid), _p(noise), int(seed), _p(out), B, n, D, None,
                                         _stream()), "vdp_euler_maruyama")
    return out


# every public op runs on the device of its tensors (see _on_tensor_device)
for _name, _fn in list(globals().items()):
    if callable(_fn) and not _name.startswith("_") and getattr(_fn, "__module__", None) == __name__ \
            and _name not in ("new_status", "drift_num_params"):
        globals()[_name] = _on_tensor_device(_fn)
del _name, _fn
