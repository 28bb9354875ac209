"""
ctypes binding of libvdp_b200.so (C ABI declared in include/vdp_b200.h).

There is NO fallback: if the shared library is missing or a call fails, an exception is raised.
Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or ``tools/build_lib.sh``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VDP_B200_LIB") or os.path.join(_HERE, "libvdp_b200.so")

c_dp = C.POINTER(C.c_double)
c_i64p = C.POINTER(C.c_int64)
i64 = C.c_int64
vp = C.c_void_p


class VdpError(RuntimeError):
    pass


class CviProblem(C.Structure):
    """Mirror of ``struct vdp_cvi_problem`` (include/vdp_b200.h)."""
    _fields_ = [
        ("B", i64), ("N", i64), ("n_obs", i64),
        ("D", C.c_int32), ("drift_id", C.c_int32),
        ("drift_params", C.c_double * 16),
        ("q", C.c_double * 9),
        ("dt", C.c_double),
        ("x0_mu", C.c_double * 3), ("x0_cov", C.c_double * 9),
        ("lik_chol", C.c_double * 9),
        ("lr_data", C.c_double), ("lr_girsanov", C.c_double),
        ("time_grid", vp),
        ("pA", vp), ("pb", vp), ("pcholQ", vp), ("pmu0", vp), ("pcholP0", vp),
        ("g1", vp), ("gd", vp), ("gs", vp),
        ("d1", vp), ("d2", vp),
        ("obs_idx", vp), ("obs_y", vp),
        ("fx_mus", vp), ("fx_covs", vp),
        ("qA", vp), ("qb", vp), ("qcholQ", vp), ("qmu0", vp), ("qcholP0", vp),
        ("elbo", vp), ("ve", vp), ("kl", vp),
        ("workspace", vp), ("workspace_bytes", C.c_size_t),
        ("status", vp),
        ("flags", C.c_int32), ("reserved", C.c_int32),
        ("lr_data_dev", vp), ("lr_girsanov_dev", vp),
        ("active_dev", vp),
    ]


class CviLoop(C.Structure):
    """Mirror of ``struct vdp_cvi_loop`` (include/vdp_b200.h)."""
    _fields_ = [
        ("max_itr", C.c_int32), ("n_test", C.c_int32),
        ("later_flags", C.c_int32), ("reserved", C.c_int32),
        ("optim_tol", C.c_double),
        ("lr_data", vp), ("lr_girsanov", vp),
        ("active", vp), ("n_itr", vp),
        ("elbo_trace", vp),
        ("test_idx", vp),
        ("test_mus", vp), ("test_covs", vp),
        ("counters", vp),
    ]


class VgpProblem(C.Structure):
    """Mirror of ``struct vdp_vgp_problem`` (include/vdp_b200.h)."""
    _fields_ = [
        ("B", i64), ("N", i64), ("n_obs", i64),
        ("D", C.c_int32), ("drift_id", C.c_int32),
        ("drift_params", C.c_double * 16),
        ("q", C.c_double * 9),
        ("dt", C.c_double),
        ("x0_mu", C.c_double * 3), ("x0_cov", C.c_double * 9),
        ("lik_chol", C.c_double * 9),
        ("lr", C.c_double), ("x0_lr", C.c_double),
        ("time_grid", vp),
        ("A", vp), ("b", vp),
        ("lam", vp), ("psi", vp),
        ("q_mean", vp), ("q_chol", vp),
        ("obs_idx", vp), ("obs_y", vp),
        ("m", vp), ("S", vp),
        ("elbo", vp),
        ("workspace", vp), ("workspace_bytes", C.c_size_t),
        ("status", vp),
        ("flags", C.c_int32), ("reserved", C.c_int32),
    ]


class DistSweep(C.Structure):
    """Mirror of ``struct vdp_dist_sweep`` (include/vdp_b200.h)."""
    _fields_ = [
        ("kind", C.c_int32), ("stage", C.c_int32),
        ("rank", C.c_int32), ("nranks", C.c_int32),
        ("N_local", i64),
        ("D", C.c_int32), ("reserved", C.c_int32),
        ("scaleD", C.c_double), ("scaleS", C.c_double),
        ("in", vp * 8),
        ("out", vp * 8),
        ("prev_totals", vp), ("prev_inits", vp),
        ("workspace", vp), ("workspace_bytes", C.c_size_t),
        ("status", vp),
    ]


SWEEP_CHOL, SWEEP_NAT2SSM, SWEEP_MARG, SWEEP_INVBAND, SWEEP_MEAN, SWEEP_INFO, SWEEP_MOMENTS = 0, 1, 2, 3, 4, 5, 6
FLAG_SKIP_DATA_UPDATE, FLAG_SKIP_ELBO, FLAG_SKIP_GIRSANOV, FLAG_POSTERIOR_SSM_ONLY = 1, 2, 4, 8
FLAG_NATURALS_CARRIED = 16
FLAG_SITES_IN_NATURALS = 32
FLAG_GH_QUADRATURE, FLAG_REFACTOR_PRECISION, FLAG_NO_FUSED_NATURALS, FLAG_SEPARATE_KL_VALUE = 64, 128, 256, 512
FLAG_NO_FUSED_FORWARD = 1024
VGP_STABILIZE = 1
VGP_UPDATE_X0 = 2
VGP_STATE_CARRIED = 16
VGP_GH_QUADRATURE = 4
VGP_REFACTOR_PRECISION = 8

# name -> (restype, argtypes); every symbol include/vdp_b200.h declares
SIGNATURES = {
    "vdp_version": (C.c_int, []),
    "vdp_config_set": (C.c_int, [C.c_char_p, C.c_int]),
    "vdp_config_get": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "vdp_last_error": (C.c_char_p, []),
    "vdp_read_status": (i64, [vp, c_i64p, vp]),
    "vdp_btd_cholesky": (C.c_int, [vp, vp, vp, vp, i64, i64, C.c_int, vp, vp]),
    "vdp_btd_inverse_band": (C.c_int, [vp, vp, vp, vp, i64, i64, C.c_int, vp]),
    "vdp_btd_inverse_band_jvp": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i64, C.c_int, vp]),
    "vdp_btd_solve": (C.c_int, [vp, vp, vp, vp, C.c_int, i64, i64, C.c_int, vp]),
    "vdp_btd_matvec": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_int, i64, i64, C.c_int, vp]),
    "vdp_btd_logdet": (C.c_int, [vp, vp, i64, i64, C.c_int, vp]),
    "vdp_btd_block_to_band": (C.c_int, [vp, vp, vp, i64, i64, C.c_int, vp]),
    "vdp_btd_band_to_block": (C.c_int, [vp, vp, vp, C.c_int, i64, i64, C.c_int, vp]),
    "vdp_btd_conditional_precisions": (C.c_int, [vp, vp, vp, vp, i64, i64, C.c_int, vp]),
    "vdp_gh_expect": (C.c_int, [C.c_int, c_dp, vp, vp, C.c_int, vp, vp, i64, C.c_int, vp, vp]),
    "vdp_linearize_to_ssm": (C.c_int, [C.c_int, c_dp, c_dp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_double,
                                       C.c_double, i64, i64, C.c_int, vp, vp]),
    "vdp_ssm_to_naturals": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, C.c_int, vp]),
    "vdp_naturals_to_ssm_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_naturals_to_ssm": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, i64, i64, C.c_int, vp, vp]),
    "vdp_ssm_marginals_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_ssm_marginals": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, i64, i64, C.c_int, vp, vp]),
    "vdp_ssm_kl_divergence_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_ssm_kl_divergence": (C.c_int, [vp] * 11 + [vp, C.c_size_t, i64, i64, C.c_int, vp, vp]),
    "vdp_ssm_log_pdf_workspace": (C.c_size_t, [i64, i64, i64]),
    "vdp_ssm_log_pdf": (C.c_int, [vp] * 7 + [vp, C.c_size_t, i64, i64, i64, C.c_int, vp]),
    "vdp_ssm_sample": (C.c_int, [vp] * 6 + [C.c_uint64, vp, i64, i64, i64, C.c_int, vp]),
    "vdp_static_scratch_bytes": (C.c_size_t, []),
    "vdp_ssm_marginals_static": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, i64, C.c_int, vp, vp]),
    "vdp_ssm_to_expectations": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i64, C.c_int, vp]),
    "vdp_expectations_to_ssm": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, C.c_int, vp, vp]),
    "vdp_expectations_to_ssm_vjp": (C.c_int, [vp] * 11 + [i64, i64, C.c_int, vp, vp]),
    "vdp_ssm_natgrad_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_ssm_natgrad_step": (C.c_int, [vp] * 10 + [C.c_double, vp, C.c_size_t, i64, i64, C.c_int, vp, vp]),
    "vdp_cvi_kl_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_cvi_kl_and_grad": (C.c_int, [C.c_int, c_dp, c_dp, C.c_double, vp, c_dp, c_dp, vp, vp, vp, vp, vp, vp, vp, vp,
                                      C.c_size_t, i64, i64, C.c_int, vp, vp]),
    "vdp_cvi_workspace": (C.c_size_t, [i64, i64, C.c_int, i64]),
    "vdp_cvi_site_update": (C.c_int, [C.POINTER(CviProblem), vp]),
    "vdp_cvi_materialize_sites": (C.c_int, [C.POINTER(CviProblem), vp]),
    "vdp_cvi_site_update_sharded": (C.c_int, [C.POINTER(CviProblem), C.c_int, i64, i64, vp]),
    "vdp_cvi_elbo": (C.c_int, [C.POINTER(CviProblem), vp]),
    "vdp_cvi_optimize_sites": (C.c_int, [C.POINTER(CviProblem), C.POINTER(CviLoop), vp]),
    "vdp_cvi_update_data_sites": (C.c_int, [C.POINTER(CviProblem), vp]),
    "vdp_cvi_full_sites": (C.c_int, [C.POINTER(CviProblem), vp, vp, vp, vp]),
    "vdp_cvi_update_girsanov_sites_range": (C.c_int, [C.POINTER(CviProblem), i64, i64, vp]),
    "vdp_cvi_elbo_terms_range": (C.c_int, [C.POINTER(CviProblem), i64, i64, vp]),
    "vdp_drift_num_params": (C.c_int, [C.c_int]),
    "vdp_cvi_grad_kl_prior_params": (C.c_int, [C.POINTER(CviProblem), vp, vp]),
    "vdp_cvi_grad_ve_prior_params": (C.c_int, [C.POINTER(CviProblem), vp, vp, C.c_int, C.c_double, C.c_double, vp,
                                               vp]),
    "vdp_drift_param_grad_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_drift_param_grad": (C.c_int, [C.c_int, c_dp, c_dp, C.c_double, C.c_double, C.c_double, C.c_double,
                                       C.c_double, vp, vp, vp, vp, i64, vp, vp, C.c_size_t, i64, i64, C.c_int, vp,
                                       vp]),
    "vdp_euler_maruyama": (C.c_int, [C.c_int, c_dp, c_dp, vp, vp, vp, C.c_uint64, vp, i64, i64, C.c_int, vp, vp]),
    "vdp_cvi_host_create": (C.c_int, [C.POINTER(vp), i64, i64, C.c_int, i64, C.c_int, C.c_int]),
    "vdp_cvi_host_set_prior": (C.c_int, [vp, i64, vp, vp, vp, vp, vp]),
    "vdp_cvi_host_site_update": (C.c_int, [vp, C.POINTER(CviProblem), c_i64p]),
    "vdp_cvi_host_destroy": (C.c_int, [vp]),
    "vdp_cvi_session_create": (C.c_int, [C.POINTER(vp), C.POINTER(CviProblem), i64, C.c_int, C.c_int]),
    "vdp_cvi_session_site_update": (C.c_int, [vp, vp, C.c_double, C.c_double, C.c_int, vp, vp, vp, vp, vp, c_i64p]),
    "vdp_cvi_session_download_state": (C.c_int, [vp, C.POINTER(CviProblem)]),
    "vdp_cvi_session_destroy": (C.c_int, [vp]),
    "vdp_linear_drift_to_ssm": (C.c_int, [vp, vp, c_dp, vp, vp, vp, vp, i64, i64, C.c_int, vp, vp]),
    "vdp_sq_drift_difference_workspace": (C.c_size_t, [i64, i64, C.c_int]),
    "vdp_sq_drift_difference": (C.c_int, [C.c_int, c_dp, c_dp, C.c_double, vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t,
                                          i64, i64, C.c_int, vp, vp]),
    "vdp_vgp_workspace": (C.c_size_t, [i64, i64, C.c_int, i64]),
    "vdp_vgp_forward_pass": (C.c_int, [C.POINTER(VgpProblem), vp]),
    "vdp_vgp_update_lagrange": (C.c_int, [C.POINTER(VgpProblem), vp]),
    "vdp_vgp_update_param": (C.c_int, [C.POINTER(VgpProblem), vp]),
    "vdp_vgp_update_initial_statistics": (C.c_int, [C.POINTER(VgpProblem), vp]),
    "vdp_vgp_elbo": (C.c_int, [C.POINTER(VgpProblem), vp]),
    "vdp_vgp_sweep": (C.c_int, [C.POINTER(VgpProblem), vp]),
    "vdp_comm_unique_id": (C.c_int, [vp]),
    "vdp_comm_create": (C.c_int, [C.POINTER(vp), vp, C.c_int, C.c_int]),
    "vdp_comm_all_gather": (C.c_int, [vp, vp, vp, C.c_size_t, vp]),
    "vdp_comm_destroy": (C.c_int, [vp]),
    "vdp_dist_workspace": (C.c_size_t, [C.c_int]),
    "vdp_dist_blob_bytes": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "vdp_dist_sweep_summary": (C.c_int, [C.POINTER(DistSweep), vp, vp, vp]),
    "vdp_dist_sweep_finish": (C.c_int, [C.POINTER(DistSweep), vp, vp, vp]),
    "vdp_test_fastmath": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, vp]),
    "vdp_test_fp64_fma": (C.c_int, [vp, i64, C.c_int, i64, vp]),
    "vdp_profile_enable": (C.c_int, [C.c_int]),
    "vdp_profile_read": (C.c_int, [c_dp, c_i64p, C.c_int]),
}

PHASES = ["data_sites", "sites_to_nat", "chol_fwd", "nat2ssm_bwd", "marg_fwd", "invband_bwd", "kl_grad", "elbo"]

_lib = None


def load():
    """Load the library once and attach signatures.  Raises VdpError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VdpError(f"{LIB_PATH} is missing: the CUDA library is not built and there is no CPU fallback. "
                       f"Run tools/build_lib.sh (needs nvcc).")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


CONFIG_KEYS = ("SWEEPS", "NSEG", "GH_QUADRATURE", "REFACTOR_PRECISION", "NO_CARRIED_NATURALS", "NO_FUSED_NATURALS",
               "SEPARATE_KL_VALUE", "NO_FUSED_FORWARD")
_SWEEPS = {None: 0, "auto": 0, "seq": 1, "tiled": 2}


def configure(**kw):
    """Process-wide tuning switches of the library (vdp_config_set): ``configure(SWEEPS="tiled", NSEG=3)``; ``None``
    restores the default (0).  Used by tests and tuning runs; per-call ``FLAG_*`` select numerics routes re-entrantly."""
    lib = load()
    for k, v in kw.items():
        k = k.upper()
        if k not in CONFIG_KEYS:
            raise VdpError(f"unknown configuration key {k}")
        if k == "SWEEPS" and not isinstance(v, int):
            v = _SWEEPS[v]
        check(lib.vdp_config_set(k.encode(), int(v or 0)), "vdp_config_set")


def config_get(key):
    out = C.c_int()
    check(load().vdp_config_get(key.upper().encode(), C.byref(out)), "vdp_config_get")
    return out.value


def check(rc: int, what: str):
    if rc != 0:
        msg = load().vdp_last_error().decode("utf-8", "replace")
        raise VdpError(f"{what} failed with code {rc}: {msg}")


def dvec(values, n):
    """host double[n] from a sequence (zero padded)."""
    arr = (C.c_double * n)()
    for i, v in enumerate(values):
        arr[i] = float(v)
    return arr
