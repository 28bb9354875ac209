"""
ctypes front end of oracle/cvi_d1.c (the C / OpenMP restatement of the D = 1 site update).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): imported by tests/ and by bench.py's CPU-baseline legs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import scalar_d1

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libcvi_d1.so")
_lib = None


def build():
    subprocess.check_call(["bash", os.path.join(_HERE, "build_c.sh")], stdout=subprocess.DEVNULL)


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = C.CDLL(LIB_PATH)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int64)
        lib.cvi_d1_site_update.restype = C.c_int
        lib.cvi_d1_site_update.argtypes = ([C.c_int, dp, C.c_double, dp, C.c_int64, C.c_int64, C.c_int64, ip, dp,
                                            C.c_double, C.c_double, C.c_double] + [dp] * 12 +
                                           [C.c_double, C.c_double, dp, dp, C.c_int, dp, C.c_int])
        lib.cvi_d1_max_threads.restype = C.c_int
        _lib = lib
    return _lib


def _dp(a):
    assert a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(C.POINTER(C.c_double))


class CBatchedCVISitesSDE1D(scalar_d1.BatchedCVISitesSDE1D):
    """Same state as the NumPy class; ``site_update`` runs in C, ``threads`` OpenMP threads over the sequences."""

    def __init__(self, *args, threads=1, **kw):
        super().__init__(*args, **kw)
        assert self.dtype == np.float64
        self.threads = threads
        x, w = self.gh20
        self._ghx = np.ascontiguousarray(x, dtype=np.float64)
        self._ghw = np.ascontiguousarray(w * np.pi ** (-0.5), dtype=np.float64)
        self.obs_idx = np.ascontiguousarray(self.obs_idx, dtype=np.int64)

    def site_update(self, data_lr, girsanov_lr):
        lib = load()
        for k in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
            setattr(self, k, np.ascontiguousarray(getattr(self, k)))
        p = {k: np.ascontiguousarray(v) for k, v in self.p.items()}
        elbo = np.empty(self.B)
        params = np.ascontiguousarray(self.params, dtype=np.float64)
        rc = lib.cvi_d1_site_update(
            int(self.drift_id), _dp(params), float(self.q), _dp(self.time_grid), self.n1, self.B,
            self.obs_idx.shape[1], self.obs_idx.ctypes.data_as(C.POINTER(C.c_int64)), _dp(self.obs_y),
            float(self.noise_std), float(self.p0_mu), float(self.p0_cov), _dp(p["A"]), _dp(p["b"]), _dp(p["cQ"]),
            _dp(p["mu0"]), _dp(p["cP0"]), _dp(self.g1), _dp(self.gd), _dp(self.gs), _dp(self.d1), _dp(self.d2),
            _dp(self.fx_mus), _dp(self.fx_covs), float(data_lr), float(girsanov_lr), _dp(self._ghx), _dp(self._ghw),
            20, _dp(elbo), int(self.threads))
        if rc != 0:
            raise MemoryError("cvi_d1_site_update")
        return elbo


def max_threads():
    return load().cvi_d1_max_threads()
