"""Error of the library's fast float64 reciprocal / sqrt / division (csrc/fastmath.cuh) in ulps."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def ulps(got, want):
    return np.abs(got - want) / np.spacing(np.abs(want))


def test_fastmath_within_one_ulp():
    from vi_diffusion_processes_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(0)
    n = 1 << 20
    x = np.concatenate([np.exp(rng.uniform(np.log(1e-30), np.log(1e30), n)), rng.uniform(0.5, 2.0, n),
                        np.array([1.0, 2.0, 4.0, 0.25, 3.0, 1e-200, 1e200, 1.0 + 2.0 ** -52, 2.0 - 2.0 ** -52])])
    y = rng.normal(size=x.shape) * np.exp(rng.uniform(-20, 20, x.shape))
    dx, dy = torch.as_tensor(x).cuda(), torch.as_tensor(y).cuda()
    out = [torch.empty_like(dx) for _ in range(4)]
    rc = lib.vdp_test_fastmath(dx.data_ptr(), dy.data_ptr(), *(o.data_ptr() for o in out), x.size,
                               C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    rcp, sq, rsq, quo = (o.cpu().numpy() for o in out)
    xl, yl = x.astype(np.longdouble), y.astype(np.longdouble)
    assert ulps(rcp, (1 / xl).astype(np.float64)).max() <= 1.0
    assert ulps(sq, np.sqrt(xl).astype(np.float64)).max() <= 1.0
    assert ulps(rsq, (1 / np.sqrt(xl)).astype(np.float64)).max() <= 2.0
    assert ulps(quo, (yl / xl).astype(np.float64)).max() <= 1.0
    # exact cases stay exact
    assert rcp[-9] == 1.0 and rcp[-8] == 0.5 and sq[-7] == 2.0 and sq[-6] == 0.5
