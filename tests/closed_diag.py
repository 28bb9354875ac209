"""(test infrastructure, not collected by pytest) Diagnostic: closed-form moments vs Gauss-Hermite loops vs the NumPy oracle, update by update (small noise)."""
import os
import sys

import numpy as np

from vi_diffusion_processes_b200 import _lib  # noqa: E402
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import models_np as M      # noqa: E402  (checker)
from oracle import sde_np as S         # noqa: E402
from vi_diffusion_processes_b200.cvi import CviEngine   # noqa: E402

noise = float(sys.argv[1]) if len(sys.argv) > 1 else 1e-3
rng = np.random.default_rng(33)
N, B, n_obs = 800, 4, 10
grid = np.round(np.arange(N + 1) * 0.005, 3)
idx = np.sort(np.stack([rng.choice(N + 1, n_obs, replace=False) for _ in range(B)]), axis=1)
y = rng.normal(size=(B, n_obs, 1)) * 0.3 + np.array([1.0, -1.0, 0.2, 1.4])[:, None, None]
params = [2.5, 0.7]
mk = lambda: CviEngine(1, params, np.ones((1, 1)), grid, idx, y, [[noise]], [1.0], [[1e-2]])
a, c = mk(), mk()
sde = S.DoubleWellSDE(q=np.ones((1, 1)), scale=params[0], c=params[1])
refs = []
try:
    for b in range(1):
        refs.append(M.CVISitesSDE(sde, grid, (grid[idx[3]], y[3]), M.MultivariateGaussian(noise * np.eye(1)),
                                  prior_initial_state=S.Gaussian(np.array([1.0]), np.array([[1e-2]]))))
except Exception as e:   # the oracle's constructor signature may differ: the GPU-vs-GPU comparison still prints
    print("oracle not built:", e)
for it in range(4):
    ea = a.site_update(0.5, 0.5).clone()                     # closed-form moments (default route)
    c.route_flags = _lib.FLAG_GH_QUADRATURE                  # the 20-node loops
    ec = c.site_update(0.5, 0.5).clone()
    out = {}
    for name in ("g1", "gd", "gs", "d1", "d2", "fx_mus", "fx_covs"):
        x, r = getattr(a, name), getattr(c, name)
        dlt = (x - r).abs()
        out[name] = (float(dlt.max()), float(r.abs().max()), int(dlt.argmax()))
    print(it, "elbo", float((ea - ec).abs().max()), float(ec.abs().max()), out)
    if refs:
        el = refs[0].site_update(0.5, 0.5) if hasattr(refs[0], "site_update") else None
        print("   oracle elbo", el, "closed", float(ea[3]), "gh", float(ec[3]))
        gref = refs[0].girsanov_sites
        for e, nm in ((a, "closed"), (c, "gh")):
            print("   ", nm, "g1 err", float(np.abs(e.g1[3].cpu().numpy() - gref.nat1).max()),
                  "gd err", float(np.abs(e.gd[3].cpu().numpy() - gref.nat2.block_diagonal).max()),
                  "gs err", float(np.abs(e.gs[3].cpu().numpy() - gref.nat2.block_sub_diagonal).max()))
        for e, nm in ((a, "closed"), (c, "gh")):
            print("   ", nm, "m err", float(np.abs(e.fx_mus[3].cpu().numpy() - refs[0].fx_mus).max()),
                  "S err", float(np.abs(e.fx_covs[3].cpu().numpy() - refs[0].fx_covs).max()))
print("min S", float(a.fx_covs.min()), float(c.fx_covs.min()))
