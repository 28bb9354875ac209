import sys, numpy as np, torch
sys.path.insert(0, '.')
from vi_diffusion_processes_b200.cvi import CviEngine
import os
os.environ.setdefault("VDP_SWEEPS", sys.argv[1] if len(sys.argv) > 1 else "tiled")
g = np.load('./tests/golden/c2_dw.npz')
D = 1
eng = CviEngine(int(g["drift_id"]), g["drift_params"], g["q"], g["time_grid"], g["obs_idx"], g["obs_y"], float(g["sigma"]) * np.eye(D), g["x0_mu"], g["x0_cov"])
lr = float(g["lr"])
el = []
for it in range(3):
    el.append(eng.site_update(lr, lr).item())
    if it == 1: eng.relinearize()
r = lambda a, b, s=None: float(np.abs(a - b).max() / (np.abs(b).max() if s is None else s))
print(os.environ["VDP_SWEEPS"], "elbo", r(np.array(el), g["cvi_elbo"]), el, g["cvi_elbo"])
print("m", r(eng.fx_mus[0].cpu().numpy(), g["cvi_m"]), "S", r(eng.fx_covs[0].cpu().numpy(), g["cvi_S"]))
th = [t[0].cpu().numpy() for t in eng._prior_naturals(eng.pA, eng.pb)]
for nm, got, want, t in (("g1", eng.g1, g["cvi_g1"], th[0]), ("gd", eng.gd, g["cvi_gd"], th[1]), ("gs", eng.gs, g["cvi_gs"], th[2])):
    print(nm, r(got[0].cpu().numpy(), want), r(got[0].cpu().numpy(), want, max(np.abs(want).max(), 1e-3 * np.abs(t + want).max())))
