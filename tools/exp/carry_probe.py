"""where does a carried time-sharded update spend its time (one GPU, 10^7 steps)"""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import c5_time_sharded as c5
from vi_diffusion_processes_b200.dist_cvi import TimeShardedCvi
from vi_diffusion_processes_b200.dist_time import Comm
N = 10_000_000
grid, idx, y = c5.make_problem(N, 10_000)
sh = TimeShardedCvi(8, [0.5], np.eye(3), grid, idx, y, 0.1 * np.eye(3), np.zeros(3), np.eye(3), comm=Comm.emulated(1))
for _ in range(2):
    sh.site_update(0.5, 0.5)
torch.cuda.synchronize()
st = sh.ranks[0]
e = st.eng
def timed(name, f, n=3):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); print(name, round((time.perf_counter() - t) / n * 1e3, 3), "ms", flush=True)
d = torch.randn_like(e.d1); d2 = torch.randn_like(e.d2) * 1e-3
timed("index_add th1", lambda: st.th1[0].index_add_(0, e.obs_idx[0], d[0]))
timed("index_add thd", lambda: st.thd[0].index_add_(0, e.obs_idx[0], d2[0]))
timed("signature", lambda: sh._th_signature())
timed("rebuild full", lambda: sh._rebuild())
bef = [(e.d1.clone(), e.d2.clone())]
timed("rebuild carried", lambda: sh._rebuild(bef))
timed("update", lambda: sh.site_update(0.5, 0.5))
print("alloc retries", torch.cuda.memory_stats()["num_alloc_retries"], "reserved GB", torch.cuda.memory_reserved() / 1e9)
sh.carry_naturals = False
timed("update no carry", lambda: sh.site_update(0.5, 0.5))
print(e.obs_idx.dtype, e.obs_idx.shape, int(e.obs_idx.min()), int(e.obs_idx.max()), st.th1.shape)
