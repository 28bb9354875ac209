import torch, time
N=10_000_000
th=torch.zeros(N,3,3,dtype=torch.float64,device='cuda'); t1=torch.zeros(N,3,dtype=torch.float64,device='cuda')
idx=torch.sort(torch.randperm(N,device='cuda')[:10000]).values
d=torch.randn(10000,3,3,dtype=torch.float64,device='cuda'); d1=torch.randn(10000,3,dtype=torch.float64,device='cuda')
for name,f in (("index_add_",lambda: (th.index_add_(0,idx,d), t1.index_add_(0,idx,d1))),
               ("index_put_acc",lambda: (th.index_put_((idx,),d,accumulate=True), t1.index_put_((idx,),d1,accumulate=True))),
               ("th[idx]+=",lambda: (th.__setitem__(idx, th[idx]+d),))):
    f(); torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(5): f()
    torch.cuda.synchronize(); print(name, (time.perf_counter()-t)/5*1e3, "ms")
