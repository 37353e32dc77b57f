import torch
for n in (35_000_000, 70_000_000, 268_435_456):
    a=torch.empty(n,dtype=torch.float64,device='cuda').normal_(); b=torch.empty_like(a)
    for _ in range(5): b.copy_(a)
    torch.cuda.synchronize()
    ev=[torch.cuda.Event(enable_timing=True) for _ in range(41)]
    ev[0].record()
    for i in range(40):
        b.copy_(a); ev[i+1].record()
    torch.cuda.synchronize()
    ts=[ev[i].elapsed_time(ev[i+1]) for i in range(40)]
    t=sum(ts)/40
    print(n, 'mean ms',t,'min',min(ts),'GB/s mean', 2*n*8/t/1e6, 'GB/s best', 2*n*8/min(ts)/1e6)
