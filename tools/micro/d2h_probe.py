import torch, time
for mb in (0.25, 1, 2, 4, 8, 32):
    n=int(mb*2**20)
    d=torch.empty(n,dtype=torch.uint8,device='cuda'); h=torch.empty(n,dtype=torch.uint8).pin_memory()
    for _ in range(5): h.copy_(d,non_blocking=True); torch.cuda.synchronize()
    t=time.perf_counter()
    for _ in range(50): h.copy_(d,non_blocking=True); torch.cuda.synchronize()
    dt=(time.perf_counter()-t)/50
    print(f"D2H {mb} MB: {dt*1e3:.3f} ms  {n/dt/1e9:.1f} GB/s")
