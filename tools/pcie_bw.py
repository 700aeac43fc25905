import torch, time
n=72_000_000
h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device='cuda')
for _ in range(3): d.copy_(h,non_blocking=True)
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): d.copy_(h,non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
print("H2D GB/s",n/dt/1e9, "ms", dt*1e3)
t=time.perf_counter()
for _ in range(20): h.copy_(d,non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
print("D2H GB/s",n/dt/1e9)
