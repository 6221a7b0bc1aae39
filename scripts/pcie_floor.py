import torch, time
T,K,N=512,4096,14336
d=torch.empty((T,N),device='cuda'); h=torch.empty((T,N)).pin_memory()
xd=torch.empty((T,K),device='cuda'); xh=torch.randn((T,K)).pin_memory()
for i in range(3): h.copy_(d,non_blocking=True); torch.cuda.synchronize()
t0=time.perf_counter()
for i in range(20): h.copy_(d,non_blocking=True); torch.cuda.synchronize()
dt=(time.perf_counter()-t0)/20; print(f"D2H 29.4MB alone: {dt*1e6:.1f} us {4*T*N/dt/1e9:.1f} GB/s")
t0=time.perf_counter()
for i in range(20): xd.copy_(xh,non_blocking=True); torch.cuda.synchronize()
dt=(time.perf_counter()-t0)/20; print(f"H2D 8.4MB alone: {dt*1e6:.1f} us {4*T*K/dt/1e9:.1f} GB/s")
s2=torch.cuda.Stream()
t0=time.perf_counter()
for i in range(20):
    h.copy_(d,non_blocking=True)
    with torch.cuda.stream(s2): xd.copy_(xh,non_blocking=True)
    torch.cuda.synchronize()
dt=(time.perf_counter()-t0)/20; print(f"both concurrently: {dt*1e6:.1f} us")
