import torch, time
x = torch.empty(1<<30, dtype=torch.uint8).pin_memory()
d = torch.empty(1<<30, dtype=torch.uint8, device='cuda')
for n in (1<<30, 1<<28, 1<<24, 1<<20):
    torch.cuda.synchronize()
    t0=time.perf_counter()
    reps = (1<<32)//n
    for i in range(reps): d[:n].copy_(x[:n], non_blocking=True)
    torch.cuda.synchronize()
    dt=time.perf_counter()-t0
    print(f"H2D chunk {n>>20} MiB x{reps}: {n*reps/dt/1e9:.1f} GB/s")
h = torch.empty(1<<30, dtype=torch.uint8).pin_memory()
torch.cuda.synchronize(); t0=time.perf_counter()
for i in range(4): h.copy_(d, non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t0
print(f"D2H 1 GiB x4: {4*(1<<30)/dt/1e9:.1f} GB/s")
# bidirectional
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
torch.cuda.synchronize(); t0=time.perf_counter()
for i in range(4):
    with torch.cuda.stream(s1): d.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2): h.copy_(d, non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t0
print(f"bidirectional: H2D {4*(1<<30)/dt/1e9:.1f} GB/s + D2H same")
