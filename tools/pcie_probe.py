import torch, time
a = torch.empty(256*1024*1024//4, dtype=torch.float32).pin_memory()
d = torch.empty_like(a, device='cuda')
b = torch.empty_like(a).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for name, fn in (("h2d", lambda: d.copy_(a, non_blocking=True)), ("d2h", lambda: b.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t=time.perf_counter()
    for _ in range(5): fn()
    torch.cuda.synchronize(); dt=(time.perf_counter()-t)/5
    print(name, "%.1f GB/s" % (a.numel()*4/dt/1e9))
d2 = torch.empty_like(d)
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1): d.copy_(a, non_blocking=True)
    with torch.cuda.stream(s2): b.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/5
print("both directions at once: %.1f GB/s each" % (a.numel()*4/dt/1e9))
