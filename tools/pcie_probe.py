import torch, time
x = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
d = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
for name, fn in (("h2d", lambda: d.copy_(x, non_blocking=True)), ("d2h", lambda: x.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(3): fn()
    torch.cuda.synchronize()
    print(name, "GB/s", 3 * 1.0737 / (time.perf_counter() - t))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
y = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True); e = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(3):
    with torch.cuda.stream(s1): d.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2): y.copy_(e, non_blocking=True)
torch.cuda.synchronize()
print("duplex each GB/s", 3 * 1.0737 / (time.perf_counter() - t))
