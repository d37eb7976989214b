import torch, time
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 4, dtype=torch.uint8, pin_memory=True); d2 = torch.empty(n // 4, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for _ in range(2): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(5): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); print("H2D alone GB/s", 5 * n / (time.perf_counter() - t) / 1e9)
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t
print("H2D with concurrent D2H (1/4 size): H2D GB/s", 5 * n / dt / 1e9, "D2H GB/s", 5 * n / 4 / dt / 1e9)
