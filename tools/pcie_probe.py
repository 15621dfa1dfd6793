import torch, time
n = 1 << 28  # 1 GiB of fp32
d = torch.empty(n, dtype=torch.float32, device="cuda")
h = torch.empty(n, dtype=torch.float32, pin_memory=True)
for name, fn in (("d2h", lambda: h.copy_(d, non_blocking=True)), ("h2d", lambda: d.copy_(h, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(5): fn()
    torch.cuda.synchronize()
    print(name, "contiguous GB/s", 5 * n * 4 / (time.perf_counter() - t) / 1e9)
# strided destination: 2002 rows x 8 KB out of 80 KB pitch (what cnf2_get_stats does for haplobase)
rows, w, pitch = 2002, 1000, 10000
d2 = torch.empty(rows, w, dtype=torch.float64, device="cuda")
h2 = torch.empty(rows, pitch, dtype=torch.float64, pin_memory=True)
for width in (1000, 4000):
    d2 = torch.empty(rows, width, dtype=torch.float64, device="cuda")
    view = h2[:, :width]
    view.copy_(d2, non_blocking=True); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(20): view.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    print("d2h 2-D rows of", width * 8, "B GB/s", 20 * rows * width * 8 / (time.perf_counter() - t) / 1e9)
