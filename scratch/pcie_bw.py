import torch, time
n = 1600 * 1000 * 1000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device='cuda')
h2 = torch.empty(n // 2, dtype=torch.uint8, pin_memory=True); d2 = torch.empty(n // 2, dtype=torch.uint8, device='cuda')
for _ in range(2): d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); d.copy_(h, non_blocking=True); e1.record(); torch.cuda.synchronize()
print('H2D 1.6 GB: %.2f ms  %.1f GB/s' % (e0.elapsed_time(e1), n / e0.elapsed_time(e1) / 1e6))
e0.record(); h2.copy_(d2, non_blocking=True); e1.record(); torch.cuda.synchronize()
print('D2H 0.8 GB: %.2f ms  %.1f GB/s' % (e0.elapsed_time(e1), n / 2 / e0.elapsed_time(e1) / 1e6))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t0 = time.perf_counter()
with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) * 1e3
print('both directions at once: %.2f ms' % dt)
