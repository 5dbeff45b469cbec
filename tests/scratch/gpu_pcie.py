import torch, time
n = 60 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device='cuda')
h2 = torch.empty(15 << 20, dtype=torch.uint8).pin_memory(); d2 = torch.empty(15 << 20, dtype=torch.uint8, device='cuda')
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for label, both in (("H2D alone", False), ("H2D + D2H concurrently", True)):
    ts = []
    for _ in range(10):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s1):
            e0.record(); d.copy_(h, non_blocking=True); e1.record()
        if both:
            with torch.cuda.stream(s2):
                h2.copy_(d2, non_blocking=True)
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"{label}: 60 MiB in {t:.3f} ms = {n / t / 1e6:.1f} GB/s")
ts = []
for _ in range(10):
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); h2.copy_(d2, non_blocking=True); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
t = sorted(ts)[5]; print(f"D2H alone: 15 MiB in {t:.3f} ms = {(15 << 20) / t / 1e6:.1f} GB/s")
# zero-copy read by a kernel: torch cannot map host memory into kernels directly; skip
