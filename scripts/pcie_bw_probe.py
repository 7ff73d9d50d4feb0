import torch, time
x = torch.empty(1<<30, dtype=torch.uint8).pin_memory()
y = torch.empty(1<<30, dtype=torch.uint8, device='cuda')
for n in (1<<30, 64<<20):
    for _ in range(2): y[:n].copy_(x[:n], non_blocking=True)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 5 if n > (1<<29) else 40
    for _ in range(reps): y[:n].copy_(x[:n], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print("H2D", n>>20, "MiB:", n*reps/ (e0.elapsed_time(e1)*1e-3)/1e9, "GB/s")
    e0.record()
    for _ in range(reps): x[:n].copy_(y[:n], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print("D2H", n>>20, "MiB:", n*reps/ (e0.elapsed_time(e1)*1e-3)/1e9, "GB/s")
