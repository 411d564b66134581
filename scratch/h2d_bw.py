import torch, time
dev = torch.device("cuda", 0)
for mb in (33.5, 134):
    n = int(mb * 1e6) // 4
    h = torch.empty(n, dtype=torch.float32).pin_memory(); d = torch.empty(n, dtype=torch.float32, device=dev)
    s = torch.cuda.Stream()
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s)
        for _ in range(10): d.copy_(h, non_blocking=True)
        e1.record(s)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"H2D {mb} MB: {ms:.3f} ms = {n*4/ms/1e6:.1f} GB/s")
    with torch.cuda.stream(s):
        e0.record(s)
        for _ in range(10): h.copy_(d, non_blocking=True)
        e1.record(s)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"D2H {mb} MB: {ms:.3f} ms = {n*4/ms/1e6:.1f} GB/s")
