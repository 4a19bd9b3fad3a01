import time, torch
for mb in (3, 26, 256):
    a = torch.empty(mb << 20, dtype=torch.uint8).pin_memory(); b = torch.empty_like(a, device="cuda")
    for _ in range(3): b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
    print(f"pinned H2D {mb} MiB: {dt*1e3:.3f} ms  {mb/1024/dt:.1f} GiB/s")
