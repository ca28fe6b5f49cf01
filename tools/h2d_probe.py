import torch, time
x = torch.empty(64, 1920, 3840, dtype=torch.uint8).pin_memory()
y = torch.empty_like(x, device="cuda")
for _ in range(2):
    y.copy_(x, non_blocking=True); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    y.copy_(x, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print("H2D one 472 MB copy: %.2f ms = %.1f GB/s" % (dt * 1e3, x.numel() / dt / 1e9))
xs = [x[i] for i in range(64)]
t0 = time.perf_counter()
for _ in range(5):
    for i in range(64):
        y[i].copy_(xs[i], non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print("H2D 64 x 7.4 MB copies: %.2f ms = %.1f GB/s" % (dt * 1e3, x.numel() / dt / 1e9))
# copies on a side stream, alone and with kernels running on another stream
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
a = torch.empty(1 << 28, dtype=torch.uint8, device="cuda")
for busy in (False, True):
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(sa):
        e0.record()
        for i in range(64):
            y[i].copy_(xs[i], non_blocking=True)
        e1.record()
    if busy:
        with torch.cuda.stream(sb):
            for _ in range(200):
                a.add_(1)           # HBM-bound kernels, ~0.1 ms each
    torch.cuda.synchronize()
    print("H2D on a side stream, kernels %s: %.2f ms" % ("running beside it" if busy else "idle", e0.elapsed_time(e1)))
