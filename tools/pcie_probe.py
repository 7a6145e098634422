"""Raw PCIe ceilings of the box (pinned host buffers): D2H alone, H2D alone, both at once, D2H split over two streams."""
import torch
dev = torch.device("cuda", 0)
n = 469 << 20
d = torch.empty(n, dtype=torch.uint8, device=dev)
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(86 << 20, dtype=torch.uint8, device=dev)
h2 = torch.empty(86 << 20, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=8):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


t = timed(lambda: h.copy_(d, non_blocking=True)); print(f"D2H alone           {n / t / 1e6:7.1f} GB/s")
t = timed(lambda: d2.copy_(h2, non_blocking=True)); print(f"H2D alone           {(86 << 20) / t / 1e6:7.1f} GB/s")


def both():
    with torch.cuda.stream(s1):
        d2.copy_(h2, non_blocking=True)
    h.copy_(d, non_blocking=True)
t = timed(both); print(f"D2H with H2D beside {n / t / 1e6:7.1f} GB/s D2H")


def split():
    half = n // 2
    with torch.cuda.stream(s1):
        h[:half].copy_(d[:half], non_blocking=True)
    with torch.cuda.stream(s2):
        h[half:].copy_(d[half:], non_blocking=True)
t = timed(split); print(f"D2H on two streams  {n / t / 1e6:7.1f} GB/s")


def chunks():
    c = 38 << 20
    for o in range(0, n - c + 1, c):
        h[o:o + c].copy_(d[o:o + c], non_blocking=True)
t = timed(chunks); print(f"D2H in 38 MB copies {(n // (38 << 20)) * (38 << 20) / t / 1e6:7.1f} GB/s")
