"""HBM write-only vs copy bandwidth (GPU box): is a pure-store kernel slower than the copy peak?"""
import torch
dev = "cuda:0"
for mb in (38, 152, 608, 2432):
    n = mb * 1024 * 1024 // 4
    bufs = [torch.empty(n, device=dev) for _ in range(4)]
    src = torch.randn(n, device=dev)
    def timeit(fn, reps=20):
        for i in range(3): fn(i)
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for i, (a, b) in enumerate(ev):
            a.record(); fn(i); b.record()
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) for a, b in ev)
        return ts[len(ts) // 2] * 1e3
    t_fill = timeit(lambda i: bufs[i % 4].zero_())
    t_copy = timeit(lambda i: bufs[i % 4].copy_(src))
    print(f"{mb:5d} MB  fill {t_fill:8.1f} us = {mb * 1.048576 / t_fill * 1e3:7.0f} GB/s written   copy {t_copy:8.1f} us = {2 * mb * 1.048576 / t_copy * 1e3:7.0f} GB/s moved")
