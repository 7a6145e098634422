"""Build time against the number of pairs (development tool): slope = steady-state time per pair, intercept = fixed cost."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cardioflow_b200 as cf
from cardioflow_b200 import _cabi
dev = torch.device("cuda:0")
lib = _cabi.lib()
H = W = int(os.environ.get("SW_HW", "32")); C = 256
prec = os.environ.get("SW_PREC", "bf16")
for P in [int(x) for x in os.environ.get("SW_P", "7,14,29,58,116,232").split(",")]:
    f1 = torch.randn(P, C, H, W, device=dev); f2 = torch.randn(P, C, H, W, device=dev)
    pyr = torch.empty(lib.cf_corr_pyramid_bytes(P, H, W, 4) // 4, device=dev)
    with torch.no_grad():
        for _ in range(3):
            cf.CorrBlock(f1, f2, precision=prec, pyramid_buffer=pyr)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            cf.CorrBlock(f1, f2, precision=prec, pyramid_buffer=pyr)
        ts = []
        for _ in range(15):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); g.replay(); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    print(f"build_sweep prec={prec} HW={H} P={P:4d} us={ts[len(ts)//2]:9.1f} us_per_pair={ts[len(ts)//2]/P:7.3f}", flush=True)
