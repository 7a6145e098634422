"""Bring-up: per-CTA phase timeline of lookup_kernel (globaltimer ns).  GPU box only; needs a build with CF_NVCC_EXTRA=-DCF_BRINGUP
into CF_BUILD_DIR (tools/r2_call26.sh shows the recipe).  Prints phase durations and the residency picture per SM."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import cardioflow_b200 as cf
from cardioflow_b200 import _cabi
lib = _cabi.lib()
lib.cf_debug_set_lookup_timestamps.argtypes = [ctypes.c_void_p]
dev = "cuda:0"
P, C, H, W = (29, 256, 32, 32) if (len(sys.argv) < 2 or sys.argv[1] == "cfg2") else (64, 256, 64, 64)
g = torch.Generator(device=dev).manual_seed(1)
f1, f2 = torch.randn(P, C, H, W, device=dev, generator=g), torch.randn(P, C, H, W, device=dev, generator=g)
ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
coords = torch.stack([xs, ys]).float()[None].repeat(P, 1, 1, 1) + torch.randn(P, 2, H, W, device=dev, generator=g) * 1.5
with torch.no_grad():
    blk = cf.CorrBlock(f1, f2, precision="bf16")
    out = torch.empty(P, 4 * 81, H, W, device=dev)
    for _ in range(5):
        blk(coords, out=out)
    torch.cuda.synchronize()
    n_cta = (P * H * W + 31) // 32 * 4
    ts = torch.zeros(n_cta * 8, dtype=torch.int64, device=dev)
    lib.cf_debug_set_lookup_timestamps(ts.data_ptr())
    blk(coords, out=out)
    torch.cuda.synchronize()
    lib.cf_debug_set_lookup_timestamps(None)
t = ts.cpu().numpy().reshape(n_cta, 8).astype(np.float64)
t0 = t[:, 0].min()
rel = (t[:, :6] - t0) / 1e3
sm = t[:, 7].astype(int)
lvl = np.arange(n_cta) % 4
print(f"{n_cta} CTAs, kernel span {rel[:, 5].max():.2f} us (first start -> last end)")
names = ["start->coords", "coords->staged(issue)", "issue->taps done", "taps->windows landed", "landed->stores issued"]
for l in range(4):
    m = lvl == l
    d = np.diff(rel[m], axis=1)
    print(f"level {l}: " + " | ".join(f"{n} {d[:, i].mean():.2f}" for i, n in enumerate(names)) + f" | lifetime {(rel[m, 5] - rel[m, 0]).mean():.2f} us")
starts = np.sort(rel[:, 0])
print("CTA starts: " + " ".join(f"{q}%:{np.percentile(starts, q):.2f}" for q in (1, 10, 25, 50, 75, 90, 99)))
ends = np.sort(rel[:, 5])
print("CTA ends:   " + " ".join(f"{q}%:{np.percentile(ends, q):.2f}" for q in (1, 10, 25, 50, 75, 90, 99)))
# residency on one SM
for s_ in (0, 73, 147):
    m = sm == s_
    ev = sorted([(a, 1) for a in rel[m, 0]] + [(b, -1) for b in rel[m, 5]])
    cur, mx, area, last = 0, 0, 0.0, 0.0
    for tt, dlt in ev:
        area += cur * (tt - last); last = tt; cur += dlt; mx = max(mx, cur)
    print(f"SM {s_}: {m.sum()} CTAs, max resident {mx}, mean resident {area / max(last, 1e-9):.2f}, busy until {last:.2f} us")
