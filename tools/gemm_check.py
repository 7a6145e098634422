"""Bring-up check of the tcgen05 correlation kernel against torch fp32 matmul (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cardioflow_b200 as cf

dev = "cuda:0"
torch.manual_seed(0)
ok = True
for (B, C, H, W) in [(1, 32, 16, 16), (1, 256, 32, 32), (2, 64, 16, 16), (3, 256, 32, 32), (1, 256, 64, 64), (2, 128, 24, 24),
                     (1, 256, 28, 28), (2, 96, 16, 40), (29, 256, 32, 32)]:
    f1 = torch.randn(B, C, H, W, device=dev)
    f2 = torch.randn(B, C, H, W, device=dev)
    ref = torch.matmul(f1.view(B, C, -1).transpose(1, 2).double(), f2.view(B, C, -1).double()) / (C ** 0.5)
    try:
        blk = cf.CorrBlock(f1, f2, precision="bf16")
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        print(f"B={B} C={C} {H}x{W}: EXC {e}")
        ok = False
        continue
    got = blk.corr_pyramid[0].view(B, H * W, H * W).double()
    err = (got - ref).abs()
    rel = err.max().item() / ref.abs().max().item()
    pool_ref = torch.nn.functional.avg_pool2d(blk.corr_pyramid[0], 2, stride=2)
    perr = (blk.corr_pyramid[1] - pool_ref).abs().max().item()
    ex = cf.CorrBlock(f1, f2, precision="fp32")
    l3 = (blk.corr_pyramid[3] - ex.corr_pyramid[3]).abs().max().item() / ex.corr_pyramid[3].abs().max().item()
    status = "ok" if rel < 1e-2 and perr == 0 and l3 < 1e-2 else "FAIL"
    ok &= status == "ok"
    print(f"B={B} C={C} {H}x{W}: {blk.used_precision} L0 rel={rel:.3e} L1-vs-pool(L0) maxabs={perr:.3e} L3 rel={l3:.3e} {status}")
    if status == "FAIL":
        bad = (err > 1e-2 * ref.abs().max()).nonzero()
        print("   bad entries:", bad.shape[0], "first:", bad[:5].tolist(), "rows:", sorted(set((bad[:, 1] // 32).tolist()))[:20],
              "cols:", sorted(set((bad[:, 2] // 32).tolist()))[:20])
# timing
B, C, H, W = 29, 256, 32, 32
f1 = torch.randn(B, C, H, W, device=dev); f2 = torch.randn(B, C, H, W, device=dev)
lib = cf._cabi.lib()
buf = torch.empty(lib.cf_corr_pyramid_bytes(B, H, W, 4) // 4, device=dev)
for prec in ("bf16", "fp32"):
    for _ in range(3):
        cf.CorrBlock(f1, f2, precision=prec, pyramid_buffer=buf)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        cf.CorrBlock(f1, f2, precision=prec, pyramid_buffer=buf)
    e1.record(); torch.cuda.synchronize()
    print(f"corr_build {prec}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per 29-pair build")
print("ALL OK" if ok else "SOME FAILED")
