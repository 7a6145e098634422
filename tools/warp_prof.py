"""ncu target: a handful of warp launches (cfg3 shapes), one per case, after one warm-up each.
    ncu --set full -k regex:warp2d -o gpurun_out/prof_warp python tools/warp_prof.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cardioflow_b200 as cf
dev = torch.device("cuda:0")
B, IH = 300, 224
torch.manual_seed(0)
for cin in (1, 5):
    img = torch.rand(B, cin, IH, IH, device=dev)
    noise = torch.randn(B, 2, IH, IH, device=dev) * 3
    smooth = torch.nn.functional.interpolate(torch.randn(B, 2, 7, 7, device=dev) * 4, size=(IH, IH), mode="bicubic",
                                             align_corners=True).contiguous()
    out = torch.empty(B, cin, IH, IH, device=dev)
    for flow in (noise, smooth):
        cf.warp(img, flow, out=out)
        torch.cuda.synchronize()
lab = torch.randint(0, 4, (B, 1, IH, IH), device=dev)
cf.warp_onehot_argmax(smooth, lab)
torch.cuda.synchronize()
