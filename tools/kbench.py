"""Per-kernel timing on the GPU box (CUDA events, rotating buffers larger than L2).  Development tool:
prints one line per kernel with its algorithmic bytes / flops and the achieved rate.

    python tools/kbench.py [cfg2|cfg4|cfg3] [--reps N]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cardioflow_b200 as cf  # noqa: E402
from cardioflow_b200 import _cabi  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("cfg", nargs="?", default="cfg2")
ap.add_argument("--reps", type=int, default=30)
ap.add_argument("--only", default="")
ap.add_argument("--graph", action="store_true")
args = ap.parse_args()
dev = torch.device("cuda:0")
if os.environ.get("KB_L2_FETCH"):
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    torch.zeros(1, device=dev)
    print("cudaDeviceSetLimit(MaxL2FetchGranularity) ->", rt.cudaDeviceSetLimit(5, ctypes.c_size_t(int(os.environ["KB_L2_FETCH"]))))
    v = ctypes.c_size_t()
    rt.cudaDeviceGetLimit(ctypes.byref(v), 5)
    print("L2 fetch granularity now", v.value)
if os.environ.get("KB_L2_PERSIST_MB"):
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    torch.zeros(1, device=dev)
    print("cudaDeviceSetLimit(PersistingL2CacheSize) ->", rt.cudaDeviceSetLimit(6, ctypes.c_size_t(int(os.environ["KB_L2_PERSIST_MB"]) << 20)))
    v = ctypes.c_size_t()
    rt.cudaDeviceGetLimit(ctypes.byref(v), 6)
    print("persisting L2 set-aside now", v.value / 2**20, "MB; max", torch.cuda.get_device_properties(0).L2_cache_size / 2**20, "MB L2")
lib = _cabi.lib()
PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
HBM, TF = PEAK["hbm_gbs"], PEAK["bf16_tflops"]


def timeit(fn, nrot, reps=args.reps):
    for i in range(3):
        fn(i % nrot)
    torch.cuda.synchronize()
    if args.graph:   # replay CUDA graphs: takes the Python / launch overhead out of small kernels' timings
        graphs = []
        for i in range(nrot):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn(i)
            graphs.append(g)
        eager, fn = fn, (lambda i: graphs[i].replay())
        fn(0)
        torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i, (a, b) in enumerate(evs):
        a.record()
        fn(i % nrot)
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def report(name, us, us_min, nbytes=None, flops=None):
    s = f"{name:34s} median {us:9.1f} us  min {us_min:9.1f} us"
    if nbytes:
        gbs = nbytes / us / 1e3
        s += f"  {nbytes / 1e6:9.1f} MB  {gbs:7.0f} GB/s = {100 * gbs / HBM:5.1f}% of measured HBM"
    if flops:
        tf = flops / us / 1e6
        s += f"  {flops / 1e9:8.1f} GFLOP {tf:7.1f} TF/s = {100 * tf / TF:5.1f}% of measured bf16"
    print(s, flush=True)


def want(k):
    return not args.only or k in args.only.split(",")


if args.cfg in ("cfg2", "cfg4", "cfg1", "ref24", "cfg2h", "cfg2q"):
    P, C, H, W, IH = {"cfg2": (29, 256, 32, 32, 256), "cfg4": (64, 256, 64, 64, 512), "cfg1": (1, 256, 28, 28, 224),
                      "ref24": (2, 256, 24, 24, 192), "cfg2h": (14, 256, 32, 32, 256), "cfg2q": (7, 256, 32, 32, 256)}[args.cfg]
    L, R = 4, 4
    nrot = 3 if args.cfg == "cfg2" else (2 if args.cfg == "cfg4" else 8)
    if os.environ.get("KB_NROT"):
        nrot = int(os.environ["KB_NROT"])
    torch.manual_seed(0)
    f1 = [torch.randn(P, C, H, W, device=dev) for _ in range(nrot)]
    f2 = [torch.randn(P, C, H, W, device=dev) for _ in range(nrot)]
    pyr_bytes = lib.cf_corr_pyramid_bytes(P, H, W, L)
    pyr = [torch.empty(pyr_bytes // 4, device=dev) for _ in range(nrot)]
    grid = cf.coords_grid(P, H, W, device=dev)
    coords = [grid + torch.randn(P, 2, H, W, device=dev) * 2.0 for _ in range(nrot)]
    outs = [torch.empty(P, L * 81, H, W, device=dev) for _ in range(nrot)]
    HWn = H * W
    blks = {}
    for prec in ("bf16", "bf16x3", "fp32"):
        if not want("build_" + prec):
            continue
        us, mn = timeit(lambda i: blks.__setitem__(i, cf.CorrBlock(f1[i], f2[i], precision=prec, pyramid_buffer=pyr[i])), nrot,
                        reps=max(5, args.reps // (1 if prec == "bf16" else 4)))
        lv = sum((H >> k) * (W >> k) for k in range(L))
        nb = P * (2 * C * HWn * 4 + 4 * HWn * lv)
        report(f"corr_build[{prec}] (gemm+pyramid)", us, mn, nb, 2.0 * P * HWn * HWn * C)
    for i in range(nrot):
        blks[i] = cf.CorrBlock(f1[i], f2[i], precision=os.environ.get("KB_LOOKUP_PREC", "bf16"), pyramid_buffer=pyr[i])
    from oracle.corr_oracle import lookup_algorithmic_bytes
    lb = lookup_algorithmic_bytes(P, H, W, L, R)
    if want("lookup"):
        us, mn = timeit(lambda i: blks[i](coords[i], out=outs[i]), nrot)
        report("lookup (noise coords)", us, mn, lb)
        us, mn = timeit(lambda i: blks[i](grid, out=outs[i]), nrot)
        report("lookup (integer grid, iter 0)", us, mn, lb)
    if want("lookup"):
        traj = [grid + torch.randn(P, 2, H, W, device=dev) * 0.5 * (k ** 0.5) for k in range(12)]
        outs12 = [torch.empty(P, L * 81, H, W, device=dev) for _ in range(12)]

        def twelve(i):
            for k in range(12):
                blks[i](traj[k], out=outs12[k])
        us, mn = timeit(twelve, nrot, reps=max(5, args.reps // 3))
        report("lookup x12 same pyramid (per call)", us / 12, mn / 12, lb)
    if want("otf"):     # K4: lookup without a materialised volume -- tensor cores (tcgen05) vs the fp32 CUDA-core variant
        flops = 0
        hh, ww = H, W
        for _ in range(L):
            flops += 2.0 * C * HWn * min(100, hh * ww)
            hh, ww = hh // 2, ww // 2
        flops *= P
        with torch.no_grad():
            for prec, label in (("bf16", "tcgen05 bf16"), ("fp32", "CUDA cores fp32")):
                if prec == "fp32" and args.cfg == "cfg4" and not os.environ.get("KB_OTF_SIMT"):
                    continue
                ob = [cf.CorrBlock(f1[i], f2[i], on_the_fly=True, precision=prec) for i in range(nrot)]
                us, mn = timeit(lambda i: ob[i](coords[i], out=outs[i]), nrot, reps=max(5, args.reps // 3))
                report(f"on-the-fly lookup [{label}]", us, mn, None, flops)
                del ob
    if want("pool"):
        us, mn = timeit(lambda i: _cabi.check(lib.cf_corr_pool(pyr[i].data_ptr(), P, H, W, L, 0, _cabi.stream())), nrot)
        lv = [(H >> k) * (W >> k) for k in range(L)]
        report("pool chain L0->L3", us, mn, P * HWn * 4 * (lv[0] + 2 * lv[1] + 2 * lv[2] + lv[3]))
    if want("warp"):
        for cin in (1, 5):
            img = [torch.rand(P, cin, IH, IH, device=dev) for _ in range(nrot)]
            flow = [torch.randn(P, 2, IH, IH, device=dev) * 3 for _ in range(nrot)]
            wout = [torch.empty(P, cin, IH, IH, device=dev) for _ in range(nrot)]
            us, mn = timeit(lambda i: cf.warp(img[i], flow[i], out=wout[i]), nrot)
            report(f"warp {P}x{cin}x{IH}^2", us, mn, 4 * P * IH * IH * (2 * cin + 2))

if args.cfg in ("cfg2", "cfg4", "cfg1") and want("bwd"):
    # backward passes: cf_corr_lookup_bwd, cf_corr_pool_bwd, cf_corr_volume_bwd, cf_warp2d_bwd
    st_ = _cabi.stream
    go = [torch.randn(P, L * 81, H, W, device=dev) for _ in range(nrot)]
    gp = [torch.zeros(pyr_bytes // 4, device=dev) for _ in range(nrot)]
    us, mn = timeit(lambda i: _cabi.check(lib.cf_corr_lookup_bwd(coords[i].data_ptr(), go[i].data_ptr(), P, H, W, L, R,
                                                                  gp[i].data_ptr(), st_())), nrot)
    report("lookup backward (accumulating)", us, mn, lb)
    us, mn = timeit(lambda i: _cabi.check(lib.cf_corr_pool_bwd(gp[i].data_ptr(), P, H, W, L, st_())), nrot)
    lv = [(H >> k) * (W >> k) for k in range(L)]
    report("pool chain backward", us, mn, P * HWn * 4 * (2 * lv[0] + 2 * lv[1] + 2 * lv[2] + lv[3]))
    gf1, gf2 = torch.empty_like(f1[0]), torch.empty_like(f2[0])
    if not os.environ.get("KB_SKIP_SIMT_BWD"):
        us, mn = timeit(lambda i: _cabi.check(lib.cf_corr_volume_bwd(f1[i].data_ptr(), f2[i].data_ptr(), gp[i].data_ptr(), P, C, H, W,
                                                                      gf1.data_ptr(), gf2.data_ptr(), st_())), nrot, reps=max(5, args.reps // 4))
        report("corr volume backward (2 GEMMs), fp32 CUDA cores", us, mn, P * (4 * HWn * HWn + 16 * C * HWn), 4.0 * P * HWn * HWn * C)
    wsb = lib.cf_corr_volume_bwd_workspace_bytes(P, C, H, W)
    if wsb:
        import ctypes
        wsv = torch.empty(wsb, dtype=torch.uint8, device=dev)
        usedp = ctypes.c_int(0)
        us, mn = timeit(lambda i: _cabi.check(lib.cf_corr_volume_bwd_ex(f1[i].data_ptr(), f2[i].data_ptr(), gp[i].data_ptr(), P, C, H, W,
                                                                         _cabi.PRECISIONS["auto"], gf1.data_ptr(), gf2.data_ptr(), wsv.data_ptr(), wsb,
                                                                         ctypes.byref(usedp), st_())), nrot, reps=max(5, args.reps // 4))
        # G is read once per product (2 x 4 HW^2), the maps are read and the gradients written
        report(f"corr volume backward (2 GEMMs), tcgen05 {_cabi.PRECISION_NAMES.get(usedp.value, usedp.value)}", us, mn,
               P * (8 * HWn * HWn + 16 * C * HWn), 4.0 * P * HWn * HWn * C)
        del wsv
    for cin in (1, 5):
        img = [torch.rand(P, cin, IH, IH, device=dev) for _ in range(nrot)]
        flow = [torch.randn(P, 2, IH, IH, device=dev) * 3 for _ in range(nrot)]
        gout = [torch.randn(P, cin, IH, IH, device=dev) for _ in range(nrot)]
        gs, gfl = torch.zeros_like(img[0]), torch.empty_like(flow[0])
        us, mn = timeit(lambda i: _cabi.check(lib.cf_warp2d_bwd(flow[i].data_ptr(), img[i].data_ptr(), gout[i].data_ptr(), P, cin, IH, IH,
                                                                 gs.data_ptr(), gfl.data_ptr(), st_())), nrot)
        report(f"warp backward {P}x{cin}x{IH}^2 (both grads)", us, mn, 4 * P * IH * IH * (4 * cin + 4))
    del go, gp

if args.cfg in ("cfg2", "cfg4") and want("upsample"):
    from oracle.upsample_oracle import algorithmic_bytes
    for cin, fn in ((2, cf.upsample_flow), (4, cf.upsample_seg)):
        fld = [torch.randn(P, cin, H, W, device=dev) for _ in range(nrot)]
        msk = [torch.randn(P, 576, H, W, device=dev) for _ in range(nrot)]
        with torch.no_grad():
            us, mn = timeit(lambda i: fn(fld[i], msk[i]), nrot)
        report(f"convex upsample {P}x{cin}x{H}^2 -> {8 * H}^2", us, mn, algorithmic_bytes(P, cin, H, W))

if args.cfg == "cfg3":
    B, IH = 300, 224
    nrot = 2
    for cin in (1, 4, 5):
        img = [torch.rand(B, cin, IH, IH, device=dev) for _ in range(nrot)]
        flow = [torch.randn(B, 2, IH, IH, device=dev) * 3 for _ in range(nrot)]
        smooth = [torch.nn.functional.interpolate(torch.randn(B, 2, 7, 7, device=dev) * 4, size=(IH, IH), mode="bicubic",
                                                  align_corners=True).contiguous() for _ in range(nrot)]
        wout = [torch.empty(B, cin, IH, IH, device=dev) for _ in range(nrot)]
        us, mn = timeit(lambda i: cf.warp(img[i], flow[i], out=wout[i]), nrot)
        report(f"warp {B}x{cin}x{IH}^2 noise flow", us, mn, 4 * B * IH * IH * (2 * cin + 2))
        us, mn = timeit(lambda i: cf.warp(img[i], smooth[i], out=wout[i]), nrot)
        report(f"warp {B}x{cin}x{IH}^2 smooth flow", us, mn, 4 * B * IH * IH * (2 * cin + 2))
    lab = [torch.randint(0, 4, (B, 1, IH, IH), device=dev) for _ in range(nrot)]
    us, mn = timeit(lambda i: cf.warp_onehot_argmax(smooth[i], lab[i]), nrot)
    report(f"onehot-argmax {B}x{IH}^2", us, mn, B * IH * IH * 24)
    if want("ncc"):   # the registration loss on the warped frames of the same batch (lib/loss.py:15-76)
        nrot = 4      # 2 x 60 MB per set: rotate over > L2
        I = [torch.rand(B, 1, IH, IH, device=dev) for _ in range(nrot)]
        J = [(I[k] + 0.05 * torch.randn(B, 1, IH, IH, device=dev)).contiguous() for k in range(nrot)]
        npx = B * IH * IH
        nws = lib.cf_ncc_workspace_bytes(B, IH, IH, 0)
        ws = torch.empty(nws // 4 + 1, device=dev)
        omap = torch.empty(B, 1, IH, IH, device=dev)
        us, mn = timeit(lambda i: _cabi.check(lib.cf_ncc(I[i].data_ptr(), J[i].data_ptr(), B, IH, IH, 9, 1e-3, None,
                                                          ws[nws // 4:].data_ptr(), ws.data_ptr(), nws, _cabi.stream())), nrot)
        report(f"ncc mean {B}x{IH}^2", us, mn, 8 * npx)
        us, mn = timeit(lambda i: _cabi.check(lib.cf_ncc(I[i].data_ptr(), J[i].data_ptr(), B, IH, IH, 9, 1e-3, omap.data_ptr(),
                                                          None, None, 0, _cabi.stream())), nrot)
        report(f"ncc map {B}x{IH}^2", us, mn, 12 * npx)
        bws = torch.empty(5 * npx, device=dev)
        g1 = torch.ones(1, device=dev)
        gI, gJ = torch.empty_like(I[0]), torch.empty_like(I[0])
        us, mn = timeit(lambda i: _cabi.check(lib.cf_ncc_bwd(I[i].data_ptr(), J[i].data_ptr(), B, IH, IH, 9, 1e-3, g1.data_ptr(), 0,
                                                              -1.0 / npx, gI.data_ptr(), gJ.data_ptr(), bws.data_ptr(), 20 * npx,
                                                              _cabi.stream())), nrot)
        report(f"ncc backward, both grads (2 launches) {B}x{IH}^2", us, mn, 16 * npx)
        us, mn = timeit(lambda i: _cabi.check(lib.cf_ncc_bwd(I[i].data_ptr(), J[i].data_ptr(), B, IH, IH, 9, 1e-3, g1.data_ptr(), 0,
                                                              -1.0 / npx, None, gJ.data_ptr(), bws.data_ptr(), 20 * npx,
                                                              _cabi.stream())), nrot)
        report(f"ncc backward, grad_J only (2 launches) {B}x{IH}^2", us, mn, 12 * npx)
        import torch.nn.functional as F
        wones = torch.ones(1, 1, 9, 9, device=dev)

        def aten_ncc(i):    # the reference's formulation on ATen CUDA kernels (5 conv2d + elementwise), for scale only
            a, b = I[i], J[i]
            s = [F.conv2d(t, wones, padding=4) for t in (a, b, a * a, b * b, a * b)]
            uI, uJ = s[0] / 81.0, s[1] / 81.0
            cross = s[4] - uJ * s[0] - uI * s[1] + uI * uJ * 81.0
            iv = s[2] - 2 * uI * s[0] + uI * uI * 81.0
            jv = s[3] - 2 * uJ * s[1] + uJ * uJ * 81.0
            return 1 - torch.clamp(cross * cross / (iv * jv + 1e-3), 0.001, 0.999).mean()
        us, mn = timeit(aten_ncc, nrot, reps=10)
        report(f"(ATen conv2d formulation, same GPU) {B}x{IH}^2", us, mn, 8 * npx)
