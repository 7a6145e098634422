"""Experiment (GPU box): does keeping a CorrBlock's pyramid L2-resident pay?  One cfg2 step (29 pairs: build +
12 lookups + warp) is split into chunks of `pb` pairs, each chunk = its own CorrBlock (build + 12 lookups), the
whole step replayed as ONE CUDA graph.  A chunk of pb pairs has a 5.57 MB * pb pyramid (29 -> 161.5 MB > 126 MB L2).

    python tools/l2exp.py [--chunks 29,15,10,8,6,4] [--reps 50]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cardioflow_b200 as cf  # noqa: E402
from cardioflow_b200 import _cabi  # noqa: E402
from bench import WORKLOADS, make_unit, ITERS, LEVELS, RADIUS  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--chunks", default="29,15,10,8,6,4")
ap.add_argument("--reps", type=int, default=50)
ap.add_argument("--workload", default="cfg2")
args = ap.parse_args()
dev = torch.device("cuda:0")
wl = WORKLOADS[args.workload]
P, H, W = wl["pairs"], wl["H"], wl["W"]
lib = _cabi.lib()
nset = 3
sets = [make_unit(wl, seed=1002 + s, device=dev) for s in range(nset)]
outs = torch.empty(ITERS, P, LEVELS * 81, H, W, device=dev)
warped = torch.empty(P, 1, wl["IH"], wl["IW"], device=dev)

for pb in [int(x) for x in args.chunks.split(",")]:
    pyr = torch.empty(lib.cf_corr_pyramid_bytes(pb, H, W, LEVELS) // 4, device=dev)

    def step(u):
        for p0 in range(0, P, pb):
            p1 = min(P, p0 + pb)
            blk = cf.CorrBlock(u["f1"][p0:p1], u["f2"][p0:p1], num_levels=LEVELS, radius=RADIUS, pyramid_buffer=pyr)
            for k in range(ITERS):
                blk(u["coords"][k, p0:p1], out=outs[k, p0:p1])
        cf.warp(u["img"], u["flow"], out=warped)

    with torch.no_grad():
        step(sets[0])
        torch.cuda.synchronize()
        graphs = []
        for u in sets:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                step(u)
            graphs.append(g)
        for i in range(5):
            graphs[i % nset].replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.reps):
            graphs[i % nset].replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.reps
    print(f"chunk {pb:3d} pairs  pyramid {pyr.numel() * 4 / 1e6:7.1f} MB  {ms * 1e3:8.1f} us/step  {P / ms * 1e3:9.0f} pairs/s", flush=True)
    del graphs
