"""Experiment: (1) lookup time vs batch size (ramp/tail intercept vs steady-state slope), (2) a cfg2 step split into
independent sub-batches ("lanes") on parallel streams inside one CUDA graph.  Prints one line per variant."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import cardioflow_b200 as cf  # noqa: E402
from cardioflow_b200 import _cabi  # noqa: E402

ITERS, LEVELS, RADIUS = bench.ITERS, bench.LEVELS, bench.RADIUS


def timed(fn, reps=60, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def lookup_sweep(dev):
    lib = _cabi.lib()
    for P in (4, 8, 15, 29, 58, 116, 232):
        wl = dict(bench.WORKLOADS["cfg2"], pairs=P)
        sets = [bench.make_unit(wl, seed=7 + s, device=dev) for s in range(3)]
        H = W = 32
        K81 = LEVELS * 81
        pyr = [torch.empty(lib.cf_corr_pyramid_bytes(P, H, W, LEVELS) // 4, dtype=torch.float32, device=dev) for _ in range(3)]
        outs = torch.empty(ITERS, P, K81, H, W, dtype=torch.float32, device=dev)
        graphs = []
        with torch.no_grad():
            for u, pb in zip(sets, pyr):
                blk = cf.CorrBlock(u["f1"], u["f2"], num_levels=LEVELS, radius=RADIUS, pyramid_buffer=pb)
                blk(u["coords"][0], out=outs[0])
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    for k in range(ITERS):
                        blk(u["coords"][k], out=outs[k])
                graphs.append((g, blk))
        i = [0]

        def run():
            graphs[i[0] % 3][0].replay()
            i[0] += 1
        ms = timed(run) / ITERS
        print(f"lookup_sweep P={P:4d} lookup_us={ms * 1e3:8.2f} us_per_pair={ms * 1e3 / P:7.3f}", flush=True)
        del graphs, pyr, outs, sets
        torch.cuda.empty_cache()


def lanes(dev, n_lanes):
    wl = bench.WORKLOADS["cfg2"]
    P, H, W, IH, IW = wl["pairs"], wl["H"], wl["W"], wl["IH"], wl["IW"]
    K81 = LEVELS * 81
    lib = _cabi.lib()
    sets = [bench.make_unit(wl, seed=1002 + s, device=dev) for s in range(3)]
    cuts = [round(P * j / n_lanes) for j in range(n_lanes + 1)]
    pyrs = [torch.empty(lib.cf_corr_pyramid_bytes(cuts[j + 1] - cuts[j], H, W, LEVELS) // 4, dtype=torch.float32, device=dev)
            for j in range(n_lanes)]
    outs = torch.empty(ITERS, P, K81, H, W, dtype=torch.float32, device=dev)
    warped = torch.empty(P, 1, IH, IW, dtype=torch.float32, device=dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(n_lanes)]

    def lane_step(u, j):
        a, b = cuts[j], cuts[j + 1]
        blk = cf.CorrBlock(u["f1"][a:b], u["f2"][a:b], num_levels=LEVELS, radius=RADIUS, pyramid_buffer=pyrs[j])
        for k in range(ITERS):
            blk(u["coords"][k, a:b], out=outs[k, a:b])
        cf.warp(u["img"][a:b], u["flow"][a:b], out=warped[a:b])
        return blk

    def step(u):
        keep = []
        if n_lanes == 1:
            return [lane_step(u, 0)]
        cur = torch.cuda.current_stream(dev)
        for j, s in enumerate(streams):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                keep.append(lane_step(u, j))
        for s in streams:
            cur.wait_stream(s)
        return keep

    with torch.no_grad():
        step(sets[0])
        torch.cuda.synchronize()
        graphs, keep = [], []
        for u in sets:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                keep.append(step(u))
            graphs.append(g)
        i = [0]

        def run():
            graphs[i[0] % 3].replay()
            i[0] += 1
        ms = timed(run, reps=200)
    print(f"lanes={n_lanes} step_ms={ms:.4f} pairs_per_s={P / ms * 1e3:.0f}", flush=True)


if __name__ == "__main__":
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    for n in (1, 2, 3, 4):
        lanes(dev, n)
    lookup_sweep(dev)
