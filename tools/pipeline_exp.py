"""Experiment: cfg2 steps issued round-robin on D streams (D steps in flight, each with its own pyramid / output buffers, each
one CUDA graph of n lanes), against the sequential replay bench.py does.  Prints one line per variant."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import cardioflow_b200 as cf  # noqa: E402
from cardioflow_b200 import _cabi  # noqa: E402

ITERS, LEVELS, RADIUS = bench.ITERS, bench.LEVELS, bench.RADIUS


def main():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _cabi.lib()
    wl = bench.WORKLOADS["cfg2"]
    P, H, W, IH, IW = wl["pairs"], wl["H"], wl["W"], wl["IH"], wl["IW"]
    K81 = LEVELS * 81
    n_sets = 3
    sets = [bench.make_unit(wl, seed=1000 + s, device=dev) for s in range(n_sets)]
    steps = int(os.environ.get("PX_STEPS", "240"))
    for depth, n_lanes in ((1, 1), (1, 2), (2, 1), (2, 2), (3, 1), (3, 2), (4, 1)):
        bounds = cf.lane_bounds(P, n_lanes)
        slots = []
        with torch.no_grad():
            for d in range(depth):
                stream = torch.cuda.Stream(device=dev)
                outs = torch.empty(ITERS, P, K81, H, W, dtype=torch.float32, device=dev)
                warped = torch.empty(P, 1, IH, IW, dtype=torch.float32, device=dev)
                pyr = [torch.empty(lib.cf_corr_pyramid_bytes(b - a, H, W, LEVELS) // 4, dtype=torch.float32, device=dev) for a, b in bounds]
                lanes = cf.Lanes(len(bounds), dev)

                def chain(u, j, a, b, outs=outs, warped=warped, pyr=pyr):
                    bk = cf.CorrBlock(u["f1"][a:b], u["f2"][a:b], num_levels=LEVELS, radius=RADIUS, precision="bf16", pyramid_buffer=pyr[j])
                    for k in range(ITERS):
                        bk(u["coords"][k, a:b], out=outs[k, a:b])
                    cf.warp(u["img"][a:b], u["flow"][a:b], out=warped[a:b])
                    return bk
                graphs, keep = [], []
                with torch.cuda.stream(stream):
                    lanes.run(P, lambda j, a, b: chain(sets[0], j, a, b))
                    torch.cuda.synchronize()
                    for u in sets:
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, stream=stream):
                            keep.append(lanes.run(P, lambda j, a, b, u=u: chain(u, j, a, b)))
                        graphs.append(g)
                slots.append((stream, graphs, keep, outs, warped, pyr, lanes))
        torch.cuda.synchronize()

        def run(n):
            for i in range(n):
                stream, graphs = slots[i % depth][0], slots[i % depth][1]
                with torch.cuda.stream(stream):
                    graphs[i % n_sets].replay()
        run(2 * depth + 3)
        torch.cuda.synchronize()
        cur = torch.cuda.current_stream(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        for s in slots:
            s[0].wait_event(e0)
        run(steps)
        for s in slots:
            cur.wait_stream(s[0])
        e1.record(cur)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        print(f"pipeline depth={depth} lanes={n_lanes}: {ms * 1e3:7.1f} us/step  {P / ms * 1e3:9.0f} pairs/s", flush=True)
        del slots
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
