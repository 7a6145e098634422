#!/usr/bin/env python
"""Benchmark of the optical-flow hot path: frame-pairs/sec for (correlation pyramid build + 12 lookups + warp).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cfg2|cfg3|cfg4|cfg5 [--patients P --slices S]] [--no-graph]

One "step" = one (patient, slice) cine unit of BASELINE.json configs[1]: 29 consecutive frame pairs with
256-channel 32x32 feature maps -> CorrBlock build (batched over the 29 pairs), 12 CorrBlock lookups, and the
29 SpatialTransformer warps of the 256x256 frames.  Data is synthetic (seed 1002).
For N > 1 (torchrun, one rank per GPU) every rank runs its own unit per step: weak scaling, no data-path
collective (the path shards by (patient, slice)); time is the max over ranks.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput; `e2e` = the same metric through the
public module API with pinned HOST buffers (H2D of every input + D2H of every output inside the timed
region); `roofline` = the dominant kernel (the lookup) against the measured HBM peak; `cpu_baseline` = the
oracle (torch CPU kernels, the ATen ops the reference's Python dispatches to) on a bounded sample.
`--impl reference` times that CPU implementation alone (the reference arm).  `--impl reference-gpu` (informational, cfg2 / cfg4)
times the same oracle with torch CUDA ops on one B200: the reference "as actually run" (SURVEY.md 8(d)).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner, warnings), so fd 1 is pointed
# at stderr for the whole run and the result line goes to a duplicate of the original stdout.
_RESULT_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)

import torch  # noqa: E402


def emit(line):
    _RESULT_OUT.write(json.dumps(line) + "\n")
    _RESULT_OUT.flush()


WORKLOADS = {
    # name: pairs, C, fmap H, W, image H, W
    "cfg1": dict(pairs=1, C=256, H=28, W=28, IH=224, IW=224,
                 desc="configs[0]: single frame pair, 256-ch fmaps 28x28 (224^2 input), 4 levels, r=4, 12 lookups, one 224^2 warp"),
    "cfg2": dict(pairs=29, C=256, H=32, W=32, IH=256, IW=256,
                 desc="configs[1]: 29 frame pairs, 256-ch fmaps 32x32 (256^2 input), 4 levels, r=4, 12 lookups, 29 warps"),
    "cfg5": dict(pairs=29, C=256, H=32, W=32, IH=256, IW=256,
                 desc="configs[4]: cohort of patients x slices cine units (29 pairs each, 32x32 fmaps, 256^2 frames + 4-class masks), "
                      "sharded by (patient, slice), one NCCL gather"),
    "cfg3": dict(items=300, pairs=300, C=1, H=28, W=28, IH=224, IW=224,
                 desc="configs[2]: SpatialTransformer warp of image + one-hot 4-class masks by dense flow, 10 slices x 30 frames at 224^2"),
    "cfg4": dict(pairs=64, C=256, H=64, W=64, IH=512, IW=512,
                 desc="configs[3]: 64 frame pairs, 256-ch fmaps 64x64, materialised volume, 12 lookups, 64 warps"),
}
ITERS, LEVELS, RADIUS = 12, 4, 4
METRIC = "frame-pairs/sec (corr+12-iter lookup+warp, 256^2)"


def lookup_algorithmic_bytes(b, h, w, num_levels=4, radius=4):
    """SURVEY.md 8(d) K3, compulsory traffic of one lookup call: out 4*L*K*HW + coords 8*HW + window footprint
    sum_l 4*HW*min((2r+2)^2, h_l*w_l) per pair (each input read once, each output written once)."""
    hw, k = h * w, (2 * radius + 1) ** 2
    foot, hl, wl = 0, h, w
    for _ in range(num_levels):
        foot += min((2 * radius + 2) ** 2, hl * wl)
        hl, wl = hl // 2, wl // 2
    return b * (4 * num_levels * k * hw + 8 * hw + 4 * hw * foot)


# ------------------------------------------------------------------------------------------------ data
def make_unit(wl, seed, device="cpu", pin=False):
    """Synthetic cine unit (SURVEY.md 8(d) cfg 2 recipe): consecutive frames are shifted/noisy copies."""
    g = torch.Generator().manual_seed(seed)
    P, C, H, W, IH, IW = wl["pairs"], wl["C"], wl["H"], wl["W"], wl["IH"], wl["IW"]
    base = torch.randn(C, H, W, generator=g)
    frames = [base]
    for t in range(P):
        frames.append(torch.roll(frames[-1], shifts=(t % 3 - 1, 1), dims=(1, 2)) + 0.1 * torch.randn(C, H, W, generator=g))
    frames = torch.stack(frames)
    f1, f2 = frames[:-1].contiguous(), frames[1:].contiguous()
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(P, 1, 1, 1)
    coords, cur = [], grid.clone()
    for k in range(ITERS):
        coords.append(cur.clone())
        cur = cur + torch.randn(P, 2, H, W, generator=g) * 0.5
        cur[:, 0].clamp_(-4, W + 4)
        cur[:, 1].clamp_(-4, H + 4)
    coords = torch.stack(coords)                                   # [12, P, 2, H, W]
    img = torch.rand(P, 1, IH, IW, generator=g)
    flow = torch.nn.functional.interpolate(torch.randn(P, 2, 9, 9, generator=g) * 3.0, size=(IH, IW),
                                           mode="bicubic", align_corners=True).contiguous()
    unit = dict(f1=f1, f2=f2, coords=coords, img=img, flow=flow)
    if pin:
        unit = {k: v.pin_memory() for k, v in unit.items()}
    if device != "cpu":
        unit = {k: v.to(device) for k, v in unit.items()}
    return unit


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.stop_flag, self.thread = [], set(), False, None
        self.max_mhz, self.ok = None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = str(e)

    def _reasons(self):
        nv = self.nv
        fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        mask = fn(self.h)
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}
        return {n for bit, n in names.items() if mask & bit}

    def _run(self):
        while not self.stop_flag:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.reasons |= self._reasons()
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.002)

    def start(self):
        if self.ok:
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join()
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no NVML samples"}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_pairs_per_sec(wl, sample_pairs, budget_s=12.0, threads=None):
    """The reference's CPU path (oracle: torch CPU kernels) on `sample_pairs` pairs of the workload."""
    from oracle import corr_oracle, warp_oracle
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    unit = make_unit(dict(wl, pairs=sample_pairs), seed=1002)
    st = warp_oracle.SpatialTransformer((wl["IH"], wl["IW"]))

    def run():
        # per pair, like the reference's frame loop (SegFlowGaussian.py:907-954): a CorrBlock per pair
        for p in range(sample_pairs):
            blk = corr_oracle.CorrBlock(unit["f1"][p:p + 1], unit["f2"][p:p + 1], num_levels=LEVELS, radius=RADIUS)
            for k in range(ITERS):
                blk(unit["coords"][k, p:p + 1])
            st(flow=unit["flow"][p:p + 1], original=unit["img"][p:p + 1])

    with torch.no_grad():
        run()
        best, t_start, reps = float("inf"), time.perf_counter(), 0
        while reps < 3 or (time.perf_counter() - t_start < budget_s and reps < 50):
            t0 = time.perf_counter()
            run()
            best = min(best, time.perf_counter() - t0)
            reps += 1
    return sample_pairs / best, threads, reps


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = 1 if args.workload in ("cfg4", "cfg1") else 4
    threads = args.cpu_threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    from oracle import corr_oracle, warp_oracle
    if args.workload == "cfg3":          # the warp alone: 30 of the 300 five-channel frames per step
        n, IH, IW = 30, wl["IH"], wl["IW"]
        g = torch.Generator().manual_seed(1003)
        src, flow = torch.rand(n, 5, IH, IW, generator=g), torch.randn(n, 2, IH, IW, generator=g) * 3
        st = warp_oracle.SpatialTransformer((IH, IW))
        with torch.no_grad():
            for _ in range(args.warmup):
                st(flow=flow, original=src)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                st(flow=flow, original=src)
            dt = time.perf_counter() - t0
        value = n * args.steps / dt
        desc = f"{n} of {wl['items']} frames per step (5-channel SpatialTransformer warp), torch CPU fp32"
        emit({"impl": "reference", "metric": "frames warped/sec (image + 4 one-hot masks, 224^2)", "value": value,
                          "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "f32", "data": "synthetic", "config": {"workload": "cfg3: " + wl["desc"], "sample": desc},
                          "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": "port", "sample": desc},
                          "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0})
        return
    unit = make_unit(dict(wl, pairs=sample), seed=1002)
    st = warp_oracle.SpatialTransformer((wl["IH"], wl["IW"]))

    def step():
        for p in range(sample):
            blk = corr_oracle.CorrBlock(unit["f1"][p:p + 1], unit["f2"][p:p + 1], num_levels=LEVELS, radius=RADIUS)
            for k in range(ITERS):
                blk(unit["coords"][k, p:p + 1])
            st(flow=unit["flow"][p:p + 1], original=unit["img"][p:p + 1])

    with torch.no_grad():
        for _ in range(args.warmup):
            step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    sample_desc = f"{sample} of {wl['pairs']} pairs per step (per-pair CorrBlock + 12 lookups + 1 warp), torch CPU fp32"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "frame-pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "pairs_per_step_per_gpu": wl["pairs"], "pairs_timed_per_step": sample,
                   "corr_precision": "fp32 (torch CPU kernels)", "sample": sample_desc,
                   "same_config_note": "same workload and metric as the GPU arm; a CPU step times `pairs_timed_per_step` of the "
                                       "`pairs_per_step_per_gpu` pairs (per-pair cost does not depend on the count) so that the run "
                                       "ends within minutes; value = pairs timed / time"},
        "cpu_baseline": {"value": value, "unit": "frame-pairs/s", "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": value, "unit": "frame-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)



def run_reference_gpu(args, wl):
    """SURVEY.md 8(d) / BASELINE.md section 4 "reference as actually run": the same oracle (the ATen ops the reference's Python
    dispatches to: matmul, avg_pool2d, grid_sample, cat/permute) executed with torch CUDA kernels on one B200, device-resident
    inputs, CUDA-event timing.  Informational "before" for the hand-written kernels; not the driver's reference arm (that is the
    CPU path, --impl reference).  Two figures: the reference's own loop structure (one CorrBlock per pair, SegFlowGaussian.py:907-954)
    and the best case for ATen (all pairs of the step in one batched CorrBlock)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if not torch.cuda.is_available():
        sys.exit("--impl reference-gpu needs a CUDA device")
    from oracle import corr_oracle, warp_oracle
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    P = wl["pairs"]
    unit = make_unit(wl, seed=1002, device=dev)
    st = warp_oracle.SpatialTransformer((wl["IH"], wl["IW"])).to(dev)

    def per_pair():
        for p in range(P):
            blk = corr_oracle.CorrBlock(unit["f1"][p:p + 1], unit["f2"][p:p + 1], num_levels=LEVELS, radius=RADIUS)
            for k in range(ITERS):
                blk(unit["coords"][k, p:p + 1])
            st(flow=unit["flow"][p:p + 1], original=unit["img"][p:p + 1])

    def batched():
        blk = corr_oracle.CorrBlock(unit["f1"], unit["f2"], num_levels=LEVELS, radius=RADIUS)
        for k in range(ITERS):
            blk(unit["coords"][k])
        st(flow=unit["flow"], original=unit["img"])

    def timed(fn, steps):
        for _ in range(max(args.warmup, 1)):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    with torch.no_grad():
        steps = max(1, min(args.steps, 20))
        ms_pair = timed(per_pair, steps)
        batch_note = ""
        try:
            ms_batch = timed(batched, steps)
        except RuntimeError as e:          # cuDNN's grid_sample refuses very large N (cfg4: 262 144 "images"): native ATen kernel
            torch.cuda.synchronize()
            batch_note = " [cuDNN grid_sample refused this batch (%s); timed with torch.backends.cudnn disabled]" % str(e).split(".")[0]
            with torch.backends.cudnn.flags(enabled=False):
                ms_batch = timed(batched, steps)
    emit({"impl": "reference-gpu", "metric": METRIC, "value": P / (ms_pair * 1e-3), "unit": "frame-pairs/s", "n_gpus": 1,
          "steps": steps, "warmup": max(args.warmup, 1), "ms_per_step": ms_pair, "higher_is_better": True, "scaling": "weak",
          "vs_baseline": None, "dtype": "f32 (ATen CUDA: matmul may use TF32 only if the user enables it; defaults kept)",
          "data": "synthetic",
          "config": {"workload": args.workload + ": " + wl["desc"],
                     "what": "torch CUDA eager ops of the oracle (matmul, /sqrt(C), 3 x avg_pool2d, per level grid_sample + cat + "
                             "permute), one CorrBlock per pair as the reference loops; device-resident inputs"},
          "batched": {"value": P / (ms_batch * 1e-3), "ms_per_step": ms_batch,
                      "what": "all pairs of the step in ONE batched CorrBlock / lookup / warp call (best case for ATen)" + batch_note},
          "gpu_launches": 0})


# ------------------------------------------------------------------------------------------------ cfg5 cohort
def make_unit_device(wl, seed, dev):
    """cfg5 unit generated ON the device from its own seed (any rank can build any unit): the cfg2 recipe plus a
    4-class label map (concentric-ring phantom) that is warped through one-hot/argmax."""
    g = torch.Generator(device=dev).manual_seed(seed)
    P, C, H, W, IH, IW = wl["pairs"], wl["C"], wl["H"], wl["W"], wl["IH"], wl["IW"]
    frames = torch.randn(P + 1, C, H, W, generator=g, device=dev)
    frames[1:] = 0.9 * frames[:-1].roll(shifts=(1, 1), dims=(2, 3)) + 0.45 * frames[1:]
    ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None]
    steps = torch.randn(ITERS, P, 2, H, W, generator=g, device=dev) * 0.5
    steps[0] = 0
    coords = grid + steps.cumsum(0)
    coords[:, :, 0].clamp_(-4, W + 4)
    coords[:, :, 1].clamp_(-4, H + 4)
    img = torch.rand(P, 1, IH, IW, generator=g, device=dev)
    flow = torch.nn.functional.interpolate(torch.randn(P, 2, 9, 9, generator=g, device=dev) * 3.0, size=(IH, IW),
                                           mode="bicubic", align_corners=True).contiguous()
    yy, xx = torch.meshgrid(torch.arange(IH, device=dev), torch.arange(IW, device=dev), indexing="ij")
    rad = ((yy - IH / 2) ** 2 + (xx - IW / 2) ** 2).float().sqrt() / (IH / 2)
    label = (rad < 0.2).long() * 3 + ((rad >= 0.2) & (rad < 0.35)).long() * 2 + ((rad >= 0.35) & (rad < 0.6)).long()
    label = label[None, None].repeat(P, 1, 1, 1).contiguous()
    return dict(f1=frames[:-1].contiguous(), f2=frames[1:].contiguous(), coords=coords.contiguous(), img=img, flow=flow,
                label=label)


def pin_to_gpu_numa(local_rank):
    """Bind this rank's host threads to the CPUs NVML reports as local to its GPU BEFORE any pinned buffer is allocated: pinned
    memory is placed on the NUMA node of the allocating thread, and eight ranks allocating and copying from one socket share
    that socket's memory controllers and PCIe root complexes (e2e efficiency 0.22 at 8 GPUs)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
        try:
            node = pynvml.nvmlDeviceGetNumaNodeId(h)
        except Exception:  # noqa: BLE001
            node = None
        return {"cpus": len(cpus), "numa_node": node}
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:80]}


def cohort_measure(args, wl, dev, rank, world, dist, patients, slices, flow_gather=False):
    """BASELINE.json configs[4]: the synthetic cohort sharded by (patient, slice) over the ranks, no data-path collective,
    ONE NCCL all-gather of the per-unit results at the end.  Results of a unit: the registered label maps of its 29 pairs
    (uint8, written by the mask-warp kernel straight into this rank's slice of the all-gather buffer: the usage pattern of
    SegFlowGaussian.py:3571-3580 `warp_linear` -> one label map per frame, collected per (patient, slice, frame) like
    compute_SSIM.py:26-104) plus a compact row (two float64 checksums, four class areas).  Returns the `cohort` object."""
    import cardioflow_b200 as cf
    from cardioflow_b200 import _cabi, shard
    lib = _cabi.lib()
    P, H, W, IH, IW = wl["pairs"], wl["H"], wl["W"], wl["IH"], wl["IW"]
    K81 = LEVELS * (2 * RADIUS + 1) ** 2
    units = shard.cohort_units(patients, slices)
    mine = shard.shard_units(units, rank, world)
    pyramid = torch.empty(lib.cf_corr_pyramid_bytes(P, H, W, LEVELS) // 4, dtype=torch.float32, device=dev)
    outs = torch.empty(ITERS, P, K81, H, W, dtype=torch.float32, device=dev)
    warped = torch.empty(P, 1, IH, IW, dtype=torch.float32, device=dev)
    # the pairs of a unit are independent: n parallel chains over contiguous sub-batches (cardioflow_b200.Lanes)
    bounds = cf.lane_bounds(P, max(1, args.lanes))
    lanes = cf.Lanes(len(bounds), dev)
    lane_pyr = ([torch.empty(lib.cf_corr_pyramid_bytes(b - a, H, W, LEVELS) // 4, dtype=torch.float32, device=dev)
                 for a, b in bounds] if len(bounds) > 1 else [pyramid])
    lane_counts = torch.zeros(len(bounds), 4, dtype=torch.int64, device=dev)
    sum_ws = torch.empty(lib.cf_sum_f64_workspace_bytes() // 8, dtype=torch.float64, device=dev)
    # result buffers: allocated as the RECEIVE buffers of the final all-gather; kernels write into this rank's slice
    maps = shard.GatherBuffer(len(units), (P, 1, IH, IW), torch.uint8, dev, rank=rank, world_size=world)
    sums = shard.GatherBuffer(len(units), (2,), torch.float64, dev, rank=rank, world_size=world)
    areas = shard.GatherBuffer(len(units), (len(bounds), 4), torch.int64, dev, rank=rank, world_size=world)
    lab_scratch = torch.empty(P, 1, IH, IW, dtype=torch.uint8, device=dev)
    sum_scratch = torch.zeros(2, dtype=torch.float64, device=dev)
    area_scratch = torch.zeros(len(bounds), 4, dtype=torch.int64, device=dev)

    def chain(u, j, a, b, lab_out, area_out):
        blk = cf.CorrBlock(u["f1"][a:b], u["f2"][a:b], num_levels=LEVELS, radius=RADIUS, precision=args.precision,
                           pyramid_buffer=lane_pyr[j])
        for k in range(ITERS):
            blk(u["coords"][k, a:b], out=outs[k, a:b])
        cf.warp(u["img"][a:b], u["flow"][a:b], out=warped[a:b])
        # registered label maps (uint8) + the lane's class areas, from ONE launch, straight into the result buffers
        cf.warp_onehot_argmax(u["flow"][a:b], u["label"][a:b], num_classes=4, out=lab_out[a:b], return_counts=True,
                              counts_out=area_out[j], counts_total=True)
        return blk

    def compute(u, lab_out, sum_out, area_out):
        lanes.run(P, lambda j, a, b: chain(u, j, a, b, lab_out, area_out))
        shard.sum_f64(outs[ITERS - 1], sum_out[0:1], sum_ws)     # one pass, fixed summation order (cf_sum_f64)
        shard.sum_f64(warped, sum_out[1:2], sum_ws)

    with torch.no_grad():
        # static input / output buffers + ONE CUDA graph of a unit's whole device work (replayed per unit); the results of a
        # replay land in fixed staging slices only because a graph bakes addresses in: one device copy per unit moves them
        # into the unit's slot (the eager path, --no-graph, writes the slots directly)
        static = make_unit_device(wl, shard.unit_seed(*units[0]), dev)
        g_lab = torch.empty(P, 1, IH, IW, dtype=torch.uint8, device=dev)
        g_sum = torch.zeros(2, dtype=torch.float64, device=dev)
        g_area = torch.zeros(len(bounds), 4, dtype=torch.int64, device=dev)
        compute(static, g_lab, g_sum, g_area)
        torch.cuda.synchronize()
        graph = None
        if not args.no_graph:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                compute(static, g_lab, g_sum, g_area)

        def run_unit(seed, lab_dst, sum_dst, area_dst):
            u = make_unit_device(wl, seed, dev)              # input synthesis: not part of the path, not timed
            for k, v in u.items():
                static[k].copy_(v)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if graph:
                graph.replay()
                lab_dst.copy_(g_lab)
                sum_dst.copy_(g_sum)
                area_dst.copy_(g_area)
            else:
                compute(static, lab_dst, sum_dst, area_dst)
            e1.record()
            return e0, e1

        for _ in range(max(args.warmup, 3)):
            run_unit(shard.unit_seed(*units[0]), lab_scratch, sum_scratch, area_scratch)
        if dist:                                             # communicator set-up is not part of the gather being timed
            warm = torch.zeros(world, 4, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(warm.view(-1), warm[rank].reshape(-1))
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        sampler = ClockSampler(dev.index)
        sampler.start()
        evs = [run_unit(shard.unit_seed(pat, sl), maps.local[i], sums.local[i], areas.local[i]) for i, (pat, sl) in enumerate(mine)]
        # The per-unit intervals exclude the (untimed) input synthesis between units, whose wall-clock cost differs from rank
        # to rank; without a barrier here that skew would be charged to the gather as waiting time.  Compute skew is already
        # in the max over ranks of compute_ms; the gather is timed from a common start.
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        t_maps, t_sums, t_areas = maps.gather(), sums.gather(), areas.gather()      # in place: send buffer = recv[rank]
        g1.record()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        compute_ms = sum(a.elapsed_time(b) for a, b in evs)
        gather_ms = g0.elapsed_time(g1)
        # the gathered results must equal what any rank computes for a unit on its own (deterministic kernels)
        check = []
        if rank == 0:
            for j in sorted({0, len(units) // 2, len(units) - 1, min(1, len(units) - 1)}):
                run_unit(shard.unit_seed(*units[j]), lab_scratch, sum_scratch, area_scratch)
                check.append(bool(torch.equal(lab_scratch, maps.unit(t_maps, j)) and torch.equal(sum_scratch, sums.unit(t_sums, j))
                                  and torch.equal(area_scratch, areas.unit(t_areas, j))
                                  and int(area_scratch.sum()) == P * IH * IW))
            torch.cuda.synchronize()
        # optional: the worst-case payload of SURVEY 8(e), the full-resolution flows (2 x 256^2 x 4 B per pair)
        flow_info = None
        if flow_gather and dist:
            fbuf = shard.GatherBuffer(len(units), (P, 2, IH, IW), torch.float32, dev, rank=rank, world_size=world)
            for i in range(len(mine)):
                fbuf.local[i].copy_(static["flow"])           # stand-in payload: the flows are the network's output, not this path's
            torch.cuda.synchronize()
            dist.barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record()
            fbuf.gather()
            f1.record()
            torch.cuda.synchronize()
            flow_info = [f0.elapsed_time(f1), fbuf.nbytes_sent]
            del fbuf
    t = torch.tensor([compute_ms, gather_ms, flow_info[0] if flow_info else 0.0], dtype=torch.float64, device=dev)
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    compute_ms, gather_ms, flow_ms = t.tolist()
    total_ms = compute_ms + gather_ms            # slowest rank's compute, then the gather from the common barrier
    n_pairs = len(units) * P
    payload = maps.nbytes_sent + sums.nbytes_sent + areas.nbytes_sent
    out = {
        "workload": "cfg5: " + WORKLOADS["cfg5"]["desc"], "scaling": "strong", "value": n_pairs / (total_ms * 1e-3),
        "unit": "frame-pairs/s", "patients": patients, "slices": slices, "units": len(units), "units_on_rank0": len(mine),
        "pairs_per_unit": P, "lanes": len(bounds), "corr_precision": args.precision,
        "compute_ms_max_rank": compute_ms, "gather_ms_max_rank": gather_ms, "clocks": clocks,
        "launch": "eager launches from Python, one unit at a time" if args.no_graph else
                  "one CUDA graph replay per unit (build + 12 lookups + warp + uint8 mask warp with class areas + 2 checksums)",
        "timing": "value = pairs / (max over ranks of the summed per-unit device intervals + max over ranks of the gather timed from a barrier)",
        "gather": {"collective": "all_gather_into_tensor (NCCL, in place: send buffer = this rank's slice of the receive buffer)"
                                 if world > 1 else "none (1 rank)",
                   "payload": "registered label maps uint8 [units, 29, 1, 256, 256] + 2 float64 checksums + class areas per unit",
                   "bytes_sent_per_rank": int(payload), "bytes_total": int(payload * world),
                   "gbytes_per_s_per_rank_received": (payload * (world - 1) / (gather_ms * 1e-3) / 1e9) if world > 1 and gather_ms > 0 else None,
                   "units_verified_on_rank0": check},
        "gpu_launches": len(mine) * ((2 + ITERS + 2) * len(bounds) + 4 + (3 if graph else 0)),
    }
    if flow_info:
        out["flow_gather_variant"] = {"what": "SURVEY 8(e) worst case: full-resolution flows fp32 [units, 29, 2, 256, 256] through the same in-place all-gather",
                                      "bytes_sent_per_rank": int(flow_info[1]), "bytes_total": int(flow_info[1] * world), "ms_max_rank": flow_ms,
                                      "gbytes_per_s_per_rank_received": flow_info[1] * (world - 1) / (flow_ms * 1e-3) / 1e9 if flow_ms > 0 else None}
    if rank == 0 and not all(check):
        out["error"] = "gathered results differ from a local recomputation"
    return out


def run_cohort(args, wl):
    """`--workload cfg5`: the cohort alone (the default run attaches the same object as `cohort` when N > 1)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        sys.exit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    co = cohort_measure(args, wl, dev, rank, world, dist, args.patients, args.slices, flow_gather=args.flow_gather)
    if rank == 0:
        line = {"metric": METRIC, "value": co["value"], "unit": "frame-pairs/s", "n_gpus": world, "steps": co["units"],
                "warmup": max(args.warmup, 3), "ms_per_step": (co["compute_ms_max_rank"] + co["gather_ms_max_rank"]) / max(co["units_on_rank0"], 1),
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "bf16 GEMM (fp32 accumulate) / f32 lookup+warp" if args.precision == "bf16" else "f32-grade GEMM / f32 lookup+warp",
                "data": "synthetic",
                "config": {k: co[k] for k in ("workload", "patients", "slices", "units", "units_on_rank0", "pairs_per_unit", "lanes", "launch", "corr_precision")},
                "clocks": co["clocks"], "compute_ms_max_rank": co["compute_ms_max_rank"], "gather_ms_max_rank": co["gather_ms_max_rank"],
                "timing": co["timing"], "gather": co["gather"], "gpu_launches": co["gpu_launches"]}
        line["config"]["l2"] = "every unit has fresh inputs and writes 623 MB (> 126 MB L2)"
        if "flow_gather_variant" in co:
            line["flow_gather_variant"] = co["flow_gather_variant"]
        if "error" in co:
            line["error"] = co["error"]
        emit(line)
    if dist:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ cfg3 warp batch
def run_warp_batch(args, wl):
    """BASELINE.json configs[2]: SpatialTransformer warp of image + one-hot 4-class masks (5 channels) by a dense smooth
    flow, 10 slices x 30 frames at 224^2 in one batch.  A step = one cf_warp2d launch over the 300 items."""
    import cardioflow_b200 as cf
    from cardioflow_b200 import _cabi
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        sys.exit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    _cabi.lib()
    B, IH, IW, C = wl["items"], wl["IH"], wl["IW"], 5

    def make(seed, pin=False):
        g = torch.Generator().manual_seed(seed)
        img = torch.rand(B, 1, IH, IW, generator=g)
        yy, xx = torch.meshgrid(torch.arange(IH), torch.arange(IW), indexing="ij")
        rad = ((yy - IH / 2) ** 2 + (xx - IW / 2) ** 2).float().sqrt() / (IH / 2)
        label = (rad < 0.2).long() * 3 + ((rad >= 0.2) & (rad < 0.35)).long() * 2 + ((rad >= 0.35) & (rad < 0.6)).long()
        onehot = torch.nn.functional.one_hot(label, 4).permute(2, 0, 1).float()[None].expand(B, 4, IH, IW)
        src = torch.cat([img, onehot], 1).contiguous()
        flow = torch.nn.functional.interpolate(torch.randn(B, 2, 7, 7, generator=g) * 4.0, size=(IH, IW), mode="bicubic",
                                               align_corners=True).contiguous()
        if pin:
            return src.pin_memory(), flow.pin_memory()
        return src.to(dev), flow.to(dev)

    n_sets = 2       # 2 x (361 + 120) MB of inputs + 361 MB written per step: every step streams > 126 MB L2
    sets = [make(1003 + 31 * rank + s) for s in range(n_sets)]
    out = torch.empty(B, C, IH, IW, dtype=torch.float32, device=dev)
    st = cf.SpatialTransformer((IH, IW)).to(dev)
    with torch.no_grad():
        for i in range(max(args.warmup, 3)):
            cf.warp(sets[i % n_sets][0], sets[i % n_sets][1], out=out)
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            cf.warp(sets[i % n_sets][0], sets[i % n_sets][1], out=out)
        e1.record()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        if dist:
            dist.barrier()
        ms_total = e0.elapsed_time(e1)
        # end to end: pinned host buffers -> module API -> pinned host result
        h_src, h_flow = make(1003 + 31 * rank, pin=True)
        h_out = torch.empty(B, C, IH, IW, dtype=torch.float32).pin_memory()
        d_src, d_flow = torch.empty_like(sets[0][0]), torch.empty_like(sets[0][1])
        e2e_steps = args.e2e_steps or min(args.steps, 10)

        def e2e_step():
            d_src.copy_(h_src, non_blocking=True)
            d_flow.copy_(h_flow, non_blocking=True)
            h_out.copy_(st(flow=d_flow, original=d_src), non_blocking=True)

        e2e_step()
        torch.cuda.synchronize()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(e2e_steps):
            e2e_step()
        b_.record()
        torch.cuda.synchronize()
        e2e_ms = a.elapsed_time(b_)
    if dist:
        t = torch.tensor([ms_total, e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, e2e_ms = t.tolist()
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak = json.load(open(peaks_path))["hbm_gbs"] if os.path.exists(peaks_path) else 6650.0
        nbytes = 4 * B * IH * IW * (2 * C + 2)
        ms = ms_total / args.steps
        line = {
            "metric": "frames warped/sec (image + 4 one-hot masks, 224^2)", "value": world * B * args.steps / (ms_total * 1e-3),
            "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg3: " + wl["desc"], "items_per_step_per_gpu": B, "channels": C,
                       "l2": f"inputs rotate over {n_sets} sets; a step reads 482 MB and writes 301 MB (> 126 MB L2)"},
            "clocks": clocks,
            "e2e": {"value": world * B * e2e_steps / (e2e_ms * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": int(h_src.numel() * 4 + h_flow.numel() * 4), "d2h_bytes_per_step": int(h_out.numel() * 4),
                    "steps": e2e_steps},
            "gpu_launches": args.steps,
            "roofline": {"kernel": "warp2d_tile_kernel", "bound": "hbm", "achieved": nbytes / (ms * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": nbytes / (ms * 1e-3) / 1e9 / peak, "traffic": None,
                         "algorithmic_bytes_per_launch": nbytes, "avg_launch_ms": ms},
        }
        emit(line)
    if dist:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "reference-gpu"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="bf16", choices=["auto", "bf16", "bf16x3", "fp32"],
                    help="correlation build.  The library default (auto) is fp32-grade like the reference (bf16x3 split on tcgen05); the "
                         "bench headline uses plain bf16 operands, the tolerance north_star states for the tensor-core correlation "
                         "(<= 1e-2 relative, fp32 accumulate); both are reported in profiles/")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from Python instead of replaying one CUDA graph per step")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer leg (default min(steps, 10))")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0, help="--impl reference: host threads (default: all; 1 = per-core figure)")
    ap.add_argument("--l2-persist-mb", type=int, default=0,
                    help="experiment: cudaLimitPersistingL2CacheSize in MB (the lookup's pyramid reads carry L2 evict_last hints)")
    ap.add_argument("--lanes", type=int, default=2,
                    help="independent sub-batches of the step's pairs run as parallel stream branches (cardioflow_b200.Lanes); "
                         "1 = one chain over the whole batch")
    ap.add_argument("--patients", type=int, default=150, help="cfg5 only")
    ap.add_argument("--slices", type=int, default=10, help="cfg5 only")
    ap.add_argument("--flow-gather", action="store_true", help="cfg5 / cohort: also time the all-gather of the full-resolution flows (22.8 GB in all)")
    ap.add_argument("--cohort-patients", type=int, default=150,
                    help="N > 1: size of the cfg5 cohort run after the main measurement and reported as `cohort` (0 = skip)")
    ap.add_argument("--on-the-fly", action="store_true",
                    help="cfg2 / cfg4: CorrBlock(on_the_fly=True): no materialised volume, window correlations recomputed per lookup on tcgen05")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        return run_reference_arm(args, wl)
    if args.impl == "reference-gpu":
        if args.workload not in ("cfg1", "cfg2", "cfg4"):
            sys.exit("--impl reference-gpu covers cfg1, cfg2 and cfg4")
        return run_reference_gpu(args, wl)
    if args.workload == "cfg5":
        return run_cohort(args, wl)
    if args.workload == "cfg3":
        return run_warp_batch(args, wl)

    import cardioflow_b200 as cf
    from cardioflow_b200 import _cabi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        sys.exit("bench.py needs a CUDA device: the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    _cabi.lib()  # fail loudly if the extension is missing
    affinity = pin_to_gpu_numa(local_rank)      # before the first pinned allocation
    if args.on_the_fly:
        if args.workload not in ("cfg2", "cfg4"):
            sys.exit("--on-the-fly applies to cfg2 / cfg4")
        args.lanes = 1                          # the on-the-fly kernel is persistent over the whole batch
    if args.l2_persist_mb:
        import ctypes
        rt = ctypes.CDLL("libcudart.so.12")
        torch.zeros(1, device=dev)
        rc = rt.cudaDeviceSetLimit(6, ctypes.c_size_t(args.l2_persist_mb << 20))      # cudaLimitPersistingL2CacheSize
        if rc:
            sys.exit(f"cudaDeviceSetLimit(PersistingL2CacheSize, {args.l2_persist_mb} MB) failed: {rc}")

    P, C, H, W, IH, IW = wl["pairs"], wl["C"], wl["H"], wl["W"], wl["IH"], wl["IW"]
    K81 = LEVELS * (2 * RADIUS + 1) ** 2
    n_sets = 3
    sets = [make_unit(wl, seed=1002 + 17 * rank + s, device=dev) for s in range(n_sets)]
    lib = _cabi.lib()
    pyr_floats = lib.cf_corr_pyramid_bytes(P, H, W, LEVELS) // 4
    pyramid = torch.empty(pyr_floats, dtype=torch.float32, device=dev)
    outs = torch.empty(ITERS, P, K81, H, W, dtype=torch.float32, device=dev)
    warped = torch.empty(P, 1, IH, IW, dtype=torch.float32, device=dev)
    st = cf.SpatialTransformer((IH, IW)).to(dev)
    ev_l0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev_l1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]

    def step(u, ev=None):
        if args.on_the_fly:
            blk = cf.CorrBlock(u["f1"], u["f2"], num_levels=LEVELS, radius=RADIUS, on_the_fly=True,
                               precision="fp32" if args.precision == "fp32" else "bf16")
        else:
            blk = cf.CorrBlock(u["f1"], u["f2"], num_levels=LEVELS, radius=RADIUS, precision=args.precision,
                               pyramid_buffer=pyramid)
        if ev:
            ev[0].record()
        for k in range(ITERS):
            blk(u["coords"][k], out=outs[k])
        if ev:
            ev[1].record()
        cf.warp(u["img"], u["flow"], out=warped)
        return blk

    def lookups_only(u, blk):
        for k in range(ITERS):
            blk(u["coords"][k], out=outs[k])

    # Lanes: the pairs of a step are independent, so the step may run as n parallel chains (build -> 12 lookups -> warp
    # each) over contiguous sub-batches; same kernels, same results, ramps and drains of one chain filled by the other.
    bounds = cf.lane_bounds(P, max(1, args.lanes))
    n_lanes = len(bounds)
    lanes = cf.Lanes(n_lanes, dev)
    lane_pyr = ([torch.empty(lib.cf_corr_pyramid_bytes(b - a, H, W, LEVELS) // 4, dtype=torch.float32, device=dev)
                 for a, b in bounds] if n_lanes > 1 else [pyramid])

    def lane_chain(u, j, a, b):
        bk = cf.CorrBlock(u["f1"][a:b], u["f2"][a:b], num_levels=LEVELS, radius=RADIUS, precision=args.precision,
                          pyramid_buffer=lane_pyr[j])
        for k in range(ITERS):
            bk(u["coords"][k, a:b], out=outs[k, a:b])
        cf.warp(u["img"][a:b], u["flow"][a:b], out=warped[a:b])
        return bk

    def step_lanes(u):
        return lanes.run(P, lambda j, a, b: lane_chain(u, j, a, b))

    with torch.no_grad():
        blk = step(sets[0])
        used_precision = blk.used_precision
        torch.cuda.synchronize()
        use_graph = not args.no_graph
        graphs = lookup_graphs = full_graphs = None
        lane_lookup_graphs = []
        if n_lanes > 1:
            step_lanes(sets[0])
            torch.cuda.synchronize()
        if use_graph:
            # the step is launch-bound from Python (15 launches of 5-60 us): replay it as ONE CUDA graph per input set
            graphs, lookup_graphs, full_graphs, blks, keep = [], [], [], [], []
            for u in sets:
                gph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gph):
                    blks.append(step(u))
                full_graphs.append(gph)
            if n_lanes > 1:
                for u in sets:
                    gph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gph):
                        keep.append(step_lanes(u))
                    graphs.append(gph)
            else:
                graphs = full_graphs
            for u, bk in zip(sets, blks):          # the 12 lookups alone, for the roofline of the dominant kernel
                gph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gph):
                    lookups_only(u, bk)
                lookup_graphs.append(gph)
            if n_lanes > 1:                        # ... and as the step runs them: one 12-lookup chain per lane, concurrently
                for u, lane_blks in zip(sets, keep):
                    gph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gph):
                        lanes.run(P, lambda j, a, b, u=u, lb=lane_blks: [lb[j](u["coords"][k, a:b], out=outs[k, a:b]) for k in range(ITERS)])
                    lane_lookup_graphs.append(gph)

        def run_step(i):
            if graphs:
                graphs[i % n_sets].replay()
            elif n_lanes > 1:
                step_lanes(sets[i % n_sets])
            else:
                step(sets[i % n_sets])

        for i in range(args.warmup):
            run_step(i)
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t_begin.record()
        for i in range(args.steps):
            run_step(i)
        t_end.record()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        if dist:
            dist.barrier()
        ms_total = t_begin.elapsed_time(t_end)
        # Roofline of the dominant kernel: the lookup launched over the WHOLE batch, alone on its stream (lane branches
        # overlap, so per-kernel durations inside a lane step are not separable).  Graph replays carry no inner events:
        # time the 12-lookup graph right after a full single-chain step of the same set (same cache state as inside a
        # step); eager: events around the 12 lookups of a single-chain step.  CUDA events on the launching stream.
        n_ev = min(args.steps, 30)
        for i in range(n_ev):
            if full_graphs:
                full_graphs[i % n_sets].replay()
                ev_l0[i].record()
                lookup_graphs[i % n_sets].replay()
                ev_l1[i].record()
            else:
                step(sets[i % n_sets], (ev_l0[i], ev_l1[i]))
        torch.cuda.synchronize()
        lookup_ms = sorted(ev_l0[i].elapsed_time(ev_l1[i]) for i in range(n_ev))[n_ev // 2] / ITERS
        # the same 12 lookups in the configuration of the step: each lane's chain on its own stream, all lanes at once
        lane_lookup_ms = None
        if lane_lookup_graphs:
            evs2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_ev)]
            for i in range(n_ev):
                graphs[i % n_sets].replay()
                evs2[i][0].record()
                lane_lookup_graphs[i % n_sets].replay()
                evs2[i][1].record()
            torch.cuda.synchronize()
            lane_lookup_ms = sorted(a.elapsed_time(b) for a, b in evs2)[n_ev // 2] / ITERS

        # ---- end-to-end leg: pinned host buffers -> public module API -> pinned host results ----
        e2e_steps = args.e2e_steps or min(args.steps, 10)
        host = make_unit(wl, seed=1002 + 17 * rank, pin=True)
        h_outs = torch.empty(ITERS, P, K81, H, W, dtype=torch.float32).pin_memory()
        h_warp = torch.empty(P, 1, IH, IW, dtype=torch.float32).pin_memory()
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        d2h = h_outs.numel() * 4 + h_warp.numel() * 4

        # Two-deep pipeline: the H2D copies of step i+1 run on a copy stream (PCIe is full duplex) while step i
        # computes and streams its results back; every step still moves all of its inputs and outputs.
        copy_stream = torch.cuda.Stream(device=dev)
        main_stream = torch.cuda.current_stream(dev)
        dbuf = [{k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in host.items()} for _ in range(2)]
        in_ready = [torch.cuda.Event() for _ in range(2)]
        in_free = [torch.cuda.Event() for _ in range(2)]

        def upload(i):
            s = i % 2
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(in_free[s])           # the step that last read this buffer set is done
                for k, v in host.items():
                    dbuf[s][k].copy_(v, non_blocking=True)
                in_ready[s].record(copy_stream)

        def e2e_step(i, prefetch=True):
            s = i % 2
            if prefetch:
                upload(i + 1)
            main_stream.wait_event(in_ready[s])
            d = dbuf[s]
            if args.on_the_fly:
                b = cf.CorrBlock(d["f1"], d["f2"], num_levels=LEVELS, radius=RADIUS, on_the_fly=True,
                                 precision="fp32" if args.precision == "fp32" else "bf16")
            else:
                b = cf.CorrBlock(d["f1"], d["f2"], num_levels=LEVELS, radius=RADIUS, precision=args.precision,
                                 pyramid_buffer=pyramid)
            for k in range(ITERS):
                b(d["coords"][k], out=outs[k])
                h_outs[k].copy_(outs[k], non_blocking=True)
            w = st(flow=d["flow"], original=d["img"])
            in_free[s].record(main_stream)
            h_warp.copy_(w, non_blocking=True)

        for ev in in_free:
            ev.record(main_stream)
        upload(0)
        e2e_step(0)
        e2e_step(1, prefetch=False)
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        # the timed region contains exactly e2e_steps uploads and e2e_steps downloads
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        upload(2)
        for i in range(2, 2 + e2e_steps):
            e2e_step(i, prefetch=(i + 1 < 2 + e2e_steps))
        e1.record()
        copy_stream.synchronize()
        torch.cuda.synchronize()
        e2e_ms = e0.elapsed_time(e1)

    if dist:
        t = torch.tensor([ms_total, e2e_ms, lookup_ms, lane_lookup_ms or 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, e2e_ms, lookup_ms, lane_lookup_ms = t.tolist()
        lane_lookup_ms = lane_lookup_ms or None
    value = world * P * args.steps / (ms_total * 1e-3)
    e2e_value = world * P * e2e_steps / (e2e_ms * 1e-3)

    # ---- roofline of the dominant kernel (lookup: 12 of the step's launches, ~80 % of its bytes) ----
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        with open(peaks_path) as f:
            peak, peak_src = json.load(f)["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (copy, read+write)"
    else:
        peak, peak_src = 6650.0, "fallback from B200_PROFILING.md"
    lookup_bytes = lookup_algorithmic_bytes(P, H, W, LEVELS, RADIUS)
    achieved = lookup_bytes / (lookup_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "lookup_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(args.workload)
    # fp32: SIMT GEMM + a pool launch per coarser level; bf16: transpose pre-pass + tcgen05 GEMM whose epilogue writes
    # levels 0-1 (all four levels when the fmap width is 32) + a pool launch per remaining level; 12 lookups; 1 warp
    if used_precision == "fp32":
        build_launches = 1 + (LEVELS - 1)
    else:
        # version 2 of the build (H % 4 == 0, W % 8 == 0): pre-pass + persistent GEMM (levels 0-2) + one pool launch per further level
        # plain bf16 at W = 32: version 1 (pre-pass + GEMM writing all four levels); else version 2
        v2 = H % 4 == 0 and W % 8 == 0 and not (used_precision == "bf16" and W == 32)
        build_launches = (2 + max(0, LEVELS - 3)) if v2 else 2 + (0 if (W == 32 and LEVELS <= 4) else LEVELS - 2)
    launches_per_step = (build_launches + ITERS + 1) * n_lanes
    if args.on_the_fly:                          # pooled fmap2 pyramid (copy + 3 pools) + 5 transposes, 12 lookups, 1 warp
        launches_per_step = (1 + (LEVELS - 1) + 1 + LEVELS) + ITERS + 1

    line = {
        "metric": METRIC, "value": value, "unit": "frame-pairs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": {"fp32": "f32", "bf16x3": "f32-grade: bf16x3 split GEMM on tcgen05 (fp32 accumulate) / f32 lookup+warp"}.get(
            used_precision, "bf16 GEMM (fp32 accumulate) / f32 lookup+warp"),
        "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "pairs_per_step_per_gpu": P,
                   "corr_precision": used_precision, "cuda_graph": bool(graphs),
                   "lanes": n_lanes, "lane_pairs": [b - a for a, b in bounds],
                   "launch": ("one CUDA graph replay per step" if graphs else "eager launches from Python")
                             + (f", {n_lanes} parallel stream branches over contiguous sub-batches of the pairs" if n_lanes > 1 else ""),
                   "l2": (lambda mb: f"inputs rotate over {n_sets} sets and each step writes {mb:.0f} MB (> 126 MB L2)" if mb > 126 else
                          f"NOT flushed: a step writes only {mb:.0f} MB, its working set stays in the 126 MB L2 (single-pair workload; "
                          "the step is launch-latency-bound, bandwidth figures are L2-assisted)")((pyr_floats * 4 + outs.numel() * 4) / 1e6)},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "frame-pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"kernel": "lookup_kernel<4>", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": lookup_bytes, "avg_launch_ms": lookup_ms,
                     "launch_timed": f"all {P} pairs in one launch, alone on its stream",
                     "frac_of_nominal_8000": achieved / 8000.0},
    }
    line["config"]["host_affinity"] = affinity
    if traffic:
        # informational: what the launch moves through DRAM (ncu, steady state: profiles/lookup_traffic.json) over the live launch time --
        # how close the kernel is to the measured HBM peak on the bytes it really moves; `frac` stays on the algorithmic bytes
        line["roofline"]["dram_traffic_over_launch_time"] = {"gbytes_per_s": traffic / (lookup_ms * 1e-3) / 1e9,
                                                             "frac_of_peak": traffic / (lookup_ms * 1e-3) / 1e9 / peak,
                                                             "traffic_over_algorithmic": traffic / lookup_bytes}
    if lane_lookup_ms:
        # informational: the same kernel as the step runs it (one chain per lane, lanes concurrent): bytes of the whole batch over the
        # elapsed time per lookup iteration.  `frac` above stays the conservative single-launch figure.
        line["roofline"]["as_run_in_the_step"] = {
            "what": f"{n_lanes} lanes x 12 lookups as parallel stream branches (the step's configuration), elapsed / 12",
            "ms_per_lookup_iteration": lane_lookup_ms, "achieved": lookup_bytes / (lane_lookup_ms * 1e-3) / 1e9,
            "frac": lookup_bytes / (lane_lookup_ms * 1e-3) / 1e9 / peak}
    if args.on_the_fly:
        # K4: the dominant kernel is the tensor-core lookup; useful work = SURVEY 8(d) K4 (window products a per-query kernel needs;
        # the bounding-box GEMMs execute several times that)
        flops, hh, ww = 0.0, H, W
        for _ in range(LEVELS):
            flops += 2.0 * C * H * W * min(100, hh * ww)
            hh, ww = hh // 2, ww // 2
        flops *= P
        tpeak = 1365.0
        if os.path.exists(peaks_path):
            with open(peaks_path) as f:
                tpeak = json.load(f).get("bf16_tflops_sustained", tpeak)
        op_bytes = lib.cf_corr_otf_operand_bytes(P, C, H, W, LEVELS)
        line["config"]["on_the_fly"] = {"operand_bytes": int(op_bytes), "materialised_pyramid_bytes": int(pyr_floats * 4),
                                        "variant": "tcgen05 (bf16 operands, fp32 accumulate)" if used_precision == "bf16" else "CUDA cores fp32"}
        line["roofline"] = {"kernel": "lookup_tc_kernel<4>", "bound": "tensor", "achieved": flops / (lookup_ms * 1e-3) / 1e12, "peak": tpeak,
                            "unit": "TFLOP/s", "frac": flops / (lookup_ms * 1e-3) / 1e12 / tpeak, "traffic": None,
                            "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained",
                            "useful_flops_per_launch": flops, "avg_launch_ms": lookup_ms,
                            "note": "useful FLOPs = sum_l 2*C*HW*min(100, h_l*w_l) per pair (SURVEY 8(d) K4); the kernel is bound by L2 operand "
                                    "streaming and its epilogue, not by the tensor pipe (DESIGN section 8)",
                            "output_bytes_per_launch": 4 * P * K81 * H * W,
                            "output_gbytes_per_s": 4 * P * K81 * H * W / (lookup_ms * 1e-3) / 1e9}
    # ---- N > 1: the path's real multi-GPU design (cfg5 cohort, strong scaling, one NCCL all-gather of the label maps) ----
    if world > 1 and args.cohort_patients > 0 and not args.on_the_fly:
        line_cohort = cohort_measure(args, WORKLOADS["cfg5"], dev, rank, world, dist, args.cohort_patients, args.slices,
                                     flow_gather=True)
        line["cohort"] = line_cohort
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = 1 if args.workload == "cfg4" else 4
        v, cores, reps = cpu_pairs_per_sec(wl, sample)
        line["cpu_baseline"] = {"value": v, "unit": "frame-pairs/s", "cores": cores, "kind": "port",
                                "sample": f"{sample} of {P} pairs (per-pair CorrBlock + 12 lookups + warp), "
                                          f"torch CPU fp32 oracle, best of {reps}"}
    if rank == 0:
        emit(line)
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
