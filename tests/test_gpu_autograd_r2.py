"""GPU parity of the autograd holes closed in round 2 (VERDICT Missing #3, Weak #3), through the C ABI.

* SpatialTransformerContour backward: the reference back-propagates an MSE point loss through it into `original` = the
  predicted flow (/root/reference/nnunet/training/network_training/SegFlowGaussian.py:1497-1503).  Oracle: torch-CPU
  autograd of oracle/warp_oracle.SpatialTransformerContour (ATen grid_sampler_2d_backward).
* nearest-mode SpatialTransformer backward: ATen semantics (gradient to the nearest texel, zero flow gradient).
* the 12 lookups of one CorrBlock share ONE gradient pyramid; repeated backward passes and gradients taken w.r.t. the
  pyramid itself must still be right.
* feature maps with H*W < 32 (ADVICE: the (b, p) split of a 32-query CTA wraps more than once).
Tolerance norm: max|a-b| / max|b| (conftest.rel_err).
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import c_oracle, corr_oracle, warp_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cf():
    import cardioflow_b200 as cf
    return cf


@pytest.mark.parametrize("mode", ["bilinear", "nearest"])
@pytest.mark.parametrize("size,chan,npts", [((40, 56), 2, 77), ((32, 32), 3, 5), ((192, 192), 2, 300)])
def test_contour_backward_matches_aten(mode, size, chan, npts):
    cf = _cf()
    g = torch.Generator().manual_seed(41)
    H, W = size
    field = torch.randn(2, chan, H, W, generator=g)
    locs = torch.stack([torch.rand(2, 1, npts, generator=g) * (W + 4) - 2, torch.rand(2, 1, npts, generator=g) * (H + 4) - 2], dim=1)
    wgt = torch.randn(2, chan, 1, npts, generator=g)
    fr = field.clone().requires_grad_()
    lr = locs.clone().requires_grad_()
    # the reference normalises its argument in place; autograd needs a non-leaf for that, as in the trainer (torch.clone(...))
    want = warp_oracle.SpatialTransformerContour(size, mode)(lr.clone(), fr)
    (want * wgt).sum().backward()

    fd = field.to(DEV).requires_grad_()
    ld = locs.to(DEV).requires_grad_()
    mod = cf.SpatialTransformerContour(size, mode).to(DEV)
    arg = ld.clone()
    got = mod(arg, fd)
    assert got.grad_fn is not None, "the point loss must back-propagate through the drop-in"
    assert rel_err(got.detach().cpu().numpy(), want.detach().numpy()) <= 1e-6
    assert float(arg.detach().abs().max()) <= 1.2          # normalised in place, like the reference
    (got * wgt.to(DEV)).sum().backward()
    assert rel_err(fd.grad.cpu().numpy(), fr.grad.numpy()) <= 1e-5      # atomics: summation order only
    if mode == "bilinear":
        assert rel_err(ld.grad.cpu().numpy(), lr.grad.numpy()) <= 1e-4
    else:
        assert not ld.grad.any() and not lr.grad.any()


def test_point_loss_pattern_of_the_trainer():
    """delta = transformer(torch.clone(points), flow.repeat(2,1,1,1)); loss = mse(points + flip(delta), target); grad -> flow."""
    cf = _cf()
    g = torch.Generator().manual_seed(42)
    flow = (torch.randn(1, 2, 48, 48, generator=g) * 2)
    pts = torch.rand(2, 2, 1, 30, generator=g) * 47
    target = pts + torch.randn(2, 2, 1, 30, generator=g)

    def loss_of(mod, fl, p, t):
        delta = mod(torch.clone(p), fl.repeat(2, 1, 1, 1))
        return torch.nn.functional.mse_loss(p + torch.flip(delta, dims=[1]), t)

    fr = flow.clone().requires_grad_()
    loss_of(warp_oracle.SpatialTransformerContour((48, 48)), fr, pts, target).backward()
    fd = flow.to(DEV).requires_grad_()
    loss = loss_of(cf.SpatialTransformerContour((48, 48)).to(DEV), fd, pts.to(DEV), target.to(DEV))
    loss.backward()
    assert fd.grad is not None and float(fd.grad.abs().max()) > 0
    assert rel_err(fd.grad.cpu().numpy(), fr.grad.numpy()) <= 1e-5


@pytest.mark.parametrize("size,chan", [((24, 24), 3), ((33, 47), 1)])
def test_nearest_warp_backward_matches_aten(size, chan):
    cf = _cf()
    g = torch.Generator().manual_seed(43)
    flow = torch.randn(2, 2, *size, generator=g) * 3 + 0.13      # away from exact .5 ties
    src = torch.rand(2, chan, *size, generator=g)
    wgt = torch.randn(2, chan, *size, generator=g)
    fr, sr = flow.clone().requires_grad_(), src.clone().requires_grad_()
    (warp_oracle.SpatialTransformer(size, "nearest")(fr, sr) * wgt).sum().backward()
    fd, sd = flow.to(DEV).requires_grad_(), src.to(DEV).requires_grad_()
    (cf.SpatialTransformer(size, "nearest").to(DEV)(fd, sd) * wgt.to(DEV)).sum().backward()
    assert rel_err(sd.grad.cpu().numpy(), sr.grad.numpy()) <= 1e-5
    assert not fd.grad.any() and not fr.grad.any()               # zero, not an exception (ATen returns zeros)


def test_lookup_gradient_buffer_is_shared_and_survives_repeated_passes():
    """12 lookups -> one zeroed gradient pyramid (not 12); a second backward pass and a gradient taken w.r.t. the pyramid
    itself (the build's backward never runs, the shared buffer is left behind) must not corrupt later passes."""
    cf = _cf()
    g = torch.Generator().manual_seed(44)
    H = W = 16
    f1, f2 = torch.randn(2, 32, H, W, generator=g), torch.randn(2, 32, H, W, generator=g)
    coords = [corr_oracle.coords_grid(2, H, W) + torch.randn(2, 2, H, W, generator=g) * (0.5 * k) for k in range(12)]
    wgts = [torch.randn(2, 324, H, W, generator=g) for _ in coords]
    r1, r2 = f1.clone().requires_grad_(), f2.clone().requires_grad_()
    ref = corr_oracle.CorrBlock(r1, r2)
    sum((ref(c) * w).sum() for c, w in zip(coords, wgts)).backward()

    d1, d2 = f1.to(DEV).requires_grad_(), f2.to(DEV).requires_grad_()
    blk = cf.CorrBlock(d1, d2, radius=4, precision="fp32")
    loss = sum((blk(c.to(DEV)) * w.to(DEV)).sum() for c, w in zip(coords, wgts))
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    (gp,) = torch.autograd.grad(loss, blk._pyramid, retain_graph=True)      # leaves the shared buffer in the block's state
    assert gp.shape == blk._pyramid.shape
    pyr_bytes = blk._pyramid.numel() * 4
    assert torch.cuda.max_memory_allocated() - base < 3.5 * pyr_bytes, "lookup backward must not allocate a pyramid per call"
    loss.backward(retain_graph=True)
    g1a, g2a = d1.grad.clone(), d2.grad.clone()
    assert rel_err(g1a.cpu().numpy(), r1.grad.numpy()) <= 2e-5 and rel_err(g2a.cpu().numpy(), r2.grad.numpy()) <= 2e-5
    d1.grad = d2.grad = None
    loss.backward()
    assert torch.equal(d1.grad, g1a) and torch.equal(d2.grad, g2a)          # deterministic, nothing stale


@pytest.mark.parametrize("hw,levels", [((4, 5), 2), ((5, 5), 2), ((2, 3), 1), ((4, 4), 1), ((3, 9), 1)])
def test_tiny_maps_below_32_queries_per_pair(hw, levels):
    """H*W < 32: a 32-query CTA spans several batch items (ADVICE: 'at most one wrap' was wrong for L <= 2)."""
    cf = _cf()
    H, W = hw
    B = 7
    g = torch.Generator().manual_seed(45)
    lv, h, w = [], H, W
    for _ in range(levels):
        lv.append(torch.randn(B * H * W, 1, h, w, generator=g))
        h, w = h // 2, w // 2
    coords = corr_oracle.coords_grid(B, H, W) + torch.randn(B, 2, H, W, generator=g) * 1.5
    for radius in (1, 4):
        want = c_oracle.lookup([t.numpy() for t in lv], coords.numpy(), radius=radius)
        blk = cf.CorrBlock.from_pyramid([t.to(DEV) for t in lv], radius=radius)
        guard = torch.full((B * levels * (2 * radius + 1) ** 2 * H * W + 64,), 777.0, device=DEV)
        out = guard[32:32 + B * levels * (2 * radius + 1) ** 2 * H * W].view(B, -1, H, W)
        blk(coords.to(DEV), out=out)
        assert np.array_equal(out.cpu().numpy(), want)
        assert bool((guard[:32] == 777.0).all()) and bool((guard[-32:] == 777.0).all())      # nothing written outside `out`
