"""GPU parity of the on-the-fly lookup on the tensor cores (csrc/lookup_tc.cu; K4 of SURVEY.md 8(d), BASELINE.json configs[3]
"materialised vs on-the-fly"), through the C ABI (cf_corr_otf_prepare / cf_corr_lookup_otf).

Contract: same output as corr_fn(coords1), /root/reference/nnunet/network_architecture/SegFlowGaussian.py:929-935.
Oracles (torch CPU): (a) oracle/corr_oracle.CorrBlock on the fp32 maps -- tolerance 1e-2 (bf16 operands; north_star);
(b) the same lookup on a pyramid built from the bf16-ROUNDED operands the kernel actually multiplies (f1 rounded, each pooled
fmap2 level rounded after fp32 pooling): products of bf16 values are exact in fp32, so only the accumulation order differs --
tolerance 2e-5.  (b) pins the indexing (chunks, bounding boxes, tap ownership) far below the bf16 noise of (a).
Norm: max|a-b| / max|b| (conftest.rel_err).
"""
import math

import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import corr_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cf():
    import cardioflow_b200 as cf
    return cf


def rounded_operand_oracle(f1, f2, levels, radius):
    B, C, H, W = f1.shape
    a = f1.bfloat16().float().flatten(2)
    pyr, cur = [], f2
    for _ in range(levels):
        h, w = cur.shape[-2:]
        bq = cur.bfloat16().float().flatten(2)
        corr = torch.einsum("bcp,bcq->bpq", a.double(), bq.double()).float() / math.sqrt(C)
        pyr.append(corr.reshape(B * H * W, 1, h, w))
        cur = torch.nn.functional.avg_pool2d(cur, 2, stride=2)
    blk = corr_oracle.CorrBlock.__new__(corr_oracle.CorrBlock)
    blk.num_levels, blk.radius, blk.corr_pyramid = levels, radius, pyr
    return blk


@pytest.mark.parametrize("hw,c,radius,levels,b", [((32, 32), 256, 4, 4, 2), ((64, 64), 64, 4, 4, 1), ((16, 16), 32, 4, 4, 3),
                                                 ((24, 48), 40, 3, 3, 2), ((8, 16), 8, 1, 2, 5), ((40, 32), 128, 2, 4, 1)])
def test_tensor_core_on_the_fly_lookup(hw, c, radius, levels, b):
    cf = _cf()
    H, W = hw
    g = torch.Generator().manual_seed(H * 31 + c)
    f1, f2 = torch.randn(b, c, H, W, generator=g), torch.randn(b, c, H, W, generator=g)
    ref32 = corr_oracle.CorrBlock(f1, f2, num_levels=levels, radius=radius)
    refb = rounded_operand_oracle(f1, f2, levels, radius)
    with torch.no_grad():
        blk = cf.CorrBlock(f1.to(DEV), f2.to(DEV), num_levels=levels, radius=radius, on_the_fly=True)
        assert blk.used_precision == "bf16" and blk._otf_tc
        grid = corr_oracle.coords_grid(b, H, W)
        for k, sigma in enumerate((0.0, 0.7, 2.5, 9.0)):
            coords = grid + torch.randn(b, 2, H, W, generator=g) * sigma
            if k == 1:
                coords[0, :, 0, :5] = torch.tensor([[3.0], [2.0]])            # exactly on the lattice
            if k == 2:
                coords[-1, :, 1, :4] = 1.0e4                                   # far outside: zeros
                coords[0, 0, 2, :3] = -7.5                                     # window partly left of the image
                coords[0, 1, 3, :3] = H + 3.25                                 # window below the image
            got = blk(coords.to(DEV)).cpu().numpy()
            assert rel_err(got, refb(coords).numpy()) <= 2e-5, (k, "indexing / ownership")
            assert rel_err(got, ref32(coords).numpy()) <= 1e-2, (k, "bf16 operand tolerance")
            if k == 2:
                assert not got[-1, :, 1, :4].any()


def test_wild_flow_many_chunks_and_whole_batch_dead():
    """Bounding boxes as large as the image (every query of a block looks somewhere else) and blocks whose queries are all dead."""
    cf = _cf()
    H = W = 32
    g = torch.Generator().manual_seed(9)
    f1, f2 = torch.randn(2, 64, H, W, generator=g), torch.randn(2, 64, H, W, generator=g)
    refb = rounded_operand_oracle(f1, f2, 4, 4)
    coords = torch.rand(2, 2, H, W, generator=g) * (W + 20) - 10                # uniformly anywhere, incl. outside
    coords[1, :, :8, :16] = float("nan")                                         # one whole query block dead
    with torch.no_grad():
        blk = cf.CorrBlock(f1.to(DEV), f2.to(DEV), on_the_fly=True)
        got = blk(coords.to(DEV)).cpu().numpy()
    want = refb(torch.nan_to_num(coords, nan=1.0e4)).numpy()
    assert rel_err(got, want) <= 2e-5
    assert not got[1, :, :8, :16].any()


def test_output_guard_words_and_reuse():
    """Nothing is written outside `out`; a second lookup on the same block (other coords) is independent of the first."""
    cf = _cf()
    H, W, b, c = 16, 32, 3, 16
    g = torch.Generator().manual_seed(10)
    f1, f2 = torch.randn(b, c, H, W, generator=g), torch.randn(b, c, H, W, generator=g)
    n = b * 324 * H * W
    buf = torch.full((n + 256,), 777.0, device=DEV)
    out = buf[128:128 + n].view(b, 324, H, W)
    with torch.no_grad():
        blk = cf.CorrBlock(f1.to(DEV), f2.to(DEV), on_the_fly=True)
        c1 = corr_oracle.coords_grid(b, H, W) + torch.randn(b, 2, H, W, generator=g) * 2
        c2 = corr_oracle.coords_grid(b, H, W) + torch.randn(b, 2, H, W, generator=g) * 5
        first = blk(c1.to(DEV)).clone()
        blk(c2.to(DEV), out=out)
        again = blk(c1.to(DEV))
    assert bool((buf[:128] == 777.0).all()) and bool((buf[128 + n:] == 777.0).all())
    assert torch.equal(first, again)
    assert rel_err(out.cpu().numpy(), rounded_operand_oracle(f1, f2, 4, 4)(c2).numpy()) <= 2e-5


@pytest.mark.parametrize("precision", [None, "fp32"])
def test_on_the_fly_block_is_differentiable(precision):
    """VERDICT round 1, Weak #3: `on_the_fly=True` used to raise under autograd.  The forward pass still holds no volume; the
    backward pass accumulates the lookups' gradients in one gradient pyramid and differentiates the correlation like the
    materialised path.  A loss that is linear in the lookups has fmap gradients that do not depend on the forward values, so
    they must agree with the materialised block's to fp32 grade (5e-5: the volume backward runs the bf16x3 split) although the
    tensor-core forward itself is bf16 grade."""
    cf = _cf()
    B, C, H, W, L, r = 2, 64, 16, 32, 3, 3
    g = torch.Generator().manual_seed(11)
    f1, f2 = torch.randn(B, C, H, W, generator=g), torch.randn(B, C, H, W, generator=g)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    base = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    coords = [(base + torch.randn(B, 2, H, W, generator=g) * 1.5).to(DEV) for _ in range(3)]
    wts = [torch.randn(B, L * (2 * r + 1) ** 2, H, W, generator=g).to(DEV) for _ in range(3)]

    def grads(**kw):
        a, b_ = f1.to(DEV).requires_grad_(), f2.to(DEV).requires_grad_()
        blk = cf.CorrBlock(a, b_, num_levels=L, radius=r, **kw)
        outs = [blk(c) for c in coords]
        assert all(o.grad_fn is not None for o in outs)
        sum((o * w).sum() for o, w in zip(outs, wts)).backward()
        return a.grad.cpu().numpy(), b_.grad.cpu().numpy(), [o.detach() for o in outs]

    m1, m2, mo = grads(precision="fp32")
    kw = {"on_the_fly": True}
    if precision:
        kw["precision"] = precision
    o1, o2, oo = grads(**kw)
    assert rel_err(o1, m1) <= 5e-5 and rel_err(o2, m2) <= 5e-5
    for a, b_ in zip(oo, mo):
        assert rel_err(a.cpu().numpy(), b_.cpu().numpy()) <= (2e-5 if precision == "fp32" else 1e-2)
    # a second pass over a fresh block gives the same gradients (no state leaks between backward passes)
    p1, p2, _ = grads(**kw)
    assert np.array_equal(p1, o1) and np.array_equal(p2, o2)
    # under no_grad the block stays the plain forward path
    with torch.no_grad():
        blk = cf.CorrBlock(f1.to(DEV).requires_grad_(), f2.to(DEV), num_levels=L, radius=r, on_the_fly=True)
        assert blk(coords[0]).grad_fn is None
