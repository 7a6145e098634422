"""Convex 8x upsampling on the GPU through the C ABI, against the reference's own
upsample_flow / upsample_seg outputs (tests/golden/upsample_reference.npz) and the CPU oracle.

Tolerance: <= 1e-5 relative (max|diff| / max|ref|), the fp32 bar of north_star; the only non-bit-exact step is
expf (CUDA libdevice vs the host libm, ~1 ulp)."""
import numpy as np
import pytest
import torch

from conftest import UPSAMPLE_CASES, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.mark.parametrize("name", UPSAMPLE_CASES)
def test_golden_fixture(golden_upsample, name):
    import cardioflow_b200 as cf
    g = golden_upsample
    mask = torch.from_numpy(g[f"{name}.mask"]).cuda()
    with torch.no_grad():
        up_flow = cf.upsample_flow(torch.from_numpy(g[f"{name}.flow"]).cuda(), mask)
        up_seg = cf.upsample_seg(torch.from_numpy(g[f"{name}.seg"]).cuda(), mask)
    assert up_flow.shape == g[f"{name}.up_flow"].shape and up_seg.shape == g[f"{name}.up_seg"].shape
    assert rel_err(up_flow.cpu().numpy(), g[f"{name}.up_flow"]) <= TOL
    assert rel_err(up_seg.cpu().numpy(), g[f"{name}.up_seg"]) <= TOL


@pytest.mark.parametrize("shape", [(1, 1, 28, 28), (3, 2, 32, 32), (2, 4, 24, 40), (1, 6, 7, 3), (2, 2, 64, 64)])
def test_shapes_against_oracle(shape):
    import cardioflow_b200 as cf
    from cardioflow_b200 import _cabi
    from oracle import upsample_oracle
    n, c, h, w = shape
    gen = torch.Generator().manual_seed(h * 100 + w + c)
    field = torch.randn(n, c, h, w, generator=gen) * 3
    mask = torch.randn(n, 576, h, w, generator=gen) * 4
    want = upsample_oracle.upsample_convex(field, mask)
    out = torch.empty(n, c, 8 * h, 8 * w, device="cuda")
    f, m = field.cuda(), mask.cuda()
    _cabi.check(_cabi.lib().cf_upsample_convex(f.data_ptr(), m.data_ptr(), n, c, h, w, 8, 8.0, out.data_ptr(), _cabi.stream()))
    assert rel_err(out.cpu().numpy(), want.numpy()) <= TOL


def test_edge_cases_and_errors():
    import cardioflow_b200 as cf
    from cardioflow_b200 import _cabi
    f = torch.randn(1, 2, 8, 8, device="cuda")
    m = torch.zeros(1, 576, 8, 8, device="cuda")
    with torch.no_grad():
        up = cf.upsample_flow(f, m)            # uniform weights: 8 * the zero-padded 3x3 box mean, constant inside a cell
    box = torch.nn.functional.avg_pool2d(8 * f, 3, stride=1, padding=1, count_include_pad=True)
    assert rel_err(up.cpu().numpy(), box.repeat_interleave(8, 2).repeat_interleave(8, 3).cpu().numpy()) <= TOL
    big = torch.full((1, 576, 8, 8), -1e4, device="cuda")
    big[:, 4 * 64:5 * 64] = 1e4                # all weight on the centre tap, extreme logits stay finite
    with torch.no_grad():
        up = cf.upsample_flow(f, big)
    assert torch.equal(up, (8 * f).repeat_interleave(8, 2).repeat_interleave(8, 3))
    with pytest.raises(ValueError):
        cf.upsample_flow(f, m[:, :500])
    with pytest.raises(RuntimeError):
        cf.upsample_flow(f.cpu(), m.cpu())      # no CPU fallback
    out = torch.empty(1, 2, 64, 64, device="cuda")
    with pytest.raises(ValueError):
        _cabi.check(_cabi.lib().cf_upsample_convex(f.data_ptr(), m.data_ptr(), 1, 2, 8, 8, 4, 4.0, out.data_ptr(), _cabi.stream()))


def test_full_size_properties_cfg2():
    """29 pairs at 32x32 -> 256^2: convexity (a constant field comes back scaled by 8 in the interior for ANY mask) and
    linearity in the field, properties that hold at any size."""
    import cardioflow_b200 as cf
    gen = torch.Generator(device="cuda").manual_seed(7)
    mask = torch.randn(29, 576, 32, 32, device="cuda", generator=gen) * 2
    a = torch.randn(29, 2, 32, 32, device="cuda", generator=gen)
    b = torch.randn(29, 2, 32, 32, device="cuda", generator=gen)
    with torch.no_grad():
        const = cf.upsample_flow(torch.full((29, 2, 32, 32), 0.5, device="cuda"), mask)
        ua, ub, uab = cf.upsample_flow(a, mask), cf.upsample_flow(b, mask), cf.upsample_flow(a + b, mask)
    assert const.shape == (29, 2, 256, 256)
    assert float((const[:, :, 8:-8, 8:-8] - 4.0).abs().max()) <= 1e-5
    assert float((uab - (ua + ub)).abs().max()) <= 1e-4 * float(uab.abs().max())


@pytest.mark.parametrize("shape", [(2, 2, 12, 12), (1, 4, 9, 33), (2, 1, 7, 3), (1, 6, 8, 8), (3, 2, 32, 32)])
def test_backward_matches_autograd_of_the_oracle(shape):
    """Gradients w.r.t. the field and the mask against torch autograd through the CPU restatement of
    SegFlowGaussian.py:846-871 (what the reference's trainer differentiates).  <= 1e-4 relative: fp32 atomics order
    and ex2.approx in the recomputed softmax."""
    import cardioflow_b200 as cf
    from oracle import upsample_oracle
    n, c, h, w = shape
    gen = torch.Generator().manual_seed(sum(shape))
    field = torch.randn(n, c, h, w, generator=gen)
    mask = torch.randn(n, 576, h, w, generator=gen) * 2
    gout = torch.randn(n, c, 8 * h, 8 * w, generator=gen)
    fr, mr = field.clone().requires_grad_(), mask.clone().requires_grad_()
    upsample_oracle.upsample_convex(fr, mr).backward(gout)
    fg, mg = field.cuda().requires_grad_(), mask.cuda().requires_grad_()
    cf.upsample_seg(fg, mg).backward(gout.cuda()) if c == 4 else cf.upsample_flow(fg, mg).backward(gout.cuda())
    assert rel_err(fg.grad.cpu().numpy(), fr.grad.numpy()) <= 1e-4
    assert rel_err(mg.grad.cpu().numpy(), mr.grad.numpy()) <= 1e-4
    # only one input needs a gradient
    fg2 = field.cuda().requires_grad_()
    cf.upsample_flow(fg2, mask.cuda()).backward(gout.cuda())
    assert rel_err(fg2.grad.cpu().numpy(), fr.grad.numpy()) <= 1e-4
