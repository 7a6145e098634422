"""GPU parity: the warp kernels (through the C ABI / module API) against the oracle and the golden vectors.

Tolerances: tap coordinates bit-exact; bilinear values bit-exact against the C oracle (same op order, fmaf as
ATen's CUDA kernel) and <= 1e-5 relative (max|diff| / max|ref|) against torch's CPU grid_sample / fixtures.
"""
import numpy as np
import pytest
import torch

import cardioflow_b200 as cf
from conftest import WARP_CASES, rel_err
from oracle import c_oracle, taps, warp_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _gpu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("name", WARP_CASES)
def test_golden_fixture(golden_warp, name):
    flow, src = golden_warp[f"{name}.flow"], golden_warp[f"{name}.src"]
    st = cf.SpatialTransformer(flow.shape[2:]).to(DEV)
    out = st(flow=_gpu(flow), original=_gpu(src)).cpu().numpy()
    assert rel_err(out, golden_warp[f"{name}.bilinear"]) <= 1e-5
    assert np.array_equal(out, c_oracle.warp2d(flow, src))
    t = cf.warp_taps(_gpu(flow)).cpu().numpy()
    nt = taps.warp_taps(flow)
    assert np.array_equal(t[:, 0], nt["uy"]) and np.array_equal(t[:, 1], nt["ux"])
    near = cf.SpatialTransformer(flow.shape[2:], mode="nearest").to(DEV)(_gpu(flow), _gpu(src)).cpu().numpy()
    assert np.array_equal(near, golden_warp[f"{name}.nearest"])


@pytest.mark.parametrize("size", [(128, 128), (192, 192), (224, 224), (256, 256), (30, 45)])
@pytest.mark.parametrize("chan", [1, 2, 4, 5])
def test_shapes_and_flows(size, chan):
    g = torch.Generator().manual_seed(size[0] * 10 + chan)
    h, w = size
    src = torch.rand(2, chan, h, w, generator=g)
    flows = {
        "zero": torch.zeros(2, 2, h, w),
        "integer": torch.randint(-5, 6, (2, 2, h, w), generator=g).float(),
        "half": torch.full((2, 2, h, w), -0.5),
        "noise": torch.randn(2, 2, h, w, generator=g) * 3,
        "large": torch.randn(2, 2, h, w, generator=g) * h,
    }
    st = cf.SpatialTransformer(size).to(DEV)
    ref_mod = warp_oracle.SpatialTransformer(size)
    for kind, flow in flows.items():
        out = st(flow.to(DEV), src.to(DEV)).cpu()
        ref = ref_mod(flow, src)
        assert rel_err(out.numpy(), ref.numpy()) <= 1e-5, kind
        if chan == 1:
            assert np.array_equal(out.numpy(), c_oracle.warp2d(flow.numpy(), src.numpy())), kind


def test_module_api_quirks():
    g = torch.Generator().manual_seed(3)
    flow, src = torch.randn(1, 2, 32, 32, generator=g) * 2, torch.rand(1, 3, 32, 32, generator=g)
    st = cf.SpatialTransformer((32, 32))
    st = st.to(DEV)
    assert st.grid.is_cuda
    a = st(flow=flow.to(DEV), original=src.to(DEV))
    b = st(flow.to(DEV), src.to(DEV), mode="nearest")           # forward's mode is ignored (integration.py:79)
    assert torch.equal(a, b)
    c = cf.warp(src.to(DEV), flow.to(DEV))                       # voxelmorph-order alias
    assert torch.equal(a, c)
    nc = src.to(DEV).permute(0, 1, 3, 2).contiguous().permute(0, 1, 3, 2)   # non-contiguous input
    assert torch.equal(st(flow.to(DEV), nc), a)
    with pytest.raises(ValueError):
        st(torch.zeros(1, 2, 16, 16, device=DEV), torch.zeros(1, 1, 16, 16, device=DEV))
    flow_before = flow.to(DEV).clone()
    fl = flow_before.clone()
    st(fl, src.to(DEV))
    assert torch.equal(fl, flow_before)                          # inputs are not modified


def test_onehot_argmax(golden_warp):
    flow, label = golden_warp["argmax.flow"], golden_warp["argmax.label"]
    out = cf.warp_onehot_argmax(_gpu(flow), _gpu(label), 4).cpu().numpy()
    assert np.array_equal(out, golden_warp["argmax.out"])
    g = torch.Generator().manual_seed(11)
    label = torch.randint(0, 4, (3, 1, 224, 224), generator=g)
    flow = torch.randn(3, 2, 224, 224, generator=g) * 4
    out = cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV)).cpu()
    assert torch.equal(out, warp_oracle.warp_onehot_argmax(flow, label))
    # and it equals the unfused module path on the device
    onehot = torch.nn.functional.one_hot(label[:, 0], 4).permute(0, 3, 1, 2).contiguous().float().to(DEV)
    unfused = torch.argmax(cf.SpatialTransformer((224, 224)).to(DEV)(flow.to(DEV), onehot), dim=1, keepdim=True)
    assert torch.equal(out, unfused.cpu())


@pytest.mark.parametrize("shape,ncls", [((3, 224, 224), 4), ((2, 37, 50), 4), ((5, 16, 33), 2), ((1, 96, 200), 7)])
def test_onehot_argmax_counts(shape, ncls):
    """cf_warp2d_onehot_argmax_counts: the label map equals the count-free entry point (and the oracle), counts equal a
    bincount of every warped map; widths that are not a multiple of the 32-column tile exercise the predicated lanes."""
    B, H, W = shape
    g = torch.Generator().manual_seed(B * 1000 + W)
    label = torch.randint(0, ncls, (B, 1, H, W), generator=g)
    flow = torch.randn(B, 2, H, W, generator=g) * 3
    plain = cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV), ncls)
    buf = torch.full((B, 1, H, W), -7, dtype=torch.int64, device=DEV)
    out, counts = cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV), ncls, out=buf, return_counts=True)
    assert out.data_ptr() == buf.data_ptr() and torch.equal(out, plain)
    assert torch.equal(out.cpu(), warp_oracle.warp_onehot_argmax(flow, label, ncls))
    want = torch.stack([torch.bincount(out[b].flatten().cpu(), minlength=ncls) for b in range(B)])
    assert torch.equal(counts.cpu(), want) and int(counts.sum()) == B * H * W
    # counts are overwritten, not accumulated, by a second call into the same storage
    out2, counts2 = cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV), ncls, return_counts=True)
    assert torch.equal(counts2, counts)
    # counts_out: a row slice of a larger table is filled in place, rows outside it stay untouched
    table = torch.full((B + 2, ncls), -1, dtype=torch.int64, device=DEV)
    _, c3 = cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV), ncls, return_counts=True, counts_out=table[1:B + 1])
    assert c3.data_ptr() == table[1:].data_ptr() and torch.equal(table[1:B + 1], counts)
    assert bool((table[0] == -1).all()) and bool((table[B + 1] == -1).all())
    with pytest.raises(ValueError):
        cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV), ncls, return_counts=True, counts_out=table)
    with pytest.raises(ValueError):
        cf.warp_onehot_argmax(flow.to(DEV), label.to(DEV), ncls, out=torch.empty(B, 1, H, W, device=DEV))


def test_full_size_properties_cfg3():
    """BASELINE cfg 3 size (300 x (1+4) x 224^2): size-independent properties instead of a CPU oracle."""
    g = torch.Generator(device=DEV).manual_seed(5)
    src = torch.rand(300, 5, 224, 224, device=DEV, generator=g)
    st = cf.SpatialTransformer((224, 224)).to(DEV)
    ident = st(torch.zeros(300, 2, 224, 224, device=DEV), src)
    # zero flow = identity up to the fp32 round trip: integer coordinates can land a few ulp(223) = 1.5e-5 off
    assert float((ident - src).abs().max()) <= 1e-4
    flow = torch.randn(300, 2, 224, 224, device=DEV, generator=g) * 3
    other = torch.rand(300, 5, 224, 224, device=DEV, generator=g)
    lin = st(flow, 2.0 * src + other)
    assert float((lin - (2.0 * st(flow, src) + st(flow, other))).abs().max()) <= 1e-5      # linear in the source
    shift = torch.zeros(300, 2, 224, 224, device=DEV)
    shift[:, 1] = 1000.0
    assert float(st(shift, src).abs().max()) == 0.0             # far out of range -> zero padding
    # the WHOLE batch against the oracle (torch CPU grid_sample behind the reference's normalisation), not a sample
    ref = warp_oracle.SpatialTransformer((224, 224))(flow.cpu(), src.cpu())
    got = st(flow, src).cpu()
    assert rel_err(got.numpy(), ref.numpy()) <= 1e-5
    assert float((got - ref).abs().max()) <= 1e-5


@pytest.mark.parametrize("name", ["uw_24", "uw_17x40", "uw_oob_16", "uw_zero_12"])
def test_utils_warp_golden(golden_utils_warp, name):
    """cf_warp2d_xy (lib/utils.py:29-59 `warp`, align_corners=False, flow ch0 = x) against the reference's own outputs."""
    import cardioflow_b200 as cf
    g = golden_utils_warp
    with torch.no_grad():
        got = cf.utils_warp(torch.from_numpy(g[f"{name}.x"]).cuda(), torch.from_numpy(g[f"{name}.flo"]).cuda())
    assert rel_err(got.cpu().numpy(), g[f"{name}.out"]) <= 1e-5


def test_utils_warp_full_size_against_oracle():
    import cardioflow_b200 as cf
    from oracle import warp_oracle
    gen = torch.Generator().manual_seed(9)
    x = torch.rand(4, 5, 224, 224, generator=gen)
    flo = torch.randn(4, 2, 224, 224, generator=gen) * 3
    with torch.no_grad():
        got = cf.utils_warp(x.cuda(), flo.cuda())
    assert rel_err(got.cpu().numpy(), warp_oracle.utils_warp(x, flo).numpy()) <= 1e-5
    with pytest.raises(RuntimeError):
        cf.utils_warp(x, flo)          # CPU tensors: no fallback


@pytest.mark.parametrize("shape", [(2, 3, 24, 24), (1, 5, 17, 40), (1, 2, 16, 16)])
def test_utils_warp_backward_matches_autograd_of_the_oracle(shape):
    """Gradients of cf_warp2d_xy w.r.t. x and flo against torch autograd through the CPU restatement of utils.py:29-59."""
    import cardioflow_b200 as cf
    from oracle import warp_oracle
    b, c, h, w = shape
    gen = torch.Generator().manual_seed(sum(shape))
    x, flo = torch.rand(b, c, h, w, generator=gen), torch.randn(b, 2, h, w, generator=gen) * 2 + 0.3
    gout = torch.randn(b, c, h, w, generator=gen)
    xr, fr = x.clone().requires_grad_(), flo.clone().requires_grad_()
    warp_oracle.utils_warp(xr, fr).backward(gout)
    xg, fg = x.cuda().requires_grad_(), flo.cuda().requires_grad_()
    cf.utils_warp(xg, fg).backward(gout.cuda())
    assert rel_err(xg.grad.cpu().numpy(), xr.grad.numpy()) <= 1e-5
    assert rel_err(fg.grad.cpu().numpy(), fr.grad.numpy()) <= 1e-4


def test_empty_batch_and_zero_channels_return_empty_like_grid_sample():
    """Edge case the ATen path of the reference accepts: B = 0 or C = 0 give empty outputs (and gradients), no kernel launch."""
    st = cf.SpatialTransformer((16, 20)).to(DEV)
    out = st(torch.zeros(0, 2, 16, 20, device=DEV), torch.zeros(0, 3, 16, 20, device=DEV))
    assert out.shape == (0, 3, 16, 20)
    flow = torch.zeros(2, 2, 16, 20, device=DEV, requires_grad=True)
    out = st(flow, torch.zeros(2, 0, 16, 20, device=DEV))
    assert out.shape == (2, 0, 16, 20)
    out.sum().backward()
    assert flow.grad is not None and not bool(flow.grad.any())
    lab, cnt = cf.warp_onehot_argmax(torch.zeros(0, 2, 16, 20, device=DEV), torch.zeros(0, 1, 16, 20, dtype=torch.int64, device=DEV),
                                     return_counts=True)
    assert lab.shape == (0, 1, 16, 20) and cnt.shape == (0, 4)
    want = warp_oracle.SpatialTransformer((16, 20))(torch.zeros(0, 2, 16, 20), torch.zeros(0, 3, 16, 20))
    assert want.shape == out.new_zeros(0, 3, 16, 20).shape
