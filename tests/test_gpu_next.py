"""GPU parity of the SURVEY.md 8(f) rows and of the on-the-fly lookup (K4), through the C ABI.

Oracles: torch-CPU autograd of the oracle modules (ATen grid_sampler_2d_backward / avg_pool2d backward / matmul
backward are the reference's own backward), the frozen reference outputs for VecInt / SpatialTransformerContour
(tests/golden/warp_reference.npz, produced by oracle/make_golden.py from the reference's integration.py).
Tolerance norm everywhere: max|a-b| / max|b| (conftest.rel_err).
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import corr_oracle, warp_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cf():
    import cardioflow_b200 as cf
    return cf


# ------------------------------------------------------------------------------------------- warp backward
@pytest.mark.parametrize("size,chan", [((32, 40), 1), ((64, 64), 4), ((30, 45), 2), ((128, 128), 5)])
def test_warp_backward_matches_aten(size, chan):
    cf = _cf()
    g = torch.Generator().manual_seed(11)
    flow = (torch.randn(2, 2, *size, generator=g) * 3).requires_grad_()
    src = torch.rand(2, chan, *size, generator=g).requires_grad_()
    wgt = torch.randn(2, chan, *size, generator=g)
    (warp_oracle.SpatialTransformer(size)(flow, src) * wgt).sum().backward()
    flow_d = flow.detach().to(DEV).requires_grad_()
    src_d = src.detach().to(DEV).requires_grad_()
    st = cf.SpatialTransformer(size).to(DEV)
    (st(flow_d, src_d) * wgt.to(DEV)).sum().backward()
    assert rel_err(src_d.grad.cpu().numpy(), src.grad.numpy()) <= 1e-5     # atomics: summation order only
    assert rel_err(flow_d.grad.cpu().numpy(), flow.grad.numpy()) <= 1e-5


def test_warp_backward_only_one_input_needs_grad():
    cf = _cf()
    g = torch.Generator().manual_seed(12)
    flow, src = torch.randn(1, 2, 24, 24, generator=g), torch.rand(1, 3, 24, 24, generator=g)
    st = cf.SpatialTransformer((24, 24)).to(DEV)
    f = flow.to(DEV).requires_grad_()
    st(f, src.to(DEV)).sum().backward()
    fr = flow.clone().requires_grad_()
    warp_oracle.SpatialTransformer((24, 24))(fr, src).sum().backward()
    assert rel_err(f.grad.cpu().numpy(), fr.grad.numpy()) <= 1e-5
    s = src.to(DEV).requires_grad_()
    st(flow.to(DEV), s).sum().backward()
    sr = src.clone().requires_grad_()
    warp_oracle.SpatialTransformer((24, 24))(flow, sr).sum().backward()
    assert rel_err(s.grad.cpu().numpy(), sr.grad.numpy()) <= 1e-5


# ------------------------------------------------------------------------------------------- VecInt / Contour
def test_vecint_matches_reference_fixture(golden_warp):
    cf = _cf()
    vec = torch.from_numpy(golden_warp["vecint.vec"]).to(DEV)
    with torch.no_grad():
        out = cf.VecInt((32, 32), 7).to(DEV)(vec)
    assert np.array_equal(out.cpu().numpy(), golden_warp["vecint.out7"])   # same fp32 op sequence: bit-exact


@pytest.mark.parametrize("nsteps", [0, 1, 2, 5, 7])
def test_vecint_steps_and_autograd_path(nsteps):
    cf = _cf()
    g = torch.Generator().manual_seed(13)
    vec = torch.randn(2, 2, 48, 40, generator=g) * 4
    want = warp_oracle.VecInt((48, 40), nsteps)(vec)
    mod = cf.VecInt((48, 40), nsteps).to(DEV)
    with torch.no_grad():
        got = mod(vec.to(DEV))
    assert np.array_equal(got.cpu().numpy(), want.numpy())
    # autograd path (composed from the differentiable warp) gives the same forward and the oracle's gradient
    v = vec.to(DEV).requires_grad_()
    out = mod(v)
    assert rel_err(out.detach().cpu().numpy(), want.numpy()) <= 1e-6
    out.sum().backward()
    vr = vec.clone().requires_grad_()
    warp_oracle.VecInt((48, 40), nsteps)(vr).sum().backward()
    assert rel_err(v.grad.cpu().numpy(), vr.grad.numpy()) <= 1e-4


def test_contour_matches_reference_fixture(golden_warp):
    cf = _cf()
    locs = torch.from_numpy(golden_warp["contour.locs"]).to(DEV)
    field = torch.from_numpy(golden_warp["contour.field"]).to(DEV)
    out = cf.SpatialTransformerContour((32, 32)).to(DEV)(locs, field)
    assert rel_err(out.cpu().numpy(), golden_warp["contour.out"]) <= 1e-6
    assert float(locs.abs().max()) <= 1.0 + 1e-6      # the argument is normalised in place, like the reference


@pytest.mark.parametrize("mode", ["bilinear", "nearest"])
def test_contour_rectangular_and_modes(mode):
    cf = _cf()
    g = torch.Generator().manual_seed(14)
    field = torch.randn(2, 3, 40, 56, generator=g)
    locs = torch.stack([torch.rand(2, 1, 77, generator=g) * 60 - 2, torch.rand(2, 1, 77, generator=g) * 44 - 2], dim=1)
    want = warp_oracle.SpatialTransformerContour((40, 56), mode)(locs.clone(), field)
    got = cf.SpatialTransformerContour((40, 56), mode).to(DEV)(locs.to(DEV), field.to(DEV))
    assert rel_err(got.cpu().numpy(), want.numpy()) <= 1e-6


# ------------------------------------------------------------------------------------------- CorrBlock backward
def _oracle_from_levels(levels, radius):
    blk = corr_oracle.CorrBlock.__new__(corr_oracle.CorrBlock)
    blk.num_levels, blk.radius, blk.corr_pyramid = len(levels), radius, levels
    return blk


@pytest.mark.parametrize("hw,radius,levels", [((16, 16), 4, 4), ((24, 20), 3, 3), ((32, 32), 4, 4), ((17, 21), 2, 2)])
def test_lookup_backward_wrt_pyramid(hw, radius, levels):
    cf = _cf()
    from cardioflow_b200.corr import _CorrLookup
    H, W = hw
    g = torch.Generator().manual_seed(21)
    B = 2
    lv = []
    h, w = H, W
    for _ in range(levels):
        lv.append(torch.randn(B * H * W, 1, h, w, generator=g).requires_grad_())
        h, w = h // 2, w // 2
    coords = corr_oracle.coords_grid(B, H, W) + torch.randn(B, 2, H, W, generator=g) * 3
    coords[0, :, 0, 0] = torch.tensor([-30.0, 5.0])    # far outside
    coords[0, :, 0, 1] = torch.tensor([3.0, 7.0])      # exactly on the integer lattice
    K = levels * (2 * radius + 1) ** 2
    wgt = torch.randn(B, K, H, W, generator=g)
    (_oracle_from_levels(lv, radius)(coords) * wgt).sum().backward()

    blk = cf.CorrBlock.from_pyramid([t.detach().to(DEV) for t in lv], radius=radius)
    packed = blk._pyramid.clone().requires_grad_()
    out = _CorrLookup.apply(packed, coords.to(DEV), blk._meta, None)
    (out * wgt.to(DEV)).sum().backward()
    from cardioflow_b200.corr import _level_views
    for got, ref in zip(_level_views(packed.grad, B, H, W, levels), lv):
        assert rel_err(got.cpu().numpy(), ref.grad.numpy()) <= 1e-5


@pytest.mark.parametrize("hw,c", [((16, 16), 32), ((24, 24), 64), ((32, 32), 256), ((17, 21), 24)])
def test_corrblock_backward_to_fmaps(hw, c):
    cf = _cf()
    H, W = hw
    g = torch.Generator().manual_seed(22)
    f1 = torch.randn(2, c, H, W, generator=g).requires_grad_()
    f2 = torch.randn(2, c, H, W, generator=g).requires_grad_()
    coords = [corr_oracle.coords_grid(2, H, W) + torch.randn(2, 2, H, W, generator=g) * k for k in (0.0, 1.5, 4.0)]
    wgts = [torch.randn(2, 324, H, W, generator=g) for _ in coords]
    ref = corr_oracle.CorrBlock(f1, f2)
    sum((ref(cd) * wg).sum() for cd, wg in zip(coords, wgts)).backward()

    d1, d2 = f1.detach().to(DEV).requires_grad_(), f2.detach().to(DEV).requires_grad_()
    blk = cf.CorrBlock(d1, d2, radius=4, precision="fp32")
    sum((blk(cd.to(DEV)) * wg.to(DEV)).sum() for cd, wg in zip(coords, wgts)).backward()
    assert rel_err(d1.grad.cpu().numpy(), f1.grad.numpy()) <= 2e-5
    assert rel_err(d2.grad.cpu().numpy(), f2.grad.numpy()) <= 2e-5


def test_pool_backward_odd_sizes():
    """7 -> 3 style levels: the trailing row / column of an odd level receives no gradient from the pool."""
    cf = _cf()
    from cardioflow_b200 import _cabi
    from cardioflow_b200.corr import _level_views
    B, H, W, L = 1, 28, 28, 4                       # 28 -> 14 -> 7 -> 3
    g = torch.Generator().manual_seed(23)
    lib = _cabi.lib()
    n = lib.cf_corr_pyramid_bytes(B, H, W, L) // 4
    gp = torch.zeros(n)
    views = _level_views(gp, B, H, W, L)
    for v in views:
        v.copy_(torch.randn(v.shape, generator=g))
    lv0 = torch.randn(B * H * W, 1, H, W, generator=g).requires_grad_()
    pyr = [lv0]
    for _ in range(L - 1):
        pyr.append(torch.nn.functional.avg_pool2d(pyr[-1], 2, stride=2))
    sum((p * v).sum() for p, v in zip(pyr, views)).backward()
    gd = gp.to(DEV)
    _cabi.check(lib.cf_corr_pool_bwd(gd.data_ptr(), B, H, W, L, torch.cuda.current_stream().cuda_stream))
    got = _level_views(gd, B, H, W, L)[0]
    assert rel_err(got.cpu().numpy(), lv0.grad.numpy()) <= 1e-6


# ------------------------------------------------------------------------------------------- on-the-fly lookup
@pytest.mark.parametrize("hw,c,radius,levels", [((32, 32), 64, 4, 4), ((24, 24), 16, 4, 3), ((17, 21), 8, 3, 2),
                                               ((28, 28), 256, 4, 4), ((16, 16), 5, 1, 4)])
def test_on_the_fly_lookup_matches_materialised(hw, c, radius, levels):
    cf = _cf()
    H, W = hw
    g = torch.Generator().manual_seed(31)
    f1, f2 = torch.randn(2, c, H, W, generator=g), torch.randn(2, c, H, W, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2, num_levels=levels, radius=radius)
    with torch.no_grad():
        blk = cf.CorrBlock(f1.to(DEV), f2.to(DEV), num_levels=levels, radius=radius, on_the_fly=True, precision="fp32")
        assert blk.used_precision == "fp32"       # the CUDA-core variant (the tensor-core one: tests/test_gpu_otf_tc.py)
        for k, sigma in enumerate((0.0, 2.0, 6.0)):
            coords = corr_oracle.coords_grid(2, H, W) + torch.randn(2, 2, H, W, generator=g) * sigma
            if k == 2:
                coords[1, :, 0, :4] = 1.0e4       # far outside: zeros
            want = ref(coords).numpy()
            got = blk(coords.to(DEV)).cpu().numpy()
            # differs from the materialised fp32 path by summation order only (pool-then-dot vs dot-then-pool)
            assert rel_err(got, want) <= 2e-5
            if k == 2:
                assert not got[1, :, 0, :4].any()


def test_on_the_fly_memory_is_linear_in_hw():
    from cardioflow_b200 import _cabi
    lib = _cabi.lib()
    assert lib.cf_fmap_pyramid_bytes(64, 256, 64, 64, 4) < 0.4e9 < 5.6e9 < lib.cf_corr_pyramid_bytes(64, 64, 64, 4)


@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 592 * 4 + 1, 1 << 20, 9_502_720 + 3])
def test_sum_f64_is_exact_enough_and_deterministic(n):
    """cf_sum_f64 (per-unit result rows of the cohort driver): equals the float64 sum of the fp32 values to double rounding,
    bit-identical across calls, writes only its output slot."""
    from cardioflow_b200 import shard
    gen = torch.Generator().manual_seed(n % 1000)
    x = (torch.randn(n, generator=gen) * 100).cuda()
    row = torch.full((3,), -7.0, dtype=torch.float64, device="cuda")
    shard.sum_f64(x, row[1:2])
    first = row.clone()
    shard.sum_f64(x, row[1:2])
    assert torch.equal(first, row) and float(row[0]) == -7.0 and float(row[2]) == -7.0
    want = x.double().sum()
    assert abs(float(row[1]) - float(want)) <= 1e-12 * float(x.double().abs().sum()) + 1e-300
    with pytest.raises(ValueError):
        shard.sum_f64(x.double(), row[1:2])
