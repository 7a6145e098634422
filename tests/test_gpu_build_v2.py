"""GPU parity of version 2 of the correlation build (csrc/corr_tc2.cu) through the C ABI (cf_corr_build_ex).

Oracle: oracle/corr_oracle.CorrBlock (torch CPU fp32 restatement of torchvision raft.py:387-392, 424-431; reference call
site /root/reference/nnunet/network_architecture/SegFlowGaussian.py:929).  Tolerances (norm: max|a-b| / max|b|):
  * precision "bf16"   (one tcgen05 MMA per K step, operands rounded to bf16)      <= 1e-2   (north_star)
  * precision "bf16x3" (hi/lo bf16 splits, hi*hi + hi*lo + lo*hi, fp32 accumulate)  <= 1e-5   -- what "auto" selects, because the
    reference computes this volume in fp32 (SegFlowGaussian.py:2909 runs without autocast)
  * pooled levels: bit-identical to avg_pool2d of the level below in every mode.
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import corr_oracle, taps

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cf():
    import cardioflow_b200 as cf
    return cf


SHAPES = [((32, 32), 256, 3), ((64, 64), 128, 2), ((28, 32), 64, 2), ((24, 24), 64, 3), ((16, 16), 256, 2), ((8, 8), 8, 2),
          ((40, 96), 32, 1), ((12, 40), 72, 2), ((64, 32), 200, 1), ((20, 8), 24, 2)]


@pytest.mark.parametrize("hw,c,b", SHAPES)
def test_split_mode_is_fp32_grade_and_auto_selects_it(hw, c, b):
    cf = _cf()
    h, w = hw
    L = 4 if min(h, w) >= 16 else (3 if min(h, w) >= 8 else 2)
    g = torch.Generator().manual_seed(c * 7 + h)
    f1, f2 = torch.randn(b, c, h, w, generator=g), torch.randn(b, c, h, w, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2, num_levels=L)
    with torch.no_grad():
        auto = cf.CorrBlock(f1.to(DEV), f2.to(DEV), num_levels=L)
        assert auto.used_precision == "bf16x3", auto.used_precision
        bf = cf.CorrBlock(f1.to(DEV), f2.to(DEV), num_levels=L, precision="bf16")
        assert bf.used_precision == "bf16"
    for blk, tol in ((auto, 1e-5), (bf, 1e-2)):
        pyr = [p.cpu().numpy() for p in blk.corr_pyramid]
        for lvl in range(L):
            assert pyr[lvl].shape == tuple(ref.corr_pyramid[lvl].shape)
            assert rel_err(pyr[lvl], ref.corr_pyramid[lvl].numpy()) <= tol, (lvl, blk.used_precision)
            if lvl:
                assert np.array_equal(pyr[lvl][:, 0], taps.avg_pool2x2(pyr[lvl - 1][:, 0])), lvl
    # the bf16 mode must really be the coarser one (guards against both modes running the same code)
    e3 = rel_err(auto.corr_pyramid[0].cpu().numpy(), ref.corr_pyramid[0].numpy())
    e1 = rel_err(bf.corr_pyramid[0].cpu().numpy(), ref.corr_pyramid[0].numpy())
    assert e1 > 20 * e3, (e1, e3)


@pytest.mark.parametrize("hw,c", [((32, 32), 256), ((64, 64), 64), ((24, 40), 48)])
def test_input_dtypes_and_layouts(hw, c):
    """fp32 / bf16, NCHW / channels_last: same volume (bf16 inputs: the volume of the rounded maps); a bf16 channels_last map is
    read by the TMA directly (no workspace)."""
    cf = _cf()
    from cardioflow_b200 import _cabi
    h, w = hw
    g = torch.Generator().manual_seed(h + c)
    f1, f2 = torch.randn(2, c, h, w, generator=g), torch.randn(2, c, h, w, generator=g)
    d1, d2 = f1.to(DEV), f2.to(DEV)
    with torch.no_grad():
        base = cf.CorrBlock(d1, d2)                                                     # fp32 NCHW, bf16x3
        cl = cf.CorrBlock(d1.contiguous(memory_format=torch.channels_last), d2.contiguous(memory_format=torch.channels_last))
        assert cl.used_precision == "bf16x3"
        for a, b_ in zip(base.corr_pyramid, cl.corr_pyramid):
            assert torch.equal(a, b_)                                                   # same operands, same MMAs
        h1, h2 = d1.bfloat16(), d2.bfloat16()
        want = corr_oracle.CorrBlock(h1.float().cpu(), h2.float().cpu())
        nchw = cf.CorrBlock(h1, h2)
        nhwc = cf.CorrBlock(h1.contiguous(memory_format=torch.channels_last), h2.contiguous(memory_format=torch.channels_last))
        assert nchw.used_precision == "bf16" and nhwc.used_precision == "bf16"          # a bf16 map has no low part
        for lvl in range(4):
            assert torch.equal(nchw.corr_pyramid[lvl], nhwc.corr_pyramid[lvl])
            # products of bf16 values are exact in fp32; only the accumulation order differs from the oracle
            assert rel_err(nhwc.corr_pyramid[lvl].cpu().numpy(), want.corr_pyramid[lvl].numpy()) <= 1e-5
    # zero-copy claim: the ABI accepts bf16 NHWC with NO workspace
    lib = _cabi.lib()
    x1 = h1.contiguous(memory_format=torch.channels_last)
    x2 = h2.contiguous(memory_format=torch.channels_last)
    pyr = torch.empty(lib.cf_corr_pyramid_bytes(2, h, w, 4) // 4, device=DEV)
    used = _cabi.ctypes.c_int(-1)
    rc = lib.cf_corr_build_ex(x1.data_ptr(), x2.data_ptr(), _cabi.DTYPE_BF16, _cabi.LAYOUT_NHWC, 2, c, h, w, 4, _cabi.PREC_AUTO, 0,
                              pyr.data_ptr(), None, 0, _cabi.ctypes.byref(used), None, torch.cuda.current_stream().cuda_stream)
    assert rc == 0 and used.value == _cabi.PREC_BF16
    from cardioflow_b200.corr import _level_views
    for got, ref_lv in zip(_level_views(pyr, 2, h, w, 4), nhwc.corr_pyramid):      # (the packed buffer has alignment gaps)
        assert torch.equal(got, ref_lv)


def test_outputs_stay_inside_the_pyramid_buffer():
    """Guard words around the packed pyramid (compute-sanitizer is closed on this pool): the persistent kernel's direct stores,
    incl. a query block that hangs over the end of the map (HW % 128 != 0) and x-blocks beyond the map width."""
    cf = _cf()
    from cardioflow_b200 import _cabi
    lib = _cabi.lib()
    for (h, w, c, b) in ((24, 24, 64, 3), (12, 40, 16, 2), (32, 32, 256, 2)):
        n = lib.cf_corr_pyramid_bytes(b, h, w, 3) // 4
        buf = torch.full((n + 128,), 777.0, device=DEV)
        g = torch.Generator().manual_seed(h)
        f1, f2 = torch.randn(b, c, h, w, generator=g).to(DEV), torch.randn(b, c, h, w, generator=g).to(DEV)
        with torch.no_grad():
            blk = cf.CorrBlock(f1, f2, num_levels=3, pyramid_buffer=buf[64:64 + n])
        torch.cuda.synchronize()
        assert bool((buf[:64] == 777.0).all()) and bool((buf[64 + n:] == 777.0).all())
        ref = corr_oracle.CorrBlock(f1.cpu(), f2.cpu(), num_levels=3)
        for lvl in range(3):
            assert rel_err(blk.corr_pyramid[lvl].cpu().numpy(), ref.corr_pyramid[lvl].numpy()) <= 1e-5


def test_version1_still_serves_shapes_version2_rejects():
    """W % 8 != 0 (cfg1: 28 x 28): "bf16" goes through version 1, "auto" through the exact fp32 kernel."""
    cf = _cf()
    g = torch.Generator().manual_seed(5)
    f1, f2 = torch.randn(1, 256, 28, 28, generator=g), torch.randn(1, 256, 28, 28, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2)
    with torch.no_grad():
        a = cf.CorrBlock(f1.to(DEV), f2.to(DEV))
        b = cf.CorrBlock(f1.to(DEV), f2.to(DEV), precision="bf16")
    assert a.used_precision == "fp32" and b.used_precision == "bf16"
    assert rel_err(a.corr_pyramid[0].cpu().numpy(), ref.corr_pyramid[0].numpy()) <= 1e-5
    assert rel_err(b.corr_pyramid[0].cpu().numpy(), ref.corr_pyramid[0].numpy()) <= 1e-2
