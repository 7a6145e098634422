"""CPU: pin the oracle (three restatements) to the golden vectors frozen from the real reference."""
import importlib.util
import os
import warnings

import numpy as np
import pytest
import torch

from conftest import CORR_CASES, WARP_CASES, rel_err
from oracle import c_oracle, corr_oracle, taps, warp_oracle

REF_FILE = "/root/reference/nnunet/network_architecture/integration.py"


@pytest.mark.parametrize("name", WARP_CASES)
def test_warp_restatements_match_reference_fixture(golden_warp, name):
    flow, src = golden_warp[f"{name}.flow"], golden_warp[f"{name}.src"]
    ref = golden_warp[f"{name}.bilinear"]
    size = flow.shape[2:]
    t = warp_oracle.SpatialTransformer(size)(flow=torch.from_numpy(flow), original=torch.from_numpy(src)).numpy()
    assert np.array_equal(t, ref)                      # torch restatement: bit-exact with the reference module
    assert rel_err(taps.warp_from_taps(flow, src), ref) < 1e-6
    c, ctaps = c_oracle.warp2d(flow, src, want_taps=True)
    assert rel_err(c, ref) < 1e-6
    nt = taps.warp_taps(flow)
    assert np.array_equal(ctaps[:, 0], nt["uy"]) and np.array_equal(ctaps[:, 1], nt["ux"])
    near = warp_oracle.SpatialTransformer(size, mode="nearest")(torch.from_numpy(flow), torch.from_numpy(src)).numpy()
    assert np.array_equal(near, golden_warp[f"{name}.nearest"])
    assert np.array_equal(c_oracle.warp2d(flow, src, mode="nearest"), golden_warp[f"{name}.nearest"])


def test_forward_mode_argument_is_ignored(golden_warp):
    flow, src = torch.from_numpy(golden_warp["noise_48.flow"]), torch.from_numpy(golden_warp["noise_48.src"])
    st = warp_oracle.SpatialTransformer((48, 48))
    assert torch.equal(st(flow, src, mode="nearest"), st(flow, src))


def test_state_dict_grid(golden_warp):
    st = warp_oracle.SpatialTransformer((6, 5))
    assert list(st.state_dict().keys()) == ["grid"]
    assert np.array_equal(st.state_dict()["grid"].numpy(), golden_warp["state.grid_6x5"])


def test_onehot_argmax_vecint_contour(golden_warp):
    flow, label = golden_warp["argmax.flow"], golden_warp["argmax.label"]
    assert np.array_equal(warp_oracle.warp_onehot_argmax(torch.from_numpy(flow), torch.from_numpy(label)).numpy(),
                          golden_warp["argmax.out"])
    assert np.array_equal(c_oracle.warp_onehot_argmax(flow, label), golden_warp["argmax.out"])
    vec = torch.from_numpy(golden_warp["vecint.vec"])
    assert np.array_equal(warp_oracle.VecInt((32, 32), 7)(vec).numpy(), golden_warp["vecint.out7"])
    locs = torch.from_numpy(golden_warp["contour.locs"]).clone()
    out = warp_oracle.SpatialTransformerContour((32, 32))(locs, torch.from_numpy(golden_warp["contour.field"]))
    assert np.array_equal(out.numpy(), golden_warp["contour.out"])
    assert float(locs.abs().max()) <= 1.0 + 1e-6     # argument normalised in place, like the reference


@pytest.mark.parametrize("name", CORR_CASES)
def test_corr_restatements_match_torchvision_fixture(golden_corr, name):
    f1, f2 = golden_corr[f"{name}.f1"], golden_corr[f"{name}.f2"]
    blk = corr_oracle.CorrBlock(torch.from_numpy(f1), torch.from_numpy(f2), num_levels=4, radius=4)
    pyr = [p.numpy() for p in blk.corr_pyramid]
    cpyr = c_oracle.build_pyramid(f1, f2)
    assert rel_err(cpyr[0], pyr[0]) < 1e-5            # different summation order than BLAS
    for lvl in (1, 2, 3):
        assert np.array_equal(pyr[lvl], golden_corr[f"{name}.pyr{lvl}"])
        assert np.array_equal(taps.avg_pool2x2(pyr[lvl - 1][:, 0]), golden_corr[f"{name}.pyr{lvl}"][:, 0])
        assert np.array_equal(c_oracle.avg_pool2x2(pyr[lvl - 1]), golden_corr[f"{name}.pyr{lvl}"])
    for k in range(4):
        coords, ref = golden_corr[f"{name}.coords{k}"], golden_corr[f"{name}.lookup{k}"]
        assert np.array_equal(blk(torch.from_numpy(coords)).numpy(), ref)
        c, ctaps, cidx = c_oracle.lookup(pyr, coords, want_taps=True)
        scale = max(np.abs(ref).max(), 1.0)
        assert np.abs(c - ref).max() / scale < 1e-6
        assert np.abs(taps.lookup_from_taps(pyr, coords) - ref).max() / scale < 1e-6
        nt = taps.lookup_taps(coords, [p.shape[-2:] for p in pyr])
        for lvl in range(4):
            assert np.array_equal(ctaps[lvl][:, 0], nt[lvl]["ux"]) and np.array_equal(ctaps[lvl][:, 1], nt[lvl]["uy"])
            assert np.array_equal(cidx[lvl][:, 0], nt[lvl]["x0"]) and np.array_equal(cidx[lvl][:, 1], nt[lvl]["y0"])
    assert not golden_corr[f"{name}.lookup3"].any()   # far out-of-range coords -> zeros


def test_round_trip_is_not_identity():
    """integer coordinates can land one ulp below the integer -> floor flips."""
    for size in (28, 64, 256):
        p = np.arange(-4, size + 4, dtype=np.float32)
        u = taps.lookup_coord(p, size)
        assert (np.floor(u) != p).any()


def test_coords_grid_and_errors():
    g = corr_oracle.coords_grid(2, 3, 5)
    assert g.shape == (2, 2, 3, 5) and g[1, 0, 2, 4] == 4 and g[1, 1, 2, 4] == 2
    with pytest.raises(ValueError):
        corr_oracle.CorrBlock(torch.zeros(1, 4, 8, 8), torch.zeros(1, 4, 8, 8))       # too small for 4 levels
    with pytest.raises(ValueError):
        corr_oracle.CorrBlock(torch.zeros(1, 4, 16, 16), torch.zeros(1, 4, 16, 24))


@pytest.mark.skipif(not os.path.exists(REF_FILE), reason="reference tree not present (GPU box)")
def test_live_reference_agrees_with_oracle():
    spec = importlib.util.spec_from_file_location("ref_integration", REF_FILE)
    ref = importlib.util.module_from_spec(spec)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        spec.loader.exec_module(ref)
        g = torch.Generator().manual_seed(7)
        flow = torch.randn(2, 2, 40, 56, generator=g) * 4
        src = torch.rand(2, 3, 40, 56, generator=g)
        a = ref.SpatialTransformer((40, 56))(flow=flow.clone(), original=src)
    b = warp_oracle.SpatialTransformer((40, 56))(flow=flow, original=src)
    assert torch.equal(a, b)
    from torchvision.models.optical_flow.raft import CorrBlock as TV
    f1, f2 = torch.randn(1, 16, 16, 20, generator=g), torch.randn(1, 16, 16, 20, generator=g)
    tv = TV(num_levels=4, radius=4)
    tv.build_pyramid(f1, f2)
    mine = corr_oracle.CorrBlock(f1, f2)
    coords = corr_oracle.coords_grid(1, 16, 20) + torch.randn(1, 2, 16, 20, generator=g) * 2
    assert torch.equal(tv.index_pyramid(coords), mine(coords))


@pytest.mark.parametrize("name", ["u_6x5", "u_12x12", "u_9x33"])
def test_upsample_restatement_matches_reference_fixture(golden_upsample, name):
    """oracle/upsample_oracle.py against the outputs of the reference's own upsample_flow / upsample_seg
    (SegFlowGaussian.py:846-871, frozen by oracle/make_golden.py with one thread): same ATen ops; ATen's softmax / sum split their work by thread count,
    so agreement is to the last ulp or two (<= 1e-6 relative), not bitwise, when the suite runs multi-threaded."""
    from oracle import upsample_oracle
    g = golden_upsample
    mask = torch.from_numpy(g[f"{name}.mask"])
    assert rel_err(upsample_oracle.upsample_flow(torch.from_numpy(g[f"{name}.flow"]), mask).numpy(), g[f"{name}.up_flow"]) <= 1e-6
    assert rel_err(upsample_oracle.upsample_seg(torch.from_numpy(g[f"{name}.seg"]), mask).numpy(), g[f"{name}.up_seg"]) <= 1e-6
    # convexity: a constant field is reproduced (scaled by 8) away from the zero-padded border
    const = torch.full((1, 2, 6, 6), 0.25)
    up = upsample_oracle.upsample_flow(const, torch.randn(1, 576, 6, 6))
    assert torch.allclose(up[:, :, 8:-8, 8:-8], torch.full_like(up[:, :, 8:-8, 8:-8], 2.0), atol=1e-6)


@pytest.mark.parametrize("name", ["uw_24", "uw_17x40", "uw_oob_16", "uw_zero_12"])
def test_utils_warp_restatement_matches_reference_fixture(golden_utils_warp, name):
    """oracle/warp_oracle.utils_warp against the reference's own lib/utils.py `warp` (frozen by oracle/make_golden.py)."""
    g = golden_utils_warp
    got = warp_oracle.utils_warp(torch.from_numpy(g[f"{name}.x"]), torch.from_numpy(g[f"{name}.flo"]))
    assert np.array_equal(got.numpy(), g[f"{name}.out"])
    if name == "uw_zero_12":      # align_corners=False: zero flow is NOT the identity (half-pixel shift towards the centre)
        assert not np.allclose(got.numpy(), g[f"{name}.x"])


@pytest.mark.parametrize("name", ["ncc_noise_40", "ncc_smooth_48x37", "ncc_shift_64", "ncc_tiny_5x7", "ncc_same_33x70"])
def test_ncc_restatement_matches_reference_fixture(golden_ncc, name):
    """oracle/ncc_oracle.NCC against the reference's own lib/loss.py `NCC` (class compiled out of its file by
    oracle/make_golden.py).  Same ATen convolutions; a multi-threaded run may differ in the last ulps of the box sums,
    which the cancelling variance terms amplify, so the bound is the reference's own distance from float64."""
    from oracle import ncc_oracle
    g = golden_ncc
    I, J = torch.from_numpy(g[f"{name}.I"]), torch.from_numpy(g[f"{name}.J"])
    ref_map = torch.from_numpy(g[f"{name}.map"])                       # 1 - cc, [T,1,1,H,W]
    exact = 1 - ncc_oracle.cc64(I, J)[:, None]
    tol = max(1e-6, 4 * float((ref_map.double() - exact).abs().max()))
    got = ncc_oracle.NCC(reduction=None)(I[:, None], J[:, None])
    assert got.shape == ref_map.shape and float((got - ref_map).abs().max()) <= tol
    assert abs(float(ncc_oracle.NCC()(I, J)) - float(g[f"{name}.loss"])) <= 1e-6
    if f"{name}.gI" in g.files:
        Ig, Jg = I.clone().requires_grad_(True), J.clone().requires_grad_(True)
        ncc_oracle.NCC()(Ig, Jg).backward()
        scale = float(np.abs(g[f"{name}.gI"]).max())
        assert float((Ig.grad - torch.from_numpy(g[f"{name}.gI"])).abs().max()) <= 1e-3 * scale
        assert float((Jg.grad - torch.from_numpy(g[f"{name}.gJ"])).abs().max()) <= 1e-3 * scale
    with pytest.raises(ValueError):
        ncc_oracle.NCC(reduction=None)(I, J)                           # the reference's 4-D branch has no T, B, C
