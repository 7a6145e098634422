"""GPU parity: CorrBlock (pyramid build + lookup) through the C ABI / module API against the oracle.

Tolerances (BASELINE.json north_star): tap indices / coordinates bit-exact; fp32 lookup <= 1e-5 relative
(max|diff| / max|ref|) -- and bit-exact against the C oracle on the same pyramid; pooling bit-exact; fp32
correlation <= 1e-5 relative (summation order); tensor-core (bf16 in, fp32 accumulate) correlation <= 1e-2.
"""
import numpy as np
import pytest
import torch

import cardioflow_b200 as cf
from conftest import CORR_CASES, rel_err
from oracle import c_oracle, corr_oracle, taps

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _gpu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _coord_sets(g, b, h, w):
    grid = corr_oracle.coords_grid(b, h, w)
    border = grid.clone()
    border[:, 0] = torch.where(grid[:, 0] < w / 2, -0.5, w - 0.5)
    return {
        "grid": grid,                                                       # iteration 0: exact integers
        "noise": grid + torch.randn(b, 2, h, w, generator=g) * 2.0,
        "border": border,
        "wide": grid + torch.randn(b, 2, h, w, generator=g) * (w / 2),      # many taps leave the image
        "far": grid + 1000.0,
        "negfar": grid - 3000.0,
    }


@pytest.mark.parametrize("name", CORR_CASES)
def test_golden_fixture(golden_corr, name):
    f1, f2 = golden_corr[f"{name}.f1"], golden_corr[f"{name}.f2"]
    blk = cf.CorrBlock(_gpu(f1), _gpu(f2), radius=4, precision="fp32")
    assert blk.used_precision == "fp32"
    ref_l0 = corr_oracle.CorrBlock(torch.from_numpy(f1), torch.from_numpy(f2)).corr_pyramid[0].numpy()
    assert rel_err(blk.corr_pyramid[0].cpu().numpy(), ref_l0) <= 1e-5
    for lvl in (1, 2, 3):
        got = blk.corr_pyramid[lvl].cpu().numpy()
        assert got.shape == golden_corr[f"{name}.pyr{lvl}"].shape
        assert rel_err(got, golden_corr[f"{name}.pyr{lvl}"]) <= 1e-5
        # pooling itself is bit-exact: level l+1 == ATen-order pooling of OUR level l
        assert np.array_equal(got[:, 0], taps.avg_pool2x2(blk.corr_pyramid[lvl - 1].cpu().numpy()[:, 0]))
    for k in range(4):
        out = blk(_gpu(golden_corr[f"{name}.coords{k}"])).cpu().numpy()
        ref = golden_corr[f"{name}.lookup{k}"]
        assert out.shape == ref.shape
        assert np.abs(out - ref).max() <= 1e-5 * max(np.abs(ref).max(), 1.0)


@pytest.mark.parametrize("hw", [(16, 16), (24, 24), (28, 28), (32, 32), (16, 40), (64, 64), (17, 21)])
def test_lookup_bit_exact_on_oracle_pyramid(hw):
    """Lookup isolated from GEMM precision: same fp32 pyramid on both sides."""
    h, w = hw
    b, c = (1, 8) if h * w > 2000 else (2, 8)
    g = torch.Generator().manual_seed(h * 100 + w)
    f1, f2 = torch.randn(b, c, h, w, generator=g), torch.randn(b, c, h, w, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2)
    pyr = [p.numpy() for p in ref.corr_pyramid]
    blk = cf.CorrBlock.from_pyramid([p.to(DEV) for p in ref.corr_pyramid], radius=4)
    for kind, coords in _coord_sets(g, b, h, w).items():
        out = blk(coords.to(DEV)).cpu().numpy()
        want_c, ctaps, cidx = c_oracle.lookup(pyr, coords.numpy(), want_taps=True)
        assert np.array_equal(out, want_c), kind                               # bit-exact vs the C oracle
        want_t = ref(coords).numpy()
        assert np.abs(out - want_t).max() <= 1e-5 * max(np.abs(want_t).max(), 1.0), kind
        t, idx = cf.lookup_taps(coords.to(DEV))
        assert np.array_equal(t.cpu().numpy(), ctaps) and np.array_equal(idx.cpu().numpy(), cidx), kind
        if kind in ("far", "negfar"):
            assert not out.any()


@pytest.mark.parametrize("radius,levels", [(1, 4), (2, 3), (3, 2), (4, 1), (4, 5)])
def test_other_radius_and_levels(radius, levels):
    g = torch.Generator().manual_seed(radius * 10 + levels)
    h = w = 32
    f1, f2 = torch.randn(1, 8, h, w, generator=g), torch.randn(1, 8, h, w, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2, num_levels=levels, radius=radius)
    blk = cf.CorrBlock.from_pyramid([p.to(DEV) for p in ref.corr_pyramid], radius=radius)
    coords = corr_oracle.coords_grid(1, h, w) + torch.randn(1, 2, h, w, generator=g) * 3
    out = blk(coords.to(DEV)).cpu().numpy()
    assert out.shape == (1, levels * (2 * radius + 1) ** 2, h, w)
    assert np.array_equal(out, c_oracle.lookup([p.numpy() for p in ref.corr_pyramid], coords.numpy(), radius=radius))
    built = cf.CorrBlock(f1.to(DEV), f2.to(DEV), num_levels=levels, radius=radius, precision="fp32")
    assert len(built.corr_pyramid) == levels
    assert np.abs(built(coords.to(DEV)).cpu().numpy() - ref(coords).numpy()).max() <= 1e-5 * 4


def test_window_orientation_anisotropic():
    """Channel i*9+j must sample (x + i - r, y + j - r): a ramp in x only tells the two axes apart."""
    h = w = 16
    lvl0 = torch.arange(w, dtype=torch.float32).view(1, 1, 1, w).repeat(h * w, 1, h, 1)
    pyr = [lvl0]
    for _ in range(3):
        pyr.append(torch.nn.functional.avg_pool2d(pyr[-1], 2, stride=2))
    blk = cf.CorrBlock.from_pyramid([p.to(DEV) for p in pyr])
    coords = torch.full((1, 2, h, w), 8.0)
    out = blk(coords.to(DEV)).cpu().view(4, 9, 9, h, w)
    lvl = out[0, :, :, 5, 5]
    assert torch.allclose(lvl[:, 0], torch.arange(4.0, 13.0), atol=1e-5)   # varies along i (x offset)
    assert torch.allclose(lvl[3, :], torch.full((9,), 7.0), atol=1e-5)     # constant along j (y offset)


@pytest.mark.parametrize("hw,c", [((32, 32), 256), ((28, 28), 256), ((24, 24), 64), ((16, 16), 256), ((64, 64), 128)])
def test_correlation_precisions(hw, c):
    h, w = hw
    g = torch.Generator().manual_seed(c + h)
    f1, f2 = torch.randn(2, c, h, w, generator=g), torch.randn(2, c, h, w, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2)
    exact = cf.CorrBlock(f1.to(DEV), f2.to(DEV), precision="fp32")
    for lvl in range(4):
        assert rel_err(exact.corr_pyramid[lvl].cpu().numpy(), ref.corr_pyramid[lvl].numpy()) <= 1e-5
    auto = cf.CorrBlock(f1.to(DEV), f2.to(DEV))      # tensor cores when the shape allows
    for lvl in range(4):
        assert rel_err(auto.corr_pyramid[lvl].cpu().numpy(), ref.corr_pyramid[lvl].numpy()) <= 1e-2
        if lvl:
            assert np.array_equal(auto.corr_pyramid[lvl].cpu().numpy()[:, 0],
                                  taps.avg_pool2x2(auto.corr_pyramid[lvl - 1].cpu().numpy()[:, 0]))


def test_twelve_iteration_trajectory_cfg1():
    """BASELINE cfg 1: 256-ch 28x28 fmaps, 4 levels, radius 4, 12 lookups along a synthetic trajectory."""
    g = torch.Generator().manual_seed(1001)
    f1, f2 = torch.randn(1, 256, 28, 28, generator=g), torch.randn(1, 256, 28, 28, generator=g)
    ref = corr_oracle.CorrBlock(f1, f2)
    blk = cf.CorrBlock(f1.to(DEV), f2.to(DEV), radius=4, precision="fp32")
    coords = corr_oracle.coords_grid(1, 28, 28)
    gpu_grid = cf.coords_grid(1, 28, 28, device=DEV)
    assert torch.equal(gpu_grid.cpu(), coords)
    for k in range(12):
        out = blk(coords.to(DEV)).cpu().numpy()
        want = ref(coords).numpy()
        assert np.abs(out - want).max() <= 2e-5 * np.abs(want).max(), k      # fp32 GEMM order + lookup
        coords = (coords + torch.randn(1, 2, 28, 28, generator=g) * 0.5 * (k + 1))
        coords[:, 0].clamp_(-4, 32)
        coords[:, 1].clamp_(-4, 32)


def test_torchvision_call_site_adaptor():
    """The reference's in-tree third-party call site (nnMTLTrainerV2Raft.py:97-118): raft_large's corr_block
    replaced by an adaptor over the new kernels must reproduce the 12-iteration flow predictions."""
    from torchvision.models.optical_flow import raft_large
    torch.manual_seed(0)
    model = raft_large(weights=None).to(DEV).eval()
    img1, img2 = torch.rand(1, 3, 128, 128, device=DEV) * 2 - 1, torch.rand(1, 3, 128, 128, device=DEV) * 2 - 1
    with torch.no_grad():
        want = model(img1, img2, num_flow_updates=12)

        class Adaptor(torch.nn.Module):
            out_channels = model.corr_block.out_channels
            num_levels, radius = 4, 4

            def build_pyramid(self, fmap1, fmap2):
                self.blk = cf.CorrBlock(fmap1, fmap2, num_levels=4, radius=4, precision="fp32")

            def index_pyramid(self, centroids_coords):
                return self.blk(centroids_coords)

        model.corr_block = Adaptor()
        got = model(img1, img2, num_flow_updates=12)
    for a, b in zip(got, want):
        assert float((a - b).abs().max()) <= 1e-3 * max(float(b.abs().max()), 1.0)


def test_full_size_properties_cfg4():
    """BASELINE cfg 4 size (64x64 fmaps, 16.7M-entry volume per pair): properties + a sampled oracle check."""
    g = torch.Generator(device=DEV).manual_seed(1004)
    b, c, h, w = 4, 256, 64, 64
    f1 = torch.randn(b, c, h, w, device=DEV, generator=g)
    f2 = torch.randn(b, c, h, w, device=DEV, generator=g)
    blk = cf.CorrBlock(f1, f2, precision="fp32")
    grid = cf.coords_grid(b, h, w, device=DEV)
    out = blk(grid).view(b, 4, 9, 9, h * w)
    # centre tap of level 0 at integer coords == corr[p, p] (diagonal of the volume), up to the round trip
    diag = torch.einsum("bcp,bcp->bp", f1.view(b, c, -1), f2.view(b, c, -1)) / 16.0
    assert float((out[:, 0, 4, 4] - diag).abs().max()) <= 1e-4 * float(diag.abs().max())
    assert not bool(blk(grid + 5000.0).any())
    # linearity in fmap2: corr(f1, a*f2 + f2') lookups == a*lookup + lookup'
    f2b = torch.randn(b, c, h, w, device=DEV, generator=g)
    coords = grid + torch.randn(b, 2, h, w, device=DEV, generator=g) * 2
    lhs = cf.CorrBlock(f1, 0.5 * f2 + f2b, precision="fp32")(coords)
    rhs = 0.5 * blk(coords) + cf.CorrBlock(f1, f2b, precision="fp32")(coords)
    assert float((lhs - rhs).abs().max()) <= 1e-4 * float(rhs.abs().max())
    ref = corr_oracle.CorrBlock(f1[:1].cpu(), f2[:1].cpu())
    want = ref(coords[:1].cpu()).numpy()
    assert np.abs(blk(coords)[:1].cpu().numpy() - want).max() <= 2e-5 * np.abs(want).max()


def test_full_batch_cfg4_whole_batch():
    """BASELINE cfg 4 at full size, the WHOLE batch (64 pairs, 256 channels, 64x64 fmaps; 5.7 GB pyramid): (1) level 0 of the
    fp32-grade tensor-core build (bf16x3) within 1e-5 of a CPU fp32 matmul for every pair; (2) pooled levels bit-identical to
    avg_pool2d; (3) lookups of all 64 x 4096 queries BIT-EXACT against the C oracle run on the pyramid the GPU built (tiled
    levels 0 and 1 un-tiled through corr_pyramid)."""
    b, c, h, w = 64, 256, 64, 64
    g = torch.Generator().manual_seed(1004)
    f1, f2 = torch.randn(b, c, h, w, generator=g), torch.randn(b, c, h, w, generator=g)
    grid = corr_oracle.coords_grid(b, h, w)
    coords = grid + torch.randn(b, 2, h, w, generator=g) * 2.0
    with torch.no_grad():
        blk = cf.CorrBlock(f1.to(DEV), f2.to(DEV))
        assert blk.used_precision == "bf16x3" and blk._tiled_mask == 3
        out = blk(coords.to(DEV)).cpu().numpy()
        pyr = [p.cpu() for p in blk.corr_pyramid]
    del blk
    torch.cuda.empty_cache()
    worst = 0.0
    for i in range(b):                                   # CPU fp32 reference, pair by pair (bounded memory)
        want = torch.matmul(f1[i].view(c, -1).t(), f2[i].view(c, -1)) / 16.0
        got = pyr[0][i * h * w:(i + 1) * h * w, 0].reshape(h * w, h * w)
        worst = max(worst, float((got - want).abs().max()) / float(want.abs().max()))
    assert worst <= 1e-5, worst
    for l in range(1, 4):
        assert torch.equal(pyr[l], torch.nn.functional.avg_pool2d(pyr[l - 1], 2, stride=2)), l
    want_c = c_oracle.lookup([p.numpy() for p in pyr], coords.numpy())
    assert np.array_equal(out, want_c)


@pytest.mark.parametrize("hw,c,want_mask", [((64, 64), 128, 3), ((32, 64), 64, 3), ((64, 32), 256, 1), ((40, 96), 32, 3), ((64, 128), 32, 3),
                                            ((36, 72), 40, 1), ((16, 16), 64, 0)])
def test_tiled_levels_generic_epilogue(monkeypatch, hw, c, want_mask):
    """Any width (version 2 of the build stores from registers, tiling is address arithmetic): level 0 is tiled when its rows
    are >= 128 B, level 1 likewise (W >= 64, H % 8 == 0); pyramids and lookups must be bit-identical to the row-major build."""
    h, w = hw
    g = torch.Generator().manual_seed(h * w + c)
    f1, f2 = torch.randn(2, c, h, w, generator=g).to(DEV), torch.randn(2, c, h, w, generator=g).to(DEV)
    monkeypatch.setenv("CF_GEMM_TILED", "3")      # level-1 tiling is off by default (it costs the build more than the lookup gains)
    monkeypatch.setenv("CARDIOFLOW_TILED", "0")
    plain = cf.CorrBlock(f1, f2, precision="bf16")
    monkeypatch.setenv("CARDIOFLOW_TILED", "1")
    tiled = cf.CorrBlock(f1, f2, precision="bf16")
    assert plain._tiled_mask == 0 and tiled._tiled_mask == want_mask
    for a, b in zip(plain.corr_pyramid, tiled.corr_pyramid):
        assert a.shape == b.shape and torch.equal(a, b)
    for s in (0.0, 1.3, 50.0):
        coords = cf.coords_grid(2, h, w, device=DEV) + s * torch.randn(2, 2, h, w, generator=g).to(DEV)
        assert torch.equal(plain(coords), tiled(coords))


def test_tiled_level0_is_invisible(monkeypatch):
    """The tensor-core builder stores level 0 as 4x4 tiles when no gradient is needed (cf_corr_build_layout /
    cf_corr_lookup_layout).  Nothing observable may change: pyramid values and every lookup are bit-identical to the
    row-major build, `corr_pyramid` still has torchvision's shapes, and autograd keeps the row-major layout."""
    g = torch.Generator().manual_seed(77)
    f1, f2 = torch.randn(3, 256, 32, 32, generator=g).to(DEV), torch.randn(3, 256, 32, 32, generator=g).to(DEV)
    coords = [(cf.coords_grid(3, 32, 32, device=DEV) + s * torch.randn(3, 2, 32, 32, generator=g).to(DEV)) for s in (0.0, 0.7, 3.0, 40.0)]
    monkeypatch.setenv("CARDIOFLOW_TILED", "0")
    plain = cf.CorrBlock(f1, f2, precision="bf16")
    assert plain._tiled_mask == 0
    monkeypatch.setenv("CARDIOFLOW_TILED", "1")
    tiled = cf.CorrBlock(f1, f2, precision="bf16")
    assert tiled._tiled_mask == 1, "W = 32 tensor-core build is expected to tile level 0"
    for a, b in zip(plain.corr_pyramid, tiled.corr_pyramid):
        assert a.shape == b.shape and torch.equal(a, b)
    for c in coords:
        assert torch.equal(plain(c), tiled(c))
    # raw storage of level 0 really is tiled: element (y, x) of image 5 sits at tile-major position
    raw = tiled._pyramid[:3 * 1024 * 1024].view(3 * 1024, 8, 8, 4, 4)
    assert float(raw[5, 2, 3, 1, 2]) == float(plain.corr_pyramid[0][5, 0, 2 * 4 + 1, 3 * 4 + 2])
    # with autograd the layout stays row-major and gradients flow
    fa, fb = f1.clone().requires_grad_(), f2.clone().requires_grad_()
    blk = cf.CorrBlock(fa, fb, precision="bf16")
    assert blk._tiled_mask == 0
    blk(coords[1]).mean().backward()
    assert fa.grad is not None and torch.isfinite(fa.grad).all()


@pytest.mark.parametrize("n_lanes,graph", [(2, False), (3, False), (2, True)])
def test_lanes_are_bit_identical_to_one_chain(n_lanes, graph):
    """Sub-batches of the pairs on parallel streams (cardioflow_b200.Lanes) give the bytes of the single full-batch chain:
    every kernel of the path treats batch items independently."""
    g = torch.Generator().manual_seed(77)
    P, C, H, W, IH = 7, 128, 32, 32, 96
    f1, f2 = (torch.randn(P, C, H, W, generator=g).to(DEV) for _ in range(2))
    coords = [(corr_oracle.coords_grid(P, H, W) + torch.randn(P, 2, H, W, generator=g) * 2.0).to(DEV) for _ in range(3)]
    img, flow = torch.rand(P, 1, IH, IH, generator=g).to(DEV), (torch.randn(P, 2, IH, IH, generator=g) * 3).to(DEV)

    def chain(a, b, outs, warped):
        blk = cf.CorrBlock(f1[a:b], f2[a:b], radius=4)
        for k, c in enumerate(coords):
            blk(c[a:b], out=outs[k, a:b])
        cf.warp(img[a:b], flow[a:b], out=warped[a:b])
        return blk

    def buffers():
        return torch.full((3, P, 324, H, W), float("nan"), device=DEV), torch.full((P, 1, IH, IH), float("nan"), device=DEV)

    with torch.no_grad():
        ref_o, ref_w = buffers()
        chain(0, P, ref_o, ref_w)
        out_o, out_w = buffers()
        lanes = cf.Lanes(n_lanes, DEV)
        torch.cuda.synchronize()
        if graph:
            lanes.run(P, lambda j, a, b: chain(a, b, out_o, out_w))      # warm-up outside capture
            torch.cuda.synchronize()
            out_o.fill_(float("nan"))
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph):
                keep = lanes.run(P, lambda j, a, b: chain(a, b, out_o, out_w))
            gph.replay()
        else:
            keep = lanes.run(P, lambda j, a, b: chain(a, b, out_o, out_w))
        torch.cuda.synchronize()
    assert len(keep) == n_lanes
    assert torch.equal(out_o, ref_o) and torch.equal(out_w, ref_w)


def test_full_batch_cfg2_production_path():
    """BASELINE cfg 2 exactly as bench.py runs it: 29 pairs, 256 channels, 32x32 fmaps, bf16 tensor-core build (level 0 stored
    as 4x4 tiles), two lanes.  (1) every lane's lookups are BIT-EXACT against the C oracle run on the pyramid the GPU built
    (the lookup adds no error of its own at full batch); (2) the pyramid is within the stated 1e-2 of the fp32 oracle;
    (3) pooling levels equal avg_pool2d of the level below, bit for bit."""
    b, c, h, w = 29, 256, 32, 32
    g = torch.Generator().manual_seed(1002)
    base = torch.randn(c, h, w, generator=g)
    frames = [base]
    for t in range(b):
        frames.append(torch.roll(frames[-1], shifts=(t % 3 - 1, 1), dims=(1, 2)) + 0.1 * torch.randn(c, h, w, generator=g))
    frames = torch.stack(frames)
    f1, f2 = frames[:-1].contiguous(), frames[1:].contiguous()
    grid = corr_oracle.coords_grid(b, h, w)
    traj = [grid, grid + torch.randn(b, 2, h, w, generator=g) * 1.2, grid + torch.randn(b, 2, h, w, generator=g) * 1.7]
    f1d, f2d = f1.to(DEV), f2.to(DEV)
    outs = [torch.empty(b, 324, h, w, device=DEV) for _ in traj]
    lanes = cf.Lanes(2, DEV)
    blocks = {}

    def chain(j, lo, hi):
        blk = cf.CorrBlock(f1d[lo:hi], f2d[lo:hi], num_levels=4, radius=4, precision="bf16")
        for k, co in enumerate(traj):
            blk(co[lo:hi].to(DEV), out=outs[k][lo:hi])
        blocks[j] = (lo, hi, blk)
    with torch.no_grad():
        lanes.run(b, chain)
    torch.cuda.synchronize()
    ref = corr_oracle.CorrBlock(f1, f2)                                        # fp32 CPU oracle, whole batch
    for lo, hi, blk in blocks.values():
        pyr = [p.cpu() for p in blk.corr_pyramid]                              # un-tiled views of what the GPU built
        for l in range(4):
            want = ref.corr_pyramid[l].view(b, h * w, *ref.corr_pyramid[l].shape[1:])[lo:hi].reshape(pyr[l].shape)
            assert rel_err(pyr[l].numpy(), want.numpy()) <= 1e-2, l
            if l:
                assert torch.equal(pyr[l], torch.nn.functional.avg_pool2d(pyr[l - 1], 2, stride=2)), l
        for k, co in enumerate(traj):
            want_c = c_oracle.lookup([p.numpy() for p in pyr], co[lo:hi].contiguous().numpy())
            assert np.array_equal(outs[k][lo:hi].cpu().numpy(), want_c), (lo, hi, k)
    assert rel_err(outs[2].cpu().numpy(), ref(traj[2]).numpy()) <= 1e-2
