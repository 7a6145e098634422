"""Guard words around every output buffer of the hot path ("Next round" #7: compute-sanitizer is closed on this
pool, so the NCC guard-word pattern is extended to the GEMM, the lookup and the warp).  Each kernel writes into a slice of a larger
buffer pre-filled with a sentinel; the sentinel words on both sides must survive, for the shapes that exercise tails (sizes off the
tile / vector widths) as well as the production shapes.  Values are checked elsewhere (test_gpu_corr / test_gpu_warp / ...).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
PAD = 64            # elements on either side: keeps fp32 slices 256-byte aligned
SENT = 777.0


def _cf():
    import cardioflow_b200 as cf
    return cf


def _guarded(n, dtype=torch.float32, sent=SENT):
    buf = torch.full((n + 2 * PAD,), sent, dtype=dtype, device=DEV)
    return buf, buf[PAD:PAD + n]


def _intact(buf, sent=SENT):
    return bool((buf[:PAD] == sent).all()) and bool((buf[-PAD:] == sent).all())


@pytest.mark.parametrize("precision", ["fp32", "auto", "bf16", "bf16x3"])
@pytest.mark.parametrize("hw,c,b", [((32, 32), 64, 2), ((28, 28), 40, 1), ((64, 64), 32, 1), ((20, 8), 24, 2), ((17, 21), 16, 1)])
def test_pyramid_build_stays_inside_its_buffer(precision, hw, c, b):
    cf = _cf()
    from cardioflow_b200 import _cabi
    h, w = hw
    if precision in ("bf16", "bf16x3") and (h % 4 or w % 8 or c % 8):
        pytest.skip("an explicit tensor-core precision is an error on shapes off the tcgen05 tiling (test_gpu_corr covers the error)")
    L = 4 if min(h, w) >= 16 else 2
    n = _cabi.lib().cf_corr_pyramid_bytes(b, h, w, L) // 4
    buf, view = _guarded(n)
    g = torch.Generator().manual_seed(h * w + c)
    f1, f2 = torch.randn(b, c, h, w, generator=g).to(DEV), torch.randn(b, c, h, w, generator=g).to(DEV)
    with torch.no_grad():
        blk = cf.CorrBlock(f1, f2, num_levels=L, precision=precision, pyramid_buffer=view)
        lv = blk.corr_pyramid
    torch.cuda.synchronize()
    assert _intact(buf), (precision, hw)
    assert lv[0].shape == (b * h * w, 1, h, w) and bool(torch.isfinite(lv[0]).all())


@pytest.mark.parametrize("hw,b,levels,radius", [((32, 32), 3, 4, 4), ((28, 28), 2, 3, 4), ((17, 21), 1, 2, 3), ((64, 64), 1, 4, 4), ((8, 8), 5, 2, 1)])
def test_lookup_stays_inside_its_buffer(hw, b, levels, radius):
    cf = _cf()
    h, w = hw
    g = torch.Generator().manual_seed(h + b)
    f1, f2 = torch.randn(b, 32, h, w, generator=g).to(DEV), torch.randn(b, 32, h, w, generator=g).to(DEV)
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    coords = (torch.stack([xs, ys]).float()[None].repeat(b, 1, 1, 1) + torch.randn(b, 2, h, w, generator=g) * 6).to(DEV)
    coords[0, :, 0, 0] = float("nan")                     # dead queries write zeros, nothing else
    coords[0, :, 0, 1] = 1e9
    K = levels * (2 * radius + 1) ** 2
    with torch.no_grad():
        blk = cf.CorrBlock(f1, f2, num_levels=levels, radius=radius)
        buf, view = _guarded(b * K * h * w)
        out = blk(coords, out=view.view(b, K, h, w))
    torch.cuda.synchronize()
    assert _intact(buf), hw
    assert bool((out[0, :, 0, 0] == 0).all()) and bool((out[0, :, 0, 1] == 0).all())


@pytest.mark.parametrize("mode", ["bilinear", "nearest"])
@pytest.mark.parametrize("size,chan,b", [((224, 224), 1, 3), ((224, 224), 5, 2), ((100, 60), 2, 2), ((37, 53), 3, 1), ((256, 256), 4, 1)])
def test_warp_stays_inside_its_buffer(mode, size, chan, b):
    cf = _cf()
    h, w = size
    g = torch.Generator().manual_seed(h + chan)
    src = torch.rand(b, chan, h, w, generator=g).to(DEV)
    flow = (torch.randn(b, 2, h, w, generator=g) * 4).to(DEV)
    flow[0, :, :3] = 500.0                                # a band pointing far outside the image
    buf, view = _guarded(b * chan * h * w)
    with torch.no_grad():
        out = cf.warp(src, flow, mode=mode, out=view.view(b, chan, h, w))
    torch.cuda.synchronize()
    assert _intact(buf), (mode, size)
    assert bool((out[0, :, :3] == 0).all())               # zero padding outside the image


@pytest.mark.parametrize("dtype", [torch.int64, torch.uint8])
@pytest.mark.parametrize("size,b", [((224, 224), 4), ((100, 60), 3), ((37, 53), 2)])
def test_label_warp_stays_inside_its_buffer(dtype, size, b):
    cf = _cf()
    h, w = size
    g = torch.Generator().manual_seed(h)
    lab = torch.randint(0, 4, (b, 1, h, w), generator=g).to(DEV)
    flow = (torch.randn(b, 2, h, w, generator=g) * 2).to(DEV)
    sent = 99
    buf, view = _guarded(b * h * w, dtype=dtype, sent=sent)
    cbuf = torch.full((b * 4 + 2 * PAD,), -5, dtype=torch.int64, device=DEV)
    out, counts = cf.warp_onehot_argmax(flow, lab, 4, out=view.view(b, 1, h, w), return_counts=True, counts_out=cbuf[PAD:PAD + b * 4].view(b, 4))
    torch.cuda.synchronize()
    assert _intact(buf, sent), (dtype, size)
    assert bool((cbuf[:PAD] == -5).all()) and bool((cbuf[-PAD:] == -5).all())
    assert int(counts.sum()) == b * h * w and int(out.max()) <= 3
