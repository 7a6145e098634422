"""NCC loss (lib/loss.py:15-76, the row directly downstream of the warp) on the GPU through the C ABI, against the reference
class's own outputs (tests/golden/ncc_reference.npz: loss, 1 - cc map, gradients) and the CPU oracle.

Tolerances.  The variance terms cancel catastrophically in fp32, so two correct fp32 evaluations differ by ~1e-6 * mean^2 / var
relative.  The bar is therefore stated through float64 (oracle/ncc_oracle.cc64): the kernel must be at least as close to
the exact value as 4x the reference's own worst distance from it (floor 1e-6 absolute on a map in [0.001, 0.999]), which also
bounds its distance from the reference map by the sum of the two.  Loss: 1e-6 absolute.  Gradients: 1e-3 of the largest
reference gradient (same bound the oracle is pinned with)."""
import numpy as np
import pytest
import torch

from conftest import NCC_CASES

pytestmark = pytest.mark.gpu


def _exact_bound(I, J, ref_map5):
    from oracle import ncc_oracle
    exact = 1 - ncc_oracle.cc64(I, J)[:, None]
    return exact, max(1e-6, 4 * float((ref_map5.double() - exact).abs().max()))


@pytest.mark.parametrize("name", NCC_CASES)
def test_golden_fixture(golden_ncc, name):
    import cardioflow_b200 as cf
    g = golden_ncc
    I, J = torch.from_numpy(g[f"{name}.I"]), torch.from_numpy(g[f"{name}.J"])
    ref_map = torch.from_numpy(g[f"{name}.map"])                                        # 1 - cc, [T,1,1,H,W]
    exact, tol = _exact_bound(I, J, ref_map)
    with torch.no_grad():
        got = cf.NCC(reduction=None)(I[:, None].cuda(), J[:, None].cuda()).cpu()
        loss = cf.NCC()(I.cuda(), J.cuda())
    assert got.shape == ref_map.shape and loss.dim() == 0
    assert float((got.double() - exact).abs().max()) <= tol
    assert float((got - ref_map).abs().max()) <= 2 * tol
    assert abs(float(loss) - float(g[f"{name}.loss"])) <= 1e-6
    if f"{name}.gI" in g.files:
        Ig, Jg = I.cuda().requires_grad_(True), J.cuda().requires_grad_(True)
        cf.NCC()(Ig, Jg).backward()
        scale = float(np.abs(g[f"{name}.gI"]).max())
        assert float((Ig.grad.cpu() - torch.from_numpy(g[f"{name}.gI"])).abs().max()) <= 1e-3 * scale
        assert float((Jg.grad.cpu() - torch.from_numpy(g[f"{name}.gJ"])).abs().max()) <= 1e-3 * scale


def _smooth(gen, n, h, w):
    a = torch.rand(n, 1, h // 4 + 2, w // 4 + 2, generator=gen)
    return torch.nn.functional.interpolate(a, size=(h, w), mode="bicubic", align_corners=True).clamp(0, 1).contiguous()


@pytest.mark.parametrize("shape", [(4, 224, 224), (2, 256, 256), (3, 57, 25), (2, 113, 100), (5, 3, 3), (1, 8, 300), (2, 9, 24)])
def test_shapes_against_oracle(shape):
    """Strip (24 columns) and band (56 rows) boundaries, images smaller than the window, ragged edges; cine-like smooth
    images (small local variance: the ill-conditioned regime) and noise."""
    import cardioflow_b200 as cf
    from oracle import ncc_oracle
    n, h, w = shape
    gen = torch.Generator().manual_seed(n * 1000 + h + w)
    for I, J in ((torch.rand(n, 1, h, w, generator=gen), torch.rand(n, 1, h, w, generator=gen)),
                 (lambda a: (a, (a + 0.03 * torch.randn(n, 1, h, w, generator=gen)).contiguous()))(_smooth(gen, n, h, w))):
        ref = 1 - ncc_oracle.cc_map(I, J)
        exact = 1 - ncc_oracle.cc64(I, J)
        tol = max(1e-6, 4 * float((ref.double() - exact).abs().max()))
        with torch.no_grad():
            got = cf.NCC(reduction=None)(I[None].cuda(), J[None].cuda()).cpu()[0]
            loss = cf.NCC()(I.cuda(), J.cuda())
        assert float((got.double() - exact).abs().max()) <= tol
        assert abs(float(loss) - float(exact.mean())) <= 1e-6        # exact = 1 - cc64


@pytest.mark.parametrize("shape", [(2, 64, 64), (3, 57, 25), (1, 120, 130)])
@pytest.mark.parametrize("mean", [True, False])
def test_backward_against_autograd_of_the_oracle(shape, mean):
    import cardioflow_b200 as cf
    from oracle import ncc_oracle
    n, h, w = shape
    gen = torch.Generator().manual_seed(77 + h)
    I = _smooth(gen, n, h, w)
    J = (I + 0.05 * torch.randn(n, 1, h, w, generator=gen)).contiguous()
    up = torch.randn(1, n, 1, h, w, generator=gen)
    Id, Jd = I.double().requires_grad_(True), J.double().requires_grad_(True)      # float64 autograd = the exact gradient
    Ig, Jg = I.cuda().requires_grad_(True), J.cuda().requires_grad_(True)
    if mean:
        ncc_oracle.NCC()(Id, Jd).backward()
        cf.NCC()(Ig, Jg).backward()
    else:
        (ncc_oracle.NCC(reduction=None)(Id[None], Jd[None]) * up.double()).sum().backward()
        (cf.NCC(reduction=None)(Ig[None], Jg[None]) * up.cuda()).sum().backward()
    for got, want in ((Ig.grad, Id.grad), (Jg.grad, Jd.grad)):
        scale = float(want.abs().max())
        assert float((got.cpu().double() - want).abs().max()) <= 1e-3 * scale
    # single-gradient launches (3 coefficient maps instead of 5) give the same numbers as the two-gradient launch
    for which in (0, 1):
        a, b = I.cuda().requires_grad_(which == 0), J.cuda().requires_grad_(which == 1)
        out = cf.NCC()(a, b) if mean else (cf.NCC(reduction=None)(a[None], b[None]) * up.cuda()).sum()
        out.backward()
        one, both = (a.grad, Ig.grad) if which == 0 else (b.grad, Jg.grad)
        assert (a.grad is None) != (b.grad is None)
        assert float((one - both).abs().max()) <= 1e-5 * float(both.abs().max())


def test_only_one_gradient_and_identical_images():
    import cardioflow_b200 as cf
    gen = torch.Generator().manual_seed(5)
    I = torch.rand(2, 1, 40, 40, generator=gen).cuda()
    J = I.clone().requires_grad_(True)
    loss = cf.NCC()(I, J)                       # cc = var^2 / (var^2 + eps) everywhere
    loss.backward()
    assert J.grad is not None and torch.isfinite(J.grad).all()
    assert 0.0 < float(loss.detach()) < 1.0
    flat = torch.full((1, 1, 20, 30), 0.25, device="cuda")
    with torch.no_grad():
        m = cf.NCC(reduction=None)(flat[None], flat[None])
    assert float(m[0, 0, 0, 10, 15]) == pytest.approx(0.999, abs=1e-7)      # zero variance inside: cc = 0 -> clamped to 0.001
    assert float(m[0, 0, 0, 0, 0]) < 0.01                                    # the zero padding gives the border variance


def test_errors_and_cabi_contract():
    import cardioflow_b200 as cf
    from cardioflow_b200 import _cabi
    a = torch.rand(2, 1, 16, 16, device="cuda")
    with pytest.raises(ValueError, match="5-D"):
        cf.NCC(reduction=None)(a, a)
    with pytest.raises(ValueError, match="single-channel"):
        cf.NCC()(torch.rand(2, 2, 16, 16, device="cuda"), torch.rand(2, 2, 16, 16, device="cuda"))
    with pytest.raises(ValueError, match="win 7"):
        cf.NCC(win=7)(a, a)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cf.NCC()(a.cpu(), a.cpu())
    lib = _cabi.lib()
    out = torch.empty(2, 16, 16, device="cuda")
    assert lib.cf_ncc(a.data_ptr(), a.data_ptr(), 2, 16, 16, 9, 1e-3, None, None, None, 0, None) == _cabi.CF_ERR_BAD_ARG
    assert lib.cf_ncc(a.data_ptr(), a.data_ptr(), 2, 16, 16, 9, 1e-3, None, out.data_ptr(), None, 0, None) == _cabi.CF_ERR_BAD_ARG
    assert b"workspace" in lib.cf_last_error()
    assert lib.cf_ncc_workspace_bytes(2, 16, 16, 1) == 5 * 2 * 16 * 16 * 4


@pytest.mark.parametrize("shape", [(2, 40, 44), (3, 17, 23), (1, 5, 7), (2, 33, 128)])
def test_outputs_stay_inside_their_buffers(shape):
    """Guard words around every output (map, partial sums, coefficient maps, gradients) stay untouched: vector path
    (W % 4 == 0) and element-wise path, strips and bands that end inside the image."""
    from cardioflow_b200 import _cabi
    lib = _cabi.lib()
    n, h, w = shape
    npx = n * h * w
    gen = torch.Generator().manual_seed(w)
    I, J = torch.rand(n, 1, h, w, generator=gen).cuda(), torch.rand(n, 1, h, w, generator=gen).cuda()
    G = 64                                               # guard floats on both sides (keeps 16-byte alignment)

    def guarded(count):
        buf = torch.full((count + 2 * G,), 12345.0, device="cuda")
        return buf, buf[G:G + count]

    def intact(buf, count):
        return bool((buf[:G] == 12345.0).all()) and bool((buf[G + count:] == 12345.0).all())

    nws = lib.cf_ncc_workspace_bytes(n, h, w, 0) // 4
    mbuf, m = guarded(npx)
    pbuf, part = guarded(nws)
    sbuf, csum = guarded(4)
    _cabi.check(lib.cf_ncc(I.data_ptr(), J.data_ptr(), n, h, w, 9, 1e-3, m.data_ptr(), csum.data_ptr(), part.data_ptr(), nws * 4,
                           _cabi.stream()))
    torch.cuda.synchronize()
    assert intact(mbuf, npx) and intact(pbuf, nws) and intact(sbuf, 4)
    assert bool((sbuf[G + 1:G + 4] == 12345.0).all()) and not bool((m == 12345.0).any())
    wbuf, ws = guarded(5 * npx)
    ibuf, gi = guarded(npx)
    jbuf, gj = guarded(npx)
    up = torch.randn(npx, generator=gen).cuda()
    for gI, gJ in ((gi, gj), (gi, None), (None, gj)):
        _cabi.check(lib.cf_ncc_bwd(I.data_ptr(), J.data_ptr(), n, h, w, 9, 1e-3, up.data_ptr(), 1, -1.0,
                                   gI.data_ptr() if gI is not None else None, gJ.data_ptr() if gJ is not None else None,
                                   ws.data_ptr(), 5 * npx * 4, _cabi.stream()))
        torch.cuda.synchronize()
        assert intact(wbuf, 5 * npx) and intact(ibuf, npx) and intact(jbuf, npx)


def test_forward_is_capturable_in_a_cuda_graph():
    """cf_ncc only enqueues on the current stream (no hidden synchronisation): a captured forward replays on new data."""
    import cardioflow_b200 as cf
    gen = torch.Generator().manual_seed(3)
    I, J = torch.rand(3, 1, 64, 72, generator=gen).cuda(), torch.rand(3, 1, 64, 72, generator=gen).cuda()
    loss_fn = cf.NCC()
    with torch.no_grad():
        eager = loss_fn(I, J).clone()
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            loss_fn(I, J)                                   # warm the allocator on the capture stream
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = loss_fn(I, J)
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager)
        I.copy_(torch.rand(3, 1, 64, 72, generator=gen))    # new data in the captured buffers
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, loss_fn(I, J)) and not torch.equal(out, eager)


def test_full_size_properties_cfg3_batch():
    """300 x 224^2 (the BASELINE cfg 3 batch, the frames the warp produces): size-independent properties of the loss."""
    import cardioflow_b200 as cf
    g = torch.Generator(device="cuda").manual_seed(1003)
    n, h, w = 300, 224, 224
    I = torch.rand(n, 1, h, w, device="cuda", generator=g)
    J = (I + 0.1 * torch.randn(n, 1, h, w, device="cuda", generator=g)).contiguous()
    with torch.no_grad():
        ncc = cf.NCC()
        lij, lji = ncc(I, J), ncc(J, I)
        assert abs(float(lij) - float(lji)) <= 1e-6                              # symmetric in its arguments
        assert abs(float(ncc(2.5 * I + 0.0, J)) - float(lij)) <= 2e-4           # cc is invariant to gain up to the eps term
        m = cf.NCC(reduction=None)(I[None], J[None])                             # [1, n, 1, h, w]
        assert m.shape == (1, n, 1, h, w) and float(m.min()) >= 0.001 - 1e-7 and float(m.max()) <= 0.999 + 1e-7
        assert abs(float(m.double().mean()) - float(lij)) <= 1e-6               # the mean path equals the mean of the map
        parts = torch.stack([ncc(I[k:k + 60], J[k:k + 60]) for k in range(0, n, 60)])
        assert abs(float(parts.double().mean()) - float(lij)) <= 1e-6           # images are independent: batch splits agree
        same = ncc(I, I)
        assert float(same) < float(lij) and float(same) < 0.05                   # identical images: cc ~ var^2 / (var^2 + eps)
        assert torch.equal(ncc(I, J), lij)                                       # deterministic reduction
