"""GPU tests of the local correlation layer CorrVolume(radius, stride)(cur, prev) (csrc/local_corr.cu) through the C ABI.

PARITY UNPINNED: the reference's definition (nnunet/lib/raft.py) is missing; the oracle restates the
standard local cost volume with the call-site contract of /root/reference/nnunet/network_architecture/SegFlowGaussian.py:259,1376
([B,C,H,W] x [B,C,H,W] -> [B,(2r+1)^2,H,W]).  Tolerance: 1e-5 relative (fp32 sums over C in a different order than torch's).
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import local_corr_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("shape,radius,stride,norm,xm", [((2, 16, 24, 40), 4, 1, "sqrt", False), ((1, 32, 48, 48), 4, 4, "sqrt", False),
                                                        ((2, 8, 17, 33), 2, 2, "mean", True), ((1, 5, 9, 7), 1, 3, "none", False),
                                                        ((1, 64, 64, 64), 4, 2, "sqrt", True)])
def test_forward_and_backward_match_the_oracle(shape, radius, stride, norm, xm):
    import cardioflow_b200 as cf
    g = torch.Generator().manual_seed(sum(shape) + radius)
    cur, prev = torch.randn(*shape, generator=g), torch.randn(*shape, generator=g)
    wgt = torch.randn(shape[0], (2 * radius + 1) ** 2, shape[2], shape[3], generator=g)
    cr, pr = cur.clone().requires_grad_(), prev.clone().requires_grad_()
    want = local_corr_oracle.corr_volume(cr, pr, radius, stride, norm, xm)
    (want * wgt).sum().backward()
    cd, pd = cur.to(DEV).requires_grad_(), prev.to(DEV).requires_grad_()
    layer = cf.CorrVolume(radius=radius, stride=stride, normalization=norm, x_major=xm)
    got = layer(cd, pd)
    assert got.shape == want.shape
    assert rel_err(got.detach().cpu().numpy(), want.detach().numpy()) <= 1e-5
    (got * wgt.to(DEV)).sum().backward()
    assert rel_err(cd.grad.cpu().numpy(), cr.grad.numpy()) <= 1e-5
    assert rel_err(pd.grad.cpu().numpy(), pr.grad.numpy()) <= 1e-5


def test_call_site_contract_and_properties():
    """81 channels at the input's resolution (SegFlowGaussian.py:256-263); the centre channel is the plain per-pixel correlation;
    displacements that leave the image read zeros; a shifted copy peaks at the matching displacement."""
    import cardioflow_b200 as cf
    g = torch.Generator().manual_seed(3)
    cur = torch.randn(2, 12, 20, 28, generator=g).to(DEV)
    with torch.no_grad():
        out = cf.CorrVolume(radius=4, stride=1)(cur, cur)
        assert out.shape == (2, 81, 20, 28)
        centre = (cur * cur).sum(1) / np.sqrt(12.0)
        assert rel_err(out[:, 40].cpu().numpy(), centre.cpu().numpy()) <= 1e-5
        assert float(out[:, 0, :4, :].abs().max()) == 0.0 and float(out[:, 0, :, :4].abs().max()) == 0.0    # dy = dx = -4
        prev = torch.roll(cur, shifts=(2, -3), dims=(2, 3))          # prev[y + 2, x - 3] = cur[y, x]
        out = cf.CorrVolume(radius=4, stride=1)(cur, prev)
        best = out[:, :, 6:-6, 6:-6].mean(dim=(0, 2, 3)).argmax().item()
        assert best == (2 + 4) * 9 + (-3 + 4)
    with pytest.raises(ValueError):
        cf.CorrVolume(radius=5)
    with pytest.raises(RuntimeError):
        cf.CorrVolume()(cur.cpu(), cur.cpu())
