"""CPU checks of the local-correlation oracle (PARITY UNPINNED: oracle/local_corr_oracle.py restates the standard local cost
volume under the call-site contract of SegFlowGaussian.py:259,1376; there is no reference source to pin it to)."""
import math

import torch

from oracle import local_corr_oracle


def test_shape_centre_channel_padding_and_orders():
    g = torch.Generator().manual_seed(1)
    cur, prev = torch.randn(2, 6, 10, 12, generator=g), torch.randn(2, 6, 10, 12, generator=g)
    out = local_corr_oracle.corr_volume(cur, prev, radius=2, stride=2)
    assert out.shape == (2, 25, 10, 12)
    assert torch.allclose(out[:, 12], (cur * prev).sum(1) / math.sqrt(6), atol=1e-6)
    assert float(out[:, 0, :4].abs().max()) == 0.0              # dy = -2 * 2 rows above the image: zero padding
    # brute force at one pixel / displacement
    b, y, x, dy, dx = 1, 5, 6, 1, -2
    want = sum(float(cur[b, c, y, x] * prev[b, c, y + 2 * dy, x + 2 * dx]) for c in range(6)) / math.sqrt(6)
    assert abs(float(out[b, (dy + 2) * 5 + (dx + 2), y, x]) - want) < 1e-5
    xm = local_corr_oracle.corr_volume(cur, prev, radius=2, stride=2, x_major=True)
    assert torch.equal(xm[:, (dx + 2) * 5 + (dy + 2)], out[:, (dy + 2) * 5 + (dx + 2)])
    mean = local_corr_oracle.corr_volume(cur, prev, radius=2, stride=2, normalization="mean")
    assert torch.allclose(mean * 6 / math.sqrt(6), out, atol=1e-6)
