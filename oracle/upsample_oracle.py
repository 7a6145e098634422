"""CPU oracle (test infrastructure) for the convex 8x upsampling -- torch fp32 restatement of
/root/reference/nnunet/network_architecture/SegFlowGaussian.py:846-857 (``upsample_flow``) and :860-871
(``upsample_seg``).  Pinned by executable source: ``oracle/make_golden.py`` extracts those two methods from the
reference file itself and freezes their outputs in ``tests/golden/upsample_reference.npz``.
"""
import torch
import torch.nn.functional as F


def upsample_convex(field, mask, scale=8.0):
    """SegFlowGaussian.py:848-857: softmax over the 9 neighbours, unfold(8 * field, 3x3, pad 1), weighted sum,
    pixel-shuffle to [N, C, 8H, 8W].  The reference hard-codes the channel count (2 / 4); this takes it from the field."""
    N, C, H, W = field.shape
    m = torch.softmax(mask.view(N, 1, 9, 8, 8, H, W), dim=2)
    up = F.unfold(scale * field, [3, 3], padding=1).view(N, C, 9, 1, 1, H, W)
    up = torch.sum(m * up, dim=2).permute(0, 1, 4, 2, 5, 3)
    return up.reshape(N, C, 8 * H, 8 * W)


def upsample_flow(flow, mask):
    return upsample_convex(flow, mask)


def upsample_seg(seg, mask):
    return upsample_convex(seg, mask)


def algorithmic_bytes(B, C, H, W):
    """Compulsory traffic of one call: mask + field read once, output written once."""
    return 4 * B * H * W * (576 + C + 64 * C)
