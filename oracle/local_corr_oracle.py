"""Oracle (test infrastructure) for the local correlation layer ``CorrVolume(radius, stride)(cur, prev)``.

PARITY UNPINNED: the reference's own definition (nnunet/lib/raft.py, imported at
/root/reference/nnunet/network_architecture/SegFlowGaussian.py:41, used :259, :1376) was never committed and nothing the
reference calls defines it.  This is a plain torch-CPU restatement of the STANDARD local cost volume with the call-site's
shape contract ([B,C,H,W] x [B,C,H,W] -> [B,(2r+1)^2,H,W]); it pins the CUDA kernel to that definition, not to the reference.
"""
import math

import torch
import torch.nn.functional as F


def corr_volume(cur, prev, radius=4, stride=1, normalization="sqrt", x_major=False):
    B, C, H, W = cur.shape
    pad = radius * stride
    pp = F.pad(prev, (pad, pad, pad, pad))
    side = 2 * radius + 1
    out = cur.new_zeros(B, side * side, H, W)
    for i in range(side):            # dy index
        for j in range(side):        # dx index
            sh = pp[:, :, i * stride:i * stride + H, j * stride:j * stride + W]
            ch = j * side + i if x_major else i * side + j
            out[:, ch] = (cur * sh).sum(1)
    scale = {"sqrt": 1.0 / math.sqrt(C), "mean": 1.0 / C, "none": 1.0}[normalization]
    return out * scale
