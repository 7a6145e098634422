"""``CorrVolume(radius, stride)(cur, prev)`` -- the local correlation layer of the reference's skip connections.

Call-site contract (/root/reference/nnunet/network_architecture/SegFlowGaussian.py): ``from nnunet.lib.raft import CorrVolume``
(:41), ``CorrVolume(radius=self.radius[idx], stride=self.stride[idx])`` (:259; config ``stride: [4, 2, 1, 1]``),
``corr = layer(skip_co_current[s], previous_skip_co[s])`` (:1376), fed to a conv block with ``(2r+1)**2`` input channels.

PARITY UNPINNED: ``nnunet/lib/raft.py`` was never committed and the reference calls no third-party twin, so the semantics
below are the standard local cost volume of that shape, not a restatement of a file we could read:

    out[b, (dy+r)*(2r+1) + (dx+r), y, x] = scale * sum_c cur[b,c,y,x] * prev[b,c, y + dy*stride, x + dx*stride]   (zeros outside)

``normalization`` ("sqrt": 1/sqrt(C) like RAFT's CorrBlock, default; "mean": 1/C like PWC-Net; "none") and ``x_major`` (channel
order (dx+r)*(2r+1) + (dy+r), the order of RAFT's lookup) let a user holding the missing file match it.
"""
import math

import torch
import torch.nn as nn

from . import _cabi


class _LocalCorr(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cur, prev, radius, stride, scale, x_major):
        a, b = _cabi.f32c(cur), _cabi.f32c(prev)
        B, C, H, W = a.shape
        out = torch.empty(B, (2 * radius + 1) ** 2, H, W, dtype=torch.float32, device=a.device)
        if out.numel():
            with torch.cuda.device(a.device):
                _cabi.check(_cabi.lib().cf_local_corr(a.data_ptr(), b.data_ptr(), B, C, H, W, radius, stride, scale,
                                                      1 if x_major else 0, out.data_ptr(), _cabi.stream()), "cf_local_corr")
        ctx.save_for_backward(a, b)
        ctx.args = (radius, stride, scale, x_major)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        a, b = ctx.saved_tensors
        radius, stride, scale, x_major = ctx.args
        B, C, H, W = a.shape
        go = grad_out.contiguous().float()
        need_a, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        ga = torch.empty_like(a) if need_a else None
        gb = torch.empty_like(b) if need_b else None
        if (need_a or need_b) and a.numel():
            with torch.cuda.device(a.device):
                _cabi.check(_cabi.lib().cf_local_corr_bwd(a.data_ptr(), b.data_ptr(), go.data_ptr(), B, C, H, W, radius, stride,
                                                          scale, 1 if x_major else 0, ga.data_ptr() if need_a else None,
                                                          gb.data_ptr() if need_b else None, _cabi.stream()), "cf_local_corr_bwd")
        return ga, gb, None, None, None, None


class CorrVolume(nn.Module):
    """Drop-in for ``nnunet.lib.raft.CorrVolume(radius, stride)`` by call-site contract (parity unpinned, see module docstring)."""

    def __init__(self, radius=4, stride=1, normalization="sqrt", x_major=False):
        super().__init__()
        if not 1 <= int(radius) <= 4:
            raise ValueError(f"radius must be in [1, 4], got {radius}")
        if int(stride) < 1:
            raise ValueError(f"stride must be >= 1, got {stride}")
        if normalization not in ("sqrt", "mean", "none"):
            raise ValueError("normalization must be 'sqrt', 'mean' or 'none'")
        self.radius, self.stride, self.normalization, self.x_major = int(radius), int(stride), normalization, bool(x_major)

    def forward(self, cur, prev):
        if cur.dim() != 4 or cur.shape != prev.shape:
            raise ValueError(f"cur and prev must be [B, C, H, W] of the same shape, got {tuple(cur.shape)} and {tuple(prev.shape)}")
        _cabi.require_cuda(cur, prev)
        C = cur.shape[1]
        scale = {"sqrt": 1.0 / math.sqrt(C) if C else 1.0, "mean": 1.0 / C if C else 1.0, "none": 1.0}[self.normalization]
        return _LocalCorr.apply(cur, prev, self.radius, self.stride, scale, self.x_major)
