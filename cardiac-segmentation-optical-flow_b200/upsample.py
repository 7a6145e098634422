"""Convex 8x upsampling with a learned mask (RAFT), the step after every lookup / update iteration.

Mirrors the two methods of the reference network
(/root/reference/nnunet/network_architecture/SegFlowGaussian.py:846-857 ``upsample_flow(flow, mask)`` and
:860-871 ``upsample_seg(seg, mask)``): same argument order, same ``8 *`` scaling of the field in both, output
``[N, C, 8H, 8W]``.  CUDA kernels behind ``cf_upsample_convex`` / ``cf_upsample_convex_bwd`` (the reference
differentiates through it in training: the loss is taken on the up-sampled flow of every iteration).
"""
import torch

from . import _cabi


class _UpsampleConvex(torch.autograd.Function):
    @staticmethod
    def forward(ctx, field, mask):
        f, m = _cabi.f32c(field), _cabi.f32c(mask)
        N, C, H, W = f.shape
        out = torch.empty(N, C, 8 * H, 8 * W, dtype=torch.float32, device=f.device)
        with torch.cuda.device(f.device):
            _cabi.check(_cabi.lib().cf_upsample_convex(f.data_ptr(), m.data_ptr(), N, C, H, W, 8, 8.0, out.data_ptr(),
                                                       _cabi.stream()), "cf_upsample_convex")
        ctx.save_for_backward(f, m)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        f, m = ctx.saved_tensors
        N, C, H, W = f.shape
        go = grad_out.contiguous().float()
        gf = torch.zeros_like(f) if ctx.needs_input_grad[0] else None
        gm = torch.empty_like(m) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(f.device):
            _cabi.check(_cabi.lib().cf_upsample_convex_bwd(f.data_ptr(), m.data_ptr(), go.data_ptr(), N, C, H, W, 8, 8.0,
                                                           gf.data_ptr() if gf is not None else None,
                                                           gm.data_ptr() if gm is not None else None, _cabi.stream()),
                        "cf_upsample_convex_bwd")
        return gf, gm


def _upsample(field, mask, what):
    if field.dim() != 4 or mask.dim() != 4:
        raise ValueError(f"{what}: field and mask must be 4-D, got {tuple(field.shape)} and {tuple(mask.shape)}")
    N, C, H, W = field.shape
    if mask.shape[0] != N or mask.shape[1] != 9 * 64 or tuple(mask.shape[2:]) != (H, W):
        raise ValueError(f"{what}: mask must be [{N}, 576, {H}, {W}] for a field {tuple(field.shape)}, got {tuple(mask.shape)}")
    _cabi.require_cuda(field, mask)
    return _UpsampleConvex.apply(field, mask)


def upsample_flow(flow, mask):
    """[N,2,H,W] flow at 1/8 resolution -> [N,2,8H,8W] (SegFlowGaussian.py:846-857)."""
    return _upsample(flow, mask, "upsample_flow")


def upsample_seg(seg, mask):
    """[N,4,H,W] segmentation logits at 1/8 resolution -> [N,4,8H,8W] (SegFlowGaussian.py:860-871)."""
    return _upsample(seg, mask, "upsample_seg")
