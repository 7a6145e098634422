"""Oracle (test infrastructure): CPU fp32 restatement of the reference's dense warp modules.

Restates ``/root/reference/nnunet/network_architecture/integration.py``:
``SpatialTransformer`` (:37-79), ``SpatialTransformerContour`` (:5-34), ``VecInt`` (:82-99).
Quirks kept on purpose:

* ``forward(flow, original, mode=...)`` -- flow first; the ``mode`` argument of forward is ignored,
  only the constructor's mode is used (:79).
* ``flow[:, 0]`` displaces rows (y), ``flow[:, 1]`` displaces columns (x) (:48-50, :74).
* fp32 sequence ``p = grid + flow ; n = 2 * (p / (S - 1) - 0.5)`` then ``grid_sample(align_corners=True)``
  with zero padding (:63-79).
* a float32 buffer ``grid`` of shape [1, 2, H, W] is part of the state dict (:59).
* the contour variant normalises its argument IN PLACE, divides channel i by ``size[~i] - 1`` and does not
  swap channels (:23-34).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


def identity_grid(size):
    axes = [torch.arange(0, s) for s in size]
    mesh = torch.meshgrid(*axes, indexing="ij")
    return torch.stack(mesh).unsqueeze(0).float()


def warp2d(flow, original, mode="bilinear", grid=None):
    """out[b,c,y,x] = sample(original[b,c], row = y + flow[b,0,y,x], col = x + flow[b,1,y,x])."""
    h, w = flow.shape[2:]
    if grid is None:
        grid = identity_grid((h, w)).to(flow.device)
    locs = grid + flow
    row_n = 2 * (locs[:, 0] / (h - 1) - 0.5)
    col_n = 2 * (locs[:, 1] / (w - 1) - 0.5)
    sample_grid = torch.stack([col_n, row_n], dim=-1)  # grid_sample wants (x, y)
    return F.grid_sample(original, sample_grid, align_corners=True, mode=mode)


class SpatialTransformer(nn.Module):
    def __init__(self, size, mode="bilinear"):
        super().__init__()
        if len(size) != 2:
            raise ValueError("oracle covers the 2-D transformer only (no 3-D caller exists in the reference)")
        self.mode = mode
        self.register_buffer("grid", identity_grid(size))

    def forward(self, flow, original, mode="bilinear"):
        return warp2d(flow, original, self.mode, self.grid)


class SpatialTransformerContour(nn.Module):
    def __init__(self, size, mode="bilinear"):
        super().__init__()
        self.mode = mode
        self.shape = size

    def forward(self, new_locs, original, mode="bilinear"):
        for i in range(2):
            new_locs[:, i, ...] = 2 * (new_locs[:, i, ...] / (self.shape[~i] - 1) - 0.5)
        return F.grid_sample(original, new_locs.permute(0, 2, 3, 1), align_corners=True, mode=self.mode)


class VecInt(nn.Module):
    """Scaling and squaring: vec <- vec / 2^n ; n times vec <- vec + warp(vec, vec)."""

    def __init__(self, inshape, nsteps):
        super().__init__()
        assert nsteps >= 0
        self.nsteps = nsteps
        self.scale = 1.0 / (2 ** nsteps)
        self.transformer = SpatialTransformer(inshape)

    def forward(self, vec):
        vec = vec * self.scale
        for _ in range(self.nsteps):
            vec = vec + self.transformer(vec, vec)
        return vec


def warp_onehot_argmax(flow, label, num_classes=4):
    """``warp_linear`` pattern of SegFlowGaussian.py:3571-3580: one_hot -> warp -> argmax(dim=1)."""
    onehot = F.one_hot(label[:, 0].long(), num_classes=num_classes).permute(0, 3, 1, 2).contiguous().float()
    return torch.argmax(warp2d(flow, onehot), dim=1, keepdim=True)


def warp_algorithmic_bytes(b, c, h, w):
    """src read + out write + flow read."""
    return 4 * b * h * w * (2 * c + 2)


def utils_warp(x, flo):
    """/root/reference/nnunet/lib/utils.py:29-59 `warp(x, flo)` without its hard-coded ``.cuda()``: flow channel 0
    displaces columns, channel 1 rows; ``grid_sample`` with its defaults (bilinear, zeros, align_corners=False)."""
    B, C, H, W = x.size()
    xx = torch.arange(0, W).view(1, -1).repeat(H, 1).view(1, 1, H, W).repeat(B, 1, 1, 1)
    yy = torch.arange(0, H).view(-1, 1).repeat(1, W).view(1, 1, H, W).repeat(B, 1, 1, 1)
    vgrid = torch.cat((xx, yy), 1).float() + flo
    vgrid[:, 0, :, :] = 2.0 * vgrid[:, 0, :, :].clone() / max(W - 1, 1) - 1.0
    vgrid[:, 1, :, :] = 2.0 * vgrid[:, 1, :, :].clone() / max(H - 1, 1) - 1.0
    return torch.nn.functional.grid_sample(x, vgrid.permute(0, 2, 3, 1), mode="bilinear", padding_mode="zeros",
                                           align_corners=False)
