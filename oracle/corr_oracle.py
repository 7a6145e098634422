"""Oracle (test infrastructure): CPU fp32 restatement of the RAFT correlation block.

Follows the call-site contract of the reference
(``/root/reference/nnunet/network_architecture/SegFlowGaussian.py:42`` import,
``:929`` ``CorrBlock(f1, f2, radius=4)``, ``:935`` ``corr_fn(coords1)``, ``:839`` ``coords_grid``) and the
arithmetic of the third-party code those names come from: public RAFT ``core/corr.py`` /
``core/utils/utils.py`` (``bilinear_sampler``, ``coords_grid``), identical to torchvision
``models/optical_flow/raft.py:387-431`` + ``_utils.py:8-26`` which the reference reaches through
``raft_large`` (``nnMTLTrainerV2Raft.py:97-118``).

Executed with torch CPU kernels (matmul / avg_pool2d / grid_sample), i.e. the very ATen ops the
reference's Python dispatches to -- that is why it is also the timed CPU baseline in ``bench.py``.
"""
import math

import torch
import torch.nn.functional as F


def coords_grid(batch, ht, wd, device="cpu"):
    """[batch, 2, ht, wd] fp32, channel 0 = x (column index), channel 1 = y (row index)."""
    ys, xs = torch.meshgrid(torch.arange(ht, device=device), torch.arange(wd, device=device), indexing="ij")
    grid = torch.stack([xs, ys], dim=0).float()
    return grid[None].repeat(batch, 1, 1, 1)


def bilinear_sampler(img, coords):
    """grid_sample with absolute pixel coordinates (x, y) in the last dim; align_corners=True, zeros."""
    h, w = img.shape[-2:]
    xgrid, ygrid = coords.split([1, 1], dim=-1)
    xgrid = 2 * xgrid / (w - 1) - 1
    ygrid = 2 * ygrid / (h - 1) - 1
    return F.grid_sample(img, torch.cat([xgrid, ygrid], dim=-1), align_corners=True)


def min_fmap_size(num_levels):
    # torchvision raft.py:368-383: the last level must keep >= 2 values per axis.
    return 2 * (2 ** (num_levels - 1))


class CorrBlock:
    """All-pairs correlation, avg-pooled pyramid, radius-r bilinear lookup."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        if fmap1.shape != fmap2.shape:
            raise ValueError(f"feature maps must have the same shape, got {tuple(fmap1.shape)} != {tuple(fmap2.shape)}")
        if any(s < min_fmap_size(num_levels) for s in fmap1.shape[-2:]):
            raise ValueError(f"feature maps too small for {num_levels} levels: {tuple(fmap1.shape[-2:])}")
        self.num_levels = num_levels
        self.radius = radius
        corr = self.corr(fmap1, fmap2)
        b, h1, w1, dim, h2, w2 = corr.shape
        corr = corr.reshape(b * h1 * w1, dim, h2, w2)
        self.corr_pyramid = [corr]
        for _ in range(num_levels - 1):
            corr = F.avg_pool2d(corr, 2, stride=2)
            self.corr_pyramid.append(corr)

    @staticmethod
    def corr(fmap1, fmap2):
        b, c, h, w = fmap1.shape
        f1 = fmap1.float().reshape(b, c, h * w)
        f2 = fmap2.float().reshape(b, c, h * w)
        corr = torch.matmul(f1.transpose(1, 2), f2)
        corr = corr.view(b, h, w, 1, h, w)
        return corr / torch.sqrt(torch.tensor(c).float())

    def __call__(self, coords):
        r = self.radius
        coords = coords.permute(0, 2, 3, 1)
        b, h, w, _ = coords.shape
        side = 2 * r + 1
        d = torch.linspace(-r, r, side)
        # first window axis displaces x, second displaces y (RAFT corr.py / torchvision raft.py:397-407)
        delta = torch.stack(torch.meshgrid(d, d, indexing="ij"), dim=-1).view(1, side, side, 2).to(coords.device)
        out = []
        for lvl, corr in enumerate(self.corr_pyramid):
            centroid = coords.reshape(b * h * w, 1, 1, 2) / 2 ** lvl
            sampled = bilinear_sampler(corr, centroid + delta)
            out.append(sampled.view(b, h, w, -1))
        return torch.cat(out, dim=-1).permute(0, 3, 1, 2).contiguous().float()


def algorithmic_flops(b, c, h, w):
    return 2.0 * b * (h * w) ** 2 * c


def pyramid_sizes(h, w, num_levels):
    sizes = [(h, w)]
    for _ in range(num_levels - 1):
        h, w = h // 2, w // 2
        sizes.append((h, w))
    return sizes


def lookup_algorithmic_bytes(b, h, w, num_levels=4, radius=4):
    """out + coords + window footprint."""
    hw = h * w
    k = (2 * radius + 1) ** 2
    foot = sum(min((2 * radius + 2) ** 2, hl * wl) for hl, wl in pyramid_sizes(h, w, num_levels))
    return b * (4 * num_levels * k * hw + 8 * hw + 4 * hw * foot)


def _self_check():
    torch.manual_seed(0)
    f1, f2 = torch.randn(1, 8, 16, 16), torch.randn(1, 8, 16, 16)
    blk = CorrBlock(f1, f2)
    out = blk(coords_grid(1, 16, 16) + torch.randn(1, 2, 16, 16))
    assert out.shape == (1, 324, 16, 16) and math.isfinite(out.sum().item())


if __name__ == "__main__":
    _self_check()
    print("ok")
