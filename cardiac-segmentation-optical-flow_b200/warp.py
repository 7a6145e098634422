"""SpatialTransformer / SpatialTransformerContour / VecInt with the reference's module API, on sm_100a kernels.

Mirrors /root/reference/nnunet/network_architecture/integration.py (:37-79, :5-34, :82-99) including its
quirks: ``forward(flow, original, mode=...)`` with flow first; forward's ``mode`` ignored
(constructor's used); flow channel 0 displaces rows and channel 1 columns; a float32 ``grid`` buffer
[1,2,H,W] in the state dict (checkpoints are loaded strictly, network_trainer.py:444).
"""
import torch
import torch.nn as nn

from . import _cabi

_MODES = {"bilinear": _cabi.MODE_BILINEAR, "nearest": _cabi.MODE_NEAREST}


class _Warp2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flow, original, mode, out=None):
        f, src = _cabi.f32c(flow), _cabi.f32c(original)
        B, C, H, W = src.shape
        if out is None:
            out = torch.empty_like(src)
        elif out.shape != src.shape or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous fp32 tensor of shape {tuple(src.shape)}")
        ctx.empty = src.numel() == 0          # empty batch / zero channels: grid_sample returns an empty tensor, so do we
        if ctx.empty:
            ctx.save_for_backward(f, src)
            ctx.mode = mode
            return out
        with torch.cuda.device(src.device):
            _cabi.check(_cabi.lib().cf_warp2d(f.data_ptr(), src.data_ptr(), B, C, H, W, mode, out.data_ptr(),
                                              _cabi.stream()), "cf_warp2d")
        ctx.save_for_backward(f, src)
        ctx.mode = mode
        return out

    @staticmethod
    def backward(ctx, grad_out):
        f, src = ctx.saved_tensors
        B, C, H, W = src.shape
        go = grad_out.contiguous().float()
        need_flow, need_src = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gsrc = torch.zeros_like(src) if need_src else None
        gflow = torch.empty_like(f) if need_flow else None
        if ctx.empty:
            return (gflow.zero_() if need_flow else None), gsrc, None, None
        if ctx.mode != _cabi.MODE_BILINEAR:
            # ATen grid_sampler_2d_backward, nearest: the output gradient goes to the nearest texel, the grid gradient is zero
            if need_src:
                with torch.cuda.device(src.device):
                    _cabi.check(_cabi.lib().cf_warp2d_nearest_bwd(f.data_ptr(), go.data_ptr(), B, C, H, W, gsrc.data_ptr(),
                                                                  _cabi.stream()), "cf_warp2d_nearest_bwd")
            return (gflow.zero_() if need_flow else None), gsrc, None, None
        with torch.cuda.device(src.device):
            _cabi.check(_cabi.lib().cf_warp2d_bwd(f.data_ptr(), src.data_ptr(), go.data_ptr(), B, C, H, W,
                                                  gsrc.data_ptr() if need_src else None,
                                                  gflow.data_ptr() if need_flow else None, _cabi.stream()),
                        "cf_warp2d_bwd")
        return gflow, gsrc, None, None


def _check_pair(flow, original, size=None):
    if flow.dim() != 4 or flow.shape[1] != 2:
        raise ValueError(f"flow must be [B, 2, H, W], got {tuple(flow.shape)}")
    if original.dim() != 4 or original.shape[0] != flow.shape[0] or original.shape[2:] != flow.shape[2:]:
        raise ValueError(f"original must be [B, C, H, W] matching flow {tuple(flow.shape)}, got {tuple(original.shape)}")
    if size is not None and tuple(flow.shape[2:]) != tuple(size):
        raise ValueError(f"flow spatial size {tuple(flow.shape[2:])} != transformer size {tuple(size)}")
    _cabi.require_cuda(flow, original)


def warp(src, flow, mode="bilinear", out=None):
    """Functional alias in voxelmorph's (src, flow) order; the module below keeps the reference's order."""
    _check_pair(flow, src)
    return _Warp2d.apply(flow, src, _MODES[mode], out)


def warp_onehot_argmax(flow, label, num_classes=4, out=None, return_counts=False, counts_out=None, counts_total=False):
    """argmax_c warp(one_hot(label), flow) without the one-hot intermediate (the `warp_linear` pattern,
    SegFlowGaussian.py:3571-3580).  label [B,1,H,W] integer; returns int64 [B,1,H,W] (written into ``out`` when given).
    ``return_counts=True`` also returns int64 [B, num_classes]: pixels per class of every warped map, from the same launch
    (written into ``counts_out`` when given: e.g. a row slice of a per-cine table that several lanes fill).
    A uint8 ``out`` selects the 1-byte label maps (cf_warp2d_onehot_argmax_u8; ``counts_total=True``: one int64 [num_classes]
    row for the whole batch instead of one per frame)."""
    if label.dim() != 4 or label.shape[1] != 1 or label.shape[0] != flow.shape[0] or label.shape[2:] != flow.shape[2:]:
        raise ValueError(f"label must be [B, 1, H, W] matching flow {tuple(flow.shape)}, got {tuple(label.shape)}")
    _cabi.require_cuda(flow, label)
    f = _cabi.f32c(flow)
    lab = label.detach().to(torch.int64).contiguous()
    B, _, H, W = lab.shape
    if out is None:
        out = torch.empty_like(lab)
    elif (out.dtype not in (torch.int64, torch.uint8) or out.shape != lab.shape or not out.is_contiguous()
          or out.device != lab.device):
        raise ValueError("out must be a contiguous int64 (or uint8) tensor of the label's shape on the same device")
    if out.dtype == torch.uint8:          # registered label maps, 1 byte per pixel (e.g. a slice of an all-gather buffer)
        if lab.numel() == 0:
            return (out, counts_out) if return_counts else out
        counts = None
        if return_counts:
            cshape = (num_classes,) if counts_total else (B, num_classes)     # class areas of the whole batch / per frame
            counts = counts_out if counts_out is not None else torch.empty(cshape, dtype=torch.int64, device=lab.device)
            if (counts.dtype != torch.int64 or tuple(counts.shape) != cshape or not counts.is_contiguous()
                    or counts.device != lab.device):
                raise ValueError(f"counts_out must be a contiguous int64 {list(cshape)} tensor on the label's device")
        with torch.cuda.device(f.device):
            _cabi.check(_cabi.lib().cf_warp2d_onehot_argmax_u8(f.data_ptr(), lab.data_ptr(), num_classes, B, H, W, out.data_ptr(),
                                                               counts.data_ptr() if counts is not None else None,
                                                               1 if counts_total else 0, _cabi.stream()),
                        "cf_warp2d_onehot_argmax_u8")
        return (out, counts) if return_counts else out
    if lab.numel() == 0:                      # empty batch: nothing to warp (argmax of an empty one-hot stack is empty)
        if return_counts:
            return out, (counts_out if counts_out is not None else torch.zeros(B, num_classes, dtype=torch.int64, device=lab.device))
        return out
    with torch.cuda.device(f.device):
        if return_counts:
            if counts_out is None:
                counts = torch.empty(B, num_classes, dtype=torch.int64, device=lab.device)
            else:
                counts = counts_out
                if (counts.dtype != torch.int64 or tuple(counts.shape) != (B, num_classes) or not counts.is_contiguous()
                        or counts.device != lab.device):
                    raise ValueError(f"counts_out must be a contiguous int64 [{B}, {num_classes}] tensor on the label's device")
            _cabi.check(_cabi.lib().cf_warp2d_onehot_argmax_counts(f.data_ptr(), lab.data_ptr(), num_classes, B, H, W,
                                                                   out.data_ptr(), counts.data_ptr(), _cabi.stream()),
                        "cf_warp2d_onehot_argmax_counts")
            return out, counts
        _cabi.check(_cabi.lib().cf_warp2d_onehot_argmax(f.data_ptr(), lab.data_ptr(), num_classes, B, H, W,
                                                        out.data_ptr(), _cabi.stream()), "cf_warp2d_onehot_argmax")
    return out


class _UtilsWarp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, flo):
        f, src = _cabi.f32c(flo), _cabi.f32c(x)
        B, C, H, W = src.shape
        out = torch.empty_like(src)
        with torch.cuda.device(src.device):
            _cabi.check(_cabi.lib().cf_warp2d_xy(f.data_ptr(), src.data_ptr(), B, C, H, W, out.data_ptr(), _cabi.stream()),
                        "cf_warp2d_xy")
        ctx.save_for_backward(f, src)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        f, src = ctx.saved_tensors
        B, C, H, W = src.shape
        go = grad_out.contiguous().float()
        need_src, need_flow = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gsrc = torch.zeros_like(src) if need_src else None
        gflow = torch.empty_like(f) if need_flow else None
        with torch.cuda.device(src.device):
            _cabi.check(_cabi.lib().cf_warp2d_xy_bwd(f.data_ptr(), src.data_ptr(), go.data_ptr(), B, C, H, W,
                                                     gsrc.data_ptr() if need_src else None,
                                                     gflow.data_ptr() if need_flow else None, _cabi.stream()),
                        "cf_warp2d_xy_bwd")
        return gsrc, gflow


def utils_warp(x, flo):
    """Drop-in for ``nnunet.lib.utils.warp(x, flo)`` (/root/reference/nnunet/lib/utils.py:29-59, called at
    SegFlowGaussian.py:1044 inside the training graph): x [B,C,H,W], flo [B,2,H,W] with channel 0 = x displacement;
    grid_sample defaults (bilinear, zeros, align_corners=False).  Differentiable w.r.t. both arguments."""
    _check_pair(flo, x)
    return _UtilsWarp.apply(x, flo)


def warp_taps(flow):
    """Parity hook: [B,2,H,W] sampling coordinates (ch0 = row, ch1 = column) the warp kernel derives."""
    _cabi.require_cuda(flow)
    f = _cabi.f32c(flow)
    B, _, H, W = f.shape
    taps = torch.empty_like(f)
    with torch.cuda.device(f.device):
        _cabi.check(_cabi.lib().cf_warp2d_taps(f.data_ptr(), B, H, W, taps.data_ptr(), _cabi.stream()), "cf_warp2d_taps")
    return taps


class SpatialTransformer(nn.Module):
    """N-D spatial transformer of the reference, 2-D only (no caller in the reference builds a 3-D one)."""

    def __init__(self, size, mode="bilinear"):
        super().__init__()
        size = tuple(int(s) for s in size)
        if len(size) != 2:
            raise ValueError("only the 2-D SpatialTransformer is on the hot path (no 3-D caller exists in the reference)")
        if mode not in _MODES:
            raise ValueError(f"mode must be 'bilinear' or 'nearest', got {mode!r}")
        self.mode = mode
        self.size = size
        ys, xs = torch.meshgrid(torch.arange(0, size[0]), torch.arange(0, size[1]), indexing="ij")
        # kept only for state-dict compatibility (integration.py:59); the kernel derives the grid from indices
        self.register_buffer("grid", torch.stack([ys, xs]).unsqueeze(0).float())

    def forward(self, flow, original, mode="bilinear"):
        _check_pair(flow, original, self.size)
        return _Warp2d.apply(flow, original, _MODES[self.mode])


class _SamplePoints(torch.autograd.Function):
    """(locs [B,2,h,P] in pixels, original [B,C,H,W]) -> [B,C,h,P]; differentiable w.r.t. both (the reference back-propagates
    its MSE point loss through this module into ``original`` = the predicted flow,
    training/network_training/SegFlowGaussian.py:1497-1503)."""

    @staticmethod
    def forward(ctx, locs, original, mode, size):
        lc, src = _cabi.f32c(locs), _cabi.f32c(original)
        if lc.data_ptr() == locs.data_ptr():
            lc = lc.clone()                     # the caller normalises its tensor in place afterwards (reference side effect)
        B, C, H, W = src.shape
        P = lc.shape[2] * lc.shape[3]
        out = torch.empty(B, C, lc.shape[2], lc.shape[3], dtype=torch.float32, device=src.device)
        with torch.cuda.device(src.device):
            _cabi.check(_cabi.lib().cf_sample_points(lc.data_ptr(), src.data_ptr(), B, C, H, W, P, size[0], size[1], mode,
                                                     out.data_ptr(), _cabi.stream()), "cf_sample_points")
        ctx.save_for_backward(lc, src)
        ctx.mode, ctx.size = mode, size
        return out

    @staticmethod
    def backward(ctx, grad_out):
        lc, src = ctx.saved_tensors
        B, C, H, W = src.shape
        P = lc.shape[2] * lc.shape[3]
        go = grad_out.contiguous().float()
        need_locs, need_src = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gsrc = torch.zeros_like(src) if need_src else None
        glocs = torch.empty_like(lc) if need_locs else None
        if need_locs or need_src:
            with torch.cuda.device(src.device):
                _cabi.check(_cabi.lib().cf_sample_points_bwd(lc.data_ptr(), src.data_ptr(), go.data_ptr(), B, C, H, W, P,
                                                             ctx.size[0], ctx.size[1], ctx.mode,
                                                             gsrc.data_ptr() if need_src else None,
                                                             glocs.data_ptr() if need_locs else None, _cabi.stream()),
                            "cf_sample_points_bwd")
        return glocs, gsrc, None, None


class SpatialTransformerContour(nn.Module):
    """Samples ``original`` at P points; ``new_locs`` [B,2,1,P] (ch0 = x, ch1 = y, reversed-size
    normalisation as integration.py:23-24).  Like the reference it normalises ``new_locs`` IN PLACE."""

    def __init__(self, size, mode="bilinear"):
        super().__init__()
        if mode not in _MODES:
            raise ValueError(f"mode must be 'bilinear' or 'nearest', got {mode!r}")
        self.mode = mode
        self.shape = tuple(int(s) for s in size)

    def forward(self, new_locs, original, mode="bilinear"):
        if new_locs.dim() != 4 or new_locs.shape[1] != 2:
            raise ValueError(f"new_locs must be [B, 2, h, P], got {tuple(new_locs.shape)}")
        _cabi.require_cuda(new_locs, original)
        if original.dim() != 4 or original.shape[0] != new_locs.shape[0]:
            raise ValueError(f"original must be [B, C, H, W] with B = {new_locs.shape[0]}, got {tuple(original.shape)}")
        out = _SamplePoints.apply(new_locs, original, _MODES[self.mode], (self.shape[-2], self.shape[-1]))
        # reference side effect (integration.py:23-24): the argument ends up normalised to [-1, 1]
        with torch.no_grad():
            for i in range(2):
                new_locs[:, i, ...] = 2 * (new_locs[:, i, ...] / (self.shape[~i] - 1) - 0.5)
        return out


class VecInt(nn.Module):
    """Scaling and squaring (integration.py:82-99): vec / 2^n, then n times vec += warp(vec, vec)."""

    def __init__(self, inshape, nsteps):
        super().__init__()
        assert nsteps >= 0, "nsteps should be >= 0, found: %d" % nsteps
        self.nsteps = nsteps
        self.scale = 1.0 / (2 ** self.nsteps)
        self.transformer = SpatialTransformer(inshape)

    def forward(self, vec):
        if vec.requires_grad and torch.is_grad_enabled():
            vec = vec * self.scale  # autograd path: compose from the differentiable warp
            for _ in range(self.nsteps):
                vec = vec + self.transformer(vec, vec)
            return vec
        _check_pair(vec, vec, self.transformer.size)
        v = _cabi.f32c(vec)
        B, _, H, W = v.shape
        out, work = torch.empty_like(v), torch.empty_like(v)
        with torch.cuda.device(v.device):
            _cabi.check(_cabi.lib().cf_vecint(v.data_ptr(), B, H, W, self.nsteps, out.data_ptr(), work.data_ptr(),
                                              _cabi.stream()), "cf_vecint")
        return out
