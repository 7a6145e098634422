"""CorrBlock / coords_grid with the reference's call-site API, on hand-written sm_100a kernels.

Call-site contract (reference = /root/reference):
  ``from nnunet.lib.raft_initial import CorrBlock, BasicUpdateBlock, coords_grid``
      (nnunet/network_architecture/SegFlowGaussian.py:42 -- the file itself was never committed)
  ``corr_fn = CorrBlock(fmap1, fmap2, radius=4)``   (:929, 1030, 1161, 1270, 1390, ...)
  ``corr = corr_fn(coords1)``                       (:935; coords [B,2,H/8,W/8], ch0 = x, ch1 = y)
  ``coords_grid(N, H//8, W//8, device=img.device)`` (:839-840)
Arithmetic: public RAFT core/corr.py == torchvision models/optical_flow/raft.py:387-431.
"""
import os

import torch

from . import _cabi


def coords_grid(batch, ht, wd, device="cuda"):
    """[batch, 2, ht, wd] fp32 pixel grid, channel 0 = x, channel 1 = y (RAFT utils.coords_grid)."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("cardioflow_b200.coords_grid needs a CUDA device (no CPU fallback)")
    out = torch.empty(batch, 2, ht, wd, dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        _cabi.check(_cabi.lib().cf_coords_grid(out.data_ptr(), batch, ht, wd, _cabi.stream()), "cf_coords_grid")
    return out


def _level_views(pyramid, B, H, W, L):
    lib = _cabi.lib()
    views = []
    for lvl in range(L):
        h, w = H >> lvl, W >> lvl
        off = lib.cf_corr_pyramid_level_offset(B, H, W, lvl) // 4
        views.append(pyramid[off:off + B * H * W * h * w].view(B * H * W, 1, h, w))
    return views


_WORKSPACES = {}


def _workspace(device, nbytes):
    """Per-device scratch for the bf16 copies of the feature maps (stream-ordered reuse: every use is enqueued
    on the current stream before the next one)."""
    if nbytes == 0:
        return None
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _WORKSPACES.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _WORKSPACES[key] = buf
    return buf


def _as_operand(t, precision):
    """(tensor, CF_DTYPE_*, CF_LAYOUT_*) the builder can read without a host-side copy where possible: bf16 and / or
    channels_last maps go in as they are (a bf16 channels_last map is already the K-major operand the TMA wants: no pre-pass
    at all); anything else becomes fp32 NCHW like RAFT's ``.float()``."""
    t = t.detach()
    if precision == _cabi.PREC_FP32:                      # the exact CUDA-core mode reads fp32 NCHW only
        return _cabi.f32c(t), _cabi.DTYPE_F32, _cabi.LAYOUT_NCHW
    if t.dtype not in (torch.float32, torch.bfloat16):
        t = t.float()
    dt = _cabi.DTYPE_BF16 if t.dtype == torch.bfloat16 else _cabi.DTYPE_F32
    if t.is_contiguous():
        return t, dt, _cabi.LAYOUT_NCHW
    if t.is_contiguous(memory_format=torch.channels_last):
        return t, dt, _cabi.LAYOUT_NHWC
    return t.contiguous(), dt, _cabi.LAYOUT_NCHW


class _CorrBuild(torch.autograd.Function):
    """(fmap1, fmap2) -> packed pyramid.  a2 + a3 of SURVEY.md 8(a)."""

    @staticmethod
    def forward(ctx, fmap1, fmap2, num_levels, precision, pyramid=None, allow_tiled=False, info=None, gstate=None):
        B, C, H, W = fmap1.shape
        lib = _cabi.lib()
        (f1, dt, lay), (f2, dt2, lay2) = _as_operand(fmap1, precision), _as_operand(fmap2, precision)
        if (dt, lay) != (dt2, lay2):          # mixed inputs: bring both to the reference's fp32 NCHW
            f1, f2, dt, lay = _cabi.f32c(fmap1), _cabi.f32c(fmap2), _cabi.DTYPE_F32, _cabi.LAYOUT_NCHW
        nfloats = lib.cf_corr_pyramid_bytes(B, H, W, num_levels) // 4
        if pyramid is None:
            pyramid = torch.empty(nfloats, dtype=torch.float32, device=f1.device)
        elif pyramid.numel() < nfloats or pyramid.dtype != torch.float32 or not pyramid.is_contiguous():
            raise ValueError(f"pyramid_buffer must be a contiguous fp32 tensor of >= {nfloats} elements")
        else:
            pyramid = pyramid.view(-1)[:nfloats]
        used = _cabi.ctypes.c_int(-1)
        tiled = _cabi.ctypes.c_uint(0)
        zero_copy = dt == _cabi.DTYPE_BF16 and lay == _cabi.LAYOUT_NHWC
        ws_bytes = 0 if zero_copy else lib.cf_corr_build_workspace_bytes(B, C, H, W, precision)
        ws = _workspace(f1.device, ws_bytes)
        with torch.cuda.device(f1.device):
            _cabi.check(lib.cf_corr_build_ex(f1.data_ptr(), f2.data_ptr(), dt, lay, B, C, H, W, num_levels, precision,
                                             _cabi.BUILD_ALLOW_TILED if allow_tiled else 0, pyramid.data_ptr(),
                                             ws.data_ptr() if ws is not None else None, ws_bytes,
                                             _cabi.ctypes.byref(used), _cabi.ctypes.byref(tiled), _cabi.stream()),
                        "cf_corr_build_ex")
        if info is not None:          # per-call report (thread-safe, unlike a class attribute)
            info["tiled_mask"], info["used_precision"] = tiled.value, used.value
        if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
            # the backward GEMMs read fp32 NCHW maps (training in the reference is fp32)
            ctx.save_for_backward(_cabi.f32c(fmap1), _cabi.f32c(fmap2))
        ctx.meta = (B, C, H, W, num_levels)
        ctx.used_precision = used.value
        ctx.gstate = gstate
        return pyramid

    @staticmethod
    def backward(ctx, grad_pyramid):
        f1, f2 = ctx.saved_tensors
        B, C, H, W, L = ctx.meta
        lib = _cabi.lib()
        shared = ctx.gstate.pop("buf", None) if ctx.gstate is not None else None
        if shared is not None and grad_pyramid.data_ptr() == shared.data_ptr():
            g = shared                          # the lookups' own accumulation buffer: pooled in place, no copy
        else:
            g = grad_pyramid.contiguous().clone()
        gf1, gf2 = torch.empty_like(f1), torch.empty_like(f2)
        with torch.cuda.device(f1.device):
            s = _cabi.stream()
            _cabi.check(lib.cf_corr_pool_bwd(g.data_ptr(), B, H, W, L, s), "cf_corr_pool_bwd")
            # fp32 forward -> fp32 CUDA-core GEMMs; bf16 / bf16x3 forward -> tcgen05 with the bf16x3 split (fp32-grade) where
            # the shape allows it (H*W % 8 == 0), reported in used
            prec = _cabi.PRECISIONS["fp32"] if ctx.used_precision == _cabi.PRECISIONS["fp32"] else _cabi.PRECISIONS["auto"]
            ws_bytes = lib.cf_corr_volume_bwd_workspace_bytes(B, C, H, W) if prec != _cabi.PRECISIONS["fp32"] else 0
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=f1.device) if ws_bytes else None
            used = _cabi.ctypes.c_int(0)
            _cabi.check(lib.cf_corr_volume_bwd_ex(f1.data_ptr(), f2.data_ptr(), g.data_ptr(), B, C, H, W, prec,
                                                  gf1.data_ptr() if ctx.needs_input_grad[0] else None,
                                                  gf2.data_ptr() if ctx.needs_input_grad[1] else None,
                                                  ws.data_ptr() if ws is not None else None, ws_bytes,
                                                  _cabi.ctypes.byref(used), s), "cf_corr_volume_bwd_ex")
        return (gf1 if ctx.needs_input_grad[0] else None), (gf2 if ctx.needs_input_grad[1] else None), None, None, None, None, None, None


class _CorrLookup(torch.autograd.Function):
    """(packed pyramid, coords) -> [B, L*(2r+1)^2, H, W].  a4 of SURVEY.md 8(a); no gradient w.r.t. coords
    (the reference detaches them before every lookup, SegFlowGaussian.py:933)."""

    @staticmethod
    def forward(ctx, pyramid, coords, meta, out=None, tiled_mask=0, gstate=None):
        B, H, W, L, r = meta
        c = _cabi.f32c(coords)
        shape = (B, L * (2 * r + 1) ** 2, H, W)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=pyramid.device)
        elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous fp32 tensor of shape {shape}")
        with torch.cuda.device(pyramid.device):
            _cabi.check(_cabi.lib().cf_corr_lookup_layout(pyramid.data_ptr(), c.data_ptr(), B, H, W, L, r, tiled_mask,
                                                          out.data_ptr(), _cabi.stream()), "cf_corr_lookup_layout")
        ctx.save_for_backward(c)
        ctx.meta = meta
        ctx.numel = pyramid.numel()
        ctx.gstate = gstate
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (c,) = ctx.saved_tensors
        B, H, W, L, r = ctx.meta
        # The 12 lookups of one CorrBlock add their gradients into ONE zeroed pyramid-sized buffer (the kernel accumulates).
        # The first backward of a pass hands that buffer to autograd as the pyramid's gradient, the others return None; the
        # build's backward runs after all of them (it depends on every lookup node) and takes the buffer out of `gstate`.
        # A buffer left over from another backward pass (e.g. gradients taken w.r.t. the pyramid itself) is never reused.
        task = torch._C._current_graph_task_id()
        first = ctx.gstate is None or ctx.gstate.get("buf") is None or ctx.gstate.get("task") != task
        if first:
            g = torch.zeros(ctx.numel, dtype=torch.float32, device=grad_out.device)
            if ctx.gstate is not None:
                ctx.gstate["buf"], ctx.gstate["task"] = g, task
        else:
            g = ctx.gstate["buf"]
        go = grad_out.contiguous().float()
        with torch.cuda.device(go.device):
            _cabi.check(_cabi.lib().cf_corr_lookup_bwd(c.data_ptr(), go.data_ptr(), B, H, W, L, r, g.data_ptr(),
                                                       _cabi.stream()), "cf_corr_lookup_bwd")
        return (g if first else None), None, None, None, None, None


class _OtfBuild(torch.autograd.Function):
    """Autograd anchor of an on-the-fly block: (fmap1, fmap2) -> a one-element token that the block's lookups depend on.  No volume
    exists in the forward pass; the BACKWARD pass materialises the gradient pyramid (the lookups' backward kernels accumulate into
    one buffer of the volume's size, 4/3 * B * (H*W)^2 floats, allocated when the first of them runs) and then differentiates
    the correlation exactly as the materialised path does (cf_corr_pool_bwd, cf_corr_volume_bwd_ex)."""

    @staticmethod
    def forward(ctx, fmap1, fmap2, meta, gstate):
        ctx.save_for_backward(_cabi.f32c(fmap1), _cabi.f32c(fmap2))
        ctx.meta = meta
        ctx.gstate = gstate
        return torch.zeros(1, dtype=torch.float32, device=fmap1.device)

    @staticmethod
    def backward(ctx, grad_token):
        f1, f2 = ctx.saved_tensors
        B, C, H, W, L = ctx.meta
        g = ctx.gstate.pop("buf", None)
        ctx.gstate.pop("task", None)
        if g is None:                           # no lookup took part in this backward pass
            return torch.zeros_like(f1), torch.zeros_like(f2), None, None
        lib = _cabi.lib()
        gf1, gf2 = torch.empty_like(f1), torch.empty_like(f2)
        with torch.cuda.device(f1.device):
            s = _cabi.stream()
            _cabi.check(lib.cf_corr_pool_bwd(g.data_ptr(), B, H, W, L, s), "cf_corr_pool_bwd")
            ws_bytes = lib.cf_corr_volume_bwd_workspace_bytes(B, C, H, W)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=f1.device) if ws_bytes else None
            used = _cabi.ctypes.c_int(0)
            _cabi.check(lib.cf_corr_volume_bwd_ex(f1.data_ptr(), f2.data_ptr(), g.data_ptr(), B, C, H, W, _cabi.PRECISIONS["auto"],
                                                  gf1.data_ptr() if ctx.needs_input_grad[0] else None,
                                                  gf2.data_ptr() if ctx.needs_input_grad[1] else None,
                                                  ws.data_ptr() if ws is not None else None, ws_bytes,
                                                  _cabi.ctypes.byref(used), s), "cf_corr_volume_bwd_ex")
        return (gf1 if ctx.needs_input_grad[0] else None), (gf2 if ctx.needs_input_grad[1] else None), None, None


class _OtfLookup(torch.autograd.Function):
    """On-the-fly lookup inside an autograd graph: forward = the block's volume-free kernel; backward = cf_corr_lookup_bwd into the
    block's shared gradient pyramid (see _OtfBuild).  No gradient w.r.t. coords (the reference detaches them, SegFlowGaussian.py:933)."""

    @staticmethod
    def forward(ctx, token, coords, block, out=None):
        res = block._lookup_on_the_fly(coords, out)
        ctx.save_for_backward(_cabi.f32c(coords))
        ctx.meta = block._meta
        ctx.gstate = block._gstate
        return res

    @staticmethod
    def backward(ctx, grad_out):
        (c,) = ctx.saved_tensors
        B, H, W, L, r = ctx.meta
        lib = _cabi.lib()
        task = torch._C._current_graph_task_id()
        if ctx.gstate.get("buf") is None or ctx.gstate.get("task") != task:
            numel = lib.cf_corr_pyramid_bytes(B, H, W, L) // 4
            ctx.gstate["buf"], ctx.gstate["task"] = torch.zeros(numel, dtype=torch.float32, device=grad_out.device), task
        go = grad_out.contiguous().float()
        with torch.cuda.device(go.device):
            _cabi.check(lib.cf_corr_lookup_bwd(c.data_ptr(), go.data_ptr(), B, H, W, L, r, ctx.gstate["buf"].data_ptr(),
                                               _cabi.stream()), "cf_corr_lookup_bwd")
        return torch.zeros(1, dtype=torch.float32, device=go.device), None, None, None


class CorrBlock:
    """Drop-in for RAFT's ``CorrBlock(fmap1, fmap2, num_levels=4, radius=4)``; instances are callable.

    ``precision``: "auto" (default; env ``CARDIOFLOW_CORR_PRECISION``) = fp32-grade like the reference: "bf16x3" on the tensor
    cores (hi/lo bf16 splits, three tcgen05 MMAs per K step, fp32 accumulate, ~1e-6) when the shape allows it, else "fp32"
    (CUDA cores); "bf16" (one MMA per K step, 2-3e-3 relative error) is opt-in.  Feature maps may be fp32 or bf16, contiguous
    or ``channels_last`` (a bf16 channels_last map needs no conversion pass at all).  ``pyramid_buffer`` / ``out=`` let a caller that loops over
    frame pairs reuse device memory (the reference allocates afresh every call).  Errors follow torchvision raft.py:368-383 (ValueError).
    """

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4, precision=None, pyramid_buffer=None, on_the_fly=False):
        if fmap1.shape != fmap2.shape:
            raise ValueError(f"Input feature maps should have the same shape, instead got {tuple(fmap1.shape)} "
                             f"(fmap1.shape) != {tuple(fmap2.shape)} (fmap2.shape)")
        if fmap1.dim() != 4:
            raise ValueError(f"feature maps must be [B, C, H, W], got {tuple(fmap1.shape)}")
        min_size = 2 * (2 ** (num_levels - 1))
        if any(s < min_size for s in fmap1.shape[-2:]):
            raise ValueError("Feature maps are too small to be down-sampled by the correlation pyramid. "
                             f"H and W of feature maps should be at least {min_size}; got: {tuple(fmap1.shape[-2:])}.")
        _cabi.require_cuda(fmap1, fmap2)
        precision = precision or os.environ.get("CARDIOFLOW_CORR_PRECISION", "auto")
        if precision not in _cabi.PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_cabi.PRECISIONS)}, got {precision!r}")
        self.num_levels = num_levels
        self.radius = radius
        B, _, H, W = fmap1.shape
        self._meta = (B, H, W, num_levels, radius)
        self.on_the_fly = bool(on_the_fly)
        if self.on_the_fly:
            # K4: keep fmap1 and the pooled fmap2 pyramid, recompute window correlations per lookup.  Under autograd the forward pass
            # still holds no volume; the backward pass materialises the gradient pyramid once (see _OtfBuild)
            lib = _cabi.lib()
            C = fmap1.shape[1]
            self._token = None
            if torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad):
                self._gstate = {}
                self._token = _OtfBuild.apply(fmap1, fmap2, (B, C, H, W, num_levels), self._gstate)
            self._f1 = _cabi.f32c(fmap1)
            f2 = _cabi.f32c(fmap2)
            self._f2_pyramid = torch.empty(lib.cf_fmap_pyramid_bytes(B, C, H, W, num_levels) // 4, dtype=torch.float32,
                                           device=f2.device)
            # tensor-core variant (tcgen05, bf16 operands) when the shape fits and fp32 was not asked for; else the fp32 CUDA-core one
            self._otf_tc = precision != "fp32" and bool(lib.cf_corr_otf_supported(B, C, H, W, num_levels))
            with torch.cuda.device(f2.device):
                if self._otf_tc:
                    self._operands = torch.empty(lib.cf_corr_otf_operand_bytes(B, C, H, W, num_levels), dtype=torch.uint8,
                                                 device=f2.device)
                    _cabi.check(lib.cf_corr_otf_prepare(self._f1.data_ptr(), f2.data_ptr(), B, C, H, W, num_levels,
                                                        self._operands.data_ptr(), self._f2_pyramid.data_ptr(),
                                                        self._f2_pyramid.numel() * 4, _cabi.stream()), "cf_corr_otf_prepare")
                    self._f2_pyramid = None        # scratch of the preparation only
                else:
                    _cabi.check(lib.cf_fmap_pyramid_build(f2.data_ptr(), B, C, H, W, num_levels,
                                                          self._f2_pyramid.data_ptr(), _cabi.stream()), "cf_fmap_pyramid_build")
            self.used_precision = "bf16" if self._otf_tc else "fp32"
            self._pyramid = None
            self._tiled_mask = 0
            return
        # Inference only: the tensor-core builder may store level 0 as 4x4 tiles (fewer DRAM atoms per lookup window).  With
        # autograd on, the pyramid stays row-major: the backward kernels and any user arithmetic on corr_pyramid expect it.
        needs_grad = torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad)
        allow_tiled = not needs_grad and os.environ.get("CARDIOFLOW_TILED", "1") != "0"
        info = {}
        self._gstate = {}      # gradient pyramid shared by the backward passes of this block's lookups
        self._pyramid = _CorrBuild.apply(fmap1, fmap2, num_levels, _cabi.PRECISIONS[precision], pyramid_buffer, allow_tiled, info,
                                         self._gstate)
        self.used_precision = _cabi.PRECISION_NAMES.get(info.get("used_precision", -1), "?")
        self._tiled_mask = info.get("tiled_mask", 0)
        self._views = None

    @property
    def corr_pyramid(self):
        """torchvision's ``corr_pyramid``: list of [B*H*W, 1, h_l, w_l] tensors.  Zero-copy views of the packed buffer for
        row-major levels; a level stored as 4x4 tiles is returned as an un-tiled copy."""
        if self._pyramid is None:
            return None
        if self._views is None:
            B, H, W, L, _ = self._meta
            views = _level_views(self._pyramid, B, H, W, L)
            for lvl in range(L):
                if (self._tiled_mask >> lvl) & 1:
                    n, _, h, w = views[lvl].shape
                    views[lvl] = views[lvl].reshape(n, h // 4, w // 4, 4, 4).permute(0, 1, 3, 2, 4).reshape(n, 1, h, w)
            if self._tiled_mask:
                return views           # copies: not cached (the packed buffer may be a caller-owned, reused buffer)
            self._views = views
        return self._views

    @classmethod
    def from_pyramid(cls, levels, radius=4):
        """Wrap an existing pyramid (list of [B*H*W, 1, h_l, w_l] tensors, level 0 first) -- used by the parity
        tests to isolate the lookup from the GEMM precision, and by the torchvision adaptor."""
        self = cls.__new__(cls)
        lvl0 = levels[0]
        H, W = lvl0.shape[-2:]
        B = lvl0.shape[0] // (H * W)
        L = len(levels)
        _cabi.require_cuda(*levels)
        lib = _cabi.lib()
        packed = torch.zeros(lib.cf_corr_pyramid_bytes(B, H, W, L) // 4, dtype=torch.float32, device=lvl0.device)
        self.num_levels, self.radius = L, radius
        self.on_the_fly = False
        self._meta = (B, H, W, L, radius)
        self._pyramid = packed
        self.used_precision = "external"
        self._tiled_mask = 0
        self._views = _level_views(packed, B, H, W, L)
        for view, src in zip(self._views, levels):
            view.copy_(src.reshape(view.shape))
        return self

    def __call__(self, coords, out=None):
        B, H, W, _, _ = self._meta
        if tuple(coords.shape) != (B, 2, H, W):
            raise ValueError(f"coords must be [{B}, 2, {H}, {W}], got {tuple(coords.shape)}")
        if self.on_the_fly:
            if getattr(self, "_token", None) is not None and torch.is_grad_enabled():
                return _OtfLookup.apply(self._token, coords, self, out)
            return self._lookup_on_the_fly(coords, out)
        _cabi.require_cuda(self._pyramid, coords)
        return _CorrLookup.apply(self._pyramid, coords, self._meta, out, self._tiled_mask, getattr(self, "_gstate", None))

    def _lookup_on_the_fly(self, coords, out=None):
        B, H, W, L, r = self._meta
        _cabi.require_cuda(self._f1, coords)
        c = _cabi.f32c(coords)
        shape = (B, L * (2 * r + 1) ** 2, H, W)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=c.device)
        elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous fp32 tensor of shape {shape}")
        with torch.cuda.device(c.device):
            if self._otf_tc:
                _cabi.check(_cabi.lib().cf_corr_lookup_otf(self._operands.data_ptr(), c.data_ptr(), B, self._f1.shape[1], H, W,
                                                           L, r, out.data_ptr(), _cabi.stream()), "cf_corr_lookup_otf")
                return out
            _cabi.check(_cabi.lib().cf_corr_lookup_fused(self._f1.data_ptr(), self._f2_pyramid.data_ptr(), c.data_ptr(),
                                                         B, self._f1.shape[1], H, W, L, r, out.data_ptr(),
                                                         _cabi.stream()), "cf_corr_lookup_fused")
        return out

    # torchvision-style aliases, so the same object can stand in for raft_large().corr_block's methods
    index_pyramid = __call__


def lookup_taps(coords, num_levels=4, radius=4):
    """Parity hook: sampling coordinates / floors the lookup kernel derives.  Returns (taps, idx) with shape
    [L, B*H*W, 2, 2r+1] (axis 2: 0 = x, 1 = y)."""
    B, _, H, W = coords.shape
    _cabi.require_cuda(coords)
    c = _cabi.f32c(coords)
    side = 2 * radius + 1
    taps = torch.empty(num_levels, B * H * W, 2, side, dtype=torch.float32, device=c.device)
    idx = torch.empty(num_levels, B * H * W, 2, side, dtype=torch.int32, device=c.device)
    with torch.cuda.device(c.device):
        _cabi.check(_cabi.lib().cf_corr_lookup_taps(c.data_ptr(), B, H, W, num_levels, radius, taps.data_ptr(),
                                                    idx.data_ptr(), _cabi.stream()), "cf_corr_lookup_taps")
    return taps, idx
