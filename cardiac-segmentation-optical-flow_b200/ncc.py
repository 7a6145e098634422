"""Local normalised cross-correlation loss on the warped frame.

Mirrors ``class NCC`` of the reference (/root/reference/nnunet/lib/loss.py:15-76; built by the trainers at
training/network_training/SegFlowGaussian.py:337, FinalFlowRaft.py:259, nnMTLTrainerV2Raft.py:275 and by
network_architecture/Optical_flow_model_successive.py:71): same constructor ``NCC(win=9, eps=1e-3, reduction='mean')``,
``forward(I, J)`` on ``[N,1,H,W]`` or ``[T,B,1,H,W]`` single-channel images, ``1 - mean(cc)`` for ``reduction='mean'`` and
the ``1 - cc`` map reshaped to the 5-D input otherwise.  The five 9x9 box sums, the cc formula, the clamp and the mean are
one CUDA launch behind ``cf_ncc`` (plus a fixed-order reduction of per-task partial sums); autograd through ``cf_ncc_bwd``.
"""
import torch

from . import _cabi


class _NCC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, I, J, win, eps, mean):
        i, j = _cabi.f32c(I), _cabi.f32c(J)
        N, _, H, W = i.shape
        lib = _cabi.lib()
        with torch.cuda.device(i.device):
            if mean:
                nws = lib.cf_ncc_workspace_bytes(N, H, W, 0)
                ws = torch.empty(nws // 4 + 1, dtype=torch.float32, device=i.device)      # [partials..., cc_sum]
                cc_sum = ws[nws // 4:]
                _cabi.check(lib.cf_ncc(i.data_ptr(), j.data_ptr(), N, H, W, win, eps, None, cc_sum.data_ptr(), ws.data_ptr(),
                                       nws, _cabi.stream()), "cf_ncc")
                out = (1 - cc_sum / float(N * H * W)).reshape(())
            else:
                out = torch.empty(N, 1, H, W, dtype=torch.float32, device=i.device)
                _cabi.check(lib.cf_ncc(i.data_ptr(), j.data_ptr(), N, H, W, win, eps, out.data_ptr(), None, None, 0,
                                       _cabi.stream()), "cf_ncc")
        ctx.save_for_backward(i, j)
        ctx.cfg = (win, eps, mean)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        i, j = ctx.saved_tensors
        win, eps, mean = ctx.cfg
        N, _, H, W = i.shape
        g = grad_out.contiguous().float()
        gi = torch.empty_like(i) if ctx.needs_input_grad[0] else None
        gj = torch.empty_like(j) if ctx.needs_input_grad[1] else None
        if gi is None and gj is None:
            return None, None, None, None, None
        lib = _cabi.lib()
        with torch.cuda.device(i.device):
            nws = lib.cf_ncc_workspace_bytes(N, H, W, 1)
            ws = torch.empty(nws // 4, dtype=torch.float32, device=i.device)
            _cabi.check(lib.cf_ncc_bwd(i.data_ptr(), j.data_ptr(), N, H, W, win, eps, g.data_ptr(), 0 if mean else 1,
                                       -1.0 / (N * H * W) if mean else -1.0,
                                       gi.data_ptr() if gi is not None else None, gj.data_ptr() if gj is not None else None,
                                       ws.data_ptr(), nws, _cabi.stream()), "cf_ncc_bwd")
        return gi, gj, None, None, None


class NCC(torch.nn.Module):
    """local (over window) normalized cross correlation -- drop-in for nnunet.lib.loss.NCC."""

    def __init__(self, win=9, eps=1e-3, reduction="mean"):
        super().__init__()
        self.win_raw = win
        self.win = win
        self.eps = eps
        self.reduction = reduction

    def forward(self, I, J):
        if I.shape != J.shape:
            raise ValueError(f"NCC: I {tuple(I.shape)} and J {tuple(J.shape)} differ")
        shape5 = None
        if I.dim() > 4:
            shape5 = I.shape
            T, B, C, H, W = I.shape
            I, J = I.reshape(T * B, C, H, W), J.reshape(T * B, C, H, W)
        if I.dim() != 4 or I.shape[1] != 1:
            raise ValueError(f"NCC: single-channel [N,1,H,W] or [T,B,1,H,W] images expected (the reference's ones kernel has one "
                             f"input channel), got {tuple(I.shape)}")
        mean = self.reduction == "mean"
        if not mean and shape5 is None:
            raise ValueError("NCC: reduction != 'mean' needs a 5-D [T,B,C,H,W] input (the reference reshapes to T, B, C, "
                             "which exist only on that branch)")
        _cabi.require_cuda(I, J)
        out = _NCC.apply(I, J, int(self.win_raw), float(self.eps), mean)
        return out if mean else out.view(shape5)
