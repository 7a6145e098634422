"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the reference's local normalised cross-correlation loss.

Follows /root/reference/nnunet/lib/loss.py:15-76 (`class NCC`): five zero-padded 9x9 box sums (there: conv2d with a ones
kernel, padding win // 2) of I, J, I*I, J*J, I*J, then per pixel

    u_I = I_sum / 81,  u_J = J_sum / 81
    cross = IJ_sum - u_J * I_sum - u_I * J_sum + u_I * u_J * 81
    I_var = I2_sum - 2 * u_I * I_sum + u_I * u_I * 81          (J_var alike)
    cc    = clamp(cross * cross / (I_var * J_var + eps), 0.001, 0.999)

and `1 - mean(cc)` (reduction='mean') or `1 - cc` reshaped to the 5-D input (any other reduction; 5-D input only, as in
the reference where T, B, C exist only on that branch).  Inputs are single-channel ([N,1,H,W] or [T,B,1,H,W]: the ones
kernel has one input channel).  Pinned against the reference class itself (tests/golden/ncc_reference.npz, generated
by oracle/make_golden.py from the class compiled out of its source file).

The variance terms cancel catastrophically in fp32 (I2_sum ~ 81 * mean^2 against I_var ~ 81 * var), so two correct fp32
evaluations with different summation orders differ by ~1e-6 * mean^2 / var relative; `cc64` is the float64 evaluation
used to bound both the CUDA kernel's and the reference's own distance from the exact value.
Only tests/, smoke() and bench.py's CPU legs may import this module.
"""
import torch
import torch.nn.functional as F


def _sums(I, J, win):
    w = torch.ones(1, 1, win, win, dtype=I.dtype)
    p = win // 2
    return tuple(F.conv2d(t, w, padding=p) for t in (I, J, I * I, J * J, I * J))


def cc_map(I, J, win=9, eps=1e-3):
    """Clamped local cc map [N,1,H,W] of two [N,1,H,W] images (loss.py:43-66)."""
    I_sum, J_sum, I2_sum, J2_sum, IJ_sum = _sums(I, J, win)
    n = float(win * win)
    u_I, u_J = I_sum / n, J_sum / n
    cross = IJ_sum - u_J * I_sum - u_I * J_sum + u_I * u_J * n
    I_var = I2_sum - 2 * u_I * I_sum + u_I * u_I * n
    J_var = J2_sum - 2 * u_J * J_sum + u_J * u_J * n
    return torch.clamp(cross * cross / (I_var * J_var + eps), 0.001, 0.999)


def cc64(I, J, win=9, eps=1e-3):
    return cc_map(I.double(), J.double(), win, eps)


class NCC(torch.nn.Module):
    """Same constructor / forward contract as loss.py:15-76."""

    def __init__(self, win=9, eps=1e-3, reduction="mean"):
        super().__init__()
        self.win, self.eps, self.reduction = win, eps, reduction

    def forward(self, I, J):
        shape5 = None
        if I.dim() > 4:
            shape5 = I.shape
            T, B, C, H, W = I.shape
            I, J = I.reshape(T * B, C, H, W), J.reshape(T * B, C, H, W)
        cc = cc_map(I, J, self.win, self.eps)
        if self.reduction == "mean":
            return 1 - torch.mean(cc)
        if shape5 is None:
            raise ValueError("reduction != 'mean' needs a 5-D [T,B,C,H,W] input (as in the reference)")
        return 1 - cc.view(shape5)
