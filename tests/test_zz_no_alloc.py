"""cf_corr_volume_bwd_noalloc: the tensor-core correlation-volume backward with its scratch in a static device array
(include/cardioflow.h; no counterpart in the reference -- added for callers that cannot
allocate: CUDA graph capture, tight serving loops).  Collected last on purpose (file name): a failure here hides no other test.

Checked through the C ABI against the existing entry point cf_corr_volume_bwd_ex (same kernel, caller-provided workspace) on two
calls in a row -- the second with a smaller problem, so the shared scratch still holds the first call's planes -- within the
tolerance of tests/test_gpu_bwd_tc.py (2e-5, norm max|a-b| / max|b|; the two paths run the same code, so they also agree bit for bit).
"""
import ctypes

import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SCRATCH_BYTES = 8 << 20          # CF_NOALLOC_SCRATCH_BYTES in include/cardioflow.h


def _lib():
    from cardioflow_b200 import _cabi
    lib = _cabi.lib()
    lib.cf_corr_volume_bwd_noalloc.restype = ctypes.c_int
    lib.cf_corr_volume_bwd_noalloc.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 4 + [ctypes.c_void_p] * 3
    return _cabi, lib


def _inputs(shape, seed):
    B, C, H, W = shape
    g = torch.Generator().manual_seed(seed)
    f1, f2 = torch.randn(B, C, H, W, generator=g), torch.randn(B, C, H, W, generator=g)
    grad = torch.randn(B, H * W, H * W, generator=g)
    return f1.to(DEV), f2.to(DEV), grad.to(DEV)


def _existing(_cabi, lib, f1, f2, grad):
    B, C, H, W = f1.shape
    wsb = lib.cf_corr_volume_bwd_workspace_bytes(B, C, H, W)
    ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
    o1, o2 = torch.empty_like(f1), torch.empty_like(f2)
    used = ctypes.c_int(-1)
    with torch.cuda.device(0):
        _cabi.check(lib.cf_corr_volume_bwd_ex(f1.data_ptr(), f2.data_ptr(), grad.data_ptr(), B, C, H, W, _cabi.PRECISIONS["bf16x3"],
                                              o1.data_ptr(), o2.data_ptr(), ws.data_ptr(), wsb, ctypes.byref(used), _cabi.stream()),
                    "cf_corr_volume_bwd_ex")
    torch.cuda.synchronize()
    return o1, o2


def test_noalloc_entry_point_matches_the_existing_one_on_two_calls_in_a_row():
    _cabi, lib = _lib()
    for shape, seed in (((3, 128, 32, 32), 1), ((2, 40, 16, 24), 2)):      # 3 MiB of planes, then 0.24 MiB: the scratch is reused
        f1, f2, grad = _inputs(shape, seed)
        B, C, H, W = shape
        want1, want2 = _existing(_cabi, lib, f1, f2, grad)
        # outputs are allocated BEFORE the call; the call itself allocates nothing (checked with torch's allocator counters below)
        o1, o2 = torch.full_like(f1, 7.0), torch.full_like(f2, 7.0)
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats(0)["allocation.all.allocated"]
        with torch.cuda.device(0):
            _cabi.check(lib.cf_corr_volume_bwd_noalloc(f1.data_ptr(), f2.data_ptr(), grad.data_ptr(), B, C, H, W,
                                                       o1.data_ptr(), o2.data_ptr(), _cabi.stream()), "cf_corr_volume_bwd_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(0)["allocation.all.allocated"] == before
        e1, e2 = rel_err(o1.cpu().numpy(), want1.cpu().numpy()), rel_err(o2.cpu().numpy(), want2.cpu().numpy())
        print(f"[no-alloc] {shape}: rel err vs cf_corr_volume_bwd_ex gf1 {e1:.2e} gf2 {e2:.2e}")
        assert e1 <= 2e-5 and e2 <= 2e-5
        assert torch.equal(o1, want1) and torch.equal(o2, want2)            # same kernel, same operands


def test_noalloc_entry_point_rejects_what_does_not_fit():
    _cabi, lib = _lib()
    B, C, H, W = 5, 256, 32, 32                                              # 4 planes x 2.5 MiB = 10 MiB > 8 MiB
    assert lib.cf_corr_volume_bwd_workspace_bytes(B, C, H, W) > SCRATCH_BYTES
    f1, f2, grad = _inputs((B, C, H, W), 3)
    o1 = torch.full_like(f1, 7.0)
    with torch.cuda.device(0):
        rc = lib.cf_corr_volume_bwd_noalloc(f1.data_ptr(), f2.data_ptr(), grad.data_ptr(), B, C, H, W, o1.data_ptr(), None, _cabi.stream())
    torch.cuda.synchronize()
    assert rc != 0 and b"scratch" in lib.cf_last_error()
    assert bool((o1 == 7.0).all())                                           # nothing was launched
    f1, f2, grad = _inputs((1, 16, 5, 5), 4)                                 # H*W = 25: no tensor-core path, hence no scratch variant
    with torch.cuda.device(0):
        rc = lib.cf_corr_volume_bwd_noalloc(f1.data_ptr(), f2.data_ptr(), grad.data_ptr(), 1, 16, 5, 5, f1.data_ptr(), None, _cabi.stream())
    assert rc != 0
