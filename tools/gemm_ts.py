"""Bring-up: per-role timeline of one CTA of the tcgen05 correlation kernel (globaltimer ns).  GPU box only."""
# needs a bring-up build: CF_NVCC_EXTRA=-DCF_BRINGUP python cardiac-segmentation-optical-flow_b200/_build.py (after touching corr_tcgen05.cu and cabi.cu);
# the shipped library does not export cf_debug_set_gemm_*
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cardioflow_b200 as cf
from cardioflow_b200 import _cabi
lib = _cabi.lib()
lib.cf_debug_set_gemm_timestamps.argtypes = [ctypes.c_void_p, ctypes.c_int]
dev = "cuda:0"
f1, f2 = torch.randn(29, 256, 32, 32, device=dev), torch.randn(29, 256, 32, 32, device=dev)
for _ in range(3):
    cf.CorrBlock(f1, f2, precision="bf16")
torch.cuda.synchronize()
for cta in [int(a) for a in sys.argv[1:]] or [0, 100, 200]:
    ts = torch.zeros(128, dtype=torch.int64, device=dev)
    lib.cf_debug_set_gemm_timestamps(ts.data_ptr(), cta)
    cf.CorrBlock(f1, f2, precision="bf16")
    torch.cuda.synchronize()
    lib.cf_debug_set_gemm_timestamps(None, 0)
    t = ts.cpu().tolist()
    t0 = t[0]
    rel = lambda x: (x - t0) / 1e3 if x else float("nan")
    print(f"CTA {cta}: A landed +{rel(t[1]):.2f} us")
    for k in range(8):
        if not t[32 + k]:
            break
        print(f"  tile {k}: prod kb0 wait done {rel(t[16+k]):6.2f} | mma start {rel(t[32+k]):6.2f} issued {rel(t[48+k]):6.2f} | "
              f"epi wait {rel(t[64+4*k]):6.2f} accfull {rel(t[65+4*k]):6.2f} bufs free {rel(t[66+4*k]):6.2f} stores issued {rel(t[67+4*k]):6.2f}")
