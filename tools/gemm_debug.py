import sys, os
# needs a bring-up build: CF_NVCC_EXTRA=-DCF_BRINGUP python cardiac-segmentation-optical-flow_b200/_build.py (after touching corr_tcgen05.cu and cabi.cu);
# the shipped library does not export cf_debug_set_gemm_*
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, ctypes
import cardioflow_b200 as cf
dev = "cuda:0"
torch.manual_seed(0)
B, C, H, W = 1, 32, 16, 16
HW = H * W
f1 = torch.randn(B, C, H, W, device=dev); f2 = torch.randn(B, C, H, W, device=dev)
lib = cf._cabi.lib()
dbg = torch.zeros(16384 + C * 128 + 128 * 32 + 16, device=dev)
lib.cf_debug_set_gemm_dump.argtypes = [ctypes.c_void_p]
lib.cf_debug_set_gemm_dump(dbg.data_ptr())
blk = cf.CorrBlock(f1, f2, precision="tf32")
torch.cuda.synchronize()
lib.cf_debug_set_gemm_dump(None)
acc = dbg[:16384].view(128, 128).cpu()
sa = dbg[16384:16384 + C * 128].cpu()
sb = dbg[16384 + C * 128:16384 + C * 128 + 4096].cpu()
print("tmem_base bits:", hex(dbg[16384 + C * 128 + 4096].view(torch.int32).item()))
ref = torch.matmul(f1.view(C, HW).t(), f2.view(C, HW)).cpu()   # unscaled
print("acc absmax", acc.abs().max().item(), "ref[:128,:128] absmax", ref[:128, :128].abs().max().item())
print("acc vs ref maxdiff", (acc - ref[:128, :128]).abs().max().item())
print("acc[0,:8]", acc[0, :8].tolist()); print("ref[0,:8]", ref[0, :8].tolist())
# expected smem A: chunk mc, row k, 16B group g, elem e  holds f1[k, mc*32 + 4*(g ^ (k%8)) + e]
A = f1.view(C, HW).cpu()
exp = torch.zeros(4, C, 32)
for mc in range(4):
    for k in range(C):
        for g in range(8):
            src = g ^ (k % 8)
            exp[mc, k, 4 * g:4 * g + 4] = A[k, mc * 32 + 4 * src: mc * 32 + 4 * src + 4]
print("smem A nonzero frac", (sa != 0).float().mean().item(), "matches swizzled expectation:", torch.equal(sa.view(4, C, 32), exp))
print("smem A matches unswizzled:", torch.equal(sa.view(4, C, 32), A[:, :128].view(C, 4, 32).permute(1, 0, 2).contiguous()))
Bm = f2.view(C, HW).cpu()
expb = torch.zeros(4, 32, 32)
for nc in range(4):
    for k in range(32):
        for g in range(8):
            src = g ^ (k % 8)
            expb[nc, k, 4 * g:4 * g + 4] = Bm[k, nc * 32 + 4 * src: nc * 32 + 4 * src + 4]
print("smem B nonzero frac", (sb != 0).float().mean().item(), "matches swizzled expectation:", torch.equal(sb.view(4, 32, 32), expb))
print("L0 absmax", blk.corr_pyramid[0].abs().max().item())
