"""ncu target: a streaming read+write kernel (b = a + 1 over 1 GiB) to calibrate what lts__throughput reads at the HBM roofline."""
import torch
a = torch.empty(256 << 20, dtype=torch.float32, device="cuda")   # 1 GiB
b = torch.empty_like(a)
for _ in range(3):
    torch.add(a, 1.0, out=b)
torch.cuda.synchronize()
