import os, sys
sys.path.insert(0, "/root/repo")
import torch
from raydp_b200 import ops
dev = torch.device("cuda", 0)
B = 131072
for K, N in ((64, 256),):
    x = torch.randn(B, K, device=dev).to(torch.bfloat16)
    w = torch.randn(N, K, device=dev).to(torch.bfloat16)
    out = torch.empty(B, N, dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        ops.gemm_tn(x, w, out)
    torch.cuda.synchronize()
