"""Short driver for ncu: a handful of launches of the tcgen05 GEMM (fc1 forward / dgrad / wgrad at 16K tokens),
the fused softmax-CE and the fused bucket kernel (world = 1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200.ops.dense import _init_sigs, gemm
from distriflow_b200.ops.softmax_ce import softmax_cross_entropy

_init_sigs()
dev = torch.device("cuda")
T, K, N = 16384, 768, 3072
x = torch.randn(T, K, device=dev).bfloat16()
dy = torch.randn(T, N, device=dev).bfloat16()
w = torch.randn(K, N, device=dev).bfloat16()
for _ in range(3):
    gemm(x, w, a_mn=False, b_mn=True, M=T, N=N, K=K)
    gemm(dy, w, a_mn=False, b_mn=False, M=T, N=K, K=N)
    gemm(x, dy, a_mn=True, b_mn=True, M=K, N=N, K=T)
logits = torch.randn(4096, 50304, device=dev).bfloat16().requires_grad_(True)
lab = torch.randint(0, 50257, (4096,), device=dev)
for _ in range(2):
    softmax_cross_entropy(logits, lab).backward()
torch.cuda.synchronize()
print("ok")
