"""Short driver for ncu: a few launches of every named hand-written kernel at flagship shapes (GPT-2 small, 32K tokens;
ResNet-50 stage shapes) so that one `ncu --set full` capture covers them all."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200 import ops
from distriflow_b200.ops.dense import _init_sigs, gemm
from distriflow_b200.ops.softmax_ce import softmax_cross_entropy

_init_sigs()
dev = torch.device("cuda")
T, K, N = 32768, 768, 3072
x = torch.randn(T, K, device=dev).bfloat16()
dy = torch.randn(T, N, device=dev).bfloat16()
w = torch.randn(K, N, device=dev).bfloat16()
for _ in range(2):
    gemm(x, w, a_mn=False, b_mn=True, M=T, N=N, K=K)            # fc1 forward  (A K-major, B MN-major)
    gemm(dy, w, a_mn=False, b_mn=False, M=T, N=K, K=N)          # fc1 dgrad    (both K-major)
    gemm(x, dy, a_mn=True, b_mn=True, M=K, N=N, K=T)            # fc1 wgrad    (both MN-major)
    gemm(x, x, a_mn=True, b_mn=True, M=K, N=K, K=T)             # proj wgrad   (split-K + reduce)
# Dense layer with bias + GELU, forward and backward: GEMM epilogue, act_kernel, colsum_kernel
xl = torch.randn(T, K, device=dev).bfloat16().requires_grad_(True)
wl = (torch.randn(K, N, device=dev) * 0.02).bfloat16().requires_grad_(True)
bl = torch.zeros(N, device=dev).bfloat16().requires_grad_(True)
ops.linear(xl, wl, bl, activation="gelu").float().mean().backward()
# attention wrapper: strided views into the library SDPA + the gradient pack kernel
qkv = torch.randn(32, 1024, 3 * 768, device=dev).bfloat16().requires_grad_(True)
ops.causal_self_attention(qkv, 12).float().mean().backward()
logits = torch.randn(8192, 50304, device=dev).bfloat16().requires_grad_(True)
lab = torch.randint(0, 50257, (8192,), device=dev)
softmax_cross_entropy(logits, lab).backward()
# fused add + LayerNorm
xs = torch.randn(T, K, device=dev).bfloat16().requires_grad_(True)
rs = torch.randn(T, K, device=dev).bfloat16().requires_grad_(True)
g = torch.ones(K, device=dev).bfloat16().requires_grad_(True)
b = torch.zeros(K, device=dev).bfloat16().requires_grad_(True)
s, y = ops.add_layer_norm(xs, rs, g, b)
(s.float().mean() + y.float().mean()).backward()
# implicit-GEMM conv (ResNet-50 stage 2: 28x28x128, 3x3) fwd / dgrad / wgrad
xc = torch.randn(64, 28, 28, 128, device=dev).bfloat16().requires_grad_(True)
wc = (torch.randn(3, 3, 128, 128, device=dev) * 0.03).bfloat16().requires_grad_(True)
ops.conv2d_nhwc(xc, wc, None, (1, 1), "same", "relu").float().mean().backward()
# fused BatchNorm + residual + ReLU, ResNet-50 stage-1 shape (64 x 56 x 56 x 256), forward and backward
xb = torch.randn(64, 56, 56, 256, device=dev).bfloat16().requires_grad_(True)
rb = torch.randn(64, 56, 56, 256, device=dev).bfloat16().requires_grad_(True)
gb = torch.ones(256, device=dev).bfloat16().requires_grad_(True)
bb = torch.zeros(256, device=dev).bfloat16().requires_grad_(True)
ops.batch_norm_nhwc(xb, gb, bb, torch.zeros(256, device=dev), torch.ones(256, device=dev), True, 0.1, 1e-5, "relu", rb).float().mean().backward()
# fused bucket kernel, world = 1 (optimizer half of the fused gradient path), 32 MB bucket, Adam
from distriflow_b200.ops import _lib
from distriflow_b200.ops._lib import check, ptr
from distriflow_b200.parallel.engine import _comm_sigs
from distriflow_b200.parallel.symm import SymmetricHeap
L = _comm_sigs()
n = 16 << 20
heap = SymmetricHeap(n * 4 + (1 << 20))
go, po = heap.alloc(n * 2), heap.alloc(n * 2)
heap.view(go, (n,), torch.bfloat16).normal_()
master, m, v = (torch.zeros(n, device=dev) for _ in range(3))
for step in (1, 2):
    check(L.dfb_sync_rs_opt_ag(heap.ptrs(go), heap.ptrs(po), None, ptr(master), ptr(m), ptr(v), None, n, n, 0, 1, 0, 1, 1, 1, 2, 1e-3,
                               0.9, 0.9, 0.999, 1e-8, 0.0, step, 592, 0, 0, _lib.stream()), "sync")
torch.cuda.synchronize()
print("ok")
