"""Small-shape pass over the hand-written kernels for compute-sanitizer (memcheck / racecheck / synccheck, one tool per run)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200 import ops
from distriflow_b200.ops.dense import _init_sigs, gemm
from distriflow_b200.ops.softmax_ce import softmax_cross_entropy

_init_sigs()
dev = torch.device("cuda")
torch.manual_seed(0)
a = torch.randn(300, 264, device=dev).bfloat16()
b = torch.randn(264, 200, device=dev).bfloat16()
ref = a.float() @ b.float()
for bn in (128, 256, 512):
    out = gemm(a, b, a_mn=False, b_mn=True, M=300, N=200, K=264, block_n=bn, splits=1)
    assert float((out.float() - ref).abs().max() / ref.abs().max()) < 1e-2
at = torch.randn(1024, 256, device=dev).bfloat16()
bt = torch.randn(1024, 384, device=dev).bfloat16()
out = gemm(at, bt, a_mn=True, b_mn=True, M=256, N=384, K=1024, block_n=512, splits=4)
assert float((out.float() - at.float().t() @ bt.float()).abs().max()) < 1.0
x32 = torch.randn(32, 784, device=dev, requires_grad=True)
w32 = torch.randn(784, 10, device=dev, requires_grad=True)
ops.linear(x32, w32, None, "relu").sum().backward()
lg = torch.randn(16, 1003, device=dev, requires_grad=True)
softmax_cross_entropy(lg, torch.randint(0, 1003, (16,), device=dev)).backward()
xs = torch.randn(64, 256, device=dev).bfloat16().requires_grad_(True)
g = torch.ones(256, device=dev).bfloat16().requires_grad_(True)
bb = torch.zeros(256, device=dev).bfloat16().requires_grad_(True)
s, y = ops.add_layer_norm(xs, xs.detach(), g, bb)
(s.float().sum() + y.float().sum()).backward()
xc = torch.randn(2, 14, 14, 64, device=dev).bfloat16().requires_grad_(True)
wc = (torch.randn(3, 3, 64, 64, device=dev) * 0.05).bfloat16().requires_grad_(True)
ops.conv2d_nhwc(xc, wc, None, (2, 2), "same", "relu").float().sum().backward()
q = torch.randn(2, 64, 3 * 2 * 64, device=dev).bfloat16().requires_grad_(True)
ops.causal_self_attention(q, 2).float().sum().backward()
# fused BatchNorm + residual + ReLU (forward, backward, inference), fused GELU epilogues, packed attention gradient
xb = torch.randn(4, 6, 6, 16, device=dev).bfloat16().requires_grad_(True)
rb = torch.randn(4, 6, 6, 16, device=dev).bfloat16().requires_grad_(True)
gb = torch.ones(16, device=dev).bfloat16().requires_grad_(True)
bb = torch.zeros(16, device=dev).bfloat16().requires_grad_(True)
rm, rv = torch.zeros(16, device=dev), torch.ones(16, device=dev)
ops.batch_norm_nhwc(xb, gb, bb, rm, rv, True, 0.1, 1e-5, "relu", rb).float().sum().backward()
with torch.no_grad():
    ops.batch_norm_nhwc(xb, gb, bb, rm, rv, False, 0.1, 1e-5, "relu", rb)
xm = torch.randn(256, 64, device=dev).bfloat16().requires_grad_(True)
w1 = (torch.randn(64, 256, device=dev) * 0.05).bfloat16().requires_grad_(True)
w2 = (torch.randn(256, 64, device=dev) * 0.05).bfloat16().requires_grad_(True)
b1 = torch.zeros(256, device=dev).bfloat16().requires_grad_(True)
b2 = torch.zeros(64, device=dev).bfloat16().requires_grad_(True)
ops.mlp(xm, w1, b1, w2, b2).float().sum().backward()
qkv = torch.randn(2, 16, 3 * 4 * 16, device=dev).bfloat16().requires_grad_(True)
ops.causal_self_attention(qkv, 4).float().sum().backward()

# collectives with two emulated ranks
from distriflow_b200.ops import _lib
from distriflow_b200.ops._lib import check, ptr
from distriflow_b200.parallel.async_engine import AsyncShardState
from distriflow_b200.parallel.engine import _comm_sigs
from distriflow_b200.parallel.symm import SymmetricHeap
from distriflow_b200.parallel.workqueue import DeviceWorkQueue
L = _comm_sigs()
n, R = 64 * 16, 2
heaps = SymmetricHeap.emulated(R, n * 8 + 65536)
go = [h.alloc(n * 2) for h in heaps][0]
po = [h.alloc(n * 2) for h in heaps][0]
for h in heaps:
    h.view(go, (n,), torch.bfloat16).normal_()
for r, h in enumerate(heaps):
    m0, m1, m2 = (torch.zeros(n // R, device=dev) for _ in range(3))
    check(L.dfb_sync_rs_opt_ag(h.ptrs(go), h.ptrs(po), None, ptr(m0), ptr(m1), ptr(m2), None, n, n // R, r, R, 0, 1, 1, 1, 2, 1e-3, 0.9,
                               0.9, 0.999, 1e-8, 0.0, 1, 4, 0, 0, _lib.stream()), "sync")
st = AsyncShardState.emulated(2, n, grad_dtype=torch.float32, param_dtype=torch.float32, lr=0.1, max_staleness=2)
st.init_weights(torch.randn(n, device=dev))
for w in range(2):
    st.pull(w)
    st.push(w, torch.randn(n, device=dev))
for o in range(2):
    st.apply(o)
wq = DeviceWorkQueue.local(5, 1, device=dev)
t = wq.take()
wq.complete(t)
torch.cuda.synchronize()
print("sanitize script ok")
