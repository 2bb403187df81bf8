import sys, os
sys.path.insert(0, os.getcwd())
import torch
from distriflow_b200.ops.dense import _init_sigs, gemm, gemm_aux, act_op
from distriflow_b200.utils.clocks import ClockSampler
sampler = ClockSampler(0).start()
_init_sigs()
dev = torch.device("cuda")
M, K, H = 32768, 768, 3072
x = torch.randn(M, K, device=dev).bfloat16(); w1 = torch.randn(K, H, device=dev).bfloat16(); b1 = torch.randn(H, device=dev).bfloat16()
dy = torch.randn(M, K, device=dev).bfloat16(); w2 = torch.randn(H, K, device=dev).bfloat16()
pre = torch.empty(M, H, device=dev, dtype=torch.bfloat16)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n*1e3
print("fwd fused (aux store)   us", t(lambda: gemm_aux(x, w1, a_mn=False, b_mn=True, M=M, N=H, K=K, aux=pre, aux_mode=1, bias=b1, act="gelu")))
print("fwd gemm + act kernel   us", t(lambda: act_op(gemm(x, w1, a_mn=False, b_mn=True, M=M, N=H, K=K, bias=b1), None, 0)))
print("fwd gemm only           us", t(lambda: gemm(x, w1, a_mn=False, b_mn=True, M=M, N=H, K=K, bias=b1)))
print("bwd fused (aux gelu')   us", t(lambda: gemm_aux(dy, w2, a_mn=False, b_mn=False, M=M, N=H, K=K, aux=pre, aux_mode=2)))
print("bwd gemm + act kernel   us", t(lambda: act_op(gemm(dy, w2, a_mn=False, b_mn=False, M=M, N=H, K=K), pre, 1)))
print("bwd gemm only           us", t(lambda: gemm(dy, w2, a_mn=False, b_mn=False, M=M, N=H, K=K)))
print("fwd gemm, no bias       us", t(lambda: gemm(x, w1, a_mn=False, b_mn=True, M=M, N=H, K=K)))
import json
print(json.dumps({"clocks": sampler.stop()}))
