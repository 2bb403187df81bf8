"""One forward + one backward of the attention kernels at the GPT-2 shape, for `ncu -k regex:attn_`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200.ops.attention import _FlashAttention

B, H, T, D = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (32, 12, 1024, 64)))
dev = torch.device("cuda")
x = torch.randn(B, T, 3 * H * D, device=dev).to(torch.bfloat16)
do = torch.randn(B, T, H * D, device=dev).to(torch.bfloat16)
for _ in range(2):
    xi = x.detach().requires_grad_(True)
    _FlashAttention.apply(xi, H).backward(do)
torch.cuda.synchronize()
print("ok")
