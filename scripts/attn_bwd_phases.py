"""Where the attention-backward kernel's compute warps spend their cycles (per-CTA counters, `dfb_attn_bwd_set_debug`)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200.ops import _lib
from distriflow_b200.ops.attention import _FlashAttention

B, H, T, D = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (32, 12, 1024, 64)))
L = _lib.lib()
L.dfb_attn_bwd_set_debug.restype = ctypes.c_int
L.dfb_attn_bwd_set_debug.argtypes = [ctypes.c_void_p]
dev = torch.device("cuda")
x = torch.randn(B, T, 3 * H * D, device=dev).to(torch.bfloat16)
do = torch.randn(B, T, H * D, device=dev).to(torch.bfloat16)
buf = torch.zeros(8 * 4096, dtype=torch.int64, device=dev)
for i in range(3):
    if i == 2:
        L.dfb_attn_bwd_set_debug(ctypes.c_void_p(buf.data_ptr()))
    xi = x.detach().requires_grad_(True)
    _FlashAttention.apply(xi, H).backward(do)
torch.cuda.synchronize()
L.dfb_attn_bwd_set_debug(None)
d = buf.view(-1, 8)[:148].double()
names = ["kernel", "item_total(loop+first)", "item_tail(drain+epilogue)", "first_tile", "tail_wait(dq+dkv)", "tiles", "items"]
for k, n in enumerate(names):
    print(f"{n:28s} mean {d[:, k].mean():12.0f}  max {d[:, k].max():12.0f}  min {d[:, k].min():12.0f}")
tiles, items = d[:, 5].mean(), d[:, 6].mean()
print(f"per tile (loop incl. first): {d[:, 1].mean() / tiles:.0f} cycles; first tile: {d[:, 3].mean() / items:.0f}; "
      f"tail per item: {d[:, 2].mean() / items:.0f} (of which waiting {d[:, 4].mean() / items:.0f})")
