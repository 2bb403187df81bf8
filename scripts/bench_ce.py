"""Softmax-CE kernel timing at the GPT-2 LM-head shape (8192 x 50304 bf16 chunk), device-timed, inputs rotated past L2."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes

import torch

from distriflow_b200.ops import _lib, softmax_ce
from distriflow_b200.ops._lib import check, ptr, stream
from distriflow_b200.utils.clocks import ClockSampler

dev = torch.device("cuda")
rows, C = 8192, 50304
bufs = [torch.randn(rows, C, device=dev).bfloat16() for _ in range(3)]
lab = torch.randint(0, 50257, (rows,), device=dev)
loss = torch.empty(rows, device=dev)
L = _lib.lib()
softmax_ce._sigs()


def run(i):
    x = bufs[i % 3]
    check(L.dfb_softmax_ce_index(ptr(x), ptr(lab), 0, ptr(loss), ptr(x), rows, C, ctypes.c_long(C), ctypes.c_long(C), 1,
                                 ctypes.c_float(1.0 / rows), ctypes.c_long(-100), stream()), "ce")


for i in range(3):
    run(i)
torch.cuda.synchronize()
sampler = ClockSampler(0, period_ms=50).start()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(240):
    run(i)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 240 * 1e3
print(f"softmax_ce fwd+bwd in place {rows}x{C} bf16: {us:.1f} us, {2 * rows * C * 2 / us / 1e3:.0f} GB/s of the 1R+1W minimum")
import json
print(json.dumps({"kernel": "softmax_ce_index", "rows": rows, "cols": C, "us": round(us, 1),
                  "gbs_of_1r1w": round(2 * rows * C * 2 / us / 1e3), "clocks": sampler.stop()}))
