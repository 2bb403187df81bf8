"""Causal attention micro-benchmark at the GPT-2 small shape (B 32, H 12, T 1024, D 64, bf16): the tcgen05 kernels of
`ops/csrc/attention_sm100.cu` on the packed QKV tensor against `F.scaled_dot_product_attention` (cuDNN / flash library
kernels on contiguous-per-head views).  CUDA events, 3 warm-up + N timed iterations over rotating inputs (3 x 151 MB > L2),
clocks + throttle reasons sampled during the run.  FLOPs counted for the causal half: fwd 2*B*H*T*T*D, bwd 2.5x."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from distriflow_b200.ops.attention import _FlashAttention
from distriflow_b200.utils.clocks import ClockSampler


def timeit(fn, n=20, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    B, H, T, D = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (32, 12, 1024, 64)))
    dev = torch.device("cuda")
    xs = [torch.randn(B, T, 3 * H * D, device=dev).to(torch.bfloat16) for _ in range(3)]
    dos = [torch.randn(B, T, H * D, device=dev).to(torch.bfloat16) for _ in range(3)]
    flops_f = 2.0 * B * H * T * T * D            # causal: half of 4*B*H*T*T*D
    sampler = ClockSampler(0).start()
    rows = {}

    def ours_fwd(i):
        with torch.no_grad():
            _FlashAttention.apply(xs[i % 3], H)

    def sdpa_fwd(i):
        with torch.no_grad():
            q, k, v = xs[i % 3].view(B, T, 3, H, D).permute(2, 0, 3, 1, 4)
            F.scaled_dot_product_attention(q, k, v, is_causal=True)

    def make_fb(kind):
        def run(i):
            x = xs[i % 3].detach().requires_grad_(True)
            if kind == "ours":
                y = _FlashAttention.apply(x, H)
            else:
                q, k, v = x.view(B, T, 3, H, D).permute(2, 0, 3, 1, 4)
                y = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, H * D)
            y.backward(dos[i % 3])
        return run

    rows["ours_fwd_ms"] = timeit(ours_fwd)
    rows["sdpa_fwd_ms"] = timeit(sdpa_fwd)
    rows["ours_fwd_bwd_ms"] = timeit(make_fb("ours"))
    rows["sdpa_fwd_bwd_ms"] = timeit(make_fb("sdpa"))
    rows["ours_bwd_ms"] = rows["ours_fwd_bwd_ms"] - rows["ours_fwd_ms"]
    rows["sdpa_bwd_ms"] = rows["sdpa_fwd_bwd_ms"] - rows["sdpa_fwd_ms"]
    out = {"shape": {"B": B, "H": H, "T": T, "D": D}, **{k: round(v, 4) for k, v in rows.items()},
           "ours_fwd_tflops": round(flops_f / rows["ours_fwd_ms"] / 1e9, 1), "sdpa_fwd_tflops": round(flops_f / rows["sdpa_fwd_ms"] / 1e9, 1),
           "ours_bwd_tflops": round(2.5 * flops_f / rows["ours_bwd_ms"] / 1e9, 1),
           "sdpa_bwd_tflops": round(2.5 * flops_f / rows["sdpa_bwd_ms"] / 1e9, 1),
           "fwd_speedup": round(rows["sdpa_fwd_ms"] / rows["ours_fwd_ms"], 3),
           "fwd_bwd_speedup": round(rows["sdpa_fwd_bwd_ms"] / rows["ours_fwd_bwd_ms"], 3),
           "note": "sdpa bwd includes the library's own packing of dq/dk/dv into the packed view's gradient",
           "clocks": sampler.stop()}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
