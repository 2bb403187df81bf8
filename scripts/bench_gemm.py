"""GEMM micro-benchmark: the tcgen05 kernel vs torch.matmul (cuBLAS) on the GPT-2 small shapes, all three layouts.
CUDA events, 3 warm-up + 10 timed launches; operands rotate through > 126 MB of buffers so L2 does not keep them."""
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200.ops.dense import _init_sigs, gemm, pick_tile
from distriflow_b200.utils.clocks import ClockSampler


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


FORCE_BN = int(os.environ.get("DFB_BENCH_BN", "0")) or None
FORCE_SPLITS = int(os.environ.get("DFB_BENCH_SPLITS", "0")) or None


def main():
    _init_sigs()
    dev = torch.device("cuda")
    tokens = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    shapes = [("qkv", 768, 2304), ("proj", 768, 768), ("fc1", 768, 3072), ("fc2", 3072, 768), ("lm_head", 768, 50304)]
    rows = []
    sampler = ClockSampler(0).start()
    for name, K, N in shapes:
        nbuf = max(2, int(200e6 // (tokens * max(K, N) * 2)) + 1)
        xs = [torch.randn(tokens, K, device=dev).bfloat16() for _ in range(nbuf)]
        dys = [torch.randn(tokens, N, device=dev).bfloat16() for _ in range(min(nbuf, 3))]
        w = torch.randn(K, N, device=dev).bfloat16()
        i = [0]

        def nxt(lst):
            i[0] += 1
            return lst[i[0] % len(lst)]

        flops = 2.0 * tokens * K * N
        for kind in ("fwd", "dgrad", "wgrad"):
            if kind == "fwd":
                ours = lambda: gemm(nxt(xs), w, a_mn=False, b_mn=True, M=tokens, N=N, K=K, block_n=FORCE_BN)
                ref = lambda: torch.matmul(nxt(xs), w)
                tile = pick_tile(tokens, N, K)
            elif kind == "dgrad":
                ours = lambda: gemm(nxt(dys), w, a_mn=False, b_mn=False, M=tokens, N=K, K=N, block_n=FORCE_BN)
                ref = lambda: torch.matmul(nxt(dys), w.t())
                tile = pick_tile(tokens, K, N)
            else:
                ours = lambda: gemm(nxt(xs), nxt(dys), a_mn=True, b_mn=True, M=K, N=N, K=tokens, block_n=FORCE_BN, splits=FORCE_SPLITS)
                ref = lambda: torch.matmul(nxt(xs).t(), nxt(dys))
                tile = pick_tile(K, N, tokens)
            t_o, t_r = timeit(ours), timeit(ref)
            rows.append({"shape": name, "kind": kind, "tokens": tokens, "K": K, "N": N, "tile": tile, "ours_ms": round(t_o, 4),
                         "cublas_ms": round(t_r, 4), "ours_tflops": round(flops / t_o / 1e9, 1),
                         "cublas_tflops": round(flops / t_r / 1e9, 1), "ratio": round(t_r / t_o, 3)})
            print(json.dumps(rows[-1]), flush=True)
        del xs, dys
    print(json.dumps({"clocks": sampler.stop()}), flush=True)        # clocks + throttle reasons while the rows above ran
    return rows


if __name__ == "__main__":
    main()
