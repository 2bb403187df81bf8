"""LayerNorm forward / backward at the GPT-2 shape (32768 x 768 bf16): device time and fraction of the measured copy
bandwidth (MEASURED_PEAKS.json), staged (bulk-copy ring) against the register-only backward.  The operands of one call
are 150-250 MB, larger than the 126 MB L2, and calls rotate over two input sets."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distriflow_b200 import ops
from distriflow_b200.ops.norm import _sigs
from distriflow_b200.utils.clocks import ClockSampler

L = _sigs()
dev = torch.device("cuda")
rows, D = (int(v) for v in (sys.argv[1:3] if len(sys.argv) >= 3 else (32768, 768)))
peak = 6532.9
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
sampler = ClockSampler(0).start()
side = torch.cuda.Stream()
torch.cuda.set_stream(side)
sets = []
for i in range(2):
    torch.manual_seed(i)
    sets.append(dict(x=torch.randn(rows, D, device=dev).bfloat16(), r=torch.randn(rows, D, device=dev).bfloat16(),
                     g=torch.randn(D, device=dev).bfloat16(), b=torch.randn(D, device=dev).bfloat16(),
                     gs=torch.randn(rows, D, device=dev).bfloat16(), gy=torch.randn(rows, D, device=dev).bfloat16()))


def timed(fn, n=20):
    """n calls captured in one CUDA graph (the Python / autograd launch path costs more than these kernels run)."""
    for i in range(3):
        fn(sets[i & 1])
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):        # backward nodes run on the stream of their forward: capture on that one
        for i in range(n):
            fn(sets[i & 1])
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def fwd(s):
    with torch.no_grad():
        return ops.add_layer_norm(s["x"], s["r"], s["g"], s["b"], 1e-5)


def make_bwd(with_res):
    graphs = []
    for s in sets:                      # backward alone: build the autograd graph once per input set, replay its backward
        x, r, g, b = (s[k].clone().requires_grad_(True) for k in ("x", "r", "g", "b"))
        if with_res:
            so, y = ops.add_layer_norm(x, r, g, b, 1e-5)
            graphs.append(([so, y], [s["gs"], s["gy"]], (x, r, g, b)))
        else:
            y = ops.layer_norm(x, g, b, 1e-5)
            graphs.append(([y], [s["gy"]], (x, g, b)))

    def run(s):
        outs, grads, leaves = graphs[0 if s is sets[0] else 1]
        torch.autograd.grad(outs, leaves, grads, retain_graph=True)
    return run


out = {"rows": rows, "D": D, "copy_peak_gbs": peak}
nbytes = rows * D * 2
t = timed(fwd)
out["add_ln_fwd_us"] = round(t, 1)
out["add_ln_fwd_frac_of_copy_peak"] = round(4 * nbytes / t / 1e3 / peak, 3)          # reads x, r; writes s, y
for with_res, name, streams in ((True, "add_ln_bwd", 4), (False, "ln_bwd", 3)):    # reads s, dy (, ds); writes dx
    run = make_bwd(with_res)
    for staged in (1, 0):
        L.dfb_layernorm_bwd_staged(staged)
        t = timed(run)
        key = f"{name}_{'staged' if staged else 'registers'}"
        out[key + "_us"] = round(t, 1)
        out[key + "_frac_of_copy_peak"] = round(streams * nbytes / t / 1e3 / peak, 3)
L.dfb_layernorm_bwd_staged(0)
out["note"] = "20 calls per CUDA-graph replay; backward rows include the partial-reduce kernel and the scratch allocation memsets"
out["clocks"] = sampler.stop()
print(json.dumps(out))
