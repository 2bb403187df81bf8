"""Quick per-kernel device-time table of one eager training step (torch.profiler / CUPTI; for guidance between ncu
captures — the ncu launch list of `scripts/step_profile.py` is the evidence of record).

    python scripts/torch_profile_step.py --model resnet50 --optimizer momentum
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="gpt2-small")
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--top", type=int, default=25)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    model, loss_fn, make, bs, _ = bench.build_workload(a, dev, torch)
    from distriflow_b200.parallel.engine import DataParallelEngine
    eng = DataParallelEngine(model, loss_fn, {"optimizer": a.optimizer, "learningRate": 3e-4}, rank=0, world=1, device=dev)
    x, y = (t.to(dev) for t in make())
    for _ in range(3):
        eng.step(x, y)
    torch.cuda.synchronize()
    steps = 2
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as p:
        for _ in range(steps):
            eng.step(x, y)
        torch.cuda.synchronize()
    ev = [e for e in p.key_averages() if e.device_time_total > 0]
    total = sum(e.device_time_total for e in ev)
    print(f"kernel time {total / steps / 1e3:.2f} ms per step, {sum(e.count for e in ev) / steps:.0f} launches per step")
    for e in sorted(ev, key=lambda e: -e.device_time_total)[:a.top]:
        print(f"  {e.key[:100]:100s} {e.count / steps:6.1f}  {e.device_time_total / steps / 1e3:8.3f} ms  {100 * e.device_time_total / total:5.1f} %")


if __name__ == "__main__":
    main()
