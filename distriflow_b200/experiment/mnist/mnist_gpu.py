"""The MNIST experiment on a B200 box: every worker is a GPU rank, the server is sharded across the ranks.

BASELINE.json config 2 — "MNIST CNN async SGD, maximumStaleness=4, 8 workers on 8xB200":

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m distriflow_b200.experiment.mnist.mnist_gpu \\
        --model cnn --mode async --max-staleness 4

The dataset (real IDX files if present, else MNIST-shaped synthetic blobs, raw 0..255 as in
`/root/reference/experiment/mnist/mnist_data.ts:41-44`) is resident on every GPU; micro-batches are dispatched by the
device work queue (FCFS tickets, re-issue of incomplete batches: `/root/reference/src/server/dataset.ts:47-67`);
gradients go to the owner shards with P2P reductions behind the staleness guard.  `--parity` runs one worker with the
reference's loss quirk and prints the loss curve next to the CPU oracle's (BASELINE "MNIST loss parity").
"""
from __future__ import annotations

import argparse
import json
import os
import time

import torch


def main(argv=None) -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="cnn", choices=["cnn", "mlp"])
    ap.add_argument("--mode", default="async", choices=["async", "sync"])
    ap.add_argument("--max-staleness", type=int, default=4)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--examples", type=int, default=8192)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--lr", type=float, default=0.001)
    ap.add_argument("--loss", default="softmaxCrossEntropy")
    ap.add_argument("--parity", action="store_true")
    ap.add_argument("--out", default="")
    a = ap.parse_args(argv)

    import torch.distributed as dist
    from ...losses import get_loss
    from ...models import mnist_cnn, mnist_mlp
    from ...oracle import MlpOracle, simulate_async, synthetic_mnist
    from ...parallel.async_engine import AsyncEngine
    from ...parallel.engine import DataParallelEngine
    from ...parallel.workqueue import DeviceWorkQueue

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    x_np, y_np = synthetic_mnist(a.examples, seed=0)
    x_all, y_all = torch.from_numpy(x_np).to(dev), torch.from_numpy(y_np).to(dev)
    loss_name = "referenceCompat" if a.parity else a.loss
    softmax_out = loss_name != "softmaxCrossEntropy"
    torch.manual_seed(0)
    if a.parity:
        oracle = MlpOracle(seed=1, loss="referenceCompat", lr=a.lr)
        net = mnist_mlp(softmax_output=True)
        with torch.no_grad():
            for p, v in zip(net.parameters(), oracle.vars):
                p.copy_(torch.from_numpy(v))
    else:
        net = mnist_cnn(10, softmax_output=softmax_out, dropout=False) if a.model == "cnn" else mnist_mlp(softmax_output=softmax_out)
    net.to(dev)
    lf = get_loss(loss_name)
    loss_fn = lambda m, x, y: lf(y, m(x))        # noqa: E731
    cfg = {"optimizer": "sgd", "learningRate": a.lr, "maximumStaleness": a.max_staleness, "dtype": "fp32", "gradDtype": "fp32"}
    batches = a.examples // a.batch
    losses = []
    torch.cuda.synchronize()
    if a.mode == "async":
        eng = AsyncEngine(net, loss_fn, cfg, rank=rank, world=world, device=dev)
        wq = DeviceWorkQueue(eng.state.heaps[0], batches, a.epochs, home_rank=0)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        steps = eng.run_dataset(wq, x_all, y_all, a.batch, on_step=(lambda n, l: losses.append(l)) if (a.parity or rank == 0) else None)
        e1.record()
        eng.drain()
        ms = e0.elapsed_time(e1)
        ctl = eng.counters()
    else:
        eng = DataParallelEngine(net, loss_fn, {**cfg, "mode": "sync"}, rank=rank, world=world, device=dev)
        per_rank = batches // world
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        steps = 0
        for ep in range(a.epochs):
            for i in range(per_rank):
                b = i * world + rank
                losses.append(eng.step(x_all[b * a.batch:(b + 1) * a.batch], y_all[b * a.batch:(b + 1) * a.batch]))
                steps += 1
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        ctl = {"version": eng.version}
    t = torch.tensor([ms, float(steps)], device=dev, dtype=torch.float64)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, total_steps = float(tmax[0]), float(tsum[1])
    else:
        total_steps = float(steps)
    result = {"model": a.model, "mode": a.mode, "world": world, "maximumStaleness": a.max_staleness, "steps_total": total_steps,
              "ms": ms, "samples_per_sec": total_steps * a.batch / (ms / 1e3), "counters": ctl,
              "loss_first": float(losses[0]) if losses else None, "loss_last": float(losses[-1]) if losses else None}
    if a.parity and rank == 0:
        ref = simulate_async(oracle, x_np, y_np, batch_size=a.batch, epochs=a.epochs, workers=1)
        got = [float(l) for l in losses[:len(ref["loss"])]]
        diffs = [abs(g - r) for g, r in zip(got, ref["loss"])]
        result["parity"] = {"steps_compared": len(diffs), "max_abs_loss_diff": max(diffs), "oracle_first": ref["loss"][0],
                            "oracle_last": ref["loss"][-1], "gpu_first": got[0], "gpu_last": got[-1],
                            "curve": [{"step": i, "gpu": got[i], "oracle": ref["loss"][i]} for i in range(0, len(diffs), max(1, len(diffs) // 16))]}
    if rank == 0:
        print(json.dumps(result), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(result, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
