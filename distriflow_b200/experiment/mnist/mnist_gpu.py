"""The MNIST experiment on a B200 box, written against `DistriServer` / `DistriWorker` only (backend "b200"): every
worker is a GPU rank, the server is sharded across the ranks.

BASELINE.json config 2 — "MNIST CNN async SGD, maximumStaleness=4, 8 workers on 8xB200":

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m distriflow_b200.experiment.mnist.mnist_gpu \\
        --model cnn --mode async --max-staleness 4

The GPU twin of `/root/reference/experiment/mnist/mnist_server.ts:24-35` + `mnist_client.ts:23-29`: the server owns the
model and the dataset (real IDX files if present, else MNIST-shaped synthetic blobs, raw 0..255 as in
`/root/reference/experiment/mnist/mnist_data.ts:41-44`), workers send batch metrics with every upload (`sendMetrics`),
the upload hook logs them.  `--parity` runs one worker with the reference's loss quirk and compares the loss curve AND the
final weights with the CPU oracle (BASELINE "MNIST loss parity"); `--normalise` feeds x/255 so that the softmax is not
saturated and the loss really falls.
"""
from __future__ import annotations

import argparse
import json
import os

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
    ap.add_argument("--updates-per-version", type=int, default=0, help="sync: minUpdatesPerVersion (0 = one per worker)")
    ap.add_argument("--parity", action="store_true")
    ap.add_argument("--normalise", action="store_true", help="inputs / 255 (the reference feeds raw 0..255)")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--out", default="")
    a = ap.parse_args(argv)

    import torch.distributed as dist
    from ... import DistriDataset, DistriServer, DistriWorker
    from ...models import mnist_cnn, mnist_mlp
    from ...oracle import CnnOracle, MlpOracle, simulate_async, synthetic_mnist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    x_np, y_np = synthetic_mnist(a.examples, seed=0)
    if a.normalise:
        x_np = x_np / 255.0
    loss_name = "referenceCompat" if a.parity else a.loss
    softmax_out = loss_name != "softmaxCrossEntropy"
    torch.manual_seed(0)
    oracle = None
    if a.parity:
        if a.model == "cnn":
            oracle = CnnOracle(seed=1, loss="referenceCompat", lr=a.lr, classes=10)
            net = mnist_cnn(10, softmax_output=True, dropout=False)
        else:
            oracle = MlpOracle(seed=1, loss="referenceCompat", lr=a.lr)
            net = mnist_mlp(softmax_output=True)
        with torch.no_grad():
            for p, v in zip(net.parameters(), oracle.vars):
                p.copy_(torch.from_numpy(v))
    else:
        net = mnist_cnn(10, softmax_output=softmax_out, dropout=False) if a.model == "cnn" else mnist_mlp(softmax_output=softmax_out)

    dataset = DistriDataset(torch.from_numpy(x_np), torch.from_numpy(y_np), {"batchSize": a.batch, "epochs": a.epochs})
    server = DistriServer(None, net, dataset, {
        "modelCompileConfig": {"loss": loss_name, "learningRate": a.lr},
        "serverHyperparams": {"maximumStaleness": a.max_staleness if a.mode == "async" else 0,
                              "minUpdatesPerVersion": a.updates_per_version or world},
        "sendMetrics": True, "verbose": a.verbose,
        "engine": {"dtype": "fp32", "gradDtype": "fp32"},
    }, mode=a.mode, backend="b200", device=dev)
    worker = DistriWorker(server, backend="b200")
    uploads = []
    # mnist_server.ts:30-32 logs metrics[0] twice; here loss and accuracy are what they say
    server.on_upload(lambda m: (uploads.append(m), server.log(f"{m.clientId} batch {m.batch}: loss {m.metrics[0]:.4f} "
                                                                f"accuracy {m.metrics[1]:.3f}" if m.metrics else "upload")))
    worker.setup()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = worker.run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms, float(out["uploads"])], device=dev, dtype=torch.float64)
    if world > 1:
        tmax, tsum = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, total = float(tmax[0]), float(tsum[1])
    else:
        total = float(out["uploads"])
    counters = server.engine.counters() if a.mode == "async" else {"version": int(server.engine.version)}
    applied = float(counters.get("applied", total)) if a.mode == "async" else total
    losses = [m.metrics[0] for m in uploads if m.metrics]
    result = {"model": a.model, "mode": a.mode, "world": world, "workerIDs": sorted(server.workerIDs),
              "maximumStaleness": a.max_staleness, "uploads_total": total, "ms": ms,
              "samples_per_sec": total * a.batch / (ms / 1e3),
              # async: a gradient refused by the staleness guard did no useful work; its batch is re-issued
              "samples_per_sec_applied": (min(applied, total) if a.mode == "async" else total) * a.batch / (ms / 1e3),
              "counters": counters, "loss_first": losses[0] if losses else None, "loss_last": losses[-1] if losses else None,
              "accuracy_last": uploads[-1].metrics[1] if uploads and uploads[-1].metrics else None}
    if a.parity and rank == 0:
        ref = simulate_async(oracle, x_np, y_np, batch_size=a.batch, epochs=a.epochs, workers=1)
        got = losses[:len(ref["loss"])]
        diffs = [abs(g - r) for g, r in zip(got, ref["loss"])]
        wdiff = max(float((w.cpu() - torch.from_numpy(v)).abs().max()) for w, v in zip(server.model.get_vars(), oracle.vars))
        wmove = max(float(abs(v - v0).max()) for v, v0 in zip(oracle.vars, ref["initial_vars"]))
        result["parity"] = {"steps_compared": len(diffs), "max_abs_loss_diff": max(diffs), "max_abs_weight_diff": wdiff,
                            "max_weight_movement": wmove, "normalised_inputs": bool(a.normalise),
                            "oracle_first": ref["loss"][0], "oracle_last": ref["loss"][-1], "gpu_first": got[0], "gpu_last": got[-1],
                            "curve": [{"step": i, "gpu": got[i], "oracle": ref["loss"][i]}
                                      for i in range(0, len(diffs), max(1, len(diffs) // 16))]}
    if rank == 0:
        print(json.dumps(result), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(result, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
