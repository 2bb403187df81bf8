"""Command-line trainer: `python -m distriflow_b200.train --model resnet20 --epochs 2 --save-dir saved-models`.

One process per GPU (`python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 -m distriflow_b200.train ...`);
a single process trains on one GPU or on the CPU (`--device cpu`, the plumbing configuration).  The data is the MNIST IDX
directory of the reference's experiment (`/root/reference/experiment/mnist/mnist_data.ts:7-15`) when `--mnist-dir` is given,
synthetic data of the model's input shape otherwise.  The reference has no CLI (configuration is constructor arguments,
SURVEY §5 "Config / flags"); this is the thin one its inventory asks for.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from typing import Any, List, Optional, Tuple

import torch


def parse(argv: Optional[List[str]] = None) -> argparse.Namespace:
    ap = argparse.ArgumentParser(prog="python -m distriflow_b200.train", description=__doc__.split("\n\n")[0])
    ap.add_argument("--model", default="mnist-mlp", choices=["mnist-mlp", "mnist-cnn", "resnet20", "resnet50", "gpt2-tiny",
                                                             "gpt2-small"])
    ap.add_argument("--device", default="auto", choices=["auto", "cuda", "cpu"])
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--batch", type=int, default=32, help="per-rank micro-batch")
    ap.add_argument("--examples", type=int, default=512, help="synthetic examples (ignored with --mnist-dir)")
    ap.add_argument("--seq", type=int, default=128, help="sequence length of the GPT-2 models")
    ap.add_argument("--mnist-dir", default=None, help="directory with the four MNIST IDX files")
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "momentum", "adam"])
    ap.add_argument("--lr", type=float, default=0.001)
    ap.add_argument("--warmup-steps", type=int, default=0, help="linear learning-rate warm-up")
    ap.add_argument("--save-dir", default=None)
    ap.add_argument("--checkpoint-every", type=int, default=0)
    ap.add_argument("--log", default=None, help="JSONL log path")
    ap.add_argument("--graph", action="store_true", help="replay the step from a CUDA graph")
    ap.add_argument("--nvls", action="store_true", help="reduce / broadcast inside the NVSwitch")
    ap.add_argument("--trace", action="store_true", help="print per-phase device times at the end")
    ap.add_argument("--watchdog", type=float, default=0.0, help="seconds without a finished step before a stall report")
    ap.add_argument("--seed", type=int, default=0)
    return ap.parse_args(argv)


def build(args: argparse.Namespace, device: torch.device) -> Tuple[torch.nn.Module, Any, torch.Tensor, torch.Tensor]:
    """(model, loss_fn, x, y) for the chosen workload; data on the host (the Trainer prefetches it)."""
    g = torch.Generator().manual_seed(args.seed + 1)
    torch.manual_seed(args.seed)
    n = args.examples
    if args.model.startswith("gpt2"):
        from .models.gpt2 import GPT2, GPT2Config, gpt2_loss
        cfg = GPT2Config.small() if args.model == "gpt2-small" else GPT2Config.tiny()
        seq = min(args.seq, cfg.n_positions)
        idx = torch.randint(0, cfg.vocab_size, (n, seq + 1), generator=g)
        return GPT2(cfg), gpt2_loss, idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    if args.model.startswith("resnet"):
        from .models.resnet import resnet20, resnet50, resnet_loss
        model, hw, classes = (resnet50(), 224, 1000) if args.model == "resnet50" else (resnet20(), 32, 10)
        return model, resnet_loss, torch.randn(n, hw, hw, 3, generator=g), torch.randint(0, classes, (n,), generator=g)
    from .losses import get_loss
    from .models import mnist_cnn, mnist_mlp
    model = mnist_cnn(10, softmax_output=False) if args.model == "mnist-cnn" else mnist_mlp(softmax_output=False)
    ce = get_loss("softmaxCrossEntropy")
    if args.mnist_dir:
        from .experiment.mnist.mnist_data import load_mnist
        x, y = load_mnist(args.mnist_dir)["train"]
        y = y.long()
    else:
        x, y = torch.rand(n, 28, 28, 1, generator=g) * 255, torch.randint(0, 10, (n,), generator=g)    # raw 0..255 (Q4)
    return model, (lambda m, a, b: ce(b, m(a))), x, y


def main(argv: Optional[List[str]] = None) -> int:
    args = parse(argv)
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = args.device == "cuda" or (args.device == "auto" and torch.cuda.is_available())
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl" if use_cuda else "gloo", **({"device_id": device} if use_cuda else {}))
    model, loss_fn, x, y = build(args, device)
    from .parallel.engine import DataParallelEngine
    from .trainer import Trainer
    half = use_cuda and not args.model.startswith("mnist")
    eng = DataParallelEngine(model.to(device), loss_fn,
                             {"optimizer": args.optimizer, "learningRate": args.lr, "cudaGraph": bool(args.graph and use_cuda),
                              "nvls": bool(args.nvls and use_cuda and world > 1), "trace": args.trace,
                              "watchdogSeconds": args.watchdog, "dtype": "bf16" if half else "fp32",
                              "gradDtype": "bf16" if half else "fp32", "backend": "p2p" if use_cuda else "gloo"},
                             rank=rank, world=world, device=device)
    sched = (lambda step: args.lr * min(1.0, (step + 1) / args.warmup_steps)) if args.warmup_steps > 0 else None
    tr = Trainer(eng, x, y, {"batchSize": args.batch, "epochs": args.epochs, "saveDir": args.save_dir,
                             "checkpointEvery": args.checkpoint_every, "logPath": args.log}, lr_schedule=sched)
    from .parallel.engine import PeerLostError
    try:
        out = tr.fit()
    except PeerLostError as e:
        # restart-based recovery: exit with EX_TEMPFAIL so a supervisor (`torch.distributed.run --max-restarts N`, or the same
        # command on the surviving GPUs) relaunches; the new run resumes from `<save-dir>/current`, the last version written
        # BEFORE the loss (Trainer.save refuses to checkpoint after it), at whatever world size it is given
        print(f"[distriflow_b200] {e}; last good version: {tr.latest()}", file=sys.stderr, flush=True)
        os._exit(75)
    if args.trace:
        out["phases_ms"] = {k: round(v["mean_ms_max_over_ranks"], 3) for k, v in eng.trace_summary().items()}
    eng.close()
    if rank == 0:
        out.update(model=args.model, world=world, device=str(device), version=eng.version)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
