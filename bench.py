#!/usr/bin/env python
"""Headline benchmark: synchronous data-parallel training throughput (samples/sec) of GPT-2 small (124M),
sequence 1024, bf16 — BASELINE.json config 5 — through `distriflow_b200.parallel.DataParallelEngine`.

    python bench.py --gpus N --steps K --warmup W            # N = 1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0 (contract in the task statement).  `--impl reference` prints the
"unavailable" line: the reference is TypeScript on tf.js/socket.io and has no Python package
(`pip install ... /root/reference` -> "Neither 'setup.py' nor 'pyproject.toml' found"; no node/npm offline).

Because that reference arm can never produce a number, the product arm measures its own comparison baseline ON THE
SAME BOX IN THE SAME PROCESSES right after its timed steps: PyTorch DDP + NCCL + cuBLAS/cuDNN on the identical model,
batch, sequence length, optimizer, bucket size, steps and warm-up (BASELINE.json "Baselines"), and fills
`vs_baseline = ours / ddp` and a `baseline` record with its own clocks (`--baseline none` skips it, `--ddp-compile`
runs the DDP arm under torch.compile).  `--impl ddp` runs only that arm; `--model` / `--mode` select the other
BASELINE configs.  The line also carries correctness evidence for the multi-GPU runs: `replicas_identical` (a 64-bit
checksum of every rank's working weights after the timed steps) and the first-step losses of both arms from the same
initial weights and batches.
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ddp"])
    ap.add_argument("--model", default="gpt2-small", choices=["gpt2-small", "gpt2-tiny", "resnet50", "resnet20", "mnist-cnn",
                                                               "mnist-mlp"])
    ap.add_argument("--mode", default="sync", choices=["sync", "async"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU micro-batch (0 = model default)")
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--bucket-mb", type=float, default=0.0,
                    help="gradient bucket size; 0 = auto: 32 MB on one GPU, one bucket for the whole model (256 MB cap) on several "
                         "(measured r2: at 2 and 8 GPUs one late bucket beats overlapped 32 MB buckets by 0.35-0.6 ms per GPT-2 "
                         "step: the bucket kernels slow the backward GEMMs they overlap by more than the NVLink time they hide)")
    ap.add_argument("--ddp-bucket-mb", type=float, default=32.0, help="bucket_cap_mb of the DDP comparison arm")
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--comm-ctas", type=int, default=148)
    ap.add_argument("--reserve-sms", type=int, default=0)
    ap.add_argument("--graph", dest="graph", action="store_true", default=True,
                    help="capture the whole step into a CUDA graph after 3 eager steps (default)")
    ap.add_argument("--nvls", nargs="?", const="on", default="auto", choices=["auto", "on", "off"],
                    help="reduce gradients / broadcast weights inside the NVSwitch (multicast heap).  auto: on from 4 GPUs when every "
                         "rank's device supports multicast objects (r2 at 8 GPUs: 34.70 vs 34.91 ms per step)")
    ap.add_argument("--no-graph", dest="graph", action="store_false", help="eager launches every step")
    ap.add_argument("--fuse-allgather", dest="fuse_ag", action="store_true", default=False,
                    help="fuse the all-gather of the first Dense kernel into its forward GEMM (off by default: the r2 A/B at 8 GPUs "
                         "measured 34.69 ms without it against 34.91 ms with it)")
    ap.add_argument("--no-fuse-allgather", dest="fuse_ag", action="store_false")
    ap.add_argument("--baseline", default="auto", choices=["auto", "none", "ddp"],
                    help="same-box comparison arm run after the product arm (auto: DDP for --impl ours)")
    ap.add_argument("--ddp-compile", action="store_true", help="DDP arm under torch.compile (slow to start; stronger baseline)")
    ap.add_argument("--close-barrier", default="step", choices=["step", "bucket"],
                    help="one closing cross-rank barrier per step (default) or per bucket kernel (round-1 protocol)")
    ap.add_argument("--trace", action="store_true", help="eager steps with device-timed forward/backward/collective_tail phases")
    return ap.parse_args()


def build_workload(args, device, torch):
    """Returns (model, loss_fn, make_host_batch(), samples_per_batch, config_dict)."""
    import distriflow_b200  # noqa: F401
    if args.model.startswith("gpt2"):
        from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
        cfg = GPT2Config.small() if args.model == "gpt2-small" else GPT2Config.tiny()
        seq = min(args.seq, cfg.n_positions)
        bs = args.batch or (32 if args.model == "gpt2-small" else 8)
        torch.manual_seed(1234)
        model = GPT2(cfg).to(device)

        def make():
            idx = torch.randint(0, cfg.vocab_size, (bs, seq + 1))
            return idx[:, :-1].contiguous().pin_memory(), idx[:, 1:].contiguous().pin_memory()

        return model, gpt2_loss, make, bs, {"model": "gpt2-small-124M" if args.model == "gpt2-small" else "gpt2-tiny",
                                            "seq_len": seq, "per_gpu_batch": bs}
    if args.model in ("resnet50", "resnet20"):
        from distriflow_b200.models.resnet import resnet20, resnet50, resnet_loss
        torch.manual_seed(1234)
        if args.model == "resnet50":
            model, hw, classes, bs = resnet50().to(device), 224, 1000, args.batch or 64
        else:
            model, hw, classes, bs = resnet20().to(device), 32, 10, args.batch or 256

        def make():
            return torch.randn(bs, hw, hw, 3).pin_memory(), torch.randint(0, classes, (bs,)).pin_memory()

        return model, resnet_loss, make, bs, {"model": args.model, "image": hw, "per_gpu_batch": bs}
    from distriflow_b200.losses import get_loss
    from distriflow_b200.models import mnist_cnn, mnist_mlp
    torch.manual_seed(1234)
    model = (mnist_cnn(10, softmax_output=False) if args.model == "mnist-cnn" else mnist_mlp(softmax_output=False)).to(device)
    bs = args.batch or 32
    ce = get_loss("softmaxCrossEntropy")

    def loss_fn(m, x, y):
        return ce(y, m(x))

    def make():
        return (torch.rand(bs, 28, 28, 1) * 255).pin_memory(), torch.randint(0, 10, (bs,)).pin_memory()

    return model, loss_fn, make, bs, {"model": args.model, "per_gpu_batch": bs}


def weights_checksum(engine, torch):
    """64-bit checksum of this rank's working weights (the bits, not the values).  The range whose all-gather is fused into
    its consumer GEMM is refreshed by the next forward pass, so it is left out."""
    total = torch.zeros((), dtype=torch.int64, device=engine.device)
    for b in getattr(engine, "buckets", []):
        bits = b.param_buf.view(torch.int16 if b.param_buf.element_size() == 2 else torch.int32)
        lo, hi = getattr(b, "nopush", (0, 0))
        for part in (bits[:lo], bits[hi:]) if hi > lo else (bits,):
            idx = torch.arange(1, part.numel() + 1, device=part.device, dtype=torch.int64)
            total = total + (part.to(torch.int64) * (idx % 8191 + 1)).sum()
    return total


def run_arm(args, impl, rank, local_rank, world, device, host_seed):
    """Build the model + step function of one arm, warm it up, time `args.steps` steps with CUDA events, then measure the
    same steps end to end from pinned host memory.  Returns a dict of raw measurements (this rank's)."""
    import torch
    import torch.distributed as dist
    from distriflow_b200.ops import _lib
    from distriflow_b200.utils.clocks import ClockSampler

    model, loss_fn, make_batch, bs, wcfg = build_workload(args, device, torch)
    engine = None
    bucket_mb = args.bucket_mb if args.bucket_mb > 0 else (32.0 if world == 1 else 256.0)
    use_nvls = False
    if impl == "ours" and world > 1 and args.mode == "sync" and args.nvls != "off" and args.model.startswith(("gpt2", "resnet")):
        # (the in-switch path reduces bf16 gradients; the fp32 MNIST models stay on the P2P kernel)
        from distriflow_b200.parallel.symm import multicast_supported
        ok = torch.tensor([1 if multicast_supported(device) else 0], device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)                 # every rank takes the same decision
        use_nvls = bool(int(ok)) and (args.nvls == "on" or world >= 4)
        if args.nvls == "on" and not use_nvls:
            raise SystemExit("--nvls on: a device of this job cannot join an NVSwitch multicast object")
    if impl == "ours":
        if args.mode == "async":
            from distriflow_b200.parallel.async_engine import AsyncEngine
            engine = AsyncEngine(model, loss_fn, {"optimizer": args.optimizer if args.optimizer != "adam" else "sgd",
                                                  "learningRate": 1e-3, "maximumStaleness": 4, "mode": "async",
                                                  "cudaGraph": args.graph},
                                 rank=rank, world=world, device=device)
        else:
            from distriflow_b200.parallel.engine import DataParallelEngine
            engine = DataParallelEngine(model, loss_fn, {"optimizer": args.optimizer, "learningRate": 3e-4,
                                                         "bucketBytes": int(bucket_mb * (1 << 20)),
                                                         "commCtas": args.comm_ctas, "reserveSMs": args.reserve_sms,
                                                         "cudaGraph": args.graph and not args.trace, "trace": args.trace,
                                                         "fuseAllGather": args.fuse_ag, "closeBarrier": args.close_barrier,
                                                         "nvls": use_nvls},
                                        rank=rank, world=world, device=device)
        step = engine.step
        close = lambda: None                                                      # noqa: E731
    else:   # PyTorch DDP + NCCL + cuBLAS baseline, identical model and schedule (bf16 weights and gradients)
        from baseline.ddp_baseline import make_ddp_step
        step, close = make_ddp_step(model, loss_fn, args.optimizer, 3e-4, world, device, int(args.ddp_bucket_mb),
                                    compile_model=args.ddp_compile)

    g = torch.Generator().manual_seed(host_seed + rank)
    torch.set_rng_state(g.get_state())                    # same synthetic batches in both arms
    host = [make_batch() for _ in range(4)]
    dev = [(x.to(device, non_blocking=True), y.to(device, non_blocking=True)) for x, y in host]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    graph = bool(args.graph and impl == "ours" and not args.trace)
    prep = 4 if graph else 0
    first_losses = []
    for i in range(max(prep, 3)):              # engine set-up (3 eager steps + the graph capture); never timed, not counted in W
        loss_i = step(*dev[i % len(dev)])
        if i < 3:
            first_losses.append(round(float(loss_i), 5))
    prep = max(prep, 3)
    for i in range(args.warmup):
        step(*dev[i % len(dev)])
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = _lib.LAUNCHES[0]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prof = os.environ.get("DFB_PROFILE_RANGE") == "1" and impl == args.impl     # ncu --profile-from-start off: timed steps only
    if prof:
        torch.cuda.profiler.start()
    e0.record()
    loss = None
    for i in range(args.steps):
        loss = step(*dev[i % len(dev)])
    e1.record()
    if prof:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    barrier()
    launches = _lib.LAUNCHES[0] - l0
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None

    # end-to-end through the public API: pinned host inputs -> H2D -> step -> loss D2H, every step
    e2e = None
    if not args.no_e2e:
        barrier()
        t0 = time.perf_counter()
        last = 0.0
        for i in range(args.steps):
            hx, hy = host[i % len(host)]
            x = hx.to(device, non_blocking=True)
            y = hy.to(device, non_blocking=True)
            last = float(step(x, y).item())
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3
        hx, hy = host[0]
        e2e = (e2e_ms, hx.numel() * hx.element_size() + hy.numel() * hy.element_size(), 4, last)

    identical = None
    if impl == "ours" and args.mode == "sync" and engine is not None and hasattr(engine, "buckets"):
        cs = weights_checksum(engine, torch).reshape(1)
        if world > 1:
            allcs = [torch.zeros_like(cs) for _ in range(world)]
            dist.all_gather(allcs, cs)
            identical = all(int(c) == int(allcs[0]) for c in allcs)
        else:
            identical = True
    async_counters = None
    if impl == "ours" and args.mode == "async" and engine is not None:
        engine.drain()
        c = engine.counters()                      # this rank's shard: every gradient is counted once on every shard
        async_counters = {"applied": int(c["applied"]), "dropped": int(c["dropped"]), "updates": int(c["version"]),
                          "max_staleness_seen": int(c["max_staleness_seen"])}
    trace = None
    if args.trace and engine is not None and getattr(engine, "tracer", None) is not None:
        trace = engine.trace_summary()

    t = torch.tensor([ms, e2e[0] if e2e else 0.0], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = {"ms": float(t[0]), "e2e_ms": float(t[1]), "e2e": e2e, "loss": float(loss), "launches": int(launches),
           "clocks": clocks, "bs": bs, "wcfg": wcfg, "prep": prep, "nvls": use_nvls, "bucket_mb": bucket_mb if impl == "ours" else args.ddp_bucket_mb, "graph": graph, "first_losses": first_losses,
           "replicas_identical": identical, "trace": trace, "async": async_counters,
           "buckets": len(engine.buckets) if engine is not None and hasattr(engine, "buckets") else None}
    close()
    del step, engine, model, dev
    gc.collect()
    torch.cuda.empty_cache()
    return out


def main():
    args = parse()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:          # launched under torchrun for N > 1: one line, from rank 0
            return 0
        print(json.dumps({"impl": "reference", "unavailable": "reference is TypeScript on tf.js/socket.io: no setup.py/pyproject "
                          "(pip: 'not installable'), and node/npm/tfjs are absent offline; see DESIGN.md"}))
        return 0
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:            # convenience: re-launch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), __file__] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); the CPU plumbing configuration is exercised by "
                         "`pytest -m 'not gpu'` and `python -m distriflow_b200.train --device cpu`")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    r = run_arm(args, args.impl, rank, local_rank, world, device, host_seed=4321)
    base = None
    if args.impl == "ours" and args.baseline != "none":
        try:
            base = run_arm(args, "ddp", rank, local_rank, world, device, host_seed=4321)
        except Exception as e:                      # the product number must not die with the comparison arm
            base = {"error": f"{type(e).__name__}: {e}"[:300]}

    if rank == 0:
        gb = r["bs"] * world
        value = gb * args.steps / (r["ms"] / 1e3)
        out = {"metric": "samples/sec sync DP" if args.mode == "sync" else "samples/sec async DP", "value": value,
               "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": r["ms"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "bf16" if args.model.startswith(("gpt2", "resnet")) else "fp32", "data": "synthetic",
               "impl": args.impl,
               "config": {**r["wcfg"], "global_batch": gb, "parallelism": f"dp{world}", "mode": args.mode,
                          "optimizer": args.optimizer, "bucket_mb": r["bucket_mb"], "buckets": r["buckets"],
                          "cuda_graph": r["graph"], "nvls": r["nvls"], "graph_prep_steps": r["prep"],
                          "fuse_allgather": bool(args.fuse_ag), "close_barrier": args.close_barrier,
                          "l2": "inputs+activations per step (GBs) far exceed the 126 MB L2; no explicit flush",
                          "loss": r["loss"]},
               "clocks": r["clocks"], "gpu_launches": r["launches"], "replicas_identical": r["replicas_identical"],
               "loss_first_steps": {"ours" if args.impl == "ours" else args.impl: r["first_losses"]}}
        if r["e2e"]:
            out["e2e"] = {"value": gb * args.steps / (r["e2e_ms"] / 1e3), "unit": "samples/s", "h2d_bytes_per_step": r["e2e"][1],
                          "d2h_bytes_per_step": r["e2e"][2], "ms_per_step": r["e2e_ms"] / args.steps}
        if r["trace"]:
            out["trace"] = r["trace"]
        if r["async"]:
            a = r["async"]
            frac = a["applied"] / max(a["applied"] + a["dropped"], 1)
            out["async"] = {**a, "applied_fraction": frac, "value_applied": value * frac,
                            "note": "value counts every gradient computed; value_applied only those the staleness guard let through"}
        if base is not None and "error" not in base:
            bval = gb * args.steps / (base["ms"] / 1e3)
            out["vs_baseline"] = value / bval
            out["baseline"] = {"impl": "PyTorch DDP + NCCL + cuBLAS/cuDNN, eager" + (" + torch.compile" if args.ddp_compile else ""),
                               "same_box_same_process": True, "value": bval, "unit": "samples/s",
                               "ms_per_step": base["ms"] / args.steps, "clocks": base["clocks"],
                               "config": {**base["wcfg"], "global_batch": base["bs"] * world, "optimizer": args.optimizer,
                                          "bucket_mb": base["bucket_mb"], "steps": args.steps, "warmup": args.warmup + base["prep"]},
                               "loss": base["loss"]}
            if base["e2e"]:
                out["baseline"]["e2e_value"] = gb * args.steps / (base["e2e_ms"] / 1e3)
                if r["e2e"]:
                    out["e2e"]["vs_baseline"] = out["e2e"]["value"] / out["baseline"]["e2e_value"]
            out["loss_first_steps"]["ddp"] = base["first_losses"]
        elif base is not None:
            out["baseline"] = base
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
