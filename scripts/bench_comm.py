"""Gradient-path micro-benchmark (BASELINE config 5 "gradient bucket sweep vs NCCL"): per bucket size, the fused
reduce-scatter + 1/N + optimizer + all-gather kernel (one launch, P2P over NVLink) against NCCL all-reduce + a PyTorch
optimizer update on the same bf16 buffers.  Device-timed with CUDA events, max over ranks, buffers rotated so that the
working set exceeds L2.  Reports grad-sync bus GB/s = 2(N-1)/N * bytes / t and the fraction of the NVLink roofline
(bytes that must cross the link per GPU / 770 GB/s measured peer bandwidth, B200_PROFILING.md).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_comm.py [--opt adam]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from distriflow_b200.ops import _lib
from distriflow_b200.ops._lib import check, ptr
from distriflow_b200.parallel.engine import _OPT, _comm_sigs
from distriflow_b200.parallel.symm import SymmetricHeap
from distriflow_b200.utils.clocks import ClockSampler

LINK_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--opt", default="sgd")
    ap.add_argument("--ctas", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="")
    ap.add_argument("--nvls", action="store_true", help="also time the in-switch variant (multimem.ld_reduce / multimem.st)")
    ap.add_argument("--sizes", default="", help="comma-separated bucket sizes in MB (default: the full sweep)")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    L = _comm_sigs()
    sizes_mb = [float(x) for x in a.sizes.split(",")] if a.sizes else [0.03125, 0.25, 1, 4, 16, 32, 64, 256]
    nbuf = 3
    max_elems = int(sizes_mb[-1] * (1 << 20)) // 2
    heap = SymmetricHeap(nbuf * max_elems * 4 + (1 << 20), rank, world, None, dev, multicast=a.nvls)
    offs = [(heap.alloc(max_elems * 2), heap.alloc(max_elems * 2)) for _ in range(nbuf)]
    for go, po in offs:
        heap.view(go, (max_elems,), torch.bfloat16).normal_()
        heap.view(po, (max_elems,), torch.bfloat16).normal_()
    epoch = [0]
    rows = []
    sampler = ClockSampler(local).start() if rank == 0 else None
    for mb in sizes_mb:
        n = int(mb * (1 << 20)) // 2
        n = (n + world * 64 - 1) // (world * 64) * (world * 64)
        shard = n // world
        master = torch.zeros(shard, device=dev)
        m = torch.zeros(shard, device=dev)
        v = torch.zeros(shard, device=dev)

        def ours(i):
            go, po = offs[i % nbuf]
            epoch[0] += 1
            check(L.dfb_sync_rs_opt_ag(heap.ptrs(go), heap.ptrs(po), heap.ptrs(0), ptr(master), ptr(m), ptr(v), None, n, shard, rank,
                                       world, epoch[0], 1, 1, 1, _OPT[a.opt], 1e-3, 0.9, 0.9, 0.999, 1e-8, 0.0, 1, a.ctas, 0, 0,
                                       _lib.stream()), "sync")

        def nvls(i):
            go, po = offs[i % nbuf]
            epoch[0] += 1
            check(L.dfb_sync_nvls(heap.mc_ptr(go), heap.mc_ptr(po), heap.ptrs(po), heap.ptrs(0), ptr(master), ptr(m), ptr(v), n, shard,
                                  rank, world, epoch[0], _OPT[a.opt], 1e-3, 0.9, 0.9, 0.999, 1e-8, 0.0, 1, a.ctas, 0, 0, None, None, None,
                                  _lib.stream()), "sync_nvls")

        gl = [torch.randn(n, device=dev).bfloat16() for _ in range(nbuf)]
        wl = [torch.randn(n, device=dev).bfloat16() for _ in range(nbuf)]
        full_master = torch.zeros(n, device=dev)

        def nccl(i):
            g = gl[i % nbuf]
            dist.all_reduce(g)
            full_master.add_(g.float(), alpha=-1e-3 / world)          # SGD on the full fp32 master (DDP-style replica)
            wl[i % nbuf].copy_(full_master)

        def timeit(fn):
            for i in range(3):
                fn(i)
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(a.iters):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)

        t_o, t_n = timeit(ours), timeit(nccl)
        nbytes = n * 2
        bus = lambda t: 2.0 * (world - 1) / world * nbytes / (t * 1e-3) / 1e9      # noqa: E731
        # per direction per GPU: inbound = the slices this rank pulls (N-1)/N*S + the weights the peers push into it (N-1)/N*S
        link_bytes = 2.0 * (world - 1) / world * nbytes
        roof_ms = link_bytes / (LINK_GBS * 1e9) * 1e3
        row = {"bucket_mb": mb, "world": world, "opt": a.opt, "ours_us": round(t_o * 1e3, 1), "nccl_plus_opt_us": round(t_n * 1e3, 1),
               "ours_bus_gbs": round(bus(t_o), 1), "nccl_bus_gbs": round(bus(t_n), 1), "speedup": round(t_n / t_o, 2),
               "nvlink_roofline_frac": round(roof_ms / t_o, 3)}
        if a.nvls:
            t_v = timeit(nvls)
            # in-switch: per GPU and direction the link carries S/N of reduced gradients in + (N-1)/N*S of pushed weights in
            row.update({"nvls_us": round(t_v * 1e3, 1), "nvls_bus_gbs": round(bus(t_v), 1), "nvls_vs_p2p": round(t_o / t_v, 2),
                        "nvls_vs_nccl": round(t_n / t_v, 2)})
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0:
        rows.append({"clocks": sampler.stop()})
        print(json.dumps(rows[-1]), flush=True)
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
