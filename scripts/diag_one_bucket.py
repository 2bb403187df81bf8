"""Diagnostic: GPT-2 tiny, 6 Adam steps, NCCL arm vs p2p arms with ONE bucket (fused all-gather on / off, both closing-barrier
protocols), per-parameter distance to the NCCL arm.  Launch with torchrun, one rank per GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
from distriflow_b200.models.gpt2 import GPT2, GPT2Config, gpt2_loss
from distriflow_b200.parallel.engine import DataParallelEngine

cfg = GPT2Config.tiny()
variants = ["nccl", "p2p-b1073741824", "p2p-b1073741824-nofuse", "p2p-b1073741824-bucketclose", "p2p-b65536", "p2p-b65536-nofuse",
            "p2p-b1073741824-sgd", "nccl-sgd"]
res, names = {}, None
for v in variants:
    torch.manual_seed(0)
    model = GPT2(cfg).to(dev)
    flags = v.split("-")
    bucket = next((int(f[1:]) for f in flags if f[:1] == "b" and f[1:].isdigit()), 1 << 18)
    eng = DataParallelEngine(model, gpt2_loss, {"optimizer": "sgd" if "sgd" in flags else "adam", "learningRate": 1e-3, "bucketBytes": bucket,
                                                "backend": flags[0], "fuseAllGather": "nofuse" not in flags,
                                                "closeBarrier": "bucket" if "bucketclose" in flags else "step"},
                             rank=rank, world=world, device=dev)
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    g = torch.Generator().manual_seed(100 + rank)
    idx = torch.randint(0, cfg.vocab_size, (4, 65), generator=g).to(dev)
    losses = [float(eng.step(idx[:, :-1].contiguous(), idx[:, 1:].contiguous())) for _ in range(6)]
    torch.cuda.synchronize()
    res[v] = (losses, [t.detach().float().cpu() for t in eng.full_precision_weights()], len(eng.buckets))
    if "init" not in res:
        torch.manual_seed(0)
        res["init"] = [p.detach().float().cpu() for p in GPT2(cfg).parameters() if p.requires_grad]
if rank == 0:
    for ref, vs in (("nccl", [v for v in variants if v.startswith("p2p") and "sgd" not in v]), ("nccl-sgd", ["p2p-b1073741824-sgd"])):
        wn = res[ref][1]
        moved = sum(float((a - b).norm()) ** 2 for a, b in zip(wn, res["init"])) ** 0.5
        for v in vs:
            wv = res[v][1]
            d = sum(float((a - b).norm()) ** 2 for a, b in zip(wv, wn)) ** 0.5
            worst = sorted(((float((a - b).norm()) / max(float((b - c).norm()), 1e-12), n) for a, b, c, n in zip(wv, wn, res["init"], names)), reverse=True)[:4]
            print(f"{v:32s} buckets {res[v][2]:3d} dist/moved {d / moved:.4f} losses {[round(x, 4) for x in res[v][0][-2:]]} vs {[round(x, 4) for x in res[ref][0][-2:]]}"
                  f" worst {[(round(r, 3), n) for r, n in worst]}", flush=True)
dist.destroy_process_group()
