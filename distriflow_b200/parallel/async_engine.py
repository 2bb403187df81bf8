"""Asynchronous SGD with bounded staleness over peer memory — no host round trip on the data path.

Reference behaviour replaced (`/root/reference/src/server/asynchronousSGD_server.ts:65-108`,
`/root/reference/src/client/asynchronousSGD_client.ts:43-84`): a worker uploads a gradient, the server
applies it at once and hands out fresh weights.  Here every rank is a worker AND owns 1/R of the server:

  pull   worker reads every owner's slice of the working weights with P2P loads and records the owners'
         version counters (`seen_version`);
  push   worker adds its gradient slice by slice into the owners' open accumulation buffers with P2P
         `red.add`, but only where `owner_version - seen_version <= maximumStaleness` (device-side guard —
         the `maximumStaleness` of `/root/reference/README.md:27`, which the reference never implemented);
  apply  owner flips its double-buffered accumulator, waits for in-flight pushers of the closed half, applies the
         accumulated gradients in place (SGD / momentum / Adam on the fp32 master shard), refreshes its slice of
         the working weights and bumps its version by the number of gradients applied.

`AsyncShardState` wraps the three kernels (and can emulate R ranks on one GPU for tests); `AsyncEngine` is the
training loop: device work-queue ticket -> gather batch -> pull -> fit -> push -> apply -> complete ticket.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional

import torch
import torch.nn as nn

from ..config import engine_config
from ..ops import _lib
from ..ops._lib import check, dt, ptr
from .engine import _OPT, _TORCH_DT, _comm_sigs, _round_up
from .symm import SymmetricHeap

__all__ = ["AsyncShardState", "AsyncEngine"]


class _Local:
    pass


class AsyncShardState:
    def __init__(self, heaps: List[SymmetricHeap], ranks: List[int], world: int, n: int, grad_dtype, param_dtype, lr: float,
                 max_staleness: int, optimizer: str = "sgd", policy: str = "drop", cfg: Optional[Dict[str, Any]] = None,
                 ctas: int = 32, version_per_apply: bool = False):
        self.L = _comm_sigs()
        self.heaps, self.ranks, self.world = heaps, ranks, world
        self.n = _round_up(n, world * 64)
        self.shard = self.n // world
        self.grad_dtype, self.param_dtype = grad_dtype, param_dtype
        self.lr, self.max_staleness, self.optimizer, self.policy = lr, max_staleness, optimizer, policy
        self.cfg = cfg or {"momentum": 0.9, "beta1": 0.9, "beta2": 0.999, "eps": 1e-8, "weightDecay": 0.0}
        self.ctas = ctas
        self.steps = 0
        # True: a shard's version counts optimizer UPDATES (all gradients that met in one accumulation buffer = one update);
        # False: it counts gradients (the literal "every upload is a version" of asynchronousSGD_server.ts:95-108)
        self.version_per_apply = bool(version_per_apply)
        self.step_dev: Optional[torch.Tensor] = None          # device step counter (Adam under CUDA-graph replay)
        psz = torch.empty((), dtype=param_dtype).element_size()
        self.psz = psz
        self.loc: Dict[int, _Local] = {}
        for h, r in zip(heaps, ranks):
            lo = _Local()
            lo.heap = h
            lo.p_off = h.alloc(self.n * psz)
            lo.a0_off = h.alloc(self.shard * 4)
            lo.a1_off = h.alloc(self.shard * 4)
            lo.params = h.view(lo.p_off, (self.n,), param_dtype)
            lo.acc0 = h.view(lo.a0_off, (self.shard,), torch.float32)
            lo.acc1 = h.view(lo.a1_off, (self.shard,), torch.float32)
            d = h.device
            lo.master = torch.zeros(self.shard, device=d)
            lo.m = torch.zeros(self.shard, device=d) if optimizer in ("momentum", "adam") else None
            lo.v = torch.zeros(self.shard, device=d) if optimizer == "adam" else None
            lo.seen = torch.zeros(8, dtype=torch.int32, device=d)
            lo.push_state = torch.zeros(16, dtype=torch.int32, device=d)
            lo.scratch = torch.zeros(2, dtype=torch.int32, device=d)
            self.loc[r] = lo

    @classmethod
    def emulated(cls, world: int, n: int, **kw) -> "AsyncShardState":
        npad = _round_up(n, world * 64)
        psz = torch.empty((), dtype=kw.get("param_dtype", torch.float32)).element_size()
        heaps = SymmetricHeap.emulated(world, npad * psz + 8 * (npad // world) + 65536)
        return cls(heaps, list(range(world)), world, n, **kw)

    # ------------------------------------------------------------------------------------------
    def init_weights(self, w: torch.Tensor) -> None:
        """Same full-precision weights on every rank; each rank keeps its slice as fp32 master."""
        for r, lo in self.loc.items():
            lo.params.zero_()
            lo.params[:w.numel()].copy_(w.reshape(-1))
            full = torch.zeros(self.n, device=w.device)
            full[:w.numel()] = w.reshape(-1).float()
            lo.master.copy_(full[r * self.shard:(r + 1) * self.shard])

    def _one(self, rank: Optional[int]) -> int:
        return self.ranks[0] if rank is None else rank

    def pull(self, rank: Optional[int] = None) -> None:
        r = self._one(rank)
        lo = self.loc[r]
        h = lo.heap
        check(self.L.dfb_async_pull(h.ptrs(lo.p_off), h.ptrs(0), ptr(lo.params), ptr(lo.seen), self.n * self.psz,
                                    self.shard * self.psz, r, self.world, self.ctas, _lib.stream()), "dfb_async_pull")

    def push(self, rank: Optional[int], grad: torch.Tensor, scale: float = 1.0, ticket: Optional[torch.Tensor] = None) -> None:
        r = self._one(rank)
        lo = self.loc[r]
        h = lo.heap
        assert grad.numel() == self.n and grad.is_contiguous(), (grad.numel(), self.n)
        check(self.L.dfb_async_push(ptr(grad), h.ptrs(lo.a0_off), h.ptrs(lo.a1_off), h.ptrs(0), ptr(lo.seen), ptr(lo.push_state),
                                    self.n, self.shard, r, self.world, self.max_staleness, scale,
                                    1 if self.policy == "scale" else 0, dt(grad), self.ctas, ptr(ticket), _lib.stream()),
              "dfb_async_push")

    def apply(self, rank: Optional[int] = None) -> None:
        r = self._one(rank)
        lo = self.loc[r]
        c = self.cfg
        self.steps += 1
        shard_view = lo.params[r * self.shard:(r + 1) * self.shard]
        check(self.L.dfb_async_apply(ptr(lo.acc0), ptr(lo.acc1), lo.heap.pad_ptr(r), ptr(lo.master), ptr(lo.m), ptr(lo.v),
                                     ptr(shard_view), ptr(lo.scratch), self.shard, dt(lo.params), _OPT[self.optimizer], self.lr,
                                     c["momentum"], c["beta1"], c["beta2"], c["eps"], c["weightDecay"], self.steps, self.ctas,
                                     ptr(self.step_dev), int(self.version_per_apply), _lib.stream()), "dfb_async_apply")

    def read_ctl(self, rank: Optional[int] = None) -> Dict[str, Any]:
        r = self._one(rank)
        pad = self.loc[r].heap.pad
        torch.cuda.synchronize()
        w = [int(x) & 0xffffffff for x in pad[32:41].tolist()]
        return {"version": w[0], "active": w[1], "inflight": [w[2], w[3]], "arrived": [w[4], w[5]], "applied": w[6],
                "dropped": w[7], "max_staleness_seen": w[8]}

    def full_master(self) -> torch.Tensor:
        if len(self.loc) == self.world:
            return torch.cat([self.loc[r].master for r in range(self.world)])
        import torch.distributed as dist
        mine = self.loc[self.ranks[0]].master
        parts = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(parts, mine)
        return torch.cat(parts)

    def local_weights(self, rank: Optional[int] = None) -> torch.Tensor:
        return self.loc[self._one(rank)].params

    def close(self) -> None:
        for h in self.heaps:
            h.close()


class AsyncEngine:
    """One rank's asynchronous training loop.  `step(*batch)` = pull -> fit -> push -> apply; with a
    `DeviceWorkQueue` + device-resident dataset, `run_dataset()` drives whole epochs without host round trips."""

    def __init__(self, model: nn.Module, loss_fn: Callable[..., torch.Tensor], config: Optional[Dict[str, Any]] = None,
                 rank: int = 0, world: int = 1, group=None, device: Optional[torch.device] = None):
        self.cfg = engine_config({**(config or {}), "mode": "async"})
        self.model, self.loss_fn, self.rank, self.world, self.group = model, loss_fn, rank, world, group
        self.device = device or next(model.parameters()).device
        self.param_dtype = _TORCH_DT[self.cfg["dtype"]]
        self.grad_dtype = _TORCH_DT[self.cfg["gradDtype"]]
        self.params: List[nn.Parameter] = []
        seen = set()
        for p in model.parameters():
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                self.params.append(p)
        self.offsets, n = [], 0
        for p in self.params:
            self.offsets.append(n)
            n += _round_up(p.numel(), 64)
        psz = torch.empty((), dtype=self.param_dtype).element_size()
        npad = _round_up(n, world * 64)
        heap = SymmetricHeap(npad * psz + 8 * (npad // world) + 65536, rank, world, group, self.device)
        self.state = AsyncShardState([heap], [rank], world, n, grad_dtype=self.grad_dtype, param_dtype=self.param_dtype,
                                     lr=float(self.cfg["learningRate"]), max_staleness=int(self.cfg["maximumStaleness"]),
                                     optimizer=self.cfg["optimizer"], policy=self.cfg["stalenessPolicy"], cfg=self.cfg,
                                     version_per_apply=self.cfg["versionCounts"] == "updates")
        flat = torch.zeros(self.state.n, device=self.device)
        for p, off in zip(self.params, self.offsets):
            flat[off:off + p.numel()] = p.data.reshape(-1).float()
        if world > 1:
            import torch.distributed as dist
            dist.broadcast(flat, src=0, group=group)
        self.state.init_weights(flat)
        lo = self.state.loc[rank]
        self.grad_flat = torch.zeros(self.state.n, dtype=self.grad_dtype, device=self.device)
        for p, off in zip(self.params, self.offsets):
            p.data = lo.params[off:off + p.numel()].view(p.shape)
            # a tied weight (two gradient producers, `_dfb_grad_parts`) keeps the plain path here: autograd accumulates both
            # parts and the hook below copies the sum (the in-place two-producer protocol is the sync engine's)
            if not getattr(p, "_dfb_shared", False) and int(getattr(p, "_dfb_grad_parts", 1)) == 1:
                p._dfb_grad_sink = self.grad_flat[off:off + p.numel()]
            p._dfb_grad_ready = lambda: None

            def hook(param, sink=self.grad_flat[off:off + p.numel()]):
                if param.grad is None:
                    return
                sink.copy_(param.grad.reshape(-1))
                param.grad = None

            p.register_post_accumulate_grad_hook(hook)
        self.version = 0
        # whole-step CUDA graph (pull -> fit -> push -> apply): every argument of the async kernels is constant from step to
        # step (versions, buffer choice and the staleness decision live in device memory), except Adam's step count
        self._graph_mode = bool(self.cfg["cudaGraph"]) and self.device.type == "cuda"
        if self._graph_mode and self.cfg["optimizer"] == "adam":
            self.state.step_dev = torch.zeros(1, dtype=torch.int32, device=self.device)    # advanced by a captured add_
        self._graph, self._static_in, self._static_loss, self._warm, self._graph_native = None, None, None, 0, 0
        if world > 1:
            import torch.distributed as dist
            dist.barrier(group=group)

    def _step_body(self, *batch: Any) -> torch.Tensor:
        if self.state.step_dev is not None:
            self.state.step_dev.add_(1)
        self.state.pull()
        loss = self.loss_fn(self.model, *batch)
        loss.backward()
        self.state.push(None, self.grad_flat)
        self.state.apply()
        return loss.detach()

    def step(self, *batch: Any) -> torch.Tensor:
        """pull -> fit -> push -> apply.  With `cudaGraph: True` (SGD / momentum) the step is captured after three eager
        steps and replayed; input shapes must then stay fixed."""
        if not self._graph_mode:
            loss = self._step_body(*batch)
            self.version += 1
            return loss
        if self._graph is None:
            if self._warm < 3:
                self._warm += 1
                loss = self._step_body(*batch)
                self.version += 1
                return loss
            self._static_in = [t.clone() if isinstance(t, torch.Tensor) else t for t in batch]
            torch.cuda.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            n0 = _lib.LAUNCHES[0]
            with torch.cuda.graph(self._graph):
                self._static_loss = self._step_body(*self._static_in)
            self._graph_native = _lib.LAUNCHES[0] - n0
            _lib.LAUNCHES[0] = n0
            self.state.steps -= 1                            # the capture pass launched nothing
        for dst, src in zip(self._static_in, batch):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src, non_blocking=True)
        self._graph.replay()
        _lib.LAUNCHES[0] += self._graph_native
        self.state.steps += 1
        self.version += 1
        return self._static_loss

    def run_dataset(self, wq, x_all: torch.Tensor, y_all: torch.Tensor, batch_size: int, max_steps: int = 1 << 30,
                    on_step: Optional[Callable[[int, torch.Tensor], None]] = None) -> int:
        """Train until the device work queue is exhausted.  `x_all` / `y_all` are device-resident; every step takes a
        ticket on the device, gathers its rows by that ticket, trains on it and marks it complete — the host never
        sees a batch index on the critical path.  It looks at a copy of the ticket a few steps late only to learn that
        the queue is done; the steps issued after that are no-ops on the device (ticket = 0xffffffff)."""
        xb = torch.empty((batch_size, *x_all.shape[1:]), dtype=x_all.dtype, device=x_all.device)
        yb = torch.empty((batch_size, *y_all.shape[1:]), dtype=y_all.dtype, device=y_all.device)
        lag = []          # (event, pinned copy of the ticket)
        n = 0
        while n < max_steps:
            t = wq.take_async()
            wq.gather(x_all, batch_size, xb)
            wq.gather(y_all, batch_size, yb)
            self.state.pull()
            loss = self.loss_fn(self.model, xb, yb)
            loss.backward()
            self.state.push(None, self.grad_flat, ticket=t)
            self.state.apply()
            wq.complete(accepted=self.state.loc[self.rank].push_state[15:16])
            host = torch.empty(2, dtype=torch.int32, pin_memory=True)
            host.copy_(t, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            lag.append((ev, host))
            n += 1
            self.version += 1
            if on_step is not None:
                on_step(n, loss.detach())
            done = False
            while lag and (lag[0][0].query() or len(lag) > 4):
                ev0, h0 = lag.pop(0)
                ev0.synchronize()
                if (int(h0[0]) & 0xFFFFFFFF) == 0xFFFFFFFF:
                    done = True
            if done:
                break
        return n

    def drain(self) -> None:
        """Apply whatever is still in flight (call on every rank after the last step, then barrier)."""
        if self.world > 1:
            import torch.distributed as dist
            torch.cuda.synchronize()
            dist.barrier(group=self.group)
        self.state.apply()
        self.state.apply()
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)
        self.state.pull()

    def counters(self) -> Dict[str, Any]:
        return self.state.read_ctl()
