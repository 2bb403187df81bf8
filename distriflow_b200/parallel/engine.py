"""DataParallelEngine: every worker is a GPU rank, the server is sharded across the ranks.

Roles of the reference mapped onto one 8xB200 box:
  * `FederatedServer.updateModel` — stack K uploads, mean, SGD step, broadcast
    (`/root/reference/src/server/federated_server.ts:92-116`) — is ONE kernel per gradient bucket
    (`csrc/comm.cu: sync_rs_opt_ag_kernel`): rank r owns slice r of every bucket (fp32 master weights +
    optimizer state), pulls that slice of all ranks' gradients over NVLink, applies the optimizer and pushes
    the refreshed working weights to everybody.  Buckets are launched on a side stream as soon as backward
    has produced them, so the collective overlaps the rest of backward.
  * `DistributedTfModel.fit` (`/root/reference/src/common/models.ts:137-142`) is the user's `loss_fn` +
    `backward`; weight-gradient GEMMs write straight into the flat gradient bucket (`_dfb_grad_sink`).
  * `model.version` / `savedGradient` / `maximumStaleness` (`/root/reference/README.md:23-30`) are
    `engine.version`, the gradient buckets, and the async engine's device-side staleness guard.

Back ends: "p2p" = own kernels over peer memory (the product); "nccl" / "gloo" = bucketed
`torch.distributed.all_reduce` + the same optimizer math in PyTorch (baseline, and CPU plumbing).
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from ..config import engine_config

__all__ = ["DataParallelEngine", "Bucket"]

_TORCH_DT = {"bf16": torch.bfloat16, "fp32": torch.float32}
_OPT = {"sgd": 0, "momentum": 1, "adam": 2, "none": 3}
_ALIGN = 64          # elements; every parameter starts 128-byte aligned inside its bucket (TMA needs 16)


class PeerLostError(RuntimeError):
    """A cross-GPU wait inside a kernel timed out: the named rank is down or hopelessly late."""


class Bucket:
    def __init__(self, index: int):
        self.index = index
        self.params: List[nn.Parameter] = []
        self.offsets: List[int] = []
        self.numel = 0            # padded to world * shard
        self.shard = 0
        self.param_off = 0        # byte offsets inside the symmetric heap
        self.grad_off = 0
        self.param_buf: Optional[torch.Tensor] = None
        self.grad_buf: Optional[torch.Tensor] = None
        self.master: Optional[torch.Tensor] = None
        self.m: Optional[torch.Tensor] = None
        self.v: Optional[torch.Tensor] = None
        self.pending = 0
        self.nopush = (0, 0)      # element range whose all-gather is fused into its consumer GEMM


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class DataParallelEngine:
    def __init__(self, model: nn.Module, loss_fn: Callable[..., torch.Tensor], config: Optional[Dict[str, Any]] = None,
                 rank: int = 0, world: int = 1, group=None, device: Optional[torch.device] = None, heap=None):
        self.cfg = engine_config(config)
        self.model, self.loss_fn = model, loss_fn
        self.rank, self.world, self.group = rank, world, group
        self.device = device or next(model.parameters()).device
        self.backend = self.cfg["backend"]
        if self.device.type != "cuda" and self.backend == "p2p":
            self.backend = "gloo"
        self.param_dtype = _TORCH_DT[self.cfg["dtype"]] if self.device.type == "cuda" else torch.float32
        self.grad_dtype = _TORCH_DT[self.cfg["gradDtype"]] if self.device.type == "cuda" else torch.float32
        self.lr = float(self.cfg["learningRate"])
        self.version = 0                    # number of optimizer steps applied (model version index)
        self.maximumStaleness = int(self.cfg["maximumStaleness"])
        self.comm_ctas = int(self.cfg["commCtas"])
        self.reserve_sms = int(self.cfg["reserveSMs"])
        self.nvls = False
        if self.cfg["nvls"] and self.world > 1 and self.backend == "p2p":
            from .symm import multicast_supported
            if self.param_dtype != torch.bfloat16 or self.grad_dtype != torch.bfloat16:
                raise ValueError("nvls: the in-switch reduction path takes bf16 gradients and bf16 working weights")
            if not multicast_supported(self.device):
                raise RuntimeError("nvls requested but this device cannot join an NVSwitch multicast object")
            self.nvls = True
        self._epoch = 0
        self._launches = 0                  # own-kernel launches issued by the engine (comm/optimizer kernels)
        self.versionCallbacks: List[Callable[[int, int], None]] = []
        self._build_buckets(heap)
        self._install_hooks()
        from ..ops import embed as _embed                  # process-wide, like the closing-barrier switch of the bucket kernels
        _embed.DETERMINISTIC = bool(self.cfg["deterministic"]) or _embed.DETERMINISTIC_ENV
        if self.cfg["fuseAllGather"]:
            first = getattr(model, "first_dense_kernel", None)
            w = first() if callable(first) else None
            if w is not None:
                self.fuse_allgather(w)
        if self.device.type == "cuda":
            self.comm_stream = torch.cuda.Stream(device=self.device)
        # one bucket: its own closing barrier is cheaper than a second launch (MNIST-sized models are one 32 KB bucket)
        self._close_per_step = self.backend == "p2p" and self.world > 1 and self.cfg["closeBarrier"] == "step" and len(self.buckets) > 1
        if self.backend == "p2p" and self.device.type == "cuda":
            _comm_sigs().dfb_sync_close_per_bucket(0 if self._close_per_step else 1)     # process-wide switch of the bucket kernels
        self._graph_mode = bool(self.cfg["cudaGraph"]) and self.device.type == "cuda" and self.backend == "p2p"
        self.tracer = None
        if self.cfg["trace"]:
            from ..utils.tracing import PhaseTimer
            self.tracer = PhaseTimer(self.device)
            self._graph_mode = False                       # events inside a captured graph cannot be timed: trace eagerly
        self.watchdog = None
        self._prog_pin = None
        if float(self.cfg["watchdogSeconds"]) > 0:
            self._start_watchdog(float(self.cfg["watchdogSeconds"]))
        self._graph, self._static_in, self._static_loss, self._warm, self._capturing, self._seq = None, None, None, 0, False, 0
        self._accumulating, self._acc_count = False, 0
        if self._graph_mode:
            self._epoch_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._step_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._lr_dev = torch.full((1,), self.lr, dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------------------------------ setup
    def _build_buckets(self, heap) -> None:
        params, seen = [], set()
        for p in self.model.parameters():
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                params.append(p)
        params.reverse()                                    # approximate backward order
        cap = max(1, int(self.cfg["bucketBytes"]) // torch.empty((), dtype=self.grad_dtype).element_size())
        buckets: List[Bucket] = []
        cur = Bucket(0)
        for p in params:
            n = _round_up(p.numel(), _ALIGN)
            if cur.params and cur.numel + n > cap:
                buckets.append(cur)
                cur = Bucket(len(buckets))
            cur.params.append(p)
            cur.offsets.append(cur.numel)
            cur.numel += n
        if cur.params:
            buckets.append(cur)
        R = self.world
        psz = torch.empty((), dtype=self.param_dtype).element_size()
        gsz = torch.empty((), dtype=self.grad_dtype).element_size()
        total = 0
        for b in buckets:
            b.numel = _round_up(b.numel, R * _ALIGN)
            b.shard = b.numel // R
            total += _round_up(b.numel * psz, 256) + _round_up(b.numel * gsz, 256)
        self.heap = None
        if self.backend == "p2p":
            from .symm import SymmetricHeap
            self.heap = heap or SymmetricHeap(total + 4096, self.rank, self.world, self.group, self.device, multicast=self.nvls)
            if self.nvls and not self.heap.mc_base:
                raise RuntimeError("nvls: the supplied symmetric heap has no multicast mapping")
            if self.world > 1 and self.device.type == "cuda":
                self.heap.set_barrier_timeout(float(self.cfg["barrierTimeoutSeconds"]))
        for b in buckets:
            flat32 = torch.zeros(b.numel, dtype=torch.float32, device=self.device)
            for p, off in zip(b.params, b.offsets):
                flat32[off:off + p.numel()].copy_(p.data.reshape(-1).float())
            if self.world > 1:
                import torch.distributed as dist
                if dist.is_available() and dist.is_initialized():
                    dist.broadcast(flat32, src=0, group=self.group)      # identical start on every rank (bootstrap)
            if self.heap is not None:
                b.param_off = self.heap.alloc(b.numel * psz)
                b.grad_off = self.heap.alloc(b.numel * gsz)
                b.param_buf = self.heap.view(b.param_off, (b.numel,), self.param_dtype)
                b.grad_buf = self.heap.view(b.grad_off, (b.numel,), self.grad_dtype)
            else:
                b.param_buf = torch.empty(b.numel, dtype=self.param_dtype, device=self.device)
                b.grad_buf = torch.zeros(b.numel, dtype=self.grad_dtype, device=self.device)
            b.param_buf.copy_(flat32)
            lo = self.rank * b.shard
            owned = (lo, lo + b.shard) if self.backend == "p2p" else (0, b.numel)
            b.master = flat32[owned[0]:owned[1]].clone()
            if self.cfg["optimizer"] in ("momentum", "adam"):
                b.m = torch.zeros_like(b.master)
            if self.cfg["optimizer"] == "adam":
                b.v = torch.zeros_like(b.master)
            for p, off in zip(b.params, b.offsets):
                p.data = b.param_buf[off:off + p.numel()].view(p.shape)
                p.grad = None
            del flat32
        self.buckets = buckets
        self._bucket_of = {id(p): b for b in buckets for p in b.params}

    def _install_hooks(self) -> None:
        for b in self.buckets:
            for p, off in zip(b.params, b.offsets):
                sink = b.grad_buf[off:off + p.numel()]
                if not getattr(p, "_dfb_shared", False):
                    p._dfb_grad_sink = sink                    # ops write the weight gradient here and call ready()
                p._dfb_grad_ready = (lambda bb=b, pp=p: self._param_ready(bb, pp))

                def hook(param, sink=sink, bb=b):              # parameters whose gradient comes from plain autograd
                    if param.grad is None:                     # autograd fires the hook even for an undefined gradient:
                        return                                 # that parameter went through the sink path already
                    sink.copy_(param.grad.reshape(-1))
                    param.grad = None
                    # a tied weight (`_dfb_grad_parts` producers) whose gradients were accumulated by autograd reports them all
                    for _ in range(max(1, int(getattr(param, "_dfb_grad_parts", 1)) - int(getattr(param, "_dfb_reported", 0)))):
                        self._param_ready(bb, param)

                p.register_post_accumulate_grad_hook(hook)

    def fuse_allgather(self, weight: nn.Parameter) -> bool:
        """Fuse the all-gather of `weight` (a Dense kernel `[in, out]`, bf16) into the GEMM that consumes it first."""
        if self.backend != "p2p" or self.world == 1 or weight.dim() != 2 or self.param_dtype != torch.bfloat16:
            return False
        b = self._bucket_of.get(id(weight))
        if b is None or b.nopush != (0, 0):
            return False
        from .ag_fusion import build_plan
        off = b.offsets[[id(p) for p in b.params].index(id(weight))]
        K, N = weight.shape
        bases = [self.heap.peer_bases[q] + b.param_off for q in range(self.world)]
        weight._dfb_ag = build_plan(K, N, 2, off, b.shard, self.rank, self.world, bases, self.device)
        b.nopush = (off, off + weight.numel())
        return True

    # ------------------------------------------------------------------------------------ step
    def on_new_version(self, cb: Callable[[int, int], None]) -> None:
        self.versionCallbacks.append(cb)

    def _param_ready(self, b: Bucket, p: Optional[nn.Parameter] = None) -> None:
        if p is not None and getattr(p, "_dfb_grad_parts", 1) > 1:
            p._dfb_reported = int(getattr(p, "_dfb_reported", 0)) + 1
        b.pending -= 1
        if self._accumulating:                                 # micro-batch of an unfinished version: nothing is launched
            return
        if b.pending == 0 and self.cfg["debugPoison"]:
            # Race detector: from the moment a bucket is reported complete its working weights may be overwritten by the
            # fused kernel, so nothing enqueued later in this backward may read them.  Overwrite them with NaN right here
            # (same stream, deterministic) and run the real kernel in the tail, which rewrites every element from the fp32
            # masters; a kernel that read a weight after reporting it turns the next loss into NaN.
            lo, hi = b.nopush
            b.param_buf[:lo].fill_(float("nan"))
            b.param_buf[hi:].fill_(float("nan"))
            return
        if b.pending == 0 and self.cfg["overlap"]:
            self._launch_bucket(b)

    def zero_pending(self) -> None:
        for b in self.buckets:
            b.pending = sum(int(getattr(p, "_dfb_grad_parts", 1)) for p in b.params)   # a tied weight reports once per producer
            for p in b.params:
                if getattr(p, "_dfb_grad_parts", 1) > 1:
                    p._dfb_reported = 0
            b.launched = False

    def step(self, *batch: Any) -> torch.Tensor:
        """One synchronous data-parallel step: forward, backward (+ overlapped bucket collectives), new version.

        With `cudaGraph: True` the first steps run eagerly (warm-up), then the whole step — forward, backward, the fused
        bucket kernels on the side stream — is captured once into a CUDA graph and replayed; launch-bound models (MNIST,
        ResNet-20) then run at device speed.  Inputs must keep their shapes; the learning rate is baked into the graph."""
        if not self._graph_mode:
            return self._step_body(*batch)
        if self._graph is None:
            if self._warm < 3:
                self._warm += 1
                return self._step_body(*batch)
            self._static_in = [t.clone() if isinstance(t, torch.Tensor) else t for t in batch]
            torch.cuda.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            self._capturing = True
            from ..ops import _lib
            n0 = _lib.LAUNCHES[0]
            with torch.cuda.graph(self._graph):
                self._static_loss = self._step_body(*self._static_in, _count=False)
            self._graph_native = _lib.LAUNCHES[0] - n0        # native kernels recorded in the graph (none ran during capture)
            _lib.LAUNCHES[0] = n0
            self._capturing = False
        for dst, src in zip(self._static_in, batch):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src, non_blocking=True)
        self._graph.replay()
        from ..ops import _lib
        _lib.LAUNCHES[0] += self._graph_native                # each replay launches every native kernel node of the graph
        old = self.version
        self.version += 1
        for cb in self.versionCallbacks:
            cb(old, self.version)
        return self._static_loss

    def _step_body(self, *batch: Any, _count: bool = True) -> torch.Tensor:
        self.zero_pending()
        self._seq = 0
        if self._graph_mode:
            self._step_dev.add_(1)
        if self.device.type == "cuda" and self.world > 1 and self.backend == "p2p" and self.cfg["overlap"] and self.reserve_sms:
            from ..ops import dense as _lin
            _lin.set_max_ctas(_lin.NUM_SMS - self.reserve_sms)  # optionally leave SMs free for the collective kernels
        tr = self.tracer
        with (tr.phase("forward") if tr else _NULL):
            loss = self.loss_fn(self.model, *batch)
        with (tr.phase("backward") if tr else _NULL):          # bucket kernels launched from hooks overlap this phase
            loss.backward()
        with (tr.phase("collective_tail") if tr else _NULL):   # what is left of the gradient path after backward ends
            for b in self.buckets:                             # anything not launched from a hook (unused params, overlap off)
                if not getattr(b, "launched", False):
                    self._launch_bucket(b)
            if self._close_per_step:
                self._close_step()
            if self.device.type == "cuda":
                torch.cuda.current_stream().wait_stream(self.comm_stream)
                from ..ops import dense as _lin
                _lin.set_max_ctas(0)
        if self._prog_pin is not None:                         # device-written progress word for the watchdog (capturable)
            self._prog_dev.add_(1)
            self._prog_pin.copy_(self._prog_dev, non_blocking=True)
        if self._graph_mode and self.world > 1:
            self._epoch_dev.add_(len(self.buckets))            # next step's barrier generations start after this step's
        self._acc_count = 0
        if _count:
            old = self.version
            self.version += 1
            for cb in self.versionCallbacks:
                cb(old, self.version)
        return loss.detach()

    def accumulate(self, *batch: Any) -> torch.Tensor:
        """Forward + backward of one more micro-batch of the CURRENT version: its gradient is added to the bucket-shaped
        `savedGradient` (fp32) and nothing is communicated or applied.  The next `step()` folds the saved sum into its own
        gradient, divides by the number of contributions and runs the fused reduce-scatter + optimizer + all-gather once.
        This is `minUpdatesPerVersion = K x workers` of the reference (`/root/reference/src/server/federated_server.ts:73-90`:
        K uploads are buffered and averaged before one update) with the buffer kept on the GPU (README `savedGradient`)."""
        if self._graph is not None:
            raise RuntimeError("accumulate() after the step has been captured into a CUDA graph: build the engine with cudaGraph=False")
        self._graph_mode = False
        self.zero_pending()
        self._accumulating = True
        try:
            loss = self.loss_fn(self.model, *batch)
            loss.backward()
        finally:
            self._accumulating = False
        for b in self.buckets:
            if getattr(b, "acc", None) is None:
                b.acc = torch.zeros(b.numel, dtype=torch.float32, device=self.device)
            if self._acc_count == 0:
                b.acc.copy_(b.grad_buf)
            else:
                b.acc.add_(b.grad_buf)
        self._acc_count += 1
        return loss.detach()

    @property
    def savedGradient(self) -> Optional[List[torch.Tensor]]:
        """The aggregated gradient of the micro-batches accumulated so far (one flat fp32 tensor per bucket), or None."""
        if self._acc_count == 0:
            return None
        return [b.acc for b in self.buckets]

    @property
    def savedCount(self) -> int:
        return self._acc_count

    def _launch_bucket(self, b: Bucket) -> None:
        b.launched = True
        if self._acc_count:                                     # fold the saved micro-batch gradients in: mean over all of them
            k = float(self._acc_count + 1)
            b.grad_buf.copy_((b.acc + b.grad_buf.float()) / k)
        if self.backend == "p2p":
            self._launch_bucket_p2p(b)
        else:
            self._launch_bucket_torch(b)

    def _close_step(self) -> None:
        """The step's ONE closing barrier, behind the last bucket kernel on the side stream: when it completes every peer has
        pushed all its refreshed slices into this rank's weight buffers (and stopped reading this rank's gradients)."""
        from ..ops import _lib
        from ..ops._lib import check, ptr
        L = _comm_sigs()
        graph = self._graph_mode
        epoch = self._seq if graph else self._epoch
        if epoch == 0:
            return
        with torch.cuda.stream(self.comm_stream):
            check(L.dfb_step_close(self.heap.ptrs(0), self.rank, self.world, epoch, ptr(self._epoch_dev) if graph else None,
                                   _lib.stream()), "dfb_step_close")
        self._launches += 1

    def _launch_bucket_p2p(self, b: Bucket) -> None:
        from ..ops import _lib
        from ..ops._lib import check, dt, ptr
        L = _comm_sigs()
        main = torch.cuda.current_stream()
        self.comm_stream.wait_stream(main)
        c = self.cfg
        if self.nvls:                 # the switch reduces the gradients and broadcasts the weights (multimem.ld_reduce / .st)
            graph = self._graph_mode
            if graph:
                self._seq += 1
                epoch = self._seq
            else:
                self._epoch += 1
                epoch = self._epoch
            with torch.cuda.stream(self.comm_stream):
                check(L.dfb_sync_nvls(self.heap.mc_ptr(b.grad_off), self.heap.mc_ptr(b.param_off), self.heap.ptrs(b.param_off),
                                      self.heap.ptrs(0), ptr(b.master), ptr(b.m), ptr(b.v), b.numel, b.shard, self.rank, self.world,
                                      epoch, _OPT[c["optimizer"]], self.lr, c["momentum"], c["beta1"], c["beta2"], c["eps"],
                                      c["weightDecay"], self.version + 1, self.comm_ctas, b.nopush[0], b.nopush[1],
                                      ptr(self._epoch_dev) if graph else None, ptr(self._step_dev) if graph else None,
                                      ptr(self._lr_dev) if graph else None, _lib.stream()), "dfb_sync_nvls")
            self._launches += 1
            return
        if self._graph_mode:          # replayable launch: generation and step count are read from device memory
            self._seq += 1
            with torch.cuda.stream(self.comm_stream):
                check(L.dfb_sync_rs_opt_ag_replayable(self.heap.ptrs(b.grad_off), self.heap.ptrs(b.param_off), self.heap.ptrs(0),
                                                      ptr(b.master), ptr(b.m), ptr(b.v), b.numel, b.shard, self.rank, self.world,
                                                      self._seq if self.world > 1 else 0, dt(b.grad_buf), dt(b.param_buf),
                                                      _OPT[c["optimizer"]], self.lr, c["momentum"], c["beta1"], c["beta2"], c["eps"],
                                                      c["weightDecay"], self.comm_ctas if self.world > 1 else 148 * 4, b.nopush[0],
                                                      b.nopush[1], ptr(self._epoch_dev), ptr(self._step_dev), ptr(self._lr_dev),
                                                      _lib.stream()),
                      "dfb_sync_rs_opt_ag_replayable")
            self._launches += 1
            return
        epoch = 0
        if self.world > 1:
            self._epoch += 1
            epoch = self._epoch
        with torch.cuda.stream(self.comm_stream):
            check(L.dfb_sync_rs_opt_ag(self.heap.ptrs(b.grad_off), self.heap.ptrs(b.param_off), self.heap.ptrs(0),
                                       ptr(b.master), ptr(b.m), ptr(b.v), None, b.numel, b.shard, self.rank, self.world,
                                       epoch, 1, dt(b.grad_buf), dt(b.param_buf), _OPT[c["optimizer"]], self.lr,
                                       c["momentum"], c["beta1"], c["beta2"], c["eps"], c["weightDecay"], self.version + 1,
                                       self.comm_ctas if self.world > 1 else 148 * 4, b.nopush[0], b.nopush[1], _lib.stream()),
                  "dfb_sync_rs_opt_ag")
        self._launches += 1

    def _launch_bucket_torch(self, b: Bucket) -> None:
        """Baseline path: all-reduce (NCCL/gloo) + the same optimizer math in PyTorch on a full fp32 master."""
        g = b.grad_buf
        if self.world > 1:
            import torch.distributed as dist
            g32 = g.float() if self.device.type != "cuda" else g
            dist.all_reduce(g32, group=self.group)
            g = g32
        g = g.float() / self.world
        c = self.cfg
        if c["weightDecay"]:
            g = g + c["weightDecay"] * b.master
        if c["optimizer"] == "sgd":
            b.master.add_(g, alpha=-self.lr)
        elif c["optimizer"] == "momentum":
            b.m.mul_(c["momentum"]).add_(g)
            b.master.add_(b.m, alpha=-self.lr)
        else:
            t = self.version + 1
            b.m.mul_(c["beta1"]).add_(g, alpha=1 - c["beta1"])
            b.v.mul_(c["beta2"]).addcmul_(g, g, value=1 - c["beta2"])
            mh = b.m / (1 - c["beta1"] ** t)
            vh = b.v / (1 - c["beta2"] ** t)
            b.master.addcdiv_(mh, vh.sqrt().add_(c["eps"]), value=-self.lr)
        b.param_buf.copy_(b.master)

    # ------------------------------------------------------------------------------------ state
    def check_health(self) -> None:
        """Raise `PeerLostError` if a kernel of this rank ever gave up waiting for a peer (`barrierTimeoutSeconds`): the steps
        since then averaged incomplete gradients.  Called wherever weights leave the engine (checkpoints, `full_precision_weights`)."""
        lost = self.peer_failure()
        if lost is not None:
            raise PeerLostError(f"rank {self.rank}: rank {lost[0]} never reached barrier generation {lost[1]} within "
                                f"{self.cfg['barrierTimeoutSeconds']} s; weights since then are not trustworthy")

    def full_precision_weights(self) -> List[torch.Tensor]:
        """fp32 master weights of every parameter, in `model.parameters()` order (gathers the shards)."""
        self.check_health()
        out: Dict[int, torch.Tensor] = {}
        for b in self.buckets:
            if self.backend == "p2p" and self.world > 1:
                import torch.distributed as dist
                parts = [torch.empty_like(b.master) for _ in range(self.world)]
                dist.all_gather(parts, b.master, group=self.group)
                flat = torch.cat(parts)
            else:
                flat = b.master if b.master.numel() == b.numel else torch.cat([b.master])
            for p, off in zip(b.params, b.offsets):
                out[id(p)] = flat[off:off + p.numel()].view(p.shape).clone()
        seen, res = set(), []
        for p in self.model.parameters():
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                res.append(out[id(p)])
        return res

    def save_checkpoint(self, directory: str) -> Optional[str]:
        """Write the tf.js LayersModel layout from the fp32 master weights (rank 0 writes)."""
        weights = self.full_precision_weights()
        if self.rank != 0:
            return None
        from ..checkpoint import save_layers_model
        live = [p for p in self.model.parameters() if p.requires_grad]
        saved = [p.data for p in live]
        try:
            for p, w in zip(live, weights):
                p.data = w
            return save_layers_model(self.model, directory)
        finally:
            for p, d in zip(live, saved):
                p.data = d

    # ------------------------------------------------------------------------------------ resume
    def _gather_flat(self, t: Optional[torch.Tensor], b: Bucket) -> Optional[torch.Tensor]:
        """Whole-bucket fp32 tensor from this rank's shard `t` (master / m / v)."""
        if t is None:
            return None
        if self.backend == "p2p" and self.world > 1:
            import torch.distributed as dist
            parts = [torch.empty_like(t) for _ in range(self.world)]
            dist.all_gather(parts, t, group=self.group)
            return torch.cat(parts)
        return t

    def state_dict(self) -> Dict[str, Any]:
        """Everything a resumed run needs: version (= optimizer step count), fp32 masters and optimizer moments per
        parameter.  Shards are gathered, so the result does not depend on the world size it was saved at (collective: call
        on every rank).  The reference checkpoints weights only (server/models.ts:132-138); optimizer state is new."""
        self.check_health()
        names = {id(p): n for n, p in self.model.named_parameters()}
        params: Dict[str, Dict[str, torch.Tensor]] = {}
        for b in self.buckets:
            flats = {k: self._gather_flat(getattr(b, k, None), b) for k in ("master", "m", "v")}
            for p, off in zip(b.params, b.offsets):
                entry = {k: f[off:off + p.numel()].view(p.shape).cpu().clone() for k, f in flats.items() if f is not None}
                params[names[id(p)]] = entry
        buffers = {n: b.detach().cpu().clone() for n, b in self.model.named_buffers()}     # BatchNorm running statistics etc.
        return {"format": "distriflow_b200.engine/1", "version": int(self.version), "optimizer": self.cfg["optimizer"],
                "learningRate": self.lr, "params": params, "buffers": buffers}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        """Inverse of `state_dict` at ANY world size: every rank takes its slice of every bucket, the working copies are
        refreshed from the masters, the step counter continues."""
        if sd.get("format") != "distriflow_b200.engine/1":
            raise ValueError("not an engine state dict")
        names = {id(p): n for n, p in self.model.named_parameters()}
        for b in self.buckets:
            flats = {"master": torch.zeros(b.numel, dtype=torch.float32, device=self.device)}
            for k in ("m", "v"):
                if getattr(b, k, None) is not None:
                    flats[k] = torch.zeros(b.numel, dtype=torch.float32, device=self.device)
            for p, off in zip(b.params, b.offsets):
                entry = sd["params"].get(names[id(p)])
                if entry is None or tuple(entry["master"].shape) != tuple(p.shape):
                    raise ValueError(f"state dict has no matching entry for parameter {names[id(p)]}")
                for k, f in flats.items():
                    if k in entry:                       # a state saved by plain SGD resumes under Adam with zero moments
                        f[off:off + p.numel()].copy_(entry[k].reshape(-1))
            lo = self.rank * b.shard
            lo, hi = (lo, lo + b.shard) if self.backend == "p2p" else (0, b.numel)
            for k, f in flats.items():
                getattr(b, k).copy_(f[lo:hi])
            b.param_buf.copy_(flats["master"])
        saved_buffers = sd.get("buffers", {})
        with torch.no_grad():
            for n, buf in self.model.named_buffers():
                if n in saved_buffers:
                    buf.copy_(saved_buffers[n].to(buf.device, buf.dtype))
        self.version = int(sd["version"])
        if self._graph_mode:
            self._step_dev.fill_(self.version)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        if self.world > 1:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.barrier(group=self.group)

    def save_state(self, directory: str) -> Optional[str]:
        """`<directory>/engine_state.pt` (rank 0 writes; collective).  Together with `save_checkpoint` (tf.js layout of the
        weights) this is the resume point of a synchronous run."""
        import os
        sd = self.state_dict()
        if self.rank != 0:
            return None
        os.makedirs(directory, exist_ok=True)
        path = os.path.join(directory, "engine_state.pt")
        tmp = path + ".tmp"
        torch.save(sd, tmp)
        os.replace(tmp, path)                            # never leaves a half-written state behind
        return path

    def load_state(self, directory: str) -> None:
        import os
        self.load_state_dict(torch.load(os.path.join(directory, "engine_state.pt"), map_location="cpu"))

    def set_learning_rate(self, lr: float) -> None:
        """Learning-rate schedules: takes effect at the next step, also when the step replays from a CUDA graph (the bucket
        kernels then read the rate from device memory)."""
        self.lr = float(lr)
        if getattr(self, "_lr_dev", None) is not None:
            self._lr_dev.fill_(self.lr)

    # ------------------------------------------------------------------------------------ observability
    def trace_summary(self) -> Dict[str, Dict[str, float]]:
        """Per-phase device time (mean over steps, max over ranks); needs `trace: True`.  Collective."""
        if self.tracer is None:
            raise RuntimeError("engine was built without trace=True")
        return self.tracer.summary(self.group)

    def _start_watchdog(self, seconds: float) -> None:
        from ..utils.tracing import ProgressWatchdog
        if self.device.type == "cuda":
            self._prog_dev = torch.zeros(1, dtype=torch.int64, device=self.device)
            self._prog_pin = torch.zeros(1, dtype=torch.int64).pin_memory()
            read = lambda: int(self._prog_pin[0])              # noqa: E731  (written by the GPU at the end of every step)
        else:
            read = lambda: int(self.version)                   # noqa: E731
        self.stall_reports: List[Tuple[int, float]] = []

        def on_stall(last: int, secs: float) -> None:
            self.stall_reports.append((last, secs))
            import sys
            lost = self.peer_failure(_side=True)
            who = f"rank {lost[0]} never reached barrier generation {lost[1]}" if lost else "a peer may be down or a collective is stuck"
            print(f"[distriflow_b200] rank {self.rank}: no step completed for {secs:.1f} s (last finished step {last}); {who}",
                  file=sys.stderr, flush=True)
            for cb in self.stallCallbacks:
                cb(self.rank, last, secs)

        self.stallCallbacks: List[Callable[[int, int, float], None]] = []
        self.watchdog = ProgressWatchdog(read, seconds, on_stall).start()

    def peer_failure(self, _side: bool = False) -> Optional[Tuple[int, int]]:
        """(rank, barrier generation) when one of this rank's kernels gave up waiting for a peer (`barrierTimeoutSeconds`)
        since the last call, else None.  The weights of such a step are not trustworthy: stop, or restore a checkpoint.
        Reference: the server drops a worker whose socket closes (abstract_server.ts:118-126); between GPUs there is no
        socket to close, so the waiter's time-out is the detector."""
        if self.heap is None or self.device.type != "cuda" or self.world == 1:
            return None
        if _side:
            if getattr(self, "_health_stream", None) is None:
                self._health_stream = torch.cuda.Stream(device=self.device)
            self.lost_peer = self.heap.barrier_failure(self._health_stream) or getattr(self, "lost_peer", None)
            return self.lost_peer
        self.lost_peer = self.heap.barrier_failure() or getattr(self, "lost_peer", None)
        return self.lost_peer

    def on_stall(self, cb: Callable[[int, int, float], None]) -> None:
        """`cb(rank, last_finished_step, seconds)` from the watchdog thread (needs `watchdogSeconds > 0`)."""
        if self.watchdog is None:
            raise RuntimeError("engine was built without watchdogSeconds > 0")
        self.stallCallbacks.append(cb)

    def close(self) -> None:
        if self.watchdog is not None:
            self.watchdog.stop()
            self.watchdog = None

    def gpu_launches_per_step(self) -> int:
        return len(self.buckets)


import contextlib as _contextlib

_NULL = _contextlib.nullcontext()


def _comm_sigs():
    import ctypes
    from ..ops import _lib
    from ..ops._lib import c_float, c_int, c_void_p
    L = _lib.lib()
    if not getattr(L, "_dfb_comm_sigs", False):
        P = ctypes.POINTER(c_void_p)
        L.dfb_sync_rs_opt_ag.restype = c_int
        L.dfb_sync_rs_opt_ag.argtypes = [P, P, P, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_long, ctypes.c_long, c_int,
                                         c_int, ctypes.c_uint, c_int, c_int, c_int, c_int, c_float, c_float, c_float, c_float,
                                         c_float, c_float, c_int, c_int, ctypes.c_long, ctypes.c_long, c_void_p]
        L.dfb_sync_rs_opt_ag_replayable.restype = c_int
        L.dfb_sync_rs_opt_ag_replayable.argtypes = [P, P, P, c_void_p, c_void_p, c_void_p, ctypes.c_long, ctypes.c_long, c_int, c_int,
                                                    ctypes.c_uint, c_int, c_int, c_int, c_float, c_float, c_float, c_float, c_float,
                                                    c_float, c_int, ctypes.c_long, ctypes.c_long, c_void_p, c_void_p, c_void_p, c_void_p]
        L.dfb_sync_nvls.restype = c_int
        L.dfb_sync_nvls.argtypes = [c_void_p, c_void_p, P, P, c_void_p, c_void_p, c_void_p, ctypes.c_long, ctypes.c_long, c_int, c_int,
                                    ctypes.c_uint, c_int, c_float, c_float, c_float, c_float, c_float, c_float, c_int, c_int,
                                    ctypes.c_long, ctypes.c_long, c_void_p, c_void_p, c_void_p, c_void_p]
        L.dfb_sync_close_per_bucket.restype = c_int
        L.dfb_sync_close_per_bucket.argtypes = [c_int]
        L.dfb_step_close.restype = c_int
        L.dfb_step_close.argtypes = [P, c_int, c_int, ctypes.c_uint, c_void_p, c_void_p]
        L.dfb_allgather_push.restype = c_int
        L.dfb_allgather_push.argtypes = [P, P, ctypes.c_long, ctypes.c_long, c_int, c_int, ctypes.c_uint, c_int, c_void_p]
        L.dfb_rank_barrier.restype = c_int
        L.dfb_rank_barrier.argtypes = [P, c_int, c_int, ctypes.c_uint, c_void_p]
        L.dfb_set_barrier_timeout_us.restype = c_int
        L.dfb_set_barrier_timeout_us.argtypes = [ctypes.c_int64, c_void_p]
        L.dfb_barrier_failures.restype = c_int
        L.dfb_barrier_failures.argtypes = [c_void_p, c_void_p, c_void_p]
        L.dfb_async_push.restype = c_int
        L.dfb_async_push.argtypes = [c_void_p, P, P, P, c_void_p, c_void_p, ctypes.c_long, ctypes.c_long, c_int, c_int, c_int,
                                     c_float, c_int, c_int, c_int, c_void_p, c_void_p]
        L.dfb_async_apply.restype = c_int
        L.dfb_async_apply.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      ctypes.c_long, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_float, c_int,
                                      c_int, c_void_p, c_int, c_void_p]
        L.dfb_async_pull.restype = c_int
        L.dfb_async_pull.argtypes = [P, P, c_void_p, c_void_p, ctypes.c_long, ctypes.c_long, c_int, c_int, c_int, c_void_p]
        L.dfb_async_read_ctl.restype = c_int
        L.dfb_async_read_ctl.argtypes = [c_void_p, c_void_p, c_void_p]
        L.dfb_wq_init.restype = c_int
        L.dfb_wq_init.argtypes = [c_void_p, ctypes.c_uint, ctypes.c_uint, c_void_p]
        L.dfb_wq_take.restype = c_int
        L.dfb_wq_take.argtypes = [c_void_p, c_void_p, c_void_p]
        L.dfb_wq_complete.restype = c_int
        L.dfb_wq_complete.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p]
        L.dfb_wq_words.restype = c_int
        L.dfb_wq_words.argtypes = [ctypes.c_uint]
        L.dfb_wq_pad_offset_words.restype = c_int
        L.dfb_wq_pad_offset_words.argtypes = []
        L.dfb_wq_pad_capacity.restype = c_int
        L.dfb_wq_pad_capacity.argtypes = []
        L.dfb_gather_batch.restype = c_int
        L.dfb_gather_batch.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_long, c_int, ctypes.c_long, c_void_p]
        L._dfb_comm_sigs = True
    return L
