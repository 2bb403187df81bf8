"""Phase timing, NVTX ranges and a progress watchdog.

Capability parity with the reference's tracing helper — `time(msg, action)` wraps an action and logs
`"<msg> took N ms"` (`/root/reference/src/server/abstract_server.ts:98-103`,
`/root/reference/src/client/abstract_client.ts:175-180`) — and with its only failure detection, the connect / upload
time-outs (`/root/reference/src/client/abstract_client.ts:12-13,148-158`).  On a GPU box wall-clock timing of an
asynchronous launch means nothing, so:

* `PhaseTimer` brackets a phase with CUDA events on the current stream (host clock on CPU), opens an NVTX range of the
  same name (visible in Nsight captures), and reports per phase the mean over calls and the MAX over ranks — the only
  meaningful multi-GPU number (B200_PROFILING.md: "time every multi-GPU number on the device as the max over ranks").
* `ProgressWatchdog` is the host-side watchdog of SURVEY §5.2 ("signal-word handshake; host-side watchdog on a pinned
  progress counter"): a thread polls a progress value (the engine's pinned step counter, written by the GPU at the end of
  every step) and calls `on_stall` when it has not moved for `timeout_s` — a rank spinning on a dead peer's flag inside a
  kernel cannot report anything itself.
"""
from __future__ import annotations

import contextlib
import threading
import time
from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple

import torch

__all__ = ["PhaseTimer", "ProgressWatchdog"]


class PhaseTimer:
    """`with timer.phase("forward"): ...` — device-timed phases with NVTX ranges; `summary()` reduces over ranks."""

    def __init__(self, device: Optional[torch.device] = None, nvtx: bool = True, log: Optional[Callable[[str], None]] = None):
        self.device = device or torch.device("cpu")
        self.cuda = self.device.type == "cuda"
        self.nvtx = bool(nvtx) and self.cuda
        self.log = log
        self._pending: List[Tuple[str, Any, Any]] = []          # (name, start event, stop event), resolved lazily
        self._ms: Dict[str, List[float]] = {}

    @contextlib.contextmanager
    def phase(self, name: str) -> Iterator[None]:
        if self.cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if self.nvtx:
                torch.cuda.nvtx.range_push(name)
            e0.record()
            try:
                yield
            finally:
                e1.record()
                if self.nvtx:
                    torch.cuda.nvtx.range_pop()
                self._pending.append((name, e0, e1))             # no synchronisation here: resolved in `collect`
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                ms = (time.perf_counter() - t0) * 1e3
                self._ms.setdefault(name, []).append(ms)
                if self.log:
                    self.log(f"{name} took {ms:.3f} ms")         # the reference's message shape

    def time(self, msg: str, action: Callable[[], Any]) -> Any:
        """The reference's `time(msg, action)`: run `action` inside a phase named `msg` and return its result."""
        with self.phase(msg):
            return action()

    def collect(self) -> None:
        """Resolve finished CUDA event pairs into milliseconds (synchronises the device once)."""
        if not self._pending:
            return
        torch.cuda.synchronize(self.device)
        for name, e0, e1 in self._pending:
            ms = e0.elapsed_time(e1)
            self._ms.setdefault(name, []).append(ms)
            if self.log:
                self.log(f"{name} took {ms:.3f} ms")
        self._pending = []

    def summary(self, group: Any = None) -> Dict[str, Dict[str, float]]:
        """`{phase: {"count", "mean_ms", "max_ms", "mean_ms_max_over_ranks"}}`; collective when torch.distributed is up."""
        self.collect()
        out: Dict[str, Dict[str, float]] = {}
        for name in sorted(self._ms):
            v = self._ms[name]
            out[name] = {"count": float(len(v)), "mean_ms": sum(v) / len(v), "max_ms": max(v)}
        try:
            import torch.distributed as dist
            up = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        except Exception:           # pragma: no cover - distributed not built
            up = False
        names = sorted(out)
        if up and names:
            dev = self.device if self.cuda and dist.get_backend(group) == "nccl" else torch.device("cpu")
            t = torch.tensor([out[n]["mean_ms"] for n in names], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            for n, val in zip(names, t.tolist()):
                out[n]["mean_ms_max_over_ranks"] = val
        else:
            for n in names:
                out[n]["mean_ms_max_over_ranks"] = out[n]["mean_ms"]
        return out

    def reset(self) -> None:
        self._pending, self._ms = [], {}


class ProgressWatchdog:
    """Calls `on_stall(last_value, seconds_without_progress)` when `get_progress()` stops changing for `timeout_s`.

    The watchdog arms itself at the first observed change (an idle engine is not a stalled one) and re-arms after every
    report, so a long stall is reported every `timeout_s`.  `stop()` joins the thread.
    """

    def __init__(self, get_progress: Callable[[], int], timeout_s: float, on_stall: Callable[[int, float], None],
                 poll_s: Optional[float] = None, arm_immediately: bool = False):
        if timeout_s <= 0:
            raise ValueError("timeout_s must be positive")
        self.get_progress, self.timeout_s, self.on_stall = get_progress, float(timeout_s), on_stall
        self.poll_s = poll_s if poll_s is not None else min(1.0, self.timeout_s / 4)
        self.stalls = 0
        self._armed = bool(arm_immediately)
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name="dfb-watchdog", daemon=True)

    def start(self) -> "ProgressWatchdog":
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread.is_alive():
            self._thread.join(timeout=5 * self.poll_s + 1)

    def pause(self) -> None:
        """Disarm until the next observed progress (call before a legitimately long host-side phase)."""
        self._armed = False

    def _run(self) -> None:
        last = self.get_progress()
        since = time.monotonic()
        while not self._stop.wait(self.poll_s):
            cur = self.get_progress()
            now = time.monotonic()
            if cur != last:
                last, since, self._armed = cur, now, True
                continue
            if self._armed and now - since >= self.timeout_s:
                self.stalls += 1
                try:
                    self.on_stall(int(last), now - since)
                finally:
                    since = now                                  # report again after another timeout_s
