"""Phase timer and progress watchdog (`utils/tracing.py`), CPU side.

Reference: `time(msg, action)` logging "<msg> took N ms" (`/root/reference/src/server/abstract_server.ts:98-103`) and the
10 s / 5 s connect / upload time-outs as the only failure detection (`/root/reference/src/client/abstract_client.ts:12-13`).
"""
import threading
import time

import pytest
import torch

from distriflow_b200.utils.tracing import PhaseTimer, ProgressWatchdog


def test_phase_timer_cpu_logs_reference_message_shape():
    lines = []
    t = PhaseTimer(log=lines.append)
    assert t.time("computing new weights", lambda: 41 + 1) == 42
    with t.phase("fit"):
        time.sleep(0.01)
    s = t.summary()
    assert set(s) == {"computing new weights", "fit"} and s["fit"]["count"] == 1 and s["fit"]["mean_ms"] >= 9
    assert s["fit"]["mean_ms_max_over_ranks"] == s["fit"]["mean_ms"]
    assert lines[0].startswith("computing new weights took ") and lines[0].endswith(" ms")
    t.reset()
    assert t.summary() == {}


def test_watchdog_reports_a_stall_and_rearms():
    progress = [0]
    reports = []
    fired = threading.Event()

    def on_stall(last, secs):
        reports.append((last, secs))
        fired.set()

    w = ProgressWatchdog(lambda: progress[0], 0.2, on_stall, poll_s=0.02).start()
    try:
        time.sleep(0.3)
        assert reports == []                   # never moved: idle, not stalled
        progress[0] = 7                        # first progress arms the watchdog
        assert fired.wait(2.0)
        assert reports[0][0] == 7 and reports[0][1] >= 0.2
        n = len(reports)
        progress[0] = 8                        # progress again: quiet until the next stall
        time.sleep(0.1)
        assert len(reports) == n
        time.sleep(0.4)
        assert len(reports) > n and reports[-1][0] == 8
    finally:
        w.stop()
    assert w.stalls == len(reports)
    with pytest.raises(ValueError):
        ProgressWatchdog(lambda: 0, 0, on_stall)


def test_engine_trace_and_watchdog_on_cpu():
    import torch.nn as nn
    import torch.nn.functional as F
    from distriflow_b200.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(6, 8), nn.ReLU(), nn.Linear(8, 3))
    eng = DataParallelEngine(net, lambda m, x, y: F.cross_entropy(m(x), y),
                             {"optimizer": "sgd", "learningRate": 0.1, "dtype": "fp32", "gradDtype": "fp32", "trace": True,
                              "watchdogSeconds": 0.2}, device=torch.device("cpu"))
    seen = []
    eng.on_stall(lambda rank, last, secs: seen.append((rank, last)))
    x, y = torch.randn(4, 6), torch.randint(0, 3, (4,))
    for _ in range(3):
        eng.step(x, y)
    s = eng.trace_summary()
    assert set(s) == {"forward", "backward", "collective_tail"} and s["forward"]["count"] == 3
    time.sleep(0.6)                            # no more steps: the watchdog reports the stall after version 3
    eng.close()
    assert seen and seen[0] == (0, 3) and eng.stall_reports[0][0] == 3


def test_throughput_meter_and_jsonl_logger_on_cpu(tmp_path):
    import json
    from distriflow_b200.utils.metrics import JsonlLogger, ThroughputMeter
    if torch.cuda.is_available():
        pytest.skip("host-clock branch is for CPU-only boxes")
    m = ThroughputMeter(samples_per_step=32, grad_bytes=1 << 20, world=2)
    m.start()
    for _ in range(4):
        time.sleep(0.005)
        m.step()
    r = m.stop()
    assert r["ms_per_step"] >= 4.5 and r["samples_per_sec"] > 0 and r["grad_bus_gbs"] > 0
    path = str(tmp_path / "log.jsonl")
    with JsonlLogger(path) as log:
        log.log(step=1, loss=0.5)
        log.log(step=2, loss=0.25, version="v2")
    rows = [json.loads(line) for line in open(path)]
    assert [r["step"] for r in rows] == [1, 2] and rows[1]["version"] == "v2" and "ts" in rows[0]


def test_device_prefetcher_is_a_pass_through_on_cpu():
    from distriflow_b200.utils.prefetch import DevicePrefetcher
    batches = [(torch.full((2, 3), float(i)), torch.tensor([i])) for i in range(5)]
    pulled = []

    def gen():
        for b in batches:
            pulled.append(int(b[1]))
            yield b

    pf = DevicePrefetcher(gen(), torch.device("cpu"), depth=2)
    first = next(pf)
    assert first[0] is batches[0][0] and pulled == [0, 1, 2]          # depth 2 staged, then refilled after handing one out
    rest = list(pf)
    assert [int(b[1]) for b in rest] == [1, 2, 3, 4] and pf.h2d_bytes == 0
    with pytest.raises(ValueError):
        DevicePrefetcher([], torch.device("cpu"), depth=0)
