"""GPU clock / throttle sampling during a timed region (the clocks line of B200_PROFILING.md).

Every performance number this repo reports (bench.py, scripts/bench_*.py) carries the record produced here: the median SM
clock under load, the maximum SM clock, and which throttle reasons were active while the kernels ran.  A number taken
under hw_slowdown / hw_thermal_slowdown / sw_thermal_slowdown is to be re-measured; sw_power_cap is normal on a 1 kW part.
"""
from __future__ import annotations

import subprocess
import threading
import time
from typing import Any, Dict, List, Optional

__all__ = ["ClockSampler"]


class ClockSampler:
    """`s = ClockSampler(gpu); s.start(); ...timed region...; record = s.stop()`."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REJECT = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.rows: List[List[str]] = []
        self.proc: Optional[subprocess.Popen] = None
        self.idx, self.period = gpu_index, period_ms

    def start(self) -> "ClockSampler":
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                                          str(self.period), "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> Dict[str, Any]:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(self.period / 1e3 + 0.05)
        self.proc.terminate()
        sm = sorted(int(r[1]) for r in self.rows if len(r) >= 9 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) >= 9 and r[2].isdigit()]
        pw = sorted(float(r[3]) for r in self.rows if len(r) >= 9 and r[3].replace(".", "", 1).isdigit())
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w": pw[len(pw) // 2] if pw else None, "reasons": sorted(reasons), "samples": len(sm),
                "rejected": any(r in reasons for r in self.REJECT)}
