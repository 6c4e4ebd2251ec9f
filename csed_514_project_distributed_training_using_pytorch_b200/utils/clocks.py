"""Sample SM clocks and throttle reasons with nvidia-smi while a timed region runs
(B200_PROFILING.md "clocks DURING the timed region")."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Optional

_Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
      "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
      "clocks_event_reasons.sw_power_cap")
_REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")


class ClockSampler:
    """Background nvidia-smi poller for one GPU index."""

    def __init__(self, gpu_index: int = 0, period_s: float = 0.1):
        self.gpu = gpu_index
        self.period = period_s
        self.samples = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._smi = shutil.which("nvidia-smi")

    def _poll_once(self):
        try:
            out = subprocess.run([self._smi, f"--query-gpu={_Q}", "--format=csv,noheader,nounits",
                                  "-i", str(self.gpu)], capture_output=True, text=True, timeout=5)
            parts = [p.strip() for p in out.stdout.strip().split(",")]
            if len(parts) >= 7:
                self.samples.append({
                    "sm": float(parts[0]), "max": float(parts[1]),
                    "power": float(parts[2]) if parts[2].replace(".", "").isdigit() else None,
                    "reasons": [r for r, v in zip(_REASONS, parts[3:7]) if v.lower() == "active"],
                })
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self._poll_once()
            self._stop.wait(self.period)

    def start(self):
        if self._smi is None:
            return self
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> dict:
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=10)
        if not self.samples and self._smi is not None:
            self._poll_once()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = sorted({r for s in self.samples for r in s["reasons"]})
        return {"sm_mhz": statistics.median(s["sm"] for s in self.samples),
                "sm_max_mhz": max(s["max"] for s in self.samples),
                "reasons": reasons, "samples": len(self.samples)}
