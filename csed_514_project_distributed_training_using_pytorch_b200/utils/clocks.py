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


# NVML bit masks of nvmlDeviceGetCurrentClocksEventReasons
_NVML_BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20,
              "hw_thermal_slowdown": 0x40}


class ClockSampler:
    """Background clock poller for one GPU index. Uses NVML in-process (pynvml: a query costs
    tens of microseconds and forks nothing, so it does not disturb a sub-millisecond timed
    region the way an `nvidia-smi` process per sample does); falls back to nvidia-smi."""

    def __init__(self, gpu_index: int = 0, period_s: float = 0.1):
        self.gpu = gpu_index
        self.period = period_s
        self.samples = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._smi = shutil.which("nvidia-smi")
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = (pynvml, pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(gpu_index)))
            self.source = "nvml"
        except Exception:
            self.source = "nvidia-smi" if self._smi else None

    @staticmethod
    def _physical_index(i: int) -> int:
        import os
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if i < len(ids) and ids[i].isdigit():
                return int(ids[i])
        return i

    def _poll_nvml(self):
        nv, h = self._nvml
        try:
            fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
            mask = int(fn(h))
            self.samples.append({
                "sm": float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)),
                "max": float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)),
                "power": nv.nvmlDeviceGetPowerUsage(h) / 1000.0,
                "reasons": [r for r in _REASONS if mask & _NVML_BITS[r]],
            })
        except Exception:
            pass

    def _poll_once(self):
        if self._nvml is not None:
            return self._poll_nvml()
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
        if self._smi is None and self._nvml is None:
            return self
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> dict:
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=10)
        if not self.samples and (self._smi is not None or self._nvml is not None):
            self._poll_once()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = sorted({r for s in self.samples for r in s["reasons"]})
        return {"sm_mhz": statistics.median(s["sm"] for s in self.samples),
                "sm_max_mhz": max(s["max"] for s in self.samples),
                "reasons": reasons, "samples": len(self.samples), "source": self.source}
