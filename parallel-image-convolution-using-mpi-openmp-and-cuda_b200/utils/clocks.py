"""SM clock / throttle-reason sampling during a timed region (B200_PROFILING.md recipe)."""
from __future__ import annotations

import statistics
import subprocess
import threading

_Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
      "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
      "clocks_event_reasons.sw_power_cap")
_NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index=0, period_s=0.1):
        self.gpu, self.period = gpu_index, period_s
        self.samples, self.max_mhz, self.reasons = [], None, set()
        self._stop = threading.Event()
        self._t = None

    def _poll_once(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={_Q}",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
            f = [x.strip() for x in out.strip().splitlines()[0].split(",")]
            self.samples.append(float(f[0]))
            self.max_mhz = float(f[1])
            for name, v in zip(_NAMES, f[3:7]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)
        except Exception:
            pass

    def _loop(self):
        while not self._stop.is_set():
            self._poll_once()
            self._stop.wait(self.period)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)
        if not self.samples:
            self._poll_once()

    def summary(self) -> dict:
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}
