"""nvidia-smi clock / throttle-reason sampler for benchmark hygiene (B200_PROFILING.md)."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples `nvidia-smi -lms 200` in a background process between start() and stop()."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 50):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._t: Optional[threading.Thread] = None

    def start(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if not exe:
            return self
        try:
            self.proc = subprocess.Popen([exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i",
                                          str(self.gpu_index), "-lms", str(self.period_ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return self

        def pump():
            assert self.proc is not None and self.proc.stdout is not None
            for line in self.proc.stdout:
                self.lines.append(line.strip())

        self._t = threading.Thread(target=pump, daemon=True)
        self._t.start()
        return self

    def stop(self) -> Dict:
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        if self._t is not None:
            self._t.join(timeout=2)
        sm, mx = [], []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": (statistics.median(sm) if sm else None), "sm_max_mhz": (max(mx) if mx else None),
                "reasons": sorted(reasons), "samples": len(sm)}
