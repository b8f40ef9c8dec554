"""nvidia-smi clock / throttle-reason sampler for benchmark hygiene (B200_PROFILING.md)."""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clock + throttle reasons between start() and stop().

    Backend ``smi`` (default): `nvidia-smi --query-gpu ... -lms <period>` in a child process (the recipe's clocks
    line).  Backend ``nvml``: an in-process thread reading two NVML counters per period; measured on 2 GPUs it
    slows a host-driven launch loop by ~13% (171.6k -> 148.4k frames/s, any period), so it is opt-in.
    ``DVC_CLOCK_BACKEND`` / ``DVC_CLOCK_PERIOD_MS`` override.
    """

    def __init__(self, gpu_index: int = 0, period_ms: Optional[int] = None, backend: Optional[str] = None,
                 pci_bus_id: Optional[str] = None):
        self.gpu_index = gpu_index
        self.period_ms = int(period_ms if period_ms is not None else os.environ.get("DVC_CLOCK_PERIOD_MS", 50))
        self.backend = backend or os.environ.get("DVC_CLOCK_BACKEND", "smi")
        self.pci_bus_id = pci_bus_id
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._t: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._nvml_samples: List[tuple] = []
        self._nvml_max = None

    # ------------------------------------------------------------------ NVML backend
    def _start_nvml(self) -> bool:
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = None
            if self.pci_bus_id:
                try:
                    h = nv.nvmlDeviceGetHandleByPciBusId(self.pci_bus_id.encode())
                except Exception:
                    h = None
            if h is None:
                idx = self.gpu_index
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                if vis and all(v.strip().isdigit() for v in vis.split(",")) and idx < len(vis.split(",")):
                    idx = int(vis.split(",")[idx])
                h = nv.nvmlDeviceGetHandleByIndex(idx)
            self._nvml_max = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            return False

        def pump():
            period = self.period_ms / 1e3
            while not self._stop.is_set():
                try:
                    self._nvml_samples.append((float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)),
                                               int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))))
                except Exception:
                    break
                self._stop.wait(period)

        self._t = threading.Thread(target=pump, daemon=True)
        self._t.start()
        return True

    def _stop_nvml(self) -> Dict:
        import pynvml as nv
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=2)
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        sm = [s for s, _ in self._nvml_samples]
        reasons = sorted({name for _, r in self._nvml_samples for name, bit in bits.items() if r & bit})
        return {"sm_mhz": (statistics.median(sm) if sm else None), "sm_max_mhz": self._nvml_max,
                "reasons": reasons, "samples": len(sm), "backend": "nvml", "period_ms": self.period_ms}

    def start(self) -> "ClockSampler":
        if self.backend == "nvml" and self._start_nvml():
            return self
        self.backend = "smi"
        self.period_ms = max(self.period_ms, 50)
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

    def mark(self) -> "ClockSampler":
        """Samples taken before this call are ignored by stop(): start the sampler early (its start-up cost stays out
        of the timed region), mark right before the region."""
        self._mark = len(self._nvml_samples) if self.backend == "nvml" else len(self.lines)
        return self

    def stop(self) -> Dict:
        if self.backend == "nvml":
            self._nvml_samples = self._nvml_samples[getattr(self, "_mark", 0):] if not self._stop.is_set() else self._nvml_samples
            return self._stop_nvml()
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
        for ln in self.lines[getattr(self, "_mark", 0):]:
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
                "reasons": sorted(reasons), "samples": len(sm), "backend": "nvidia-smi", "period_ms": self.period_ms}
