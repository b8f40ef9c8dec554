"""Counters, latency histograms and a JSONL sink (the reference only prints a wall-clock job
time, worker.py:219,234; see SURVEY.md §5 "Metrics / logging / observability")."""
from __future__ import annotations

import json
import threading
import time
from typing import Dict, List, Optional


class Metrics:
    def __init__(self, jsonl_path: Optional[str] = None):
        self.lock = threading.Lock()
        self.counters: Dict[str, float] = {}
        self.samples: Dict[str, List[float]] = {}
        self.jsonl_path = jsonl_path
        self.t0 = time.time()

    def inc(self, name: str, v: float = 1.0):
        with self.lock:
            self.counters[name] = self.counters.get(name, 0.0) + v

    def observe(self, name: str, v: float):
        with self.lock:
            self.samples.setdefault(name, []).append(float(v))

    def percentile(self, name: str, q: float) -> Optional[float]:
        with self.lock:
            xs = sorted(self.samples.get(name, ()))
        if not xs:
            return None
        k = min(len(xs) - 1, max(0, int(round(q / 100.0 * (len(xs) - 1)))))
        return xs[k]

    def summary(self) -> Dict:
        out = {"uptime_s": round(time.time() - self.t0, 3), "counters": dict(self.counters)}
        for name in list(self.samples):
            out[name] = {"n": len(self.samples[name]), "p50": self.percentile(name, 50), "p99": self.percentile(name, 99)}
        return out

    def emit(self, event: str, **fields):
        if not self.jsonl_path:
            return
        rec = {"t": round(time.time(), 6), "event": event, **fields}
        with self.lock, open(self.jsonl_path, "a") as f:
            f.write(json.dumps(rec) + "\n")


class Logger:
    """`log(msg)` prints iff verbose (server.py:26-28, worker.py:42-44)."""

    def __init__(self, verbose: bool = False, prefix: str = ""):
        self.verbose = verbose
        self.prefix = prefix

    def log(self, message):
        if self.verbose:
            print(f"{self.prefix}{message}", flush=True)

    def info(self, message):
        print(message, flush=True)
