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


    def prometheus_text(self, prefix: str = "dvc_", extra_gauges: Optional[Dict[str, float]] = None) -> str:
        """Counters, p50/p99 summaries and caller-supplied gauges in the Prometheus text exposition format."""
        def clean(n: str) -> str:
            return prefix + "".join(ch if ch.isalnum() else "_" for ch in n)
        lines = []
        with self.lock:
            counters = dict(self.counters)
            names = list(self.samples)
        for name, v in sorted(counters.items()):
            lines += [f"# TYPE {clean(name)}_total counter", f"{clean(name)}_total {v:g}"]
        for name in sorted(names):
            n = len(self.samples.get(name, ()))
            lines.append(f"# TYPE {clean(name)} summary")
            for q in (50, 99):
                p = self.percentile(name, q)
                if p is not None:
                    lines.append(f'{clean(name)}{{quantile="{q / 100:g}"}} {p:g}')
            lines.append(f"{clean(name)}_count {n}")
        for name, v in sorted((extra_gauges or {}).items()):
            lines += [f"# TYPE {clean(name)} gauge", f"{clean(name)} {float(v):g}"]
        return "\n".join(lines) + "\n"

    def serve_http(self, port: int, host: str = "127.0.0.1", gauges=None):
        """Serve `/metrics` (Prometheus text) and `/summary` (JSON) from a daemon thread; returns the server
        (call `.shutdown()`), whose `.server_address[1]` is the bound port (pass 0 to pick a free one)."""
        import http.server

        metrics = self

        class Handler(http.server.BaseHTTPRequestHandler):
            def do_GET(self):          # noqa: N802
                if self.path.startswith("/metrics"):
                    body, ctype = metrics.prometheus_text(extra_gauges=gauges() if gauges else None), "text/plain; version=0.0.4"
                elif self.path.startswith("/summary"):
                    body, ctype = json.dumps(metrics.summary()), "application/json"
                else:
                    self.send_error(404)
                    return
                data = body.encode()
                self.send_response(200)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(data)))
                self.end_headers()
                self.wfile.write(data)

            def log_message(self, *a):   # quiet
                pass

        srv = http.server.ThreadingHTTPServer((host, port), Handler)
        threading.Thread(target=srv.serve_forever, daemon=True).start()
        return srv


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
