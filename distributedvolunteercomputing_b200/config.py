"""One config dataclass with the reference's hard-coded defaults.

The reference has no flags: positional argv only (server.py:166-169, worker.py:341-344) and
class attributes (server.py:9-24, worker.py:14-40, args dict worker.py:186).  Every one of
those constants lives here with the same default; each can be overridden by an environment
variable ``DVC_<FIELD>`` or a ``--field value`` CLI flag *in addition to* the positional args.
"""
from __future__ import annotations

import dataclasses
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple


@dataclass
class Config:
    # ---- network endpoints (server.py:11,18 ; worker.py:23,25)
    server_ip: str = "localhost"
    own_ip: str = "localhost"
    control_port: int = 9999
    my_port: int = 5554                 # reference: fixed '5554' (one client per host); 0 = auto-pick a free port
    data_port_lo: int = 5555
    data_port_hi: int = 5599            # reference pool 5555..5599; the pool grows past it on demand
    transport: str = "req_rep"          # reference req_rep=True; the PUB/SUB branch is dead code there
    control_retry_s: float = 10.0       # worker.py:56 select timeout
    control_poll_s: float = 2.0         # server.py:98 socket timeout
    data_timeout_s: float = 5.0         # new: REQ/REP never blocks forever
    # ---- chunking / buffers (worker.py:16-17 ; server.py:12)
    number_of_frames_in_chunk: int = 100
    max_buffer_frames: int = 4000
    max_buffer_chunks: int = 40
    # ---- model / inference (worker.py:186-197)
    prototxt: str = "MobileNetSSD_deploy.prototxt.txt"
    model: str = "MobileNetSSD_deploy.caffemodel"
    confidence: float = 0.2
    consider: Tuple[str, ...] = ("person",)
    mean: Tuple[float, float, float] = (127.5, 0.0, 0.0)   # channel-0-only quirk; canonical = (127.5,)*3
    scale: float = 0.007843
    out_width: int = 400                # worker.py:108,243
    writer_fps: int = 30                # worker.py:110
    # ---- behaviour switches (Appendix C)
    compat_drop_first: bool = False     # reference consumes frame 0 for sizing (worker.py:107,112)
    compat_drop_tail: bool = False      # reference never sends the last partial chunk (worker.py:125-126)
    requester_works: bool = False       # reference excludes the requester from the pool (server.py:84-87)
    dispatch_policy: str = "least_loaded"   # or "round_robin" (reference, server.py:89-90)
    worker_timeout_s: float = 10.0      # in-flight batches older than this are re-dispatched
    verbose: bool = False
    # ---- fault injection (the reference has none)
    pipelined_worker: bool = True       # GPU volunteers overlap upload / kernels / read-back of consecutive chunks (process_stream)
    inject_stall_after: int = 0         # worker: after this many chunks, keep accepting chunks but never answer
    inject_leave_after: int = 0         # worker: after this many chunks, vanish without sending `end` (process death)
    # ---- B200 engine
    device: str = "auto"                # "auto" | "cpu" | "cuda:N"
    batch: int = 64                     # frames per GPU batch (BASELINE.json config 2)
    sub_batch: int = 0                  # 0 = whole batch per launch; >0 runs conv0..conv6 in sub-batches of this size
    use_cuda_graph: bool = True
    max_cached_states: int = 64         # per-(batch, frame size, requester) device states kept by one engine (LRU)
    metrics_jsonl: Optional[str] = None
    metrics_port: int = -1              # >= 0: serve /metrics (Prometheus text) and /summary on this port (0 = any free port)

    def validate(self) -> "Config":
        """Reject settings that would silently do something else than asked."""
        if self.transport != "req_rep":
            # the reference's REQ_REP=False branch connects to tcp://*:port, which ZeroMQ refuses (server.py:44-46,117-123;
            # worker.py:65-67,165-167): it never worked there, so there is no behaviour to be compatible with
            raise ValueError(f"transport={self.transport!r}: only 'req_rep' exists (the reference's PUB/SUB branch is dead code)")
        if self.dispatch_policy not in ("least_loaded", "round_robin"):
            raise ValueError(f"dispatch_policy={self.dispatch_policy!r}: use 'least_loaded' or 'round_robin'")
        if self.number_of_frames_in_chunk < 1 or self.batch < 1:
            raise ValueError("number_of_frames_in_chunk and batch must be positive")
        return self

    @staticmethod
    def from_env(**overrides) -> "Config":
        cfg = Config()
        for f in dataclasses.fields(Config):
            env = os.environ.get("DVC_" + f.name.upper())
            if env is not None:
                setattr(cfg, f.name, _coerce(env, getattr(cfg, f.name)))
        for k, v in overrides.items():
            if v is not None:
                setattr(cfg, k, v)
        return cfg

    @staticmethod
    def from_argv(argv: Sequence[str], positional: Sequence[str]) -> "Config":
        """Positional args keep working exactly as in the reference; ``--name value`` pairs and
        ``--flag`` booleans may follow or precede them."""
        cfg = Config.from_env()
        pos: List[str] = []
        i = 0
        names = {f.name: f for f in dataclasses.fields(Config)}
        while i < len(argv):
            a = argv[i]
            if a.startswith("--"):
                key = a[2:].replace("-", "_")
                if key not in names:
                    raise SystemExit(f"unknown option {a}")
                cur = getattr(cfg, key)
                if isinstance(cur, bool):
                    setattr(cfg, key, True)
                    i += 1
                else:
                    if i + 1 >= len(argv):
                        raise SystemExit(f"option {a} needs a value")
                    setattr(cfg, key, _coerce(argv[i + 1], cur))
                    i += 2
            else:
                pos.append(a)
                i += 1
        for name, val in zip(positional, pos):
            setattr(cfg, name, _coerce(val, getattr(cfg, name)))
        return cfg

    @property
    def addr(self) -> str:
        return f"{self.own_ip}:{self.my_port}"


def _coerce(text: str, like):
    if isinstance(like, bool):
        return text.lower() in ("1", "true", "yes", "on")
    if isinstance(like, int):
        return int(text)
    if isinstance(like, float):
        return float(text)
    if isinstance(like, tuple):
        parts = [p for p in text.replace(";", ",").split(",") if p != ""]
        if like and isinstance(like[0], float):
            return tuple(float(p) for p in parts)
        return tuple(parts)
    if like is None:
        return text
    return type(like)(text)
