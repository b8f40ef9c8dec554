"""Wire formats of the public surface, kept byte-compatible.

Control plane  : UDP text datagrams to the server's port 9999
                 ``join||addr`` ``request||addr`` ``stop||addr`` ``end||addr`` -> ``ok||<port>`` / ``ok``
                 (reference: server.py:100-156, worker.py:50-68,287-298,305-316,321-332).
Data plane     : ZeroMQ REQ/REP, 2-part messages: JSON ``{"msg","dtype","shape"}`` + raw array
                 bytes, answered with ``b"OK"`` (the imagezmq wire the reference rides on,
                 server.py:43-62,73,89; worker.py:64,157,164-180).
Chunk header   : ``id||cmd||n1-n2-...||H||W`` with the k frames side by side in one
                 ``[H, k*W, 3]`` uint8 image (worker.py:136-157,172-177).

Differences from the reference are confined to robustness: every blocking socket call has a
timeout (the reference hangs forever on a dead peer and on quit), and ports are configurable.
"""
from __future__ import annotations

import socket
import threading
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

SEP = "||"
CONTROL_COMMANDS = ("join", "request", "stop", "end")   # matched by substring, in this order (server.py:104-141)
DATA_COMMANDS = ("request", "processed", "dropped")   # "dropped": gap marker, server -> requester (not in the reference)


# ------------------------------------------------------------------------------ control plane
def make_control(cmd: str, addr: str) -> bytes:
    assert cmd in CONTROL_COMMANDS
    return f"{cmd}{SEP}{addr}".encode("utf-8")


def parse_control(data: bytes) -> Tuple[Optional[str], str]:
    """Return (command, addr). Command matching follows the reference: first substring hit in
    the order join, request, stop, end; ``None`` if nothing matches."""
    msg = data.decode("utf-8", "replace")
    cmd = next((c for c in CONTROL_COMMANDS if c in msg), None)
    parts = msg.split(SEP)
    addr = parts[1] if len(parts) > 1 else ""
    return cmd, addr


def control_roundtrip(sock: socket.socket, server: Tuple[str, int], payload: bytes, retry_s: float = 10.0,
                      max_tries: Optional[int] = None, log=None) -> str:
    """Send a control datagram and wait for the reply, resending every ``retry_s`` seconds
    (worker.py:50-58).  ``max_tries=None`` retries forever like the reference."""
    import select
    i = 0
    while max_tries is None or i < max_tries:
        if log:
            log(f"sending message trial {i}...")
        i += 1
        sock.sendto(payload, server)
        ready = select.select([sock], [], [], retry_s)
        if ready[0]:
            data, _ = sock.recvfrom(4096)
            return data.decode("utf-8")
    raise TimeoutError(f"no reply from {server} to {payload!r}")


# ------------------------------------------------------------------------------ chunk header
@dataclass
class Header:
    requester: str
    cmd: str
    frames: List[int]
    height: int
    width: int

    def encode(self) -> str:
        return SEP.join([self.requester, self.cmd, "-".join(str(n) for n in self.frames), str(self.height), str(self.width)])

    def tag(self, n: int) -> str:
        """Per-frame tag used inside processes: id||cmd||n (worker.py:115,176,280)."""
        return SEP.join([self.requester, self.cmd, str(n)])


def parse_header(text: str) -> Header:
    p = text.split(SEP)
    if len(p) < 5:
        raise ValueError(f"bad chunk header {text!r}")
    frames = [int(x) for x in p[2].split("-") if x != ""]
    return Header(p[0], p[1], frames, int(p[3]), int(p[4]))


def pack_chunk(frames: Sequence[np.ndarray]) -> np.ndarray:
    """k frames [H,W,3] -> one [H, k*W, 3] image (worker.py:141-156 does this with PIL paste)."""
    return np.ascontiguousarray(np.concatenate(list(frames), axis=1))


def unpack_chunk(img: np.ndarray, k: int, height: int, width: int) -> List[np.ndarray]:
    """Inverse of pack_chunk: frame i = img[0:H, i*W:(i+1)*W] (worker.py:175)."""
    return [img[0:height, i * width:(i + 1) * width] for i in range(k)]


# ------------------------------------------------------------------------------ data plane
class ImageSender:
    """imagezmq.ImageSender(connect_to, REQ_REP=True) equivalent with a reply timeout."""

    def __init__(self, connect_to: str, timeout_s: float = 5.0):
        import zmq
        self._zmq = zmq
        self.ctx = zmq.Context.instance()
        self.connect_to = connect_to
        self.timeout_s = timeout_s
        self.abort = threading.Event()      # set by another thread (quit): a pending reply wait gives up within 250 ms
        self._open()

    def _open(self):
        zmq = self._zmq
        self.sock = self.ctx.socket(zmq.REQ)
        self.sock.setsockopt(zmq.LINGER, 0)
        self.sock.connect(self.connect_to)

    def send_image(self, msg: str, image: np.ndarray, timeout_s: Optional[float] = None) -> bytes:
        zmq = self._zmq
        image = np.ascontiguousarray(image)
        md = dict(msg=msg, dtype=str(image.dtype), shape=image.shape)
        self.sock.send_json(md, zmq.SNDMORE)
        self.sock.send(image, copy=False)
        t = self.timeout_s if timeout_s is None else timeout_s
        left_ms = int(t * 1000)
        while left_ms > 0 and not self.abort.is_set():      # sliced so `quit` never waits for the whole timeout
            step = min(250, left_ms)
            if self.sock.poll(step, zmq.POLLIN):
                return self.sock.recv()
            left_ms -= step
        # a REQ socket that missed its reply is wedged: rebuild it so the caller can retry elsewhere
        self.sock.close(0)
        self._open()
        raise TimeoutError(f"no reply from {self.connect_to} within {t}s")

    def close(self):
        try:
            self.sock.close(0)
        except Exception:
            pass


class BadMessage(ValueError):
    """A message that cannot be decoded (bad JSON, dtype, shape or header).  The hub may - and must - still reply."""


class ImageHub:
    """imagezmq.ImageHub(open_port, REQ_REP=True) equivalent; recv_image takes a timeout."""

    def __init__(self, open_port: str = "tcp://*:5555"):
        import zmq
        self._zmq = zmq
        self.ctx = zmq.Context.instance()
        self.sock = self.ctx.socket(zmq.REP)
        self.sock.setsockopt(zmq.LINGER, 0)
        self.sock.bind(open_port)
        self.open_port = open_port

    def recv_image(self, timeout_s: Optional[float] = None):
        zmq = self._zmq
        if timeout_s is not None and not self.sock.poll(int(timeout_s * 1000), zmq.POLLIN):
            return None
        # take the whole message off the socket first: a REP socket that has received must reply before it can receive
        # again, so a malformed message has to leave us in the "may reply" state, not half-read
        parts = self.sock.recv_multipart(copy=False)
        try:
            import json
            md = json.loads(bytes(parts[0].buffer))
            img = np.frombuffer(parts[1].buffer, dtype=md["dtype"]).reshape(md["shape"])
            return md["msg"], img
        except Exception as e:
            raise BadMessage(f"malformed imagezmq message ({len(parts)} part(s)): {e!r}") from e

    def send_reply(self, reply: bytes = b"OK"):
        self.sock.send(reply)

    def close(self):
        try:
            self.sock.close(0)
        except Exception:
            pass


def free_tcp_port(host: str = "127.0.0.1") -> int:
    s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    s.bind((host, 0))
    port = s.getsockname()[1]
    s.close()
    return port
