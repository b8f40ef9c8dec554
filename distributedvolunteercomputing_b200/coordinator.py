"""`coordinator`: the server process (public surface of the reference's server.py).

Same wire behaviour as the reference class of the same name (server.py:7-163) - UDP membership
on port 9999, one REP data port per client from the 5555.. pool, chunks relayed as imagezmq
messages - but the distribution loop is `parallel.scheduler.Scheduler`:

  * chunks wait in per-requester queues until a worker is eligible (the reference drops them,
    server.py:82,88);
  * every dispatched chunk stays in the in-flight table until all of its frames have come back;
    when its owner leaves (`end`) or stops answering, the chunk is re-dispatched, so a job no
    longer hangs when a volunteer quits mid-stream;
  * `join` is idempotent, threads block on queues/conditions instead of spinning, every socket
    call has a timeout so `quit` actually terminates.
"""
from __future__ import annotations

import queue
import socket
import threading
import time
from typing import Dict, List, Optional, Set

import numpy as np

from . import protocol as proto
from .config import Config
from .parallel.scheduler import Scheduler
from .utils.metrics import Logger, Metrics


class coordinator:  # noqa: N801  (name kept from the reference's public surface)
    def __init__(self, ip: str = "localhost", cfg: Optional[Config] = None, start: bool = True):
        self.cfg = (cfg or Config.from_env()).validate()
        self.my_ip = f"{ip}:{self.cfg.control_port}"
        self.logger = Logger(self.cfg.verbose, "[server] ")
        self.log = self.logger.log
        self.metrics = Metrics(self.cfg.metrics_jsonl)
        self.sched = Scheduler(policy=self.cfg.dispatch_policy, requester_works=False,
                               worker_timeout_s=self.cfg.worker_timeout_s)
        self.sched.on_drop = self._on_drop          # a chunk given up after max_attempts is reported, not lost silently
        self.lock = threading.RLock()
        self.running = False
        self.free_ports: List[int] = list(range(self.cfg.data_port_lo, self.cfg.data_port_hi + 1))
        self.port_of: Dict[str, int] = {}
        self.senders: Dict[str, proto.ImageSender] = {}
        self.sender_locks: Dict[str, threading.Lock] = {}
        self.ingest_stop: Dict[str, threading.Event] = {}
        self.outbox: Dict[str, "queue.Queue"] = {}      # per-worker dispatch queue: one slow link does not stall the rest
        self.results: Dict[str, "queue.Queue"] = {}
        self.reply_stop: Dict[str, threading.Event] = {}
        # frame bookkeeping: requester -> frame number -> batch id ; batch id -> frames still missing
        self.frame_owner: Dict[str, Dict[int, int]] = {}
        self.missing: Dict[int, Set[int]] = {}
        self.threads: List[threading.Thread] = []
        if start:
            self.start()

    # ------------------------------------------------------------------ lifecycle
    def start(self):
        self.running = True
        self.sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        self.sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self.sock.bind(("", self.cfg.control_port))
        self.sock.settimeout(min(self.cfg.control_poll_s, 0.5))
        print("\nlistening on %s port %s" % ("", self.cfg.control_port), flush=True)
        for target in (self.manager, self.send_request, self.watchdog):
            t = threading.Thread(target=target, daemon=True)
            t.start()
            self.threads.append(t)
        self.metrics_http = None
        if self.cfg.metrics_port >= 0:
            self.metrics_http = self.metrics.serve_http(self.cfg.metrics_port, gauges=self._gauges)
            print("metrics on http://127.0.0.1:%d/metrics" % self.metrics_http.server_address[1], flush=True)

    def _gauges(self) -> Dict[str, float]:
        snap = self.sched.snapshot()
        return {"workers_alive": sum(1 for w in snap["workers"].values() if w["alive"]), "membership_epoch": snap["epoch"],
                "chunks_pending": snap["pending"], "chunks_in_flight": snap["inflight"],
                "chunks_redispatched": snap["stats"].get("redispatched", 0)}

    def exit_threads(self):
        """`quit` (server.py:160-163) - and the threads really stop."""
        self.running = False
        for snd in list(self.senders.values()):
            snd.abort.set()                    # sender threads blocked on a reply give up within 250 ms
        if getattr(self, "metrics_http", None) is not None:
            self.metrics_http.shutdown()
            self.metrics_http.server_close()
            self.metrics_http = None
        for ev in list(self.ingest_stop.values()) + list(self.reply_stop.values()):
            ev.set()
        with self.sched.cv:
            self.sched.cv.notify_all()
        for t in self.threads:
            t.join(timeout=2.0)
        try:
            self.sock.close()
        except Exception:
            pass
        for s in list(self.senders.values()):
            s.close()

    # ------------------------------------------------------------------ control plane
    def manager(self):
        while self.running:
            try:
                message, client_address = self.sock.recvfrom(4096)
            except socket.timeout:
                continue
            except OSError:
                break
            cmd, address = proto.parse_control(message)
            try:
                reply = self.handle_control(cmd, address)
            except Exception as e:  # never let one bad datagram kill membership
                self.log(f"control error {cmd} {address}: {e!r}")
                reply = None
            if reply is not None:
                self.sock.sendto(reply.encode("utf-8"), client_address)
        print("manager terminated.", flush=True)

    def handle_control(self, cmd: Optional[str], address: str) -> Optional[str]:
        if cmd == "join":
            with self.lock:
                known = address in self.port_of
            if known and not any(w.addr == address for w in self.sched.members.alive()):
                # a worker we had declared dead (send timeout / watchdog) is back, e.g. restarted on the same
                # ip:port: drop its old port, sockets and threads and start from a clean slate
                self._remove_client(address)
            with self.lock:
                _, new = self.sched.join(address, capacity=2)
                if not new and address in self.port_of:
                    return "ok||" + str(self.port_of[address])          # idempotent retry
                port = self._alloc_port()
                self.port_of[address] = port
                host, cport = address.split(":")
                print("sending to : tcp://" + host + ":" + cport, flush=True)
                self.senders[address] = proto.ImageSender("tcp://" + host + ":" + cport, timeout_s=self.cfg.data_timeout_s)
                self.sender_locks[address] = threading.Lock()
                ev = threading.Event()
                self.ingest_stop[address] = ev
                t = threading.Thread(target=self.server, args=["", port, address, ev], daemon=True)
                t.start()
                self.threads.append(t)
                self.outbox[address] = queue.Queue()
                t = threading.Thread(target=self.worker_sender, args=[address, ev], daemon=True)
                t.start()
                self.threads.append(t)
            self.metrics.inc("joins")
            return "ok||" + str(port)
        if cmd == "request":
            if address not in self.port_of:
                return None
            self.sched.become_requester(address)
            with self.lock:
                self.frame_owner.setdefault(address, {})
                if address not in self.results:
                    self.results[address] = queue.Queue()
                    ev = threading.Event()
                    self.reply_stop[address] = ev
                    t = threading.Thread(target=self.send_reply, args=[address, ev], daemon=True)
                    t.start()
                    self.threads.append(t)
            return "ok"
        if cmd == "stop":
            self.sched.stop_requesting(address)
            return "ok"
        if cmd == "end":
            self._remove_client(address)
            return "ok"
        return None

    def _alloc_port(self) -> int:
        if self.free_ports:
            return self.free_ports.pop(0)
        return proto.free_tcp_port()          # the reference stops at 45 clients (server.py:18); we keep going

    def _remove_client(self, address: str):
        requeued = self.sched.leave(address)
        if requeued:
            self.metrics.inc("redispatched_chunks", len(requeued))
            self.log(f"{address} left with {len(requeued)} chunk(s) in flight: re-dispatching")
        with self.lock:
            port = self.port_of.pop(address, None)
            if port is not None and self.cfg.data_port_lo <= port <= self.cfg.data_port_hi:
                self.free_ports.append(port)
            ev = self.ingest_stop.pop(address, None)
            if ev:
                ev.set()
            ev = self.reply_stop.pop(address, None)
            if ev:
                ev.set()
            s = self.senders.pop(address, None)
            slock = self.sender_locks.pop(address, None)
            self.outbox.pop(address, None)
            self.results.pop(address, None)
        if s is not None:
            # ZeroMQ sockets are not thread-safe: worker_sender / send_reply may be inside send_image on this one, so it
            # is closed under the same lock they hold (outside self.lock: a send may take its full time-out)
            if slock is not None:
                with slock:
                    s.close()
            else:
                s.close()
        if address in self.frame_owner:
            self.sched.drop_requester(address)
            with self.lock:
                owner = self.frame_owner.pop(address, None) or {}
                for bid in set(owner.values()):
                    self.missing.pop(bid, None)

    def _on_drop(self, batch):
        """The scheduler gave a chunk up (max_attempts): tell the requester which frames will never come, so its
        in-order writer can move on instead of waiting for ever, and forget the chunk."""
        self.metrics.inc("dropped_chunks")
        self.log(f"chunk {batch.batch_id} of {batch.requester} dropped after {batch.attempts} attempts: frames {batch.frame_ids[:3]}...")
        with self.lock:
            self.missing.pop(batch.batch_id, None)
            owner = self.frame_owner.get(batch.requester, {})
            for n in batch.frame_ids:
                owner.pop(n, None)
            q = self.results.get(batch.requester)
        if q is not None:
            gap = proto.Header(batch.requester, "dropped", tuple(batch.frame_ids), 1, 1)
            q.put((gap.encode(), np.zeros((1, len(batch.frame_ids), 3), np.uint8)))

    # ------------------------------------------------------------------ data plane: ingest
    def server(self, ip: str, port: int, address: str, stop: threading.Event):
        print("receiving at : tcp://*:" + str(port), flush=True)
        hub = proto.ImageHub("tcp://*:" + str(port))
        try:
            while self.running and not stop.is_set():
                # one bad message (undecodable JSON / shape, malformed header) must not end this client's ingest thread:
                # log it, count it, answer it (the peer blocks on the reply) and keep serving
                try:
                    got = hub.recv_image(timeout_s=0.25)
                except proto.BadMessage as e:
                    self.metrics.inc("bad_messages")
                    self.log(f"{address}: {e}")
                    hub.send_reply(b"OK")
                    continue
                if got is None:
                    continue
                info, frame = got
                self.log(info)
                try:
                    self._on_chunk(info, frame, address)
                except Exception as e:
                    self.metrics.inc("bad_messages")
                    self.log(f"{address}: chunk {info[:80]!r} rejected: {e!r}")
                finally:
                    hub.send_reply(b"OK")
        finally:
            hub.close()
            print("server terminated.", flush=True)

    def _on_chunk(self, info: str, frame: np.ndarray, sender_addr: str):
        hdr = proto.parse_header(info)
        self.sched.members.touch(sender_addr)
        if self.sched.revive(sender_addr):              # presumed dead (silence / send timeout) but evidently not
            self.metrics.inc("workers_revived")
            self.log(f"{sender_addr} is back: accepting work again")
        if hdr.cmd == "request":
            while self.running and self.sched.pending() > self.cfg.max_buffer_chunks:
                time.sleep(0.02)                                        # backpressure (server.py:54)
            b = self.sched.submit(hdr.requester, hdr.frames, payload=(info, np.array(frame, copy=True)))
            with self.lock:
                owner = self.frame_owner.setdefault(hdr.requester, {})
                for n in hdr.frames:
                    owner[n] = b.batch_id
                self.missing[b.batch_id] = set(hdr.frames)
            self.metrics.inc("frames_in", len(hdr.frames))
        elif hdr.cmd == "processed":
            fresh = False
            finished: List[int] = []
            with self.lock:
                owner = self.frame_owner.get(hdr.requester, {})
                for n in hdr.frames:
                    bid = owner.get(n)
                    if bid is None:
                        continue
                    miss = self.missing.get(bid)
                    if miss is not None and n in miss:
                        miss.discard(n)
                        fresh = True
                        if not miss:
                            finished.append(bid)
                            del self.missing[bid]
                    if miss is None or n not in miss:
                        owner.pop(n, None)          # delivered: the table holds in-flight frames only (`request live`)
                q = self.results.get(hdr.requester)
            for bid in finished:
                self.sched.complete(bid, sender_addr)
            if fresh and q is not None:
                q.put((info, np.array(frame, copy=True)))
                self.metrics.inc("frames_out", len(hdr.frames))
            elif not fresh:
                self.metrics.inc("duplicate_chunks")

    # ------------------------------------------------------------------ data plane: dispatch / return
    def send_request(self):
        """Dispatcher (server.py:77-91): takes scheduling decisions and hands each chunk to its worker's own sender
        thread, so a slow or dead link never blocks dispatch to the others (the reference's single loop blocks in
        `send_image`)."""
        while self.running:
            d = self.sched.wait_dispatch(timeout=0.2)
            if d is None:
                continue
            batch, worker, epoch = d
            box = self.outbox.get(worker.addr)
            if box is None:
                self.sched.fail(batch.batch_id, "no_sender")
                continue
            box.put(batch)
        print("send_request terminated.", flush=True)

    def worker_sender(self, address: str, stop: threading.Event):
        """One per worker: blocking REQ/REP sends of the chunks the dispatcher assigned to `address`."""
        while self.running and not stop.is_set():
            box = self.outbox.get(address)
            if box is None:
                break
            try:
                batch = box.get(timeout=0.25)
            except queue.Empty:
                continue
            info, frame = batch.payload
            sender = self.senders.get(address)
            lock = self.sender_locks.get(address)
            if sender is None or lock is None:
                self.sched.fail(batch.batch_id, "no_sender")
                continue
            try:
                with lock:
                    sender.send_image(info, frame)
                self.sched.touch(address)          # the worker acknowledged the chunk: it is alive as of now
                self.metrics.inc("chunks_dispatched")
            except TimeoutError:
                self.log(f"worker {address} did not acknowledge a chunk: presuming it dead")
                self.metrics.inc("dispatch_timeouts")
                self.sched.leave(address, reason="timeout")     # re-queues this batch and anything else it held
                while True:                                     # ...including what still waits in this outbox
                    try:
                        box.get_nowait()
                    except queue.Empty:
                        break

    def send_reply(self, address: str, stop: threading.Event):
        while self.running and not stop.is_set():
            q = self.results.get(address)
            if q is None:
                break
            try:
                info, frame = q.get(timeout=0.25)
            except queue.Empty:
                continue
            sender = self.senders.get(address)
            lock = self.sender_locks.get(address)
            if sender is None or lock is None:
                break
            for attempt in range(3):
                try:
                    with lock:
                        sender.send_image(info, frame)
                    break
                except TimeoutError:
                    self.metrics.inc("reply_timeouts")
        print("send_reply terminated.", flush=True)

    def watchdog(self):
        while self.running:
            time.sleep(0.5)
            re = self.sched.check_timeouts()
            if re:
                self.metrics.inc("redispatched_chunks", len(re))
                self.log(f"timeout: re-dispatching {len(re)} chunk(s)")


def main(argv: Optional[List[str]] = None):
    """`python server.py [ip]` then `quit` on stdin (server.py:165-175)."""
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    cfg = Config.from_argv(argv, positional=("server_ip",))
    c = coordinator(cfg.server_ip, cfg)
    while True:
        try:
            a = input("\nEnter quit to exit\n")
        except EOFError:
            a = "quit"
        if a == "quit":
            c.exit_threads()
            print("done.", flush=True)
            break
